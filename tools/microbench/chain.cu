// chain.cu — floors of a lone digest chain on one SM sub-partition (the few-large-messages regime).
//
//   issue rates   one warp, independent instructions: ALU only, FMA-pipe only (IMAD), ALU/FMA alternating — how fast can
//                 ONE warp issue, whatever the pipe?
//   md5 step      the dependent LOP3 -> IADD3 -> LEA.HI triple of an MD5 step, isolated: cycles per step
//   compress      the shipped compression functions (digest_device.cuh) on register-resident / shared-memory-resident
//                 input, no staging at all: cycles per 64-byte block for MD5, SHA-1, SHA-256 and for the chain-warp
//                 halves (Assist<>::rounds on a pre-expanded schedule in shared memory)
// Compare with the kernels' measured cycles per block (bench.py: ms_per_step * 1.965e6 / blocks) to see what the
// staging / hand-over around the chain costs.   nvcc -arch=sm_100a -O3 -o chain chain.cu && ./chain
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#include "../../downloader_b200/csrc/digest_device.cuh"

using namespace th;
#define NBLK 4096

__device__ __forceinline__ void opaque(uint32_t& x) { asm volatile("" : "+r"(x)); }

template <int MODE>
__global__ void issue_k(uint32_t* out, uint32_t one, uint32_t seed, long long* cyc) {
  uint32_t a[8];
#pragma unroll
  for (int i = 0; i < 8; i++) a[i] = seed * (2 * i + 3) + threadIdx.x;
  const uint32_t b = seed ^ 0x1234567u, c = seed * 77u;
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < NBLK; it++) {
#pragma unroll
    for (int r = 0; r < 4; r++) {
#pragma unroll
      for (int i = 0; i < 8; i++) {
        if (MODE == 0) a[i] = lop3<LUT_XOR3>(a[i], b, c);                        // 8 independent ALU chains
        if (MODE == 1) a[i] = add_fma(a[i], b, one);                              // 8 independent IMAD chains
        if (MODE == 2) a[i] = (i & 1) ? add_fma(a[i], b, one) : lop3<LUT_XOR3>(a[i], b, c);  // alternate pipes
        if (MODE == 3) a[i] = (i & 1) ? rotl(a[i], 7) : lop3<LUT_XOR3>(a[i], b, c);          // SHF / LOP3, both ALU
      }
    }
  }
  const long long t1 = clock64();
  uint32_t x = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) x ^= a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = x;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// the dependent part of an MD5 step, nothing else: f = F(b,c,d); t = a + f + k; b' = b + rotl(t, s)
__global__ void md5_step_k(uint32_t* out, uint32_t one, uint32_t seed, long long* cyc) {
  uint32_t a = seed + threadIdx.x, b = seed * 3, c = seed * 5, d = seed * 7;
  const uint32_t k = seed ^ 0x5a5a5a5au;
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < NBLK; it++) {
#pragma unroll
    for (int r = 0; r < 16; r++) {
      uint32_t t;
      t = a + k + lop3<LUT_CH>(b, c, d); a = b + rotl(t, 7);
      t = d + k + lop3<LUT_CH>(a, b, c); d = a + rotl(t, 12);
      t = c + k + lop3<LUT_CH>(d, a, b); c = d + rotl(t, 17);
      t = b + k + lop3<LUT_CH>(c, d, a); b = c + rotl(t, 22);
    }
  }
  const long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = a ^ b ^ c ^ d;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  (void)one;
}

// MODE 0 md5 compress, 1 sha1 compress, 2 sha256 compress (message words in registers, made opaque every block);
// 10+V sha1 chain-warp rounds variant V, 20+V sha256 chain-warp rounds variant V (pre-expanded W+K read from shared
// memory with LDS.128); 30 sha1 helper expansion, 40+V sha256 helper expansion variant V (written to shared memory),
// 50+V the same after an MD5 compression of the block (one warp doing both: the two costs simply add up)
template <int MODE>
__global__ void compress_k(uint32_t* out, uint32_t one, uint32_t seed, long long* cyc) {
  __shared__ uint4 ring[80 * 32];
  const int lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 80 * 32; i += blockDim.x) ring[i] = make_uint4(seed + i, seed * i, seed ^ i, i);
  __syncthreads();
  uint32_t h[8], w[16];
#pragma unroll
  for (int i = 0; i < 8; i++) h[i] = seed * (i + 1) + lane;
#pragma unroll
  for (int i = 0; i < 16; i++) w[i] = seed * (i + 11) ^ lane;
  const uint32_t dst = smem_u32(ring) + lane * 16;
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < NBLK; it++) {
    if (MODE <= 2 || MODE >= 30) {
#pragma unroll
      for (int i = 0; i < 16; i++) opaque(w[i]);
    }
    if (MODE == 0) Md5::compress(h, w, one);
    if (MODE == 1) { uint32_t v[16]; for (int i = 0; i < 16; i++) v[i] = w[i]; Sha1::compress(h, v, one); }
    if (MODE == 2) { uint32_t v[16]; for (int i = 0; i < 16; i++) v[i] = w[i]; Sha256::compress(h, v, one); }
    if (MODE >= 10 && MODE < 20) Assist<Sha1>::rounds<MODE - 10, 80>(h, ring + lane, one);
    if (MODE >= 20 && MODE < 30) Assist<Sha256>::rounds<MODE - 20, 64, 1>(h, ring + lane, one);
    if (MODE >= 220 && MODE < 230) Assist<Sha256>::rounds<MODE - 220, 64, 4>(h, ring + lane, one);  // 4 trips of 16 rounds
    if (MODE >= 320 && MODE < 330) Assist<Sha256>::rounds<MODE - 320, 64, 2>(h, ring + lane, one);  // 2 trips of 32 rounds
    if (MODE >= 110 && MODE < 120) Assist<Sha1>::rounds<MODE - 110, 0>(h, ring + lane, one);   // + K in the chain warp
    if (MODE >= 120 && MODE < 130) Assist<Sha256>::rounds<MODE - 120, 0, 1>(h, ring + lane, one);
    // (the expanded window is fed back into w[0]: without a loop-carried value the compiler hoists the whole expansion)
    if (MODE == 30) { uint32_t v[16]; for (int i = 0; i < 16; i++) v[i] = bswap(w[i]); Assist<Sha1>::expand<80>(v, dst, one); w[0] += v[15]; }
    if (MODE == 130) { uint32_t v[16]; for (int i = 0; i < 16; i++) v[i] = bswap(w[i]); Assist<Sha1>::expand<0>(v, dst, one); w[0] += v[15]; }
    if (MODE >= 140 && MODE < 150) { uint32_t v[16]; for (int i = 0; i < 16; i++) v[i] = bswap(w[i]); Assist<Sha256>::expand<MODE - 140, 0>(v, dst, one); w[0] += v[15]; }
    if (MODE >= 40 && MODE < 50) { uint32_t v[16]; for (int i = 0; i < 16; i++) v[i] = bswap(w[i]); Assist<Sha256>::expand<MODE - 40, 64>(v, dst, one); w[0] += v[15]; }
    if (MODE >= 50 && MODE < 60) {  // MD5 block, then the expansion, as two calls (ptxas is left to interleave them)
      uint32_t v[16];
      for (int i = 0; i < 16; i++) v[i] = bswap(w[i]);
      Md5::compress(h, w, one);
      Assist<Sha256>::expand<MODE - 50, 64>(v, dst, one);
      w[0] += v[15];
    }
  }
  const long long t1 = clock64();
  uint32_t x = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) x ^= h[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = x ^ ring[lane].x;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

static uint32_t* g_out;
static long long* g_cyc;

template <class K>
static double run(K kern, int threads) {
  kern<<<148, threads>>>(g_out, 1u, 99u, g_cyc);
  kern<<<148, threads>>>(g_out, 1u, 99u, g_cyc);
  cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, g_cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < 148; i++) avg += (double)h[i];
  return avg / 148;
}

int main() {
  cudaMalloc(&g_out, 148 * 1024 * 4);
  cudaMalloc(&g_cyc, 148 * 8);
  // warps per sub-partition: block of 32 = one warp on one sub-partition; 160 threads = warps 0..4, 0 and 4 share one
  const int shapes[3] = {32, 128, 256};
  const char* shape_name[3] = {"1 warp / SM (alone on its sub-partition)", "4 warps / SM (1 per sub-partition)", "8 warps / SM (2 per sub-partition)"};
  for (int s = 0; s < 3; s++) {
    printf("== %s\n", shape_name[s]);
    const double n = NBLK * 32.0;
    printf("issue: LOP3 only            %.2f cycles per warp instruction\n", run(issue_k<0>, shapes[s]) / n);
    printf("issue: IMAD only            %.2f\n", run(issue_k<1>, shapes[s]) / n);
    printf("issue: LOP3 / IMAD          %.2f\n", run(issue_k<2>, shapes[s]) / n);
    printf("issue: LOP3 / SHF           %.2f\n", run(issue_k<3>, shapes[s]) / n);
    printf("md5 step (LOP3->IADD3->LEA.HI dependent)   %.2f cycles per step\n", run(md5_step_k, shapes[s]) / (NBLK * 64.0));
    printf("Md5::compress               %.1f cycles per block (%.2f per step)\n", run(compress_k<0>, shapes[s]) / NBLK, run(compress_k<0>, shapes[s]) / NBLK / 64);
    printf("Sha1::compress              %.1f cycles per block\n", run(compress_k<1>, shapes[s]) / NBLK);
    printf("Sha256::compress            %.1f cycles per block\n", run(compress_k<2>, shapes[s]) / NBLK);
    printf("Assist<Sha1>::rounds<0>     %.1f cycles per block (chain warp, schedule from shared memory; x = IADD3)\n", run(compress_k<10>, shapes[s]) / NBLK);
    printf("Assist<Sha1>::rounds<1>     %.1f (x = two IMADs)\n", run(compress_k<11>, shapes[s]) / NBLK);
    printf("Assist<Sha256>::rounds<0>   %.1f cycles per block (13 ALU + 2 FMA per round)\n", run(compress_k<20>, shapes[s]) / NBLK);
    printf("Assist<Sha256>::rounds<1>   %.1f (11 ALU + 6 FMA)\n", run(compress_k<21>, shapes[s]) / NBLK);
    printf("Assist<Sha256>::rounds<2>   %.1f (10 ALU + 8 FMA)\n", run(compress_k<22>, shapes[s]) / NBLK);
    printf("Assist<Sha256>::rounds<3>   %.1f (10 ALU + 6 FMA)\n", run(compress_k<23>, shapes[s]) / NBLK);
    printf("Assist<Sha256>::rounds<4>   %.1f (11 ALU + 4 FMA)\n", run(compress_k<24>, shapes[s]) / NBLK);
    printf("Assist<Sha256>::rounds<1> as 4 trips of 16 rounds (4.5 KB of code)  %.1f   as 2 trips of 32 (8.7 KB)  %.1f\n",
           run(compress_k<221>, shapes[s]) / NBLK, run(compress_k<321>, shapes[s]) / NBLK);
    printf("Assist<Sha1>::expand        %.1f cycles per block (helper warp)\n", run(compress_k<30>, shapes[s]) / NBLK);
    printf("Assist<Sha256>::expand<0>   %.1f (adds on the ALU pipe)\n", run(compress_k<40>, shapes[s]) / NBLK);
    printf("Assist<Sha256>::expand<1>   %.1f (adds as IMAD)\n", run(compress_k<41>, shapes[s]) / NBLK);
    printf("expand<0> + Md5::compress   %.1f (one warp doing both)\n", run(compress_k<50>, shapes[s]) / NBLK);
    printf("expand<1> + Md5::compress   %.1f\n", run(compress_k<51>, shapes[s]) / NBLK);
    printf("-- round constants added by the chain warp instead of the helper:\n");
    printf("Assist<Sha1>::rounds<0,K>   %.1f   rounds<1,K> %.1f   expand<K> %.1f\n", run(compress_k<110>, shapes[s]) / NBLK, run(compress_k<111>, shapes[s]) / NBLK, run(compress_k<130>, shapes[s]) / NBLK);
    printf("Assist<Sha256>::rounds<0,K> %.1f   <1,K> %.1f   <4,K> %.1f   expand<0,K> %.1f   expand<1,K> %.1f\n",
           run(compress_k<120>, shapes[s]) / NBLK, run(compress_k<121>, shapes[s]) / NBLK, run(compress_k<124>, shapes[s]) / NBLK,
           run(compress_k<140>, shapes[s]) / NBLK, run(compress_k<141>, shapes[s]) / NBLK);
  }
  return 0;
}
