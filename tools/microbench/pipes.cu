// pipes.cu — measures the integer issue rates the hashing roofline rests on (SURVEY.md §8d asks for
// the 64 lane-ops/clk/SM alu-pipe figure to be validated once).  Each kernel runs a register-only
// loop of independent chains; rate = warp-instructions / SM / clk from clock64().
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu && ./pipes
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 4096

template <int MODE>
__global__ void k(uint32_t* out, uint32_t one, uint32_t seed, long long* cyc) {
  uint32_t a0 = seed + threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, a4 = a0 * 11, a5 = a0 * 13, a6 = a0 * 17, a7 = a0 * 19;
  uint32_t b0 = a0 ^ 1, b1 = a1 ^ 2, b2 = a2 ^ 3, b3 = a3 ^ 4, b4 = a4 ^ 5, b5 = a5 ^ 6, b6 = a6 ^ 7, b7 = a7 ^ 8;
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < ITERS; i++) {
#define LOP(x, y) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x) : "r"(y), "r"(seed))
#define SHFO(x) asm volatile("shf.l.wrap.b32 %0, %0, %0, 5;" : "+r"(x))
#define MAD(x, y) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x) : "r"(one), "r"(y))
#define WIDE(x, y) asm volatile("{ .reg .u64 t; mul.wide.u32 t, %0, %1; mov.b64 {%0, %1}, t; }" : "+r"(x), "+r"(y))
#define ADD3(x, y) asm volatile("{ .reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2; }" : "+r"(x) : "r"(y), "r"(seed))
    if (MODE == 0) { LOP(a0, b0); LOP(a1, b1); LOP(a2, b2); LOP(a3, b3); LOP(a4, b4); LOP(a5, b5); LOP(a6, b6); LOP(a7, b7); }
    if (MODE == 1) { SHFO(a0); SHFO(a1); SHFO(a2); SHFO(a3); SHFO(a4); SHFO(a5); SHFO(a6); SHFO(a7); }
    if (MODE == 2) { MAD(a0, b0); MAD(a1, b1); MAD(a2, b2); MAD(a3, b3); MAD(a4, b4); MAD(a5, b5); MAD(a6, b6); MAD(a7, b7); }
    if (MODE == 3) { LOP(a0, b0); MAD(a1, b1); LOP(a2, b2); MAD(a3, b3); LOP(a4, b4); MAD(a5, b5); LOP(a6, b6); MAD(a7, b7); }
    if (MODE == 4) { LOP(a0, b0); LOP(a1, b1); MAD(a2, b2); LOP(a3, b3); LOP(a4, b4); MAD(a5, b5); LOP(a6, b6); LOP(a7, b7); }  // 6 alu : 2 fma
    if (MODE == 5) { LOP(a0, b0); SHFO(a1); MAD(a2, b2); LOP(a3, b3); SHFO(a4); MAD(a5, b5); LOP(a6, b6); MAD(a7, b7); }      // 5 alu : 3 fma
    if (MODE == 7) { WIDE(a0, b0); WIDE(a1, b1); WIDE(a2, b2); WIDE(a3, b3); WIDE(a4, b4); WIDE(a5, b5); WIDE(a6, b6); WIDE(a7, b7); }
    if (MODE == 8) { LOP(a0, b0); WIDE(a1, b1); LOP(a2, b2); WIDE(a3, b3); LOP(a4, b4); WIDE(a5, b5); LOP(a6, b6); WIDE(a7, b7); }
    if (MODE == 9) { LOP(a0, b0); LOP(a1, b1); WIDE(a2, b2); LOP(a3, b3); LOP(a4, b4); WIDE(a5, b5); LOP(a6, b6); MAD(a7, b7); }  // 5 alu : 2 wide : 1 mad
    if (MODE == 6) { ADD3(a0, b0); ADD3(a1, b1); ADD3(a2, b2); ADD3(a3, b3); ADD3(a4, b4); ADD3(a5, b5); ADD3(a6, b6); ADD3(a7, b7); }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 ^ a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7 ^ b0 ^ b1 ^ b2 ^ b3 ^ b4 ^ b5 ^ b6 ^ b7;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, int warps_per_sm, int active_lanes) {
  int threads = warps_per_sm * 32, blocks = 148;
  uint32_t* out; long long* cyc;
  cudaMalloc(&out, blocks * threads * 4); cudaMalloc(&cyc, blocks * 8);
  (void)active_lanes;
  k<MODE><<<blocks, threads>>>(out, 1, 12345, cyc);
  k<MODE><<<blocks, threads>>>(out, 1, 12345, cyc);
  cudaDeviceSynchronize();
  long long h[148]; cudaMemcpy(h, cyc, blocks * 8, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < blocks; i++) avg += h[i]; avg /= blocks;
  double winst = (double)ITERS * 8 * warps_per_sm;
  printf("%-28s warps/SM %2d  cycles %9.0f  warp-inst/clk/SM %.3f  lane-ops/clk/SM %.1f\n", name, warps_per_sm, avg, winst / avg, winst * 32 / avg);
  cudaFree(out); cudaFree(cyc);
}

int main() {
  for (int w : {12, 16}) {
    run<0>("LOP3 only (alu)", w, 32);
    run<1>("SHF only (alu)", w, 32);
    run<6>("ADD x2 -> IADD3? (alu)", w, 32);
    run<2>("IMAD only (fma)", w, 32);
    run<3>("LOP3:IMAD 4:4", w, 32);
    run<4>("LOP3:IMAD 6:2", w, 32);
    run<5>("LOP3+SHF:IMAD 5:3", w, 32);
    run<7>("IMAD.WIDE only", w, 32);
    run<8>("LOP3:IMAD.WIDE 4:4", w, 32);
    run<9>("LOP3:WIDE:IMAD 5:2:1", w, 32);
  }
  return 0;
}
