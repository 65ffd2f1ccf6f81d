// latency.cu — dependent-chain latency (cycles per op, one warp per SM) of the integer ops the digests use.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define N 4096
template <int MODE>
__global__ void k(uint32_t* out, uint32_t one, uint32_t seed, long long* cyc) {
  uint32_t a = seed + threadIdx.x, b = seed * 3 + 1, c = seed ^ 0x55;
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < N; i++) {
#pragma unroll
    for (int j = 0; j < 8; j++) {
      if (MODE == 0) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a) : "r"(b), "r"(c));
      if (MODE == 1) asm volatile("shf.l.wrap.b32 %0, %0, %0, 5;" : "+r"(a));
      if (MODE == 2) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a) : "r"(one), "r"(b));
      if (MODE == 3) asm volatile("add.u32 %0, %0, %1;" : "+r"(a) : "r"(b));
      if (MODE == 4) { asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a) : "r"(b), "r"(c)); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a) : "r"(one), "r"(b)); }
      if (MODE == 5) { asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a) : "r"(b), "r"(c)); asm volatile("add.u32 %0, %0, %1;" : "+r"(a) : "r"(b)); }
      if (MODE == 6) { uint32_t t; asm volatile("shf.l.wrap.b32 %0, %1, %1, 5;" : "=r"(t) : "r"(a)); asm volatile("add.u32 %0, %1, %2;" : "=r"(a) : "r"(t), "r"(b)); }
    }
  }
  long long t1 = clock64();
  out[threadIdx.x] = a;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
template <int MODE> void run(const char* name, int ops) {
  uint32_t* out; long long* cyc; cudaMalloc(&out, 4096); cudaMalloc(&cyc, 8);
  k<MODE><<<1, 32>>>(out, 1, 99, cyc); k<MODE><<<1, 32>>>(out, 1, 99, cyc); cudaDeviceSynchronize();
  long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-34s %.2f cycles per op\n", name, (double)h / (N * 8.0 * ops));
}
int main() {
  run<0>("LOP3 -> LOP3", 1); run<1>("SHF -> SHF", 1); run<2>("IMAD -> IMAD (3 reg)", 1); run<3>("ADD -> ADD", 1);
  run<4>("LOP3 -> IMAD -> LOP3 (avg)", 2); run<5>("LOP3 -> ADD -> LOP3 (avg)", 2); run<6>("SHF+ADD (LEA.HI?) -> ...", 1);
  return 0;
}
