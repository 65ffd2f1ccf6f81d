// lanes.cu — does a warp instruction with fewer active lanes leave the ALU pipe sooner?
// Same dependent-free LOP3/SHF/IMAD loop, 1 or 2 warps per SM sub-partition, with 32 / 16 / 8 / 1
// active lanes (the others exit at once).  If 16 active lanes halved the pipe time, few-message
// batches could be spread over more, narrower warps.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define ITERS 8192
__global__ void k(uint32_t* out, uint32_t one, uint32_t seed, long long* cyc, int active, int mode) {
  const int lane = threadIdx.x & 31;
  if (mode == 0 && lane >= active) return;                 // contiguous low lanes
  if (mode == 1 && (lane % (32 / active)) != 0) return;    // spread lanes
  uint32_t a0 = seed + threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, a4 = a0 * 11, a5 = a0 * 13, a6 = a0 * 17, a7 = a0 * 19;
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < ITERS; i++) {
    a0 = __funnelshift_l(a0, a0, 5) ^ a1; a1 = (a1 ^ a2 ^ a3); a2 = __funnelshift_l(a2, a2, 7) + a3; a3 = a3 ^ a4 ^ a5;
    a4 = __funnelshift_l(a4, a4, 9) ^ a5; a5 = (a5 ^ a6 ^ a7); a6 = __funnelshift_l(a6, a6, 11) + a7; a7 = a7 ^ a0 ^ seed;
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 ^ a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main() {
  uint32_t* out; long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
  for (int warps : {4, 8, 16})
    for (int mode : {0, 1})
      for (int active : {32, 16, 8, 1}) {
        k<<<148, warps * 32>>>(out, 1, 7, cyc, active, mode);
        k<<<148, warps * 32>>>(out, 1, 7, cyc, active, mode);
        cudaDeviceSynchronize();
        long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        double avg = 0; for (int i = 0; i < 148; i++) avg += h[i]; avg /= 148;
        printf("warps/SM %2d  %s lanes %2d  cycles/iter %.2f\n", warps, mode ? "spread    " : "contiguous", active, avg / ITERS);
      }
  return 0;
}
