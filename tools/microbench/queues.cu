// queues.cu — how many long kernels of different streams run at once, and what queued behind them costs.
// N streams, one ~50 ms single-warp kernel each; variants: nothing after the kernel / an event record after it /
// an event record and a small D2H copy after it / the same behind a wait on a copy stream's event / the same with
// a padded grid of 148 CTAs asking for 112 KiB of shared memory each / variant 3 with a timing-enabled event.  Wall time ~50 ms means all N overlapped.
// Build: nvcc -arch=sm_100a -O3 -o queues queues.cu ; run with CUDA_DEVICE_MAX_CONNECTIONS=8|32.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

__global__ void spin(long long cycles, int* out) {
  extern __shared__ unsigned char sm[];
  if (blockIdx.x != 0) return;  // padded grids: only CTA 0 works, like the library's small launches
  const long long t0 = clock64();
  while (clock64() - t0 < cycles) {}
  if (threadIdx.x == 0) out[0] = 1 + (sm != nullptr);
}

int main() {
  setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const long long cycles = (long long)khz * 50;  // 50 ms
  int* d;
  cudaMalloc(&d, 4096 * sizeof(int));
  int* h;
  cudaHostAlloc(&h, 4096 * sizeof(int), cudaHostAllocDefault);
  cudaFuncSetAttribute(spin, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
  cudaStream_t s_copy;
  cudaStreamCreateWithFlags(&s_copy, cudaStreamNonBlocking);
  char* big_h;
  char* big_d;
  cudaHostAlloc(&big_h, 64 << 20, cudaHostAllocDefault);
  cudaMalloc(&big_d, 64 << 20);
  for (int variant = 0; variant < 6; variant++) {
    for (int n : {16, 32, 64, 128, 256}) {
      std::vector<cudaStream_t> st(n);
      std::vector<cudaEvent_t> ev(n);
      for (int i = 0; i < n; i++) {
        cudaStreamCreateWithFlags(&st[i], cudaStreamNonBlocking);
        if (variant == 5) cudaEventCreate(&ev[i]);  // timing-enabled
        else cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming);
      }
      spin<<<1, 32, 0, st[0]>>>(1000, d);
      cudaDeviceSynchronize();
      const auto t0 = std::chrono::steady_clock::now();
      for (int i = 0; i < n; i++) {
        if (variant >= 3) {  // inputs arrive on a copy stream; the kernel's stream waits for them
          cudaMemcpyAsync(big_d, big_h, 4 << 20, cudaMemcpyHostToDevice, s_copy);
          cudaEvent_t in;
          cudaEventCreateWithFlags(&in, cudaEventDisableTiming);
          cudaEventRecord(in, s_copy);
          cudaStreamWaitEvent(st[i], in, 0);
          cudaEventDestroy(in);
        }
        if (variant == 4) spin<<<148, 64, 112 * 1024, st[i]>>>(cycles, d + i);
        else spin<<<1, 32, 0, st[i]>>>(cycles, d + i);
        if (variant >= 1) cudaEventRecord(ev[i], st[i]);
        if (variant >= 2) cudaMemcpyAsync(h + i, d + i, 4, cudaMemcpyDeviceToHost, st[i]);
      }
      cudaDeviceSynchronize();
      const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
      printf("variant %d (%s) streams %3d: %.1f ms (%.1f kernels deep)\n", variant,
             variant == 0 ? "kernel only" : variant == 1 ? "kernel + event record" : variant == 2 ? "kernel + event + D2H" : variant == 3 ? "wait(copy) + kernel + event + D2H" : variant == 4 ? "same, grid 148 x 112 KiB smem" : "variant 3 with a TIMING-enabled event after the kernel", n, ms, ms / 50.0);
      for (int i = 0; i < n; i++) {
        cudaStreamDestroy(st[i]);
        cudaEventDestroy(ev[i]);
      }
    }
  }
  return 0;
}
