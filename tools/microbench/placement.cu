// placement.cu — where do the warps of small CTAs land?  Prints, for a few CTAs, %smid and %warpid of each warp
// (hardware warp slot; slot % 4 = SM sub-partition) for 32-, 64- and 128-thread CTAs.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(int* out) {
  if ((threadIdx.x & 31) == 0) {
    unsigned smid, wid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    asm volatile("mov.u32 %0, %%warpid;" : "=r"(wid));
    int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    out[2 * w] = smid; out[2 * w + 1] = wid;
  }
  // keep the CTA resident for a while so that all CTAs coexist
  long long t = clock64(); while (clock64() - t < 2000000) {}
}
int main() {
  int* d; cudaMalloc(&d, 1 << 20);
  for (int threads : {32, 64, 128}) {
    for (int blocks : {128, 296, 592}) {
      int warps = blocks * threads / 32;
      k<<<blocks, threads, 40 * 1024>>>(d);
      cudaDeviceSynchronize();
      int* h = new int[2 * warps]; cudaMemcpy(h, d, 8 * warps, cudaMemcpyDeviceToHost);
      int cnt[160][4] = {};
      for (int w = 0; w < warps; w++) cnt[h[2 * w]][h[2 * w + 1] & 3]++;
      int hist[16] = {}; int maxper = 0;
      for (int s = 0; s < 148; s++) for (int q = 0; q < 4; q++) { hist[cnt[s][q] < 15 ? cnt[s][q] : 15]++; if (cnt[s][q] > maxper) maxper = cnt[s][q]; }
      printf("threads %3d blocks %3d: warps per sub-partition histogram:", threads, blocks);
      for (int i = 0; i <= maxper; i++) printf(" [%d]=%d", i, hist[i]);
      printf("   first CTAs (smid,warpid):");
      for (int w = 0; w < 6; w++) printf(" (%d,%d)", h[2 * w], h[2 * w + 1]);
      printf("\n");
      delete[] h;
    }
  }
  return 0;
}
