// probe: which launch configurations accept a 2-CTA cluster on this box?
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(320, 1) k(int* out) {
  extern __shared__ unsigned char sm[];
  unsigned r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  if (threadIdx.x == 0) { out[blockIdx.x] = (int)r; }
}
int main() {
  int* d; cudaMalloc(&d, 4096);
  int sizes[] = {0, 100 * 1024, 200 * 1024, 226 * 1024, 232448 - 2048, 232448 - 1024, 232448};
  for (int s : sizes) {
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, s);
    cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3(148); cfg.blockDim = dim3(320); cfg.dynamicSmemBytes = s;
    cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int nclusters = -1; cudaError_t eo = cudaOccupancyMaxActiveClusters(&nclusters, k, &cfg);
    cudaError_t el = cudaLaunchKernelEx(&cfg, k, d);
    cudaError_t es = cudaDeviceSynchronize();
    printf("smem %6d: setattr %s | maxActiveClusters %d (%s) | launch %s | sync %s\n", s, cudaGetErrorName(e), nclusters, cudaGetErrorName(eo), cudaGetErrorName(el), cudaGetErrorName(es));
    cudaGetLastError();
  }
  return 0;
}
