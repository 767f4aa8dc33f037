// Launch-configuration probe: which combinations of cluster shape, dynamic smem and PDL does cudaLaunchKernelEx accept?
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(256, 1) k(int* out) {
  extern __shared__ unsigned char sm[];
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  if (threadIdx.x == 0) { sm[0] = (unsigned char)r; out[blockIdx.y * gridDim.x + blockIdx.x] = (int)r; }
}
static void tryit(int cx, int cy, int smem, bool pdl, dim3 grid) {
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  int* out; cudaMalloc(&out, 4096 * 4);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[2]; int na = 0;
  if (pdl) { at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[na].val.programmaticStreamSerializationAllowed = 1; ++na; }
  at[na].id = cudaLaunchAttributeClusterDimension; at[na].val.clusterDim.x = cx; at[na].val.clusterDim.y = cy; at[na].val.clusterDim.z = 1; ++na;
  cfg.attrs = at; cfg.numAttrs = na;
  int ncl = -1;
  cudaError_t eo = cudaOccupancyMaxActiveClusters(&ncl, k, &cfg);
  cudaError_t e = cudaLaunchKernelEx(&cfg, k, out);
  cudaError_t e2 = cudaDeviceSynchronize();
  printf("cluster %dx%d smem %d pdl %d grid %dx%d: occ=%d (%s) launch=%s sync=%s\n", cx, cy, smem, (int)pdl, grid.x, grid.y, ncl,
         cudaGetErrorString(eo), cudaGetErrorString(e), cudaGetErrorString(e2));
  cudaGetLastError();
  cudaFree(out);
}
int main() {
  tryit(1, 2, 1024, false, dim3(8, 18));
  tryit(1, 2, 199424, false, dim3(8, 18));
  tryit(1, 2, 199424, true, dim3(8, 18));
  tryit(2, 1, 199424, true, dim3(8, 18));
  tryit(2, 1, 199424, true, dim3(18, 8));
  tryit(1, 2, 230000, true, dim3(8, 18));
  return 0;
}
