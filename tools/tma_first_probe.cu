// Development probe (not product code): latency of the FIRST TMA tile load of a kernel, by where the tensor map lives.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o build/tma_first_probe tools/tma_first_probe.cu -lcuda
// mode 0: __grid_constant__ kernel parameter, prefetch.tensormap at kernel start (what the GEMM kernels do)
// mode 1: same without the prefetch
// mode 2: tensor map in global memory (uploaded once, stays in L2 across launches), prefetch at kernel start
// Reports cycles from kernel start to the first box landed, and the latency of a second, later load.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <cuda.h>
#include "../multimodal_diffusion_b200/csrc/avd_ptx.cuh"
using namespace avd;

template <int MODE>
__global__ void __launch_bounds__(128) probe(const __grid_constant__ CUtensorMap tm_param, const CUtensorMap* tm_glob,
                                             unsigned long long* out, int delay) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 32768);
  const CUtensorMap* tm = MODE == 2 ? tm_glob : &tm_param;
  unsigned long long t0, t1, t2, t3;
  asm volatile("mov.u64 %0, %%clock64;" : "=l"(t0));
  if (threadIdx.x == 0) {
    if (MODE != 1) tma_prefetch_desc(tm);
    mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); fence_barrier_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 0; i < delay; ++i) asm volatile("nanosleep.u32 100;");
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t1));
    mbar_expect_tx(&bar[0], 16384);
    tma_load_3d(smem, tm, &bar[0], 0, blockIdx.x * 128, 0);
    mbar_wait(&bar[0], 0);
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t2));
    mbar_expect_tx(&bar[1], 16384);
    tma_load_3d(smem + 16384, tm, &bar[1], 64, blockIdx.x * 128, 0);
    mbar_wait(&bar[1], 0);
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t3));
    out[blockIdx.x * 4 + 0] = t1 - t0;
    out[blockIdx.x * 4 + 1] = t2 - t1;
    out[blockIdx.x * 4 + 2] = t3 - t2;
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int MODE>
void run(const CUtensorMap& tm, const CUtensorMap* tmg, unsigned long long* d, int delay, const char* name) {
  const int smem = 32768 + 256, ctas = 136;
  cudaFuncSetAttribute(probe<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  std::vector<double> first, second, pre;
  for (int rep = 0; rep < 5; ++rep) {
    probe<MODE><<<ctas, 128, smem>>>(tm, tmg, d, delay);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
    if (rep == 0) continue;
    std::vector<unsigned long long> h(ctas * 4);
    cudaMemcpy(h.data(), d, ctas * 4 * 8, cudaMemcpyDeviceToHost);
    for (int c = 0; c < ctas; ++c) { pre.push_back(h[c * 4]); first.push_back(h[c * 4 + 1]); second.push_back(h[c * 4 + 2]); }
  }
  auto med = [](std::vector<double>& v) { std::sort(v.begin(), v.end()); return v[v.size() / 2]; };
  printf("%-44s delay %3d | setup %6.0f  first load %6.0f  second load %6.0f cycles\n", name, delay, med(pre), med(first), med(second));
}

int main() {
  const int K = 512, R = 136 * 128;
  void* buf; cudaMalloc(&buf, (size_t)R * K * 2); cudaMemset(buf, 0, (size_t)R * K * 2);
  unsigned long long* d; cudaMalloc(&d, 1 << 16);
  void* fn = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  CUtensorMap tm;
  cuuint64_t dims[3] = {(cuuint64_t)K, (cuuint64_t)R, 1}; cuuint64_t strides[2] = {(cuuint64_t)K * 2, (cuuint64_t)R * K * 2};
  cuuint32_t box[3] = {64, 128, 1}, es[3] = {1, 1, 1};
  CUresult r = ((EncodeTiledFn)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, buf, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
  CUtensorMap* tmg; cudaMalloc(&tmg, 1024); cudaMemcpy(tmg, &tm, sizeof(tm), cudaMemcpyHostToDevice);
  for (int delay : {0, 20}) {
    run<0>(tm, tmg, d, delay, "param tensor map + prefetch.tensormap");
    run<1>(tm, tmg, d, delay, "param tensor map, no prefetch");
    run<2>(tm, tmg, d, delay, "global-memory tensor map + prefetch");
  }
  return 0;
}
