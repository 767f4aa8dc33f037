// Development probe (not product code): L2 -> SM throughput of TMA tile loads.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o build/tma_probe tools/tma_probe.cu -lcuda
// Each CTA runs `nprod` producer threads; each keeps `depth` 16 KB boxes (128 rows x 64 bf16, 128B swizzle) in flight
// and loads `iters` boxes.  `share` CTAs read the same rows (share = 1: all distinct).
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <cuda.h>
#include "../multimodal_diffusion_b200/csrc/avd_ptx.cuh"
using namespace avd;

struct Args { int iters, depth, nprod, share, K; unsigned long long* out; };

__global__ void __launch_bounds__(128) probe(const __grid_constant__ CUtensorMap tm, Args a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 192 * 1024);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { for (int i = 0; i < 16; ++i) mbar_init(&bar[i], 1); fence_barrier_init(); tma_prefetch_desc(&tm); }
  __syncthreads();
  if (warp < a.nprod && lane == 0) {
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t0));
    const int kblocks = a.K >> 6;
    const int group = blockIdx.x >> a.share;
    uint64_t* mybar = bar + warp * 4;
    uint8_t* mybuf = smem + warp * (a.depth * 16384);
    int st = 0, kb = warp % kblocks, row = group * 8 * 128;
    uint32_t ph = 0;
    const int depth = a.depth, nprod = a.nprod, iters = a.iters;
    for (int it = 0; it < iters; ++it) {
      if (it >= depth) mbar_wait(&mybar[st], ph ^ 1);
      mbar_expect_tx(&mybar[st], 16384);
      tma_load_3d(mybuf + st * 16384, &tm, &mybar[st], kb * 64, row, 0);
      kb += nprod;
      if (kb >= kblocks) { kb -= kblocks; row += 128; if (row >= (group + 1) * 8 * 128) row = group * 8 * 128; }
      if (++st == depth) { st = 0; ph ^= 1; }
    }
    for (int i = 0; i < depth; ++i) { mbar_wait(&mybar[st], ph ^ 1); if (++st == depth) { st = 0; ph ^= 1; } }
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t1));
    a.out[blockIdx.x * 4 + warp] = t1 - t0;
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  const int K = 2048, rows = 148 * 8 * 128;     // 151552 rows x 2048 bf16 = 620 MB?  too big: use fewer rows below
  const int R = 148 * 8 * 128 / 8;              // 18944 rows x 4 KB = 77 MB (L2-resident after the first pass)
  (void)rows;
  void* buf; cudaMalloc(&buf, (size_t)R * K * 2); cudaMemset(buf, 0, (size_t)R * K * 2);
  unsigned long long* d; cudaMalloc(&d, 1 << 20);
  void* fn = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  CUtensorMap tm;
  cuuint64_t dims[3] = {(cuuint64_t)K, (cuuint64_t)R, 1}; cuuint64_t strides[2] = {(cuuint64_t)K * 2, (cuuint64_t)R * K * 2};
  cuuint32_t box[3] = {64, 128, 1}, es[3] = {1, 1, 1};
  CUresult r = ((EncodeTiledFn)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, buf, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
  const int smem = 192 * 1024 + 256;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  printf("%6s %6s %6s %6s | %10s %12s %10s\n", "ctas", "nprod", "depth", "share", "cyc/box", "B/cyc/SM", "KB/cyc chip");
  for (int ctas : {40, 148})
    for (int share : {0, 3, 4})
      for (int nprod : {1, 2, 4})
        for (int depth : {1, 2, 3}) {
          Args a{256, depth, nprod, share, K, d};     // share = log2 of the CTAs reading the same rows
          probe<<<ctas, 128, smem>>>(tm, a);
          probe<<<ctas, 128, smem>>>(tm, a);
          cudaError_t e = cudaDeviceSynchronize();
          if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
          std::vector<unsigned long long> h(ctas * 4);
          cudaMemcpy(h.data(), d, ctas * 4 * 8, cudaMemcpyDeviceToHost);
          std::vector<double> t;
          for (int c = 0; c < ctas; ++c) for (int w = 0; w < nprod; ++w) t.push_back((double)h[c * 4 + w]);
          std::sort(t.begin(), t.end());
          const double cyc = t[t.size() / 2];
          printf("%6d %6d %6d %6d | %10.1f %12.1f %10.2f\n", ctas, nprod, depth, a.share, cyc / a.iters,
                 16384.0 * a.iters * nprod / cyc, 16384.0 * a.iters * nprod * ctas / cyc / 1024.0);
        }
  return 0;
}
