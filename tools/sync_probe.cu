// Development probe (not product code): cost of a producer -> consumer hand-off between CTAs of one persistent kernel
// through global memory: 136 CTAs in 17 groups of 8.  Per round every CTA stores `bytes` of output (coalesced 16-byte
// stores), fences, bumps its group's counter; then waits until its group's counter shows all 8 arrivals, issues a
// 24 KB TMA-less plain load of the data the group produced (one 16-byte load per thread), and starts the next round.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o build/sync_probe tools/sync_probe.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256, 1) probe(uint4* buf, int* cnt, int rounds, int bytes, unsigned long long* out) {
  const int group = blockIdx.x / 8, tid = threadIdx.x;
  uint4* mine = buf + (size_t)blockIdx.x * (bytes / 16);
  unsigned long long t0 = clock64();
  uint4 acc = make_uint4(0, 0, 0, 0);
  for (int r = 0; r < rounds; ++r) {
    for (int i = tid; i < bytes / 16; i += 256) mine[i] = make_uint4(r, i, tid, acc.x);
    __syncthreads();
    if (tid == 0) {
      __threadfence();
      atomicAdd(&cnt[group * 32], 1);
      int spins = 0;
      while (true) {
        int v;
        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(&cnt[group * 32]));
        if (v >= 8 * (r + 1)) break;
        if (++spins > (1 << 24)) { asm volatile("trap;"); }
      }
    }
    __syncthreads();
    // consume something the group produced (another CTA's data)
    const uint4* other = buf + (size_t)(group * 8 + ((blockIdx.x + 1) & 7)) * (bytes / 16);
    uint4 v = other[tid];
    acc.x += v.x + v.y;
  }
  unsigned long long t1 = clock64();
  if (tid == 0) { out[blockIdx.x * 2] = t1 - t0; out[blockIdx.x * 2 + 1] = acc.x; }
}

int main(int argc, char** argv) {
  const int nctas = argc > 1 ? atoi(argv[1]) : 136;
  uint4* buf; int* cnt; unsigned long long* out;
  cudaMalloc(&buf, 136 * 65536); cudaMalloc(&cnt, 17 * 32 * 4); cudaMalloc(&out, 136 * 16);
  for (int bytes : {256 * 16, 16384, 49152}) {
    const int rounds = 200;
    cudaMemset(cnt, 0, 17 * 32 * 4);
    probe<<<nctas, 256>>>(buf, cnt, rounds, bytes, out);
    cudaMemset(cnt, 0, 17 * 32 * 4);
    probe<<<nctas, 256>>>(buf, cnt, rounds, bytes, out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    std::vector<unsigned long long> h(136 * 2);
    cudaMemcpy(h.data(), out, 136 * 16, cudaMemcpyDeviceToHost);
    std::vector<double> t;
    for (int c = 0; c < nctas; ++c) t.push_back((double)h[c * 2]);
    std::sort(t.begin(), t.end());
    printf("bytes/CTA/round %6d : %8.0f cycles per round (median CTA)\n", bytes, t[nctas / 2] / rounds);
  }
  return 0;
}
