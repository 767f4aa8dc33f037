// Development probe (not product code): issue rate of tcgen05.mma from one thread, as a function of N, of how
// many CTAs share the SM, of how many accumulators are interleaved, and of how many warps issue.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o build/mma_probe tools/mma_probe.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include "../multimodal_diffusion_b200/csrc/avd_ptx.cuh"
using namespace avd;

struct Args { int N; int n_mma; int n_acc; int issuers; unsigned long long* out; int mode; };

template <int MODE, int KB>
__global__ void __launch_bounds__(128) probe(Args a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 96 * 1024);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 8);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); mbar_init(&bar[2], 1); mbar_init(&bar[3], 1); bar[6] = 0; fence_barrier_init(); mbar_arrive(&bar[2]); }
  if (warp == 0) tmem_alloc<256>(slot);
  // zero the operand tiles so the MMAs are numerically harmless
  for (int i = threadIdx.x; i < 96 * 1024 / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *slot;
  const uint32_t hi = smem_desc_hi_sw128(1024);
  const uint32_t a_lo = smem_desc_lo_sw128(smem_u32(smem), 16);
  const uint32_t b_lo = smem_desc_lo_sw128(smem_u32(smem + 16384), 16);
  const uint32_t idesc = make_idesc_bf16(128, a.N, 0, 0);
  if (warp < a.issuers && lane == 0) {
    unsigned long long t0, t1, t2;
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t0));
    const int acc_base = warp * (a.n_acc * a.N);
    const uint32_t acc1 = (a.n_acc == 2) ? a.N : 0;
    for (int i = 0; i < a.n_mma; i += KB) {
      const uint32_t so = ((i / KB) & 1) * 3072;
      if (MODE & 2) { mbar_wait(&bar[2], 0); }          // an already-completed barrier (phase 0 done)
      if (MODE & 8) {                                    // same with the non-suspending mbarrier.test_wait
        uint32_t ok = 0;
        while (!ok)
          asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}\n"
                       : "=r"(ok) : "r"(smem_u32(&bar[2])), "r"(0) : "memory");
      }
      if (MODE & 16) {                                   // a plain shared-memory load + dependent branch instead
        volatile uint32_t* f = reinterpret_cast<volatile uint32_t*>(bar + 6);
        while (*f != 0) {}
      }
      if (MODE & 4) tc_fence_after();
#pragma unroll
      for (int k = 0; k < KB; ++k)
        umma_bf16_lh(tmem + acc_base + ((k & 1) ? acc1 : 0), a_lo + so + 2 * (k & 3), b_lo + so + 2 * (k & 3), hi, idesc, i > 0 ? 1u : 0u);
      if (MODE & 1) umma_commit(&bar[3]);               // per-k-block commit (nobody waits on it)
    }
    umma_commit(&bar[warp]);
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t1));
    mbar_wait(&bar[warp], 0);
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t2));
    a.out[(blockIdx.x * 2 + warp) * 2] = t1 - t0;
    a.out[(blockIdx.x * 2 + warp) * 2 + 1] = t2 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc<256>(tmem); }
}

template <int MODE, int KB>
void run(int N, int ctas, unsigned long long* d) {
  const int smem = 96 * 1024 + 256;
  cudaFuncSetAttribute(probe<MODE, KB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  Args a{N, 256, 1, 1, d, MODE};
  probe<MODE, KB><<<ctas, 128, smem>>>(a);
  probe<MODE, KB><<<ctas, 128, smem>>>(a);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
  std::vector<unsigned long long> h(ctas * 4);
  cudaMemcpy(h.data(), d, ctas * 4 * 8, cudaMemcpyDeviceToHost);
  std::vector<double> iss, don;
  for (int c = 0; c < ctas; ++c) { iss.push_back(h[(c * 2) * 2]); don.push_back(h[(c * 2) * 2 + 1]); }
  std::sort(iss.begin(), iss.end()); std::sort(don.begin(), don.end());
  printf("%5d %6d %5d %3d | %12.1f %12.1f\n", N, ctas, MODE, KB, iss[iss.size() / 2] / a.n_mma, don[don.size() / 2] / a.n_mma);
}

int main() {
  unsigned long long* d;
  cudaMalloc(&d, 1 << 20);
  printf("%5s %6s %5s %3s | %12s %12s   (mode bits: 1 commit / 2 try_wait / 4 fence per KB MMAs)\n", "N", "ctas", "mode", "KB", "issue cyc/mma", "done cyc/mma");
  for (int N : {64, 128}) {
    run<0, 4>(N, 148, d); run<1, 4>(N, 148, d); run<2, 4>(N, 148, d); run<4, 4>(N, 148, d); run<7, 4>(N, 148, d);
    run<7, 8>(N, 148, d); run<7, 4>(N, 296, d); run<7, 8>(N, 296, d);
    run<8, 4>(N, 148, d); run<13, 4>(N, 148, d); run<16, 4>(N, 148, d); run<3, 4>(N, 148, d);
  }
  return 0;
}
