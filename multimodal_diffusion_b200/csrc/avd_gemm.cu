// avd_gemm.cu -- K4: the tcgen05 GEMM of the MMDT blocks, the noise head and (via operand splitting) the
// fp32-parity mode.
//
//   out = epilogue( A[M,K] x W[N,K]^T ),   A, W bf16 K-major, fp32 accumulation in TMEM.
//
// Warp roles (192 threads, one 128 x BN output tile per CTA):
//   warp 0      TMA producer: cp.async.bulk.tensor (3-D maps {K, rows, segment}) into a STAGES-deep ring of
//               128B-swizzled smem tiles, mbarrier complete_tx
//   warp 1      TMEM allocator + MMA issuer: one elected lane issues tcgen05.mma.cta_group::1.kind::f16
//               (M=128, N=BN, K=16) x 4 per 64-wide k-block; tcgen05.commit releases the smem slot / signals
//               the epilogue
//   warps 2..5  epilogue: tcgen05.ld 32 lanes x 32 columns -> registers; thread == output row
//               v = acc * rinv(row) + bias (+ resid) [gelu]; stores fp32 and/or bf16 segments, and the
//               per-row partial sums of squares that the NEXT GEMM turns into its RMSNorm row scale.
//
// Precision modes are realised by the k-loop: for mode p the operands are stored as hi/(mid/)lo bf16
// segments and the loop runs over a list of (A segment, W segment) pairs, smallest products first, all
// accumulating into the same TMEM tile (1, 3 or 6 passes over K).
#include "avd_internal.h"
#include "avd_ptx.cuh"

namespace avd {

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;            // 64 bf16 = 128 B = one swizzle atom row
constexpr int GEMM_THREADS = 192;

struct GemmParams {
  int M, N, K;
  int n_pairs;
  uint32_t pairs;            // 4 bits per pair: (a_seg << 2) | w_seg, pair i in bits [4i, 4i+4)
  const float* bias;
  const float* ssq_in;
  int ssq_parts;
  float inv_sqrt_rms_dim;    // 1/sqrt(rms_dim)
  float rms_eps;
  const float* resid;
  int act;
  float* out_f32;
  __nv_bfloat16* out_bf16;
  int out_nseg;
  float* ssq_out;
};

template <int BN, int STAGES>
struct GemmSmem {
  static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
  static constexpr int B_BYTES = BN * GEMM_BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES;
  static constexpr int TOTAL = BAR_OFFSET + 256;  // tiles + barriers
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(GEMM_THREADS) gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA,
                                                               const __grid_constant__ CUtensorMap tmB,
                                                               const GemmParams p) {
  using S = GemmSmem<BN, STAGES>;
  constexpr uint32_t TMEM_COLS = BN < 32 ? 32 : BN;
  extern __shared__ __align__(1024) uint8_t gemm_smem[];
  uint8_t* smem = gemm_smem;
  if ((smem_u32(smem) & 1023u) != 0) __trap();   // 128B-swizzle tiles need 1024-B aligned bases
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::BAR_OFFSET);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* acc_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * GEMM_BM;
  const int n0 = blockIdx.x * BN;
  const int kblocks = p.K / GEMM_BK;
  const int iters = kblocks * p.n_pairs;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(acc_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      for (int it = 0; it < iters; ++it) {
        const int st = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        mbar_wait(&empty_bar[st], ph ^ 1);
        const int pi = it / kblocks, kb = it - pi * kblocks;
        const uint32_t pr = (p.pairs >> (4 * pi)) & 0xF;
        const int sa = pr >> 2, sw = pr & 3;
        uint8_t* a_dst = smem + st * S::STAGE_BYTES;
        uint8_t* b_dst = a_dst + S::A_BYTES;
        mbar_expect_tx(&full_bar[st], S::STAGE_BYTES);
        tma_load_3d(a_dst, &tmA, &full_bar[st], kb * GEMM_BK, m0, sa);
        tma_load_3d(b_dst, &tmB, &full_bar[st], kb * GEMM_BK, n0, sw);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(GEMM_BM, BN, 0, 0);
      for (int it = 0; it < iters; ++it) {
        const int st = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        mbar_wait(&full_bar[st], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + st * S::STAGE_BYTES);
        const uint32_t b_addr = a_addr + S::A_BYTES;
#pragma unroll
        for (int k = 0; k < GEMM_BK / 16; ++k) {
          const uint64_t da = make_smem_desc_sw128(a_addr + k * 32, 16, 1024);
          const uint64_t db = make_smem_desc_sw128(b_addr + k * 32, 16, 1024);
          umma_bf16(tmem_acc, da, db, idesc, (it > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(&empty_bar[st]);   // smem slot reusable once these MMAs have read it
      }
      umma_commit(acc_bar);            // accumulator complete
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..5)
    const int quad = warp & 3;                 // TMEM lane quadrant this warp may access
    const int row = m0 + quad * 32 + lane;
    const bool live = row < p.M;
    float rinv = 1.0f;
    if (p.ssq_in != nullptr && live) {
      float ss = 0.0f;
      for (int q = 0; q < p.ssq_parts; ++q) ss += p.ssq_in[(int64_t)row * p.ssq_parts + q];
      rinv = 1.0f / (sqrtf(ss) * p.inv_sqrt_rms_dim + p.rms_eps);
    }
    mbar_wait(acc_bar, 0);
    tc_fence_after();
    const int64_t seg_stride = (int64_t)p.M * p.N;
    float ssq_acc = 0.0f;
#pragma unroll 1
    for (int ch = 0; ch < BN / 32; ++ch) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_acc + ((uint32_t)(quad * 32) << 16) + ch * 32, r);
      tmem_ld_wait();
      const int col0 = n0 + ch * 32;
      if (live) {
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * rinv;
        if (p.bias != nullptr) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + j));
            v[j] += b4.x; v[j + 1] += b4.y; v[j + 2] += b4.z; v[j + 3] += b4.w;
          }
        }
        if (p.resid != nullptr) {
          const float* rp = p.resid + (int64_t)row * p.N + col0;
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            float4 x4 = *reinterpret_cast<const float4*>(rp + j);
            v[j] += x4.x; v[j + 1] += x4.y; v[j + 2] += x4.z; v[j + 3] += x4.w;
          }
        }
        if (p.act == 1) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
        }
        if (p.out_f32 != nullptr) {
          float* op = p.out_f32 + (int64_t)row * p.N + col0;
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            *reinterpret_cast<float4*>(op + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        }
        if (p.out_bf16 != nullptr) {
          __nv_bfloat16* op = p.out_bf16 + (int64_t)row * p.N + col0;
          float rem[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) rem[j] = v[j];
          for (int sg = 0; sg < p.out_nseg; ++sg) {
            uint32_t w[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              __nv_bfloat16 h0 = __float2bfloat16_rn(rem[2 * j]);
              __nv_bfloat16 h1 = __float2bfloat16_rn(rem[2 * j + 1]);
              rem[2 * j] -= __bfloat162float(h0);
              rem[2 * j + 1] -= __bfloat162float(h1);
              w[j] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
            }
            uint4* o4 = reinterpret_cast<uint4*>(op + sg * seg_stride);
#pragma unroll
            for (int j = 0; j < 4; ++j) o4[j] = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
          }
        }
        if (p.ssq_out != nullptr) {
#pragma unroll
          for (int j = 0; j < 32; ++j) ssq_acc = fmaf(v[j], v[j], ssq_acc);
          if (ch & 1) {  // a 64-column part is complete
            p.ssq_out[(int64_t)row * (p.N / 64) + (col0 >> 6)] = ssq_acc;
            ssq_acc = 0.0f;
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<TMEM_COLS>(tmem_acc);
  }
}

template <int BN, int STAGES>
static int launch_gemm_bn(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, cudaStream_t s) {
  using S = GemmSmem<BN, STAGES>;
  static bool attr_set = false;
  if (!attr_set) {
    AVD_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
    attr_set = true;
  }
  dim3 grid(p.N / BN, (p.M + GEMM_BM - 1) / GEMM_BM);
  LaunchScope ls(KIND_GEMM, s);
  gemm_tc_kernel<BN, STAGES><<<grid, GEMM_THREADS, S::TOTAL, s>>>(tmA, tmB, p);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

int launch_gemm(const avd_gemm_args& g, cudaStream_t s) {
  const int nseg = nseg_of(g.precision);
  AVD_REQUIRE(nseg > 0, "precision must be 1, 3 or 6");
  AVD_REQUIRE(g.M > 0 && g.N > 0 && g.K > 0, "empty GEMM");
  AVD_REQUIRE(g.K % GEMM_BK == 0, "K must be a multiple of 64");
  AVD_REQUIRE(g.N % 32 == 0, "N must be a multiple of 32");
  AVD_REQUIRE(g.A != nullptr && g.W != nullptr, "null operand");
  AVD_REQUIRE(g.ssq_out == nullptr || g.N % 64 == 0, "ssq_out needs N % 64 == 0");
  AVD_REQUIRE(g.out_f32 != nullptr || g.out_bf16 != nullptr, "no output");

  // tile width: widest BN that divides N and still gives >= ~1 wave of CTAs on 148 SMs
  const int mt = (g.M + GEMM_BM - 1) / GEMM_BM;
  int bn = 32;
  if (g.N % 128 == 0 && (int64_t)mt * (g.N / 128) >= 120) bn = 128;
  else if (g.N % 64 == 0) bn = 64;
  if (g.ssq_out != nullptr && bn < 64) bn = 64;

  GemmParams p{};
  p.M = g.M; p.N = g.N; p.K = g.K;
  // smallest partial products first so they are not absorbed by the large ones
  if (g.precision == 1) { p.n_pairs = 1; p.pairs = 0x0; }
  else if (g.precision == 3) { p.n_pairs = 3; p.pairs = (0x4u) | (0x1u << 4) | (0x0u << 8); }  // (lo,hi) (hi,lo) (hi,hi)
  else {
    // segments 0=hi 1=mid 2=lo: (1,1) (2,0) (0,2) (1,0) (0,1) (0,0)
    p.n_pairs = 6;
    p.pairs = (0x5u) | (0x8u << 4) | (0x2u << 8) | (0x4u << 12) | (0x1u << 16) | (0x0u << 20);
  }
  p.bias = g.bias;
  p.ssq_in = g.ssq_in;
  p.ssq_parts = g.ssq_parts;
  p.inv_sqrt_rms_dim = g.rms_dim > 0 ? 1.0f / sqrtf((float)g.rms_dim) : 1.0f;
  p.rms_eps = g.rms_eps;
  p.resid = g.resid;
  p.act = g.act;
  p.out_f32 = g.out_f32;
  p.out_bf16 = static_cast<__nv_bfloat16*>(g.out_bf16);
  p.out_nseg = g.out_bf16 != nullptr ? nseg_of(g.out_precision) : 0;
  AVD_REQUIRE(g.out_bf16 == nullptr || p.out_nseg > 0, "out_precision must be 1, 3 or 6");
  p.ssq_out = g.ssq_out;

  CUtensorMap tmA, tmB;
  AVD_TRY(get_tensor_map_bf16_3d(g.A, g.K, g.M, nseg, (uint64_t)g.K * 2, (uint64_t)g.M * g.K * 2, GEMM_BK, GEMM_BM,
                                 &tmA));
  AVD_TRY(get_tensor_map_bf16_3d(g.W, g.K, g.N, nseg, (uint64_t)g.K * 2, (uint64_t)g.N * g.K * 2, GEMM_BK, bn, &tmB));

  if (bn == 128) return launch_gemm_bn<128, 3>(tmA, tmB, p, s);
  if (bn == 64) return launch_gemm_bn<64, 4>(tmA, tmB, p, s);
  return launch_gemm_bn<32, 4>(tmA, tmB, p, s);
}

}  // namespace avd
