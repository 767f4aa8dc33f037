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

#include <cstdlib>

namespace avd {

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;            // 64 bf16 = 128 B = one swizzle atom row
constexpr int GEMM_THREADS = 192;
constexpr int GEMM_BOX = 4096;         // one epilogue staging box: 32 rows x 128 B (32 fp32 or 64 bf16 columns)

struct GemmParams {
  int M, N, K;
  int n_pairs;
  uint32_t pairs;            // 4 bits per pair: (a_seg << 2) | w_seg, pair i in bits [4i, 4i+4)
  const float* bias;
  const float* ssq_in;
  int ssq_parts;
  float inv_sqrt_rms_dim;    // 1/sqrt(rms_dim)
  float rms_eps;
  int has_resid;             // residual tile comes in through tmR (and the fp32 result is written over it in smem)
  int act;
  int has_f32;               // fp32 result leaves through tmF
  int out_nseg;              // bf16 segments leaving through tmO (0 = none)
  float* ssq_out;
  int cx, cy;                // cluster dims: A tiles are multicast across x (N tiles), W tiles across y (M tiles)
  int dbg;                   // debug: 1 = MMA loop only (loads stop after the first fill), 2 = loads only (no MMA)
  unsigned long long* trace; // debug: per-CTA phase timestamps (avd_debug_trace), nullptr in production
};

__device__ __forceinline__ void trace_mark(const GemmParams& p, int slot) {
  if (p.trace != nullptr) {
    const int cta = blockIdx.y * gridDim.x + blockIdx.x;
    unsigned long long g, c;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(c));
    p.trace[(cta * 16 + slot) * 2] = g;
    p.trace[(cta * 16 + slot) * 2 + 1] = c;
  }
}

template <int BN, int STAGES>
struct GemmSmem {
  static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
  static constexpr int B_BYTES = BN * GEMM_BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int PIPE_BYTES = STAGES * STAGE_BYTES;
  static constexpr int F32_TILE = GEMM_BM * BN * 4;
  static constexpr int BF16_TILE = GEMM_BM * BN * 2;
  static constexpr int TAIL_BYTES = 256 + BN * 4;          // mbarriers + TMEM slot, bias tile
  static int total(bool resid) { return PIPE_BYTES + (resid ? F32_TILE : 0) + TAIL_BYTES; }
};

__device__ __forceinline__ float4 lds128(const uint8_t* p) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(smem_u32(p)));
  return v;
}
__device__ __forceinline__ void sts128(uint8_t* p, float a, float b, float c, float d) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(smem_u32(p)), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void sts128u(uint8_t* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(smem_u32(p)), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// smem map:  [ STAGES x (A 128x64 | W BNx64) ]  [ residual / fp32 result tile, only with has_resid ]  [ barriers | bias ]
// After the last MMA has retired the pipeline region is reused as epilogue staging: fp32 result boxes first (when
// there is no residual tile to overwrite), then out_nseg bf16 tiles.  Every staging box is 32 rows x 128 B with the
// 128-byte swizzle the TMA store expects, written by the warp that owns those 32 TMEM lanes.
template <int BN, int STAGES>
__global__ void __launch_bounds__(GEMM_THREADS) gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA,
                                                               const __grid_constant__ CUtensorMap tmB,
                                                               const __grid_constant__ CUtensorMap tmR,
                                                               const __grid_constant__ CUtensorMap tmF,
                                                               const __grid_constant__ CUtensorMap tmO,
                                                               const GemmParams p) {
  using S = GemmSmem<BN, STAGES>;
  constexpr uint32_t TMEM_COLS = BN < 32 ? 32 : BN;
  extern __shared__ __align__(1024) uint8_t gemm_smem[];
  uint8_t* smem = gemm_smem;
  if ((smem_u32(smem) & 1023u) != 0) __trap();   // 128B-swizzle tiles need 1024-B aligned bases
  uint8_t* res_tile = smem + S::PIPE_BYTES;
  uint8_t* tail = res_tile + (p.has_resid ? S::F32_TILE : 0);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tail);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* acc_bar = empty_bar + STAGES;
  uint64_t* res_bar = acc_bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tail + 192);
  float* bias_s = reinterpret_cast<float*>(tail + 256);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * GEMM_BM;
  const int n0 = blockIdx.x * BN;
  const int kblocks = p.K / GEMM_BK;
  const int iters = kblocks * p.n_pairs;
  const int csize = p.cx * p.cy;
  const int rx = blockIdx.x % p.cx, ry = blockIdx.y % p.cy;      // position inside the cluster
  const uint16_t a_mask = (uint16_t)(((1u << p.cx) - 1u) << (ry * p.cx));
  uint16_t w_mask = 0;
  for (int y = 0; y < p.cy; ++y) w_mask |= (uint16_t)(1u << (rx + y * p.cx));

  if (threadIdx.x == 0) trace_mark(p, 0);
  if (warp == 0 && lane == 0) {
    if (csize > 1 && (int)cluster_ctarank() != ry * p.cx + rx) __trap();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (p.has_resid) tma_prefetch_desc(&tmR);
    if (p.has_f32) tma_prefetch_desc(&tmF);
    if (p.out_nseg) tma_prefetch_desc(&tmO);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], p.cx + p.cy - 1);   // every CTA that reads a slice of this slot's multicasts
    }
    mbar_init(acc_bar, 1);
    mbar_init(res_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<TMEM_COLS>(tmem_slot);
  if (warp >= 2) {
    const int t = threadIdx.x - 64;
    if (t < BN) bias_s[t] = p.bias != nullptr ? p.bias[n0 + t] : 0.0f;   // weights: not produced by the previous kernel
  }
  tc_fence_before();
  if (csize > 1) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot;
  pdl_launch_dependents();
  if (threadIdx.x == 0) trace_mark(p, 1);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      const int a_rows = GEMM_BM / p.cx, w_rows = BN / p.cy;    // this CTA's slice of the A / W tile
      auto load_w = [&](int it) {
        const int st = it % STAGES;
        const int pi = it / kblocks, kb = it - pi * kblocks;
        const int sw = (p.pairs >> (4 * pi)) & 3;
        uint8_t* b_dst = smem + st * S::STAGE_BYTES + S::A_BYTES + ry * w_rows * 128;
        if (p.cy > 1) tma_load_3d_mc(b_dst, &tmB, &full_bar[st], kb * GEMM_BK, n0 + ry * w_rows, sw, w_mask);
        else tma_load_3d(b_dst, &tmB, &full_bar[st], kb * GEMM_BK, n0, sw);
      };
      auto load_a = [&](int it) {
        const int st = it % STAGES;
        const int pi = it / kblocks, kb = it - pi * kblocks;
        const int sa = ((p.pairs >> (4 * pi)) & 0xF) >> 2;
        uint8_t* a_dst = smem + st * S::STAGE_BYTES + rx * a_rows * 128;
        if (p.cx > 1) tma_load_3d_mc(a_dst, &tmA, &full_bar[st], kb * GEMM_BK, m0 + rx * a_rows, sa, a_mask);
        else tma_load_3d(a_dst, &tmA, &full_bar[st], kb * GEMM_BK, m0, sa);
      };
      const int pre = iters < STAGES ? iters : STAGES;
      for (int it = 0; it < pre; ++it) {          // weights first: they do not depend on the previous kernel
        mbar_expect_tx(&full_bar[it], S::STAGE_BYTES);
        load_w(it);
      }
      pdl_wait();
      for (int it = 0; it < pre; ++it) load_a(it);
      if (p.has_resid) {
        mbar_expect_tx(res_bar, S::F32_TILE);
        for (int ch = 0; ch < BN / 32; ++ch)
          for (int q = 0; q < 4; ++q)
            tma_load_3d(res_tile + (ch * 4 + q) * GEMM_BOX, &tmR, res_bar, n0 + ch * 32, m0 + q * 32, 0);
      }
      for (int it = pre; it < iters && p.dbg != 1 && p.dbg != 3; ++it) {
        const int st = it % STAGES;
        mbar_wait(&empty_bar[st], ((it / STAGES) & 1) ^ 1);
        mbar_expect_tx(&full_bar[st], S::STAGE_BYTES);
        load_w(it);
        load_a(it);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(GEMM_BM, BN, 0, 0);
      const uint16_t c_mask = a_mask | w_mask;
      const uint32_t desc_hi = smem_desc_hi_sw128(1024);
      const uint32_t a_lo0 = smem_desc_lo_sw128(smem_u32(smem), 16);
      int st = 0;
      uint32_t ph = 0;
      for (int it = 0; it < iters; ++it) {
        if ((p.dbg != 1 && p.dbg != 3) || it < STAGES) mbar_wait(&full_bar[st], ph);
        tc_fence_after();
        if (it == 0) trace_mark(p, 2);
        const uint32_t a_lo = a_lo0 + st * (S::STAGE_BYTES >> 4);
        const uint32_t b_lo = a_lo + (S::A_BYTES >> 4);
        if (p.dbg != 2) {
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k)     // +32 bytes (2 x 16-byte units) per 16-wide k step
            umma_bf16_lh(tmem_acc, a_lo + 2 * k, b_lo + 2 * k, desc_hi, idesc, (it > 0 || k > 0) ? 1u : 0u);
        }
        // smem slot reusable (cluster-wide, for the multicast writers) once these MMAs have read it
        if (it + STAGES < iters) {
          if (csize > 1) umma_commit_mc(&empty_bar[st], c_mask); else umma_commit(&empty_bar[st]);
        }
        if (++st == STAGES) { st = 0; ph ^= 1; }
      }
      umma_commit(acc_bar);            // accumulator complete
      trace_mark(p, 3);
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..5)
    const int quad = warp & 3;                 // TMEM lane quadrant this warp may access
    const int row = m0 + quad * 32 + lane;
    const bool live = row < p.M;
    const bool box_live = m0 + quad * 32 < p.M;
    pdl_wait();
    float rinv = 1.0f;
    if (p.ssq_in != nullptr && live) {
      float ss = 0.0f;
      for (int q = 0; q < p.ssq_parts; ++q) ss += p.ssq_in[(int64_t)row * p.ssq_parts + q];
      rinv = 1.0f / (sqrtf(ss) * p.inv_sqrt_rms_dim + p.rms_eps);
    }
    named_bar_sync(1, 128);                    // bias tile visible to all four epilogue warps
    if (p.has_resid) mbar_wait(res_bar, 0);
    if (threadIdx.x == 64) trace_mark(p, 4);
    mbar_wait(acc_bar, 0);
    tc_fence_after();
    if (threadIdx.x == 64) trace_mark(p, 5);
    // staging: the fp32 result overwrites the residual tile in place, or (no residual) starts the reused pipeline smem
    uint8_t* f32_stage = p.has_resid ? res_tile : smem;
    uint8_t* bf_stage = smem + ((p.has_f32 && !p.has_resid) ? S::F32_TILE : 0);
    const int sw = lane & 7;
    float ssq_acc = 0.0f;
#pragma unroll 1
    for (int ch = 0; ch < BN / 32; ++ch) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_acc + ((uint32_t)(quad * 32) << 16) + ch * 32, r);
      tmem_ld_wait();
      if (threadIdx.x == 64 && ch == 0) trace_mark(p, 8);
      float v[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = fmaf(__uint_as_float(r[j]), rinv, bias_s[ch * 32 + j]);
      uint8_t* fbox = f32_stage + (ch * 4 + quad) * GEMM_BOX + lane * 128;
      if (p.has_resid) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float4 x4 = lds128(fbox + ((c ^ sw) << 4));
          v[4 * c] += x4.x; v[4 * c + 1] += x4.y; v[4 * c + 2] += x4.z; v[4 * c + 3] += x4.w;
        }
      }
      if (p.act == 1) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
      }
      if (threadIdx.x == 64 && ch == 0) trace_mark(p, 9);
      if (p.has_f32) {
#pragma unroll
        for (int c = 0; c < 8; ++c) sts128(fbox + ((c ^ sw) << 4), v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
      }
      if (p.ssq_out != nullptr) {
#pragma unroll
        for (int j = 0; j < 32; ++j) ssq_acc = fmaf(v[j], v[j], ssq_acc);
        if (ch & 1) {  // a 64-column part is complete
          if (live) p.ssq_out[(int64_t)row * (p.N / 64) + ((n0 + ch * 32) >> 6)] = ssq_acc;
          ssq_acc = 0.0f;
        }
      }
      if (p.out_nseg > 0) {
        for (int sg = 0; sg < p.out_nseg; ++sg) {
          uint32_t w[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const __nv_bfloat16 h0 = __float2bfloat16_rn(v[2 * j]);
            const __nv_bfloat16 h1 = __float2bfloat16_rn(v[2 * j + 1]);
            v[2 * j] -= __bfloat162float(h0);
            v[2 * j + 1] -= __bfloat162float(h1);
            w[j] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
          }
          uint8_t* bbox = bf_stage + ((sg * (BN / 64) + (ch >> 1)) * 4 + quad) * GEMM_BOX + lane * 128;
          const int c0 = (ch & 1) * 4;
#pragma unroll
          for (int c = 0; c < 4; ++c)
            sts128u(bbox + (((c0 + c) ^ sw) << 4), w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
        }
      }
      if (threadIdx.x == 64 && ch == 0) trace_mark(p, 10);
      fence_async_smem();      // staging writes (generic proxy) -> visible to the TMA store (async proxy)
      __syncwarp();
      if (threadIdx.x == 64 && ch == 0) trace_mark(p, 11);
      if (lane == 0 && box_live) {
        if (p.has_f32) tma_store_3d(&tmF, f32_stage + (ch * 4 + quad) * GEMM_BOX, n0 + ch * 32, m0 + quad * 32, 0);
        if (p.out_nseg > 0 && (ch & 1)) {
          for (int sg = 0; sg < p.out_nseg; ++sg)
            tma_store_3d(&tmO, bf_stage + ((sg * (BN / 64) + (ch >> 1)) * 4 + quad) * GEMM_BOX, n0 + (ch >> 1) * 64,
                         m0 + quad * 32, sg);
        }
        tma_store_commit();
      }
      if (threadIdx.x == 64 && ch == 0) trace_mark(p, 12);
      if (threadIdx.x == 64 && ch == BN / 32 - 1) trace_mark(p, 13);
    }
    if (lane == 0) tma_store_wait_read();      // smem must stay alive until the bulk stores have read it
    if (threadIdx.x == 64) trace_mark(p, 6);
  }

  tc_fence_before();
  if (csize > 1) cluster_sync_all(); else __syncthreads();   // peers may still multicast into / arrive on this CTA
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<TMEM_COLS>(tmem_acc);
  }
  if (threadIdx.x == 0) trace_mark(p, 7);
}

unsigned long long* g_dbg_trace = nullptr;
int64_t g_dbg_trace_slots = 0;

static int env_int(const char* name, int dflt) {
  const char* v = std::getenv(name);
  return v ? std::atoi(v) : dflt;
}

template <int BN, int STAGES>
static int launch_gemm_bn(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmR, const CUtensorMap& tmF,
                          const CUtensorMap& tmO, const GemmParams& p, cudaStream_t s) {
  using S = GemmSmem<BN, STAGES>;
  static bool attr_set = false;
  if (!attr_set) {
    AVD_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  S::total(true)));
    attr_set = true;
  }
  const int mt = (p.M + GEMM_BM - 1) / GEMM_BM;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(p.N / BN, (mt + p.cy - 1) / p.cy * p.cy);
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = S::total(p.has_resid != 0);
  cfg.stream = s;
  cudaLaunchAttribute attrs[2];
  int na = 0;
  if (p.cx * p.cy > 1) {
    attrs[na].id = cudaLaunchAttributeClusterDimension;
    attrs[na].val.clusterDim.x = p.cx;
    attrs[na].val.clusterDim.y = p.cy;
    attrs[na].val.clusterDim.z = 1;
    ++na;
  }
  if (pdl_enabled()) {
    attrs[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attrs;
  cfg.numAttrs = na;
  LaunchScope ls(KIND_GEMM, s);
  AVD_CUDA(cudaLaunchKernelEx(&cfg, gemm_tc_kernel<BN, STAGES>, tmA, tmB, tmR, tmF, tmO, p));
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
  AVD_REQUIRE(g.out_bf16 == nullptr || g.N % 64 == 0, "bf16 output needs N % 64 == 0");
  AVD_REQUIRE(g.resid == nullptr || g.out_f32 != nullptr, "a residual needs the fp32 output");
  const int out_nseg = g.out_bf16 != nullptr ? nseg_of(g.out_precision) : 0;
  AVD_REQUIRE(g.out_bf16 == nullptr || out_nseg > 0, "out_precision must be 1, 3 or 6");

  // tile width: widest BN that divides N, still gives >= ~1 wave of CTAs on 148 SMs, and whose epilogue staging
  // (fp32 tile unless it overwrites the residual tile, plus out_nseg bf16 tiles) fits the reused pipeline smem
  const int mt = (g.M + GEMM_BM - 1) / GEMM_BM;
  auto staging = [&](int bn) {
    return ((g.out_f32 != nullptr && g.resid == nullptr) ? GEMM_BM * bn * 4 : 0) + out_nseg * GEMM_BM * bn * 2;
  };
  int bn = 32;
  if (g.N % 128 == 0 && (int64_t)mt * (g.N / 128) >= 120 && staging(128) <= GemmSmem<128, 3>::PIPE_BYTES) bn = 128;
  else if (g.N % 64 == 0 && staging(64) <= GemmSmem<64, 4>::PIPE_BYTES) bn = 64;
  if ((g.ssq_out != nullptr || g.out_bf16 != nullptr) && bn < 64) bn = 64;
  const int force_bn = env_int("AVD_GEMM_BN", 0);
  if (force_bn == 128 && g.N % 128 == 0 && staging(128) <= GemmSmem<128, 3>::PIPE_BYTES) bn = 128;
  if (force_bn == 64 && g.N % 64 == 0) bn = 64;
  AVD_REQUIRE(staging(bn) <= (bn == 128 ? GemmSmem<128, 3>::PIPE_BYTES
                                        : bn == 64 ? GemmSmem<64, 4>::PIPE_BYTES : GemmSmem<32, 4>::PIPE_BYTES),
              "epilogue staging does not fit: use fewer outputs per GEMM");

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
  p.has_resid = g.resid != nullptr;
  p.act = g.act;
  p.has_f32 = g.out_f32 != nullptr;
  p.out_nseg = out_nseg;
  p.ssq_out = g.ssq_out;
  p.dbg = env_int("AVD_GEMM_DBG", 0);

  // cluster shape: A tiles multicast across cx N-tiles, W tiles across cy M-tiles (cx * cy <= 8)
  const int nt = g.N / bn;
  int cx = env_int("AVD_GEMM_CX", 1), cy = env_int("AVD_GEMM_CY", 1);
  while (cx > 1 && (nt % cx != 0 || GEMM_BM / cx % 8 != 0)) cx >>= 1;
  while (cy > 1 && (cy > mt || bn / cy % 8 != 0)) cy >>= 1;
  while (cx * cy > 8) { if (cx >= cy) cx >>= 1; else cy >>= 1; }
  p.cx = cx < 1 ? 1 : cx;
  p.cy = cy < 1 ? 1 : cy;
  const int64_t n_ctas = (int64_t)nt * ((mt + p.cy - 1) / p.cy * p.cy);
  p.trace = (g_dbg_trace != nullptr && n_ctas * 32 <= g_dbg_trace_slots) ? g_dbg_trace : nullptr;

  CUtensorMap tmA, tmB, tmR, tmF, tmO;
  AVD_TRY(get_tensor_map_3d(g.A, 2, g.K, g.M, nseg, (uint64_t)g.K * 2, (uint64_t)g.M * g.K * 2, GEMM_BK,
                            GEMM_BM / p.cx, &tmA));
  AVD_TRY(get_tensor_map_3d(g.W, 2, g.K, g.N, nseg, (uint64_t)g.K * 2, (uint64_t)g.N * g.K * 2, GEMM_BK, bn / p.cy,
                            &tmB));
  tmR = tmA; tmF = tmA; tmO = tmA;   // placeholders when unused (never dereferenced)
  if (g.resid != nullptr)
    AVD_TRY(get_tensor_map_3d(g.resid, 4, g.N, g.M, 1, (uint64_t)g.N * 4, (uint64_t)g.M * g.N * 4, 32, 32, &tmR));
  if (g.out_f32 != nullptr)
    AVD_TRY(get_tensor_map_3d(g.out_f32, 4, g.N, g.M, 1, (uint64_t)g.N * 4, (uint64_t)g.M * g.N * 4, 32, 32, &tmF));
  if (g.out_bf16 != nullptr)
    AVD_TRY(get_tensor_map_3d(g.out_bf16, 2, g.N, g.M, out_nseg, (uint64_t)g.N * 2, (uint64_t)g.M * g.N * 2, 64, 32,
                              &tmO));

  const int deep = env_int("AVD_GEMM_DEEP", 0);
  if (bn == 128 && deep) return launch_gemm_bn<128, 5>(tmA, tmB, tmR, tmF, tmO, p, s);
  if (bn == 64 && deep == 1) return launch_gemm_bn<64, 6>(tmA, tmB, tmR, tmF, tmO, p, s);
  if (bn == 64 && deep == 2) return launch_gemm_bn<64, 7>(tmA, tmB, tmR, tmF, tmO, p, s);
  if (bn == 128) return launch_gemm_bn<128, 3>(tmA, tmB, tmR, tmF, tmO, p, s);
  if (bn == 64) return launch_gemm_bn<64, 4>(tmA, tmB, tmR, tmF, tmO, p, s);
  return launch_gemm_bn<32, 4>(tmA, tmB, tmR, tmF, tmO, p, s);
}

}  // namespace avd
