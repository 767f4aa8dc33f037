// avd_gemm_ps2.cu -- K4, persistent form on CTA PAIRS (tcgen05 cta_group::2): out = epilogue( A[M,K] x W[N,K]^T ).
//
// Why: the persistent kernel (avd_gemm_ps.cu) is bound by the shared-memory port -- per 64-wide k-block one SM's port
// carries the TMA writes AND the tensor core's reads of (128 + BN) x 128 B, i.e. 768 cycles at BN = 256 against 512
// cycles of MMA.  With cta_group::2 the two SMs of a TPC compute ONE 256 x 256 tile: each stages its own 128 rows of A
// and HALF of the weight tile (128 rows), the pair's M = 256 MMA reads both halves, and each SM keeps its 128
// accumulator rows in its own TMEM.  Per SM and k-block the port then moves 2 x (16 + 16) KB = 512 cycles for the same
// 512 cycles of MMA: port and tensor pipe are balanced instead of 3:2.
//
// One cluster of two CTAs per TPC (74 pairs on 148 SMs), 384 threads per CTA, pair tiles t = pair, pair + n_pairs, ...:
//   warps 0,1   TMA producers of BOTH CTAs: k-block iteration `it` belongs to producer it & 1 and ring slot it % stages;
//               A box (own 128 rows) + W box (own 128 of the tile's 256 weight rows) complete on the LEADER's full
//               barrier (cp.async.bulk.tensor ... .cta_group::2 with the barrier's shared::cluster address); the leader
//               expects the bytes of both CTAs.  A slot is refilled after tcgen05.commit.cta_group::2 has arrived on
//               this CTA's own empty barrier (multicast to both CTAs).
//   warp 2      (leader CTA only, one lane) MMA issuer: tcgen05.mma.cta_group::2.kind::f16 M=256 N=256 K=16, four per
//               k-block = 512 cycles of tensor pipe per round, which one issuing thread sustains (tools/mma_probe.cu:
//               ~400-500 cycles per {wait, 4 MMAs, commit} round).  TMEM is double-buffered (2 x 256 columns per CTA);
//               before reusing a buffer the issuer waits for the epilogue warps of BOTH CTAs (16 arrivals, the peer's
//               through mbarrier.arrive.shared::cluster).
//   warps 4..11 epilogue of this CTA's 128 x 256 accumulator: the epilogue of gemm_ps_kernel (avd_gemm_ps_epilogue.inc).
#include "avd_gemm_common.cuh"

namespace avd {

constexpr int P2_THREADS = 384;
constexpr int P2_MAX_STAGES = 8;
constexpr int P2_BN = 256;
constexpr int P2_TAIL_BYTES = 3072;          // as PS_TAIL_BYTES: barriers, ssq exchange, two bias tiles
constexpr int P2_SMEM_MAX = 232448;

struct P2Plan {
  unsigned int* dbg;
  int stages;
  int tiles_n, tiles_total;        // pair tiles (256 x 256)
  int f32_off, bf_off, tail_off;
};

// (relaxed: the arrival hands back a TMEM buffer whose reads were ordered by tcgen05.wait::ld + tcgen05.fence; a
// release at cluster scope makes the warp wait for its outstanding global stores first, once per tile and warp)
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(P2_THREADS, 1) gemm_ps2_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                 const __grid_constant__ CUtensorMap tmB,
                                                                 const __grid_constant__ CUtensorMap tmR,
                                                                 const __grid_constant__ CUtensorMap tmO,
                                                                 const GemmParams p, const P2Plan plan) {
  constexpr int BN = P2_BN;
  constexpr bool CS = true;                            // one accumulator per tile (the epilogue's name for it)
  constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
  constexpr int B_BYTES = (BN / 2) * GEMM_BK * 2;      // this CTA's half of the weight tile
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr uint32_t TMEM_COLS = 512;                  // 2 buffers x 256 columns
  constexpr int NCH = BN / 32;
  extern __shared__ __align__(1024) uint8_t p2_smem[];
  uint8_t* smem = p2_smem;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* tail = smem + plan.tail_off;
  uint64_t* full = reinterpret_cast<uint64_t*>(tail);          // used in the leader only
  uint64_t* empty = full + P2_MAX_STAGES;
  uint64_t* acc_full = empty + P2_MAX_STAGES;
  uint64_t* acc_empty = acc_full + 2;                          // used in the leader only: 16 arrivals
  uint64_t* res_bar = acc_empty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_bar + 16);
  float* ssq_x = reinterpret_cast<float*>(tail + 512);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();                     // 0 = leader
  const int pair = (int)(blockIdx.x >> 1);
  auto wait = [&](uint64_t* bar, uint32_t parity, unsigned tag, int counter) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
      if (++spins > (1u << 26)) {
        unsigned int* dbg = plan.dbg;
        const unsigned idx = atomicAdd_system(&dbg[0], 1u);
        if (idx < 30) {
          unsigned int* r = dbg + 8 + idx * 8;
          r[0] = 200u + tag; r[1] = blockIdx.x; r[2] = threadIdx.x; r[3] = (unsigned)counter;
          r[4] = parity; r[5] = (unsigned)p.N; r[6] = (unsigned)p.K; r[7] = gridDim.x * 1000u + (unsigned)BN;
          __threadfence_system();
        }
        __nanosleep(4000000);
        __trap();
      }
    }
  };
  const int stages = plan.stages;
  const int G = (int)(gridDim.x >> 1);                         // pairs
  const int kblocks = p.K / GEMM_BK;
  const int ipt = kblocks * p.n_pairs;
  const int my_tiles = (plan.tiles_total - pair + G - 1) / G;
  const int total_it = my_tiles * ipt;
  unsigned long long* const trace = p.trace;

  if (threadIdx.x == 0) trace_mark(trace, 0);
  if (warp == 0) {
    if (lane < 2 * P2_MAX_STAGES) mbar_init(&full[lane], 1);    // full (expect_tx by the leader), empty (one commit)
    if (lane < 2) mbar_init(&acc_full[lane], 1);                // one commit
    if (lane < 2) mbar_init(&acc_empty[lane], 16);              // the epilogue warps of both CTAs
    if (lane < 16) mbar_init(&res_bar[lane], 1);
    fence_barrier_init();
  } else if (warp == 1 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (MODE == 0 ? (p.has_resid != 0) : (MODE == 3)) tma_prefetch_desc(&tmR);
  }
  if (warp == 2) tmem_alloc_2cta<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  cluster_sync_all();            // the peer must see initialised barriers before it signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  if (threadIdx.x == 0) trace_mark(trace, 1);
  const uint32_t full_leader = mapa_u32(smem_u32(full), 0);
  const uint32_t acc_empty_leader = mapa_u32(smem_u32(acc_empty), 0);

  if (warp < 2) {
    // ------------------------------------------------------------------ TMA producer h: iterations it = h (mod 2)
    if (lane == 0) {
      const int h = warp;
      const uint32_t pairs = p.pairs;
      const int tiles_n = plan.tiles_n;
      const int n_pairs = p.n_pairs;
      struct Cur { int tl, pi, kb, m0, n0; };
      auto tile_origin = [&](Cur& c) {
        const int t = pair + c.tl * G;
        const int mt = t / tiles_n;
        c.m0 = mt * 2 * GEMM_BM + (int)rank * GEMM_BM;          // this CTA's 128 rows of the pair tile
        c.n0 = (t - mt * tiles_n) * BN + (int)rank * (BN / 2);  // this CTA's 128 of the tile's 256 weight rows
      };
      auto advance = [&](Cur& c) {
        if (++c.kb == kblocks) {
          c.kb = 0;
          if (++c.pi == n_pairs) { c.pi = 0; ++c.tl; tile_origin(c); }
        }
      };
      Cur cb{0, 0, 0, 0, 0};
      tile_origin(cb);
      if (h == 1) advance(cb);
      Cur ca = cb;
      const int pre_end = total_it < stages ? total_it : stages;
      for (int it = h; it < pre_end; it += 2) {          // weights of the first ring pass run ahead of the dependency
        if (rank == 0) mbar_expect_tx(&full[it], 2 * STAGE_BYTES);
        tma_load_3d_2sm(smem + it * STAGE_BYTES + A_BYTES, &tmB, full_leader + 8u * it, cb.kb * GEMM_BK, cb.n0,
                        (pairs >> (4 * cb.pi)) & 3);
        advance(cb); advance(cb);
      }
      pdl_wait();
      int it = h;
      for (; it < pre_end; it += 2) {
        tma_load_3d_2sm(smem + it * STAGE_BYTES, &tmA, full_leader + 8u * it, ca.kb * GEMM_BK, ca.m0,
                        (pairs >> (4 * ca.pi + 2)) & 3);
        advance(ca); advance(ca);
      }
      int slot = it % stages;
      uint32_t ph = ((it / stages) & 1) ^ 1;             // parity of the consumption this refill waits for
      for (; it < total_it; it += 2) {
        wait(&empty[slot], ph, 1, it);
        const uint32_t pr = pairs >> (4 * ca.pi);
        uint8_t* dst = smem + slot * STAGE_BYTES;
        if (rank == 0) mbar_expect_tx(&full[slot], 2 * STAGE_BYTES);
        tma_load_3d_2sm(dst, &tmA, full_leader + 8u * slot, ca.kb * GEMM_BK, ca.m0, (pr >> 2) & 3);
        tma_load_3d_2sm(dst + A_BYTES, &tmB, full_leader + 8u * slot, ca.kb * GEMM_BK, ca.n0, pr & 3);
        advance(ca); advance(ca);
        slot += 2;
        if (slot >= stages) { slot -= stages; ph ^= 1; }
      }
      if (h == 0) trace_mark(trace, 10);
    }
  } else if (warp == 2) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA only)
    if (lane == 0 && rank == 0) {
      const uint32_t idesc = make_idesc_bf16(2 * GEMM_BM, BN, 0, 0, p.f16_in);
      const uint32_t desc_hi = smem_desc_hi_sw128(1024);
      const uint32_t a_lo0 = smem_desc_lo_sw128(smem_u32(smem), 16);
      int slot = 0;
      uint32_t ph = 0;
      for (int i = 0; i < my_tiles; ++i) {
        const int b = i & 1;
        wait(&acc_empty[b], ((i >> 1) & 1) ^ 1, 2, i);     // both CTAs have drained this buffer (tile i - 2)
        tc_fence_after();
        const uint32_t acc = tmem_base + b * BN;
        for (int j = 0; j < ipt; ++j) {
          wait(&full[slot], ph, 3, i * ipt + j);
          tc_fence_after();
          if (i == 0 && j == 0) trace_mark(trace, 2);
          const uint32_t a_lo = a_lo0 + slot * (STAGE_BYTES >> 4);
          const uint32_t b_lo = a_lo + (A_BYTES >> 4);
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k)
            umma_2cta_lh(acc, a_lo + 2 * k, b_lo + 2 * k, desc_hi, idesc, (j > 0 || k > 0) ? 1u : 0u);
          umma_commit_2cta(&empty[slot]);                  // frees the slot in both CTAs
          if (++slot == stages) { slot = 0; ph ^= 1; }
        }
        umma_commit_2cta(&acc_full[b]);
        if (i == my_tiles - 1) trace_mark(trace, 3);
      }
    }
#define PS_EPI_TILE(i) (pair + (i) * G)
#define PS_EPI_ROW0(mt) ((mt) * 2 * GEMM_BM + (int)rank * GEMM_BM)      /* this CTA's 128 of the pair's 256 rows */
#define PS_EPI_ACC_EMPTY(b) mbar_arrive_cluster(acc_empty_leader + 8u * (b))   /* the LEADER's issuer waits for both CTAs */
#include "avd_gemm_ps_epilogue.inc"
#undef PS_EPI_TILE
#undef PS_EPI_ROW0
#undef PS_EPI_ACC_EMPTY

  __syncwarp();   // lanes 1..31 of the role warps wait here for their lane 0
  tc_fence_before();
  cluster_sync_relaxed();   // neither CTA of a pair may retire while the other still signals its barriers or runs MMAs on its smem
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_2cta<TMEM_COLS>(tmem_base);
  }
  if (threadIdx.x == 0) trace_mark(trace, 7);
}

template <int MODE_DUMMY = 0>
static int launch_p2(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmR, const CUtensorMap& tmO,
                     const GemmParams& p, const P2Plan& plan, int smem_bytes, int grid, int mode, cudaStream_t s) {
  using KernelFn = void (*)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const CUtensorMap, const GemmParams,
                            const P2Plan);
  static const KernelFn kernels[4] = {gemm_ps2_kernel<0>, gemm_ps2_kernel<1>, gemm_ps2_kernel<2>, gemm_ps2_kernel<3>};
  AVD_TRY(ensure_dyn_smem(reinterpret_cast<const void*>(kernels[mode]), P2_SMEM_MAX));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(P2_THREADS);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = s;
  cudaLaunchAttribute attrs[2];
  int na = 0;
  if (pdl_enabled()) {
    attrs[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  attrs[na].id = cudaLaunchAttributeClusterDimension;
  attrs[na].val.clusterDim.x = 2;
  attrs[na].val.clusterDim.y = 1;
  attrs[na].val.clusterDim.z = 1;
  ++na;
  cfg.attrs = attrs;
  cfg.numAttrs = na;
  if (kind_skipped(KIND_GEMM)) return 0;
  LaunchScope ls(KIND_GEMM, s);
  AVD_CUDA(cudaLaunchKernelEx(&cfg, kernels[mode], tmA, tmB, tmR, tmO, p, plan));
  return 0;
}

// 1 = launched on CTA pairs, 0 = this GEMM does not qualify (the caller uses the single-CTA persistent kernel), < 0 error
int launch_gemm_ps2(const avd_gemm_args& g, GemmParams p, int nseg, int mode, int sms, cudaStream_t s) {
  if (g.N % P2_BN != 0 || p.n_pairs < 1) return 0;
  const int out_nseg = p.out_nseg;
  const int stage = (GEMM_BM + P2_BN / 2) * GEMM_BK * 2;
  const int f32_bytes = (g.out_f32 != nullptr || g.resid != nullptr) ? 16 * GEMM_BOX : 0;
  const bool fast16 = g.out_bf16 != nullptr && out_nseg == 1 && g.out_f32 == nullptr && g.resid == nullptr &&
                      g.ssq_out == nullptr && (g.act == 0 || g.act == 2);
  const int bf_bytes = out_nseg * 4 * GEMM_BOX * (fast16 ? 2 : 1);   // MODE 1 / 2 double-buffer their staging boxes
  int stages = (P2_SMEM_MAX - f32_bytes - bf_bytes - P2_TAIL_BYTES) / stage;
  if (stages > P2_MAX_STAGES) stages = P2_MAX_STAGES;
  const int cap = options().ps_max_stages;
  if (cap > 0 && stages > cap) stages = cap;
  if (stages < 3) return 0;
  P2Plan plan{};
  plan.stages = stages;
  plan.f32_off = stages * stage;
  plan.bf_off = plan.f32_off + f32_bytes;
  plan.tail_off = plan.bf_off + bf_bytes;
  const int smem_bytes = plan.tail_off + P2_TAIL_BYTES;
  plan.dbg = ps_debug_device_ptr();
  AVD_REQUIRE(plan.dbg != nullptr, "cannot allocate the barrier-timeout record");
  const int mt2 = (g.M + 2 * GEMM_BM - 1) / (2 * GEMM_BM);
  plan.tiles_n = g.N / P2_BN;
  plan.tiles_total = mt2 * plan.tiles_n;
  const int max_pairs = sms / 2;
  const int n_pairs_grid = plan.tiles_total < max_pairs ? plan.tiles_total : max_pairs;
  const int grid = 2 * n_pairs_grid;
  p.trace = (g_dbg_trace != nullptr && (int64_t)grid * 32 <= g_dbg_trace_slots) ? g_dbg_trace : nullptr;
  CUtensorMap tmA, tmB, tmR;
  AVD_TRY(get_tensor_map_3d(g.A, 2, g.K, g.M, nseg, (uint64_t)g.K * 2, (uint64_t)g.M * g.K * 2, GEMM_BK, GEMM_BM, &tmA));
  AVD_TRY(get_tensor_map_3d(g.W, 2, g.K, g.N, nseg, (uint64_t)g.K * 2, (uint64_t)g.N * g.K * 2, GEMM_BK, P2_BN / 2, &tmB));
  tmR = tmA;
  if (g.resid != nullptr)
    AVD_TRY(get_tensor_map_3d(g.resid, 4, g.N, g.M, 1, (uint64_t)g.N * 4, (uint64_t)g.M * g.N * 4, 32, 32, &tmR));
  CUtensorMap tmO = tmA;
  if (g.out_bf16 != nullptr && out_nseg == 1)
    AVD_TRY(get_tensor_map_3d(g.out_bf16, 2, g.N, g.M, 1, (uint64_t)g.N * 2, (uint64_t)g.M * g.N * 2, 64, 32, &tmO));
#ifndef AVD_DEBUG_HOOKS
  if (p.trace != nullptr) mode = 0;
#endif
  AVD_TRY(launch_p2<>(tmA, tmB, tmR, tmO, p, plan, smem_bytes, grid, mode, s));
  return 1;
}

}  // namespace avd
