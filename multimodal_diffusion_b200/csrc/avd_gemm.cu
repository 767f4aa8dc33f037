// avd_gemm.cu -- K4: the tcgen05 GEMM of the MMDT blocks, the noise head and (via operand splitting) the
// fp32-parity mode.
//
//   out = epilogue( A[M,K] x W[N,K]^T ),   A, W bf16 K-major, fp32 accumulation in TMEM.
//
// One CTA = one 128 x BN output tile, 256 threads, TWO independent load/MMA pipelines that split the k-blocks
// (pipeline h takes k-block iterations h, h+2, ...) and accumulate into two TMEM tiles that the epilogue adds.
// Why two: on B200 a single thread can only push one {mbarrier wait, 4 x tcgen05.mma, commit} round per ~500
// cycles and one {wait, expect_tx, 2 x TMA} round per ~300 cycles (tools/mma_probe.cu, tools/tma_probe.cu), while a
// 128x64x64 block of MMAs executes in ~220 cycles; two issuing warps hide each other's barrier latency and bring
// the tensor pipe to its floor (64 cycles per M128 N128 K16 MMA, 48 at N64).
//
// Warp roles:
//   warps 0,1   TMA producers of pipeline 0 / 1: cp.async.bulk.tensor (3-D maps {K, rows, segment}) into a ring of
//               128B-swizzled smem stages, mbarrier complete_tx.  Weight tiles of the first ring pass are issued
//               BEFORE griddepcontrol.wait (they do not depend on the previous kernel), activations after it.
//   warps 2,3   MMA issuers of pipeline 0 / 1 (warp 2 also owns the TMEM allocation): one lane issues
//               tcgen05.mma.cta_group::1.kind::f16 (M=128, N=BN, K=16) x 4 per 64-wide k-block; tcgen05.commit
//               releases the smem stage / signals the epilogue
//   all 8 warps epilogue: tcgen05.ld 32 lanes x 32 columns of both accumulators -> registers; thread == output row
//               v = (acc0+acc1) * rinv(row) + bias (+ resid) [gelu]; results are staged in 128B-swizzled smem boxes
//               and leave with 16-byte coalesced stores (fp32 and/or bf16 segments; bulk tensor stores were measured
//               equal here); per-row partial sums of squares feed the NEXT GEMM's RMSNorm row scale.
//
// Precision modes are realised by the k-loop: for mode p the operands are stored as hi/(mid/)lo bf16
// segments and the loop runs over a list of (A segment, W segment) pairs, smallest products first, all
// accumulating into the same TMEM tiles (1, 3 or 6 passes over K).
#include "avd_gemm_common.cuh"

#include <cstdlib>

namespace avd {

// STAGES = ring depth of EACH of the two pipelines.  PAIR: the CTA is one half of a cta_group::2 pair and stages only
// half of the weight tile's rows (the other half lives in the peer's shared memory).
template <int BN, int STAGES, bool PAIR = false>
struct GemmSmem {
  static constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
  static constexpr int B_BYTES = (PAIR ? BN / 2 : BN) * GEMM_BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int PIPE_BYTES = 2 * STAGES * STAGE_BYTES;
  static constexpr int F32_TILE = GEMM_BM * BN * 4;
  static constexpr int BF16_TILE = GEMM_BM * BN * 2;
  static constexpr int TAIL_BYTES = 256 + BN * 4 + 1024 + 512;   // mbarriers + TMEM slot, bias tile, ssq exchange, rinv
  static constexpr int total(bool resid) { return PIPE_BYTES + (resid ? F32_TILE : 0) + TAIL_BYTES; }
};

// smem map:  [ 2 pipelines x STAGES x (A 128x64 | W BNx64) ]  [ residual / fp32 result tile, only with has_resid ]
//            [ barriers | bias ]
// After the last MMA has retired the pipeline region is reused as epilogue staging: fp32 result boxes first (when
// there is no residual tile to overwrite), then out_nseg bf16 tiles.  Every staging box is 32 rows x 128 B with the
// 128-byte swizzle the TMA store expects, written by the warp that owns those 32 TMEM lanes.
// MODE specialises the epilogue at compile time (the runtime-flag version executes ~1.5x the instructions per chunk, and
// the epilogue is instruction-issue bound): 0 generic; 1 one 16-bit output segment only (QKV); 2 the same with the
// packed GELU (fc1); 3 residual + fp32 + one 16-bit segment + row sums of squares (out-proj, fc2).
//
// PAIR: two CTAs with consecutive blockIdx.x (cluster 2x1x1, the two SMs of a TPC) compute a 256 x BN tile with
// tcgen05.mma.cta_group::2.  Each CTA loads its own 128 rows of A and BN/2 of the W rows, every TMA completes on the
// LEADER's full barrier, the leader's two issuer threads run the M=256 MMAs and their commits arrive on both CTAs'
// barriers; each CTA then drains its own 128 accumulator rows exactly as in the single-CTA kernel.  Per SM the smem port
// moves (16 KB + BN/2 x 128 B) x 2 per k-block instead of (16 KB + BN x 128 B) x 2.
template <int BN, int STAGES, int MODE, bool PAIR>
__global__ void __launch_bounds__(GEMM_THREADS, 1) gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                  const __grid_constant__ CUtensorMap tmB,
                                                                  const __grid_constant__ CUtensorMap tmR,
                                                                  const GemmParams p) {
  using S = GemmSmem<BN, STAGES, PAIR>;
  constexpr uint32_t TMEM_COLS = BN <= 32 ? 64 : BN <= 64 ? 128 : BN <= 128 ? 256 : 512;   // two accumulators
  extern __shared__ __align__(1024) uint8_t gemm_smem[];
  uint8_t* smem = gemm_smem;
  if ((smem_u32(smem) & 1023u) != 0) __trap();   // 128B-swizzle tiles need 1024-B aligned bases
  uint8_t* res_tile = smem + S::PIPE_BYTES;
  const bool k_resid = MODE == 0 ? (p.has_resid != 0) : (MODE == 3);
  uint8_t* tail = res_tile + (k_resid ? S::F32_TILE : 0);
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail);
  // per pipeline h: full[h*STAGES + s], empty[(2 + h)*STAGES + s]; then acc[2], res
  uint64_t* acc_bar = bars + 4 * STAGES;
  uint64_t* res_bar = acc_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tail + 240);
  float* bias_s = reinterpret_cast<float*>(tail + 256);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // PAIR grids put the row tiles on x: a CTA pair is two CTAs whose cluster rank differs in the last bit, and the
  // launch is rejected ("cluster misconfiguration") unless the cluster's x extent is even
  const int m0 = (PAIR ? blockIdx.x : blockIdx.y) * GEMM_BM;
  const int n0 = (PAIR ? blockIdx.y : blockIdx.x) * BN;
  const uint32_t rank = PAIR ? cluster_ctarank() : 0u;   // 0 = the pair's leader (issues the MMAs, owns the full barriers)
  const int n0b = n0 + (PAIR ? (int)rank * (BN / 2) : 0);   // first W row this CTA stages
  const int kblocks = p.K / GEMM_BK;
  const int iters = kblocks * p.n_pairs;
  unsigned long long* const trace = p.trace;

  if (threadIdx.x == 0) trace_mark(trace, 0);
  if (warp == 0) {                       // one barrier per lane instead of a serial loop on the critical setup path
    if (lane < 4 * STAGES + 3) mbar_init(&bars[lane], 1);
    fence_barrier_init();
    if (MODE == 0 && lane == 0) trace_mark(trace, 14);
  } else if (warp == 1 && lane == 0) {   // descriptor prefetches off the initialising warp
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (k_resid) tma_prefetch_desc(&tmR);
  }
  if (warp == 2) {
    if (PAIR) tmem_alloc_2cta<TMEM_COLS>(tmem_slot); else tmem_alloc<TMEM_COLS>(tmem_slot);
    if (MODE == 0 && lane == 0) trace_mark(trace, 15);
  }
  // bias (weights: not produced by the previous kernel): the global loads are issued here but only land in smem after
  // the setup barrier, so their latency is off the path to the first TMA / MMA
  float bias_r[2] = {0.0f, 0.0f};
  if (warp >= 4 && p.bias != nullptr) {
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int t = threadIdx.x - 128 + u * 128;
      if (t < BN) bias_r[u] = __ldg(p.bias + n0 + t);
    }
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all(); else __syncthreads();   // the peer must see initialised barriers before it signals them
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot;
  pdl_launch_dependents();
  if (threadIdx.x == 0) trace_mark(trace, 1);

  if (warp < 2) {
    // ------------------------------------------------------------------ TMA producer of pipeline h
    // (!PAIR: the warp runs converged, an elected lane issues the expect_tx / TMA instructions; PAIR: lane 0)
    if (PAIR ? lane == 0 : true) {
      const int h = warp;
      const int count = (iters - h + 1) >> 1;            // iterations h, h+2, ...
      uint64_t* full = bars + h * STAGES;
      uint64_t* empty = bars + (2 + h) * STAGES;
      uint8_t* ring = smem + h * STAGES * S::STAGE_BYTES;
      const uint32_t pairs = p.pairs;
      int kb = h, pi = 0;
      while (kb >= kblocks) { kb -= kblocks; ++pi; }
      int kb_w = kb, pi_w = pi;                          // the weight loads of the first ring pass run ahead
      const int pre = count < STAGES ? count : STAGES;
      // PAIR: both CTAs' boxes complete on the leader's barrier, which expects the bytes of both
      const uint32_t full0 = PAIR ? mapa_u32(smem_u32(full), 0) : 0u;
      auto expect = [&](int j) {
        if (!PAIR) mbar_expect_tx_elect(&full[j], S::STAGE_BYTES);
        else if (rank == 0) mbar_expect_tx(&full[j], 2 * S::STAGE_BYTES);
      };
      auto load = [&](void* dst, const CUtensorMap* tm, int j, int c0, int c1, int c2) {
        if (PAIR) tma_load_3d_2sm(dst, tm, full0 + 8u * j, c0, c1, c2);
        else tma_load_3d_elect(dst, tm, &full[j], c0, c1, c2);
      };
      for (int j = 0; j < pre; ++j) {
        expect(j);
        load(ring + j * S::STAGE_BYTES + S::A_BYTES, &tmB, j, kb_w * GEMM_BK, n0b, (pairs >> (4 * pi_w)) & 3);
        kb_w += 2;
        while (kb_w >= kblocks) { kb_w -= kblocks; ++pi_w; }
      }
      pdl_wait();
      for (int j = 0; j < pre; ++j) {
        load(ring + j * S::STAGE_BYTES, &tmA, j, kb * GEMM_BK, m0, (pairs >> (4 * pi + 2)) & 3);
        kb += 2;
        while (kb >= kblocks) { kb -= kblocks; ++pi; }
      }
      if (h == 0 && k_resid) {
        if (PAIR) {
          mbar_expect_tx(res_bar, S::F32_TILE);
          for (int ch = 0; ch < BN / 32; ++ch)
            for (int q = 0; q < 4; ++q)
              tma_load_3d(res_tile + (ch * 4 + q) * GEMM_BOX, &tmR, res_bar, n0 + ch * 32, m0 + q * 32, 0);
        } else {
          mbar_expect_tx_elect(res_bar, S::F32_TILE);
          for (int ch = 0; ch < BN / 32; ++ch)
            for (int q = 0; q < 4; ++q)
              tma_load_3d_elect(res_tile + (ch * 4 + q) * GEMM_BOX, &tmR, res_bar, n0 + ch * 32, m0 + q * 32, 0);
        }
      }
      int st = 0;
      uint32_t ph = 0;                                   // parity of the ring pass being REFILLED
      for (int j = pre; j < count; ++j) {
        mbar_wait(&empty[st], ph);
        uint8_t* dst = ring + st * S::STAGE_BYTES;
        const uint32_t pr = pairs >> (4 * pi);
        expect(st);
        load(dst, &tmA, st, kb * GEMM_BK, m0, (pr >> 2) & 3);
        load(dst + S::A_BYTES, &tmB, st, kb * GEMM_BK, n0b, pr & 3);
        kb += 2;
        while (kb >= kblocks) { kb -= kblocks; ++pi; }
        if (++st == STAGES) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp < 4) {
    // ------------------------------------------------------------------ MMA issuer of pipeline h
    // !PAIR: the whole warp runs the loop converged and an elected lane issues (umma_*_elect) -- under a divergent
    // `lane == 0` every tcgen05 instruction sits in an R2UR / vote / branch loop.  PAIR: lane 0 of the leader CTA.
    if ((PAIR ? lane == 0 : true) && rank == 0) {
      const int h = warp - 2;
      const int count = (iters - h + 1) >> 1;
      uint64_t* full = bars + h * STAGES;
      uint64_t* empty = bars + (2 + h) * STAGES;
      const uint32_t idesc = make_idesc_bf16(PAIR ? 2 * GEMM_BM : GEMM_BM, BN, 0, 0, p.f16_in);
      const uint32_t desc_hi = smem_desc_hi_sw128(1024);
      const uint32_t a_lo0 = smem_desc_lo_sw128(smem_u32(smem + h * STAGES * S::STAGE_BYTES), 16);
      const uint32_t acc = tmem_acc + h * BN;
      int st = 0;
      uint32_t ph = 0;
      uint32_t a_lo = a_lo0;
      for (int j = 0; j < count; ++j) {
        mbar_wait(&full[st], ph);
        tc_fence_after();
        const uint32_t b_lo = a_lo + (S::A_BYTES >> 4);
#pragma unroll
        for (int k = 0; k < GEMM_BK / 16; ++k) {   // +32 bytes (2 x 16-byte units) per 16-wide k step
          if (PAIR) umma_2cta_lh(acc, a_lo + 2 * k, b_lo + 2 * k, desc_hi, idesc, (j > 0 || k > 0) ? 1u : 0u);
          else umma_bf16_lh_elect(acc, a_lo + 2 * k, b_lo + 2 * k, desc_hi, idesc, (j > 0 || k > 0) ? 1u : 0u);
        }
        if (j + STAGES < count) {                  // stage reusable (in both CTAs) once these MMAs have read it
          if (PAIR) umma_commit_2cta(&empty[st]); else umma_commit_elect(&empty[st]);
        }
        a_lo += S::STAGE_BYTES >> 4;
        if (++st == STAGES) { st = 0; ph ^= 1; a_lo = a_lo0; }
      }
      if (PAIR) umma_commit_2cta(&acc_bar[h]);     // (the host only pairs CTAs when both pipelines have work)
      else if (count > 0) umma_commit_elect(&acc_bar[h]);   // accumulator h complete
      else if (lane == 0) mbar_arrive(&acc_bar[h]);   // nothing to do for this pipeline (a single k-block)
      if (h == 0 && lane == 0) trace_mark(trace, 3);
    }
  }
  __syncwarp();   // lanes 1..31 of the role warps sleep here (no polling next to the issuing lane) until lane 0 is done

  {
    // ------------------------------------------------------------------ epilogue (all 8 warps)
    // warp w and warp w+4 share TMEM lane quadrant w%4; within each 64-column pair of chunks, warp w takes the
    // even 32-column chunk and warp w+4 the odd one.
    const int quad = warp & 3;                 // TMEM lane quadrant this warp may access
    const int hf = warp >> 2;
#ifdef AVD_DEBUG_HOOKS
    constexpr bool TRACE_CH = true;            // per-chunk phase stamps in every epilogue mode
    const int dbg = p.dbg;                     // timing experiments: 1 skip tcgen05.ld, 2 skip global stores, 4 skip staging
#else
    constexpr bool TRACE_CH = MODE == 0;
    constexpr int dbg = 0;
#endif
    const int row = m0 + quad * 32 + lane;
    const bool live = row < p.M;
    const bool two_acc = iters > 1;
    const int act = MODE == 0 ? p.act : (MODE == 2 ? 2 : 0);
    const int out_nseg = MODE == 0 ? p.out_nseg : 1;
    const bool f16_out = p.f16_out != 0;
    const float out_scale = p.out_scale;
    const bool has_resid = k_resid, has_f32 = MODE == 0 ? (p.out_f32 != nullptr) : (MODE == 3);
    const bool has_ssq = MODE == 0 ? (p.ssq_out != nullptr) : (MODE == 3);
    float* const ssq_out = p.ssq_out;
    float* ssq_x = bias_s + BN;                // [2][128] partial row sums handed from warp w+4 to warp w
    float* rinv_s = ssq_x + 256;               // [128] RMSNorm row scales, computed by warps 4..7 during the mainloop
    pdl_wait();
    float rinv = 1.0f;
    if (hf == 1) {
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int t = threadIdx.x - 128 + u * 128;
        if (t < BN) bias_s[t] = bias_r[u];
      }
      if (p.ssq_in != nullptr && live) {
        float ss = 0.0f;
        for (int q = 0; q < p.ssq_parts; ++q) ss += p.ssq_in[(int64_t)row * p.ssq_parts + q];
        rinv = 1.0f / (sqrtf(ss) * p.inv_sqrt_rms_dim + p.rms_eps);
      }
      rinv *= p.a_scale;
      rinv_s[quad * 32 + lane] = rinv;
      named_bar_sync(1, 256);                  // (not arrive: these warps read the bias tile the others wrote)
    } else {
      named_bar_sync(1, 256);
      rinv = rinv_s[quad * 32 + lane];
    }
    if (has_resid) mbar_wait(res_bar, 0);
    if (threadIdx.x == 128) trace_mark(trace, 4);
    mbar_wait(&acc_bar[0], 0);
    mbar_wait(&acc_bar[1], 0);
    tc_fence_after();
    if (threadIdx.x == 128) trace_mark(trace, 5);
    // staging: the fp32 result overwrites the residual tile in place, or (no residual) starts the reused pipeline smem
    uint8_t* f32_stage = has_resid ? res_tile : smem;
    uint8_t* bf_stage = smem + ((has_f32 && !has_resid) ? S::F32_TILE : 0);
    const int sw = lane & 7;
    const uint32_t t_row = tmem_acc + ((uint32_t)(quad * 32) << 16);
    uint32_t r[32], r1[32];
    if (hf < BN / 32 && !(dbg & 1)) {          // prefetch this warp's first chunk
      tmem_ld_32x32(t_row + hf * 32, r);
      if (two_acc) tmem_ld_32x32(t_row + BN + hf * 32, r1);
    }
#pragma unroll 1
    for (int ch = hf; ch < BN / 32; ch += 2) {
      float v[32];
      tmem_ld_wait();
      if (two_acc) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) + __uint_as_float(r1[j]);
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
      }
      if (ch + 2 < BN / 32 && !(dbg & 1)) {    // next chunk's TMEM loads fly while this one is processed
        tmem_ld_32x32(t_row + (ch + 2) * 32, r);
        if (two_acc) tmem_ld_32x32(t_row + BN + (ch + 2) * 32, r1);
      }
      if (TRACE_CH && threadIdx.x == 128 && ch == hf) trace_mark(trace, 8);
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        const float4 b4 = *reinterpret_cast<const float4*>(bias_s + ch * 32 + j);
        v[j] = fmaf(v[j], rinv, b4.x); v[j + 1] = fmaf(v[j + 1], rinv, b4.y);
        v[j + 2] = fmaf(v[j + 2], rinv, b4.z); v[j + 3] = fmaf(v[j + 3], rinv, b4.w);
      }
      uint8_t* fbox = f32_stage + (ch * 4 + quad) * GEMM_BOX + lane * 128;
      if (has_resid) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float4 x4 = lds128(fbox + ((c ^ sw) << 4));
          v[4 * c] += x4.x; v[4 * c + 1] += x4.y; v[4 * c + 2] += x4.z; v[4 * c + 3] += x4.w;
        }
      }
      const bool packed_gelu = (act == 2 && out_nseg == 1 && !has_f32 && !has_ssq);
      if (act == 1) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
      } else if (act == 2 && !packed_gelu) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = gelu_fast(v[j]);
      }
      if (TRACE_CH && threadIdx.x == 128 && ch == hf) trace_mark(trace, 9);
      if (has_f32) {
#pragma unroll
        for (int c = 0; c < 8; ++c) sts128(fbox + ((c ^ sw) << 4), v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
      }
      float ssq_acc = 0.0f;
      if (has_ssq) {
#pragma unroll
        for (int j = 0; j < 32; ++j) ssq_acc = fmaf(v[j], v[j], ssq_acc);
        if (hf == 1) ssq_x[((ch >> 1) & 1) * 128 + quad * 32 + lane] = ssq_acc;
      }
      const int c0 = (ch & 1) * 4;
      if (packed_gelu) {
        uint8_t* bbox = bf_stage + ((ch >> 1) * 4 + quad) * GEMM_BOX + lane * 128;
if (f16_out) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
          sts128u(bbox + (((c0 + c) ^ sw) << 4), gelu_pair_f16x2(v[8 * c], v[8 * c + 1]),
                  gelu_pair_f16x2(v[8 * c + 2], v[8 * c + 3]), gelu_pair_f16x2(v[8 * c + 4], v[8 * c + 5]),
                  gelu_pair_f16x2(v[8 * c + 6], v[8 * c + 7]));
          } else {
#pragma unroll
        for (int c = 0; c < 4; ++c)
          sts128u(bbox + (((c0 + c) ^ sw) << 4), gelu_pair_bf16x2(v[8 * c], v[8 * c + 1]),
                  gelu_pair_bf16x2(v[8 * c + 2], v[8 * c + 3]), gelu_pair_bf16x2(v[8 * c + 4], v[8 * c + 5]),
                  gelu_pair_bf16x2(v[8 * c + 6], v[8 * c + 7]));
          }
      } else if (out_nseg == 1) {
        uint8_t* bbox = bf_stage + ((ch >> 1) * 4 + quad) * GEMM_BOX + lane * 128;
if (f16_out) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
          sts128u(bbox + (((c0 + c) ^ sw) << 4), pack16(true, out_scale, v[8 * c], v[8 * c + 1]), pack16(true, out_scale, v[8 * c + 2], v[8 * c + 3]),
                  pack16(true, out_scale, v[8 * c + 4], v[8 * c + 5]), pack16(true, out_scale, v[8 * c + 6], v[8 * c + 7]));
          } else {
#pragma unroll
        for (int c = 0; c < 4; ++c)
          sts128u(bbox + (((c0 + c) ^ sw) << 4), pack_bf16x2(v[8 * c], v[8 * c + 1]), pack_bf16x2(v[8 * c + 2], v[8 * c + 3]),
                  pack_bf16x2(v[8 * c + 4], v[8 * c + 5]), pack_bf16x2(v[8 * c + 6], v[8 * c + 7]));
          }
      } else if (out_nseg > 1) {
        for (int sg = 0; sg < out_nseg; ++sg) {
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
#pragma unroll
          for (int c = 0; c < 4; ++c)
            sts128u(bbox + (((c0 + c) ^ sw) << 4), w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
        }
      }
      if (TRACE_CH && threadIdx.x == 128 && ch == hf) trace_mark(trace, 10);
      const bool paired = (out_nseg > 0 || has_ssq);   // the two warps of a quadrant share 64-column units
      if (paired) named_bar_sync(2 + quad, 64); else __syncwarp();
      if (TRACE_CH && threadIdx.x == 128 && ch == hf) trace_mark(trace, 11);
      if (hf == 0 && has_ssq && live)
        ssq_out[(int64_t)row * (p.N / 64) + ((n0 + ch * 32) >> 6)] = ssq_acc + ssq_x[((ch >> 1) & 1) * 128 + quad * 32 + lane];
      // staged boxes -> global, coalesced: 8 lanes cover one 128-byte row, one warp instruction covers 4 rows
      const int rr = lane >> 3, cc = lane & 7;
      if (has_f32 && !(dbg & 2)) {
        const uint8_t* box = f32_stage + (ch * 4 + quad) * GEMM_BOX;
        float* gp = p.out_f32 + (int64_t)(m0 + quad * 32) * p.N + n0 + ch * 32 + cc * 4;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int rw = it * 4 + rr;
          const float4 x4 = lds128(box + rw * 128 + ((cc ^ (rw & 7)) << 4));
          if (m0 + quad * 32 + rw < p.M) *reinterpret_cast<float4*>(gp + (int64_t)rw * p.N) = x4;
        }
      }
      if (out_nseg > 0 && !(dbg & 2)) {      // the pair's 64-column bf16 box: this warp stores rows [16 hf, 16 hf + 16)
        for (int sg = 0; sg < out_nseg; ++sg) {
          const uint8_t* box = bf_stage + ((sg * (BN / 64) + (ch >> 1)) * 4 + quad) * GEMM_BOX;
          __nv_bfloat16* gp = p.out_bf16 + (int64_t)sg * p.M * p.N + (int64_t)(m0 + quad * 32) * p.N + n0 + (ch >> 1) * 64 + cc * 8;
#pragma unroll
          for (int it = 0; it < 4; ++it) {
            const int rw = hf * 16 + it * 4 + rr;
            const float4 x4 = lds128(box + rw * 128 + ((cc ^ (rw & 7)) << 4));
            if (m0 + quad * 32 + rw < p.M) *reinterpret_cast<float4*>(gp + (int64_t)rw * p.N) = x4;
          }
        }
      }
      if (TRACE_CH && threadIdx.x == 128 && ch == hf) trace_mark(trace, 12);
    }
    if (threadIdx.x == 128) trace_mark(trace, 6);
  }

  tc_fence_before();
  if (PAIR) cluster_sync_relaxed(); else __syncthreads();   // neither CTA of a pair may retire while the other still drains
  if (warp == 2) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_2cta<TMEM_COLS>(tmem_acc); else tmem_dealloc<TMEM_COLS>(tmem_acc);
  }
  if (threadIdx.x == 0) trace_mark(trace, 7);
}

unsigned long long* g_dbg_trace = nullptr;
int64_t g_dbg_trace_slots = 0;

template <int BN, int STAGES, bool PAIR = false>
static int launch_gemm_bn(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmR, const GemmParams& p,
                          int mode, cudaStream_t s) {
  using S = GemmSmem<BN, STAGES, PAIR>;
  using KernelFn = void (*)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const GemmParams);
  static const KernelFn kernels[4] = {gemm_tc_kernel<BN, STAGES, 0, PAIR>, gemm_tc_kernel<BN, STAGES, 1, PAIR>,
                                      gemm_tc_kernel<BN, STAGES, 2, PAIR>, gemm_tc_kernel<BN, STAGES, 3, PAIR>};
  AVD_TRY(ensure_dyn_smem(reinterpret_cast<const void*>(kernels[mode]), S::total(BN <= 64)));   // once per device
  cudaLaunchConfig_t cfg{};
  const int mt = (p.M + GEMM_BM - 1) / GEMM_BM;
  // a pair = two consecutive row tiles (the last may be empty)
  cfg.gridDim = PAIR ? dim3((mt + 1) & ~1, p.N / BN) : dim3(p.N / BN, mt);
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = S::total(p.has_resid != 0);
  cfg.stream = s;
  cudaLaunchAttribute attrs[2];
  int na = 0;
  if (pdl_enabled()) {
    attrs[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (PAIR) {
    attrs[na].id = cudaLaunchAttributeClusterDimension;
    attrs[na].val.clusterDim.x = 2;
    attrs[na].val.clusterDim.y = 1;
    attrs[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = attrs;
  cfg.numAttrs = na;
  if (kind_skipped(KIND_GEMM)) return 0;
  LaunchScope ls(KIND_GEMM, s);
  AVD_CUDA(cudaLaunchKernelEx(&cfg, kernels[mode], tmA, tmB, tmR, p));
  return 0;
}

static int pipe_bytes_of(int bn) {
  switch (bn) {
    case 256: return GemmSmem<256, 2>::PIPE_BYTES;
    case 192: return GemmSmem<192, 2>::PIPE_BYTES;
    case 128: return GemmSmem<128, 3>::PIPE_BYTES;
    case 64: return GemmSmem<64, 3>::PIPE_BYTES;
    default: return GemmSmem<32, 3>::PIPE_BYTES;
  }
}

// Validation, kernel parameters, tile width and epilogue mode of one GEMM (shared by launch_gemm and the chained launch).
int plan_gemm(const avd_gemm_args& g, GemmPlan* plan) {
  const int nseg = nseg_of(g.precision);
  AVD_REQUIRE(nseg > 0, "precision must be 1 (bf16), 2 (f16), 3 or 6");
  AVD_REQUIRE(g.M > 0 && g.N > 0 && g.K > 0, "empty GEMM");
  AVD_REQUIRE(g.K % GEMM_BK == 0, "K must be a multiple of 64");
  AVD_REQUIRE(g.N % 32 == 0, "N must be a multiple of 32");
  AVD_REQUIRE(g.A != nullptr && g.W != nullptr, "null operand");
  AVD_REQUIRE(g.ssq_out == nullptr || g.N % 64 == 0, "ssq_out needs N % 64 == 0");
  AVD_REQUIRE(g.out_f32 != nullptr || g.out_bf16 != nullptr, "no output");
  AVD_REQUIRE(g.out_bf16 == nullptr || g.N % 64 == 0, "bf16 output needs N % 64 == 0");
  AVD_REQUIRE(g.resid == nullptr || g.out_f32 != nullptr, "a residual needs the fp32 output");
  AVD_REQUIRE(g.act >= 0 && g.act <= 2, "act must be 0 (none), 1 (GELU erf) or 2 (GELU fast erf)");
  const int out_nseg = g.out_bf16 != nullptr ? nseg_of(g.out_precision) : 0;
  AVD_REQUIRE(g.out_bf16 == nullptr || out_nseg > 0, "out_precision must be 1, 2, 3 or 6");

  GemmParams p{};
  p.M = g.M; p.N = g.N; p.K = g.K;
  // smallest partial products first so they are not absorbed by the large ones
  if (g.precision == 1 || g.precision == 2) { p.n_pairs = 1; p.pairs = 0x0; }
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
  p.out_f32 = g.out_f32;
  p.out_bf16 = static_cast<__nv_bfloat16*>(g.out_bf16);
  p.out_nseg = out_nseg;
  p.ssq_out = g.ssq_out;
  p.f16_in = is_f16(g.precision) ? 1 : 0;
  p.f16_out = (g.out_bf16 != nullptr && is_f16(g.out_precision)) ? 1 : 0;
  p.a_scale = g.a_scale != 0.0f ? g.a_scale : 1.0f;
  p.out_scale = g.out_scale != 0.0f ? g.out_scale : 1.0f;
  // Tile width: the widest BN that divides N, still gives >= ~0.8 waves of CTAs on 148 SMs, and whose epilogue staging
  // (fp32 tile unless it overwrites the residual tile, plus out_nseg bf16 tiles) fits the reused pipeline smem.
  // A residual tile needs dedicated smem, which only the narrow tiles can spare.
  const int mt = (g.M + GEMM_BM - 1) / GEMM_BM;
  auto staging = [&](int bn) {
    return ((g.out_f32 != nullptr && g.resid == nullptr) ? GEMM_BM * bn * 4 : 0) + out_nseg * GEMM_BM * bn * 2;
  };
  auto usable = [&](int bn) {
    if (g.N % bn != 0) return false;
    if (g.resid != nullptr && bn > 64) return false;
    if ((g.ssq_out != nullptr || g.out_bf16 != nullptr) && bn < 64) return false;
    return staging(bn) <= pipe_bytes_of(bn);
  };
  const int cands[5] = {256, 192, 128, 64, 32};
  int bn = 0;
  const int force_bn = options().gemm_bn;
  if (force_bn > 0 && usable(force_bn)) bn = force_bn;
  for (int i = 0; i < 5 && bn == 0; ++i)
    if (usable(cands[i]) && (int64_t)mt * (g.N / cands[i]) >= 120) bn = cands[i];
  for (int i = 3; i < 5 && bn == 0; ++i)            // small problems: narrow tiles, more CTAs
    if (usable(cands[i])) bn = cands[i];
  for (int i = 0; i < 5 && bn == 0; ++i)
    if (usable(cands[i])) bn = cands[i];
  // Two kernels.  While the tiles chosen above fit one wave of CTAs the one-tile-per-CTA kernel of this file wins (wide
  // tiles, every warp on the epilogue); beyond one wave the persistent kernel (avd_gemm_ps.cu) hides the epilogue under
  // the next tile's mainloop (+45 % at M = 32768, tools/kbench.py --M).  AVD_GEMM_V = 3 / 4 forces one of them.
  const int gemm_v = options().gemm_v;
  plan->p = p;
  plan->nseg = nseg;
  plan->force_bn = force_bn;
  plan->bn = bn;
  plan->mode = 0;
  {
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    plan->persistent = gemm_v == 4 || (gemm_v != 3 && (bn == 0 || (int64_t)mt * (g.N / bn) > sms));
    if (plan->persistent) return 0;
  }
  AVD_REQUIRE(bn > 0, "no tile shape fits this combination of outputs: use fewer outputs per GEMM");
  // epilogue specialisation (see the kernel's MODE); every other combination uses the generic one
  const bool one16 = g.out_bf16 != nullptr && out_nseg == 1;
  if (one16 && g.out_f32 == nullptr && g.resid == nullptr && g.ssq_out == nullptr && g.act == 0) plan->mode = 1;
  else if (one16 && g.out_f32 == nullptr && g.resid == nullptr && g.ssq_out == nullptr && g.act == 2) plan->mode = 2;
  else if (one16 && g.out_f32 != nullptr && g.resid != nullptr && g.ssq_out != nullptr && g.act == 0) plan->mode = 3;
  if (bn < 64 || options().gemm_generic) plan->mode = 0;
  return 0;
}

int launch_gemm(const avd_gemm_args& g, cudaStream_t s) {
  GemmPlan plan;
  AVD_TRY(plan_gemm(g, &plan));
  GemmParams p = plan.p;
  const int nseg = plan.nseg, bn = plan.bn;
  const int mt = (g.M + GEMM_BM - 1) / GEMM_BM;
  if (plan.persistent) return launch_gemm_ps(g, p, nseg, plan.force_bn, s);

  const int64_t n_ctas = (int64_t)(g.N / bn) * mt;
  p.trace = (g_dbg_trace != nullptr && n_ctas * 32 <= g_dbg_trace_slots) ? g_dbg_trace : nullptr;

  CUtensorMap tmA, tmB, tmR;
  AVD_TRY(get_tensor_map_3d(g.A, 2, g.K, g.M, nseg, (uint64_t)g.K * 2, (uint64_t)g.M * g.K * 2, GEMM_BK, GEMM_BM, &tmA));
  AVD_TRY(get_tensor_map_3d(g.W, 2, g.K, g.N, nseg, (uint64_t)g.K * 2, (uint64_t)g.N * g.K * 2, GEMM_BK, bn, &tmB));
  CUtensorMap tmBh = tmB;   // box of bn/2 weight rows: what one CTA of a pair stages
  if (bn >= 64)
    AVD_TRY(get_tensor_map_3d(g.W, 2, g.K, g.N, nseg, (uint64_t)g.K * 2, (uint64_t)g.N * g.K * 2, GEMM_BK, bn / 2, &tmBh));
  tmR = tmA;   // placeholder when unused (never dereferenced)
  if (g.resid != nullptr)
    AVD_TRY(get_tensor_map_3d(g.resid, 4, g.N, g.M, 1, (uint64_t)g.N * 4, (uint64_t)g.M * g.N * 4, 32, 32, &tmR));
  // the phase trace uses the generic epilogue (except in debug-hooks builds, which stamp every mode)
  int mode = plan.mode;
#ifdef AVD_DEBUG_HOOKS
  p.dbg = options().gemm_dbg;
#else
  if (p.trace != nullptr) mode = 0;
#endif
  // CTA pairs (cta_group::2): each SM stages half of the weight tile.  Needs both pipelines busy in the leader, a
  // specialised epilogue, and the padded grid (an even number of row tiles) still within one wave.
  const int pair_on = options().gemm_pair;
  int sms_now = 148;
  {
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms_now, cudaDevAttrMultiProcessorCount, dev);
  }
  const bool pair = pair_on != 0 && (mode != 0 || pair_on == 2) && mt >= 2 && (g.K / GEMM_BK) * p.n_pairs >= 2 &&
                    (int64_t)((mt + 1) & ~1) * (g.N / bn) <= sms_now;
  if (pair) {
    switch (bn) {
      case 256: return launch_gemm_bn<256, 3, true>(tmA, tmBh, tmR, p, mode, s);
      case 192: return launch_gemm_bn<192, 3, true>(tmA, tmBh, tmR, p, mode, s);
      case 128: return launch_gemm_bn<128, 4, true>(tmA, tmBh, tmR, p, mode, s);
      case 64: return launch_gemm_bn<64, 4, true>(tmA, tmBh, tmR, p, mode, s);
      default: break;
    }
  }
  switch (bn) {
    case 256: return launch_gemm_bn<256, 2>(tmA, tmB, tmR, p, mode, s);
    case 192: return launch_gemm_bn<192, 2>(tmA, tmB, tmR, p, mode, s);
    case 128: return launch_gemm_bn<128, 3>(tmA, tmB, tmR, p, mode, s);
    case 64: return launch_gemm_bn<64, 3>(tmA, tmB, tmR, p, mode, s);
    default: return launch_gemm_bn<32, 3>(tmA, tmB, tmR, p, mode, s);
  }
}

}  // namespace avd
