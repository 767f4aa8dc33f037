// avd_attn_fa.cu -- K5, multi-block case (N > 192 keys): attention with P and O resident in TENSOR MEMORY.
//
// What the per-warp timeline of attn_tc_kernel's multi-block path showed (tools/attn_wtrace.py, 16 x 2048 keys): a key
// block of 2 x 128 x 192 scores costs 6900 cycles against 3072 of MUFU; the 12 P V MMAs (N = 64) take ~96 cycles each
// instead of 32 because both operands come through the shared-memory port (P was staged there with 48 KB of STS per tile
// and block, next to the S MMAs' operands and the TMA writes), the tile's issuer lane sits in a softmax warp and makes it
// late, and every block reads O back (tcgen05.ld + 32 FMA per thread) after waiting for P V.  Here (the split
// FlashAttention-4 describes, sized for head_dim 64):
//
//   * P never touches shared memory: the softmax threads write it with tcgen05.st into TMEM columns next to S, and
//     O += P V is a tcgen05.mma with the A operand IN TMEM (32 cycles per K = 16 step, only V's 2 KB from smem);
//   * O stays in TMEM across the key blocks (the MMA accumulates).  A row's reference max only moves when the block's max
//     exceeds it by more than 2^8; then -- rarely after the first blocks -- the two warps that own the row rescale their
//     32 + 32 O columns in place (tcgen05.ld / st).  No per-block O read-out, no wait for P V in the softmax loop;
//   * 128-key blocks, ONE pass over S per block: the exponentials are taken against the row's carried reference max in
//     the same pass that finds the block max (no pair exchange between the tcgen05.ld and the MUFU work); the pass is
//     repeated only when the reference has to move;
//   * one control WARP (warp 16, converged; an elected lane issues -- umma_*_elect) issues every TMA load (4-stage K / V
//     ring, two blocks ahead) and every MMA, in the fixed order  S_0(j+1), P V_0(j), S_1(j+1), P V_1(j): S(j+1) goes as soon
//     as S(j) has been read (its own barrier), P V(j) once P(j) is stored; while one tile's softmax runs, the tensor pipe
//     works for the other.
//
// TMEM per query tile t (256 columns): S [0, 128) fp32 | P [128, 192) 16-bit pairs | O [192, 256) fp32.
// One CTA = one (sequence, head, PAIR of 128-query tiles); warps 0..7 softmax of tile 0, 8..15 of tile 1 (warp w: TMEM lanes
// 32 (w % 4).., the two warps of a lane quadrant split the key columns and the output columns in halves).
#include "avd_attn_common.cuh"

#include <type_traits>

namespace avd {

constexpr int FA_BKV = 128;
constexpr int FA_STAGES = 4;
constexpr int FA_KV_BYTES = FA_BKV * ATT_HD * 2;            // 16 KB: [128 keys][64] 16-bit, 128B swizzle
constexpr int FA_THREADS = 17 * 32;                          // SW = 8: two softmax warps per TMEM lane quadrant and tile
constexpr int FA_THREADS_ROW = 9 * 32;                       // SW = 4: one softmax thread per query row
constexpr int FA_XCHG_BYTES = (8 + 4) * 128 * 4;            // row max [2 bufs][2 tiles][2 halves][128], row sum [2][2][128]
constexpr int FA_SMEM = 2 * ATT_Q_BYTES + FA_STAGES * 2 * FA_KV_BYTES + FA_XCHG_BYTES + 256;
// A row's reference max moves when the block max is 2^FA_RESCALE_LOG2 above it; until then P = exp2(s - ref) may reach
// 2^8 (fp16 keeps its full mantissa there; the row sum and O are fp32).  2^12 was measured too: no difference (235 vs 232 us),
// a repeated pass is rare enough at 2^8.
#ifndef AVD_FA_RESCALE_LOG2
#define AVD_FA_RESCALE_LOG2 8
#endif
constexpr float FA_RESCALE_LOG2 = (float)AVD_FA_RESCALE_LOG2;
enum { FB_Q = 0, FB_KF = 1 /*+stage*/, FB_VF = 5 /*+stage*/, FB_KVE = 9 /*+stage*/, FB_S = 13 /*+tile*/, FB_P = 15 /*+tile*/,
       FB_O = 17 /*+tile*/, FB_R = 19 /*+tile*/, FB_COUNT = 21 };

template <bool F16, bool MASKED, int SW>
__global__ void __launch_bounds__((2 * SW + 1) * 32, 1) attn_fa_kernel(const __grid_constant__ CUtensorMap tmQKV,
                                                                const AttnParams p) {
  extern __shared__ __align__(1024) uint8_t fa_smem[];
  uint8_t* smem = fa_smem;
  if ((smem_u32(smem) & 1023u) != 0) __trap();   // 128B-swizzle tiles need 1024-B aligned bases
  uint8_t* sQ = smem;                                       // 2 query tiles (later: output staging)
  uint8_t* sKV = sQ + 2 * ATT_Q_BYTES;                      // [stage]{K, V}
  float* xmax = reinterpret_cast<float*>(sKV + FA_STAGES * 2 * FA_KV_BYTES);
  float* xsum = xmax + 8 * 128;
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(xmax) + FA_XCHG_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + FB_COUNT);

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int head = blockIdx.y, seq = blockIdx.z;
  const int row_base = seq * p.N;
  const int q_first = blockIdx.x * 2 * ATT_BQ;
  const int n_tiles = (p.N - q_first > ATT_BQ) ? 2 : 1;
  const int colQ = head * ATT_HD, colK = p.d + head * ATT_HD, colV = 2 * p.d + head * ATT_HD;
  const int nkv = (p.N + FA_BKV - 1) / FA_BKV;
  unsigned long long* const trace = p.trace;   // per-warp stamps of -DAVD_ATT_WTRACE builds (ATT_WMARK), else unused
  (void)trace;

  if (warp == 1) {
    if (lane < FB_COUNT) {
      uint32_t count = 1;
      if ((lane >= FB_P && lane < FB_O) || lane >= FB_R) {  // arrivals: the warps of the tile that own valid query rows
        const int rows = min(ATT_BQ, p.N - q_first - ((lane - FB_P) & 1) * ATT_BQ);
        count = rows > 0 ? (SW * 8) * ((rows + 31) >> 5) : 1;
      }
      mbar_init(&bars[lane], count);
    }
    fence_barrier_init();
  } else if (warp == 2 && lane == 0) {
    tma_prefetch_desc(&tmQKV);
  }
  if (warp == 0) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);   // warp-uniform for the compiler too
  pdl_launch_dependents();

  if (warp == 2 * SW) {
    // ------------------------------------------------------------------ control warp: TMA loads and every MMA
    // The whole warp runs this code converged (all lanes wait on the barriers); one elected lane issues the tcgen05
    // instructions and the K / V TMA loads (*_elect); lane 0 only the one-off Q load.
    auto load_kv = [&](int b) {
      const int s = b & (FA_STAGES - 1);
      uint8_t* k = sKV + s * 2 * FA_KV_BYTES;
      mbar_expect_tx_elect(&bars[FB_KF + s], FA_KV_BYTES);
      tma_load_3d_elect(k, &tmQKV, &bars[FB_KF + s], colK, row_base + b * FA_BKV, 0);
      mbar_expect_tx_elect(&bars[FB_VF + s], FA_KV_BYTES);
      tma_load_3d_elect(k + FA_KV_BYTES, &tmQKV, &bars[FB_VF + s], colV, row_base + b * FA_BKV, 0);
    };
    if (lane == 0) {
      pdl_wait();                                           // qkv comes from the previous kernel
      mbar_expect_tx(&bars[FB_Q], n_tiles * ATT_Q_BYTES);
      for (int t = 0; t < n_tiles; ++t)
        tma_load_3d(sQ + t * ATT_Q_BYTES, &tmQKV, &bars[FB_Q], colQ, row_base + q_first + t * ATT_BQ, 0);
    }
    __syncwarp();
    for (int b = 0; b < min(2, nkv); ++b) load_kv(b);

    const uint32_t hi = smem_desc_hi_sw128(1024);
    const uint32_t idesc_o = make_idesc_bf16(ATT_BQ, ATT_HD, 0, 1, F16);         // B = V is MN-major
    const uint32_t q_lo0 = smem_desc_lo_sw128(smem_u32(sQ), 16);
    const uint32_t kv_lo0 = smem_desc_lo_sw128(smem_u32(sKV), 16);
    auto issue_s = [&](int t, int b) {                      // S_t(b) = Q_t K_b^T -> TMEM columns [256 t, 256 t + 128)
      const int kvn = (min(FA_BKV, p.N - b * FA_BKV) + 15) & ~15;
      const uint32_t idesc_s = make_idesc_bf16(ATT_BQ, kvn, 0, 0, F16);
      const uint32_t q_lo = q_lo0 + t * (ATT_Q_BYTES >> 4);
      const uint32_t k_lo = kv_lo0 + (b & (FA_STAGES - 1)) * (2 * FA_KV_BYTES >> 4);
#pragma unroll
      for (int k = 0; k < ATT_HD / 16; ++k)
        umma_bf16_lh_elect(tmem_base + t * 256, q_lo + 2 * k, k_lo + 2 * k, hi, idesc_s, k > 0 ? 1u : 0u);
      umma_commit_elect(&bars[FB_S + t]);
    };
    mbar_wait(&bars[FB_Q], 0);
    mbar_wait(&bars[FB_KF + 0], 0);
    tc_fence_after();
    for (int t = 0; t < n_tiles; ++t) issue_s(t, 0);

    for (int j = 0; j < nkv; ++j) {
      const int s = j & (FA_STAGES - 1);
      if (j + 2 < nkv) {                                    // the stage of block j - 2 was handed back two blocks ago
        const int b = j + 2;
        if (b >= FA_STAGES) mbar_wait(&bars[FB_KVE + (b & (FA_STAGES - 1))], ((b / FA_STAGES) - 1) & 1);
        load_kv(b);
      }
      const int kvn = (min(FA_BKV, p.N - j * FA_BKV) + 15) & ~15;
      const uint32_t v_lo = kv_lo0 + ((s * 2 * FA_KV_BYTES + FA_KV_BYTES) >> 4);
      for (int t = 0; t < n_tiles; ++t) {
        if (j + 1 < nkv) {                                  // S_t(j+1) first: the tile's next softmax pass waits for it,
          mbar_wait(&bars[FB_R + t], j & 1);                // and it only needs S_t(j) READ (P stores may still be in flight)
          if (t == 0) mbar_wait(&bars[FB_KF + ((j + 1) & (FA_STAGES - 1))], ((j + 1) / FA_STAGES) & 1);
          tc_fence_after();
          ATT_WMARK(t * 3);
          issue_s(t, j + 1);
          ATT_WMARK(t * 3 + 1);
        }
        mbar_wait(&bars[FB_P + t], j & 1);                  // P_t(j) is in TMEM (and O_t rescaled where a row's max moved)
        if (t == 0) mbar_wait(&bars[FB_VF + s], (j / FA_STAGES) & 1);
        tc_fence_after();
        const uint32_t d_o = tmem_base + t * 256 + 192, a_p = tmem_base + t * 256 + 128;
        if (kvn == FA_BKV) {
#pragma unroll
          for (int ks = 0; ks < FA_BKV / 16; ++ks)          // O_t += P_t(j) V_j: A from TMEM, 8 columns per 16 keys
            umma_bf16_ts_elect(d_o, a_p + ks * 8, v_lo + ks * 128, hi, idesc_o, (j > 0 || ks > 0) ? 1u : 0u);
        } else {
          for (int ks = 0; ks < (kvn >> 4); ++ks)
            umma_bf16_ts_elect(d_o, a_p + ks * 8, v_lo + ks * 128, hi, idesc_o, (j > 0 || ks > 0) ? 1u : 0u);
        }
        umma_commit_elect(&bars[FB_O + t]);                 // P_t(j) has been read, O_t is current (phase j)
        if (t == n_tiles - 1) umma_commit_elect(&bars[FB_KVE + s]);   // K_j / V_j have been read by every tile
        ATT_WMARK(t * 3 + 2);
      }
    }
  } else if (SW == 4 && (warp >> 2) < n_tiles && q_first + (warp >> 2) * ATT_BQ + (warp & 3) * 32 < p.N) {
    // ------------------------------------------------------------------ softmax + output of query tile t, one THREAD per row
    // 128 scores of a block live in this thread's registers: the row max needs no exchange, the exponentials run as one
    // interleaved stream (MUFU c, sum / pack of c - 8), and with two warps per scheduler the two tiles alternate between
    // their MUFU-bound pass and the wait for their MMAs.
    const int t = warp >> 2, quad = warp & 3;
    const uint32_t tmem_T = tmem_base + t * 256 + ((uint32_t)(quad * 32) << 16);
    const float sc = p.scale_log2e;
    float m_run = -INFINITY, l_run = 0.0f;
    for (int j = 0; j < nkv; ++j) {
      const uint32_t ph = j & 1;
      const int kv_valid = min(FA_BKV, p.N - j * FA_BKV);
      ATT_WMARK(0);
      mbar_wait(&bars[FB_S + t], ph);
      tc_fence_after();
      ATT_WMARK(1);
      bool pv_seen = (j == 0);
      auto pv_done = [&]() {                                // P V_t(j-1) has read P and updated O (see the SW = 8 branch)
        if (!pv_seen) {
          mbar_wait(&bars[FB_O + t], ph ^ 1);
          tc_fence_after();
          pv_seen = true;
        }
      };
      // one pass over the row's (up to) four 32-score chunks, the next chunk's tcgen05.ld in flight under the current one's
      // arithmetic: block max over the valid keys and, with EXP, the exponentials against `neg_m_sc` as ONE interleaved
      // stream (MUFU c, max of score c, sum / pack of exponential c - 8) -> P columns, partial row sum
      auto pass = [&](auto exp_tag, float neg_m_sc, float& l_blk) -> float {
        constexpr bool EXP = decltype(exp_tag)::value;
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        float l0 = 0.0f, l1 = 0.0f;
        uint32_t x0[32], x1[32], x2[32];                     // chunk i lives in x{i % 3}: exponentials / raw scores / in flight
        auto mask = [&](int i, uint32_t (&x)[32]) {          // padded / out-of-sequence keys: score -inf, probability 0
          const int lim = kv_valid - i * 32;                 // (<= 0: a chunk past the keys of a ragged last block -- stale TMEM)
          uint32_t vb = lim >= 32 ? 0xffffffffu : (lim > 0 ? ((1u << lim) - 1u) : 0u);
          if (MASKED && lim > 0) vb &= ~p.kmask[(size_t)blockIdx.z * p.kmask_words + j * (FA_BKV / 32) + i];
          if (vb != 0xffffffffu) {
#pragma unroll
            for (int c = 0; c < 32; ++c)
              if (!((vb >> c) & 1u)) x[c] = 0xff800000u;
          }
        };
        // max of raw score c and its exponential (in place)
        auto expo = [&](uint32_t (&x)[32], int c) {
          mx[c & 3] = fmaxf(mx[c & 3], __uint_as_float(x[c]));
          if (EXP) {
            const float a = fmaf(__uint_as_float(x[c]), sc, neg_m_sc);
            float e;
            asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(a));
            x[c] = __float_as_uint(e);
          }
        };
        // row sum and 16-bit pack of exponentials c, c + 1 of a chunk whose MUFU work finished a chunk ago
        auto post = [&](const uint32_t (&e)[32], int c, uint32_t (&w)[16]) {
          l0 += __uint_as_float(e[c]);
          l1 += __uint_as_float(e[c + 1]);
          w[c >> 1] = F16 ? pack_f16x2(__uint_as_float(e[c]), __uint_as_float(e[c + 1]))
                          : pack_bf16x2(__uint_as_float(e[c]), __uint_as_float(e[c + 1]));
        };
        // chunk i (exponentials in `cur`): its sums / packs run between the exponentials of chunk i + 1 (raw scores landing
        // in `nxt`) while chunk i + 2 is fetched into `pre` -- nothing in the stream waits for a MUFU result less than a
        // chunk old
        // All four chunks are always processed (straight-line code keeps the three buffers in registers): in a ragged last
        // block the chunks past the keys read stale TMEM, are masked to -inf and store zeros that P V (K = round16(keys)) never reads.
        auto step = [&](int i, uint32_t (&cur)[32], uint32_t (&nxt)[32], uint32_t (&pre)[32]) {
          uint32_t w[16];
          if (i + 1 < 4) {
            tmem_ld_wait();                                  // chunk i + 1 has landed
            if (i + 2 < 4) tmem_ld_32x32(tmem_T + (i + 2) * 32, pre);
            mask(i + 1, nxt);
#pragma unroll
            for (int c = 0; c < 32; c += 2) {
              expo(nxt, c);
              expo(nxt, c + 1);
              if (EXP) post(cur, c, w);
            }
          } else if (EXP) {
#pragma unroll
            for (int c = 0; c < 32; c += 2) post(cur, c, w);
          }
          if (EXP) {
            if (i == 0) pv_done();
            tmem_st_32x16(tmem_T + 128 + i * 16, w);
          }
        };
        tmem_ld_32x32(tmem_T, x0);
        tmem_ld_wait();
        tmem_ld_32x32(tmem_T + 32, x1);
        mask(0, x0);
#pragma unroll
        for (int c = 0; c < 32; ++c) expo(x0, c);
        step(0, x0, x1, x2);
        step(1, x1, x2, x0);
        step(2, x2, x0, x1);
        step(3, x0, x1, x2);
        l_blk = l0 + l1;
        return fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      };
      // speculate that the reference max stays (exponentials in the same pass that finds the block max); a warp with a
      // row that has no reference yet finds the max first
      const bool spec = !__any_sync(0xffffffffu, m_run == -INFINITY);
      float l_blk = 0.0f;
      const float m_row = spec ? pass(std::true_type{}, -m_run * sc, l_blk) : pass(std::false_type{}, 0.0f, l_blk);
      ATT_WMARK(2);
      // the reference max moves only when it must: first live block of the row, or the block max is 2^FA_RESCALE_LOG2 above it
      const bool need = m_row != -INFINITY && (m_run == -INFINITY || (m_row - m_run) * sc > FA_RESCALE_LOG2);
      bool have_p = spec;
      if (__any_sync(0xffffffffu, need)) {
        float alpha = 1.0f;
        if (need) {
          alpha = ex2_approx((m_run - m_row) * sc);
          m_run = m_row;
          l_run *= alpha;
        }
        if (j > 0) {
          pv_done();
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            uint32_t o[32];
            tmem_ld_32x32(tmem_T + 192 + h * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int c = 0; c < 32; ++c) o[c] = __float_as_uint(__uint_as_float(o[c]) * alpha);
            tmem_st_32x32(tmem_T + 192 + h * 32, o);
          }
        }
        have_p = false;
      }
      ATT_WMARK(3);
      if (!have_p) {
        const float neg_m_sc = m_run == -INFINITY ? 0.0f : -m_run * sc;
        (void)pass(std::true_type{}, neg_m_sc, l_blk);
      }
      tc_fence_before();
      mbar_arrive(&bars[FB_R + t]);                         // S_t(j) will not be read again: S_t(j+1) may overwrite it
      l_run += l_blk;
      ATT_WMARK(4);
      tmem_st_wait();
      tc_fence_before();    // S has been read, P (and a rescaled O) written: the tile's next MMAs may go
      mbar_arrive(&bars[FB_P + t]);
      ATT_WMARK(5);
    }

    // ---- O_t complete: normalise this thread's row, stage the warp's 32 rows x 128 B in the retired Q tile, coalesced stores
    mbar_wait(&bars[FB_O + t], (nkv - 1) & 1);
    tc_fence_after();
    float rem[64];
    {
      const float inv_l = 1.0f / l_run;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t o[32];
        tmem_ld_32x32(tmem_T + 192 + h * 32, o);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 32; ++c) rem[h * 32 + c] = __uint_as_float(o[c]) * inv_l;
      }
    }
    const int64_t seg_stride = p.M * (int64_t)p.d;
    uint8_t* stage = sQ + t * ATT_Q_BYTES + quad * 32 * 128;
    const int q_tile0 = q_first + t * ATT_BQ + quad * 32;          // first query row of this warp
    for (int sg = 0; sg < p.out_nseg; ++sg) {
#pragma unroll
      for (int v8 = 0; v8 < 8; ++v8) {
        uint32_t w[4];
        if (p.out_nseg == 1) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            w[k] = F16 ? pack_f16x2(rem[v8 * 8 + k * 2], rem[v8 * 8 + k * 2 + 1])
                       : pack_bf16x2(rem[v8 * 8 + k * 2], rem[v8 * 8 + k * 2 + 1]);
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int c = v8 * 8 + k * 2;
            const __nv_bfloat16 h0 = __float2bfloat16_rn(rem[c]);
            const __nv_bfloat16 h1 = __float2bfloat16_rn(rem[c + 1]);
            rem[c] -= __bfloat162float(h0);
            rem[c + 1] -= __bfloat162float(h1);
            w[k] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
          }
        }
        *reinterpret_cast<uint4*>(stage + lane * 128 + ((v8 ^ (lane & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
      }
      __syncwarp();
      const int rr = lane >> 3, cc = lane & 7;
      __nv_bfloat16* gp = p.out + sg * seg_stride + ((int64_t)row_base + q_tile0) * p.d + head * ATT_HD + cc * 8;
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const int rw = it * 4 + rr;
        const uint4 x4 = *reinterpret_cast<const uint4*>(stage + rw * 128 + ((cc ^ (rw & 7)) << 4));
        if (q_tile0 + rw < p.N) *reinterpret_cast<uint4*>(gp + (int64_t)rw * p.d) = x4;
      }
      if (sg + 1 < p.out_nseg) __syncwarp();
    }
  } else if (SW == 8 && (warp >> 3) < n_tiles && q_first + (warp >> 3) * ATT_BQ + (warp & 3) * 32 < p.N) {
    // ------------------------------------------------------------------ softmax + output of query tile t, two warps per row group
    const int t = warp >> 3, sp = (warp >> 2) & 1, quad = warp & 3;
    const int r = quad * 32 + lane;                         // row inside the tile == TMEM lane
    const uint32_t tmem_T = tmem_base + t * 256 + ((uint32_t)(quad * 32) << 16);
    const uint32_t tmem_S = tmem_T + sp * 64;               // this thread's 64 score columns
    const uint32_t tmem_P = tmem_T + 128 + sp * 32;         // ... their 32 columns of 16-bit pairs
    const uint32_t tmem_O = tmem_T + 192 + sp * 32;         // this thread's 32 of the 64 output columns
    const uint32_t pair_bar = 1 + t * 4 + quad;             // named barrier of the two warps that share these rows
    const float sc = p.scale_log2e;
    float m_run = -INFINITY, l_run = 0.0f;
    auto valid_bits = [&](int j, int ch, int lim) -> uint32_t {
      uint32_t v = lim >= 32 ? 0xffffffffu : (lim > 0 ? ((1u << lim) - 1u) : 0u);
      if (MASKED && lim > 0) v &= ~p.kmask[(size_t)blockIdx.z * p.kmask_words + j * (FA_BKV / 32) + ch];
      return v;
    };
    for (int j = 0; j < nkv; ++j) {
      const uint32_t ph = j & 1;
      const int kv_valid = min(FA_BKV, p.N - j * FA_BKV);
      float* xm = xmax + ((ph * 2 + t) * 2) * 128;
      // one pass over this thread's (up to) two 32-score chunks: block max over the valid keys and, with EXP, the
      // exponentials against the reference max `neg_m_sc` -> P columns (tcgen05.st) and the partial row sum
      // P V_t(j-1) reads the P columns and updates O_t behind S_t(j) on the tensor pipe: before this block's first P store
      // (or O rescale) it must have completed -- by then it normally has
      bool pv_seen = (j == 0);
      auto pv_done = [&]() {
        if (!pv_seen) {
          mbar_wait(&bars[FB_O + t], ph ^ 1);
          tc_fence_after();
          pv_seen = true;
        }
      };
      auto pass = [&](auto exp_tag, float neg_m_sc, float& l_blk) -> float {
        constexpr bool EXP = decltype(exp_tag)::value;
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        l_blk = 0.0f;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int lim = kv_valid - (sp * 2 + i) * 32;
          if (lim > 0) {
            uint32_t x[32];
            tmem_ld_32x32(tmem_S + i * 32, x);
            tmem_ld_wait();
            if (EXP) ATT_WMARK(6 + 3 * i);
            const uint32_t vb = valid_bits(j, sp * 2 + i, lim);
            if (vb != 0xffffffffu) {                        // padded / out-of-sequence keys: score -inf, probability 0
#pragma unroll
              for (int c = 0; c < 32; ++c)
                if (!((vb >> c) & 1u)) x[c] = 0xff800000u;
            }
            float e[32];
            if (EXP) {
              ex2_chunk(x, sc, neg_m_sc, e);
              ATT_WMARK_DEP(7 + 3 * i, e[31]);
            }
#pragma unroll
            for (int c = 0; c < 32; ++c) mx[c & 3] = fmaxf(mx[c & 3], __uint_as_float(x[c]));
            if (EXP) {
              float l0 = 0.0f, l1 = 0.0f;
              uint32_t w[16];
#pragma unroll
              for (int c = 0; c < 32; c += 2) {
                l0 += e[c];
                l1 += e[c + 1];
                w[c >> 1] = F16 ? pack_f16x2(e[c], e[c + 1]) : pack_bf16x2(e[c], e[c + 1]);
              }
              l_blk += l0 + l1;
              pv_done();
              tmem_st_32x16(tmem_P + i * 16, w);
              ATT_WMARK_DEP(8 + 3 * i, l_blk);
            }
          }
        }
        return fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      };
      ATT_WMARK(0);
      mbar_wait(&bars[FB_S + t], ph);
      tc_fence_after();
      ATT_WMARK(1);
      // Speculate that the reference max stays: exponentials against the carried m_run in the same pass that finds the
      // block max (no pair exchange between the TMEM load and the MUFU work).  A warp with a row that has no reference yet
      // (first block, or only padded keys so far) finds the max first.  Both warps of a pair hold the same rows and the same
      // m_run, so they take the same branches.
      const bool spec = !__any_sync(0xffffffffu, m_run == -INFINITY);
      float l_blk = 0.0f, m_part;
      if (spec) m_part = pass(std::true_type{}, -m_run * sc, l_blk);
      else m_part = pass(std::false_type{}, 0.0f, l_blk);
      ATT_WMARK(2);
      xm[sp * 128 + r] = m_part;
      named_bar_sync(pair_bar, 64);
      ATT_WMARK(3);
      const float m_row = fmaxf(m_part, xm[(sp ^ 1) * 128 + r]);
      // the reference max moves only when it must: first live block of the row, or the block max is 2^FA_RESCALE_LOG2 above it (P stays
      // inside fp16, the row sum well conditioned)
      const bool need = m_row != -INFINITY && (m_run == -INFINITY || (m_row - m_run) * sc > FA_RESCALE_LOG2);
      bool have_p = spec;
      if (__any_sync(0xffffffffu, need)) {
        float alpha = 1.0f;
        if (need) {
          alpha = ex2_approx((m_run - m_row) * sc);         // m_run = -inf on the first live block -> 0
          m_run = m_row;
          l_run *= alpha;
        }
        if (j > 0) {
          pv_done();                                        // O_t is current
          uint32_t o[32];
          tmem_ld_32x32(tmem_O, o);
          tmem_ld_wait();
#pragma unroll
          for (int c = 0; c < 32; ++c) o[c] = __float_as_uint(__uint_as_float(o[c]) * alpha);
          tmem_st_32x32(tmem_O, o);
        }
        have_p = false;
      }
      if (!have_p) {
        const float neg_m_sc = m_run == -INFINITY ? 0.0f : -m_run * sc;   // every key so far padded: exponents stay finite
        (void)pass(std::true_type{}, neg_m_sc, l_blk);
      }
      tc_fence_before();
      mbar_arrive(&bars[FB_R + t]);                         // S_t(j) will not be read again: S_t(j+1) may overwrite it
      ATT_WMARK(4);
      l_run += l_blk;
      tmem_st_wait();
      tc_fence_before();    // S has been read, P (and a rescaled O) written: the tile's next MMAs may go
      mbar_arrive(&bars[FB_P + t]);
      ATT_WMARK(5);
    }

    // ---- O_t complete; total row sum = the two halves' partial sums; normalise this thread's 32 output columns
    mbar_wait(&bars[FB_O + t], (nkv - 1) & 1);
    tc_fence_after();
    float rem[32];
    {
      uint32_t o[32];
      tmem_ld_32x32(tmem_O, o);
      tmem_ld_wait();
      float* xs = xsum + (t * 2) * 128;
      xs[sp * 128 + r] = l_run;
      named_bar_sync(pair_bar, 64);
      const float inv_l = 1.0f / (l_run + xs[(sp ^ 1) * 128 + r]);
#pragma unroll
      for (int c = 0; c < 32; ++c) rem[c] = __uint_as_float(o[c]) * inv_l;
    }
    // ---- stage the 32 rows x 128 B of this warp pair in the retired Q tile; coalesced 128-byte row stores
    const int64_t seg_stride = p.M * (int64_t)p.d;
    uint8_t* stage = sQ + t * ATT_Q_BYTES + quad * 32 * 128;
    const int q_tile0 = q_first + t * ATT_BQ + quad * 32;          // first query row of this warp pair
    for (int sg = 0; sg < p.out_nseg; ++sg) {
#pragma unroll
      for (int v4 = 0; v4 < 4; ++v4) {
        uint32_t w[4];
        if (p.out_nseg == 1) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            w[k] = F16 ? pack_f16x2(rem[v4 * 8 + k * 2], rem[v4 * 8 + k * 2 + 1])
                       : pack_bf16x2(rem[v4 * 8 + k * 2], rem[v4 * 8 + k * 2 + 1]);
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int c = v4 * 8 + k * 2;
            const __nv_bfloat16 h0 = __float2bfloat16_rn(rem[c]);
            const __nv_bfloat16 h1 = __float2bfloat16_rn(rem[c + 1]);
            rem[c] -= __bfloat162float(h0);
            rem[c + 1] -= __bfloat162float(h1);
            w[k] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
          }
        }
        *reinterpret_cast<uint4*>(stage + lane * 128 + (((sp * 4 + v4) ^ (lane & 7)) << 4)) =
            make_uint4(w[0], w[1], w[2], w[3]);
      }
      named_bar_sync(pair_bar, 64);
      const int rr = lane >> 3, cc = lane & 7;
      __nv_bfloat16* gp = p.out + sg * seg_stride + ((int64_t)row_base + q_tile0) * p.d + head * ATT_HD + cc * 8;
#pragma unroll
      for (int it = 0; it < 4; ++it) {                       // this warp stores rows [16 sp, 16 sp + 16) of the pair's 32
        const int rw = sp * 16 + it * 4 + rr;
        const uint4 x4 = *reinterpret_cast<const uint4*>(stage + rw * 128 + ((cc ^ (rw & 7)) << 4));
        if (q_tile0 + rw < p.N) *reinterpret_cast<uint4*>(gp + (int64_t)rw * p.d) = x4;
      }
      if (sg + 1 < p.out_nseg) named_bar_sync(pair_bar, 64);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

// =====================================================================================================
// attn_fa2_kernel: the same data path with S DOUBLE-BUFFERED in tensor memory -- the softmax threads never wait for an MMA.
//
// What is left in attn_fa_kernel's block period (3460 cycles per 2 x 128 x 128 scores against 2048 of MUFU) is the serial
// hand-over  P(j) staged -> control warp sees it -> S(j+1) issued -> S(j+1) complete -> next pass: ~1000 cycles in which the
// tile's softmax warps idle.  With 96-key blocks a tile's TMEM quarter holds two S buffers and O (96 + 96 + 64 = 256 columns)
// when P(j) is written IN PLACE over the first 48 columns of S(j): one thread owns a whole row (its 96 scores are in registers
// before any P column is stored, so the row max is exact and nothing is speculative).  S(j+1) was issued a block ago; the
// control warp answers P(j) with P V(j) and then S(j+2) into the buffer P V(j) is reading, in order on the tensor pipe.
//   TMEM per query tile t: S / P buffer 0 [0, 96) | buffer 1 [96, 192) | O [192, 256)
//   warps 0..3 softmax of tile 0 (thread = query row), 4..7 tile 1, warp 8 control (TMA ring of four 96-key K / V stages,
//   three blocks ahead; every MMA, elected lane)
// =====================================================================================================
#ifndef AVD_F2_BKV
#define AVD_F2_BKV 96
#endif
constexpr int F2_BKV = AVD_F2_BKV;                          // 64 or 96
constexpr int F2_NCH = F2_BKV / 32;
#ifndef AVD_F2_STAGGER_NS
#define AVD_F2_STAGGER_NS 0
#endif
constexpr unsigned F2_STAGGER_NS = AVD_F2_STAGGER_NS;
constexpr int F2_STAGES = 4;
constexpr int F2_KV_BYTES = F2_BKV * ATT_HD * 2;            // 12 KB
constexpr int F2_XCHG_BYTES = (8 + 4) * 128 * 4;            // SW = 8: row max [2 bufs][2 tiles][2 halves][128], row sum [2][2][128]
constexpr int F2_SMEM = 2 * ATT_Q_BYTES + F2_STAGES * 2 * F2_KV_BYTES + F2_XCHG_BYTES + 512;
enum { GB_Q = 0, GB_KF = 1 /*+stage*/, GB_VF = 5 /*+stage*/, GB_KVE = 9 /*+stage*/, GB_S = 13 /*+2 tile + buffer*/,
       GB_P = 17 /*+tile*/, GB_O = 19 /*+tile*/, GB_COUNT = 21 };

template <bool F16, bool MASKED, int SW>   // SW = 4: one softmax thread per query row; SW = 8 (96-key blocks only): two warps per
                                           // row group, each with 48 of the block's scores in registers
__global__ void __launch_bounds__((2 * SW + 1) * 32, 1) attn_fa2_kernel(const __grid_constant__ CUtensorMap tmQ,
                                                                 const __grid_constant__ CUtensorMap tmKV,
                                                                 const AttnParams p) {
  extern __shared__ __align__(1024) uint8_t fa_smem[];
  uint8_t* smem = fa_smem;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* sQ = smem;                                       // 2 query tiles (later: output staging)
  uint8_t* sKV = sQ + 2 * ATT_Q_BYTES;                      // [stage]{K, V}
  float* xmax = reinterpret_cast<float*>(sKV + F2_STAGES * 2 * F2_KV_BYTES);
  float* xsum = xmax + 8 * 128;
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(xmax) + F2_XCHG_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + GB_COUNT);
  static_assert(SW == 4 || (SW == 8 && F2_BKV == 96), "the two-warps-per-row-group form splits a 96-key block 48 / 48");

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int head = blockIdx.y, seq = blockIdx.z;
  const int row_base = seq * p.N;
  const int q_first = blockIdx.x * 2 * ATT_BQ;
  const int n_tiles = (p.N - q_first > ATT_BQ) ? 2 : 1;
  const int colQ = head * ATT_HD, colK = p.d + head * ATT_HD, colV = 2 * p.d + head * ATT_HD;
  const int nkv = (p.N + F2_BKV - 1) / F2_BKV;
  unsigned long long* const trace = p.trace;
  (void)trace;

  if (warp == 1) {
    if (lane < GB_COUNT) {
      uint32_t count = 1;
      if (lane >= GB_P && lane < GB_O) {                    // arrivals: the warps of the tile that own valid query rows
        const int rows = min(ATT_BQ, p.N - q_first - (lane - GB_P) * ATT_BQ);
        count = rows > 0 ? (SW * 8) * ((rows + 31) >> 5) : 1;
      }
      mbar_init(&bars[lane], count);
    }
    fence_barrier_init();
  } else if (warp == 2 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmKV);
  }
  if (warp == 0) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  pdl_launch_dependents();

  if (warp == 2 * SW) {
    // ------------------------------------------------------------------ control warp (converged; elected lane issues)
    auto load_kv = [&](int b) {
      const int s = b & (F2_STAGES - 1);
      uint8_t* k = sKV + s * 2 * F2_KV_BYTES;
      mbar_expect_tx_elect(&bars[GB_KF + s], F2_KV_BYTES);
      tma_load_3d_elect(k, &tmKV, &bars[GB_KF + s], colK, row_base + b * F2_BKV, 0);
      mbar_expect_tx_elect(&bars[GB_VF + s], F2_KV_BYTES);
      tma_load_3d_elect(k + F2_KV_BYTES, &tmKV, &bars[GB_VF + s], colV, row_base + b * F2_BKV, 0);
    };
    if (lane == 0) {
      pdl_wait();
      mbar_expect_tx(&bars[GB_Q], n_tiles * ATT_Q_BYTES);
      for (int t = 0; t < n_tiles; ++t)
        tma_load_3d(sQ + t * ATT_Q_BYTES, &tmQ, &bars[GB_Q], colQ, row_base + q_first + t * ATT_BQ, 0);
    }
    __syncwarp();
    for (int b = 0; b < min(3, nkv); ++b) load_kv(b);

    const uint32_t hi = smem_desc_hi_sw128(1024);
    const uint32_t idesc_o = make_idesc_bf16(ATT_BQ, ATT_HD, 0, 1, F16);         // B = V is MN-major
    const uint32_t q_lo0 = smem_desc_lo_sw128(smem_u32(sQ), 16);
    const uint32_t kv_lo0 = smem_desc_lo_sw128(smem_u32(sKV), 16);
    auto issue_s = [&](int t, int b) {                      // S_t(b) = Q_t K_b^T -> the tile's buffer b & 1
      const int kvn = (min(F2_BKV, p.N - b * F2_BKV) + 15) & ~15;
      const uint32_t idesc_s = make_idesc_bf16(ATT_BQ, kvn, 0, 0, F16);
      const uint32_t q_lo = q_lo0 + t * (ATT_Q_BYTES >> 4);
      const uint32_t k_lo = kv_lo0 + (b & (F2_STAGES - 1)) * (2 * F2_KV_BYTES >> 4);
#pragma unroll
      for (int k = 0; k < ATT_HD / 16; ++k)
        umma_bf16_lh_elect(tmem_base + t * 256 + (b & 1) * F2_BKV, q_lo + 2 * k, k_lo + 2 * k, hi, idesc_s, k > 0 ? 1u : 0u);
      umma_commit_elect(&bars[GB_S + t * 2 + (b & 1)]);
    };
    mbar_wait(&bars[GB_Q], 0);
    for (int b = 0; b < min(2, nkv); ++b) {
      mbar_wait(&bars[GB_KF + b], 0);
      tc_fence_after();
      for (int t = 0; t < n_tiles; ++t) issue_s(t, b);
    }
    for (int j = 0; j < nkv; ++j) {
      const int s = j & (F2_STAGES - 1);
      const int kvn = (min(F2_BKV, p.N - j * F2_BKV) + 15) & ~15;
      const uint32_t v_lo = kv_lo0 + ((s * 2 * F2_KV_BYTES + F2_KV_BYTES) >> 4);
      for (int t = 0; t < n_tiles; ++t) {
        mbar_wait(&bars[GB_P + t], j & 1);                  // P_t(j) is in its buffer (and O_t rescaled where a row's max moved)
        if (t == 0) mbar_wait(&bars[GB_VF + s], (j / F2_STAGES) & 1);
        tc_fence_after();
        ATT_WMARK(t * 3);
        const uint32_t d_o = tmem_base + t * 256 + 192, a_p = tmem_base + t * 256 + (j & 1) * F2_BKV;
        // O_t += P_t(j) V_j: A from TMEM, 8 columns per 16 keys.  SW = 8: the two warps of a row group each wrote the P of
        // their 48 keys over the start of their own 48 score columns -- k-steps 3.. read from column 48 on.
        auto p_col = [&](int ks) { return (SW == 8 && ks >= 3) ? ks * 8 + 24 : ks * 8; };
        if (kvn == F2_BKV) {
#pragma unroll
          for (int ks = 0; ks < F2_BKV / 16; ++ks)
            umma_bf16_ts_elect(d_o, a_p + p_col(ks), v_lo + ks * 128, hi, idesc_o, (j > 0 || ks > 0) ? 1u : 0u);
        } else {
          for (int ks = 0; ks < (kvn >> 4); ++ks)
            umma_bf16_ts_elect(d_o, a_p + p_col(ks), v_lo + ks * 128, hi, idesc_o, (j > 0 || ks > 0) ? 1u : 0u);
        }
        umma_commit_elect(&bars[GB_O + t]);                 // O_t is current through block j (phase j)
        if (t == n_tiles - 1) umma_commit_elect(&bars[GB_KVE + s]);   // K_j / V_j have been read by every tile
        ATT_WMARK(t * 3 + 1);
        if (j + 2 < nkv) {                                  // S_t(j+2) into the buffer P V_t(j) is reading: in order behind it
          if (t == 0) {
            mbar_wait(&bars[GB_KF + ((j + 2) & (F2_STAGES - 1))], ((j + 2) / F2_STAGES) & 1);
            tc_fence_after();
          }
          issue_s(t, j + 2);
          ATT_WMARK(t * 3 + 2);
        }
      }
      if (j + 3 < nkv) {                                    // the stage of block j - 1 was handed back a block ago
        const int b = j + 3;
        if (b >= F2_STAGES) mbar_wait(&bars[GB_KVE + (b & (F2_STAGES - 1))], ((b / F2_STAGES) - 1) & 1);
        load_kv(b);
      }
    }
  } else if (SW == 4 && (warp >> 2) < n_tiles && q_first + (warp >> 2) * ATT_BQ + (warp & 3) * 32 < p.N) {
    // ------------------------------------------------------------------ softmax + output of query tile t, one thread per row
    const int t = warp >> 2, quad = warp & 3;
    const uint32_t tmem_T = tmem_base + t * 256 + ((uint32_t)(quad * 32) << 16);
    const float sc = p.scale_log2e;
    float m_run = -INFINITY, l_run = 0.0f;
    // Nothing couples the two tiles any more, and started together they stay in step: both in their MUFU-bound exponentials,
    // then both in the MUFU-free part of a block.  The second tile starts half a block late.
    if (F2_STAGGER_NS > 0 && t == 1) __nanosleep(F2_STAGGER_NS);
    for (int j = 0; j < nkv; ++j) {
      const int kv_valid = min(F2_BKV, p.N - j * F2_BKV);
      const uint32_t tmem_S = tmem_T + (j & 1) * F2_BKV;
      ATT_WMARK(0);
      mbar_wait(&bars[GB_S + t * 2 + (j & 1)], (j >> 1) & 1);
      tc_fence_after();
      ATT_WMARK(1);
      // The whole block body exists twice: RAGGED (a key_padding_mask, or the last block of a sequence that is not a
      // multiple of the block) carries the per-score selects, the common form has none (if-converted, they cost 96 SEL +
      // predicate set-up per block on every block).
      float l0 = 0.0f, l1 = 0.0f;
      bool o_seen = false;
      auto block = [&](auto ragged_tag) {
        constexpr bool RAGGED = decltype(ragged_tag)::value;
        // all chunks, always (straight-line code keeps them in registers): in a ragged last block the chunks past the
        // keys read stale TMEM, are masked to -inf and store zeros that P V (K = round16(keys)) never reads
        uint32_t x[F2_NCH][32];
#pragma unroll
        for (int i = 0; i < F2_NCH; ++i) tmem_ld_32x32(tmem_S + i * 32, x[i]);
        tmem_ld_wait();
        ATT_WMARK(2);
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        auto mask_max = [&](int i, uint32_t (&x)[32]) {       // padded / out-of-sequence keys: score -inf, probability 0
          const int lim = kv_valid - i * 32;
          uint32_t vb = lim >= 32 ? 0xffffffffu : (lim > 0 ? ((1u << lim) - 1u) : 0u);
          if (MASKED && lim > 0) vb &= ~p.kmask[(size_t)blockIdx.z * p.kmask_words + j * (F2_BKV / 32) + i];
          if (RAGGED && vb != 0xffffffffu) {
#pragma unroll
            for (int c = 0; c < 32; ++c)
              if (!((vb >> c) & 1u)) x[c] = 0xff800000u;
          }
#pragma unroll
          for (int c = 0; c < 32; ++c) mx[c & 3] = fmaxf(mx[c & 3], __uint_as_float(x[c]));
        };
#pragma unroll
        for (int i = 0; i < F2_NCH; ++i) mask_max(i, x[i]);
        const float m_row = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
        // the reference max moves only when it must: first live block of the row, or the block max is 2^FA_RESCALE_LOG2 above it (P stays
        // inside fp16, the row sum well conditioned)
        const bool need = m_row != -INFINITY && (m_run == -INFINITY || (m_row - m_run) * sc > FA_RESCALE_LOG2);
        if (__any_sync(0xffffffffu, need)) {
          float alpha = 1.0f;
          if (need) {
            alpha = ex2_approx((m_run - m_row) * sc);         // m_run = -inf on the first live block -> 0
            m_run = m_row;
            l_run *= alpha;
          }
          if (j > 0) {
            mbar_wait(&bars[GB_O + t], (j - 1) & 1);          // P V_t(j-1) complete: O_t is current
            tc_fence_after();
            o_seen = true;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              uint32_t o[32];
              tmem_ld_32x32(tmem_T + 192 + h * 32, o);
              tmem_ld_wait();
#pragma unroll
              for (int c = 0; c < 32; ++c) o[c] = __float_as_uint(__uint_as_float(o[c]) * alpha);
              tmem_st_32x32(tmem_T + 192 + h * 32, o);
            }
          }
        }
        ATT_WMARK(3);
        const float neg_m_sc = m_run == -INFINITY ? 0.0f : -m_run * sc;   // every key so far padded: exponents stay finite
        auto expo = [&](uint32_t (&x)[32], int c) {
          const float a = fmaf(__uint_as_float(x[c]), sc, neg_m_sc);
          float e;
          asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(a));
          x[c] = __float_as_uint(e);
        };
        auto post = [&](const uint32_t (&e)[32], int c, uint32_t (&w)[16]) {
          l0 += __uint_as_float(e[c]);
          l1 += __uint_as_float(e[c + 1]);
          w[c >> 1] = F16 ? pack_f16x2(__uint_as_float(e[c]), __uint_as_float(e[c + 1]))
                          : pack_bf16x2(__uint_as_float(e[c]), __uint_as_float(e[c + 1]));
        };
        // the exponentials of chunk i + 1 between the sums / packs of chunk i: nothing waits for a MUFU result less than a
        // chunk old.  P(j) goes IN PLACE over the first 48 columns of S(j) -- every score of the row is in registers.
        {
          uint32_t w[16];
#pragma unroll
          for (int c = 0; c < 32; ++c) expo(x[0], c);
#pragma unroll
          for (int i = 0; i < F2_NCH; ++i) {
            if (i + 1 < F2_NCH) {
#pragma unroll
              for (int c = 0; c < 32; c += 2) {
                expo(x[i + 1], c);
                expo(x[i + 1], c + 1);
                post(x[i], c, w);
              }
            } else {
#pragma unroll
              for (int c = 0; c < 32; c += 2) post(x[i], c, w);
            }
            tmem_st_32x16(tmem_S + i * 16, w);
          }
        }
      };
      if (MASKED || kv_valid < F2_BKV) block(std::true_type{});
      else block(std::false_type{});
      l_run += l0 + l1;
      ATT_WMARK(4);
      // S(j+1) was issued a block ago, so nothing else keeps this warp from running a block ahead of the control warp:
      // P(j) is only announced once P V(j-1) has completed, i.e. the control warp has consumed P(j-1) -- the P barrier is
      // never two phases ahead of its waiter (every block consumes one phase of the O barrier, in order)
      if (j > 0 && !o_seen) {
        mbar_wait(&bars[GB_O + t], (j - 1) & 1);
        tc_fence_after();
      }
      tmem_st_wait();
      tc_fence_before();    // S has been read, P (and a rescaled O) written: the tile's next MMAs may go
      mbar_arrive(&bars[GB_P + t]);
      ATT_WMARK(5);
    }

    // ---- O_t complete: normalise this thread's row, stage the warp's 32 rows x 128 B in the retired Q tile, coalesced stores
    mbar_wait(&bars[GB_O + t], (nkv - 1) & 1);
    tc_fence_after();
    float rem[64];
    {
      const float inv_l = 1.0f / l_run;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t o[32];
        tmem_ld_32x32(tmem_T + 192 + h * 32, o);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 32; ++c) rem[h * 32 + c] = __uint_as_float(o[c]) * inv_l;
      }
    }
    const int64_t seg_stride = p.M * (int64_t)p.d;
    uint8_t* stage = sQ + t * ATT_Q_BYTES + quad * 32 * 128;
    const int q_tile0 = q_first + t * ATT_BQ + quad * 32;          // first query row of this warp
    for (int sg = 0; sg < p.out_nseg; ++sg) {
#pragma unroll
      for (int v8 = 0; v8 < 8; ++v8) {
        uint32_t w[4];
        if (p.out_nseg == 1) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            w[k] = F16 ? pack_f16x2(rem[v8 * 8 + k * 2], rem[v8 * 8 + k * 2 + 1])
                       : pack_bf16x2(rem[v8 * 8 + k * 2], rem[v8 * 8 + k * 2 + 1]);
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int c = v8 * 8 + k * 2;
            const __nv_bfloat16 h0 = __float2bfloat16_rn(rem[c]);
            const __nv_bfloat16 h1 = __float2bfloat16_rn(rem[c + 1]);
            rem[c] -= __bfloat162float(h0);
            rem[c + 1] -= __bfloat162float(h1);
            w[k] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
          }
        }
        *reinterpret_cast<uint4*>(stage + lane * 128 + ((v8 ^ (lane & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
      }
      __syncwarp();
      const int rr = lane >> 3, cc = lane & 7;
      __nv_bfloat16* gp = p.out + sg * seg_stride + ((int64_t)row_base + q_tile0) * p.d + head * ATT_HD + cc * 8;
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const int rw = it * 4 + rr;
        const uint4 x4 = *reinterpret_cast<const uint4*>(stage + rw * 128 + ((cc ^ (rw & 7)) << 4));
        if (q_tile0 + rw < p.N) *reinterpret_cast<uint4*>(gp + (int64_t)rw * p.d) = x4;
      }
      if (sg + 1 < p.out_nseg) __syncwarp();
    }
  }
  else if (SW == 8 && warp < 16 && (warp >> 3) < n_tiles && q_first + (warp >> 3) * ATT_BQ + (warp & 3) * 32 < p.N) {
    // ------------------------------------------------------------------ softmax + output of query tile t, two warps per row group
    // Each warp keeps 48 of the block's 96 scores of its rows in registers (exact block max after one exchange with its
    // partner, nothing speculative) and writes their P over the first 24 of its own 48 score columns.
    const int t = warp >> 3, sp = (warp >> 2) & 1, quad = warp & 3;
    const int r = quad * 32 + lane;
    const uint32_t tmem_T = tmem_base + t * 256 + ((uint32_t)(quad * 32) << 16);
    const uint32_t tmem_O = tmem_T + 192 + sp * 32;
    const uint32_t pair_bar = 1 + t * 4 + quad;
    const float sc = p.scale_log2e;
    float m_run = -INFINITY, l_run = 0.0f;
    if (F2_STAGGER_NS > 0 && t == 1) __nanosleep(F2_STAGGER_NS);   // the second tile half a block behind the first
    for (int j = 0; j < nkv; ++j) {
      const int kv_valid = min(F2_BKV, p.N - j * F2_BKV);
      const uint32_t tmem_S = tmem_T + (j & 1) * F2_BKV + sp * 48;
      float* xm = xmax + (((j & 1) * 2 + t) * 2) * 128;
      ATT_WMARK(0);
      mbar_wait(&bars[GB_S + t * 2 + (j & 1)], (j >> 1) & 1);
      tc_fence_after();
      ATT_WMARK(1);
      float l0 = 0.0f, l1 = 0.0f;
      bool o_seen = false;
      auto block = [&](auto ragged_tag) {
        constexpr bool RAGGED = decltype(ragged_tag)::value;
        uint32_t xa[32], xb[16];
        tmem_ld_32x32(tmem_S, xa);
        tmem_ld_32x16(tmem_S + 32, xb);
        tmem_ld_wait();
        ATT_WMARK(2);
        if (RAGGED) {                                       // padded / out-of-sequence keys: score -inf, probability 0
          const int lim = kv_valid - sp * 48;               // my key k (0..47) is inside the sequence iff k < lim
          uint64_t vb = lim >= 48 ? 0xFFFFFFFFFFFFull : (lim > 0 ? ((1ull << lim) - 1ull) : 0ull);
          if (MASKED) {
            const uint32_t* km = p.kmask + (size_t)blockIdx.z * p.kmask_words;
            const int w0 = j * 3;
            const uint32_t m0 = w0 < p.kmask_words ? km[w0] : 0u, m1 = w0 + 1 < p.kmask_words ? km[w0 + 1] : 0u,
                           m2 = w0 + 2 < p.kmask_words ? km[w0 + 2] : 0u;
            const uint64_t pad = sp == 0 ? ((uint64_t)m0 | ((uint64_t)(m1 & 0xffffu) << 32))
                                         : ((uint64_t)(m1 >> 16) | ((uint64_t)m2 << 16));
            vb &= ~pad;
          }
#pragma unroll
          for (int c = 0; c < 32; ++c)
            if (!((vb >> c) & 1ull)) xa[c] = 0xff800000u;
#pragma unroll
          for (int c = 0; c < 16; ++c)
            if (!((vb >> (32 + c)) & 1ull)) xb[c] = 0xff800000u;
        }
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int c = 0; c < 32; ++c) mx[c & 3] = fmaxf(mx[c & 3], __uint_as_float(xa[c]));
#pragma unroll
        for (int c = 0; c < 16; ++c) mx[c & 3] = fmaxf(mx[c & 3], __uint_as_float(xb[c]));
        const float m_part = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
        xm[sp * 128 + r] = m_part;
        named_bar_sync(pair_bar, 64);
        const float m_row = fmaxf(m_part, xm[(sp ^ 1) * 128 + r]);
        const bool need = m_row != -INFINITY && (m_run == -INFINITY || (m_row - m_run) * sc > FA_RESCALE_LOG2);
        if (__any_sync(0xffffffffu, need)) {
          float alpha = 1.0f;
          if (need) {
            alpha = ex2_approx((m_run - m_row) * sc);       // m_run = -inf on the first live block -> 0
            m_run = m_row;
            l_run *= alpha;
          }
          if (j > 0) {
            mbar_wait(&bars[GB_O + t], (j - 1) & 1);        // P V_t(j-1) complete: O_t is current
            tc_fence_after();
            o_seen = true;
            uint32_t o[32];
            tmem_ld_32x32(tmem_O, o);
            tmem_ld_wait();
#pragma unroll
            for (int c = 0; c < 32; ++c) o[c] = __float_as_uint(__uint_as_float(o[c]) * alpha);
            tmem_st_32x32(tmem_O, o);
          }
        }
        ATT_WMARK(3);
        const float neg_m_sc = m_run == -INFINITY ? 0.0f : -m_run * sc;
        auto expo = [&](uint32_t& x) {
          const float a = fmaf(__uint_as_float(x), sc, neg_m_sc);
          float e;
          asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(a));
          x = __float_as_uint(e);
        };
        auto post = [&](uint32_t e0, uint32_t e1) -> uint32_t {
          l0 += __uint_as_float(e0);
          l1 += __uint_as_float(e1);
          return F16 ? pack_f16x2(__uint_as_float(e0), __uint_as_float(e1))
                     : pack_bf16x2(__uint_as_float(e0), __uint_as_float(e1));
        };
        uint32_t w[16], w2[8];
#pragma unroll
        for (int c = 0; c < 32; ++c) expo(xa[c]);
#pragma unroll
        for (int i = 0; i < 8; ++i) {                       // the last 16 exponentials between the sums / packs of the first 32
          expo(xb[2 * i]);
          expo(xb[2 * i + 1]);
          w[2 * i] = post(xa[4 * i], xa[4 * i + 1]);
          w[2 * i + 1] = post(xa[4 * i + 2], xa[4 * i + 3]);
        }
        tmem_st_32x16(tmem_S, w);
#pragma unroll
        for (int i = 0; i < 8; ++i) w2[i] = post(xb[2 * i], xb[2 * i + 1]);
        tmem_st_32x8(tmem_S + 16, w2);
      };
      if (MASKED || kv_valid < F2_BKV) block(std::true_type{});
      else block(std::false_type{});
      l_run += l0 + l1;
      ATT_WMARK(4);
      if (j > 0 && !o_seen) {                               // the P barrier is never two phases ahead of the control warp
        mbar_wait(&bars[GB_O + t], (j - 1) & 1);
        tc_fence_after();
      }
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(&bars[GB_P + t]);
      ATT_WMARK(5);
    }

    // ---- O_t complete; total row sum = the two halves' partial sums; normalise this thread's 32 output columns
    mbar_wait(&bars[GB_O + t], (nkv - 1) & 1);
    tc_fence_after();
    float rem[32];
    {
      uint32_t o[32];
      tmem_ld_32x32(tmem_O, o);
      tmem_ld_wait();
      float* xs = xsum + (t * 2) * 128;
      xs[sp * 128 + r] = l_run;
      named_bar_sync(pair_bar, 64);
      const float inv_l = 1.0f / (l_run + xs[(sp ^ 1) * 128 + r]);
#pragma unroll
      for (int c = 0; c < 32; ++c) rem[c] = __uint_as_float(o[c]) * inv_l;
    }
    const int64_t seg_stride = p.M * (int64_t)p.d;
    uint8_t* stage = sQ + t * ATT_Q_BYTES + quad * 32 * 128;
    const int q_tile0 = q_first + t * ATT_BQ + quad * 32;
    for (int sg = 0; sg < p.out_nseg; ++sg) {
#pragma unroll
      for (int v4 = 0; v4 < 4; ++v4) {
        uint32_t w[4];
        if (p.out_nseg == 1) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            w[k] = F16 ? pack_f16x2(rem[v4 * 8 + k * 2], rem[v4 * 8 + k * 2 + 1])
                       : pack_bf16x2(rem[v4 * 8 + k * 2], rem[v4 * 8 + k * 2 + 1]);
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int c = v4 * 8 + k * 2;
            const __nv_bfloat16 h0 = __float2bfloat16_rn(rem[c]);
            const __nv_bfloat16 h1 = __float2bfloat16_rn(rem[c + 1]);
            rem[c] -= __bfloat162float(h0);
            rem[c + 1] -= __bfloat162float(h1);
            w[k] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
          }
        }
        *reinterpret_cast<uint4*>(stage + lane * 128 + (((sp * 4 + v4) ^ (lane & 7)) << 4)) =
            make_uint4(w[0], w[1], w[2], w[3]);
      }
      named_bar_sync(pair_bar, 64);
      const int rr = lane >> 3, cc = lane & 7;
      __nv_bfloat16* gp = p.out + sg * seg_stride + ((int64_t)row_base + q_tile0) * p.d + head * ATT_HD + cc * 8;
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const int rw = sp * 16 + it * 4 + rr;
        const uint4 x4 = *reinterpret_cast<const uint4*>(stage + rw * 128 + ((cc ^ (rw & 7)) << 4));
        if (q_tile0 + rw < p.N) *reinterpret_cast<uint4*>(gp + (int64_t)rw * p.d) = x4;
      }
      if (sg + 1 < p.out_nseg) named_bar_sync(pair_bar, 64);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

int launch_attention_fa(const void* qkv16, const AttnParams& p, int n_seq, bool f16, cudaStream_t s) {
  if (p.N <= FA_BKV) return 0;
  const int d = p.d;
  CUtensorMap tm;
  AVD_TRY(get_tensor_map_3d(qkv16, 2, (uint64_t)3 * d, (uint64_t)p.M, 1, (uint64_t)3 * d * 2, (uint64_t)p.M * 3 * d * 2,
                            ATT_HD, ATT_BQ, &tm));
  if (options().attn_fa >= 3) {                             // S double-buffered, 96-key blocks (attn_fa2_kernel)
    CUtensorMap tmkv;
    AVD_TRY(get_tensor_map_3d(qkv16, 2, (uint64_t)3 * d, (uint64_t)p.M, 1, (uint64_t)3 * d * 2, (uint64_t)p.M * 3 * d * 2,
                              ATT_HD, F2_BKV, &tmkv));
    using Fn2 = void (*)(const CUtensorMap, const CUtensorMap, const AttnParams);
    static const Fn2 k2[2][2][2] = {
        {{attn_fa2_kernel<false, false, 4>, attn_fa2_kernel<false, true, 4>},
         {attn_fa2_kernel<true, false, 4>, attn_fa2_kernel<true, true, 4>}},
#if AVD_F2_BKV == 96
        {{attn_fa2_kernel<false, false, 8>, attn_fa2_kernel<false, true, 8>},
         {attn_fa2_kernel<true, false, 8>, attn_fa2_kernel<true, true, 8>}}};
#else
        {{attn_fa2_kernel<false, false, 4>, attn_fa2_kernel<false, true, 4>},
         {attn_fa2_kernel<true, false, 4>, attn_fa2_kernel<true, true, 4>}}};
#endif
    const bool two = options().attn_fa == 4 && F2_BKV == 96;   // two softmax warps per row group
    const Fn2 f2 = k2[two ? 1 : 0][f16 ? 1 : 0][p.kmask != nullptr ? 1 : 0];
    AVD_TRY(ensure_dyn_smem(reinterpret_cast<const void*>(f2), F2_SMEM));
    cudaLaunchConfig_t c2{};
    c2.gridDim = dim3((p.N + 2 * ATT_BQ - 1) / (2 * ATT_BQ), p.n_heads, n_seq);
    c2.blockDim = dim3(two ? 17 * 32 : 9 * 32);
    c2.dynamicSmemBytes = F2_SMEM;
    c2.stream = s;
    cudaLaunchAttribute a2[1];
    int n2 = 0;
    if (pdl_enabled()) {
      a2[n2].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      a2[n2].val.programmaticStreamSerializationAllowed = 1;
      ++n2;
    }
    c2.attrs = a2;
    c2.numAttrs = n2;
    AttnParams q2 = p;
#ifndef AVD_ATT_WTRACE
    q2.trace = nullptr;
#endif
    AVD_CUDA(cudaLaunchKernelEx(&c2, f2, tm, tmkv, q2));
    return 1;
  }
  using Fn = void (*)(const CUtensorMap, const AttnParams);
  static const Fn kernels[2][2][2] = {
      {{attn_fa_kernel<false, false, 8>, attn_fa_kernel<false, true, 8>},
       {attn_fa_kernel<true, false, 8>, attn_fa_kernel<true, true, 8>}},
      {{attn_fa_kernel<false, false, 4>, attn_fa_kernel<false, true, 4>},
       {attn_fa_kernel<true, false, 4>, attn_fa_kernel<true, true, 4>}}};
  const bool row = options().attn_fa == 2;                  // one softmax thread per query row
  const Fn fn = kernels[row ? 1 : 0][f16 ? 1 : 0][p.kmask != nullptr ? 1 : 0];
  AVD_TRY(ensure_dyn_smem(reinterpret_cast<const void*>(fn), FA_SMEM));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((p.N + 2 * ATT_BQ - 1) / (2 * ATT_BQ), p.n_heads, n_seq);
  cfg.blockDim = dim3(row ? FA_THREADS_ROW : FA_THREADS);
  cfg.dynamicSmemBytes = FA_SMEM;
  cfg.stream = s;
  cudaLaunchAttribute attrs[1];
  int na = 0;
  if (pdl_enabled()) {
    attrs[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attrs;
  cfg.numAttrs = na;
  AttnParams q = p;
#ifndef AVD_ATT_WTRACE
  q.trace = nullptr;
#endif
  AVD_CUDA(cudaLaunchKernelEx(&cfg, fn, tm, q));
  return 1;
}

}  // namespace avd
