// avd_attn.cu -- K5: joint video+audio self-attention (nn.MultiheadAttention core, mmdt.py:58-61).
//
// attn_tc_kernel (bf16 mode): flash-style.  One CTA = one (sequence, head, PAIR of 128-query tiles), 512 threads:
//   warps 0..7   softmax / output of query tile 0: warp w works on TMEM lanes (= query rows) 32 (w%4) .. +31, and the
//                two warps that share a lane quadrant split the key columns (and the 64 output columns) in halves
//   warps 8..15  the same for query tile 1
//   lane 0 of the first warp of each tile is the tile's "leader": between the softmax phases (when its warp would
//   only wait) it issues the tile's tcgen05.mma; the leader of tile 0 also issues every TMA load.  (A 17th issuer
//   warp would cap the kernel at 96 registers per thread: 5 warps on one SM sub-partition.)
// so the sequence's K/V tiles are loaded once for both query tiles and the two tiles' softmax phases run in
// parallel.  Per key block j (up to 192 keys; N = 133 is ONE block of 144 key columns):
//     S_t = Q_t K_j^T     tcgen05.mma  M=128, N=round16(keys), K=64     -> TMEM cols [256 t, 256 t + 192)
//     online softmax      tcgen05.ld rows of S, fp32 max / exp2 / sum in registers, P (bf16) -> 128B-swizzled smem
//     O_t,j = P_t V_j     tcgen05.mma  M=128, N=64, K=round16(keys)     -> TMEM cols [256 t + 192, 256 t + 256)
//                         (V consumed MN-major straight from the [keys, 64] TMA tile)
//     O_t = O_t * alpha + O_t,j   in registers
// K_{j+1} is fetched as soon as the S MMAs of block j have retired, V_{j+1} once the PV MMAs have.
// No mask: the reference's sampler passes none, and null-branch prompt rows stay in the sequence as keys/values
// (SURVEY F8).  Key columns past the sequence end are given probability 0; the rows TMA fetched for them belong to
// the next sequence (or are zero fill) and never reach the result.
//
// attn_f32_kernel (fp32-parity mode): CUDA-core fp32, one warp per query row, scores staged in smem.
#include "avd_attn_common.cuh"

#include <cstdlib>

namespace avd {

constexpr int ATT_BKV = 192;                         // key block: S_t needs ATT_BKV TMEM columns per query tile
constexpr int ATT_THREADS = 512;
constexpr int ATT_KV_BYTES = ATT_BKV * ATT_HD * 2;   // 24 KB
constexpr int ATT_P_BYTES = 3 * ATT_Q_BYTES;         // three 64-key swizzle atoms of [128 rows][128 B]
constexpr int ATT_XCHG_BYTES = (8 + 4) * 128 * 4;    // row-max exchange [2 bufs][2 tiles][2 halves][128], row-sum [2][2][128]
constexpr int ATT_THIN_ROWS = 8;                     // a second query tile with at most this many rows is "thin"
constexpr int ATT_THIN_BYTES = ATT_THIN_ROWS * ATT_BKV * 4 + 64;   // its raw scores [8][192] fp32 + row sums
constexpr int ATT_SMEM = 2 * ATT_Q_BYTES + 2 * ATT_KV_BYTES + 2 * ATT_P_BYTES + ATT_XCHG_BYTES + 256 + ATT_THIN_BYTES;
constexpr uint32_t ATT_TMEM_COLS = 512;

__device__ __forceinline__ void att_mark(unsigned long long* trace, int slot) {
#ifdef AVD_ATT_WTRACE
  return;                                  // the per-warp stamps below own the buffer
#endif
  if (trace != nullptr) {
    const int cta = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    unsigned long long g, c;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(c));
    trace[(cta * 16 + slot) * 2] = g;
    trace[(cta * 16 + slot) * 2 + 1] = c;
  }
}

// Softmax of ONE row of a thin query tile by one whole warp (lane = key column mod 32), from the raw scores the tile's
// TMEM-reading warps dumped to smem.  A thin tile (the 5 rows 128..132 of the 133-token joint sequence) would otherwise
// cost its two TMEM-lane-bound warps a full 160 MUFU instructions for 5 live lanes, all on scheduler 0: the longest
// chain in the kernel.  Writes P (16-bit, K-major 128B-swizzled 64-key atoms, zero past the valid keys) and the row sum.
template <bool F16>
__device__ __forceinline__ void thin_row_softmax(const float* srow, uint8_t* sPt, float* l_out, int i, int lane,
                                                 int kv_valid, int kv_n, float sc, const uint32_t* kmask) {
  float mx = -INFINITY;
  uint32_t dead = 0;                                  // bit k: this lane's k-th column (lane + 32 k) is a padded key
  if (kmask != nullptr)
    for (int c = lane, k = 0; c < kv_valid; c += 32, ++k) dead |= ((kmask[c >> 5] >> lane) & 1u) << k;
  for (int c = lane, k = 0; c < kv_valid; c += 32, ++k)
    if (!((dead >> k) & 1u)) mx = fmaxf(mx, srow[c]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  const float m_sc = mx * sc;
  float l = 0.0f;
  for (int c = lane, k = 0; c < kv_n; c += 32, ++k) {
    const float pv = (c < kv_valid && !((dead >> k) & 1u)) ? ex2_approx(fmaf(srow[c], sc, -m_sc)) : 0.0f;
    l += pv;
    uint8_t* dst = sPt + (c >> 6) * ATT_Q_BYTES + i * 128 + ((((c & 63) >> 3) ^ (i & 7)) << 4) + (c & 7) * 2;
    if (F16) *reinterpret_cast<__half*>(dst) = __float2half_rn(pv);
    else *reinterpret_cast<__nv_bfloat16*>(dst) = __float2bfloat16_rn(pv);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
  if (lane == 0) l_out[i] = l;
  fence_async_smem();   // P (generic-proxy stores) -> visible to the tensor core's async-proxy reads
}

// barrier slots
enum { AB_Q = 0, AB_K = 1, AB_V = 2, AB_S = 3 /*+t*/, AB_P = 5 /*+t*/, AB_O = 7 /*+t*/, AB_R = 9 /*+t*/, AB_COUNT = 11 };

#ifndef AVD_LAZY_MAX
#define AVD_LAZY_MAX 1   // multi-block softmax: one pass against the carried max (0: always max pass + exp pass; 16 x 2048 keys 348 -> 318 us)
#endif
template <bool ONE_BLOCK, bool F16, bool MASKED>   // MASKED: key_padding_mask bits present (its own instantiation: the single-block kernel sits at its 128-register limit). F16: q/k/v, P and the output are fp16 (AVD_PREC_F16) instead of bf16.  ONE_BLOCK: N <= 192 keys, the whole score row stays in registers (single softmax pass)
__global__ void __launch_bounds__(ATT_THREADS, 1) attn_tc_kernel(const __grid_constant__ CUtensorMap tmQ,
                                                                 const __grid_constant__ CUtensorMap tmKV,
                                                                 const AttnParams p) {
  extern __shared__ __align__(1024) uint8_t att_smem[];
  uint8_t* smem = att_smem;
  if ((smem_u32(smem) & 1023u) != 0) __trap();   // 128B-swizzle tiles need 1024-B aligned bases
  uint8_t* sQ = smem;                            // 2 query tiles (later: output staging)
  uint8_t* sK = sQ + 2 * ATT_Q_BYTES;
  uint8_t* sV = sK + ATT_KV_BYTES;
  uint8_t* sP = sV + ATT_KV_BYTES;               // 2 tiles x 3 atoms
  float* xmax = reinterpret_cast<float*>(sP + 2 * ATT_P_BYTES);      // [2][2 tiles][2 halves][128]
  float* xsum = xmax + 8 * 128;                                      // [2 tiles][2 halves][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * ATT_P_BYTES + ATT_XCHG_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + AB_COUNT);

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  unsigned long long* const trace = p.trace;
  if (tid == 0) att_mark(trace, 0);
  const int head = blockIdx.y;
  const int seq = blockIdx.z;
  const int row_base = seq * p.N;           // first row of this sequence in the [M, 3d] matrix
  const int q_first = blockIdx.x * 2 * ATT_BQ;
  const int n_tiles = (p.N - q_first > ATT_BQ) ? 2 : 1;          // active query tiles of this CTA
  const int colQ = head * ATT_HD, colK = p.d + head * ATT_HD, colV = 2 * p.d + head * ATT_HD;
  const int nkv = (p.N + ATT_BKV - 1) / ATT_BKV;
  const bool POLY_EXP = p.poly_exp != 0;
  // (the mask pointer is read from the kernel parameters where it is used: keeping it live costs the single-block
  // kernel, which sits at its 128-register limit, a spill)
  // bit c set = column c of 32-key chunk `ch` of key block `j` takes part in the softmax (inside the sequence, not padded)
  auto valid_bits = [&](int j, int ch, int lim) -> uint32_t {
    uint32_t v = lim >= 32 ? 0xffffffffu : (lim > 0 ? ((1u << lim) - 1u) : 0u);
    if (MASKED && lim > 0) v &= ~p.kmask[(size_t)blockIdx.z * p.kmask_words + j * (ATT_BKV / 32) + ch];
    return v;
  };
  // thin second tile: its softmax is spread over the tile's 8 warps (one row each) through smem instead of running
  // thread-per-row on the two warps that may touch its TMEM lanes
  const int rows1 = p.N - q_first - ATT_BQ;
  const bool thin = ONE_BLOCK && n_tiles == 2 && rows1 <= ATT_THIN_ROWS;
  float* thin_sc = reinterpret_cast<float*>(sP + 2 * ATT_P_BYTES + ATT_XCHG_BYTES + 256);   // [8][ATT_BKV]
  float* thin_l = thin_sc + ATT_THIN_ROWS * ATT_BKV;                                        // [8]

  if (warp == 1) {                       // one barrier per lane
    if (lane < AB_COUNT) {
      const int i = lane;
      uint32_t count = 1;
      if ((i >= AB_P && i < AB_O) || i >= AB_R) {       // arrivals: the warp pairs of the tile that own valid query rows
        const int rows = min(ATT_BQ, p.N - q_first - ((i - AB_P) & 1) * ATT_BQ);
        count = rows > 0 ? 64 * ((rows + 31) >> 5) : 1;
      }
      mbar_init(&bars[i], count);
    }
    fence_barrier_init();
  } else if (warp == 2 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmKV);
  }
  if (warp == 0) tmem_alloc<ATT_TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  if (tid == 0) att_mark(trace, 1);

  if ((warp >> 3) < n_tiles && q_first + (warp >> 3) * ATT_BQ + (warp & 3) * 32 < p.N) {
    // ------------------------------------------------------------------ softmax + output of query tile t
    const int t = warp >> 3, sp = (warp >> 2) & 1, quad = warp & 3;
    const int r = quad * 32 + lane;                 // row inside the tile == TMEM lane
    const uint32_t tmem_S = tmem_base + t * 256 + ((uint32_t)(quad * 32) << 16);
    const uint32_t tmem_O = tmem_S + ATT_BKV + sp * 32;       // this thread's 32 of the 64 output columns
    const uint32_t pair_bar = 1 + t * 4 + quad;     // named barrier of the two warps that share these rows
    uint8_t* sPt = sP + t * ATT_P_BYTES;
    float o_acc[32];
    if constexpr (!ONE_BLOCK) {
#pragma unroll
      for (int c = 0; c < 32; ++c) o_acc[c] = 0.0f;
    }
    float m_run = -INFINITY, l_run = 0.0f;
    const float sc = p.scale_log2e;
    const bool mark = (tid == 0);
    const bool leader = (warp & 7) == 0 && lane == 0;      // first lane of the tile's issuing warp
    // The tile's first warp issues its MMAs CONVERGED: all lanes wait on the barriers, an elected lane issues
    // (umma_*_elect).  Under a divergent `if (leader)` every tcgen05 instruction sits in an R2UR / vote / branch loop.
    const bool lead_warp = (warp & 7) == 0;
    const bool loader = leader && t == 0;                  // requests K_{j+1} / V_{j+1}
    const uint32_t hi = smem_desc_hi_sw128(1024);
    const uint32_t q_lo = smem_desc_lo_sw128(smem_u32(sQ + t * ATT_Q_BYTES), 16);
    const uint32_t k_lo = smem_desc_lo_sw128(smem_u32(sK), 16);
    const uint32_t v_lo = smem_desc_lo_sw128(smem_u32(sV), 16);
    const uint32_t p_lo = smem_desc_lo_sw128(smem_u32(sPt), 16);
    if (tid == 0) {                                         // the leader of tile 0 owns the TMA loads
      pdl_wait();                                           // qkv comes from the previous kernel
      mbar_expect_tx(&bars[AB_Q], n_tiles * ATT_Q_BYTES);
      for (int tt = 0; tt < n_tiles; ++tt)
        tma_load_3d(sQ + tt * ATT_Q_BYTES, &tmQ, &bars[AB_Q], colQ, row_base + q_first + tt * ATT_BQ, 0);
      mbar_expect_tx(&bars[AB_K], ATT_KV_BYTES);
      tma_load_3d(sK, &tmKV, &bars[AB_K], colK, row_base, 0);
      mbar_expect_tx(&bars[AB_V], ATT_KV_BYTES);
      tma_load_3d(sV, &tmKV, &bars[AB_V], colV, row_base, 0);
    }

    for (int j = 0; j < nkv; ++j) {
      const uint32_t ph = j & 1;
      const int kv_valid = min(ATT_BKV, p.N - j * ATT_BKV);
      const int nch = (kv_valid + 31) >> 5;         // 32-column chunks holding valid keys (<= 6)
      const int ch0 = sp * 3;                       // this thread's chunks: [ch0, ch1)
      const int ch1 = min(nch, ch0 + 3);
      float* xm = xmax + ((ph * 2 + t) * 2) * 128;
      const int kv_n = (kv_valid + 15) & ~15;
      ATT_WMARK(0);
      if (lead_warp && j == 0) {                    // S_t = Q_t K_0^T (later blocks: issued right behind P V of the block before)
        mbar_wait(&bars[AB_Q], 0);
        mbar_wait(&bars[AB_K], ph);
        tc_fence_after();
        if (mark) att_mark(trace, 2);
        const uint32_t idesc_s = make_idesc_bf16(ATT_BQ, kv_n, 0, 0, F16);
#pragma unroll
        for (int k = 0; k < ATT_HD / 16; ++k)
          umma_bf16_lh_elect(tmem_base + t * 256, q_lo + 2 * k, k_lo + 2 * k, hi, idesc_s, k > 0 ? 1u : 0u);
        umma_commit_elect(&bars[AB_S + t]);
        if (mark) att_mark(trace, 3);
      }
      __syncwarp();
      mbar_wait(&bars[AB_S + t], ph);
      tc_fence_after();
      ATT_WMARK(1);
      if (loader && j + 1 < nkv) {                  // K_{j+1} may land as soon as every tile's S MMAs have read K_j
        for (int tt = 0; tt < n_tiles; ++tt)
          if (tt != t) mbar_wait(&bars[AB_S + tt], ph);
        mbar_expect_tx(&bars[AB_K], ATT_KV_BYTES);
        tma_load_3d(sK, &tmKV, &bars[AB_K], colK, row_base + (j + 1) * ATT_BKV, 0);
      }
      if (mark && j == 0) att_mark(trace, 5);
      float l_blk = 0.0f, alpha;
      if constexpr (ONE_BLOCK) {
        // ---- single pass: up to 3 chunks (96 scores) of this row live in registers
        uint32_t x[3][32];
#pragma unroll
        for (int i = 0; i < 3; ++i)
          if (i == 0 || ch0 + i < nch) tmem_ld_32x32(tmem_S + (ch0 + i) * 32, x[i]);   // columns past the keys are masked
        tmem_ld_wait();
        if (thin && t == 1) {
          if (r < rows1) {                           // raw scores of the live rows -> smem
#pragma unroll
            for (int i = 0; i < 3; ++i)
              if (ch0 + i < nch) {
                float4* dst = reinterpret_cast<float4*>(thin_sc + r * ATT_BKV + (ch0 + i) * 32);
#pragma unroll
                for (int c = 0; c < 8; ++c)
                  dst[c] = make_float4(__uint_as_float(x[i][4 * c]), __uint_as_float(x[i][4 * c + 1]),
                                       __uint_as_float(x[i][4 * c + 2]), __uint_as_float(x[i][4 * c + 3]));
              }
          }
          named_bar_sync(9, 256);                    // all 8 warps of the tile: scores are in smem
          if ((warp & 7) < rows1)
            thin_row_softmax<F16>(thin_sc + (warp & 7) * ATT_BKV, sPt, thin_l, warp & 7, lane, kv_valid, kv_n, sc,
                                  MASKED ? p.kmask + (size_t)blockIdx.z * p.kmask_words : nullptr);
          named_bar_sync(10, 256);                   // P rows and row sums written
          alpha = 0.0f;
          l_blk = (sp == 0 && r < rows1) ? thin_l[r] : 0.0f;
        } else {
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          const uint32_t vb = valid_bits(j, ch0 + i, kv_valid - (ch0 + i) * 32);
          if (vb == 0xffffffffu) {                   // whole chunk of valid keys: no masking
#pragma unroll
            for (int c = 0; c < 32; ++c) mx[c & 3] = fmaxf(mx[c & 3], __uint_as_float(x[i][c]));
          } else if (vb != 0u) {
#pragma unroll
            for (int c = 0; c < 32; ++c)
              mx[c & 3] = fmaxf(mx[c & 3], ((vb >> c) & 1u) ? __uint_as_float(x[i][c]) : -INFINITY);
          }
        }
        const float m_part = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
        xm[sp * 128 + r] = m_part;
        named_bar_sync(pair_bar, 64);
        const float m_new = fmaxf(m_part, xm[(sp ^ 1) * 128 + r]);
        if (mark && j == 0) att_mark(trace, 6);
        alpha = 0.0f;
        const float m_sc = m_new * sc;
        m_run = m_new;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          const int ch = ch0 + i;
          const int lim = kv_valid - ch * 32;
          if (lim > -16) {                           // chunks wholly past round16(keys) are never read by the PV MMA
            float l0 = 0.0f, l1 = 0.0f;
            uint32_t w[16];
            const uint32_t vb = valid_bits(j, ch, lim);
            float e[32];
            ex2_chunk(x[i], sc, -m_sc, e);
            if (vb == 0xffffffffu) {                 // whole chunk of valid keys: no masking
#pragma unroll
              for (int c = 0; c < 32; c += 2) {
                const float p0 = e[c];
                const float p1 = e[c + 1];
                l0 += p0;
                l1 += p1;
                w[c >> 1] = F16 ? pack_f16x2(p0, p1) : pack_bf16x2(p0, p1);
              }
            } else {
#pragma unroll
              for (int c = 0; c < 32; c += 2) {
                float p0 = e[c];
                float p1 = e[c + 1];
                p0 = ((vb >> c) & 1u) ? p0 : 0.0f;
                p1 = ((vb >> (c + 1)) & 1u) ? p1 : 0.0f;
                l0 += p0;
                l1 += p1;
                w[c >> 1] = F16 ? pack_f16x2(p0, p1) : pack_bf16x2(p0, p1);
              }
            }
            l_blk += l0 + l1;
            uint8_t* atom = sPt + (ch >> 1) * ATT_Q_BYTES + r * 128;
            const int q_base = (ch & 1) * 4;  // 16-byte chunk index of this 32-column group inside the 128-byte row
#pragma unroll
            for (int q = 0; q < 4; ++q)
              *reinterpret_cast<uint4*>(atom + (((q_base + q) ^ (r & 7)) << 4)) =
                  make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
          }
        }
        }   // !thin
      } else if (AVD_LAZY_MAX && j > 0) {
        // ---- lazy reference max: exponentials against the carried m_run in ONE pass over the S columns while the block's
        // own max is tracked; the pass is repeated with the new max only when some row of the warp moved up by more than
        // 2^8 (P must stay inside fp16 / the row sum well conditioned).  Both warps of a pair hold the same 32 rows and
        // see the same exchanged max, so they take the same branch.
        auto pass = [&](float m_sc, float& m_part) {
          float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
          l_blk = 0.0f;
          for (int ch = ch0; ch < ch1; ++ch) {
            uint32_t x[32];
            tmem_ld_32x32(tmem_S + ch * 32, x);
            tmem_ld_wait();
            const uint32_t vm = valid_bits(j, ch, kv_valid - ch * 32);
            float l0 = 0.0f, l1 = 0.0f;
            uint32_t w[16];
            if (vm == 0xffffffffu) {
              float e[32];
              ex2_chunk(x, sc, -m_sc, e);
#pragma unroll
              for (int c = 0; c < 32; c += 2) {
                mx[c & 3] = fmaxf(mx[c & 3], __uint_as_float(x[c]));
                mx[(c + 1) & 3] = fmaxf(mx[(c + 1) & 3], __uint_as_float(x[c + 1]));
                l0 += e[c];
                l1 += e[c + 1];
                w[c >> 1] = F16 ? pack_f16x2(e[c], e[c + 1]) : pack_bf16x2(e[c], e[c + 1]);
              }
            } else {
#pragma unroll
              for (int c = 0; c < 32; c += 2) {
                const bool v0 = (vm >> c) & 1u, v1 = (vm >> (c + 1)) & 1u;
                const float s0 = __uint_as_float(x[c]), s1 = __uint_as_float(x[c + 1]);
                mx[c & 3] = fmaxf(mx[c & 3], v0 ? s0 : -INFINITY);
                mx[(c + 1) & 3] = fmaxf(mx[(c + 1) & 3], v1 ? s1 : -INFINITY);
                const float p0 = v0 ? ex2_approx(fmaf(s0, sc, -m_sc)) : 0.0f;
                const float p1 = v1 ? ex2_approx(fmaf(s1, sc, -m_sc)) : 0.0f;
                l0 += p0;
                l1 += p1;
                w[c >> 1] = F16 ? pack_f16x2(p0, p1) : pack_bf16x2(p0, p1);
              }
            }
            l_blk += l0 + l1;
            uint8_t* atom = sPt + (ch >> 1) * ATT_Q_BYTES + r * 128;
            const int q_base = (ch & 1) * 4;
#pragma unroll
            for (int q = 0; q < 4; ++q)
              *reinterpret_cast<uint4*>(atom + (((q_base + q) ^ (r & 7)) << 4)) =
                  make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
          }
          m_part = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
        };
        const float m_ref = m_run == -INFINITY ? 0.0f : m_run;
        float m_part;
        pass(m_ref * sc, m_part);
        ATT_WMARK(2);
        xm[sp * 128 + r] = m_part;
        named_bar_sync(pair_bar, 64);
        const float m_row = fmaxf(m_part, xm[(sp ^ 1) * 128 + r]);
        ATT_WMARK(3);
        const bool need = m_row != -INFINITY && (m_run == -INFINITY || (m_row - m_run) * sc > 8.0f);
        alpha = 1.0f;
        if (__any_sync(0xffffffffu, need)) {
          const float m_new = fmaxf(m_run, m_row);
          const float m_use = m_new == -INFINITY ? 0.0f : m_new;
          alpha = ex2_approx((m_run - m_use) * sc);
          m_run = m_new;
          float dummy;
          pass(m_use * sc, dummy);
          ATT_WMARK(4);
        }
      } else {
        // ---- pass 1: row max over this thread's valid keys
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        for (int ch = ch0; ch < ch1; ++ch) {
          uint32_t x[32];
          tmem_ld_32x32(tmem_S + ch * 32, x);
          tmem_ld_wait();
          const uint32_t vm = valid_bits(j, ch, kv_valid - ch * 32);
#pragma unroll
          for (int c = 0; c < 32; ++c) mx[c & 3] = fmaxf(mx[c & 3], ((vm >> c) & 1u) ? __uint_as_float(x[c]) : -INFINITY);
        }
        const float m_part = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
        xm[sp * 128 + r] = m_part;
        named_bar_sync(pair_bar, 64);
        const float m_new = fmaxf(m_run, fmaxf(m_part, xm[(sp ^ 1) * 128 + r]));
        const float m_use = m_new == -INFINITY ? 0.0f : m_new;   // every key so far padded: keep the exponents finite
        alpha = ex2_approx((m_run - m_use) * sc);   // m_run = -inf on the first block -> 0
        const float m_sc = m_use * sc;
        m_run = m_new;
        // ---- pass 2: p = exp2(s*c - m*c), row sum, P -> smem (K-major, 128B swizzle, 64-key atoms)
        for (int ch = ch0; ch < ch1; ++ch) {
          uint32_t x[32];
          tmem_ld_32x32(tmem_S + ch * 32, x);
          tmem_ld_wait();
          const uint32_t vm = valid_bits(j, ch, kv_valid - ch * 32);
          float l0 = 0.0f, l1 = 0.0f;
          uint32_t w[16];
          if (vm == 0xffffffffu && POLY_EXP) {               // whole chunk of valid keys, MUFU and FMA pipes share the exponentials
#pragma unroll
            for (int c = 0; c < 32; c += 2) {
              const float p0 = ex2_approx(fmaf(__uint_as_float(x[c]), sc, -m_sc));
              const float p1 = ex2_poly(fmaf(__uint_as_float(x[c + 1]), sc, -m_sc));
              l0 += p0;
              l1 += p1;
              w[c >> 1] = F16 ? pack_f16x2(p0, p1) : pack_bf16x2(p0, p1);
            }
          } else if (vm == 0xffffffffu) {
            float e[32];
            ex2_chunk(x, sc, -m_sc, e);
#pragma unroll
            for (int c = 0; c < 32; c += 2) {
              const float p0 = e[c];
              const float p1 = e[c + 1];
              l0 += p0;
              l1 += p1;
              w[c >> 1] = F16 ? pack_f16x2(p0, p1) : pack_bf16x2(p0, p1);
            }
          } else {
#pragma unroll
            for (int c = 0; c < 32; c += 2) {
              float p0 = ex2_approx(fmaf(__uint_as_float(x[c]), sc, -m_sc));
              float p1 = ex2_approx(fmaf(__uint_as_float(x[c + 1]), sc, -m_sc));
              p0 = ((vm >> c) & 1u) ? p0 : 0.0f;
              p1 = ((vm >> (c + 1)) & 1u) ? p1 : 0.0f;
              l0 += p0;
              l1 += p1;
              w[c >> 1] = F16 ? pack_f16x2(p0, p1) : pack_bf16x2(p0, p1);
            }
          }
          l_blk += l0 + l1;
          uint8_t* atom = sPt + (ch >> 1) * ATT_Q_BYTES + r * 128;
          const int q_base = (ch & 1) * 4;
#pragma unroll
          for (int q = 0; q < 4; ++q)
            *reinterpret_cast<uint4*>(atom + (((q_base + q) ^ (r & 7)) << 4)) =
                make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
        }
      }
      l_run = l_run * alpha + l_blk;                // partial row sum over this thread's key columns
      fence_async_smem();   // P (generic-proxy stores) -> visible to the tensor core's async-proxy reads
      tc_fence_before();    // the S columns have been read: the next S MMA may overwrite them
      mbar_arrive(&bars[AB_P + t]);
      ATT_WMARK(5);
      if (mark && j == 0) att_mark(trace, 7);
      if (lead_warp) {                              // O_t,j = P_t V_j
        mbar_wait(&bars[AB_P + t], ph);
        mbar_wait(&bars[AB_V], ph);
        if (j > 0) mbar_wait(&bars[AB_R + t], ph ^ 1);     // O_t,j-1 has been read out of TMEM
        tc_fence_after();
        ATT_WMARK(9);
        const uint32_t idesc_o = make_idesc_bf16(ATT_BQ, ATT_HD, 0, 1, F16);   // B = V is MN-major
        const int ksteps = kv_n >> 4;
        for (int ks = 0; ks < ksteps; ++ks) {
          const uint32_t pa = p_lo + (((ks >> 2) * ATT_Q_BYTES + (ks & 3) * 32) >> 4);
          const uint32_t va = v_lo + ((ks * 16 * 128) >> 4);            // 16 keys = 16 rows of 128 B
          umma_bf16_lh_elect(tmem_base + t * 256 + ATT_BKV, pa, va, hi, idesc_o, ks > 0 ? 1u : 0u);
        }
        umma_commit_elect(&bars[AB_O + t]);
        if (mark && j == 0) att_mark(trace, 4);
        ATT_WMARK(10);
        if (!ONE_BLOCK && j + 1 < nkv) {
          // S_t(j+1) = Q_t K_{j+1}^T goes in right behind P V: the S columns were drained when P_j was staged (AB_P), and
          // K_{j+1} was requested when the S MMAs of block j retired -- the tensor pipe runs it under this tile's O update
          // instead of after it
          const int kvn1 = (min(ATT_BKV, p.N - (j + 1) * ATT_BKV) + 15) & ~15;
          mbar_wait(&bars[AB_K], ph ^ 1);
          tc_fence_after();
          ATT_WMARK(11);
          const uint32_t idesc_s = make_idesc_bf16(ATT_BQ, kvn1, 0, 0, F16);
#pragma unroll
          for (int k = 0; k < ATT_HD / 16; ++k)
            umma_bf16_lh_elect(tmem_base + t * 256, q_lo + 2 * k, k_lo + 2 * k, hi, idesc_s, k > 0 ? 1u : 0u);
          umma_commit_elect(&bars[AB_S + t]);
        }
      }
      __syncwarp();
      ATT_WMARK(6);
      mbar_wait(&bars[AB_O + t], ph);
      tc_fence_after();
      ATT_WMARK(7);
      if (mark && j == 0) att_mark(trace, 8);
      {   // ---- O = O * alpha + O_j on this thread's 32 output columns
        uint32_t x[32];
        tmem_ld_32x32(tmem_O, x);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          if constexpr (ONE_BLOCK) o_acc[c] = __uint_as_float(x[c]);
          else o_acc[c] = fmaf(o_acc[c], alpha, __uint_as_float(x[c]));
        }
      }
      ATT_WMARK(8);
      if (j + 1 < nkv) {
        tc_fence_before();
        mbar_arrive(&bars[AB_R + t]);
        if (loader) {                               // V_{j+1} once every tile's PV MMAs have read V_j
          for (int tt = 0; tt < n_tiles; ++tt)
            if (tt != t) mbar_wait(&bars[AB_O + tt], ph);
          mbar_expect_tx(&bars[AB_V], ATT_KV_BYTES);
          tma_load_3d(sV, &tmKV, &bars[AB_V], colV, row_base + (j + 1) * ATT_BKV, 0);
        }
      }
    }
    if (mark) att_mark(trace, 9);

    // ---- total row sum = the two halves' partial sums; normalise this thread's 32 output columns
    float* xs = xsum + (t * 2) * 128;
    xs[sp * 128 + r] = l_run;
    named_bar_sync(pair_bar, 64);
    const float inv_l = 1.0f / (l_run + xs[(sp ^ 1) * 128 + r]);
    // ---- stage the 32 rows x 128 B of this warp pair in the retired Q tile; coalesced 128-byte row stores
    const int64_t seg_stride = p.M * (int64_t)p.d;
    uint8_t* stage = sQ + t * ATT_Q_BYTES + quad * 32 * 128;
    const int q_tile0 = q_first + t * ATT_BQ + quad * 32;        // first query row of this warp pair
    float rem[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) rem[c] = o_acc[c] * inv_l;
    for (int sg = 0; sg < p.out_nseg; ++sg) {
#pragma unroll
      for (int v4 = 0; v4 < 4; ++v4) {
        uint32_t w[4];
        if (p.out_nseg == 1) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            w[k] = F16 ? pack_f16x2(rem[v4 * 8 + k * 2], rem[v4 * 8 + k * 2 + 1]) : pack_bf16x2(rem[v4 * 8 + k * 2], rem[v4 * 8 + k * 2 + 1]);
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
      for (int it = 0; it < 4; ++it) {               // this warp stores rows [16 sp, 16 sp + 16) of the pair's 32
        const int rw = sp * 16 + it * 4 + rr;
        const uint4 x4 = *reinterpret_cast<const uint4*>(stage + rw * 128 + ((cc ^ (rw & 7)) << 4));
        if (q_tile0 + rw < p.N) *reinterpret_cast<uint4*>(gp + (int64_t)rw * p.d) = x4;
      }
      if (sg + 1 < p.out_nseg) named_bar_sync(pair_bar, 64);
    }
    if (mark) att_mark(trace, 10);
  } else if (thin && (warp >> 3) == 1) {
    // ------------------------------------------------------------------ the other six warps of a thin tile: one row each
    named_bar_sync(9, 256);
    if ((warp & 7) < rows1)
      thin_row_softmax<F16>(thin_sc + (warp & 7) * ATT_BKV, sP + ATT_P_BYTES, thin_l, warp & 7, lane, p.N,
                            (p.N + 15) & ~15, p.scale_log2e,
                            MASKED ? p.kmask + (size_t)blockIdx.z * p.kmask_words : nullptr);
    named_bar_sync(10, 256);
  }

  tc_fence_before();
  __syncthreads();
  if (tid == 0) att_mark(trace, 11);
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc<ATT_TMEM_COLS>(tmem_base);
  }
}

// =====================================================================================================
// attn_pipe_kernel: the multi-block case (N > 192 keys) with the three stages of a key block running concurrently.
//
// attn_tc_kernel's multi-block path is a serial chain per query tile -- S = Q K^T, row max (first pass over S in TMEM),
// exp / P (second pass), O_j = P V, O update -- 8400 cycles per 2 x 128 x 192 block against 3072 of MUFU and 1920 of
// MMA.  Here one CTA owns ONE 128-query tile, which frees TMEM for two S buffers and registers
// for a whole 96-score half row per thread (one tcgen05.ld pass instead of two):
//   warp 8 (one lane)  issuer: TMA loads of Q and of the K / V blocks into double buffers; S(j+2) = Q K_{j+2}^T is issued
//                      as soon as softmax(j) has pulled S(j) out of its TMEM buffer, O_j = P_j V_j as soon as P_j is staged
//   warps 0..7         softmax of block j -- thread = query row, the two warps of a lane quadrant split the 192 key
//                      columns -- while the tensor pipe runs S(j+1) / S(j+2) and P(j-1) V(j-1); the read-back of O(j-1)
//                      is deferred to the start of block j (P V of block j-1 had the whole S wait to finish)
// Per block the softmax warps then see: wait S (already there), 3 x tcgen05.ld, max, 96 exponentials, P store; the
// MUFU pipe (16 ex2 per clock per SM) is the floor: 128 x 192 / 16 = 1536 cycles.
//
// MEASURED (tools/kbench.py --M 32768, 16 x 2048 tokens): 474 us against 358 us for attn_tc_kernel's multi-block path --
// 6000 cycles per 128 x 192 block.  With ONE query tile per SM only 8 softmax warps (2 per scheduler) are resident, and
// each of them walks a dependent chain per block (tcgen05.ld round trip, O read-back, pair exchange, 96 MUFU ops issued
// from one warp, P stores) with nobody to fill its bubbles; attn_tc_kernel's two tiles x 8 warps hide each other's.  The
// overlap this kernel adds (S(j+2) and P V under the softmax) is worth less than the second tile's warps.  It is therefore
// OFF by default (option attn_pipe); a variant of attn_tc_kernel reading S once per 128-key block was measured too
// (380 us): the multi-block path is not bound by tcgen05.ld volume either.  What is left is an FA4-style split (two
// softmax warpgroups on two tiles, O kept in TMEM with lazy rescaling, a dedicated correction group), DESIGN.md section 5.
// =====================================================================================================
constexpr int AP_THREADS = 288;
constexpr int AP_XCHG_BYTES = (4 + 2) * 128 * 4;      // row max [2 bufs][2 halves][128], row sum [2 halves][128]
constexpr int AP_SMEM = ATT_Q_BYTES + 4 * ATT_KV_BYTES + 2 * ATT_P_BYTES + AP_XCHG_BYTES + 256;
enum { PB_Q = 0, PB_K = 1 /*+b*/, PB_V = 3 /*+b*/, PB_S = 5 /*+b*/, PB_SF = 7 /*+b*/, PB_P = 9 /*+b*/, PB_O = 11, PB_OF = 12,
       PB_COUNT = 13 };

template <bool F16, bool MASKED>
__global__ void __launch_bounds__(AP_THREADS, 1) attn_pipe_kernel(const __grid_constant__ CUtensorMap tmQ,
                                                                  const __grid_constant__ CUtensorMap tmKV,
                                                                  const AttnParams p) {
  extern __shared__ __align__(1024) uint8_t ap_smem[];
  uint8_t* smem = ap_smem;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* sQ = smem;                                  // later: output staging
  uint8_t* sK = sQ + ATT_Q_BYTES;                      // [2]
  uint8_t* sV = sK + 2 * ATT_KV_BYTES;                 // [2]
  uint8_t* sP = sV + 2 * ATT_KV_BYTES;                 // [2] x 3 atoms
  float* xmax = reinterpret_cast<float*>(sP + 2 * ATT_P_BYTES);
  float* xsum = xmax + 4 * 128;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * ATT_P_BYTES + AP_XCHG_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + PB_COUNT);

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int head = blockIdx.y;
  const int seq = blockIdx.z;
  const int row_base = seq * p.N;
  const int q0 = blockIdx.x * ATT_BQ;
  const int colQ = head * ATT_HD, colK = p.d + head * ATT_HD, colV = 2 * p.d + head * ATT_HD;
  const int nkv = (p.N + ATT_BKV - 1) / ATT_BKV;

  if (warp == 0) {
    if (lane < PB_COUNT) {
      const int i = lane;
      const bool warps8 = (i >= PB_SF && i < PB_O) || i == PB_OF;      // arrivals: one per softmax warp
      mbar_init(&bars[i], warps8 ? 8 : 1);
    }
    fence_barrier_init();
  } else if (warp == 1 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmKV);
  }
  if (warp == 8) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  const uint32_t tmem_O = tmem_base + 2 * ATT_BKV;     // S buffers at columns [0,192) and [192,384), O at [384,448)

  if (warp == 8) {
    // ------------------------------------------------------------------ issuer
    if (lane == 0) {
      pdl_wait();                                      // qkv comes from the previous kernel
      mbar_expect_tx(&bars[PB_Q], ATT_Q_BYTES);
      tma_load_3d(sQ, &tmQ, &bars[PB_Q], colQ, row_base + q0, 0);
      for (int b = 0; b < 2 && b < nkv; ++b) {
        mbar_expect_tx(&bars[PB_K + b], ATT_KV_BYTES);
        tma_load_3d(sK + b * ATT_KV_BYTES, &tmKV, &bars[PB_K + b], colK, row_base + b * ATT_BKV, 0);
        mbar_expect_tx(&bars[PB_V + b], ATT_KV_BYTES);
        tma_load_3d(sV + b * ATT_KV_BYTES, &tmKV, &bars[PB_V + b], colV, row_base + b * ATT_BKV, 0);
      }
      const uint32_t hi = smem_desc_hi_sw128(1024);
      const uint32_t q_lo = smem_desc_lo_sw128(smem_u32(sQ), 16);
      auto issue_S = [&](int j) {                      // S(j) = Q K_j^T into S buffer j & 1
        const int b = j & 1;
        const int kv_valid = min(ATT_BKV, p.N - j * ATT_BKV);
        const int kv_n = (kv_valid + 15) & ~15;
        const uint32_t idesc_s = make_idesc_bf16(ATT_BQ, kv_n, 0, 0, F16);
        const uint32_t k_lo = smem_desc_lo_sw128(smem_u32(sK + b * ATT_KV_BYTES), 16);
#pragma unroll
        for (int k = 0; k < ATT_HD / 16; ++k)
          umma_bf16_lh(tmem_base + b * ATT_BKV, q_lo + 2 * k, k_lo + 2 * k, hi, idesc_s, k > 0 ? 1u : 0u);
        umma_commit(&bars[PB_S + b]);
      };
      mbar_wait(&bars[PB_Q], 0);
      mbar_wait(&bars[PB_K], 0);
      tc_fence_after();
      issue_S(0);
      if (nkv > 1) {
        mbar_wait(&bars[PB_K + 1], 0);
        tc_fence_after();
        issue_S(1);
      }
      const uint32_t idesc_o = make_idesc_bf16(ATT_BQ, ATT_HD, 0, 1, F16);    // B = V is MN-major
      for (int j = 0; j < nkv; ++j) {
        const int b = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        const bool more = j + 2 < nkv;
        if (more) {                                    // K buffer b is free once the S(j) MMAs have executed
          mbar_wait(&bars[PB_S + b], ph);
          mbar_expect_tx(&bars[PB_K + b], ATT_KV_BYTES);
          tma_load_3d(sK + b * ATT_KV_BYTES, &tmKV, &bars[PB_K + b], colK, row_base + (j + 2) * ATT_BKV, 0);
        }
        // O_j = P_j V_j: P staged, V landed, O(j-1) read out of TMEM
        mbar_wait(&bars[PB_P + b], ph);
        mbar_wait(&bars[PB_V + b], ph);
        if (j > 0) mbar_wait(&bars[PB_OF], (j - 1) & 1);
        tc_fence_after();
        {
          const int kv_valid = min(ATT_BKV, p.N - j * ATT_BKV);
          const int ksteps = ((kv_valid + 15) & ~15) >> 4;
          const uint32_t p_lo = smem_desc_lo_sw128(smem_u32(sP + b * ATT_P_BYTES), 16);
          const uint32_t v_lo = smem_desc_lo_sw128(smem_u32(sV + b * ATT_KV_BYTES), 16);
          for (int ks = 0; ks < ksteps; ++ks) {
            const uint32_t pa = p_lo + (((ks >> 2) * ATT_Q_BYTES + (ks & 3) * 32) >> 4);
            const uint32_t va = v_lo + ((ks * 16 * 128) >> 4);                  // 16 keys = 16 rows of 128 B
            umma_bf16_lh(tmem_O, pa, va, hi, idesc_o, ks > 0 ? 1u : 0u);
          }
          umma_commit(&bars[PB_O]);
        }
        if (more) {                                    // S(j+2): S buffer b has been read by softmax(j) (it staged P_j after)
          mbar_wait(&bars[PB_SF + b], ph);
          mbar_wait(&bars[PB_K + b], ph ^ 1);
          tc_fence_after();
          issue_S(j + 2);
          mbar_wait(&bars[PB_O], j & 1);               // V buffer b is free once the P V MMAs of block j have executed
          mbar_expect_tx(&bars[PB_V + b], ATT_KV_BYTES);
          tma_load_3d(sV + b * ATT_KV_BYTES, &tmKV, &bars[PB_V + b], colV, row_base + (j + 2) * ATT_BKV, 0);
        }
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ softmax + output
    const int sp = warp >> 2, quad = warp & 3;
    const int r = quad * 32 + lane;                    // row inside the tile == TMEM lane
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const uint32_t pair_bar = 1 + quad;
    const float sc = p.scale_log2e;
    float o_acc[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) o_acc[c] = 0.0f;
    float m_run = -INFINITY, l_run = 0.0f, alpha_pend = 0.0f;
    auto valid_bits = [&](int j, int ch, int lim) -> uint32_t {
      uint32_t v = lim >= 32 ? 0xffffffffu : (lim > 0 ? ((1u << lim) - 1u) : 0u);
      if (MASKED && lim > 0) v &= ~p.kmask[(size_t)blockIdx.z * p.kmask_words + j * (ATT_BKV / 32) + ch];
      return v;
    };
    auto drain_O = [&](int j) {                        // o_acc = o_acc * alpha_j + O_j  (this thread's 32 of the 64 columns)
      mbar_wait(&bars[PB_O], j & 1);
      tc_fence_after();
      uint32_t xo[32];
      tmem_ld_32x32(tmem_O + lane_base + sp * 32, xo);
      tmem_ld_wait();
#pragma unroll
      for (int c = 0; c < 32; ++c) o_acc[c] = fmaf(o_acc[c], alpha_pend, __uint_as_float(xo[c]));
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars[PB_OF]);
    };
    for (int j = 0; j < nkv; ++j) {
      const int b = j & 1;
      const uint32_t ph = (j >> 1) & 1;
      const int kv_valid = min(ATT_BKV, p.N - j * ATT_BKV);
      const int nch = (kv_valid + 31) >> 5;
      const int ch0 = sp * 3;
      mbar_wait(&bars[PB_S + b], ph);
      tc_fence_after();
      uint32_t x[3][32];
#pragma unroll
      for (int i = 0; i < 3; ++i)
        if (i == 0 || ch0 + i < nch) tmem_ld_32x32(tmem_base + b * ATT_BKV + lane_base + (ch0 + i) * 32, x[i]);
      tmem_ld_wait();
      tc_fence_before();                               // S buffer b may be overwritten by S(j+2)
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars[PB_SF + b]);
      if (j > 0) drain_O(j - 1);
      // ---- row max over this thread's valid keys, exchanged with the other half's warp
      float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const uint32_t vb = valid_bits(j, ch0 + i, kv_valid - (ch0 + i) * 32);
        if (vb == 0xffffffffu) {
#pragma unroll
          for (int c = 0; c < 32; ++c) mx[c & 3] = fmaxf(mx[c & 3], __uint_as_float(x[i][c]));
        } else if (vb != 0u) {
#pragma unroll
          for (int c = 0; c < 32; ++c)
            mx[c & 3] = fmaxf(mx[c & 3], ((vb >> c) & 1u) ? __uint_as_float(x[i][c]) : -INFINITY);
        }
      }
      const float m_part = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
      float* xm = xmax + (b * 2) * 128;
      xm[sp * 128 + r] = m_part;
      named_bar_sync(pair_bar, 64);
      const float m_new = fmaxf(m_run, fmaxf(m_part, xm[(sp ^ 1) * 128 + r]));
      const float m_use = m_new == -INFINITY ? 0.0f : m_new;   // every key so far padded: keep the exponents finite
      const float alpha = ex2_approx((m_run - m_use) * sc);    // m_run = -inf on the first block -> 0
      const float m_sc = m_use * sc;
      m_run = m_new;
      // ---- p = exp2(s*c - m*c), row sum, P -> smem (K-major, 128B swizzle, 64-key atoms)
      uint8_t* sPb = sP + b * ATT_P_BYTES;
      float l_blk = 0.0f;
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const int ch = ch0 + i;
        const int lim = kv_valid - ch * 32;
        if (lim > -16) {                               // chunks wholly past round16(keys) are never read by the PV MMA
          float l0 = 0.0f, l1 = 0.0f;
          uint32_t w[16];
          const uint32_t vb = valid_bits(j, ch, lim);
          float e[32];
          ex2_chunk(x[i], sc, -m_sc, e);
          if (vb == 0xffffffffu) {
#pragma unroll
            for (int c = 0; c < 32; c += 2) {
              l0 += e[c];
              l1 += e[c + 1];
              w[c >> 1] = F16 ? pack_f16x2(e[c], e[c + 1]) : pack_bf16x2(e[c], e[c + 1]);
            }
          } else {
#pragma unroll
            for (int c = 0; c < 32; c += 2) {
              const float p0 = ((vb >> c) & 1u) ? e[c] : 0.0f;
              const float p1 = ((vb >> (c + 1)) & 1u) ? e[c + 1] : 0.0f;
              l0 += p0;
              l1 += p1;
              w[c >> 1] = F16 ? pack_f16x2(p0, p1) : pack_bf16x2(p0, p1);
            }
          }
          l_blk += l0 + l1;
          uint8_t* atom = sPb + (ch >> 1) * ATT_Q_BYTES + r * 128;
          const int q_base = (ch & 1) * 4;
#pragma unroll
          for (int q = 0; q < 4; ++q)
            *reinterpret_cast<uint4*>(atom + (((q_base + q) ^ (r & 7)) << 4)) =
                make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
        }
      }
      l_run = l_run * alpha + l_blk;
      alpha_pend = alpha;                              // applied to o_acc when O_j is added
      fence_async_smem();                              // P (generic-proxy stores) -> visible to the tensor core
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars[PB_P + b]);
    }
    drain_O(nkv - 1);

    // ---- total row sum = the two halves' partial sums; normalise and store (as attn_tc_kernel)
    xsum[sp * 128 + r] = l_run;
    named_bar_sync(pair_bar, 64);
    const float inv_l = 1.0f / (l_run + xsum[(sp ^ 1) * 128 + r]);
    const int64_t seg_stride = p.M * (int64_t)p.d;
    uint8_t* stage = sQ + quad * 32 * 128;             // Q is retired: every S MMA has executed
    const int q_tile0 = q0 + quad * 32;
    float rem[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) rem[c] = o_acc[c] * inv_l;
    for (int sg = 0; sg < p.out_nseg; ++sg) {
#pragma unroll
      for (int v4 = 0; v4 < 4; ++v4) {
        uint32_t w[4];
        if (p.out_nseg == 1) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            w[k] = F16 ? pack_f16x2(rem[v4 * 8 + k * 2], rem[v4 * 8 + k * 2 + 1]) : pack_bf16x2(rem[v4 * 8 + k * 2], rem[v4 * 8 + k * 2 + 1]);
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
  if (warp == 8) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

// =====================================================================================================
// attn_small_kernel: one key block, N <= 136 (the 133-token joint sequence of the mvp configs), TWO CTAs per SM.
//
// attn_tc_kernel's single-block path is a latency chain (TMA round trip, S MMA, softmax, P V MMA, store: ~9700 cycles
// with the tensor pipe active 9 % of the time) and its 182 KB of shared memory, 512 TMEM columns and 512 threads at 128
// registers allow ONE CTA per SM.  Beyond one wave of (sequence, head) pairs -- A2V batch 64 is 1024 of them -- the SMs
// then run 6.9 waves of chains back to back.  This kernel fits twice on an SM (111 KB, 256 TMEM columns, 256 threads), so
// two chains overlap.  What made the footprint is the second query tile, which at N = 133 holds 5 rows: here those rows
// do not go through the tensor core at all.  Warp w computes row 128 + w on CUDA cores -- q . K^T against the swizzled K
// tile while the S MMA of the full tile is in flight, the softmax in registers, p . V against the V tile while the P V
// MMA runs -- 5 x 133 x 64 MACs twice, in fp32.
//   warps 0..7  softmax / output of the 128-row tile exactly as tile 0 of attn_tc_kernel (thread = row, the two warps of a
//               TMEM lane quadrant split the key columns); lane 0 of warp 0 issues the TMA loads and the two MMAs
// =====================================================================================================
constexpr int AS_THREADS = 256;
constexpr int AS_KV_ROWS = 144;                        // round16(136)
constexpr int AS_KV_BYTES = AS_KV_ROWS * ATT_HD * 2;   // 18 KB
constexpr int AS_MAX_THIN = 8;
constexpr int AS_XCHG_BYTES = 4 * 128 * 4;             // row max [2 halves][128], row sum [2 halves][128]
constexpr int AS_THIN_BYTES = AS_MAX_THIN * 160 * 4;   // the thin rows' probabilities [8][160] fp32
constexpr int AS_SMEM = ATT_Q_BYTES + 2 * AS_KV_BYTES + ATT_P_BYTES + 1024 + AS_XCHG_BYTES + AS_THIN_BYTES + 256;
enum { SB_Q = 0, SB_K = 1, SB_V = 2, SB_S = 3, SB_P = 4, SB_O = 5, SB_COUNT = 6 };

template <bool F16>
__device__ __forceinline__ float2 cvt16x2(uint32_t w) {
  if (F16) return __half22float2(*reinterpret_cast<const __half2*>(&w));
  return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
}

template <bool F16, bool MASKED>
__global__ void __launch_bounds__(AS_THREADS, 2) attn_small_kernel(const __grid_constant__ CUtensorMap tmQ,
                                                                   const __grid_constant__ CUtensorMap tmQ8,
                                                                   const __grid_constant__ CUtensorMap tmKV,
                                                                   const AttnParams p) {
  extern __shared__ __align__(1024) uint8_t as_smem[];
  uint8_t* smem = as_smem;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* sQ = smem;                                  // 128 rows (later: output staging)
  uint8_t* sK = sQ + ATT_Q_BYTES;
  uint8_t* sV = sK + AS_KV_BYTES;
  uint8_t* sP = sV + AS_KV_BYTES;                      // 3 atoms of [128 rows][64 keys]
  uint8_t* sQ1 = sP + ATT_P_BYTES;                     // the thin rows: [8][64] 16-bit, 128B-swizzled
  float* xmax = reinterpret_cast<float*>(sQ1 + 1024);
  float* xsum = xmax + 2 * 128;
  float* thin_p = xsum + 2 * 128;                      // [8][160]
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(thin_p) + AS_THIN_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + SB_COUNT);

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int head = blockIdx.y;
  const int seq = blockIdx.z;
  const int row_base = seq * p.N;
  const int colQ = head * ATT_HD, colK = p.d + head * ATT_HD, colV = 2 * p.d + head * ATT_HD;
  const int rows1 = p.N > ATT_BQ ? p.N - ATT_BQ : 0;   // thin rows (<= 8)
  const int kv_valid = p.N;
  const int kv_n = (kv_valid + 15) & ~15;
  const int nch = (kv_valid + 31) >> 5;                // <= 5
  const float sc = p.scale_log2e;
  const uint32_t* kmask = MASKED ? p.kmask + (size_t)blockIdx.z * p.kmask_words : nullptr;

  if (warp == 1) {
    if (lane < SB_COUNT) mbar_init(&bars[lane], lane == SB_P ? 8 : 1);
    fence_barrier_init();
  } else if (warp == 2 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmKV);
    tma_prefetch_desc(&tmQ8);
  }
  if (warp == 0) tmem_alloc<256>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();

  const int sp = warp >> 2, quad = warp & 3;
  const int r = quad * 32 + lane;
  const uint32_t tmem_S = tmem_base + ((uint32_t)(quad * 32) << 16);
  const uint32_t tmem_O = tmem_S + ATT_BKV + sp * 32;
  const uint32_t pair_bar = 1 + quad;
  const bool leader = tid == 0;
  const uint32_t hi = smem_desc_hi_sw128(1024);
  if (leader) {
    pdl_wait();
    mbar_expect_tx(&bars[SB_Q], ATT_Q_BYTES + (rows1 > 0 ? 1024 : 0));
    tma_load_3d(sQ, &tmQ, &bars[SB_Q], colQ, row_base, 0);
    if (rows1 > 0) tma_load_3d(sQ1, &tmQ8, &bars[SB_Q], colQ, row_base + ATT_BQ, 0);
    mbar_expect_tx(&bars[SB_K], AS_KV_BYTES);
    tma_load_3d(sK, &tmKV, &bars[SB_K], colK, row_base, 0);
    mbar_expect_tx(&bars[SB_V], AS_KV_BYTES);
    tma_load_3d(sV, &tmKV, &bars[SB_V], colV, row_base, 0);
  }
  __syncwarp();
  if (warp == 0) {                                     // the MMAs are issued by the converged warp (elected lane)
    mbar_wait(&bars[SB_Q], 0);
    mbar_wait(&bars[SB_K], 0);
    tc_fence_after();
    const uint32_t idesc_s = make_idesc_bf16(ATT_BQ, kv_n, 0, 0, F16);
    const uint32_t q_lo = smem_desc_lo_sw128(smem_u32(sQ), 16);
    const uint32_t k_lo = smem_desc_lo_sw128(smem_u32(sK), 16);
#pragma unroll
    for (int k = 0; k < ATT_HD / 16; ++k)
      umma_bf16_lh_elect(tmem_base, q_lo + 2 * k, k_lo + 2 * k, hi, idesc_s, k > 0 ? 1u : 0u);
    umma_commit_elect(&bars[SB_S]);
  }

  // ---- thin rows, part 1 (under the S MMA): s = q . K^T on CUDA cores, softmax in registers, p -> smem
  float thin_inv_l = 0.0f;
  const bool thin_row = warp < rows1;
  if (thin_row) {
    mbar_wait(&bars[SB_Q], 0);
    mbar_wait(&bars[SB_K], 0);
    float2 q[32];                                      // the row's 64 query values
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const uint4 u = *reinterpret_cast<const uint4*>(sQ1 + warp * 128 + ((c ^ (warp & 7)) << 4));
      q[4 * c] = cvt16x2<F16>(u.x); q[4 * c + 1] = cvt16x2<F16>(u.y);
      q[4 * c + 2] = cvt16x2<F16>(u.z); q[4 * c + 3] = cvt16x2<F16>(u.w);
    }
    float sv[5];
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
      const int k = lane + 32 * i;
      float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f, a3 = 0.0f;   // four independent chains: the dot product is latency, not issue, bound
      bool ok = k < kv_valid;
      if (MASKED && ok) ok = !((kmask[k >> 5] >> (k & 31)) & 1u);
      if (k < kv_n) {
        const uint8_t* kr = sK + k * 128;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const uint4 u = *reinterpret_cast<const uint4*>(kr + ((c ^ (k & 7)) << 4));
          const float2 a = cvt16x2<F16>(u.x), b = cvt16x2<F16>(u.y), c2 = cvt16x2<F16>(u.z), d = cvt16x2<F16>(u.w);
          a0 = fmaf(q[4 * c].x, a.x, a0); a1 = fmaf(q[4 * c].y, a.y, a1);
          a2 = fmaf(q[4 * c + 1].x, b.x, a2); a3 = fmaf(q[4 * c + 1].y, b.y, a3);
          a0 = fmaf(q[4 * c + 2].x, c2.x, a0); a1 = fmaf(q[4 * c + 2].y, c2.y, a1);
          a2 = fmaf(q[4 * c + 3].x, d.x, a2); a3 = fmaf(q[4 * c + 3].y, d.y, a3);
        }
      }
      const float acc = (a0 + a1) + (a2 + a3);
      sv[i] = ok ? acc : -INFINITY;
      mx = fmaxf(mx, sv[i]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    const float m_use = mx == -INFINITY ? 0.0f : mx;
    float l = 0.0f;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
      const int k = lane + 32 * i;
      const float pv = sv[i] == -INFINITY ? 0.0f : ex2_approx((sv[i] - m_use) * sc);
      l += pv;
      thin_p[warp * 160 + k] = pv;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
    thin_inv_l = 1.0f / l;
    __syncwarp();
  }

  // ---- the 128-row tile: single-pass softmax of this thread's chunks (as attn_tc_kernel's one-block path)
  mbar_wait(&bars[SB_S], 0);
  tc_fence_after();
  float l_blk = 0.0f;
  {
    const int ch0 = sp * 3;
    auto valid_bits = [&](int ch, int lim) -> uint32_t {
      uint32_t v = lim >= 32 ? 0xffffffffu : (lim > 0 ? ((1u << lim) - 1u) : 0u);
      if (MASKED && lim > 0) v &= ~kmask[ch];
      return v;
    };
    uint32_t x[3][32];
#pragma unroll
    for (int i = 0; i < 3; ++i)
      if (i == 0 || ch0 + i < nch) tmem_ld_32x32(tmem_S + (ch0 + i) * 32, x[i]);
    tmem_ld_wait();
    float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const uint32_t vb = valid_bits(ch0 + i, kv_valid - (ch0 + i) * 32);
      if (vb == 0xffffffffu) {
#pragma unroll
        for (int c = 0; c < 32; ++c) mx[c & 3] = fmaxf(mx[c & 3], __uint_as_float(x[i][c]));
      } else if (vb != 0u) {
#pragma unroll
        for (int c = 0; c < 32; ++c)
          mx[c & 3] = fmaxf(mx[c & 3], ((vb >> c) & 1u) ? __uint_as_float(x[i][c]) : -INFINITY);
      }
    }
    const float m_part = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
    xmax[sp * 128 + r] = m_part;
    named_bar_sync(pair_bar, 64);
    const float m_new = fmaxf(m_part, xmax[(sp ^ 1) * 128 + r]);
    const float m_sc = (m_new == -INFINITY ? 0.0f : m_new) * sc;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int ch = ch0 + i;
      const int lim = kv_valid - ch * 32;
      if (lim > -16) {
        float l0 = 0.0f, l1 = 0.0f;
        uint32_t w[16];
        const uint32_t vb = valid_bits(ch, lim);
        float e[32];
        ex2_chunk(x[i], sc, -m_sc, e);
        if (vb == 0xffffffffu) {
#pragma unroll
          for (int c = 0; c < 32; c += 2) {
            l0 += e[c];
            l1 += e[c + 1];
            w[c >> 1] = F16 ? pack_f16x2(e[c], e[c + 1]) : pack_bf16x2(e[c], e[c + 1]);
          }
        } else {
#pragma unroll
          for (int c = 0; c < 32; c += 2) {
            const float p0 = ((vb >> c) & 1u) ? e[c] : 0.0f;
            const float p1 = ((vb >> (c + 1)) & 1u) ? e[c + 1] : 0.0f;
            l0 += p0;
            l1 += p1;
            w[c >> 1] = F16 ? pack_f16x2(p0, p1) : pack_bf16x2(p0, p1);
          }
        }
        l_blk += l0 + l1;
        uint8_t* atom = sP + (ch >> 1) * ATT_Q_BYTES + r * 128;
        const int q_base = (ch & 1) * 4;
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4)
          *reinterpret_cast<uint4*>(atom + (((q_base + q4) ^ (r & 7)) << 4)) =
              make_uint4(w[4 * q4], w[4 * q4 + 1], w[4 * q4 + 2], w[4 * q4 + 3]);
      }
    }
  }
  fence_async_smem();
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(&bars[SB_P]);
  if (warp == 0) {                                     // O = P V
    mbar_wait(&bars[SB_P], 0);
    mbar_wait(&bars[SB_V], 0);
    tc_fence_after();
    const uint32_t idesc_o = make_idesc_bf16(ATT_BQ, ATT_HD, 0, 1, F16);   // B = V is MN-major
    const uint32_t p_lo = smem_desc_lo_sw128(smem_u32(sP), 16);
    const uint32_t v_lo = smem_desc_lo_sw128(smem_u32(sV), 16);
    const int ksteps = kv_n >> 4;
    for (int ks = 0; ks < ksteps; ++ks) {
      const uint32_t pa = p_lo + (((ks >> 2) * ATT_Q_BYTES + (ks & 3) * 32) >> 4);
      const uint32_t va = v_lo + ((ks * 16 * 128) >> 4);
      umma_bf16_lh_elect(tmem_base + ATT_BKV, pa, va, hi, idesc_o, ks > 0 ? 1u : 0u);
    }
    umma_commit_elect(&bars[SB_O]);
  }
  __syncwarp();

  // ---- thin rows, part 2 (under the P V MMA): o = p . V, two output columns per lane, straight to global
  if (thin_row) {
    mbar_wait(&bars[SB_V], 0);
    float oa[4] = {0.0f, 0.0f, 0.0f, 0.0f}, ob[4] = {0.0f, 0.0f, 0.0f, 0.0f};    // four independent chains over the keys
    const float* pr = thin_p + warp * 160;                 // (entries past kv_valid are zero probabilities)
    for (int k = 0; k < kv_n; k += 4) {
#pragma unroll
      for (int u4 = 0; u4 < 4; ++u4) {
        const int kk = k + u4;
        const float pk = pr[kk];
        const uint32_t u = *reinterpret_cast<const uint32_t*>(sV + kk * 128 + (((lane >> 2) ^ (kk & 7)) << 4) + (lane & 3) * 4);
        const float2 v2 = cvt16x2<F16>(u);
        oa[u4] = fmaf(pk, v2.x, oa[u4]);
        ob[u4] = fmaf(pk, v2.y, ob[u4]);
      }
    }
    float o0 = ((oa[0] + oa[1]) + (oa[2] + oa[3])) * thin_inv_l;
    float o1 = ((ob[0] + ob[1]) + (ob[2] + ob[3])) * thin_inv_l;
    const int64_t seg_stride = p.M * (int64_t)p.d;
    const int64_t off = ((int64_t)row_base + ATT_BQ + warp) * p.d + head * ATT_HD + 2 * lane;
    for (int sg = 0; sg < p.out_nseg; ++sg) {
      uint32_t w;
      if (p.out_nseg == 1) {
        w = F16 ? pack_f16x2(o0, o1) : pack_bf16x2(o0, o1);
      } else {
        const __nv_bfloat16 h0 = __float2bfloat16_rn(o0), h1 = __float2bfloat16_rn(o1);
        o0 -= __bfloat162float(h0);
        o1 -= __bfloat162float(h1);
        w = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
      }
      *reinterpret_cast<uint32_t*>(p.out + sg * seg_stride + off) = w;
    }
  }

  // ---- O of the 128-row tile: normalise, stage, store
  mbar_wait(&bars[SB_O], 0);
  tc_fence_after();
  float o_acc[32];
  {
    uint32_t xo[32];
    tmem_ld_32x32(tmem_O, xo);
    tmem_ld_wait();
#pragma unroll
    for (int c = 0; c < 32; ++c) o_acc[c] = __uint_as_float(xo[c]);
  }
  xsum[sp * 128 + r] = l_blk;
  named_bar_sync(pair_bar, 64);
  const float inv_l = 1.0f / (l_blk + xsum[(sp ^ 1) * 128 + r]);
  {
    const int64_t seg_stride = p.M * (int64_t)p.d;
    uint8_t* stage = sQ + quad * 32 * 128;             // Q is retired: the S MMA has executed
    const int q_tile0 = quad * 32;
    float rem[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) rem[c] = o_acc[c] * inv_l;
    for (int sg = 0; sg < p.out_nseg; ++sg) {
#pragma unroll
      for (int v4 = 0; v4 < 4; ++v4) {
        uint32_t w[4];
        if (p.out_nseg == 1) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            w[k] = F16 ? pack_f16x2(rem[v4 * 8 + k * 2], rem[v4 * 8 + k * 2 + 1]) : pack_bf16x2(rem[v4 * 8 + k * 2], rem[v4 * 8 + k * 2 + 1]);
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
    tmem_dealloc<256>(tmem_base);
  }
}

int launch_attention_tc(const void* qkv_bf16, void* out, int out_nseg, int n_seq, int N, int n_heads, int hd,
                        cudaStream_t s, bool f16, const uint32_t* kmask) {
  AVD_REQUIRE(hd == ATT_HD, "tcgen05 attention kernel is built for head_dim 64");
  AVD_REQUIRE(n_seq > 0 && N > 0 && n_heads > 0, "empty attention");
  AVD_REQUIRE(out_nseg >= 1 && out_nseg <= 3, "bad output precision");
  const int d = n_heads * hd;
  const int64_t M = (int64_t)n_seq * N;
  CUtensorMap tmQ, tmKV;
  AVD_TRY(get_tensor_map_3d(qkv_bf16, 2, (uint64_t)3 * d, (uint64_t)M, 1, (uint64_t)3 * d * 2, (uint64_t)M * 3 * d * 2,
                            ATT_HD, ATT_BQ, &tmQ));
  AVD_TRY(get_tensor_map_3d(qkv_bf16, 2, (uint64_t)3 * d, (uint64_t)M, 1, (uint64_t)3 * d * 2, (uint64_t)M * 3 * d * 2,
                            ATT_HD, ATT_BKV, &tmKV));
  using AttnFn = void (*)(const CUtensorMap, const CUtensorMap, const AttnParams);
  // [one_block][f16][masked]
  static const AttnFn kernels[2][2][2] = {
      {{attn_tc_kernel<false, false, false>, attn_tc_kernel<false, false, true>},
       {attn_tc_kernel<false, true, false>, attn_tc_kernel<false, true, true>}},
      {{attn_tc_kernel<true, false, false>, attn_tc_kernel<true, false, true>},
       {attn_tc_kernel<true, true, false>, attn_tc_kernel<true, true, true>}}};
  AttnParams p{};
  p.N = N; p.d = d; p.n_heads = n_heads; p.M = M;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.out_nseg = out_nseg;
  p.scale_log2e = (1.0f / sqrtf((float)hd)) * 1.4426950408889634f;
  p.kmask = kmask;
  p.kmask_words = (N + 31) / 32;
  p.poly_exp = options().attn_poly;      // measured 12 % slower on B200: off unless asked for
  {
    const int64_t n_ctas = (int64_t)((N + 2 * ATT_BQ - 1) / (2 * ATT_BQ)) * n_heads * n_seq;
    p.trace = (g_dbg_trace != nullptr && n_ctas * 32 <= g_dbg_trace_slots) ? g_dbg_trace : nullptr;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((N + 2 * ATT_BQ - 1) / (2 * ATT_BQ), n_heads, n_seq);
  cfg.blockDim = dim3(ATT_THREADS);
  cfg.dynamicSmemBytes = ATT_SMEM;
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
  if (kind_skipped(KIND_ATTN_TC)) return 0;
  LaunchScope ls(KIND_ATTN_TC, s);
  AVD_REQUIRE(!f16 || out_nseg == 1, "the fp16 attention output has one segment");
  // single block, at most 8 rows in the second query tile, more (sequence, head) pairs than SMs: two CTAs per SM
  const int small_opt = options().attn_small;
  int sms = 148;
  {
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  if (small_opt != 0 && N <= ATT_BQ + AS_MAX_THIN && ((int64_t)n_heads * n_seq > sms || small_opt == 2)) {
    using SmallFn = void (*)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const AttnParams);
    static const SmallFn small_kernels[2][2] = {{attn_small_kernel<false, false>, attn_small_kernel<false, true>},
                                                {attn_small_kernel<true, false>, attn_small_kernel<true, true>}};
    const SmallFn sf = small_kernels[f16 ? 1 : 0][kmask != nullptr ? 1 : 0];
    CUtensorMap tmQ8, tmKVs;
    AVD_TRY(get_tensor_map_3d(qkv_bf16, 2, (uint64_t)3 * d, (uint64_t)M, 1, (uint64_t)3 * d * 2, (uint64_t)M * 3 * d * 2,
                              ATT_HD, AS_MAX_THIN, &tmQ8));
    AVD_TRY(get_tensor_map_3d(qkv_bf16, 2, (uint64_t)3 * d, (uint64_t)M, 1, (uint64_t)3 * d * 2, (uint64_t)M * 3 * d * 2,
                              ATT_HD, AS_KV_ROWS, &tmKVs));
    AVD_TRY(ensure_dyn_smem(reinterpret_cast<const void*>(sf), AS_SMEM));
    cfg.gridDim = dim3(1, n_heads, n_seq);
    cfg.blockDim = dim3(AS_THREADS);
    cfg.dynamicSmemBytes = AS_SMEM;
    p.trace = nullptr;
    AVD_CUDA(cudaLaunchKernelEx(&cfg, sf, tmQ, tmQ8, tmKVs, p));
    return 0;
  }
  if (N > ATT_BKV && options().attn_fa != 0 && options().attn_pipe == 0) {   // multi-block: P and O resident in tensor memory (avd_attn_fa.cu)
    const int rc = launch_attention_fa(qkv_bf16, p, n_seq, f16, s);
    if (rc != 0) return rc < 0 ? rc : 0;
  }
  if (N > ATT_BKV && options().attn_pipe != 0) {     // multi-block: the pipelined kernel, one 128-query tile per CTA
    static const AttnFn pipe_kernels[2][2] = {{attn_pipe_kernel<false, false>, attn_pipe_kernel<false, true>},
                                              {attn_pipe_kernel<true, false>, attn_pipe_kernel<true, true>}};
    const AttnFn pf = pipe_kernels[f16 ? 1 : 0][kmask != nullptr ? 1 : 0];
    AVD_TRY(ensure_dyn_smem(reinterpret_cast<const void*>(pf), AP_SMEM));
    cfg.gridDim = dim3((N + ATT_BQ - 1) / ATT_BQ, n_heads, n_seq);
    cfg.blockDim = dim3(AP_THREADS);
    cfg.dynamicSmemBytes = AP_SMEM;
    p.trace = nullptr;
    AVD_CUDA(cudaLaunchKernelEx(&cfg, pf, tmQ, tmKV, p));
    return 0;
  }
  const AttnFn fn = kernels[(N <= ATT_BKV && options().attn_one) ? 1 : 0][f16 ? 1 : 0][kmask != nullptr ? 1 : 0];
  AVD_TRY(ensure_dyn_smem(reinterpret_cast<const void*>(fn), ATT_SMEM));   // once per device
  AVD_CUDA(cudaLaunchKernelEx(&cfg, fn, tmQ, tmKV, p));
  return 0;
}


// =====================================================================================================
// fp32 CUDA-core attention (parity mode): one warp per query row.
// =====================================================================================================
constexpr int AF_WARPS = 4;
constexpr int AF_ROWS = 16;   // query rows per CTA

__global__ void __launch_bounds__(AF_WARPS * 32) attn_f32_kernel(const float* __restrict__ qkv,
                                                                 __nv_bfloat16* __restrict__ out, int out_nseg, int N,
                                                                 int d, int hd, int64_t M, float scale,
                                                                 const uint32_t* __restrict__ kmask, int kmask_words) {
  extern __shared__ float af_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* sc = af_smem + (size_t)warp * N;                 // scores of this warp's current row
  float* qs = af_smem + (size_t)AF_WARPS * N + warp * hd; // its query vector
  const int head = blockIdx.y, seq = blockIdx.z;
  const int64_t row0 = (int64_t)seq * N;
  const int ld = 3 * d;
  const float* Kp = qkv + row0 * ld + d + head * hd;
  const float* Vp = qkv + row0 * ld + 2 * d + head * hd;
  const int64_t seg_stride = M * (int64_t)d;

  for (int qi = warp; qi < AF_ROWS; qi += AF_WARPS) {
    const int q = blockIdx.x * AF_ROWS + qi;
    if (q >= N) break;
    const float* Qp = qkv + (row0 + q) * ld + head * hd;
    for (int c = lane; c < hd; c += 32) qs[c] = Qp[c];
    __syncwarp();
    float mx = -INFINITY;
    for (int j = lane; j < N; j += 32) {
      const float* kr = Kp + (int64_t)j * ld;
      float acc = 0.0f;
      for (int c = 0; c < hd; c += 4) {
        float4 k4 = *reinterpret_cast<const float4*>(kr + c);
        acc = fmaf(qs[c], k4.x, acc); acc = fmaf(qs[c + 1], k4.y, acc);
        acc = fmaf(qs[c + 2], k4.z, acc); acc = fmaf(qs[c + 3], k4.w, acc);
      }
      acc *= scale;
      if (kmask != nullptr && ((kmask[(size_t)seq * kmask_words + (j >> 5)] >> (j & 31)) & 1u)) acc = -INFINITY;
      sc[j] = acc;                               // padded key: exp(-inf - mx) = 0 below
      mx = fmaxf(mx, acc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.0f;
    for (int j = lane; j < N; j += 32) {
      float e = expf(sc[j] - mx);
      sc[j] = e;
      sum += e;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    __syncwarp();
    const float inv = 1.0f / sum;
    for (int c = lane; c < hd; c += 32) {
      float acc = 0.0f;
      for (int j = 0; j < N; ++j) acc = fmaf(sc[j], Vp[(int64_t)j * ld + c], acc);
      float v = acc * inv;
      int64_t idx = (row0 + q) * d + head * hd + c;
      __nv_bfloat16 h0 = __float2bfloat16_rn(v);
      out[idx] = h0;
      if (out_nseg >= 2) {
        float r1 = v - __bfloat162float(h0);
        __nv_bfloat16 h1 = __float2bfloat16_rn(r1);
        out[seg_stride + idx] = h1;
        if (out_nseg >= 3) out[2 * seg_stride + idx] = __float2bfloat16_rn(r1 - __bfloat162float(h1));
      }
    }
    __syncwarp();
  }
}

int launch_attention_f32(const float* qkv, void* out, int out_nseg, int n_seq, int N, int n_heads, int hd,
                         cudaStream_t s, const uint32_t* kmask) {
  AVD_REQUIRE(hd % 4 == 0 && hd <= 256, "head_dim must be a multiple of 4, <= 256");
  AVD_REQUIRE(n_seq > 0 && N > 0 && n_heads > 0, "empty attention");
  const int d = n_heads * hd;
  size_t smem = ((size_t)AF_WARPS * N + AF_WARPS * hd) * sizeof(float);
  AVD_REQUIRE(smem <= 200 * 1024, "sequence too long for the fp32 parity attention kernel");
  if (smem > 48 * 1024) {
    AVD_CUDA(cudaFuncSetAttribute(attn_f32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  dim3 grid((N + AF_ROWS - 1) / AF_ROWS, n_heads, n_seq);
  if (kind_skipped(KIND_ATTN_F32)) return 0;
  LaunchScope ls(KIND_ATTN_F32, s);
  attn_f32_kernel<<<grid, AF_WARPS * 32, smem, s>>>(qkv, static_cast<__nv_bfloat16*>(out), out_nseg, N, d, hd,
                                                    (int64_t)n_seq * N, 1.0f / sqrtf((float)hd), kmask, (N + 31) / 32);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace avd
