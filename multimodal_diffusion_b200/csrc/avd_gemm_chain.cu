// avd_gemm_chain.cu -- K4, chained form: up to four DEPENDENT GEMMs of one MMDT block in ONE persistent launch
//
//   out-proj (+residual, row sums)  ->  fc1 (+GELU)  ->  fc2 (+residual, row sums)  ->  QKV of the next block
//
// (mmdt.py:96-99 with the attention call between QKV and out-proj left as its own kernel).  At the sampler's shapes
// (M = 2B x 133 = 2128 rows, 136 tiles per GEMM on 148 SMs) each of these GEMMs is one wave of 128-row tiles, and a
// kernel boundary between two of them costs the launch gap, the CTA setup (barrier init, TMEM allocation, cold
// instruction fetch: ~0.4 us) and the wait for ALL 136 CTAs of the predecessor although a tile only needs the 8 tiles of
// its own 128-row group (profiles/r2*_kbench_phase_trace.txt).  Here one CTA per SM keeps its barriers, its TMEM
// allocation and its tensor maps across the phases and runs ONE tile of every phase:
//
//   * phase p tile (r, c) needs every tile (r, *) of phase p-1 (its K dimension is their N dimension).  The producing
//     CTAs count themselves into flag[p-1][r] (release) after their epilogue's stores; the consuming CTA's producer
//     thread polls that counter (acquire), crosses into the async proxy (fence.proxy.async) and only then issues the TMA
//     loads of the activation tiles.  Waiters check out of the counter and the last one resets it, so the flags are
//     zero again when the launch ends (they are zeroed once, when the workspace is created).
//   * the WEIGHT tiles of phase p do not depend on anything: they are fetched into the ring while the epilogue of phase
//     p-1 is still running and while the flag is polled -- that is most of what the boundary used to cost.
//
// Roles (384 threads, as in gemm_ps_kernel): warps 0,1 TMA producers, warps 2,3 tcgen05.mma issuers (alternating
// k-blocks into two accumulators, or column-split halves of one 256-wide accumulator), warps 4..11 epilogue.  The ring
// is re-partitioned per phase (its stage size depends on the tile width); slot s and its two mbarriers always belong to
// the producer / issuer pair s & 1, and each of them keeps one parity bit per slot, so a barrier's uses are seen in
// order by its one waiter whatever the ring depth of the phase is.
//
// All CTAs of the launch must be co-resident (grid <= SM count, one CTA per SM by its shared memory) -- the launcher
// guarantees it; every spin is bounded and leaves a host-visible record before trapping (avd_check_errors).
#include "avd_gemm_common.cuh"

namespace avd {

constexpr int CH_THREADS = 384;
constexpr int CH_MAX_STAGES = 8;
constexpr int CH_MAX_PHASES = 4;
constexpr int CH_A_BYTES = GEMM_BM * GEMM_BK * 2;
constexpr int CH_F32_BYTES = 16 * GEMM_BOX;            // 8 epilogue warps x 2 fp32 boxes (residual in, result out)
constexpr int CH_BF_BYTES = 4 * GEMM_BOX;              // one 16-bit box per TMEM lane quadrant
constexpr int CH_TAIL_BYTES = 2048;
constexpr int CH_SMEM = 231424;                        // 226 KB
constexpr int CH_RING_SAFE = CH_SMEM - CH_F32_BYTES - CH_BF_BYTES - CH_TAIL_BYTES;   // 144 KB never used for staging
constexpr int CH_RING_WIDE = CH_RING_SAFE + CH_F32_BYTES;   // ring of a phase whose epilogue needs no fp32 boxes

struct alignas(64) ChainPhase {
  CUtensorMap tmA, tmB, tmR;
  GemmParams p;
  int tiles_n, tiles_total;
  int stages;               // ring depth of this phase (even)
  int safe_slots;           // slots that end inside CH_RING_SAFE: usable while the previous phase's epilogue still stages
  unsigned int* wait_flag;  // [row tile] counters of the phase this one consumes (nullptr: the previous KERNEL, via PDL)
  int wait_target;          // tiles per row group of the producing phase
  int wait_total;           // wait_target + this phase's tiles per row group: the last waiter to check out resets
  unsigned int* signal_flag;   // [row tile] counters this phase's tiles count themselves into (nullptr: last phase)
};
struct ChainArgs {
  ChainPhase ph[CH_MAX_PHASES];
  unsigned int* dbg;        // host-visible timeout records (shared with the persistent kernel)
  unsigned long long* trace;   // debug-hooks builds: [cta][phase][16] clock64 stamps (avd_debug_trace), else nullptr
};

struct ChainBars {
  uint64_t* full;       // [8]  count 1 (expect_tx)
  uint64_t* empty;      // [8]  count 2 (both issuers in a column-split phase; commit + plain arrive otherwise)
  uint64_t* mma_done;   // count 2: both issuers, after their last MMA of a phase
  uint64_t* epi_done;   // count 8: the epilogue warps, after their last staging / TMEM access of a phase
  uint64_t* dep;        // count 1: producer 0 has seen the phase's dependency flag
  uint64_t* res;        // [16] residual boxes, two per epilogue warp
};

struct ChainCtx {
  uint8_t* smem;
  ChainBars b;
  unsigned int* dbg;
  uint32_t tmem_base;
  unsigned long long* trace;
};
// stamps: 0 producer enters the phase, 1 first-pass weights issued, 2 dependency seen, 3 last load issued,
//         4 epilogue: accumulator complete, 5 epilogue: stores issued, 6 signalled; first chunk of warp 4: 7 tcgen05.ld
//         done, 8 residual added, 9 staged, 10 pair barrier passed, 11 fp32 stores issued, 12 16-bit stores issued
__device__ __forceinline__ void chain_mark(const ChainCtx& c, int phase, int slot) {
#ifdef AVD_DEBUG_HOOKS
  if (c.trace != nullptr) {
    unsigned long long g;
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(g));   // per-SM cycle counter: deltas inside one CTA are exact
    c.trace[((size_t)blockIdx.x * CH_MAX_PHASES + phase) * 16 + slot] = g;
  }
#endif
}

__device__ __forceinline__ void chain_timeout(const ChainCtx& c, unsigned tag, int phase, uint32_t parity) {
  unsigned int* dbg = c.dbg;
  const unsigned idx = atomicAdd_system(&dbg[0], 1u);
  if (idx < 30) {
    unsigned int* r = dbg + 8 + idx * 8;
    r[0] = 100u + tag; r[1] = blockIdx.x; r[2] = threadIdx.x; r[3] = (unsigned)phase;
    r[4] = parity; r[5] = 0u; r[6] = 0u; r[7] = gridDim.x * 1000u;
    __threadfence_system();
  }
  __nanosleep(4000000);
  __trap();
}
__device__ __forceinline__ void chain_wait(const ChainCtx& c, uint64_t* bar, uint32_t parity, unsigned tag, int phase) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity))
    if (++spins > (1u << 26)) chain_timeout(c, tag, phase, parity);
}
// Phase-boundary waits last microseconds while the epilogue warps of the same SM sub-partitions are at work: a bare
// try_wait loop floods the shared-memory pipe the epilogue's LDS / STS go through (the first trace of this kernel showed
// 2000 cycles for 8 LDS + 32 FADD).  These waits back off between polls.
__device__ __forceinline__ void chain_wait_idle(const ChainCtx& c, uint64_t* bar, uint32_t parity, unsigned tag, int phase) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    __nanosleep(128);
    if (++spins > (1u << 24)) chain_timeout(c, tag, phase, parity);
  }
}
__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// per-role bookkeeping that survives the phases
struct ChainState {
  // Ring bookkeeping that EVERY role thread derives from the phase parameters alone (never from what it happened to wait
  // on): a slot is consumed by issuer s & 1 in an alternating phase but by both issuers in a column-split phase, so a
  // thread's own wait history says nothing about a barrier's phase.  fills: parity of the number of times slot s has
  // been filled in the phases before the current one; used: slots filled at least once before the current phase.
  uint32_t fills = 0;
  uint32_t used = 0;
  int n_done = 0;       // phases this CTA has executed so far (parity of mma_done / epi_done)
  int n_dep = 0;        // executed phases that waited on a flag (parity of dep)
  int qr = 0;           // epilogue: residual boxes consumed so far (parity of res[])
  // a phase of ipt k-block iterations on a ring of `stages` slots has filled slot s  ipt / stages + (s < ipt % stages)  times
  __device__ __forceinline__ void ring_advance(int ipt, int stages) {
    const int q = ipt / stages, rem = ipt - q * stages;
    for (int sl = 0; sl < stages; ++sl) {
      const int n = q + (sl < rem ? 1 : 0);
      if (n & 1) fills ^= 1u << sl;
      if (n > 0) used |= 1u << sl;
    }
  }
};

// ------------------------------------------------------------------------------------------------ producer
template <int BN>
__device__ __forceinline__ void chain_produce(const ChainPhase& ph, int pidx, int h, const ChainCtx& c, ChainState& st) {
  constexpr int STAGE_BYTES = CH_A_BYTES + BN * GEMM_BK * 2;
  if ((int)blockIdx.x >= ph.tiles_total) return;
  const int t = blockIdx.x;
  const int mt = t / ph.tiles_n;
  const int m0 = mt * GEMM_BM, n0 = (t - mt * ph.tiles_n) * BN;
  const int kblocks = ph.p.K / GEMM_BK;
  const int ipt = kblocks * ph.p.n_pairs;
  const int stages = ph.stages;
  const uint32_t pairs = ph.p.pairs;
  uint8_t* const smem = c.smem;
  if (h == 0) chain_mark(c, pidx, 0);
  // the ring is laid out anew for this tile width: every MMA of the previous phase must have read its operands
  if (st.n_done > 0) chain_wait_idle(c, c.b.mma_done, (st.n_done - 1) & 1, 1, pidx);
  bool epi_waited = st.n_done == 0;
  auto claim = [&](int slot, int pass) {         // fill number `pass` of this slot in this phase: its last use has retired
    if (!epi_waited && slot >= ph.safe_slots) {  // this slot overlaps the fp32 staging boxes of the previous epilogue
      chain_wait_idle(c, c.b.epi_done, (st.n_done - 1) & 1, 2, pidx);
      epi_waited = true;
    }
    if (pass > 0 || ((st.used >> slot) & 1u))    // empty[slot] has completed (fills so far) phases once that fill is consumed
      chain_wait(c, &c.b.empty[slot], (((st.fills >> slot) & 1u) + (uint32_t)pass - 1u) & 1u, 3, pidx);
  };
  const int pre_end = ipt < stages ? ipt : stages;
  {
    int kb = h, pi = 0;
    while (kb >= kblocks) { kb -= kblocks; ++pi; }
    for (int j = h; j < pre_end; j += 2) {       // weights of the first ring pass: no dependency
      claim(j, 0);
      mbar_expect_tx(&c.b.full[j], STAGE_BYTES);
      tma_load_3d(smem + j * STAGE_BYTES + CH_A_BYTES, &ph.tmB, &c.b.full[j], kb * GEMM_BK, n0, (pairs >> (4 * pi)) & 3);
      kb += 2;
      while (kb >= kblocks) { kb -= kblocks; ++pi; }
    }
  }
  if (h == 0) chain_mark(c, pidx, 1);
  // dependency of the activation tiles
  if (ph.wait_flag == nullptr) {
    pdl_wait();
  } else {
    if (h == 0) {
      const unsigned int* f = ph.wait_flag + mt;
      uint32_t spins = 0;
      while (ld_acquire_gpu(f) < (unsigned)ph.wait_target) {
        __nanosleep(64);
        if (++spins > (1u << 22)) chain_timeout(c, 4, pidx, 0);
      }
      mbar_arrive(c.b.dep);                      // the other waiters of this CTA (producer 1, epilogue warps)
      const unsigned old = atomicAdd(ph.wait_flag + mt, 1u);     // check out; the last waiter leaves the flag zero
      if (old + 1u == (unsigned)ph.wait_total) atomicExch(ph.wait_flag + mt, 0u);
    } else {
      chain_wait_idle(c, c.b.dep, st.n_dep & 1, 5, pidx);
    }
    fence_proxy_async_all();                     // the tiles were written through the generic proxy; TMA reads them
  }
  if (h == 0) chain_mark(c, pidx, 2);
  int kb = h, pi = 0;
  while (kb >= kblocks) { kb -= kblocks; ++pi; }
  int j = h;
  for (; j < pre_end; j += 2) {
    tma_load_3d(smem + j * STAGE_BYTES, &ph.tmA, &c.b.full[j], kb * GEMM_BK, m0, (pairs >> (4 * pi + 2)) & 3);
    kb += 2;
    while (kb >= kblocks) { kb -= kblocks; ++pi; }
  }
  int slot = j % stages, pass = j / stages;
  for (; j < ipt; j += 2) {
    claim(slot, pass);
    const uint32_t pr = pairs >> (4 * pi);
    uint8_t* dst = smem + slot * STAGE_BYTES;
    mbar_expect_tx(&c.b.full[slot], STAGE_BYTES);
    tma_load_3d(dst, &ph.tmA, &c.b.full[slot], kb * GEMM_BK, m0, (pr >> 2) & 3);
    tma_load_3d(dst + CH_A_BYTES, &ph.tmB, &c.b.full[slot], kb * GEMM_BK, n0, pr & 3);
    kb += 2;
    while (kb >= kblocks) { kb -= kblocks; ++pi; }
    slot += 2;
    if (slot >= stages) { slot -= stages; ++pass; }
  }
  if (h == 0) chain_mark(c, pidx, 3);
  st.ring_advance(ipt, stages);
  ++st.n_done;
  if (ph.wait_flag != nullptr) ++st.n_dep;
}

// ------------------------------------------------------------------------------------------------ MMA issuer
template <int BN>
__device__ __forceinline__ void chain_issue(const ChainPhase& ph, int pidx, int h, const ChainCtx& c, ChainState& st) {
  constexpr bool CS = BN == 256;
  constexpr int STAGE_BYTES = CH_A_BYTES + BN * GEMM_BK * 2;
  static_assert(CS || 2 * BN <= 512, "two accumulators of BN columns must fit the TMEM allocation");
  if ((int)blockIdx.x >= ph.tiles_total) return;
  const int kblocks = ph.p.K / GEMM_BK;
  const int ipt = kblocks * ph.p.n_pairs;
  const int stages = ph.stages;
  const uint32_t idesc = make_idesc_bf16(GEMM_BM, CS ? 128 : BN, 0, 0, ph.p.f16_in);
  const uint32_t desc_hi = smem_desc_hi_sw128(1024);
  const uint32_t a_lo0 = smem_desc_lo_sw128(smem_u32(c.smem), 16);
  // the accumulator columns are the ones the previous phase's epilogue reads
  if (st.n_done > 0) chain_wait_idle(c, c.b.epi_done, (st.n_done - 1) & 1, 6, pidx);
  tc_fence_after();
  const uint32_t acc = c.tmem_base + (CS ? h * 128 : h * BN);
  bool any = false;
  int slot = CS ? 0 : h, pass = 0;
  for (int j = CS ? 0 : h; j < ipt; j += CS ? 1 : 2) {
    // (the first k-block of a phase arrives only after the row group's flag: an idle wait; the rest are on the critical path)
    if (!any) chain_wait_idle(c, &c.b.full[slot], (((st.fills >> slot) & 1u) + (uint32_t)pass) & 1u, 7, pidx);
    else chain_wait(c, &c.b.full[slot], (((st.fills >> slot) & 1u) + (uint32_t)pass) & 1u, 7, pidx);
    tc_fence_after();
    const uint32_t a_lo = a_lo0 + slot * (STAGE_BYTES >> 4);
    const uint32_t b_lo = a_lo + (CH_A_BYTES >> 4) + (CS ? h * ((128 * GEMM_BK * 2) >> 4) : 0);
#pragma unroll
    for (int k = 0; k < GEMM_BK / 16; ++k)
      umma_bf16_lh(acc, a_lo + 2 * k, b_lo + 2 * k, desc_hi, idesc, (any || k > 0) ? 1u : 0u);
    any = true;
    umma_commit(&c.b.empty[slot]);               // slot reusable once these MMAs have read it ...
    if (!CS) mbar_arrive(&c.b.empty[slot]);      // ... (the barrier counts two arrivals: both issuers when column-split)
    slot += CS ? 1 : 2;
    if (slot >= stages) { slot -= stages; ++pass; }
  }
  if (any) umma_commit(c.b.mma_done); else mbar_arrive(c.b.mma_done);
  st.ring_advance(ipt, stages);
  ++st.n_done;
}

// ------------------------------------------------------------------------------------------------ epilogue
// MODE as in gemm_tc_kernel: 1 one 16-bit segment; 2 the same with the packed GELU; 3 residual + fp32 + one 16-bit
// segment + row sums of squares.
template <int BN, int MODE>
__device__ __forceinline__ void chain_epilogue(const ChainPhase& ph, int pidx, const ChainCtx& c, ChainState& st) {
  constexpr bool CS = BN == 256;
  constexpr int NCH = BN / 32;
  static_assert(MODE >= 1 && MODE <= 3 && BN >= 64, "chained phases use the specialised epilogues");
  if ((int)blockIdx.x >= ph.tiles_total) return;
  const GemmParams& p = ph.p;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int e = warp - 4;
  const int quad = warp & 3;
  const int hf = e >> 2;
  constexpr bool has_resid = MODE == 3, has_f32 = MODE == 3, has_ssq = MODE == 3;
  const bool f16_out = p.f16_out != 0;
  const float out_scale = p.out_scale;
  float* const ssq_out = p.ssq_out;
  uint8_t* const smem = c.smem;
  uint8_t* const fbox0 = smem + CH_RING_SAFE + e * 2 * GEMM_BOX;
  uint8_t* const bf_quad = smem + CH_RING_SAFE + CH_F32_BYTES + quad * GEMM_BOX;
  float* const ssq_x = reinterpret_cast<float*>(smem + CH_RING_SAFE + CH_F32_BYTES + CH_BF_BYTES + 512);
  const int sw = lane & 7;
  const int rr = lane >> 3, cc = lane & 7;
  const int t = blockIdx.x;
  const int mt = t / ph.tiles_n;
  const int m0 = mt * GEMM_BM, n0 = (t - mt * ph.tiles_n) * BN;
  const int row = m0 + quad * 32 + lane;
  const bool live = row < p.M;
  const int ipt = (p.K / GEMM_BK) * p.n_pairs;
  constexpr int CPW = NCH / 2;                   // chunks per warp

  // weights first (no dependency): lane l keeps bias[n0 + 32 ch + l] of this warp's chunks
  float bias_r[CPW];
#pragma unroll
  for (int i = 0; i < CPW; ++i) bias_r[i] = p.bias != nullptr ? __ldg(p.bias + n0 + (hf + 2 * i) * 32 + lane) : 0.0f;
  // dependency: row sums / residual of this row group come from the previous phase (or the previous kernel)
  if (ph.wait_flag == nullptr) {
    pdl_wait();
  } else {
    if (lane == 0) chain_wait_idle(c, c.b.dep, st.n_dep & 1, 8, pidx);
    __syncwarp();
  }
  auto issue_resid = [&](int i) {                // residual box of this warp's i-th chunk
    if (!has_resid || i >= CPW) return;
    __syncwarp();
    if (lane == 0) {
      fence_proxy_async_all();                   // earlier generic accesses (smem box reads, global writes of x) before the async copy
      uint64_t* bar = &c.b.res[e * 2 + ((st.qr + i) & 1)];
      mbar_expect_tx(bar, GEMM_BOX);
      tma_load_3d(fbox0 + ((st.qr + i) & 1) * GEMM_BOX, &ph.tmR, bar, n0 + (hf + 2 * i) * 32, m0 + quad * 32, 0);
    }
  };
  issue_resid(0);
  float rinv = 1.0f;
  if (p.ssq_in != nullptr && live) {
    float ss = 0.0f;
    for (int q = 0; q < p.ssq_parts; ++q) ss += __ldcg(p.ssq_in + (int64_t)row * p.ssq_parts + q);
    rinv = 1.0f / (sqrtf(ss) * p.inv_sqrt_rms_dim + p.rms_eps);
  }
  rinv *= p.a_scale;
  if (lane == 0) {                               // one polling lane per warp, backing off
    uint32_t spins = 0;
    while (!mbar_try_wait(c.b.mma_done, st.n_done & 1)) {
      __nanosleep(64);
      if (++spins > (1u << 24)) chain_timeout(c, 9, pidx, st.n_done & 1);
    }
  }
  __syncwarp();
  tc_fence_after();
  if (warp == 4 && lane == 0) chain_mark(c, pidx, 4);
  const bool two_acc = !CS && ipt > 1;
  const uint32_t t_row = c.tmem_base + ((uint32_t)(quad * 32) << 16);
  uint32_t r[32], r1[32];
  tmem_ld_32x32(t_row + hf * 32, r);
  if (two_acc) tmem_ld_32x32(t_row + BN + hf * 32, r1);
#pragma unroll 1
  for (int i = 0; i < CPW; ++i) {
    const int ch = hf + 2 * i;
    float v[32];
    tmem_ld_wait();
    const bool mk = warp == 4 && lane == 0 && i == 0;
    if (mk) chain_mark(c, pidx, 7);
    if (two_acc) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) + __uint_as_float(r1[j]);
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
    }
    if (i + 1 < CPW) {
      tmem_ld_32x32(t_row + (ch + 2) * 32, r);
      if (two_acc) tmem_ld_32x32(t_row + BN + (ch + 2) * 32, r1);
    }
    {
      float bl = bias_r[0];
#pragma unroll
      for (int u = 1; u < CPW; ++u) bl = (i == u) ? bias_r[u] : bl;
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = fmaf(v[j], rinv, __shfl_sync(0xffffffffu, bl, j));
    }
    uint8_t* const fbox = fbox0 + (has_resid ? ((st.qr + i) & 1) * GEMM_BOX : 0);
    uint8_t* frow = fbox + lane * 128;
    if (has_resid) {
      issue_resid(i + 1);
      if (lane == 0) chain_wait(c, &c.b.res[e * 2 + ((st.qr + i) & 1)], ((st.qr + i) >> 1) & 1, 10, pidx);
      __syncwarp();
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float4 x4 = lds128(frow + ((q ^ sw) << 4));
        v[4 * q] += x4.x; v[4 * q + 1] += x4.y; v[4 * q + 2] += x4.z; v[4 * q + 3] += x4.w;
      }
    }
    if (mk) chain_mark(c, pidx, 8);
    if (has_f32) {
#pragma unroll
      for (int q = 0; q < 8; ++q) sts128(frow + ((q ^ sw) << 4), v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
    }
    float ssq_acc = 0.0f;
    if (has_ssq) {
#pragma unroll
      for (int j = 0; j < 32; ++j) ssq_acc = fmaf(v[j], v[j], ssq_acc);
      if (hf == 1) ssq_x[quad * 32 + lane] = ssq_acc;
    }
    const int c0 = (ch & 1) * 4;
    uint8_t* brow = bf_quad + lane * 128;
    if (MODE == 2) {
      if (f16_out) {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          sts128u(brow + (((c0 + q) ^ sw) << 4), gelu_pair_f16x2(v[8 * q], v[8 * q + 1]),
                  gelu_pair_f16x2(v[8 * q + 2], v[8 * q + 3]), gelu_pair_f16x2(v[8 * q + 4], v[8 * q + 5]),
                  gelu_pair_f16x2(v[8 * q + 6], v[8 * q + 7]));
      } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          sts128u(brow + (((c0 + q) ^ sw) << 4), gelu_pair_bf16x2(v[8 * q], v[8 * q + 1]),
                  gelu_pair_bf16x2(v[8 * q + 2], v[8 * q + 3]), gelu_pair_bf16x2(v[8 * q + 4], v[8 * q + 5]),
                  gelu_pair_bf16x2(v[8 * q + 6], v[8 * q + 7]));
      }
    } else {
      if (f16_out) {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          sts128u(brow + (((c0 + q) ^ sw) << 4), pack16(true, out_scale, v[8 * q], v[8 * q + 1]),
                  pack16(true, out_scale, v[8 * q + 2], v[8 * q + 3]), pack16(true, out_scale, v[8 * q + 4], v[8 * q + 5]),
                  pack16(true, out_scale, v[8 * q + 6], v[8 * q + 7]));
      } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          sts128u(brow + (((c0 + q) ^ sw) << 4), pack_bf16x2(v[8 * q], v[8 * q + 1]), pack_bf16x2(v[8 * q + 2], v[8 * q + 3]),
                  pack_bf16x2(v[8 * q + 4], v[8 * q + 5]), pack_bf16x2(v[8 * q + 6], v[8 * q + 7]));
      }
    }
    if (mk) chain_mark(c, pidx, 9);
    named_bar_sync(2 + quad, 64);
    if (mk) chain_mark(c, pidx, 10);
    if (hf == 0 && has_ssq && live)
      ssq_out[(int64_t)row * (p.N / 64) + ((n0 + ch * 32) >> 6)] = ssq_acc + ssq_x[quad * 32 + lane];
    if (has_f32) {
      float* gp = p.out_f32 + (int64_t)(m0 + quad * 32) * p.N + n0 + ch * 32 + cc * 4;
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const int rw = it * 4 + rr;
        const float4 x4 = lds128(fbox + rw * 128 + ((cc ^ (rw & 7)) << 4));
        if (m0 + quad * 32 + rw < p.M) *reinterpret_cast<float4*>(gp + (int64_t)rw * p.N) = x4;
      }
    }
    if (mk) chain_mark(c, pidx, 11);
    {
      __nv_bfloat16* gp = p.out_bf16 + (int64_t)(m0 + quad * 32) * p.N + n0 + (ch >> 1) * 64 + cc * 8;
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const int rw = hf * 16 + it * 4 + rr;
        const float4 x4 = lds128(bf_quad + rw * 128 + ((cc ^ (rw & 7)) << 4));
        if (m0 + quad * 32 + rw < p.M) *reinterpret_cast<float4*>(gp + (int64_t)rw * p.N) = x4;
      }
    }
    if (mk) chain_mark(c, pidx, 12);
    named_bar_sync(2 + quad, 64);                // the pair's shared boxes are rewritten by the next chunk
  }
  if (has_resid) st.qr += CPW;
  if (warp == 4 && lane == 0) chain_mark(c, pidx, 5);
  // this warp is done with TMEM and the staging boxes of this phase
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(c.b.epi_done);
  // tile complete: every epilogue warp's global stores are ordered before ONE release increment of the row-group counter
  if (ph.signal_flag != nullptr) {
    named_bar_sync(1, 256);
    if (warp == 4 && lane == 0) {
      __threadfence();
      fence_proxy_async_all();                   // consumers read these bytes through TMA
      atomicAdd(ph.signal_flag + mt, 1u);
      chain_mark(c, pidx, 6);
    }
  }
  ++st.n_done;
  if (ph.wait_flag != nullptr) ++st.n_dep;
}

template <int BN0, int MD0, int BN1, int MD1, int BN2, int MD2, int BN3, int MD3>
__global__ void __launch_bounds__(CH_THREADS, 1) gemm_chain_kernel(const __grid_constant__ ChainArgs args) {
  extern __shared__ __align__(1024) uint8_t ch_smem[];
  uint8_t* smem = ch_smem;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* tail = smem + CH_RING_SAFE + CH_F32_BYTES + CH_BF_BYTES;
  ChainCtx c;
  c.smem = smem;
  c.b.full = reinterpret_cast<uint64_t*>(tail);
  c.b.empty = c.b.full + CH_MAX_STAGES;
  c.b.mma_done = c.b.empty + CH_MAX_STAGES;
  c.b.epi_done = c.b.mma_done + 1;
  c.b.dep = c.b.epi_done + 1;
  c.b.res = c.b.dep + 1;                         // [16]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(c.b.res + 16);
  c.dbg = args.dbg;
  c.trace = args.trace;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0) {
    if (lane < CH_MAX_STAGES) mbar_init(&c.b.full[lane], 1);
    else if (lane < 2 * CH_MAX_STAGES) mbar_init(&c.b.full[lane], 2);      // empty[]
    if (lane == 16) mbar_init(c.b.mma_done, 2);
    if (lane == 17) mbar_init(c.b.epi_done, 8);
    if (lane == 18) mbar_init(c.b.dep, 1);
    if (lane < 16) mbar_init(&c.b.res[lane], 1);
    fence_barrier_init();
  } else if (warp == 1 && lane == 0) {
    tma_prefetch_desc(&args.ph[0].tmA);
    tma_prefetch_desc(&args.ph[0].tmB);
    if (MD0 == 3) tma_prefetch_desc(&args.ph[0].tmR);
  } else if (warp == 3 && lane == 0) {
    if (BN1 > 0) { tma_prefetch_desc(&args.ph[1].tmA); tma_prefetch_desc(&args.ph[1].tmB); }
    if (BN2 > 0) { tma_prefetch_desc(&args.ph[2].tmA); tma_prefetch_desc(&args.ph[2].tmB); }
    if (BN3 > 0) { tma_prefetch_desc(&args.ph[3].tmA); tma_prefetch_desc(&args.ph[3].tmB); }
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  c.tmem_base = *tmem_slot;
  pdl_launch_dependents();

  ChainState st;
  if (warp < 2) {
    if (lane == 0) {
      chain_produce<BN0>(args.ph[0], 0, warp, c, st);
      if constexpr (BN1 > 0) chain_produce<BN1>(args.ph[1], 1, warp, c, st);
      if constexpr (BN2 > 0) chain_produce<BN2>(args.ph[2], 2, warp, c, st);
      if constexpr (BN3 > 0) chain_produce<BN3>(args.ph[3], 3, warp, c, st);
    }
  } else if (warp < 4) {
    if (lane == 0) {
      chain_issue<BN0>(args.ph[0], 0, warp - 2, c, st);
      if constexpr (BN1 > 0) chain_issue<BN1>(args.ph[1], 1, warp - 2, c, st);
      if constexpr (BN2 > 0) chain_issue<BN2>(args.ph[2], 2, warp - 2, c, st);
      if constexpr (BN3 > 0) chain_issue<BN3>(args.ph[3], 3, warp - 2, c, st);
    }
  } else {
    chain_epilogue<BN0, MD0>(args.ph[0], 0, c, st);
    if constexpr (BN1 > 0) chain_epilogue<BN1, MD1>(args.ph[1], 1, c, st);
    if constexpr (BN2 > 0) chain_epilogue<BN2, MD2>(args.ph[2], 2, c, st);
    if constexpr (BN3 > 0) chain_epilogue<BN3, MD3>(args.ph[3], 3, c, st);
  }
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(c.tmem_base);
  }
}

// ------------------------------------------------------------------------------------------------ host
namespace {
struct Recipe { int bn[CH_MAX_PHASES], md[CH_MAX_PHASES]; const void* fn; };

template <int BN0, int MD0, int BN1, int MD1, int BN2, int MD2, int BN3, int MD3>
Recipe make_recipe() {
  return Recipe{{BN0, BN1, BN2, BN3}, {MD0, MD1, MD2, MD3},
                reinterpret_cast<const void*>(&gemm_chain_kernel<BN0, MD0, BN1, MD1, BN2, MD2, BN3, MD3>)};
}
}  // namespace

// The chains the MMDT block produces at one-wave sizes: {out-proj, fc1, fc2, next QKV}, the same without the QKV (last
// block), and width variants the tile chooser picks for smaller d_mlp / d_model.
static const Recipe* chain_recipes(int* n) {
  static const Recipe r[] = {
      make_recipe<64, 3, 256, 2, 64, 3, 192, 1>(),
      make_recipe<64, 3, 256, 2, 64, 3, 0, 0>(),
      make_recipe<64, 3, 128, 2, 64, 3, 128, 1>(),
      make_recipe<64, 3, 128, 2, 64, 3, 0, 0>(),
  };
  *n = (int)(sizeof(r) / sizeof(r[0]));
  return r;
}

// Returns 1 when the chain was launched, 0 when this combination is not chainable (the caller launches the GEMMs one by
// one), negative on error.  flags: CH_MAX_PHASES x max_row_tiles zero-initialised counters owned by the caller.
int launch_gemm_chain(const avd_gemm_args* gs, int n, unsigned int* flags, int flags_row_tiles, cudaStream_t s) {
  if (n < 2 || n > CH_MAX_PHASES || flags == nullptr) return 0;
  int sms = 148;
  {
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  ChainArgs args{};
  int bn[CH_MAX_PHASES] = {0, 0, 0, 0}, md[CH_MAX_PHASES] = {0, 0, 0, 0};
  const int mt = (gs[0].M + GEMM_BM - 1) / GEMM_BM;
  if (mt > flags_row_tiles) return 0;
  int grid = 0;
  for (int i = 0; i < n; ++i) {
    GemmPlan pl;
    const int rc = plan_gemm(gs[i], &pl);
    if (rc < 0) return rc;
    if (pl.persistent || pl.mode < 1 || pl.mode > 3 || pl.bn < 64 || gs[i].M != gs[0].M) return 0;
    if (i > 0 && gs[i].K != gs[i - 1].N) return 0;           // phase i consumes the whole row of phase i-1
    if (i > 0 && gs[i].N / pl.bn != args.ph[0].tiles_n) return 0;   // a CTA keeps its 128-row group through the phases
    bn[i] = pl.bn; md[i] = pl.mode;
    ChainPhase& ph = args.ph[i];
    ph.p = pl.p;
    ph.p.trace = nullptr;
    ph.tiles_n = gs[i].N / pl.bn;
    ph.tiles_total = mt * ph.tiles_n;
    if (ph.tiles_total > sms) return 0;
    grid = ph.tiles_total > grid ? ph.tiles_total : grid;
    const int stage = CH_A_BYTES + pl.bn * GEMM_BK * 2;
    int stages = (pl.mode == 3 ? CH_RING_SAFE : CH_RING_WIDE) / stage;
    if (stages > CH_MAX_STAGES) stages = CH_MAX_STAGES;
    stages &= ~1;
    if (stages < 2) return 0;
    ph.stages = stages;
    ph.safe_slots = CH_RING_SAFE / stage;
    const int nseg = pl.nseg;
    AVD_TRY(get_tensor_map_3d(gs[i].A, 2, gs[i].K, gs[i].M, nseg, (uint64_t)gs[i].K * 2, (uint64_t)gs[i].M * gs[i].K * 2,
                              GEMM_BK, GEMM_BM, &ph.tmA));
    AVD_TRY(get_tensor_map_3d(gs[i].W, 2, gs[i].K, gs[i].N, nseg, (uint64_t)gs[i].K * 2, (uint64_t)gs[i].N * gs[i].K * 2,
                              GEMM_BK, pl.bn, &ph.tmB));
    ph.tmR = ph.tmA;
    if (gs[i].resid != nullptr)
      AVD_TRY(get_tensor_map_3d(gs[i].resid, 4, gs[i].N, gs[i].M, 1, (uint64_t)gs[i].N * 4, (uint64_t)gs[i].M * gs[i].N * 4,
                                32, 32, &ph.tmR));
    ph.wait_flag = i > 0 ? flags + (i - 1) * flags_row_tiles : nullptr;
    ph.wait_target = i > 0 ? args.ph[i - 1].tiles_n : 0;
    ph.wait_total = ph.wait_target + ph.tiles_n;
    ph.signal_flag = i + 1 < n ? flags + i * flags_row_tiles : nullptr;
  }
  int nr = 0;
  const Recipe* recipes = chain_recipes(&nr);
  const void* fn = nullptr;
  for (int k = 0; k < nr && fn == nullptr; ++k) {
    bool ok = true;
    for (int i = 0; i < CH_MAX_PHASES; ++i) ok = ok && recipes[k].bn[i] == bn[i] && recipes[k].md[i] == md[i];
    if (ok) fn = recipes[k].fn;
  }
  if (fn == nullptr) return 0;
  args.trace = (g_dbg_trace != nullptr && (int64_t)grid * CH_MAX_PHASES * 16 <= g_dbg_trace_slots) ? g_dbg_trace : nullptr;
#ifdef AVD_DEBUG_HOOKS
  if (args.trace != nullptr) g_dbg_trace = nullptr;   // one traced launch per avd_debug_trace call: the stamps of later launches would overwrite it
#endif
  args.dbg = ps_debug_device_ptr();
  if (args.dbg == nullptr) return -1;
  AVD_TRY(ensure_dyn_smem(fn, CH_SMEM));
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(CH_THREADS);
  cfg.dynamicSmemBytes = CH_SMEM;
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
  if (kind_skipped(KIND_GEMM)) return 1;
  LaunchScope ls(KIND_GEMM, s);
  void* kargs[1] = {&args};
  AVD_CUDA(cudaLaunchKernelExC(&cfg, fn, kargs));
  return 1;
}

}  // namespace avd
