// avd_gemm_ps.cu -- K4, persistent form: out = epilogue( A[M,K] x W[N,K]^T ) with the epilogue of one output tile
// running UNDER the mainloop of the next one.
//
// Why: with d_model = 512 every output element carries only 2K = 1024 flops, so at tensor-core speed an SM must
// retire ~8 epilogue outputs per cycle; in the one-tile-per-CTA kernel (avd_gemm.cu) the tensor pipe idles during the
// whole epilogue (40 % of the kernel at mvp shapes, and the reason the large-M sweep plateaus near 30 % of peak).
//
// One CTA per SM, 384 threads, tiles t = blockIdx.x, +gridDim.x, ... (n fastest):
//   warps 0,1   TMA producers: k-block iteration `it` (counted over ALL tiles of this CTA) belongs to producer it & 1
//               and ring slot it % stages (stages is even, so a slot never changes hands: the ring is two interleaved
//               rings, one per {producer h, issuer h}); two issuing threads because one thread sustains only ~one
//               {wait, expect_tx, 2 x TMA} round per 300 cycles (tools/tma_probe.cu).  The weight boxes of the first
//               ring pass are issued before griddepcontrol.wait (they do not depend on the previous kernel).
//   warps 2,3   MMA issuers (warp 2 also owns TMEM): tcgen05.mma.cta_group::1.kind::f16 M=128 N=BN K=16, 4 per k-block.
//               Issuer h takes the iterations it = h (mod 2) and accumulates its share of a tile into its own TMEM
//               tile [buffer tile & 1][h] (4 x BN columns in all; the epilogue adds the two halves): one thread needs
//               ~400 cycles per {mbarrier wait, 4 MMAs, commit} round whatever N is (tools/mma_probe.cu: any
//               dependent wait between MMA batches costs ~180 cycles), more than the 220-256 cycles the pipe needs for
//               the round at N = 64/128, so two threads alternate.  tcgen05.commit frees the ring slot and, after an
//               issuer's last k-block of the tile, counts one arrival on acc_full.
//   warps 4..11 epilogue: warp w reads TMEM lanes 32 (w % 4) .. +31; the two warps of a quadrant take the even / odd
//               32-column chunks.  v = acc * rinv(row) + bias (+ residual) [gelu] -> fp32 and/or bf16 segments,
//               staged through per-warp 128B-swizzled smem boxes and stored with 16-byte coalesced accesses; row
//               partial sums of squares for the next RMSNorm.  After its last tcgen05.ld of a tile the warp arrives on
//               acc_empty, which lets the issuer start tile + 2 in that buffer.
// Residual tiles come in per 32 x 32 chunk through TMA into the warp's fp32 staging box (the result overwrites it).
#include "avd_gemm_common.cuh"

#include <cstdlib>

namespace avd {

constexpr int PS_THREADS = 384;
constexpr int PS_MAX_STAGES = 8;
constexpr int PS_TAIL_BYTES = 3072;          // barriers + TMEM slot (512 B), ssq exchange (512 B), two bias tiles (2 x 1 KB)
constexpr int PS_SMEM_MAX = 232448;          // the 227 KB opt-in limit (the dynamic region starts 1024-byte aligned)

// Debug records of barrier waits that timed out: {tag, cta, thread, counter, parity, N, K, grid * 1000 + BN}.
// The record lives in mapped pinned host memory so that it survives a device-side trap.
static unsigned int* g_ps_dbg_host = nullptr;
static unsigned int* g_ps_dbg_dev = nullptr;

struct PsPlan {
  unsigned int* dbg;               // host-visible record of barrier waits that timed out (avd_debug_gemm_timeout)
  int stages;
  int tiles_n, tiles_total;
  int f32_off, bf_off, tail_off;   // byte offsets of the staging regions and the barrier block
};

// MODE: compile-time epilogue specialisation, as in gemm_tc_kernel (0 generic, 1 one 16-bit segment, 2 + packed GELU,
// 3 residual + fp32 + one 16-bit segment + row sums of squares).
template <int BN, int MODE>
__global__ void __launch_bounds__(PS_THREADS, 1) gemm_ps_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                const __grid_constant__ CUtensorMap tmB,
                                                                const __grid_constant__ CUtensorMap tmR,
                                                                const __grid_constant__ CUtensorMap tmO,
                                                                const GemmParams p, const PsPlan plan) {
  constexpr int A_BYTES = GEMM_BM * GEMM_BK * 2;
  constexpr int B_BYTES = BN * GEMM_BK * 2;
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  // BN = 256 ("column split", CS): the two issuers share every k-block and each owns 128 of the tile's 256 columns
  // (two M128 N128 MMAs per k-step instead of one per issuer on alternating k-blocks), so a tile needs ONE 256-column
  // accumulator and two tiles fit the 512 TMEM columns.  Why: the smem port carries 2 x (16 KB A + BN x 128 B W) per
  // k-block; at BN = 128 that is 512 cycles for 256 cycles of MMA, at BN = 256 it is 768 for 512.
  constexpr bool CS = BN == 256;
  constexpr uint32_t TMEM_COLS = BN <= 32 ? 128 : BN <= 64 ? 256 : 512;   // 2 buffers x 2 issuers x BN columns (CS: 2 x 256)
  static_assert(BN <= 128 || CS, "four accumulators of BN columns must fit the 512 TMEM columns");
  constexpr int NCH = BN / 32;
  extern __shared__ __align__(1024) uint8_t ps_smem[];
  uint8_t* smem = ps_smem;
  if ((smem_u32(smem) & 1023u) != 0) __trap();   // 128B-swizzle tiles need 1024-B aligned bases
  uint8_t* tail = smem + plan.tail_off;
  uint64_t* full = reinterpret_cast<uint64_t*>(tail);
  uint64_t* empty = full + PS_MAX_STAGES;
  uint64_t* acc_full = empty + PS_MAX_STAGES;
  uint64_t* acc_empty = acc_full + 2;
  uint64_t* res_bar = acc_empty + 2;           // two per epilogue warp (double-buffered residual boxes)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_bar + 16);
  float* ssq_x = reinterpret_cast<float*>(tail + 512);   // [128] partial row sums handed from warp w+4 to warp w

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // Bounded barrier wait: a protocol bug leaves a record in host-visible memory and traps instead of hanging the GPU.
  auto wait = [&](uint64_t* bar, uint32_t parity, unsigned tag, int counter) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
      if (++spins > (1u << 26)) {
        unsigned int* dbg = plan.dbg;      // [0] = number of records, then 8-word records from dbg[8]
        const unsigned idx = atomicAdd_system(&dbg[0], 1u);
        if (idx < 30) {
          unsigned int* r = dbg + 8 + idx * 8;
          r[0] = tag; r[1] = blockIdx.x; r[2] = threadIdx.x; r[3] = (unsigned)counter;
          r[4] = parity; r[5] = (unsigned)p.N; r[6] = (unsigned)p.K; r[7] = gridDim.x * 1000u + (unsigned)BN;
          __threadfence_system();
        }
        __nanosleep(4000000);              // let the other stuck waiters record themselves before the trap
        __trap();
      }
    }
  };
  const int stages = plan.stages;
  const int G = gridDim.x;
  const int kblocks = p.K / GEMM_BK;
  const int ipt = kblocks * p.n_pairs;                                   // k-block iterations per tile
  const int my_tiles = (plan.tiles_total - (int)blockIdx.x + G - 1) / G;
  const int total_it = my_tiles * ipt;

  unsigned long long* const trace = p.trace;
  if (threadIdx.x == 0) trace_mark(trace, 0);
  if (warp == 0) {                       // one barrier (or two) per lane instead of a serial loop
    if (lane < 2 * PS_MAX_STAGES) mbar_init(&full[lane], (CS && lane >= PS_MAX_STAGES) ? 2 : 1);   // full, empty (CS: both issuers release a slot)
    if (lane < 2) mbar_init(&acc_full[lane], 2);                              // one arrival per issuer
    if (lane < 2) mbar_init(&acc_empty[lane], 8);                             // one arrival per epilogue warp
    if (lane < 16) mbar_init(&res_bar[lane], 1);
    fence_barrier_init();
  } else if (warp == 1 && lane == 0) {   // descriptor prefetches off the initialising warp
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (MODE == 0 ? (p.has_resid != 0) : (MODE == 3)) tma_prefetch_desc(&tmR);
  }
  if (warp == 2) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  if (threadIdx.x == 0) trace_mark(trace, 1);

  if (warp < 2) {
    // ------------------------------------------------------------------ TMA producer h: iterations it = h (mod 2)
    if (lane == 0) {
      const int h = warp;
      const uint32_t pairs = p.pairs;
      const int tiles_n = plan.tiles_n;
      const int n_pairs = p.n_pairs;
      // cursor over this CTA's k-block iterations (tile, operand-segment pair, k-block), advanced without divisions
      struct Cur { int tl, pi, kb, m0, n0; };
      auto tile_origin = [&](Cur& c) {
        const int t = (int)blockIdx.x + c.tl * G;
        const int mt = t / tiles_n;
        c.m0 = mt * GEMM_BM;
        c.n0 = (t - mt * tiles_n) * BN;
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
        mbar_expect_tx(&full[it], STAGE_BYTES);
        tma_load_3d(smem + it * STAGE_BYTES + A_BYTES, &tmB, &full[it], cb.kb * GEMM_BK, cb.n0, (pairs >> (4 * cb.pi)) & 3);
        advance(cb); advance(cb);
      }
      pdl_wait();
      int it = h;
      for (; it < pre_end; it += 2) {
        tma_load_3d(smem + it * STAGE_BYTES, &tmA, &full[it], ca.kb * GEMM_BK, ca.m0, (pairs >> (4 * ca.pi + 2)) & 3);
        advance(ca); advance(ca);
      }
      // steady state: `it` continues from the first ring pass; stages is even, so slot parity == h
      int slot = it % stages;
      uint32_t ph = ((it / stages) & 1) ^ 1;             // parity of the consumption this refill waits for
      for (; it < total_it; it += 2) {
        wait(&empty[slot], ph, 1, it);
        const uint32_t pr = pairs >> (4 * ca.pi);
        uint8_t* dst = smem + slot * STAGE_BYTES;
        mbar_expect_tx(&full[slot], STAGE_BYTES);
        tma_load_3d(dst, &tmA, &full[slot], ca.kb * GEMM_BK, ca.m0, (pr >> 2) & 3);
        tma_load_3d(dst + A_BYTES, &tmB, &full[slot], ca.kb * GEMM_BK, ca.n0, pr & 3);
        advance(ca); advance(ca);
        slot += 2;
        if (slot >= stages) { slot -= stages; ph ^= 1; }
      }
      if (h == 0) trace_mark(trace, 10);
    }
  } else if (warp < 4) {
    // ------------------------------------------------------------------ MMA issuer h: iterations j = h (mod 2) of each tile
    if (lane == 0) {
      const int h = warp - 2;
      const uint32_t idesc = make_idesc_bf16(GEMM_BM, CS ? 128 : BN, 0, 0, p.f16_in);
      const uint32_t desc_hi = smem_desc_hi_sw128(1024);
      const uint32_t a_lo0 = smem_desc_lo_sw128(smem_u32(smem), 16);
      if (CS) {
        // column split: every k-block of every tile, this issuer's 128 columns (W rows 128 h .. 128 h + 127 of the stage)
        int slot = 0;
        uint32_t ph = 0;
        for (int i = 0; i < my_tiles; ++i) {
          const int b = i & 1;
          wait(&acc_empty[b], ((i >> 1) & 1) ^ 1, 2, i);
          tc_fence_after();
          const uint32_t acc = tmem_base + b * BN + h * 128;
          for (int j = 0; j < ipt; ++j) {
            wait(&full[slot], ph, 3, i * ipt + j);
            tc_fence_after();
            if (h == 0 && i == 0 && j == 0) trace_mark(trace, 2);
            const uint32_t a_lo = a_lo0 + slot * (STAGE_BYTES >> 4);
            const uint32_t b_lo = a_lo + (A_BYTES >> 4) + h * ((128 * GEMM_BK * 2) >> 4);
#pragma unroll
            for (int k = 0; k < GEMM_BK / 16; ++k)
              umma_bf16_lh(acc, a_lo + 2 * k, b_lo + 2 * k, desc_hi, idesc, (j > 0 || k > 0) ? 1u : 0u);
            umma_commit(&empty[slot]);                     // second of the two arrivals frees the slot
            if (++slot == stages) { slot = 0; ph ^= 1; }
          }
          umma_commit(&acc_full[b]);
          if (h == 0 && i == my_tiles - 1) trace_mark(trace, 3);
        }
      } else
      for (int i = 0; i < my_tiles; ++i) {
        const int b = i & 1;
        const int base = i * ipt;                          // global iteration of this tile's first k-block
        const int j0 = (h ^ base) & 1;                     // this issuer's first k-block iteration of the tile
        const int pass0 = (base + j0) / stages;
        int slot = base + j0 - pass0 * stages;             // ring position of global iteration base + j
        uint32_t ph = pass0 & 1;
        wait(&acc_empty[b], ((i >> 1) & 1) ^ 1, 2, i);     // the epilogue has drained this buffer (tile i - 2)
        tc_fence_after();
        const uint32_t acc = tmem_base + (b * 2 + h) * BN;
        for (int j = j0; j < ipt; j += 2) {
          wait(&full[slot], ph, 3, base + j);
          tc_fence_after();
          if (h == 0 && i == 0 && j == j0) trace_mark(trace, 2);
          const uint32_t a_lo = a_lo0 + slot * (STAGE_BYTES >> 4);
          const uint32_t b_lo = a_lo + (A_BYTES >> 4);
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k)           // +32 bytes (2 x 16-byte units) per 16-wide k step
            umma_bf16_lh(acc, a_lo + 2 * k, b_lo + 2 * k, desc_hi, idesc, (j > j0 || k > 0) ? 1u : 0u);
          umma_commit(&empty[slot]);                       // slot reusable once these MMAs have read it
          slot += 2;
          if (slot >= stages) { slot -= stages; ph ^= 1; }
        }
        if (j0 < ipt) umma_commit(&acc_full[b]);           // this issuer's half of tile i is complete
        else mbar_arrive(&acc_full[b]);                    // a single k-block: nothing for this issuer
        if (h == 0 && i == my_tiles - 1) trace_mark(trace, 3);
      }
    }
#define PS_EPI_TILE(i) ((int)blockIdx.x + (i) * G)
#define PS_EPI_ROW0(mt) ((mt) * GEMM_BM)
#define PS_EPI_ACC_EMPTY(b) mbar_arrive(&acc_empty[b])
#include "avd_gemm_ps_epilogue.inc"
#undef PS_EPI_TILE
#undef PS_EPI_ROW0
#undef PS_EPI_ACC_EMPTY

  __syncwarp();   // lanes 1..31 of the role warps wait here for their lane 0
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<TMEM_COLS>(tmem_base);
  }
  if (threadIdx.x == 0) trace_mark(trace, 7);
}

int ps_debug_read(unsigned int* out8) {       // host copy of the timeout record; clears it
  for (int i = 0; i < 256; ++i) {
    out8[i] = g_ps_dbg_host != nullptr ? g_ps_dbg_host[i] : 0u;
    if (g_ps_dbg_host != nullptr) g_ps_dbg_host[i] = 0u;
  }
  return 0;
}

unsigned int* ps_debug_device_ptr() {
  if (g_ps_dbg_host == nullptr) {
    if (cudaHostAlloc(reinterpret_cast<void**>(&g_ps_dbg_host), 1024, cudaHostAllocMapped) != cudaSuccess) return nullptr;
    for (int i = 0; i < 256; ++i) g_ps_dbg_host[i] = 0u;
    if (cudaHostGetDevicePointer(reinterpret_cast<void**>(&g_ps_dbg_dev), g_ps_dbg_host, 0) != cudaSuccess) return nullptr;
  }
  return g_ps_dbg_dev;
}

static int ps_num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
  }
  return n;
}

template <int BN>
static int launch_ps_bn(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmR, const CUtensorMap& tmO,
                        const GemmParams& p, const PsPlan& plan, int smem_bytes, int grid, int mode, cudaStream_t s) {
  using KernelFn = void (*)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const CUtensorMap, const GemmParams,
                            const PsPlan);
  static const KernelFn kernels[4] = {gemm_ps_kernel<BN, 0>, gemm_ps_kernel<BN, 1>, gemm_ps_kernel<BN, 2>,
                                      gemm_ps_kernel<BN, 3>};
  AVD_TRY(ensure_dyn_smem(reinterpret_cast<const void*>(kernels[mode]), PS_SMEM_MAX));   // once per device
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(PS_THREADS);
  cfg.dynamicSmemBytes = smem_bytes;
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
  if (kind_skipped(KIND_GEMM)) return 0;
  LaunchScope ls(KIND_GEMM, s);
  AVD_CUDA(cudaLaunchKernelEx(&cfg, kernels[mode], tmA, tmB, tmR, tmO, p, plan));
  return 0;
}

// smem plan of a tile width; returns false when fewer than 2 ring stages fit
static bool ps_plan(const avd_gemm_args& g, int out_nseg, int bn, PsPlan* plan, int* smem_bytes) {
  const int stage = (GEMM_BM + bn) * GEMM_BK * 2;
  const int f32_bytes = (g.out_f32 != nullptr || g.resid != nullptr) ? 16 * GEMM_BOX : 0;
  // the 16-bit-only epilogues (MODE 1 / 2) double-buffer their staging boxes for the bulk stores
  const bool fast16 = g.out_bf16 != nullptr && out_nseg == 1 && g.out_f32 == nullptr && g.resid == nullptr &&
                      g.ssq_out == nullptr && (g.act == 0 || g.act == 2);
  const int bf_bytes = out_nseg * 4 * GEMM_BOX * (fast16 ? 2 : 1);
  const int fixed = f32_bytes + bf_bytes + PS_TAIL_BYTES;
  int stages = (PS_SMEM_MAX - fixed) / stage;
  if (stages > PS_MAX_STAGES) stages = PS_MAX_STAGES;
  const int cap = options().ps_max_stages;
  if (cap > 0 && stages > cap) stages = cap;
  // even: global iteration `it` then always lands in a ring slot of its own parity, so every slot (and its two barriers)
  // has ONE producer and ONE issuer, who see its uses in order.  The 256-wide column-split tiles do not need that: both
  // issuers consume every slot in order, and a producer's refill of pass n waits for the consumption of pass n - 1, which
  // cannot complete before the OTHER producer's fill of pass n - 1 -- no waiter is ever two phases away from its
  // barrier.  An odd ring is what lets the residual GEMMs (fp32 staging boxes: room for three 48 KB stages, not four)
  // take the 256-wide tiles: smem port 768 cycles per k-block against 512 of MMA instead of 512 against 256.
  if (bn != 256) stages &= ~1;
  if (stages < 2) return false;
  plan->stages = stages;
  plan->f32_off = stages * stage;
  plan->bf_off = plan->f32_off + f32_bytes;
  plan->tail_off = plan->bf_off + bf_bytes;
  *smem_bytes = plan->tail_off + PS_TAIL_BYTES;
  return true;
}

int launch_gemm_ps(const avd_gemm_args& g, GemmParams p, int nseg, int force_bn, cudaStream_t s) {
  const int out_nseg = p.out_nseg;
  const int mt = (g.M + GEMM_BM - 1) / GEMM_BM;
  const int sms = ps_num_sms();
  const int ipt = (g.K / GEMM_BK) * p.n_pairs;
  // Tile width by a cycle model of the persistent schedule (constants from the phase traces of tools/kbench.py):
  //   mainloop of a tile  = k-block iterations x max(TMA ingest at ~110 B/cycle/SM, 4 MMAs of max(BN/2, 48) cycles)
  //   epilogue of a tile  = (BN / 64) chunk rounds of the 8 epilogue warps
  //   time = first mainloop + (waves - 1) x max(mainloop, epilogue) + last epilogue
  //   (epilogue constants: 1500 cycles per 64-column round for the generic / residual epilogues, 1100 for the specialised
  //   16-bit-only ones, GELU +10 %: with the old 1500 / +30 % the model kept fc1 on 128-wide tiles, 12 % slower than
  //   256-wide at M = 4256 .. 17024, tools/kbench.py --M with AVD_GEMM_BN=256)
  const int c_epi_env = options().ps_epi_cycles;
  const bool fast16 = g.out_bf16 != nullptr && out_nseg == 1 && g.out_f32 == nullptr && g.resid == nullptr &&
                      g.ssq_out == nullptr;
  const int c_epi = c_epi_env > 0 ? c_epi_env : (fast16 ? 1100 : 1500);
  auto usable = [&](int bn) {
    if (g.N % bn != 0) return false;
    if ((g.ssq_out != nullptr || g.out_bf16 != nullptr) && bn < 64) return false;
    PsPlan pl; int sm;
    return ps_plan(g, out_nseg, bn, &pl, &sm) && pl.stages >= 3;
  };
  auto cost = [&](int bn) {
    const int64_t tiles = (int64_t)mt * (g.N / bn);
    const int64_t waves = (tiles + sms - 1) / sms;
    const double ingest = 2.0 * (GEMM_BM + bn);   // smem port (128 B/cycle) carries the TMA writes AND the MMA operand reads
    const double mma = 4.0 * (bn / 2 > 52 ? bn / 2 : 52);
    const double ml = ipt * (ingest > mma ? ingest : mma);
    double epi = (bn >= 64 ? bn / 64 : 1) * (double)c_epi;
    if (g.act != 0) epi *= fast16 ? 1.1 : 1.3;
    if (out_nseg > 1) epi *= 1.0 + 0.3 * (out_nseg - 1);
    return ml + (waves - 1) * (ml > epi ? ml : epi) + epi;
  };
  const int cands[4] = {256, 128, 64, 32};
  int bn = 0;
  if (force_bn > 0 && usable(force_bn)) bn = force_bn;
  if (bn == 0) {
    double best = 0.0;
    for (int i = options().ps_no256 ? 1 : 0; i < 4; ++i)
      if (usable(cands[i])) {
        const double c = cost(cands[i]);
        if (bn == 0 || c < best) { bn = cands[i]; best = c; }
      }
  }
  AVD_REQUIRE(bn > 0, "no tile shape fits this combination of outputs: use fewer outputs per GEMM");

  PsPlan plan{};
  int smem_bytes = 0;
  ps_plan(g, out_nseg, bn, &plan, &smem_bytes);
  plan.dbg = ps_debug_device_ptr();
  AVD_REQUIRE(plan.dbg != nullptr, "cannot allocate the barrier-timeout record");
  plan.tiles_n = g.N / bn;
  plan.tiles_total = mt * plan.tiles_n;
  const int grid = plan.tiles_total < sms ? plan.tiles_total : sms;
  {
    const int64_t n_ctas = grid;
    p.trace = (g_dbg_trace != nullptr && n_ctas * 32 <= g_dbg_trace_slots) ? g_dbg_trace : nullptr;
  }

  CUtensorMap tmA, tmB, tmR;
  AVD_TRY(get_tensor_map_3d(g.A, 2, g.K, g.M, nseg, (uint64_t)g.K * 2, (uint64_t)g.M * g.K * 2, GEMM_BK, GEMM_BM, &tmA));
  AVD_TRY(get_tensor_map_3d(g.W, 2, g.K, g.N, nseg, (uint64_t)g.K * 2, (uint64_t)g.N * g.K * 2, GEMM_BK, bn, &tmB));
  tmR = tmA;   // placeholder when unused (never dereferenced)
  if (g.resid != nullptr)
    AVD_TRY(get_tensor_map_3d(g.resid, 4, g.N, g.M, 1, (uint64_t)g.N * 4, (uint64_t)g.M * g.N * 4, 32, 32, &tmR));
  CUtensorMap tmO = tmA;   // the 16-bit output as {64 columns, 32 rows} boxes (bulk stores of the MODE 1 / 2 epilogues)
  if (g.out_bf16 != nullptr && out_nseg == 1)
    AVD_TRY(get_tensor_map_3d(g.out_bf16, 2, g.N, g.M, 1, (uint64_t)g.N * 2, (uint64_t)g.M * g.N * 2, 64, 32, &tmO));
  int mode = 0;
  const bool one16 = g.out_bf16 != nullptr && out_nseg == 1;
  if (one16 && g.out_f32 == nullptr && g.resid == nullptr && g.ssq_out == nullptr && g.act == 0) mode = 1;
  else if (one16 && g.out_f32 == nullptr && g.resid == nullptr && g.ssq_out == nullptr && g.act == 2) mode = 2;
  else if (one16 && g.out_f32 != nullptr && g.resid != nullptr && g.ssq_out != nullptr && g.act == 0) mode = 3;
#ifdef AVD_DEBUG_HOOKS
  if (bn < 64 || options().gemm_generic) mode = 0;   // debug-hooks builds stamp every epilogue mode
#else
  if (p.trace != nullptr || bn < 64 || options().gemm_generic) mode = 0;
#endif
  // CTA pairs (avd_gemm_ps2.cu): the 256-wide tiles on two SMs each, half of the weight tile per SM
  // Measured (tools/kbench.py --M 32768, AVD_PS_PAIR = 0 / 1): QKV 51.7 -> 47.2 us (1092 TFLOP/s), fc1 83.1 -> 77.5,
  // fc2 91.9 -> 83.9, out-proj 41.2 -> 40.6.  (With a RELEASE at cluster scope on the peer's buffer hand-back the pair
  // was 7 % slower than the single-CTA kernel at K = 512: every epilogue warp then waited for its outstanding global
  // stores once per tile; the arrival is relaxed now, avd_gemm_ps2.cu.)
  const int pair_opt = options().ps_pair;
  if (pair_opt != 0 && g.N % 256 == 0 && (bn == 256 || pair_opt == 2)) {
    const int rc = launch_gemm_ps2(g, p, nseg, mode, sms, s);
    if (rc != 0) return rc < 0 ? rc : 0;
  }
  switch (bn) {
    case 256: return launch_ps_bn<256>(tmA, tmB, tmR, tmO, p, plan, smem_bytes, grid, mode, s);
    case 128: return launch_ps_bn<128>(tmA, tmB, tmR, tmO, p, plan, smem_bytes, grid, mode, s);
    case 64: return launch_ps_bn<64>(tmA, tmB, tmR, tmO, p, plan, smem_bytes, grid, mode, s);
    default: return launch_ps_bn<32>(tmA, tmB, tmR, tmO, p, plan, smem_bytes, grid, mode, s);
  }
}

}  // namespace avd
