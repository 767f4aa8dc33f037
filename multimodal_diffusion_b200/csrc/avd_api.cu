// avd_api.cu -- the C ABI of libavdiff_b200.so (include/avdiff_b200.h): argument validation, launch
// accounting / event profiler, TMA descriptor cache, and the host-side sequencing of the MMDT blocks, the
// noise head and the DDIM + CFG sampling loop.  Nothing here allocates device memory or synchronises.
#include "avd_internal.h"
namespace avd { int ps_debug_read(unsigned int* out8); }

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <set>
#include <unordered_map>
#include <utility>
#include <vector>

namespace avd {

// ---------------------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------------------
static thread_local std::string g_last_error;

void set_error(const std::string& msg) { g_last_error = msg; }

int fail(const char* file, int line, const std::string& msg) {
  const char* base = std::strrchr(file, '/');
  g_last_error = std::string(base ? base + 1 : file) + ":" + std::to_string(line) + ": " + msg;
  return -1;
}

// ---------------------------------------------------------------------------------------------------
// launch accounting + optional CUDA-event profiler
// ---------------------------------------------------------------------------------------------------
static const char* kKindNames[KIND_COUNT] = {"tokenize", "build_x",  "gemm_tcgen05", "attn_tcgen05", "attn_f32",
                                             "norm",     "ln_act",   "cfg_untok_ddim", "pack",       "misc"};
struct ProfRecord {
  int kind;
  cudaEvent_t beg, end;
};
static std::mutex g_prof_mu;
static bool g_prof_on = false;
static std::vector<ProfRecord> g_prof;
static std::vector<cudaEvent_t> g_event_pool;
static int64_t g_launches = 0;

static cudaEvent_t take_event() {
  if (!g_event_pool.empty()) {
    cudaEvent_t e = g_event_pool.back();
    g_event_pool.pop_back();
    return e;
  }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}

// Inside a stream capture the events become event-record nodes of the graph (cudaEventRecordExternal), so a replay
// of the captured step times every kernel on the device without host launch gaps in between.
static void record_event(cudaEvent_t e, cudaStream_t s) {
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(s, &st);
  if (st == cudaStreamCaptureStatusActive) cudaEventRecordWithFlags(e, s, cudaEventRecordExternal);
  else cudaEventRecord(e, s);
}

LaunchScope::LaunchScope(Kind k, cudaStream_t s) : kind(k), stream(s), slot(-1) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  ++g_launches;
  if (g_prof_on) {
    ProfRecord r{(int)k, take_event(), take_event()};
    record_event(r.beg, s);
    g_prof.push_back(r);
    slot = (int)g_prof.size() - 1;
  }
}

LaunchScope::~LaunchScope() {
  if (slot >= 0) {
    std::lock_guard<std::mutex> lk(g_prof_mu);
    record_event(g_prof[slot].end, stream);
  }
}

// ---------------------------------------------------------------------------------------------------
// TMA descriptor cache
// ---------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;

struct MapKey {
  const void* base;
  uint64_t d0, d1, d2, s1, s2;
  uint32_t b0, b1, dt;
  bool operator==(const MapKey& o) const {
    return base == o.base && d0 == o.d0 && d1 == o.d1 && d2 == o.d2 && s1 == o.s1 && s2 == o.s2 && b0 == o.b0 &&
           b1 == o.b1 && dt == o.dt;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    uint64_t h = reinterpret_cast<uint64_t>(k.base);
    auto mix = [&](uint64_t v) { h ^= v + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2); };
    mix(k.d0); mix(k.d1); mix(k.d2); mix(k.s1); mix(k.s2); mix(k.b0); mix(k.b1); mix(k.dt);
    return (size_t)h;
  }
};
static std::mutex g_map_mu;
static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> g_maps;
static std::deque<MapKey> g_map_order;          // insertion order: the oldest half is evicted when the cache is full
constexpr size_t MAP_CACHE_MAX = 8192;

// [d2][d1][d0] tensor (d0 contiguous) of bf16 (elem_bytes 2) or fp32 (4); box [1][box1][box0] with
// box0 * elem_bytes == 128 (one 128-byte swizzle row).
int get_tensor_map_3d(const void* base, int elem_bytes, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t stride1_bytes,
                      uint64_t stride2_bytes, uint32_t box0, uint32_t box1, CUtensorMap* out) {
  std::lock_guard<std::mutex> lk(g_map_mu);
  if (g_encode == nullptr) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || fn == nullptr || qres != cudaDriverEntryPointSuccess)
      AVD_FAIL("cuTensorMapEncodeTiled not available from the CUDA driver");
    g_encode = reinterpret_cast<EncodeTiledFn>(fn);
  }
  AVD_REQUIRE(elem_bytes == 2 || elem_bytes == 4, "tensor maps are built for bf16 or fp32");
  MapKey key{base, d0, d1, d2, stride1_bytes, stride2_bytes, box0, box1, (uint32_t)elem_bytes};
  auto it = g_maps.find(key);
  if (it != g_maps.end()) {
    *out = it->second;
    return 0;
  }
  AVD_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base must be 16-byte aligned");
  AVD_REQUIRE(stride1_bytes % 16 == 0 && stride2_bytes % 16 == 0, "TMA strides must be multiples of 16 bytes");
  AVD_REQUIRE(box0 * (uint32_t)elem_bytes == 128, "inner box must span one 128-byte swizzle row");
  AVD_REQUIRE(box1 >= 1 && box1 <= 256, "box rows out of range");
  cuuint64_t dims[3] = {d0, d1, d2};
  cuuint64_t strides[2] = {stride1_bytes, stride2_bytes};
  cuuint32_t box[3] = {box0, box1, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUtensorMap m;
  CUresult r = g_encode(&m, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3,
                        const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) AVD_FAIL("cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
  if (g_maps.size() >= MAP_CACHE_MAX) {          // a sampler holds a few hundred maps; evict oldest-first, never wholesale
    for (size_t i = 0; i < MAP_CACHE_MAX / 2 && !g_map_order.empty(); ++i) {
      g_maps.erase(g_map_order.front());
      g_map_order.pop_front();
    }
  }
  g_maps.emplace(key, m);
  g_map_order.push_back(key);
  *out = m;
  return 0;
}

int get_tensor_map_bf16_3d(const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t stride1_bytes,
                           uint64_t stride2_bytes, uint32_t box0, uint32_t box1, CUtensorMap* out) {
  return get_tensor_map_3d(base, 2, d0, d1, d2, stride1_bytes, stride2_bytes, box0, box1, out);
}

thread_local int t_skip_kinds = 0;

struct OptionDesc { const char* name; const char* env; int Options::*field; int dflt; };
static const OptionDesc kOptions[] = {
    {"pdl", "AVD_PDL", &Options::pdl, 1},
    {"gemm_bn", "AVD_GEMM_BN", &Options::gemm_bn, 0},
    {"gemm_v", "AVD_GEMM_V", &Options::gemm_v, 0},
    {"gemm_generic", "AVD_GEMM_GENERIC", &Options::gemm_generic, 0},
    {"gemm_dbg", "AVD_GEMM_DBG", &Options::gemm_dbg, 0},
    {"gemm_pair", "AVD_GEMM_PAIR", &Options::gemm_pair, 0},
    {"gemm_chain", "AVD_GEMM_CHAIN", &Options::gemm_chain, 0},
    {"ps_max_stages", "AVD_PS_MAX_STAGES", &Options::ps_max_stages, 0},
    {"ps_epi_cycles", "AVD_PS_EPI_CYCLES", &Options::ps_epi_cycles, 0},
    {"ps_pair", "AVD_PS_PAIR", &Options::ps_pair, 1},
    {"ps_no256", "AVD_PS_NO256", &Options::ps_no256, 0},
    {"attn_poly", "AVD_ATTN_POLY", &Options::attn_poly, 0},
    {"attn_small", "AVD_ATTN_SMALL", &Options::attn_small, 1},
    {"attn_pipe", "AVD_ATTN_PIPE", &Options::attn_pipe, 0},
    {"attn_fa", "AVD_ATTN_FA", &Options::attn_fa, 1},
    {"attn_one", "AVD_ATTN_ONE", &Options::attn_one, 1},
    {"chains", "AVD_CHAINS", &Options::chains, 1},
    {"bx_wide", "AVD_BX_WIDE", &Options::bx_wide, -1},
};

Options& options() {
  static Options o = [] {
    Options v{};
    for (const OptionDesc& d : kOptions) {
      const char* e = std::getenv(d.env);
      v.*(d.field) = (e != nullptr && *e != 0) ? std::atoi(e) : d.dflt;
    }
    return v;
  }();
  return o;
}

bool pdl_enabled() { return options().pdl != 0; }

int ensure_dyn_smem(const void* fn, int bytes) {
  static std::mutex mu;
  static std::set<std::pair<int, const void*>> done;
  int dev = 0;
  AVD_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(mu);
  if (done.count({dev, fn})) return 0;
  AVD_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  done.insert({dev, fn});
  return 0;
}

static inline int64_t align_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

// ---------------------------------------------------------------------------------------------------
// sampler workspace carving (shared by avd_sampler_workspace_bytes and the step functions)
// ---------------------------------------------------------------------------------------------------
struct SamplerLayout {
  int Nv, Na, N, Nt, Dt, Dv, Da, d, d_adapt, n_seq, Mt, parts, S_core, S_head;
  int row_off_v, row_off_a;
  int64_t M;
  // offsets in bytes
  int64_t x_init, xa_init, ssq_init, x, xa, ssq, qkv, att, hmid, h, act_a, act_b, y, eps, bx_tok, bx_y, chain_flags, total;
};

static int sampler_layout(const avd_sampler* s, SamplerLayout* L) {
  AVD_REQUIRE(s != nullptr, "null sampler");
  AVD_REQUIRE(s->target == AVD_TARGET_VIDEO || s->target == AVD_TARGET_AUDIO, "bad target");
  AVD_REQUIRE(s->B > 0, "B must be positive");
  AVD_REQUIRE(s->tt > 0 && s->th > 0 && s->tw > 0 && s->T % s->tt == 0 && s->H % s->th == 0 && s->W % s->tw == 0,
              "tube sizes must divide latent dims");
  AVD_REQUIRE(s->chunk_len > 0 && s->chunk_stride > 0 && s->Fa >= s->chunk_len, "bad audio chunking");
  AVD_REQUIRE(s->core.n_layers > 0 && s->core.n_layers <= AVD_MAX_LAYERS, "bad layer count");
  AVD_REQUIRE(s->head.n_shared >= 0 && s->head.n_shared <= AVD_MAX_SHARED, "bad shared-layer count");
  L->S_core = nseg_of(s->core.precision);
  L->S_head = nseg_of(s->head.precision);
  AVD_REQUIRE(L->S_core > 0 && L->S_head > 0, "bad precision");
  L->Nv = (s->T / s->tt) * (s->H / s->th) * (s->W / s->tw);
  L->Na = (s->Fa - s->chunk_len) / s->chunk_stride + 1;
  L->N = L->Nv + L->Na;
  L->Dv = s->Cv * s->tt * s->th * s->tw;
  L->Da = s->Ca * s->chunk_len;
  L->d = s->core.d_model;
  L->d_adapt = L->d - s->tstep_dim;
  AVD_REQUIRE(s->tstep_dim >= 0 && L->d_adapt > 0, "tstep_dim must be in [0, d_model)");
  AVD_REQUIRE(s->tstep_dim == 0 || (s->temb_steps != nullptr && s->temb_zero != nullptr), "missing timestep-embedding tables");
  AVD_REQUIRE(L->d % 64 == 0, "d_model must be a multiple of 64");
  AVD_REQUIRE(s->head.d_in == L->d, "head input width must equal d_model");
  L->row_off_v = 0;
  L->row_off_a = L->Nv;
  L->Nt = s->target == AVD_TARGET_VIDEO ? L->Nv : L->Na;
  L->Dt = s->target == AVD_TARGET_VIDEO ? L->Dv : L->Da;
  AVD_REQUIRE(s->head.d_out == L->Dt, "head output width must equal the target token width");
  L->n_seq = 2 * s->B;
  L->M = (int64_t)L->n_seq * L->N;
  L->Mt = L->n_seq * L->Nt;
  L->parts = L->d / 64;
  const int64_t d = L->d, M = L->M, Mt = L->Mt, dm = s->core.d_mlp, hid = s->head.hidden;
  int64_t off = 0;
  auto take = [&](int64_t bytes) { int64_t o = off; off = align_up(off + bytes, 1024); return o; };
  L->x_init = take(M * d * 4);
  L->xa_init = take((int64_t)L->S_core * M * d * 2);
  L->ssq_init = take(M * L->parts * 4);
  L->x = take(M * d * 4);
  L->xa = take((int64_t)L->S_core * M * d * 2);
  L->ssq = take(M * L->parts * 4);
  L->qkv = take(M * 3 * d * ((s->core.precision == AVD_PREC_BF16 || is_f16(s->core.precision)) ? 2 : 4));
  L->att = take((int64_t)L->S_core * M * d * 2);
  L->hmid = take((int64_t)L->S_core * M * dm * 2);
  // the head runs over ALL rows of the joint sequences (the target rows are picked by the CFG/DDIM kernel): with one
  // 128-row tile per CTA the extra prompt rows cost no time, and the final RMSNorm folds into the input projection
  L->h = take((int64_t)L->S_head * M * d * 2);
  L->act_a = take((int64_t)L->S_head * M * hid * 2);
  L->act_b = take((int64_t)L->S_head * M * hid * 2);
  L->y = take(M * hid * 4);
  L->eps = take(M * (int64_t)L->Dt * 4);
  // tensor-core input builder (target rows): token matrix as up to 3 bf16 segments, adapter output fp32
  L->bx_tok = take((int64_t)3 * s->B * L->Nt * L->Dt * 2);
  L->bx_y = take((int64_t)s->B * L->Nt * L->d_adapt * 4);
  L->chain_flags = take(GEMM_CHAIN_FLAG_BYTES);   // row-group counters of the chained GEMM launches (zero between launches)
  (void)Mt;
  L->total = off;
  return 0;
}

// One fork/join stream set per device for the "chains" schedule (see core_blocks).  The library owns these few streams
// and events; they carry no state between calls.
struct ChainStreams {
  cudaStream_t side[3] = {nullptr, nullptr, nullptr};
  cudaEvent_t fork = nullptr;
  cudaEvent_t join[3] = {nullptr, nullptr, nullptr};
};
static int chain_streams(ChainStreams** out) {
  static std::mutex mu;
  static std::unordered_map<int, ChainStreams> per_dev;
  int dev = 0;
  AVD_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(mu);
  ChainStreams& c = per_dev[dev];
  if (c.fork == nullptr) {
    AVD_CUDA(cudaEventCreateWithFlags(&c.fork, cudaEventDisableTiming));
    for (int i = 0; i < 3; ++i) {
      AVD_CUDA(cudaStreamCreateWithFlags(&c.side[i], cudaStreamNonBlocking));
      AVD_CUDA(cudaEventCreateWithFlags(&c.join[i], cudaEventDisableTiming));
    }
  }
  *out = &c;
  return 0;
}

// The blocks for rows [0, n_seq * N) of the given buffers on ONE stream.
// last_out / last_prec: when given, the last block's residual stream leaves as bf16 segments of that precision in
// last_out (the noise head's operand) instead of ws->xa.
static int core_blocks_rows(const avd_core_weights* w, const float* x_in, const void* xa_in, const float* ssq_in,
                            const avd_core_workspace* ws, int n_seq, int N, void* last_out, int last_prec, cudaStream_t st) {
  const int d = w->d_model, S = nseg_of(w->precision);
  const int M = n_seq * N, parts = d / 64, hd = d / w->n_heads;
  const bool f16 = is_f16(w->precision);
  const bool tc_attn = (w->precision == AVD_PREC_BF16) || f16;
  // fp16 mode: the residual-stream operand xa holds x * AVD_F16_XSCALE; its consumers (QKV, fc1) undo the scale
  // norm="layernorm": LayerNorm subtracts the row mean, so it cannot ride the GEMM as a row scale like RMSNorm does; it
  // is its own pass over the fp32 residual stream (layernorm_act_kernel) writing the operand of the QKV / fc1 GEMM
  const bool ln = w->norm_kind == 1;
  const float x_unscale = (f16 && !ln) ? 1.0f / AVD_F16_XSCALE : 0.0f, x_scale = f16 ? AVD_F16_XSCALE : 0.0f;
  // the four GEMMs of block l as argument blocks (mmdt.py:96-99); qkv(l) reads the residual stream block l-1 left
  auto qkv_args = [&](int l) {
    const avd_block_weights& b = w->blocks[l];
    avd_gemm_args g{};   // qkv = RMS1(x) W_in^T + b_in        (mmdt.py:97 -> :60)
    g.M = M; g.N = 3 * d; g.K = d; g.precision = w->precision;
    g.A = (l == 0 || ln) ? (ln ? ws->xa : xa_in) : ws->xa; g.W = b.w_qkv; g.bias = b.b_qkv;
    g.ssq_in = ln ? nullptr : ((l == 0) ? ssq_in : ws->ssq); g.ssq_parts = parts; g.rms_dim = d; g.rms_eps = w->rms_eps;
    g.a_scale = x_unscale;
    if (tc_attn) { g.out_bf16 = ws->qkv; g.out_precision = f16 ? AVD_PREC_F16 : AVD_PREC_BF16; }
    else g.out_f32 = static_cast<float*>(ws->qkv);
    return g;
  };
  auto oproj_args = [&](int l) {
    const avd_block_weights& b = w->blocks[l];
    avd_gemm_args g{};   // x = x + a W_o^T + b_o               (mmdt.py:97)
    g.M = M; g.N = d; g.K = d; g.precision = w->precision;
    g.A = ws->att; g.W = b.w_o; g.bias = b.b_o; g.resid = (l == 0) ? x_in : ws->x;
    g.out_f32 = ws->x; g.out_bf16 = ws->xa; g.out_precision = w->precision; g.ssq_out = ws->ssq;
    g.out_scale = x_scale;
    return g;
  };
  auto fc1_args = [&](int l) {
    const avd_block_weights& b = w->blocks[l];
    avd_gemm_args g{};   // u = gelu(RMS2(x) W_1^T + b_1)       (mmdt.py:98 -> :78-79)
    g.M = M; g.N = w->d_mlp; g.K = d; g.precision = w->precision;
    g.A = ws->xa; g.W = b.w_fc1; g.bias = b.b_fc1;
    g.ssq_in = ln ? nullptr : ws->ssq; g.ssq_parts = parts; g.rms_dim = d; g.rms_eps = w->rms_eps;
    g.a_scale = x_unscale;
    g.act = tc_attn ? 2 : 1;   // bf16 mode: fast erf (|err| < 2e-7, far below the bf16 rounding of the result)
    g.out_bf16 = ws->hmid; g.out_precision = w->precision;
    return g;
  };
  auto fc2_args = [&](int l) {
    const avd_block_weights& b = w->blocks[l];
    avd_gemm_args g{};   // x = x + u W_2^T + b_2               (mmdt.py:98 -> :81)
    g.M = M; g.N = d; g.K = w->d_mlp; g.precision = w->precision;
    g.A = ws->hmid; g.W = b.w_fc2; g.bias = b.b_fc2; g.resid = ws->x;
    g.out_f32 = ws->x; g.out_bf16 = ws->xa; g.out_precision = w->precision; g.ssq_out = ws->ssq;
    g.out_scale = x_scale;
    if (l == w->n_layers - 1 && last_out != nullptr) {   // the head's operand: pre-scaled too when the head runs in fp16
      g.out_bf16 = last_out; g.out_precision = last_prec; g.out_scale = is_f16(last_prec) ? AVD_F16_XSCALE : 0.0f;
    }
    return g;
  };
  // Chained launch (avd_gemm_chain.cu): out-proj -> fc1 -> fc2 -> next block's QKV in ONE persistent kernel with
  // row-group flags instead of three kernel boundaries, when every one of them is a single wave of tiles.
  const bool may_chain = options().gemm_chain != 0 && ws->chain_flags != nullptr && !ln && tc_attn &&
                         ws->f16_sat_count == nullptr;
  bool qkv_launched = false;                      // block l's QKV already ran as the tail of block l-1's chain
  for (int l = 0; l < w->n_layers; ++l) {
    const avd_block_weights& b = w->blocks[l];
    const float* r_in = (l == 0) ? x_in : ws->x;
    if (ln) {
      AVD_REQUIRE(b.ln1_w != nullptr && b.ln1_b != nullptr && b.ln2_w != nullptr && b.ln2_b != nullptr, "missing LayerNorm weights");
      AVD_TRY(launch_layernorm_act(r_in, b.ln1_w, b.ln1_b, M, d, w->ln_eps, 0, nullptr, ws->xa,
                                   nseg_arg(w->precision, false), st));
    }
    if (!qkv_launched) AVD_TRY(launch_gemm(qkv_args(l), st));
    qkv_launched = false;
    // a = softmax(q k^T / sqrt(hd)) v
    if (tc_attn) AVD_TRY(launch_attention_tc(ws->qkv, ws->att, S, n_seq, N, w->n_heads, hd, st, f16, ws->key_mask));
    else AVD_TRY(launch_attention_f32(static_cast<const float*>(ws->qkv), ws->att, S, n_seq, N, w->n_heads, hd, st,
                                      ws->key_mask));
    if (may_chain) {
      avd_gemm_args gs[4] = {oproj_args(l), fc1_args(l), fc2_args(l), avd_gemm_args{}};
      int n = 3;
      if (l + 1 < w->n_layers) gs[n++] = qkv_args(l + 1);
      const int rc = launch_gemm_chain(gs, n, ws->chain_flags, GEMM_CHAIN_ROW_TILES, st);
      if (rc < 0) return rc;
      if (rc == 1) { qkv_launched = n == 4; continue; }
    }
    AVD_TRY(launch_gemm(oproj_args(l), st));
    if (ln) AVD_TRY(launch_layernorm_act(ws->x, b.ln2_w, b.ln2_b, M, d, w->ln_eps, 0, nullptr, ws->xa,
                                         nseg_arg(w->precision, false), st));
    AVD_TRY(launch_gemm(fc1_args(l), st));
    const avd_gemm_args g2 = fc2_args(l);
    AVD_TRY(launch_gemm(g2, st));
    if (f16 && ws->f16_sat_count != nullptr) {           // range check of this block's 16-bit operands (not in the timed path)
      AVD_TRY(launch_f16_range_scan(ws->qkv, (int64_t)M * 3 * d, ws->f16_sat_count, st));
      AVD_TRY(launch_f16_range_scan(ws->att, (int64_t)M * d, ws->f16_sat_count, st));
      AVD_TRY(launch_f16_range_scan(ws->hmid, (int64_t)M * w->d_mlp, ws->f16_sat_count, st));
      if (g2.out_bf16 != last_out || is_f16(last_prec))
        AVD_TRY(launch_f16_range_scan(g2.out_bf16, (int64_t)M * d, ws->f16_sat_count, st));
    }
  }
  return 0;
}

// "chains" schedule (options().chains = C > 1, 16-bit single-segment modes): the n_seq sequences are independent, so they
// are split into C groups whose block sequences run on forked streams and join before the head.  Kernels of different
// groups then overlap on the device (one group's epilogue / launch latency under another group's mainloop) at the cost of
// narrower grids per launch.  Off by default; bench.py --chains measures it.
static int core_blocks(const avd_core_weights* w, const float* x_in, const void* xa_in, const float* ssq_in,
                       const avd_core_workspace* ws, int n_seq, int N, void* last_out, int last_prec, cudaStream_t st) {
  AVD_REQUIRE(w != nullptr && ws != nullptr, "null argument");
  const int d = w->d_model, S = nseg_of(w->precision);
  AVD_REQUIRE(S > 0, "bad precision");
  AVD_REQUIRE(d % 64 == 0 && w->n_heads > 0 && d % w->n_heads == 0, "bad d_model / n_heads");
  AVD_REQUIRE(w->n_layers > 0 && w->n_layers <= AVD_MAX_LAYERS, "bad layer count");
  AVD_REQUIRE(w->norm_kind == 0 || last_out == nullptr, "a layernorm core hands the fp32 residual stream to its final norm");
  int C = options().chains;
  if (C > 4) C = 4;
  const bool can_chain = C > 1 && S == 1 && n_seq % C == 0 && n_seq / C >= 2 && ws->key_mask == nullptr &&
                         (last_out == nullptr || nseg_of(last_prec) == 1);
  if (!can_chain) return core_blocks_rows(w, x_in, xa_in, ssq_in, ws, n_seq, N, last_out, last_prec, st);
  ChainStreams* cs = nullptr;
  AVD_TRY(chain_streams(&cs));
  const int seq_c = n_seq / C;
  const int64_t rows = (int64_t)seq_c * N, parts = d / 64;
  AVD_CUDA(cudaEventRecord(cs->fork, st));
  for (int c = 0; c < C; ++c) {
    cudaStream_t sc = c == 0 ? st : cs->side[c - 1];
    if (c > 0) AVD_CUDA(cudaStreamWaitEvent(sc, cs->fork, 0));
    const int64_t r0 = c * rows;
    avd_core_workspace wc = *ws;
    wc.x = ws->x + r0 * d;
    wc.xa = static_cast<uint8_t*>(ws->xa) + r0 * d * 2;
    wc.ssq = ws->ssq + r0 * parts;
    wc.qkv = static_cast<uint8_t*>(ws->qkv) + r0 * 3 * d * 2;
    wc.att = static_cast<uint8_t*>(ws->att) + r0 * d * 2;
    wc.hmid = static_cast<uint8_t*>(ws->hmid) + r0 * w->d_mlp * 2;
    wc.chain_flags = nullptr;                  // one flag block cannot serve concurrent groups
    void* lo = last_out != nullptr ? static_cast<uint8_t*>(last_out) + r0 * d * 2 : nullptr;
    AVD_TRY(core_blocks_rows(w, x_in + r0 * d, static_cast<const uint8_t*>(xa_in) + r0 * d * 2, ssq_in + r0 * parts, &wc,
                             seq_c, N, lo, last_prec, sc));
    if (c > 0) {
      AVD_CUDA(cudaEventRecord(cs->join[c - 1], sc));
      AVD_CUDA(cudaStreamWaitEvent(st, cs->join[c - 1], 0));
    }
  }
  return 0;
}

// ssq_in (optional): row sums of squares of h; then the input projection applies the RMSNorm row scale
// 1/(sqrt(ssq/rms_dim) + rms_eps) in its epilogue (the norm's `scale` vector is folded into w_in by the packer).
// h_a_scale: undoes the scale carried by h (1/AVD_F16_XSCALE when h is the pre-scaled fp16 residual stream; 0 = none).
static int head_forward(const avd_head_weights* w, const void* h, int M, void* act_a, void* act_b, float* y,
                        float* eps, const float* ssq_in, int ssq_parts, int rms_dim, float rms_eps, float h_a_scale,
                        cudaStream_t st) {
  AVD_REQUIRE(w != nullptr, "null head");
  AVD_REQUIRE(w->n_shared >= 0 && w->n_shared <= AVD_MAX_SHARED, "bad shared-layer count");
  AVD_REQUIRE(w->activation >= 0 && w->activation <= 3, "activation must be 0/1 (GELU), 2 (LeakyReLU 0.1) or 3 (ReLU)");
  const int S = nseg_of(w->precision);
  AVD_REQUIRE(S > 0, "bad precision");
  avd_gemm_args g{};
  // x = input_proj[m](h)                  (noise_heads.py:205)
  g.M = M; g.N = w->hidden; g.K = w->d_in; g.precision = w->precision;
  g.A = h; g.W = w->w_in; g.bias = w->b_in; g.out_bf16 = act_a; g.out_precision = w->precision;
  g.ssq_in = ssq_in; g.ssq_parts = ssq_parts; g.rms_dim = rms_dim; g.rms_eps = rms_eps;
  g.a_scale = h_a_scale;
  AVD_TRY(launch_gemm(g, st));
  void* cur = act_a;
  void* nxt = act_b;
  for (int i = 0; i < w->n_shared; ++i) {   // x = gelu(LN(Linear(x)))   (noise_heads.py:139-146, :208)
    g = avd_gemm_args{};
    g.M = M; g.N = w->hidden; g.K = w->hidden; g.precision = w->precision;
    g.A = cur; g.W = w->w_s[i]; g.bias = w->b_s[i]; g.out_f32 = y;
    AVD_TRY(launch_gemm(g, st));
    AVD_TRY(launch_layernorm_act(y, w->ln_w[i], w->ln_b[i], M, w->hidden, w->ln_eps, w->activation <= 1 ? 1 : w->activation,
                                 nullptr, nxt, nseg_arg(w->precision, false), st));
    void* t = cur; cur = nxt; nxt = t;
  }
  g = avd_gemm_args{};                       // eps = out_proj[m](x)   (noise_heads.py:219)
  g.M = M; g.N = w->d_out; g.K = w->hidden; g.precision = w->precision;
  g.A = cur; g.W = w->w_out; g.bias = w->b_out; g.out_f32 = eps;
  AVD_TRY(launch_gemm(g, st));
  return 0;
}

static BuildXArgs make_build_args(const avd_sampler* s, const SamplerLayout& L, int modality, bool cond, bool null) {
  BuildXArgs a{};
  a.modality = modality;
  a.B = s->B;
  a.C = modality == 0 ? s->Cv : s->Ca;
  a.T = s->T; a.H = s->H; a.W = s->W; a.tt = s->tt; a.th = s->th; a.tw = s->tw;
  a.L = s->Fa; a.length = s->chunk_len; a.stride = s->chunk_stride;
  a.n_tok = modality == 0 ? L.Nv : L.Na;
  a.d_tok = modality == 0 ? L.Dv : L.Da;
  a.d_adapt = L.d_adapt;
  a.tstep_dim = s->tstep_dim;
  a.d = L.d;
  a.N = L.N;
  a.row_off = modality == 0 ? L.row_off_v : L.row_off_a;
  a.n_seq_total = L.n_seq;
  a.write_cond = cond; a.write_null = null;
  a.nseg = nseg_arg(s->core.precision, true);
  a.ssq_parts = L.parts;
  a.M = L.M;
  return a;
}

}  // namespace avd

using namespace avd;

#define AVD_STREAM(s) static_cast<cudaStream_t>(s)

extern "C" {

const char* avd_last_error(void) { return g_last_error.c_str(); }
int avd_version(void) { return 100; }
int avd_nseg(int precision) { return nseg_of(precision); }

int avd_struct_size(int which) {
  switch (which) {
    case 0: return (int)sizeof(avd_cfg_ddim_args);
    case 1: return (int)sizeof(avd_gemm_args);
    case 2: return (int)sizeof(avd_block_weights);
    case 3: return (int)sizeof(avd_core_weights);
    case 4: return (int)sizeof(avd_core_workspace);
    case 5: return (int)sizeof(avd_head_weights);
    case 6: return (int)sizeof(avd_sampler);
    case 7: return (int)sizeof(avd_vae_weights);
    default: return -1;
  }
}

int avd_device_check(void) {
  int dev = 0;
  AVD_CUDA(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  AVD_CUDA(cudaGetDeviceProperties(&prop, dev));
  if (prop.major != 10) AVD_FAIL(std::string("device is sm_") + std::to_string(prop.major * 10 + prop.minor) +
                                 ", this library is built for sm_100a only");
  return 0;
}

int avd_prof_enable(int on) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof_on = on != 0;
  return 0;
}
int avd_prof_reset(void) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  for (auto& r : g_prof) { g_event_pool.push_back(r.beg); g_event_pool.push_back(r.end); }
  g_prof.clear();
  return 0;
}
int avd_prof_num_kinds(void) { return KIND_COUNT; }
const char* avd_prof_kind_name(int kind) { return (kind >= 0 && kind < KIND_COUNT) ? kKindNames[kind] : ""; }
int avd_prof_collect(int kind, double* total_ms, int64_t* launches) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  double tot = 0.0;
  int64_t n = 0;
  for (auto& r : g_prof) {
    if (r.kind != kind) continue;
    AVD_CUDA(cudaEventSynchronize(r.end));
    float ms = 0.0f;
    AVD_CUDA(cudaEventElapsedTime(&ms, r.beg, r.end));
    tot += ms;
    ++n;
  }
  if (total_ms) *total_ms = tot;
  if (launches) *launches = n;
  return 0;
}
int64_t avd_launch_count(void) { return g_launches; }
int avd_debug_gemm_timeout(unsigned int* out8) {
  AVD_REQUIRE(out8 != nullptr, "null output");
  return ps_debug_read(out8);
}

int avd_debug_trace(void* dev_u64, int64_t n_u64) {
#ifdef AVD_DEBUG_HOOKS
  g_dbg_trace = static_cast<unsigned long long*>(dev_u64);
  g_dbg_trace_slots = dev_u64 ? n_u64 : 0;
  return 0;
#else
  (void)dev_u64; (void)n_u64;
  AVD_FAIL("phase tracing is compiled out of this build (python -m multimodal_diffusion_b200.build --debug-hooks)");
#endif
}

int avd_check_errors(void) {
  unsigned int rec[256];
  AVD_TRY(ps_debug_read(rec));
  if (rec[0] != 0) {
    std::string msg = "persistent GEMM: " + std::to_string(rec[0]) + " barrier wait(s) timed out; first record {tag " +
                      std::to_string(rec[8]) + ", cta " + std::to_string(rec[9]) + ", thread " + std::to_string(rec[10]) +
                      ", counter " + std::to_string(rec[11]) + ", parity " + std::to_string(rec[12]) + ", N " +
                      std::to_string(rec[13]) + ", K " + std::to_string(rec[14]) + ", grid*1000+BN " +
                      std::to_string(rec[15]) + "}";
    AVD_FAIL(msg);
  }
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) AVD_FAIL(std::string("CUDA error pending on this device: ") + cudaGetErrorString(e));
  return 0;
}

int avd_set_option(const char* name, int value) {
  AVD_REQUIRE(name != nullptr, "null option name");
  for (const OptionDesc& d : kOptions)
    if (std::strcmp(d.name, name) == 0) {
      options().*(d.field) = value;
      return 0;
    }
  AVD_FAIL(std::string("unknown option '") + name + "'");
}
int avd_get_option(const char* name, int* value) {
  AVD_REQUIRE(name != nullptr && value != nullptr, "null argument");
  for (const OptionDesc& d : kOptions)
    if (std::strcmp(d.name, name) == 0) {
      *value = options().*(d.field);
      return 0;
    }
  AVD_FAIL(std::string("unknown option '") + name + "'");
}

// ---- tokenisers
int avd_tube_patch(const float* z, float* tok, int B, int C, int T, int H, int W, int t, int h, int w, void* stream) {
  AVD_REQUIRE(B >= 0 && C > 0 && T > 0 && H > 0 && W > 0, "bad latent shape");
  AVD_REQUIRE(t > 0 && h > 0 && w > 0 && T % t == 0 && H % h == 0 && W % w == 0, "tube sizes must divide latent dims");
  return launch_tube_patch(z, tok, B, C, T, H, W, t, h, w, false, AVD_STREAM(stream));
}
int avd_tube_unpatch(const float* tok, float* z, int B, int C, int T, int H, int W, int t, int h, int w, void* stream) {
  AVD_REQUIRE(B >= 0 && C > 0 && T > 0 && H > 0 && W > 0, "bad latent shape");
  AVD_REQUIRE(t > 0 && h > 0 && w > 0 && T % t == 0 && H % h == 0 && W % w == 0, "tube sizes must divide latent dims");
  return launch_tube_patch(tok, z, B, C, T, H, W, t, h, w, true, AVD_STREAM(stream));
}
int avd_chunk1d(const float* z, float* tok, int B, int C, int L, int length, int stride, void* stream) {
  AVD_REQUIRE(length > 0 && stride > 0 && L >= length, "chunk_1d needs 0 < length <= L and stride > 0");
  return launch_chunk1d(z, tok, B, C, L, length, stride, AVD_STREAM(stream));
}
int avd_unchunk1d(const float* tok, float* z, int B, int C, int N, int length, int stride, int L_out, void* stream) {
  AVD_REQUIRE(length > 0 && stride > 0 && N > 0 && L_out >= 0, "bad overlap-add shape");
  return launch_unchunk1d(tok, z, B, C, N, length, stride, L_out, AVD_STREAM(stream));
}

// ---- streaming driver stitching
int avd_crossfade_audio(const float* chunks, const float* w, float* out, int N, int L, int hop, void* stream) {
  return launch_crossfade_audio(chunks, w, out, N, L, hop, AVD_STREAM(stream));
}
int avd_overlap_add_win(const float* windows, const float* win, float* y, int rows, int N, int W, int stride,
                        float norm_floor, void* stream) {
  return launch_overlap_add_win(windows, win, y, rows, N, W, stride, norm_floor, AVD_STREAM(stream));
}
int avd_crossfade_video(const uint8_t* chunks, const float* w, uint8_t* out, int N, int T, int64_t frame_bytes, int hop,
                        void* stream) {
  return launch_crossfade_video(chunks, w, out, N, T, frame_bytes, hop, AVD_STREAM(stream));
}

// ---- schedules
int avd_timestep_embedding(const float* t, float* emb, int B, int dim, float max_period, void* stream) {
  AVD_REQUIRE(dim > 0 && max_period > 0, "bad embedding shape");
  return launch_timestep_embedding(t, emb, B, dim, max_period, AVD_STREAM(stream));
}
int avd_ddim_step(const float* x_t, const float* eps_hat, const float* noise, float* x_prev, const int64_t* t_now,
                  const int64_t* t_prev, const float* alpha_bar, int T_train, int B, int64_t per_batch, float eta,
                  void* stream) {
  AVD_REQUIRE(eta == 0.0f || noise != nullptr, "eta > 0 needs the noise tensor");
  AVD_REQUIRE(B <= 65535, "batch too large for one launch");
  return launch_ddim_step(x_t, eps_hat, noise, x_prev, t_now, t_prev, alpha_bar, T_train, B, per_batch, eta,
                          AVD_STREAM(stream));
}
int avd_philox_normal(float* out, uint64_t seed, int64_t clip0, int step, int B, int64_t per_clip, void* stream) {
  AVD_REQUIRE(out != nullptr || B * per_clip == 0, "null output");
  return launch_philox_normal(out, seed, clip0, step, B, per_clip, AVD_STREAM(stream));
}
int avd_f16_range_scan(const void* buf, int64_t n, unsigned int* counter, void* stream) {
  return launch_f16_range_scan(buf, n, counter, AVD_STREAM(stream));
}
int avd_cfg_untok_ddim(const avd_cfg_ddim_args* a, const float* eps_cond, const float* eps_null, const float* z_t,
                       float* z_prev, const float* alpha_bar, void* stream) {
  AVD_REQUIRE(a != nullptr, "null args");
  AVD_REQUIRE(a->t_now < a->T_train && a->t_prev < a->T_train, "timestep out of range");
  AVD_REQUIRE(a->eta >= 0.0f, "eta must be >= 0");
  if (a->target == AVD_TARGET_VIDEO) {
    AVD_REQUIRE(a->tt > 0 && a->th > 0 && a->tw > 0 && a->T % a->tt == 0 && a->H % a->th == 0 && a->W % a->tw == 0,
                "tube sizes must divide latent dims");
  } else {
    AVD_REQUIRE(a->target == AVD_TARGET_AUDIO && a->length > 0 && a->stride > 0 && a->Nt > 0, "bad audio args");
  }
  return launch_cfg_untok_ddim(*a, eps_cond, eps_null, z_t, z_prev, alpha_bar, 0, AVD_STREAM(stream));
}

// ---- operand preparation
int avd_pack_weight(const float* W, const float* col_scale, void* out_bf16, int N, int K, int precision, void* stream) {
  AVD_REQUIRE(nseg_of(precision) > 0, "bad precision");
  return launch_pack_weight(W, col_scale, out_bf16, N, K, nseg_arg(precision, false), AVD_STREAM(stream));
}
int avd_transpose_f32(const float* W, float* Wt, int N, int K, void* stream) {
  return launch_transpose(W, Wt, N, K, AVD_STREAM(stream));
}
int avd_rows_prepare(const float* x, void* xa_bf16, float* ssq, int M, int d, int precision, void* stream) {
  AVD_REQUIRE(nseg_of(precision) > 0, "bad precision");
  AVD_REQUIRE(d % 64 == 0, "d must be a multiple of 64");
  return launch_rows_prepare(x, xa_bf16, ssq, M, d, nseg_arg(precision, true), AVD_STREAM(stream));
}
int avd_rmsnorm_rows(const float* x, const float* scale, int n_seq, int N, int row_off, int n_rows, int d, float eps,
                     float* out_f32, void* out_bf16, int precision, void* stream) {
  AVD_REQUIRE(d % 4 == 0, "d must be a multiple of 4");
  AVD_REQUIRE(row_off >= 0 && n_rows >= 0 && row_off + n_rows <= N, "row window outside the sequence");
  AVD_REQUIRE(out_bf16 == nullptr || nseg_of(precision) > 0, "bad precision");
  return launch_rmsnorm_rows(x, scale, n_seq, N, row_off, n_rows, d, eps, out_f32, out_bf16,
                             out_bf16 ? nseg_arg(precision, false) : 0, AVD_STREAM(stream));
}
int avd_layernorm_act(const float* y, const float* w, const float* b, int M, int d, float eps, int apply_gelu,
                      float* out_f32, void* out_bf16, int precision, void* stream) {
  AVD_REQUIRE(d % 4 == 0, "d must be a multiple of 4");
  AVD_REQUIRE(out_bf16 == nullptr || nseg_of(precision) > 0, "bad precision");
  return launch_layernorm_act(y, w, b, M, d, eps, apply_gelu, out_f32, out_bf16, out_bf16 ? nseg_arg(precision, false) : 0,
                              AVD_STREAM(stream));
}

// ---- GEMM / attention
int avd_gemm(const avd_gemm_args* g, void* stream) {
  AVD_REQUIRE(g != nullptr, "null args");
  return launch_gemm(*g, AVD_STREAM(stream));
}
int avd_attention_masked(const void* qkv, int qkv_is_f32, void* out_bf16, int out_precision, int n_seq, int N,
                         int n_heads, int head_dim, const unsigned int* key_mask, void* stream) {
  AVD_REQUIRE(nseg_of(out_precision) > 0, "bad precision");
  if (qkv_is_f32)
    return launch_attention_f32(static_cast<const float*>(qkv), out_bf16, nseg_of(out_precision), n_seq, N, n_heads,
                                head_dim, AVD_STREAM(stream), key_mask);
  return launch_attention_tc(qkv, out_bf16, nseg_of(out_precision), n_seq, N, n_heads, head_dim, AVD_STREAM(stream),
                             is_f16(out_precision), key_mask);
}
int avd_attention(const void* qkv, int qkv_is_f32, void* out_bf16, int out_precision, int n_seq, int N, int n_heads,
                  int head_dim, void* stream) {
  return avd_attention_masked(qkv, qkv_is_f32, out_bf16, out_precision, n_seq, N, n_heads, head_dim, nullptr, stream);
}

// ---- core / head
int avd_mmdt_blocks(const avd_core_weights* w, const float* x_in, const void* xa_in, const float* ssq_in,
                    const avd_core_workspace* ws, int n_seq, int N, void* stream) {
  AVD_REQUIRE(ws != nullptr, "null workspace");
  SkipKindsScope sk(ws->profile_skip_kinds);
  return core_blocks(w, x_in, xa_in, ssq_in, ws, n_seq, N, nullptr, 0, AVD_STREAM(stream));
}
int avd_head_forward(const avd_head_weights* w, const void* h, int M, void* act_a, void* act_b, float* y, float* eps,
                     void* stream) {
  // fp16 mode: h is a residual-stream operand as avd_rows_prepare writes it (pre-scaled by AVD_F16_XSCALE)
  return head_forward(w, h, M, act_a, act_b, y, eps, nullptr, 0, 0, 0.0f,
                      (w != nullptr && is_f16(w->precision)) ? 1.0f / AVD_F16_XSCALE : 0.0f, AVD_STREAM(stream));
}

// ---- sampler
int64_t avd_sampler_workspace_bytes(const avd_sampler* s) {
  SamplerLayout L;
  if (sampler_layout(s, &L) != 0) return -1;
  return L.total;
}

int avd_sample_prepare(const avd_sampler* s, const float* z_prompt, void* stream) {
  SamplerLayout L;
  AVD_TRY(sampler_layout(s, &L));
  AVD_REQUIRE(s->workspace != nullptr && s->workspace_bytes >= L.total, "workspace too small");
  SkipKindsScope sk(s->profile_skip_kinds);
  cudaStream_t st = AVD_STREAM(stream);
  uint8_t* ws = static_cast<uint8_t*>(s->workspace);
  // null-branch prompt rows are all-zero rows that stay in the sequence (sample_clip.py:338 / 377)
  AVD_CUDA(cudaMemsetAsync(ws + L.x_init, 0, (size_t)(L.xa_init - L.x_init), st));
  AVD_CUDA(cudaMemsetAsync(ws + L.xa_init, 0, (size_t)(L.ssq_init - L.xa_init), st));
  AVD_CUDA(cudaMemsetAsync(ws + L.ssq_init, 0, (size_t)(L.x - L.ssq_init), st));
  AVD_CUDA(cudaMemsetAsync(ws + L.chain_flags, 0, GEMM_CHAIN_FLAG_BYTES, st));   // (also clears what a failed launch left)
  const int prompt_mod = s->target == AVD_TARGET_VIDEO ? 1 : 0;   // A2V: audio prompt; V2A: video prompt
  BuildXArgs a = make_build_args(s, L, prompt_mod, /*cond=*/true, /*null=*/false);
  const float* wt = prompt_mod == 0 ? s->wv_t : s->wa_t;
  const float* bias = prompt_mod == 0 ? s->bv : s->ba;
  return launch_build_x(a, z_prompt, wt, bias, s->temb_zero, reinterpret_cast<float*>(ws + L.x_init),
                        ws + L.xa_init, reinterpret_cast<float*>(ws + L.ssq_init), st,
                        prompt_mod == 0 ? s->add_v : s->add_a, s->tadd_zero);
}

int avd_sample_step(const avd_sampler* s, int step, float* z, float* eps_tap, void* stream) {
  SamplerLayout L;
  AVD_TRY(sampler_layout(s, &L));
  AVD_REQUIRE(s->workspace != nullptr && s->workspace_bytes >= L.total, "workspace too small");
  AVD_REQUIRE(step >= 0 && step < s->steps, "step out of range");
  AVD_REQUIRE(s->eta >= 0.0f, "eta must be >= 0");
  SkipKindsScope sk(s->profile_skip_kinds);
  cudaStream_t st = AVD_STREAM(stream);
  uint8_t* ws = static_cast<uint8_t*>(s->workspace);
  float* x_init = reinterpret_cast<float*>(ws + L.x_init);
  float* ssq_init = reinterpret_cast<float*>(ws + L.ssq_init);
  const bool ln = s->core.norm_kind == 1;
  // (1) target rows of the joint sequence, identical in the cond and null sequences
  const int tgt_mod = s->target == AVD_TARGET_VIDEO ? 0 : 1;
  BuildXArgs a = make_build_args(s, L, tgt_mod, true, true);
  const float* temb_row = s->tstep_dim > 0 ? s->temb_steps + (int64_t)step * s->tstep_dim : nullptr;
  const float* tadd_row = s->tadd_steps != nullptr ? s->tadd_steps + (int64_t)step * L.d_adapt : nullptr;
  const void* w_pack = tgt_mod == 0 ? s->wv_pack : s->wa_pack;
  const int bx_opt = options().bx_wide;
  const bool bx_tc = w_pack != nullptr && L.Dt % 64 == 0 && L.d_adapt % 32 == 0 &&
                     (s->adapt_precision == AVD_PREC_BF16X3 || s->adapt_precision == AVD_PREC_FP32) &&
                     (bx_opt == 2 || bx_opt < 0);   // measured faster at every batch (B = 1: 289 -> 258 us/step)
  if (bx_tc)
    AVD_TRY(launch_build_x_tc(a, z, w_pack, s->adapt_precision, tgt_mod == 0 ? s->bv : s->ba, temb_row, x_init,
                              ws + L.xa_init, ssq_init, ws + L.bx_tok, reinterpret_cast<float*>(ws + L.bx_y), st,
                              tgt_mod == 0 ? s->add_v : s->add_a, tadd_row));
  else
    AVD_TRY(launch_build_x(a, z, tgt_mod == 0 ? s->wv_t : s->wa_t, tgt_mod == 0 ? s->bv : s->ba, temb_row, x_init,
                           ws + L.xa_init, ssq_init, st, tgt_mod == 0 ? s->add_v : s->add_a, tadd_row));
  // (2) 2B co-batched sequences through the blocks
  avd_core_workspace cw{};                       // key_mask = NULL: the sampler never pads (sample_clip.py:341)
  cw.x = reinterpret_cast<float*>(ws + L.x);
  cw.xa = ws + L.xa;
  cw.ssq = reinterpret_cast<float*>(ws + L.ssq);
  cw.qkv = ws + L.qkv;
  cw.att = ws + L.att;
  cw.hmid = ws + L.hmid;
  cw.f16_sat_count = s->f16_sat_count;
  cw.chain_flags = reinterpret_cast<unsigned int*>(ws + L.chain_flags);
  float* eps = reinterpret_cast<float*>(ws + L.eps);
  if (!ln) {
    AVD_TRY(core_blocks(&s->core, x_init, ws + L.xa_init, ssq_init, &cw, L.n_seq, L.N, ws + L.h, s->head.precision, st));
    // (3)+(4) final RMSNorm folded into the input projection of the target modality's noise head, over all rows (F9:
    // the non-target head is never evaluated; the target rows are selected by the CFG/DDIM kernel below)
    AVD_TRY(head_forward(&s->head, ws + L.h, (int)L.M, ws + L.act_a, ws + L.act_b, reinterpret_cast<float*>(ws + L.y), eps,
                         cw.ssq, L.parts, L.d, s->core.rms_eps,
                         is_f16(s->head.precision) ? 1.0f / AVD_F16_XSCALE : 0.0f, st));
  } else {
    // norm="layernorm" (mmdt.py:45-46): the final nn.LayerNorm is its own pass over the fp32 residual stream and writes
    // the head's (unscaled) operand; the head's input projection then needs no row scale
    AVD_REQUIRE(s->core.final_ln_w != nullptr && s->core.final_ln_b != nullptr, "missing final_norm.weight / .bias");
    AVD_TRY(core_blocks(&s->core, x_init, ws + L.xa_init, ssq_init, &cw, L.n_seq, L.N, nullptr, 0, st));
    AVD_TRY(launch_layernorm_act(cw.x, s->core.final_ln_w, s->core.final_ln_b, (int)L.M, L.d, s->core.ln_eps, 0, nullptr,
                                 ws + L.h, nseg_arg(s->head.precision, false), st));
    AVD_TRY(head_forward(&s->head, ws + L.h, (int)L.M, ws + L.act_a, ws + L.act_b, reinterpret_cast<float*>(ws + L.y), eps,
                         nullptr, 0, 0, 0.0f, 0.0f, st));
  }
  if (is_f16(s->head.precision) && s->f16_sat_count != nullptr) {
    AVD_TRY(launch_f16_range_scan(ws + L.act_a, L.M * s->head.hidden, s->f16_sat_count, st));
    if (s->head.n_shared > 0) AVD_TRY(launch_f16_range_scan(ws + L.act_b, L.M * s->head.hidden, s->f16_sat_count, st));
  }
  const int row_off = s->target == AVD_TARGET_VIDEO ? L.row_off_v : L.row_off_a;
  const float* eps_c = eps + (int64_t)row_off * L.Dt;                          // rows of clip b: sequence b
  const float* eps_n = eps + ((int64_t)s->B * L.N + row_off) * L.Dt;           //                 sequence B + b
  if (eps_tap != nullptr)      // compact (eps_cond ; eps_null) [2B, Nt, Dt]
    AVD_CUDA(cudaMemcpy2DAsync(eps_tap, (size_t)L.Nt * L.Dt * 4, eps_c, (size_t)L.N * L.Dt * 4, (size_t)L.Nt * L.Dt * 4,
                               (size_t)L.n_seq, cudaMemcpyDeviceToDevice, st));
  // (5) CFG + un-tokenise + DDIM
  avd_cfg_ddim_args c{};
  c.target = s->target; c.B = s->B;
  c.C = s->target == AVD_TARGET_VIDEO ? s->Cv : s->Ca;
  c.T = s->T; c.H = s->H; c.W = s->W; c.tt = s->tt; c.th = s->th; c.tw = s->tw;
  c.L = s->Fa; c.length = s->chunk_len; c.stride = s->chunk_stride; c.Nt = L.Na;
  c.guidance = s->guidance;
  c.t_now = (int)s->timesteps_host[step];
  c.t_prev = (int)s->timesteps_host[step + 1];
  c.T_train = s->T_train;
  c.eta = s->eta; c.step = step; c.seed = s->seed; c.clip0 = s->clip_index0;
  AVD_REQUIRE(c.t_now >= 0 && c.t_now < s->T_train && c.t_prev < s->T_train, "timestep out of range");
  return launch_cfg_untok_ddim(c, eps_c, eps_n, z, z, s->alpha_bar, (int64_t)L.N * L.Dt, st);
}

int avd_sample_loop(const avd_sampler* s, const float* z_prompt, float* z, void* stream) {
  AVD_TRY(avd_sample_prepare(s, z_prompt, stream));
  for (int i = 0; i < s->steps; ++i) AVD_TRY(avd_sample_step(s, i, z, nullptr, stream));
  return 0;
}

}  // extern "C"
