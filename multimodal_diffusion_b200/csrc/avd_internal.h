// avd_internal.h -- declarations shared by the translation units of libavdiff_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <cuda.h>
#include <stdint.h>
#include <string>

#include "../../include/avdiff_b200.h"

namespace avd {

// ---- error plumbing -----------------------------------------------------------------------------
void set_error(const std::string& msg);
int fail(const char* file, int line, const std::string& msg);

#define AVD_FAIL(msg) return ::avd::fail(__FILE__, __LINE__, (msg))
#define AVD_REQUIRE(cond, msg) \
  do {                         \
    if (!(cond)) AVD_FAIL(std::string("requirement failed: " #cond " -- ") + (msg)); \
  } while (0)
#define AVD_CUDA(expr)                                                                        \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess) AVD_FAIL(std::string(#expr " -> ") + cudaGetErrorString(_e));      \
  } while (0)
#define AVD_TRY(expr)          \
  do {                         \
    int _rc = (expr);          \
    if (_rc != 0) return _rc;  \
  } while (0)

// ---- launch accounting / profiler ---------------------------------------------------------------
enum Kind : int {
  KIND_TOKENIZE = 0,
  KIND_BUILD_X,
  KIND_GEMM,
  KIND_ATTN_TC,
  KIND_ATTN_F32,
  KIND_NORM,
  KIND_LN_ACT,
  KIND_CFG_DDIM,
  KIND_PACK,
  KIND_MISC,
  KIND_COUNT
};

struct LaunchScope {  // brackets one kernel launch; counts it and (optionally) times it with CUDA events
  LaunchScope(Kind k, cudaStream_t s);
  ~LaunchScope();
  Kind kind;
  cudaStream_t stream;
  int slot;
};

inline int nseg_of(int precision) {
  return (precision == 1 || precision == 2) ? 1 : precision == 3 ? 2 : precision == 6 ? 3 : -1;
}
inline bool is_f16(int precision) { return precision == AVD_PREC_F16; }
// `nseg` argument of the operand-writing kernels (avd_elementwise.cu store_split): > 0 bf16 segments; < 0 ONE fp16
// value: NSEG_F16_X carries the residual-stream scale AVD_F16_XSCALE, NSEG_F16 is unscaled (weights, normalised values)
constexpr int NSEG_F16_X = -1;
constexpr int NSEG_F16 = -2;
inline int nseg_arg(int precision, bool residual_stream) {
  return is_f16(precision) ? (residual_stream ? NSEG_F16_X : NSEG_F16) : nseg_of(precision);
}

// ---- kernels (launchers) ------------------------------------------------------------------------
// elementwise / data movement (avd_elementwise.cu)
int launch_tube_patch(const float* src, float* dst, int B, int C, int T, int H, int W, int t, int h, int w, bool inverse,
                      cudaStream_t s);
int launch_chunk1d(const float* z, float* tok, int B, int C, int L, int length, int stride, cudaStream_t s);
int launch_unchunk1d(const float* tok, float* z, int B, int C, int N, int length, int stride, int L_out,
                     cudaStream_t s);
int launch_crossfade_audio(const float* chunks, const float* w, float* out, int N, int L, int hop, cudaStream_t s);
int launch_overlap_add_win(const float* chunks, const float* w, float* out, int rows, int N, int L, int hop,
                           float norm_floor, cudaStream_t s);
int launch_crossfade_video(const uint8_t* chunks, const float* w, uint8_t* out, int N, int T, int64_t F, int hop,
                           cudaStream_t s);
int launch_timestep_embedding(const float* t, float* emb, int B, int dim, float max_period, cudaStream_t s);
int launch_ddim_step(const float* x_t, const float* eps, const float* noise, float* x_prev, const int64_t* t_now,
                     const int64_t* t_prev, const float* alpha_bar, int T_train, int B, int64_t per_batch, float eta,
                     cudaStream_t s);
// eps_clip_stride: elements between the eps rows of consecutive clips (0 = compact [B, Nt, Dt])
int launch_cfg_untok_ddim(const avd_cfg_ddim_args& a, const float* eps_c, const float* eps_n, const float* z_t,
                          float* z_prev, const float* alpha_bar, int64_t eps_clip_stride, cudaStream_t s);
int launch_pack_weight(const float* W, const float* col_scale, void* out, int N, int K, int nseg, cudaStream_t s);
int launch_transpose(const float* W, float* Wt, int N, int K, cudaStream_t s);
int launch_rows_prepare(const float* x, void* xa, float* ssq, int M, int d, int nseg, cudaStream_t s);
int launch_rmsnorm_rows(const float* x, const float* scale, int n_seq, int N, int row_off, int n_rows, int d, float eps,
                        float* out_f32, void* out_bf16, int nseg, cudaStream_t s);
int launch_layernorm_act(const float* y, const float* w, const float* b, int M, int d, float eps, int gelu,
                         float* out_f32, void* out_bf16, int nseg, cudaStream_t s);

// Fused input builder: tokenise a latent, adapter GEMV, concat timestep embedding; writes rows of the joint
// sequence into x (fp32), xa (bf16 segments) and ssq.  See avd_elementwise.cu.
struct BuildXArgs {
  int modality;            // 0 video latent [B,C,T,H,W] with tube, 1 audio latent [B,C,L] with chunk
  int B;
  int C, T, H, W, tt, th, tw;
  int L, length, stride;
  int n_tok;               // tokens per clip of this modality
  int d_tok;               // raw token width (C*t*h*w or C*length)
  int d_adapt;             // adapter output width (d - tstep_dim)
  int tstep_dim;
  int d;                   // d_adapt + tstep_dim
  int N;                   // joint sequence length
  int row_off;             // first row of this modality inside a sequence
  int n_seq_total;         // 2B
  int write_cond, write_null;  // write rows of sequence b (cond) and/or B+b (null)
  int nseg;
  int ssq_parts;
  int64_t M;               // n_seq_total * N
};
// add_rows: optional [n_tok][d_adapt] rows added after the adapter (positional + modality embeddings); tadd: optional
// [d_adapt] additive time embedding of this step
int launch_build_x(const BuildXArgs& a, const float* z, const float* wt, const float* bias, const float* temb,
                   float* x, void* xa, float* ssq, cudaStream_t s, const float* add_rows = nullptr,
                   const float* tadd = nullptr);
// the same through the tcgen05 GEMM (w_pack: adapter weight as split-bf16 segments [nseg][d_adapt][d_tok], avd_pack_weight);
// tok_scratch: nseg * B*n_tok*d_tok bf16, y_scratch: B*n_tok*d_adapt fp32
int launch_build_x_tc(const BuildXArgs& a, const float* z, const void* w_pack, int w_precision, const float* bias,
                      const float* temb, float* x, void* xa, float* ssq, void* tok_scratch, float* y_scratch,
                      cudaStream_t s, const float* add_rows = nullptr, const float* tadd = nullptr);
int launch_philox_normal(float* out, uint64_t seed, int64_t clip0, int step, int B, int64_t per_clip, cudaStream_t s);
int launch_f16_range_scan(const void* buf, int64_t n, unsigned int* counter, cudaStream_t s);

// tcgen05 GEMM (avd_gemm.cu)
int launch_gemm(const avd_gemm_args& g, cudaStream_t s);
// chained launch of dependent one-wave GEMMs (avd_gemm_chain.cu): 1 launched, 0 not chainable (launch them one by one),
// < 0 error.  flags: GEMM_CHAIN_MAX x flags_row_tiles zero-initialised counters (avd_core_workspace.chain_flags).
int launch_gemm_chain(const avd_gemm_args* gs, int n, unsigned int* flags, int flags_row_tiles, cudaStream_t s);
constexpr int GEMM_CHAIN_MAX = 4;            // phases per chained launch
constexpr int GEMM_CHAIN_ROW_TILES = 160;    // row tiles a flag block provides for (>= SM count: one wave)
constexpr int GEMM_CHAIN_FLAG_BYTES = GEMM_CHAIN_MAX * GEMM_CHAIN_ROW_TILES * 4;
static_assert(GEMM_CHAIN_FLAG_BYTES == AVD_CHAIN_FLAG_BYTES, "include/avdiff_b200.h states the size of the flag block");

// Measurement aid scoped to ONE avd_sample_* call (avd_sampler.profile_skip_kinds): bit k set -> launches of Kind k are
// skipped while that call runs on this thread.  There is no process-wide switch.
extern thread_local int t_skip_kinds;
inline bool kind_skipped(int k) { return (t_skip_kinds >> k) & 1; }
struct SkipKindsScope {
  explicit SkipKindsScope(int mask) : prev(t_skip_kinds) { t_skip_kinds = mask; }
  ~SkipKindsScope() { t_skip_kinds = prev; }
  int prev;
};

// Tuning / A-B switches.  Every AVD_* environment variable is read ONCE, when the library is first used; afterwards the
// values only change through avd_set_option (tools and tests), never through the environment.
struct Options {
  int pdl;             // AVD_PDL            programmatic dependent launch between the kernels of a step (default 1)
  int gemm_bn;         // AVD_GEMM_BN        force a tile width (0 = model)
  int gemm_v;          // AVD_GEMM_V         3 = one-tile kernel, 4 = persistent kernel, 0 = by wave count
  int gemm_generic;    // AVD_GEMM_GENERIC   1 = generic epilogue instead of the specialised ones
  int gemm_dbg;        // AVD_GEMM_DBG       debug-hooks builds: epilogue timing experiments (bit 0 no tcgen05.ld, bit 1 no global stores)
  int gemm_chain;      // AVD_GEMM_CHAIN     1 = out-proj -> fc1 -> fc2 -> next QKV as one chained persistent launch (experimental,
                       //                    measured slower than four PDL-linked launches: DESIGN.md section 5)
  int gemm_pair;       // AVD_GEMM_PAIR      cta_group::2 variant of the one-tile kernel (0 off, 1 specialised, 2 all)
  int ps_max_stages;   // AVD_PS_MAX_STAGES  ring depth cap of the persistent kernel (0 = its maximum)
  int ps_epi_cycles;   // AVD_PS_EPI_CYCLES  epilogue constant of the tile-width model (0 = built in)
  int ps_pair;         // AVD_PS_PAIR        persistent kernel on CTA pairs (cta_group::2, avd_gemm_ps2.cu): 0 off, 1 (default) wherever
                       //                    the tile-width model picks 256-wide tiles, 2 for every GEMM whose N is a multiple of 256
  int ps_no256;        // AVD_PS_NO256       1 = no 256-wide column-split tiles
  int attn_poly;       // AVD_ATTN_POLY      polynomial exp2 share of the multi-block attention path
  int attn_small;      // AVD_ATTN_SMALL     1 (default) = two-CTAs-per-SM kernel (attn_small_kernel) for N <= 136 when there are more
                       //                    (sequence, head) pairs than SMs; 2 = always when N <= 136; 0 = never
  int attn_fa;         // AVD_ATTN_FA        multi-block attention (N > 192) with P / O in tensor memory (avd_attn_fa.cu): 1 (default)
                       //                    two softmax warps per row group, 2 one thread per row, 3 / 4 S double-buffered (96-key
                       //                    blocks; thread per row / two warps per row group) -- all measured slower than 1;
                       //                    0 = attn_tc_kernel's two-tile path
  int attn_pipe;       // AVD_ATTN_PIPE      1 (default) = pipelined multi-block attention kernel (attn_pipe_kernel) for N > 192
  int attn_one;        // AVD_ATTN_ONE       0 = multi-block attention kernel even for N <= 192
  int chains;          // AVD_CHAINS         independent sequence groups of a sampler step run on forked streams (1 = off)
  int bx_wide;         // AVD_BX_WIDE        -1 = tensor-core adapter when packed weights are given, else by size; 0 / 1 = the
                       //                    4- / 32-token FMA build_x kernel; 2 = tensor-core route
};
Options& options();
bool pdl_enabled();

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) once per (device, kernel): the attribute is per device
int ensure_dyn_smem(const void* fn, int bytes);
extern unsigned long long* g_dbg_trace;   // debug phase trace of the GEMM kernel (avd_debug_trace; needs a build with AVD_DEBUG_HOOKS)
extern int64_t g_dbg_trace_slots;

// attention (avd_attn.cu)
int launch_attention_tc(const void* qkv_bf16, void* out, int out_nseg, int n_seq, int N, int n_heads, int hd,
                        cudaStream_t s, bool f16 = false, const uint32_t* kmask = nullptr);
int launch_attention_f32(const float* qkv, void* out, int out_nseg, int n_seq, int N, int n_heads, int hd,
                         cudaStream_t s, const uint32_t* kmask = nullptr);

// ---- TMA descriptor cache (avd_api.cu) ----------------------------------------------------------
// bf16 tensor [d2][d1][d0] (d0 contiguous), box [1][box1][box0], SWIZZLE_128B (box0 * 2 bytes must be 128).
int get_tensor_map_3d(const void* base, int elem_bytes, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t stride1_bytes,
                      uint64_t stride2_bytes, uint32_t box0, uint32_t box1, CUtensorMap* out);
int get_tensor_map_bf16_3d(const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t stride1_bytes,
                           uint64_t stride2_bytes, uint32_t box0, uint32_t box1, CUtensorMap* out);

}  // namespace avd
