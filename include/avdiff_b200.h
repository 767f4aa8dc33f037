/*
 * avdiff_b200.h -- C ABI of libavdiff_b200.so: the B200 (sm_100a) implementation of avdiff's
 * DDIM + classifier-free-guidance sampling hot path.
 *
 * The reference (mauruszach/multimodal_diffusion, "avdiff") is pure Python/PyTorch and has no FFI
 * layer; its "plugin surface" for this path is a set of Python callables.  Each entry point below
 * names the reference callable it stands in for (file:line relative to the reference root).  The
 * Python package `multimodal_diffusion_b200.avdiff` binds these with ctypes and re-exposes them
 * under the reference's own module/function names.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in `_host`;
 *   - the caller owns and allocates every buffer, including workspaces; the library keeps no state
 *     besides a cache of TMA descriptors keyed by (pointer, shape) and the last error string;
 *   - `stream` is a `cudaStream_t` passed as `void*`; all work is enqueued on it, nothing
 *     synchronises, everything is CUDA-graph capturable;
 *   - return value 0 = success, negative = error (message via avd_last_error());
 *   - no CPU fallback exists anywhere in this library.
 *
 * Precision modes for the tensor-core GEMMs (bf16 operands, fp32 accumulation in TMEM):
 *   AVD_PREC_BF16   (1)  plain bf16 operands;
 *   AVD_PREC_F16    (2)  plain fp16 operands (3 more mantissa bits than bf16 at the same tcgen05 kind::f16 rate;
 *                        operands that may exceed the fp16 range are pre-scaled, see AVD_F16_XSCALE);
 *   AVD_PREC_BF16X3 (3)  operands split hi+lo, 3 partial products  (~16 mantissa bits);
 *   AVD_PREC_FP32   (6)  operands split hi+mid+lo, 6 partial products (~fp32 products).
 * An activation or weight matrix for mode p is stored as avd_nseg(p) bf16 matrices back to back
 * ("segments": hi[,mid],lo).
 */
#ifndef AVDIFF_B200_H_
#define AVDIFF_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AVD_PREC_BF16 1
#define AVD_PREC_F16 2
/* AVD_PREC_F16 operands that hold the residual stream x (the A operand of the QKV and fc1 GEMMs) carry this fixed
 * power-of-two scale so that |x| up to 1.6e7 stays finite in fp16; the consuming GEMM undoes it through a_scale. */
#define AVD_F16_XSCALE 0.00390625f
#define AVD_PREC_BF16X3 3
#define AVD_PREC_FP32 6

#define AVD_TARGET_VIDEO 0 /* A2V: audio prompt, video generated (sample_clip.py:354-393) */
#define AVD_TARGET_AUDIO 1 /* V2A: video prompt, audio generated (sample_clip.py:312-352) */

#define AVD_MAX_LAYERS 64
#define AVD_MAX_SHARED 8

/* ---------------------------------------------------------------- housekeeping */
const char* avd_last_error(void);
int avd_version(void);
int avd_nseg(int precision);                 /* 1 -> 1, 3 -> 2, 6 -> 3, else -1 */
int avd_device_check(void);                  /* 0 iff the current device is sm_100 */
/* sizeof() of the ABI structs as this library was compiled (binding self-check):
 * 0 avd_cfg_ddim_args, 1 avd_gemm_args, 2 avd_block_weights, 3 avd_core_weights, 4 avd_core_workspace,
 * 5 avd_head_weights, 6 avd_sampler, 7 avd_vae_weights; -1 for any other id */
int avd_struct_size(int which);

/* Per-launch profiler (CUDA events on `stream` around every kernel this library launches).
 * Used by bench.py for the roofline object.  When the launches are being captured into a CUDA graph the events
 * become event-record nodes, so one replay of that graph times every kernel on the device back to back. */
int avd_prof_enable(int on);
int avd_prof_reset(void);
int avd_prof_num_kinds(void);
const char* avd_prof_kind_name(int kind);
/* synchronises the recorded events; returns total ms and launch count of `kind` */
int avd_prof_collect(int kind, double* total_ms, int64_t* launches);
int64_t avd_launch_count(void);              /* kernels launched by this library so far */
/* Debug only, compiled in with -DAVD_DEBUG_HOOKS (`python -m multimodal_diffusion_b200.build --debug-hooks`); the default
 * build returns -1.  GEMM CTAs write 8 x (globaltimer ns, clock64) phase stamps into dev_u64[cta*16 ..] when the grid
 * fits n_u64; NULL switches it off. */
int avd_debug_trace(void* dev_u64, int64_t n_u64);
/* out256[0] = number of barrier waits of the persistent GEMM that timed out since the last call (each of them also
 * traps, so the launch fails), then from out256[8] up to 30 records {tag, cta, thread, counter, parity, N, K,
 * grid*1000+BN}; the call clears them. */
int avd_debug_gemm_timeout(unsigned int* out256);
/* Asynchronous failures.  A device-side protocol timeout (bounded mbarrier / flag waits) traps, which fails the launch;
 * it also leaves a record in host-mapped memory.  avd_check_errors() is the host's hard-error check: it returns 0 when
 * no record exists and the device has no sticky error, else -1 with the details in avd_last_error().  It does not
 * synchronise: call it after the caller's own stream/device synchronise (FusedSampler does). */
int avd_check_errors(void);
/* Tuning switches (tile width, kernel variant, PDL ...).  Each has an AVD_<NAME> environment variable that is read ONCE
 * when the library is first used; afterwards only these calls change them.  Unknown name: -1. */
int avd_set_option(const char* name, int value);
int avd_get_option(const char* name, int* value);

/* ---------------------------------------------------------------- K1: tokenisers (bit-exact fp32 moves) */
/* ops.tube_patch_video (avdiff/utils/ops.py:100-119): z[B,C,T,H,W] -> tok[B,(T/t)(H/h)(W/w),C*t*h*w] */
int avd_tube_patch(const float* z, float* tok, int B, int C, int T, int H, int W, int t, int h, int w, void* stream);
/* ops.tube_unpatch_video (ops.py:122-144) */
int avd_tube_unpatch(const float* tok, float* z, int B, int C, int T, int H, int W, int t, int h, int w, void* stream);
/* ops.chunk_1d + permute (ops.py:17-45, sample_clip.py:184-188, tokenizers.py:83-88):
 * z[B,C,L] -> tok[B,N,C*length], N = (L-length)/stride+1 */
int avd_chunk1d(const float* z, float* tok, int B, int C, int L, int length, int stride, void* stream);
/* ops.overlap_add_1d + crop/zero-pad (ops.py:48-93, sample_clip.py:191-215, tokenizers.py:90-114):
 * tok[B,N,C*length] -> z[B,C,L_out]; windows summed in order i=0..N-1, divided by max(count,1e-8) */
int avd_unchunk1d(const float* tok, float* z, int B, int C, int N, int length, int stride, int L_out, void* stream);

/* ---------------------------------------------------------------- streaming driver: stitching
 * stream_infer.crossfade_audio (avdiff/models/infer/stream_infer.py:82-113): chunks[N,L] fp32, one fade window w[L]
 * (all ones for fade_s <= 0; the cosine fade otherwise, computed by the caller with the reference's numpy formula) ->
 * out[(N-1)*hop + L] = sum_i chunk_i*w / max(sum_i w, 1e-6), windows added in order i = 0..N-1.  Bit-exact. */
int avd_crossfade_audio(const float* chunks, const float* w, float* out, int N, int L, int hop, void* stream);
/* ops.overlap_add_1d with a window (avdiff/utils/ops.py:82-92, apply_hann=True): windows[rows,N,W] * win[W] summed in
 * order i = 0..N-1 into y[rows,(N-1)*stride+W], divided by max(sum of win, norm_floor) (1e-8 in ops.py).  Bit-exact;
 * win is computed by the caller (torch.hann_window).  rows <= 65535. */
int avd_overlap_add_win(const float* windows, const float* win, float* y, int rows, int N, int W, int stride,
                        float norm_floor, void* stream);
/* stream_infer.crossfade_video (stream_infer.py:116-143): chunks[N,T,frame_bytes] uint8 (frame_bytes = H*W*3), one
 * weight per frame w[T] -> out[(N-1)*hop + T, frame_bytes] uint8 = trunc(clip(sum_i (u8/255)*w / max(sum w,1e-6),0,1)*255).
 * Bit-exact. */
int avd_crossfade_video(const uint8_t* chunks, const float* w, uint8_t* out, int N, int T, int64_t frame_bytes, int hop,
                        void* stream);

/* ---------------------------------------------------------------- schedules */
/* su.timestep_embedding (avdiff/utils/schedule_utils.py:64-86): t[B] (float) -> emb[B,dim] = [cos | sin] (+0 pad if odd) */
int avd_timestep_embedding(const float* t, float* emb, int B, int dim, float max_period, void* stream);
/* su.ddim_step (schedule_utils.py:146-200) / ModalitySchedule.ddim_step (models/schedules.py:65-74).
 * x_t, eps_hat, x_prev: [B, per_batch] fp32; t_now/t_prev: int64 [B]; alpha_bar: fp32 [T_train];
 * noise: the z ~ N(0,1) tensor of the eta>0 branch (may be NULL when eta == 0). */
int avd_ddim_step(const float* x_t, const float* eps_hat, const float* noise, float* x_prev, const int64_t* t_now,
                  const int64_t* t_prev, const float* alpha_bar, int T_train, int B, int64_t per_batch, float eta,
                  void* stream);
/* The z ~ N(0,1) of ddim_step's eta > 0 branch (schedule_utils.py:197, torch.randn_like) as a counter-based generator:
 * out[b][e] = normal(Philox4x32-10(key = seed, counter = (e / 4, clip0 + b, step, 0)))[e % 4], Box-Muller on pairs.
 * A clip's noise depends only on (seed, its GLOBAL index, step, element): results do not depend on how clips are
 * batched or sharded over GPUs.  The fused loop (avd_sampler.eta > 0) draws exactly these values. */
int avd_philox_normal(float* out, uint64_t seed, int64_t clip0, int step, int B, int64_t per_clip, void* stream);

/* ---------------------------------------------------------------- K7: fused CFG + un-tokenise + DDIM */
/* sample_clip.py:342-348 (audio) / 381-389 (video): eps = eps_n + g (eps_c - eps_n) -> latent layout -> ddim_step(eta=0).
 * eps_cond/eps_null: token layout [B, Nt, Dt]; z_t/z_prev latent layout (may alias).
 * video: C,T,H,W + tube (t,h,w); audio: C, L=Fa + chunk (length, stride), Nt windows. */
typedef struct {
  int target;                 /* AVD_TARGET_* */
  int B;
  int C, T, H, W, tt, th, tw; /* video latent + tube (target == video) */
  int L, length, stride, Nt;  /* audio latent length Fa, chunk, window count (target == audio) */
  float guidance;
  int t_now, t_prev;          /* integer timesteps (t_prev may be -1) */
  int T_train;
  float eta;                  /* > 0: adds sigma * z (schedule_utils.py:185-197), z as avd_philox_normal draws it */
  int step;                   /* Philox counter word (the sampler's step index) */
  uint64_t seed;
  int64_t clip0;              /* global index of clip 0 of this batch */
} avd_cfg_ddim_args;
int avd_cfg_untok_ddim(const avd_cfg_ddim_args* a, const float* eps_cond, const float* eps_null, const float* z_t,
                       float* z_prev, const float* alpha_bar, void* stream);

/* ---------------------------------------------------------------- operand preparation */
/* nn.Linear weight [N,K] fp32 (optionally scaled per input column, used to fold a norm's `scale`)
 * -> avd_nseg(precision) bf16 segments [nseg][N][K]. */
int avd_pack_weight(const float* W, const float* col_scale, void* out_bf16, int N, int K, int precision, void* stream);
/* transpose W[N,K] -> Wt[K,N] fp32 (adapter weights for the fused input builder) */
int avd_transpose_f32(const float* W, float* Wt, int N, int K, void* stream);
/* x[M,d] fp32 -> bf16 segments [nseg][M][d] + per-row partial sums of squares ssq[M][d/64] */
int avd_rows_prepare(const float* x, void* xa_bf16, float* ssq, int M, int d, int precision, void* stream);
/* RMSNorm (mmdt.py:39-42) of rows [row_off, row_off+n_rows) of each of n_seq sequences of N rows:
 * y = scale * x / (||x||/sqrt(d) + eps).  Output compact [n_seq*n_rows, d] as fp32 and/or bf16 segments. */
int avd_rmsnorm_rows(const float* x, const float* scale, int n_seq, int N, int row_off, int n_rows, int d, float eps,
                     float* out_f32, void* out_bf16, int precision, void* stream);
/* 16-bit operand range check (fp16 mode): counts the elements of buf[n] (fp16 bit patterns) that are +-65504 (what the
 * saturating packs produce on overflow), inf or NaN into *counter (atomicAdd).  Zero cost unless called. */
int avd_f16_range_scan(const void* buf, int64_t n, unsigned int* counter, void* stream);
/* nn.LayerNorm(d, eps) + activation (noise_heads.py:139-146): y[M,d] fp32 -> bf16 segments and/or fp32.
 * apply_gelu: 0 plain LayerNorm, 1 GELU(erf), 2 LeakyReLU(0.1), 3 ReLU (noise_heads._make_activation, :22-31). */
int avd_layernorm_act(const float* y, const float* w, const float* b, int M, int d, float eps, int apply_gelu,
                      float* out_f32, void* out_bf16, int precision, void* stream);

/* ---------------------------------------------------------------- K4: tcgen05 GEMM with fused prologue/epilogue */
/* out = epilogue( A[M,K] * W[N,K]^T ).  Stands in for nn.Linear.forward (mmdt.py:54-60,70-71,
 * noise_heads.py:131-177, sample_clip.py:52) plus the elementwise ops around it:
 *   v = acc * rinv(row) + bias[col] (+ resid[row,col]);  v = gelu_erf(v) if act;
 *   rinv(row) = 1/(sqrt(sum_p ssq_in[row,p]) / sqrt(rms_dim) + rms_eps)   -- the RMSNorm prologue of the
 *   reference, applied after the MMA (its `scale` is folded into W's columns by avd_pack_weight). */
typedef struct {
  int M, N, K;
  int precision;
  const void* A;            /* bf16 [nseg][M][K] */
  const void* W;            /* bf16 [nseg][N][K] */
  const float* bias;        /* [N] or NULL */
  const float* ssq_in;      /* [M][ssq_parts] or NULL (no row scaling) */
  int ssq_parts;
  int rms_dim;
  float rms_eps;
  const float* resid;       /* fp32 [M][N] or NULL */
  int act;                  /* 0 none, 1 GELU(erff), 2 GELU with a 1.5e-7-accurate erf (bf16 mode) */
  float* out_f32;           /* fp32 [M][N] or NULL */
  void* out_bf16;           /* bf16 [out_nseg][M][N] or NULL */
  int out_precision;        /* precision of the consumer of out_bf16 (selects the segment count) */
  float* ssq_out;           /* [M][N/64] partial sums of squares of v, or NULL */
  float a_scale;            /* the accumulator is multiplied by this before everything else (0 means 1): undoes the
                             * scale carried by a pre-scaled fp16 A operand */
  float out_scale;          /* the 16-bit output (only it) is multiplied by this before rounding (0 means 1) */
} avd_gemm_args;
int avd_gemm(const avd_gemm_args* g, void* stream);

/* ---------------------------------------------------------------- K5: joint video+audio self-attention */
/* nn.MultiheadAttention core (mmdt.py:58-61): per sequence and head softmax(q k^T / sqrt(hd)) v, full
 * bidirectional, no mask.  qkv: [n_seq*N, 3*d] (columns q | k | v, head j = columns 64j..64j+63 of each),
 * bf16 when qkv_is_f32 == 0 (tcgen05 flash kernel), fp32 otherwise (fp32 CUDA-core kernel for fp32 mode).
 * out: bf16 segments [nseg(out_precision)][n_seq*N][d]. */
int avd_attention(const void* qkv, int qkv_is_f32, void* out_bf16, int out_precision, int n_seq, int N, int n_heads,
                  int head_dim, void* stream);
/* the same with nn.MultiheadAttention's key_padding_mask (mmdt.py:58-60) as bit words [n_seq][ceil(N/32)] (bit set =
 * key ignored; layout of avd_core_workspace.key_mask); key_mask == NULL is avd_attention */
int avd_attention_masked(const void* qkv, int qkv_is_f32, void* out_bf16, int out_precision, int n_seq, int N,
                         int n_heads, int head_dim, const unsigned int* key_mask, void* stream);

/* ---------------------------------------------------------------- K6: MMDT core + noise head */
typedef struct {
  const void* w_qkv;  const float* b_qkv;   /* [3d,d] with norm1.scale folded */
  const void* w_o;    const float* b_o;     /* [d,d] */
  const void* w_fc1;  const float* b_fc1;   /* [d_mlp,d] with norm2.scale folded */
  const void* w_fc2;  const float* b_fc2;   /* [d,d_mlp] */
  /* norm="layernorm" (mmdt.py:45-46, nn.LayerNorm): blocks.i.norm{1,2}.{weight,bias} [d]; w_qkv / w_fc1 are then packed
   * WITHOUT a folded scale.  NULL for rmsnorm. */
  const float* ln1_w; const float* ln1_b; const float* ln2_w; const float* ln2_b;
} avd_block_weights;

typedef struct {
  int d_model, n_heads, n_layers, d_mlp;
  int precision;
  float rms_eps;
  avd_block_weights blocks[AVD_MAX_LAYERS];
  const float* final_scale;                 /* final_norm.scale [d] */
  int norm_kind;                            /* 0 = rmsnorm (folded into the following GEMM), 1 = layernorm (own pass) */
  float ln_eps;                             /* nn.LayerNorm eps (1e-5) */
  const float* final_ln_w;                  /* norm="layernorm": final_norm.weight / .bias [d] (used by the fused sampler) */
  const float* final_ln_b;
} avd_core_weights;

#define AVD_CHAIN_FLAG_BYTES 2560   /* 4 phases x 160 row groups x u32, see avd_core_workspace.chain_flags */

/* Caller-allocated work buffers for M = n_seq*N rows (S = avd_nseg(precision)):
 *   x fp32 [M,d]; xa bf16 [S][M,d]; ssq fp32 [M, d/64]; qkv bf16 [M,3d] (fp32 in AVD_PREC_FP32 mode);
 *   att bf16 [S][M,d]; hmid bf16 [S][M,d_mlp]. */
typedef struct {
  float* x; void* xa; float* ssq; void* qkv; void* att; void* hmid;
  /* MMDiT.forward's key_padding_mask (mmdt.py:58-60, 134-147) packed as bits: word [seq * ceil(N/32) + k/32], bit k%32
   * set = key k of sequence seq is padding and takes no part in any softmax.  NULL = no mask (what the sampler uses). */
  const unsigned int* key_mask;
  /* fp16 mode range check: when not NULL, every 16-bit operand buffer is scanned after each block (avd_f16_range_scan)
   * and saturated / non-finite values are counted here.  NULL (the timed configuration) = no scan kernels. */
  unsigned int* f16_sat_count;
  /* AVD_CHAIN_FLAG_BYTES of ZERO-INITIALISED device memory, or NULL.  When given (and every GEMM of a block is one wave
   * of 128-row tiles) out-proj -> fc1 -> fc2 -> the next block's QKV run as ONE persistent launch whose tiles hand their
   * 128-row groups to the next GEMM through these counters; the launch leaves them zero again.  Zero them once (and
   * again after a failed launch); never share them between two avd_mmdt_blocks / sampler calls running concurrently. */
  unsigned int* chain_flags;
  /* measurement aid for THIS call only: bit k set = kernels of profiler kind k are not launched (results meaningless) */
  int profile_skip_kinds;
} avd_core_workspace;

/* MMDiT.forward minus the final norm (mmdt.py:146-147): runs n_layers blocks.  Reads the input residual
 * stream from (x_in, xa_in, ssq_in) -- which it never writes -- and leaves the result in ws->x/xa/ssq. */
int avd_mmdt_blocks(const avd_core_weights* w, const float* x_in, const void* xa_in, const float* ssq_in,
                    const avd_core_workspace* ws, int n_seq, int N, void* stream);

typedef struct {
  int d_in, hidden, d_out, n_shared;
  int precision;
  float ln_eps;
  const void* w_in;  const float* b_in;                      /* input_proj[m]  */
  /* the trunk: shared[i] followed by the modality-specific blocks spec[m][j] (or shared_specific_trunk[j]); all are
   * Linear -> LayerNorm -> activation (noise_heads.py:139-172), n_shared counts both kinds */
  const void* w_s[AVD_MAX_SHARED]; const float* b_s[AVD_MAX_SHARED];    /* block[i][0]   */
  const float* ln_w[AVD_MAX_SHARED]; const float* ln_b[AVD_MAX_SHARED]; /* block[i][1]   */
  const void* w_out; const float* b_out;                     /* out_proj[m]    */
  int activation;                                            /* 0 / 1 GELU(erf), 2 LeakyReLU(0.1), 3 ReLU (noise_heads.py:22-31) */
} avd_head_weights;

/* MultiModalNoiseHead.forward for one modality (noise_heads.py:201-220).  h: bf16 segments [S][M][d_in];
 * act_a/act_b: bf16 [S][M][hidden] scratch; y: fp32 [M][hidden] scratch; eps: fp32 [M][d_out].
 * AVD_PREC_F16: h is fp16 pre-scaled by AVD_F16_XSCALE (what avd_rows_prepare writes); scratch holds fp16. */
int avd_head_forward(const avd_head_weights* w, const void* h, int M, void* act_a, void* act_b, float* y, float* eps,
                     void* stream);

/* ---------------------------------------------------------------- K8: the sampling loop */
typedef struct {
  int target;                       /* AVD_TARGET_* */
  int B;                            /* clips in this batch; cond+null are co-batched as 2B sequences */
  int Cv, T, H, W, tt, th, tw;      /* video latent [B,Cv,T,H,W], tube (sampler uses tube.h for h and w) */
  int Ca, Fa, chunk_len, chunk_stride;
  int tstep_dim;                    /* width of the concatenated timestep embedding */
  int steps;                        /* sampler steps S */
  int T_train;
  float guidance;
  const int64_t* timesteps_host;    /* HOST int64 [S+1] (su.make_sampling_schedule) */
  const float* temb_steps;          /* [S][tstep_dim]: su.timestep_embedding(t_now(step)) */
  const float* temb_zero;           /* [tstep_dim]: embedding of t = 0 (prompt rows) */
  const float* alpha_bar;           /* [T_train] */
  const float* wv_t; const float* bv;   /* video adapter, transposed [Dv][d-tstep_dim], bias */
  const float* wa_t; const float* ba;   /* audio adapter, transposed [Da][d-tstep_dim], bias */
  avd_core_weights core;
  avd_head_weights head;            /* the TARGET modality's head */
  void* workspace;                  /* >= avd_sampler_workspace_bytes() */
  int64_t workspace_bytes;
  /* ---- generality; all zero / NULL = the shipped sampler ---- */
  float eta;                        /* ddim_eta > 0: stochastic DDIM, z from Philox (avd_philox_normal) */
  uint64_t seed;
  int64_t clip_index0;              /* global index of clip 0 (shard offset): keeps eta > 0 results shard-invariant */
  /* per-token additive rows applied after the adapter: [Nv][d - tstep_dim] / [Na][d - tstep_dim], e.g. positional +
   * modality embeddings (adapters.py:41-127) summed by the caller; NULL = none */
  const float* add_v; const float* add_a;
  /* ADDITIVE time embedding (the trainer's variant, trainer.py:45-49: tokens + emb): [S][d - tstep_dim] per step for
   * the target rows and [d - tstep_dim] for t = 0 (prompt rows); NULL = none.  tstep_dim may then be 0. */
  const float* tadd_steps; const float* tadd_zero;
  int profile_skip_kinds;           /* measurement aid, see avd_core_workspace.profile_skip_kinds */
  unsigned int* f16_sat_count;      /* fp16 range check of every operand buffer after each block / the head; NULL = off */
  /* adapter weights as split-bf16 GEMM operands [nseg][d - tstep_dim][D] (avd_pack_weight, precision adapt_precision =
   * AVD_PREC_BF16X3 or AVD_PREC_FP32), or NULL.  When given and the TARGET modality's token width allows a GEMM
   * (D % 64 == 0: the video tubes) the per-step input builder runs the adapter on tcgen05 (tokenise -> GEMM ->
   * assemble) instead of CUDA-core FMAs (option bx_wide = 0 / 1 keeps the 4- / 32-token FMA kernels). */
  const void* wv_pack; const void* wa_pack;
  int adapt_precision;
} avd_sampler;

int64_t avd_sampler_workspace_bytes(const avd_sampler* s);
/* loop-invariant part (F10): tokenise + adapt the prompt latent, t=0 embedding, zeroed null-branch prompt rows */
int avd_sample_prepare(const avd_sampler* s, const float* z_prompt, void* stream);
/* one DDIM+CFG step `step` in [0,S): z (target latent) is updated in place.  Optional taps (may be NULL):
 * eps_tap [2B,Nt,Dt] receives the per-branch head output (cond rows first, then null). */
int avd_sample_step(const avd_sampler* s, int step, float* z, float* eps_tap, void* stream);
/* sample_one_direction's loop (sample_clip.py:317-348 / 359-389) for a batch: prepare + S steps. */
int avd_sample_loop(const avd_sampler* s, const float* z_prompt, float* z, void* stream);

/* ---------------------------------------------------------------- encoders / decoders
 * AudioCodec.encode (avdiff/models/encoders/audio_codec.py:184-199): wav[B,1,L] -> z[B,Ca,Fa].
 *   pre[0][0] Conv1d(1->H,k9)+GELU, pre[1][0] Conv1d(H->H,k9)+GELU, right-pad / crop to Fa*hop, avg_pool1d(hop),
 *   to_lat 1x1.  (Fa, hop) are what AudioCodec._avgpool_frames derives (:149-176): Fa = frames_per_clip with
 *   hop = _compute_exact_pool_params(L, Fa), or Fa = ceil(L / hop_samples) with hop = hop_samples.
 *   Weights in the module's layouts: w_pre0[H,1,9], w_pre1[H,H,9], w_lat[Ca,H,1]; H a multiple of 8, <= 64. */
int avd_audio_encode(const float* wav, float* z, int B, int L, int H, int Ca, int Fa, int hop, const float* w_pre0,
                     const float* b_pre0, const float* w_pre1, const float* b_pre1, const float* w_lat,
                     const float* b_lat, void* stream);
/* AudioCodec.decode (audio_codec.py:201-214): z[B,Ca,Fa] -> wav[B,1,Fa*hop]: from_lat 1x1, nearest upsample (ATen's
 * fp32 index rule), smooth[0] / smooth[2] Conv1d(H->H,K)+GELU, smooth[4] Conv1d(H->1,K), tanh.  K in {3,5,7,9,11}.
 * scratch: >= avd_audio_decode_scratch_bytes() device bytes (one [B,H,Fa*hop] fp32 activation). */
int64_t avd_audio_decode_scratch_bytes(int B, int H, int Fa, int hop);
int avd_audio_decode(const float* z, float* wav, float* scratch, int B, int Ca, int Fa, int H, int hop, int K,
                     const float* w_from, const float* b_from, const float* w_s0, const float* b_s0, const float* w_s2,
                     const float* b_s2, const float* w_s4, const float* b_s4, void* stream);

/* avd_audio_decode without caller scratch and without any allocation: the intermediate activation lives in a
 * file-scope device array of AVD_AUDIO_DECODE_STATIC_MAX_BYTES that every GPU loading the library holds until exit.
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Calls no cudaMalloc / cudaFree /
 * allocating library routine.  Returns -1 with a message when B*H*Fa*hop*4 exceeds the limit.  All calls on a device
 * share the one array: calls that overlap on the GPU (different streams) race -- serialise them.  The array keeps the
 * previous call's contents; nothing in it has to start at zero (the first kernel writes all that the second reads). */
#define AVD_AUDIO_DECODE_STATIC_MAX_BYTES (16 * 1024 * 1024)
int avd_audio_decode_static(const float* z, float* wav, int B, int Ca, int Fa, int H, int hop, int K,
                            const float* w_from, const float* b_from, const float* w_s0, const float* b_s0,
                            const float* w_s2, const float* b_s2, const float* w_s4, const float* b_s4, void* stream);

#define AVD_VAE_MAX_BLOCKS 8
/* VideoVAE parameters (avdiff/models/encoders/vae_video3d.py:87-127), fp32 device pointers in the module's layouts:
 * conv weights [64,Cin,3,3,3], GroupNorm weight / bias [64], 1x1x1 convs [Cout,Cin,1,1,1]. */
typedef struct {
  int in_ch, lat_ch, base;           /* 3, Cv, 64 (enc_base = dec_base = 64: the kernels' trunk width) */
  int t_down, s_down;
  int enc_blocks, dec_blocks;
  int out_tanh;                      /* 0 = sigmoid (default), 1 = tanh (out_activation) */
  float gn_eps;                      /* nn.GroupNorm eps (1e-5) */
  const float* enc_w[AVD_VAE_MAX_BLOCKS]; const float* enc_b[AVD_VAE_MAX_BLOCKS];        /* enc_net[i][0] */
  const float* enc_gn_w[AVD_VAE_MAX_BLOCKS]; const float* enc_gn_b[AVD_VAE_MAX_BLOCKS];  /* enc_net[i][2] */
  const float* to_lat_w; const float* to_lat_b;                                          /* to_lat (or to_mu) */
  const float* from_lat_w; const float* from_lat_b;
  const float* dec_w[AVD_VAE_MAX_BLOCKS]; const float* dec_b[AVD_VAE_MAX_BLOCKS];        /* dec_net[i][0] */
  const float* dec_gn_w[AVD_VAE_MAX_BLOCKS]; const float* dec_gn_b[AVD_VAE_MAX_BLOCKS];  /* dec_net[i][2] */
  const float* to_img_w; const float* to_img_b;
} avd_vae_weights;
/* device scratch for one encode or decode call at full resolution (two [B,64,T,H,W] fp32 activations + statistics) */
int64_t avd_vae_scratch_bytes(int B, int T, int H, int W);
/* VideoVAE.encode (vae_video3d.py:164-190, eval: z = mu when variational): x[B,3,T,H,W] -> z[B,Cv,T/td,H/sd,W/sd].
 * T, H, W must already be multiples of (t_down, s_down, s_down) (the mirror center-crops like :139-162). */
int avd_video_encode(const avd_vae_weights* w, const float* x, float* z, void* scratch, int B, int T, int H, int W,
                     void* stream);
/* VideoVAE.decode (vae_video3d.py:196-214): z[B,Cv,Tl,Hl,Wl] -> x_hat[B,3,T,H,W] (trilinear, align_corners=False). */
int avd_video_decode(const avd_vae_weights* w, const float* z, float* x, void* scratch, int B, int Tl, int Hl, int Wl, int T,
                     int H, int W, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* AVDIFF_B200_H_ */
