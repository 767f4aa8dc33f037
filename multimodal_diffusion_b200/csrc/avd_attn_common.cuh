// avd_attn_common.cuh -- pieces shared by the tcgen05 attention kernels (avd_attn.cu: single-block / two-pass multi-block /
// small kernels; avd_attn_fa.cu: the multi-block kernel with P and O resident in tensor memory).
#pragma once
#include "avd_internal.h"
#include "avd_ptx.cuh"

namespace avd {

constexpr int ATT_BQ = 128;
constexpr int ATT_HD = 64;
constexpr int ATT_Q_BYTES = ATT_BQ * ATT_HD * 2;     // 16 KB: [128 rows][64 bf16], 128B swizzle

struct AttnParams {
  int N;            // tokens per sequence
  int d;            // model width (row stride of out)
  int n_heads;
  int64_t M;        // total rows (n_seq * N)
  __nv_bfloat16* out;
  int out_nseg;
  float scale_log2e;  // (1/sqrt(hd)) * log2(e)
  unsigned long long* trace;   // debug phase stamps (avd_debug_trace)
  int poly_exp;     // multi-block path: every second exponential on the FMA pipe (ex2_poly)
  const uint32_t* kmask;   // key_padding_mask as bits [n_seq][kmask_words] (bit = 1: key ignored) or NULL
  int kmask_words;         // ceil(N / 32)
};

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// The exponentials of a 32-score chunk, issued back to back.  Left to itself the compiler interleaves each MUFU.EX2 with
// the FADD / F2FP that consumes the previous one, and with two softmax warps per scheduler every such pair then waits out
// the MUFU latency (the SASS showed MUFU, FADD, MUFU, FADD ...).  The volatile asm keeps the 32 independent MUFUs together;
// sums and packs follow.  What remains is the pipe itself: 16 exponentials per cycle per SM, i.e. 8 cycles per warp
// instruction on each scheduler -- 128 rows x 160 key columns = 1280 cycles per CTA, which is what the phase trace shows.
__device__ __forceinline__ void ex2_chunk(const uint32_t (&x)[32], float sc, float neg_m_sc, float (&e)[32]) {
#pragma unroll
  for (int c = 0; c < 32; ++c) {
    const float a = fmaf(__uint_as_float(x[c]), sc, neg_m_sc);
    asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(e[c]) : "f"(a));
  }
}

// 2^x for x <= 0 on the FMA / ALU pipes: round-to-integer through the 1.5 * 2^23 constant, a degree-4 near-minimax
// polynomial for 2^f on [-0.5, 0.5] (max relative error 2.9e-6 in fp32, below the 16-bit rounding P gets), and the integer
// part added into the exponent field.  Used for every second exponential of the MULTI-BLOCK path only: with 16 softmax
// warps per SM and 49 152 exponentials per 192-key block that path is bound by the 16-lane MUFU pipe (the split
// FlashAttention-4 describes); in the single-block kernel (8 warps, latency bound) the same change cost 24 %.
__device__ __forceinline__ float ex2_poly(float x) {
  x = fmaxf(x, -125.0f);                       // keeps the result a normal number (scores this far down round to 0 in P)
  const float xf = x + 12582912.0f;
  const float n = xf - 12582912.0f;
  const float f = x - n;
  float q = 0.009582850150763988f;
  q = fmaf(q, f, 0.055906426161527634f);
  q = fmaf(q, f, 0.24024099111557007f);
  q = fmaf(q, f, 0.6931241750717163f);
  q = fmaf(q, f, 1.0f);
  return __int_as_float(__float_as_int(q) + (__float_as_int(xf) << 23));
}

// Per-warp timeline of one CTA (experimental builds with -DAVD_ATT_WTRACE; tools/attn_wtrace.py): lane 0 of every warp of
// CTA 0 stamps clock64 at slot `slot` of key blocks 3..5 into the avd_debug_trace buffer (which must hold 32 u64 per CTA).
#ifdef AVD_ATT_WTRACE
#define ATT_WMARK(slot)                                                                                         \
  do {                                                                                                          \
    if (trace != nullptr && lane == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && j >= 3 && j < 6) { \
      unsigned long long c_;                                                                                    \
      asm volatile("mov.u64 %0, %%clock64;" : "=l"(c_));                                                        \
      trace[(warp * 3 + (j - 3)) * 12 + (slot)] = c_;                                                           \
    }                                                                                                           \
  } while (0)
// the same, issued only once `dep` (a float) is available: a stamp after the instruction that produces it
#define ATT_WMARK_DEP(slot, dep)                                                                                \
  do {                                                                                                          \
    if (trace != nullptr && lane == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && j >= 3 && j < 6) { \
      unsigned long long c_;                                                                                    \
      asm volatile("{\n\t.reg .f32 t_;\n\tmov.f32 t_, %1;\n\tmov.u64 %0, %%clock64;\n\t}" : "=l"(c_) : "f"(dep));      \
      trace[(warp * 3 + (j - 3)) * 12 + (slot)] = c_;                                                           \
    }                                                                                                           \
  } while (0)
#else
#define ATT_WMARK(slot) do { } while (0)
#define ATT_WMARK_DEP(slot, dep) do { } while (0)
#endif

// multi-block attention with P / O in tensor memory (avd_attn_fa.cu): 1 launched, 0 not applicable, < 0 error
int launch_attention_fa(const void* qkv16, const AttnParams& p, int n_seq, bool f16, cudaStream_t s);

}  // namespace avd
