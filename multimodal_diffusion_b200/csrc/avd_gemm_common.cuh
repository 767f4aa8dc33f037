// avd_gemm_common.cuh -- pieces shared by the two tcgen05 GEMM kernels (avd_gemm.cu: one tile per CTA;
// avd_gemm_ps.cu: persistent CTAs with double-buffered TMEM accumulators).
#pragma once
#include "avd_internal.h"
#include "avd_ptx.cuh"

#include <cuda_fp16.h>

namespace avd {

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;            // 64 bf16 = 128 B = one swizzle atom row
constexpr int GEMM_THREADS = 256;
constexpr int GEMM_BOX = 4096;         // one epilogue staging box: 32 rows x 128 B (32 fp32 or 64 bf16 columns)

struct GemmParams {
  int M, N, K;
  int n_pairs;
  uint32_t pairs;            // 4 bits per pair: (a_seg << 2) | w_seg, pair i in bits [4i, 4i+4)
  const float* bias;
  const float* ssq_in;
  int ssq_parts;
  float inv_sqrt_rms_dim;    // 1/sqrt(rms_dim)
  float rms_eps;
  int has_resid;             // residual tile comes in through tmR (and the fp32 result is written over it in smem)
  int act;                   // 0 none, 1 GELU (erff), 2 GELU (fast erf, |err| < 2e-7)
  float* out_f32;            // fp32 result [M,N] (may alias the residual) or nullptr
  __nv_bfloat16* out_bf16;   // bf16 result segments [out_nseg][M][N] or nullptr
  int out_nseg;
  float* ssq_out;
  int f16_in;                // operands are fp16 (AVD_PREC_F16) instead of bf16
  int f16_out;               // out_bf16 holds fp16 (one segment)
  float a_scale;             // multiplies the accumulator (folded into the row scale)
  float out_scale;           // multiplies the 16-bit output only
  unsigned long long* trace; // debug: per-CTA phase timestamps (avd_debug_trace), nullptr in production
  int dbg;                   // debug-hooks builds only: epilogue timing experiments (option gemm_dbg)
};

__device__ __forceinline__ void trace_mark(unsigned long long* trace, int slot) {
  if (trace != nullptr) {
    const int cta = blockIdx.y * gridDim.x + blockIdx.x;
    unsigned long long g, c;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g));
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(c));
    trace[(cta * 16 + slot) * 2] = g;
    trace[(cta * 16 + slot) * 2 + 1] = c;
  }
}

__device__ __forceinline__ float4 lds128(const uint8_t* p) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(smem_u32(p)));
  return v;
}
__device__ __forceinline__ void sts128(uint8_t* p, float a, float b, float c, float d) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(smem_u32(p)), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void sts128u(uint8_t* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(smem_u32(p)), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// GELU of two values in f16x2 arithmetic, tanh form with the hardware tanh.approx.f16x2, packed as bf16x2.  Used only
// where the result is stored as ONE bf16 segment (fc1 of the bf16 mode): |tanh form - erf form| <= 4.8e-4 and the
// f16 arithmetic (2^-11) are both below the bf16 quantum (2^-9) of the stored value; rms over N(0,1) inputs 3e-4 vs
// 1.1e-3 for the bf16 rounding itself.  ~10 instructions + 1 MUFU per PAIR instead of ~16 + 2 MUFU per element.
// the same arithmetic, result kept as fp16x2 (fp16 mode).  There the tanh-form error (<= 4.8e-4 absolute) is of the order
// of the fp16 quantum of the stored value, not below it; the hot path measured WITH this activation: per-branch eps
// 0.9-1.1e-3, CFG-combined 1.7-2.4e-3 against the 1e-2 tolerance (tests/test_gpu_model.py).  The
// mirror MLP.forward uses the same activation path in the 16-bit modes, so module-level and fused numerics agree.
__device__ __forceinline__ uint32_t gelu_pair_f16x2(float a, float b) {
  // saturating conversion (one F2FP.SATFINITE, like every other fp16 pack of the library): a pre-activation beyond
  // +-65504 stays finite.  x*x may still reach +inf in f16 for |x| > 255; then inner = +inf, w = +-inf, tanh = +-1 and
  // the result is the correct limit (x or 0), never NaN (x = 0 gives u = 0).
  const uint32_t xi = pack_f16x2(a, b);
  const __half2 x = *reinterpret_cast<const __half2*>(&xi);
  const __half2 u = __hmul2(x, x);
  const __half2 inner = __hfma2(__float2half2_rn(0.0356774081f), u, __float2half2_rn(0.7978845608f));
  const __half2 w = __hmul2(x, inner);
  uint32_t wi = *reinterpret_cast<const uint32_t*>(&w), ti;
  asm("tanh.approx.f16x2 %0, %1;" : "=r"(ti) : "r"(wi));
  const __half2 t = *reinterpret_cast<const __half2*>(&ti);
  const __half2 h = __hmul2(x, __float2half2_rn(0.5f));
  const __half2 o = __hfma2(h, t, h);
  return *reinterpret_cast<const uint32_t*>(&o);
}

__device__ __forceinline__ uint32_t gelu_pair_bf16x2(float a, float b) {
  const __half2 x = __floats2half2_rn(a, b);
  const __half2 u = __hmul2(x, x);
  const __half2 inner = __hfma2(__float2half2_rn(0.0356774081f), u, __float2half2_rn(0.7978845608f));
  const __half2 w = __hmul2(x, inner);
  uint32_t wi = *reinterpret_cast<const uint32_t*>(&w), ti;
  asm("tanh.approx.f16x2 %0, %1;" : "=r"(ti) : "r"(wi));
  const __half2 t = *reinterpret_cast<const __half2*>(&ti);
  const __half2 h = __hmul2(x, __float2half2_rn(0.5f));
  const float2 o = __half22float2(__hfma2(h, t, h));
  return pack_bf16x2(o.x, o.y);
}


// 16-bit output conversion of the single-segment modes: bf16, or fp16 with the output scale and saturation
__device__ __forceinline__ uint32_t pack16(bool f16, float scale, float a, float b) {
  return f16 ? pack_f16x2(a * scale, b * scale) : pack_bf16x2(a, b);
}
__device__ __forceinline__ uint32_t gelu_pair16(bool f16, float a, float b) {
  return f16 ? gelu_pair_f16x2(a, b) : gelu_pair_bf16x2(a, b);
}

// what launch_gemm decides before it launches (avd_gemm.cu: plan_gemm); persistent: more than one wave of tiles, the
// persistent kernel picks its own tile width
struct GemmPlan {
  GemmParams p;
  int nseg, bn, mode, force_bn;
  bool persistent;
};
int plan_gemm(const avd_gemm_args& g, GemmPlan* plan);
unsigned int* ps_debug_device_ptr();         // host-visible barrier-timeout record shared by the persistent kernels

// persistent kernel (avd_gemm_ps.cu); bn = 0 lets it choose the tile width
int launch_gemm_ps(const avd_gemm_args& g, GemmParams p, int nseg, int force_bn, cudaStream_t s);
// persistent kernel on CTA pairs (avd_gemm_ps2.cu): 1 launched, 0 does not qualify, < 0 error
int launch_gemm_ps2(const avd_gemm_args& g, GemmParams p, int nseg, int mode, int sms, cudaStream_t s);
int ps_debug_read(unsigned int* out8);

}  // namespace avd
