// avd_elementwise.cu -- the HBM-bound kernels of the hot path: tokenisers (bit-exact gathers),
// timestep embedding, DDIM update, fused CFG + un-tokenise + DDIM, operand preparation
// (bf16 segment splitting, row sums of squares), RMSNorm / LayerNorm+GELU rows, and the fused
// input builder (tokenise -> adapter -> concat timestep embedding).
//
// Every arithmetic expression that the reference evaluates in fp32 elementwise torch ops is written
// with __f*_rn intrinsics in the reference's operation order so that nvcc cannot contract it into
// FMAs: those kernels are bit-exact against the CPU reference (see tests/test_gpu_elementwise.py).
#include "avd_internal.h"
#include "avd_ptx.cuh"
#include <cuda_fp16.h>

namespace avd {

static inline int cdiv(int64_t a, int64_t b) { return static_cast<int>((a + b - 1) / b); }

// Launch with the programmatic-dependent-launch attribute: the kernel may start while its predecessor drains; every
// kernel launched this way calls pdl_wait() before it touches memory (and so also orders the chain transitively).
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// =====================================================================================================
// K1: tube patch / unpatch   (ops.py:100-119, 122-144)
//   token flat index o = (((((((b*Tt+ti)*Hh+hj)*Ww+wk)*C+c)*t+dt)*h+dh)*w+dw
//   latent flat index i = (((b*C+c)*T + ti*t+dt)*H + hj*h+dh)*W + wk*w+dw
// One thread moves VEC consecutive dw elements (contiguous in both layouts when w % VEC == 0).
// =====================================================================================================
struct TubeDims {
  int B, C, T, H, W, t, h, w, Tt, Hh, Ww;
};

template <typename I>
__device__ __forceinline__ I tube_latent_index_from_token(const TubeDims& d, I o) {
  int dw = o % d.w; o /= d.w;
  int dh = o % d.h; o /= d.h;
  int dt = o % d.t; o /= d.t;
  int c = o % d.C;  o /= d.C;
  int wk = o % d.Ww; o /= d.Ww;
  int hj = o % d.Hh; o /= d.Hh;
  int ti = o % d.Tt; o /= d.Tt;
  I b = o;
  return (((b * d.C + c) * d.T + (ti * d.t + dt)) * d.H + (hj * d.h + dh)) * (I)d.W + (wk * d.w + dw);
}

template <typename I>
__device__ __forceinline__ I tube_token_index_from_latent(const TubeDims& d, I i) {
  int x = i % d.W; i /= d.W;
  int y = i % d.H; i /= d.H;
  int tt = i % d.T; i /= d.T;
  int c = i % d.C;  i /= d.C;
  I b = i;
  int wk = x / d.w, dw = x % d.w;
  int hj = y / d.h, dh = y % d.h;
  int ti = tt / d.t, dt = tt % d.t;
  return ((((((b * d.Tt + ti) * d.Hh + hj) * d.Ww + wk) * d.C + c) * d.t + dt) * d.h + dh) * (I)d.w + dw;
}

template <int VEC> struct VecOf { using type = float; };
template <> struct VecOf<2> { using type = float2; };
template <> struct VecOf<4> { using type = float4; };
constexpr int EW_UNROLL = 4;   // independent vectors in flight per thread (HBM latency needs the memory parallelism)

// I = uint32_t when every element index fits 32 bits (the divisions are then 32-bit), int64_t otherwise.
template <int VEC, bool INVERSE, typename I>
__global__ void __launch_bounds__(256) tube_kernel(const float* __restrict__ src, float* __restrict__ dst, TubeDims d,
                                                   int64_t n_vec) {
  using V = typename VecOf<VEC>::type;
  const int64_t v0 = (blockIdx.x * (int64_t)EW_UNROLL) * blockDim.x + threadIdx.x;
  V val[EW_UNROLL];
#pragma unroll
  for (int u = 0; u < EW_UNROLL; ++u) {
    const int64_t v = v0 + (int64_t)u * blockDim.x;
    if (v < n_vec) {
      const I o = (I)(v * VEC);      // index in the DESTINATION layout
      const I sidx = INVERSE ? tube_token_index_from_latent<I>(d, o) : tube_latent_index_from_token<I>(d, o);
      val[u] = __ldg(reinterpret_cast<const V*>(src + sidx));
    }
  }
#pragma unroll
  for (int u = 0; u < EW_UNROLL; ++u) {
    const int64_t v = v0 + (int64_t)u * blockDim.x;
    if (v < n_vec) *reinterpret_cast<V*>(dst + v * VEC) = val[u];
  }
}

template <int VEC, bool INVERSE>
static void launch_tube_vec(const float* src, float* dst, const TubeDims& d, int64_t total, cudaStream_t s) {
  const int64_t n_vec = total / VEC;
  const int blocks = cdiv(n_vec, 256 * EW_UNROLL);
  if (total < (int64_t(1) << 31)) tube_kernel<VEC, INVERSE, uint32_t><<<blocks, 256, 0, s>>>(src, dst, d, n_vec);
  else tube_kernel<VEC, INVERSE, int64_t><<<blocks, 256, 0, s>>>(src, dst, d, n_vec);
}

int launch_tube_patch(const float* src, float* dst, int B, int C, int T, int H, int W, int t, int h, int w, bool inverse,
                      cudaStream_t s) {
  TubeDims d{B, C, T, H, W, t, h, w, T / t, H / h, W / w};
  int64_t total = (int64_t)B * C * T * H * W;
  if (total == 0) return 0;
  bool al16 = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0;
  bool al8 = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 7) == 0;
  int vec = (w % 4 == 0 && al16) ? 4 : (w % 2 == 0 && al8) ? 2 : 1;
  LaunchScope ls(KIND_TOKENIZE, s);
  if (!inverse) {
    if (vec == 4) launch_tube_vec<4, false>(src, dst, d, total, s);
    else if (vec == 2) launch_tube_vec<2, false>(src, dst, d, total, s);
    else launch_tube_vec<1, false>(src, dst, d, total, s);
  } else {
    if (vec == 4) launch_tube_vec<4, true>(src, dst, d, total, s);
    else if (vec == 2) launch_tube_vec<2, true>(src, dst, d, total, s);
    else launch_tube_vec<1, true>(src, dst, d, total, s);
  }
  AVD_CUDA(cudaGetLastError());
  return 0;
}

// =====================================================================================================
// K1: audio chunk / un-chunk   (ops.py:17-45, 48-93; sample_clip.py:184-215)
//   tok[b,n,c*l+j] = z[b,c,n*s+j]
// =====================================================================================================
template <int VEC, typename I>
__global__ void __launch_bounds__(256) chunk1d_kernel(const float* __restrict__ z, float* __restrict__ tok, int C, int L,
                                                      int N, int length, int stride, int64_t n_vec) {
  using V = typename VecOf<VEC>::type;
  const int64_t v0 = (blockIdx.x * (int64_t)EW_UNROLL) * blockDim.x + threadIdx.x;
  V val[EW_UNROLL];
#pragma unroll
  for (int u = 0; u < EW_UNROLL; ++u) {
    const int64_t v = v0 + (int64_t)u * blockDim.x;
    if (v < n_vec) {
      I r = (I)(v * VEC);
      int j = r % length; r /= length;
      int c = r % C; r /= C;
      int n = r % N; r /= N;
      const I sidx = (r * C + c) * (I)L + (I)n * stride + j;
      val[u] = __ldg(reinterpret_cast<const V*>(z + sidx));
    }
  }
#pragma unroll
  for (int u = 0; u < EW_UNROLL; ++u) {
    const int64_t v = v0 + (int64_t)u * blockDim.x;
    if (v < n_vec) *reinterpret_cast<V*>(tok + v * VEC) = val[u];
  }
}

int launch_chunk1d(const float* z, float* tok, int B, int C, int L, int length, int stride, cudaStream_t s) {
  int N = (L - length) / stride + 1;
  int64_t total = (int64_t)B * N * C * length;
  if (total <= 0) return 0;
  bool al8 = ((reinterpret_cast<uintptr_t>(z) | reinterpret_cast<uintptr_t>(tok)) & 7) == 0;
  int vec = (length % 2 == 0 && stride % 2 == 0 && L % 2 == 0 && al8) ? 2 : 1;
  int64_t n_vec = total / vec;
  const bool small = (int64_t)B * C * L < (int64_t(1) << 31) && total < (int64_t(1) << 31);
  const int blocks = cdiv(n_vec, 256 * EW_UNROLL);
  LaunchScope ls(KIND_TOKENIZE, s);
  if (vec == 2 && small) chunk1d_kernel<2, uint32_t><<<blocks, 256, 0, s>>>(z, tok, C, L, N, length, stride, n_vec);
  else if (vec == 2) chunk1d_kernel<2, int64_t><<<blocks, 256, 0, s>>>(z, tok, C, L, N, length, stride, n_vec);
  else if (small) chunk1d_kernel<1, uint32_t><<<blocks, 256, 0, s>>>(z, tok, C, L, N, length, stride, n_vec);
  else chunk1d_kernel<1, int64_t><<<blocks, 256, 0, s>>>(z, tok, C, L, N, length, stride, n_vec);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

// Gather form of the reference's sequential overlap-add: position p receives windows i with
// i*stride <= p < i*stride + length, summed in increasing i (the order of the reference's `+=` loop,
// starting from 0.0f), then divided by max(count, 1e-8); positions past the last window are zero padding.
__device__ __forceinline__ float overlap_add_gather(const float* __restrict__ tok_b, int c, int p, int N, int C,
                                                    int length, int stride) {
  int L_ola = (N - 1) * stride + length;
  if (p >= L_ola) return 0.0f;
  if (stride == length) {     // disjoint windows: exactly one term, 0.0f + v then / 1.0f as the reference computes it
    const int i = p / length;
    return __fdiv_rn(__fadd_rn(0.0f, __ldg(tok_b + ((int64_t)i * C + c) * length + (p - i * length))), 1.0f);
  }
  int i_hi = p / stride;
  if (i_hi > N - 1) i_hi = N - 1;
  int i_lo = (p - length + stride) / stride;  // ceil((p-length+1)/stride)
  if (p - length + 1 <= 0) i_lo = 0;
  float acc = 0.0f, cnt = 0.0f;
  for (int i = i_lo; i <= i_hi; ++i) {
    int j = p - i * stride;
    if (j >= 0 && j < length) {
      acc = __fadd_rn(acc, __ldg(tok_b + ((int64_t)i * C + c) * length + j));
      cnt = __fadd_rn(cnt, 1.0f);
    }
  }
  return __fdiv_rn(acc, fmaxf(cnt, 1e-8f));
}

__global__ void unchunk1d_kernel(const float* __restrict__ tok, float* __restrict__ z, int C, int N, int length,
                                 int stride, int L_out, int64_t total) {
  int64_t o = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (o >= total) return;
  int p, c;
  int64_t b;
  if (total < (int64_t(1) << 31)) {
    const uint32_t o32 = (uint32_t)o, r = o32 / (uint32_t)L_out;
    p = (int)(o32 - r * (uint32_t)L_out);
    b = r / (uint32_t)C;
    c = (int)(r - (uint32_t)b * (uint32_t)C);
  } else {
    p = o % L_out;
    int64_t r = o / L_out;
    c = r % C;
    b = r / C;
  }
  z[o] = overlap_add_gather(tok + b * (int64_t)N * C * length, c, p, N, C, length, stride);
}

// Disjoint windows of 4 (the shipped audio tokeniser: chunk 4 / stride 4): one thread moves one window = one 16-byte
// token read and two 8-byte latent stores (rows of L_out floats are only 8-byte aligned); window index N is the zero
// padding past the last window.  Same arithmetic as the gather (0.0f + v, / 1.0f), so it stays bit-exact.
__global__ void __launch_bounds__(256) unchunk1d_w4_kernel(const float* __restrict__ tok, float* __restrict__ z, int C,
                                                           int N, int L_out, uint32_t total_w) {
  const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= total_w) return;
  const int wpr = (L_out + 3) >> 2;                 // windows per (b, c) row, including the padding ones
  const uint32_t row = w / wpr;                     // b * C + c
  const int n = (int)(w - row * wpr);
  const uint32_t b = row / C;
  const int c = (int)(row - b * C);
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (n < N) {
    v = __ldg(reinterpret_cast<const float4*>(tok + (((int64_t)b * N + n) * C + c) * 4));
    v.x = __fdiv_rn(__fadd_rn(0.0f, v.x), 1.0f); v.y = __fdiv_rn(__fadd_rn(0.0f, v.y), 1.0f);
    v.z = __fdiv_rn(__fadd_rn(0.0f, v.z), 1.0f); v.w = __fdiv_rn(__fadd_rn(0.0f, v.w), 1.0f);
  }
  float* zr = z + (int64_t)row * L_out + 4 * n;
  const int left = L_out - 4 * n;                   // elements of this window inside the row (L_out is even)
  *reinterpret_cast<float2*>(zr) = make_float2(v.x, v.y);
  if (left >= 4) *reinterpret_cast<float2*>(zr + 2) = make_float2(v.z, v.w);
}

int launch_unchunk1d(const float* tok, float* z, int B, int C, int N, int length, int stride, int L_out,
                     cudaStream_t s) {
  int64_t total = (int64_t)B * C * L_out;
  if (total <= 0) return 0;
  LaunchScope ls(KIND_TOKENIZE, s);
  const int64_t total_w = (int64_t)B * C * ((L_out + 3) / 4);
  const bool w4 = length == 4 && stride == 4 && L_out % 2 == 0 && total_w < (int64_t(1) << 31) &&
                  (reinterpret_cast<uintptr_t>(tok) & 15) == 0 && (reinterpret_cast<uintptr_t>(z) & 7) == 0;
  if (w4) unchunk1d_w4_kernel<<<cdiv(total_w, 256), 256, 0, s>>>(tok, z, C, N, L_out, (uint32_t)total_w);
  else unchunk1d_kernel<<<cdiv(total, 256), 256, 0, s>>>(tok, z, C, N, length, stride, L_out, total);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

// =====================================================================================================
// timestep embedding   (schedule_utils.py:64-86)
// =====================================================================================================
__global__ void temb_kernel(const float* __restrict__ t, float* __restrict__ emb, int B, int dim, float neg_log_mp) {
  int half = dim / 2;
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * dim) return;
  int b = idx / dim, k = idx % dim;
  if (k >= 2 * half) { emb[idx] = 0.0f; return; }  // odd dim: one zero pad column
  int kk = k < half ? k : k - half;
  float f = expf(__fdiv_rn(__fmul_rn(neg_log_mp, (float)kk), (float)half));
  float a = __fmul_rn(t[b], f);
  emb[idx] = k < half ? cosf(a) : sinf(a);
}

int launch_timestep_embedding(const float* t, float* emb, int B, int dim, float max_period, cudaStream_t s) {
  if (B * dim <= 0) return 0;
  float neg_log_mp = (float)(-log((double)max_period));
  LaunchScope ls(KIND_MISC, s);
  temb_kernel<<<cdiv((int64_t)B * dim, 256), 256, 0, s>>>(t, emb, B, dim, neg_log_mp);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

// =====================================================================================================
// Counter-based N(0,1) for ddim_step's eta > 0 branch (schedule_utils.py:185-197: sigma * randn_like(x_t)).
// Philox4x32-10 (Salmon et al., SC'11; the generator behind curand / torch's CUDA randn) keyed by the sampler seed, with
// the counter (element / 4, global clip index, step, 0): a clip's noise does not depend on its position in a batch or on
// the GPU that samples it.  Box-Muller turns the four words into four normals ((u0,u1) -> n0,n1; (u2,u3) -> n2,n3).
// =====================================================================================================
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1,
                                              uint32_t (&out)[4]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

__device__ __forceinline__ float philox_normal(uint64_t seed, int64_t clip, int step, int64_t elem) {
  uint32_t w[4];
  const uint64_t blk = (uint64_t)elem >> 2;
  philox4x32_10((uint32_t)blk, (uint32_t)clip, (uint32_t)step, (uint32_t)(blk >> 32) ^ (uint32_t)((uint64_t)clip >> 32),
                (uint32_t)seed, (uint32_t)(seed >> 32), w);
  const int lane = (int)(elem & 3);
  // u in (0, 1]: (w + 1) * 2^-32, so log(u) is finite
  const float u_r = __fmul_rn(__fadd_rn((float)(w[lane & 2] >> 8), 1.0f), 5.9604644775390625e-8f);      // 24 bits
  const float u_a = __fmul_rn((float)(w[(lane & 2) + 1] >> 8), 5.9604644775390625e-8f);
  const float rad = sqrtf(-2.0f * logf(u_r));
  float sn, cs;
  sincospif(2.0f * u_a, &sn, &cs);
  return (lane & 1) ? rad * sn : rad * cs;
}

__global__ void philox_normal_kernel(float* __restrict__ out, uint64_t seed, int64_t clip0, int step, int64_t per_clip) {
  const int b = blockIdx.y;
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < per_clip; e += (int64_t)gridDim.x * blockDim.x)
    out[(int64_t)b * per_clip + e] = philox_normal(seed, clip0 + b, step, e);
}

int launch_philox_normal(float* out, uint64_t seed, int64_t clip0, int step, int B, int64_t per_clip, cudaStream_t s) {
  if (B <= 0 || per_clip <= 0) return 0;
  AVD_REQUIRE(B <= 65535, "batch too large for one launch");
  LaunchScope ls(KIND_MISC, s);
  philox_normal_kernel<<<dim3(std::min<int64_t>(cdiv(per_clip, 256), 1024), B), 256, 0, s>>>(out, seed, clip0, step, per_clip);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

// fp16 operand range check: counts +-65504 (the saturating packs' overflow value), inf and NaN bit patterns
__global__ void f16_range_scan_kernel(const uint16_t* __restrict__ buf, int64_t n, unsigned int* __restrict__ counter) {
  unsigned int bad = 0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    bad += ((buf[i] & 0x7FFFu) >= 0x7BFFu) ? 1u : 0u;
  bad = __reduce_add_sync(0xffffffffu, bad);
  if ((threadIdx.x & 31) == 0 && bad != 0) atomicAdd(counter, bad);
}

int launch_f16_range_scan(const void* buf, int64_t n, unsigned int* counter, cudaStream_t s) {
  if (n <= 0) return 0;
  AVD_REQUIRE(buf != nullptr && counter != nullptr, "null argument");
  LaunchScope ls(KIND_MISC, s);
  f16_range_scan_kernel<<<std::min<int64_t>(cdiv(n, 256 * 8), 2048), 256, 0, s>>>(static_cast<const uint16_t*>(buf), n, counter);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

// =====================================================================================================
// DDIM   (schedule_utils.py:146-200)
// =====================================================================================================
struct DdimCoef {
  float sqrt_omb_t, denom, sqrt_a_prev, coeff_eps, sigma;
};

__device__ __forceinline__ DdimCoef ddim_coef(const float* __restrict__ alpha_bar, int64_t t_now, int64_t t_prev,
                                              float eta) {
  DdimCoef c;
  float a_t = alpha_bar[t_now > 0 ? t_now : 0];
  float a_prev = t_prev >= 0 ? alpha_bar[t_prev] : 1.0f;
  c.sqrt_omb_t = __fsqrt_rn(fmaxf(__fsub_rn(1.0f, a_t), 0.0f));
  c.denom = fmaxf(__fsqrt_rn(a_t), 1e-8f);
  c.sqrt_a_prev = __fsqrt_rn(a_prev);
  if (eta > 0.0f) {
    float frac = fmaxf(__fdiv_rn(__fsub_rn(1.0f, a_prev), fmaxf(__fsub_rn(1.0f, a_t), 1e-8f)), 0.0f);
    float omr = fmaxf(__fsub_rn(1.0f, __fdiv_rn(a_t, fmaxf(a_prev, 1e-8f))), 0.0f);
    c.sigma = __fmul_rn(eta, __fsqrt_rn(__fmul_rn(frac, omr)));
  } else {
    c.sigma = 0.0f;
  }
  c.coeff_eps = __fsqrt_rn(fmaxf(__fsub_rn(__fsub_rn(1.0f, a_prev), __fmul_rn(c.sigma, c.sigma)), 0.0f));
  return c;
}

__device__ __forceinline__ float ddim_update(const DdimCoef& c, float x, float e, float zn) {
  float x0 = __fdiv_rn(__fsub_rn(x, __fmul_rn(c.sqrt_omb_t, e)), c.denom);
  float r = __fadd_rn(__fmul_rn(c.sqrt_a_prev, x0), __fmul_rn(c.coeff_eps, e));
  return __fadd_rn(r, __fmul_rn(c.sigma, zn));
}

template <int VEC>
__global__ void __launch_bounds__(256) ddim_step_kernel(const float* __restrict__ x_t, const float* __restrict__ eps,
                                                        const float* __restrict__ noise, float* __restrict__ x_prev,
                                                        const int64_t* __restrict__ t_now, const int64_t* __restrict__ t_prev,
                                                        const float* __restrict__ alpha_bar, int64_t per_batch, float eta) {
  const int b = blockIdx.y;
  const DdimCoef c = ddim_coef(alpha_bar, t_now[b], t_prev[b], eta);
  const int64_t base = (int64_t)b * per_batch;
  const int64_t step = (int64_t)gridDim.x * blockDim.x;
  if (VEC == 4) {
    const int64_t n4 = per_batch / 4;
    const float4* x4 = reinterpret_cast<const float4*>(x_t + base);
    const float4* e4 = reinterpret_cast<const float4*>(eps + base);
    const float4* z4 = noise != nullptr ? reinterpret_cast<const float4*>(noise + base) : nullptr;
    float4* o4 = reinterpret_cast<float4*>(x_prev + base);
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += 2 * step) {
      const int64_t i1 = i + step;
      const bool two = i1 < n4;
      const float4 xa = x4[i], ea = e4[i];
      const float4 za = z4 != nullptr ? z4[i] : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 xb = xa, eb = ea, zb = za;
      if (two) { xb = x4[i1]; eb = e4[i1]; if (z4 != nullptr) zb = z4[i1]; }
      float4 r;
      r.x = ddim_update(c, xa.x, ea.x, za.x); r.y = ddim_update(c, xa.y, ea.y, za.y);
      r.z = ddim_update(c, xa.z, ea.z, za.z); r.w = ddim_update(c, xa.w, ea.w, za.w);
      o4[i] = r;
      if (two) {
        r.x = ddim_update(c, xb.x, eb.x, zb.x); r.y = ddim_update(c, xb.y, eb.y, zb.y);
        r.z = ddim_update(c, xb.z, eb.z, zb.z); r.w = ddim_update(c, xb.w, eb.w, zb.w);
        o4[i1] = r;
      }
    }
  } else {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < per_batch; i += step) {
      float zn = (noise != nullptr) ? noise[base + i] : 0.0f;
      x_prev[base + i] = ddim_update(c, x_t[base + i], eps[base + i], zn);
    }
  }
}

int launch_ddim_step(const float* x_t, const float* eps, const float* noise, float* x_prev, const int64_t* t_now,
                     const int64_t* t_prev, const float* alpha_bar, int T_train, int B, int64_t per_batch, float eta,
                     cudaStream_t s) {
  (void)T_train;
  if (B <= 0 || per_batch <= 0) return 0;
  const uintptr_t al = reinterpret_cast<uintptr_t>(x_t) | reinterpret_cast<uintptr_t>(eps) |
                       reinterpret_cast<uintptr_t>(noise) | reinterpret_cast<uintptr_t>(x_prev);
  const bool v4 = per_batch % 4 == 0 && (al & 15) == 0;
  LaunchScope ls(KIND_CFG_DDIM, s);
  if (v4) {
    dim3 grid(std::min<int64_t>(cdiv(per_batch / 4, 512), 1024), B);
    ddim_step_kernel<4><<<grid, 256, 0, s>>>(x_t, eps, noise, x_prev, t_now, t_prev, alpha_bar, per_batch, eta);
  } else {
    dim3 grid(std::min<int64_t>(cdiv(per_batch, 256), 1024), B);
    ddim_step_kernel<1><<<grid, 256, 0, s>>>(x_t, eps, noise, x_prev, t_now, t_prev, alpha_bar, per_batch, eta);
  }
  AVD_CUDA(cudaGetLastError());
  return 0;
}

// =====================================================================================================
// K7: fused CFG combine + un-tokenise + DDIM(eta=0)   (sample_clip.py:342-348 / 381-389)
//   16 B per target-latent element: reads eps_cond, eps_null (token layout) and z_t, writes z_prev.
// =====================================================================================================
struct NoiseArgs {      // eta > 0 (schedule_utils.py:185-197): z of element e of clip b = philox_normal(seed, clip0 + b, step, e)
  float eta;
  int step;
  uint64_t seed;
  int64_t clip0;
};

__device__ __forceinline__ float cfg_mix(float ec, float en, float g) {
  return __fadd_rn(en, __fmul_rn(g, __fsub_rn(ec, en)));
}

template <int VEC, typename I>
__global__ void __launch_bounds__(256) cfg_ddim_video_kernel(const float* __restrict__ eps_c, const float* __restrict__ eps_n,
                                                             const float* __restrict__ z_t, float* __restrict__ z_prev,
                                                             TubeDims d, float g, int t_now, int t_prev,
                                                             const float* __restrict__ alpha_bar, int64_t n_vec,
                                                             int64_t clip_elems, int64_t clip_stride, NoiseArgs nz) {
  constexpr int U = 2;                              // vectors per thread: 6 independent 16-byte loads in flight
  pdl_wait();
  pdl_launch_dependents();
  const int64_t v0 = (blockIdx.x * (int64_t)U) * blockDim.x + threadIdx.x;
  const DdimCoef c = ddim_coef(alpha_bar, t_now, t_prev, VEC == 1 ? nz.eta : 0.0f);   // eta > 0 takes the scalar path
  if (VEC == 4) {
    float4 ec[U], en[U], x[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t v = v0 + (int64_t)u * blockDim.x;
      if (v < n_vec) {
        const I i = (I)(v * 4);                                       // latent index
        I o = tube_token_index_from_latent<I>(d, i);                  // token index of the same element
        o += (I)((i / (I)clip_elems) * (I)(clip_stride - clip_elems));  // eps rows of a clip may sit clip_stride apart
        ec[u] = __ldg(reinterpret_cast<const float4*>(eps_c + o));
        en[u] = __ldg(reinterpret_cast<const float4*>(eps_n + o));
        x[u] = *reinterpret_cast<const float4*>(z_t + i);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t v = v0 + (int64_t)u * blockDim.x;
      if (v < n_vec) {
        float4 r;
        r.x = ddim_update(c, x[u].x, cfg_mix(ec[u].x, en[u].x, g), 0.0f);
        r.y = ddim_update(c, x[u].y, cfg_mix(ec[u].y, en[u].y, g), 0.0f);
        r.z = ddim_update(c, x[u].z, cfg_mix(ec[u].z, en[u].z, g), 0.0f);
        r.w = ddim_update(c, x[u].w, cfg_mix(ec[u].w, en[u].w, g), 0.0f);
        *reinterpret_cast<float4*>(z_prev + v * 4) = r;
      }
    }
  } else {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t v = v0 + (int64_t)u * blockDim.x;
      if (v < n_vec) {
        const I i = (I)v;
        I o = tube_token_index_from_latent<I>(d, i);
        o += (I)((i / (I)clip_elems) * (I)(clip_stride - clip_elems));
        const float zn = nz.eta > 0.0f ? philox_normal(nz.seed, nz.clip0 + (int64_t)(i / (I)clip_elems), nz.step,
                                                       (int64_t)(i % (I)clip_elems)) : 0.0f;
        z_prev[i] = ddim_update(c, z_t[i], cfg_mix(__ldg(eps_c + o), __ldg(eps_n + o), g), zn);
      }
    }
  }
}

// Audio: CFG per token element first, then the overlap-add gather (sum order i ascending), then DDIM.
__global__ void cfg_ddim_audio_kernel(const float* __restrict__ eps_c, const float* __restrict__ eps_n,
                                      const float* __restrict__ z_t, float* __restrict__ z_prev, int C, int L, int N,
                                      int length, int stride, float g, int t_now, int t_prev,
                                      const float* __restrict__ alpha_bar, int64_t total, int64_t clip_stride,
                                      NoiseArgs nz) {
  pdl_wait();
  pdl_launch_dependents();
  int64_t o = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (o >= total) return;
  DdimCoef cf = ddim_coef(alpha_bar, t_now, t_prev, nz.eta);
  int p, c;
  int64_t b;
  if (total < (int64_t(1) << 31)) {
    const uint32_t o32 = (uint32_t)o, r = o32 / (uint32_t)L;
    p = (int)(o32 - r * (uint32_t)L);
    b = r / (uint32_t)C;
    c = (int)(r - (uint32_t)b * (uint32_t)C);
  } else {
    p = o % L;
    int64_t r = o / L;
    c = r % C;
    b = r / C;
  }
  const float* ec_b = eps_c + b * clip_stride;
  const float* en_b = eps_n + b * clip_stride;
  float e = 0.0f;
  int L_ola = (N - 1) * stride + length;
  if (p < L_ola && stride == length) {   // disjoint windows: one term (0.0f + mix, / 1.0f as the reference computes it)
    const int i = p / length;
    const int64_t k = ((int64_t)i * C + c) * length + (p - i * length);
    e = __fdiv_rn(__fadd_rn(0.0f, cfg_mix(__ldg(ec_b + k), __ldg(en_b + k), g)), 1.0f);
  } else if (p < L_ola) {
    int i_hi = min(p / stride, N - 1);
    int i_lo = (p - length + 1 <= 0) ? 0 : (p - length + stride) / stride;
    float acc = 0.0f, cnt = 0.0f;
    for (int i = i_lo; i <= i_hi; ++i) {
      int j = p - i * stride;
      if (j >= 0 && j < length) {
        int64_t k = ((int64_t)i * C + c) * length + j;
        acc = __fadd_rn(acc, cfg_mix(__ldg(ec_b + k), __ldg(en_b + k), g));
        cnt = __fadd_rn(cnt, 1.0f);
      }
    }
    e = __fdiv_rn(acc, fmaxf(cnt, 1e-8f));
  }
  const float zn = nz.eta > 0.0f ? philox_normal(nz.seed, nz.clip0 + b, nz.step, (int64_t)c * L + p) : 0.0f;
  z_prev[o] = ddim_update(cf, z_t[o], e, zn);
}

// Same for disjoint windows of 4 (see unchunk1d_w4_kernel): per thread two 16-byte eps reads, z_t / z_prev as 8-byte pairs.
__global__ void __launch_bounds__(256) cfg_ddim_audio_w4_kernel(const float* __restrict__ eps_c, const float* __restrict__ eps_n,
                                                                const float* __restrict__ z_t, float* __restrict__ z_prev,
                                                                int C, int L, int N, float g, int t_now, int t_prev,
                                                                const float* __restrict__ alpha_bar, uint32_t total_w,
                                                                int64_t clip_stride) {
  pdl_wait();
  pdl_launch_dependents();
  const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= total_w) return;
  const DdimCoef cf = ddim_coef(alpha_bar, t_now, t_prev, 0.0f);
  const int wpr = (L + 3) >> 2;
  const uint32_t row = w / wpr;
  const int n = (int)(w - row * wpr);
  const uint32_t b = row / C;
  const int c = (int)(row - b * C);
  float4 e = make_float4(0.f, 0.f, 0.f, 0.f);
  if (n < N) {
    const int64_t k = (int64_t)b * clip_stride + ((int64_t)n * C + c) * 4;
    const float4 ec = __ldg(reinterpret_cast<const float4*>(eps_c + k));
    const float4 en = __ldg(reinterpret_cast<const float4*>(eps_n + k));
    e.x = __fdiv_rn(__fadd_rn(0.0f, cfg_mix(ec.x, en.x, g)), 1.0f);
    e.y = __fdiv_rn(__fadd_rn(0.0f, cfg_mix(ec.y, en.y, g)), 1.0f);
    e.z = __fdiv_rn(__fadd_rn(0.0f, cfg_mix(ec.z, en.z, g)), 1.0f);
    e.w = __fdiv_rn(__fadd_rn(0.0f, cfg_mix(ec.w, en.w, g)), 1.0f);
  }
  const int64_t o = (int64_t)row * L + 4 * n;
  const int left = L - 4 * n;
  const float2 x0 = *reinterpret_cast<const float2*>(z_t + o);
  *reinterpret_cast<float2*>(z_prev + o) = make_float2(ddim_update(cf, x0.x, e.x, 0.0f), ddim_update(cf, x0.y, e.y, 0.0f));
  if (left >= 4) {
    const float2 x1 = *reinterpret_cast<const float2*>(z_t + o + 2);
    *reinterpret_cast<float2*>(z_prev + o + 2) = make_float2(ddim_update(cf, x1.x, e.z, 0.0f), ddim_update(cf, x1.y, e.w, 0.0f));
  }
}

int launch_cfg_untok_ddim(const avd_cfg_ddim_args& a, const float* eps_c, const float* eps_n, const float* z_t,
                          float* z_prev, const float* alpha_bar, int64_t eps_clip_stride, cudaStream_t s) {
  if (a.target == AVD_TARGET_VIDEO) {
    const int64_t clip_elems = (int64_t)a.C * a.T * a.H * a.W;
    const int64_t clip_stride = eps_clip_stride > 0 ? eps_clip_stride : clip_elems;
    TubeDims d{a.B, a.C, a.T, a.H, a.W, a.tt, a.th, a.tw, a.T / a.tt, a.H / a.th, a.W / a.tw};
    int64_t total = (int64_t)a.B * a.C * a.T * a.H * a.W;
    if (total <= 0) return 0;
    uintptr_t al = reinterpret_cast<uintptr_t>(eps_c) | reinterpret_cast<uintptr_t>(eps_n) |
                   reinterpret_cast<uintptr_t>(z_t) | reinterpret_cast<uintptr_t>(z_prev);
    const NoiseArgs nz{a.eta, a.step, a.seed, a.clip0};
    bool v4 = (a.tw % 4 == 0) && ((al & 15) == 0) && !(a.eta > 0.0f);
    if (kind_skipped(KIND_CFG_DDIM)) return 0;
    LaunchScope ls(KIND_CFG_DDIM, s);
    // 32-bit index math when every eps / latent index fits (the index map is a chain of divisions)
    const bool small = (int64_t)a.B * clip_stride < (int64_t(1) << 31) && total < (int64_t(1) << 31);
    if (v4) {
      int64_t n_vec = total / 4;
      if (small)
        AVD_CUDA(launch_pdl(cfg_ddim_video_kernel<4, uint32_t>, dim3(cdiv(n_vec, 512)), dim3(256), 0, s, eps_c, eps_n, z_t,
                            z_prev, d, a.guidance, a.t_now, a.t_prev, alpha_bar, n_vec, clip_elems, clip_stride, nz));
      else
        AVD_CUDA(launch_pdl(cfg_ddim_video_kernel<4, int64_t>, dim3(cdiv(n_vec, 512)), dim3(256), 0, s, eps_c, eps_n, z_t,
                            z_prev, d, a.guidance, a.t_now, a.t_prev, alpha_bar, n_vec, clip_elems, clip_stride, nz));
    } else {
      AVD_CUDA(launch_pdl(cfg_ddim_video_kernel<1, int64_t>, dim3(cdiv(total, 512)), dim3(256), 0, s, eps_c, eps_n, z_t,
                          z_prev, d, a.guidance, a.t_now, a.t_prev, alpha_bar, total, clip_elems, clip_stride, nz));
    }
    AVD_CUDA(cudaGetLastError());
  } else {
    int64_t total = (int64_t)a.B * a.C * a.L;
    if (total <= 0) return 0;
    if (kind_skipped(KIND_CFG_DDIM)) return 0;
    LaunchScope ls(KIND_CFG_DDIM, s);
    const int64_t clip_stride = eps_clip_stride > 0 ? eps_clip_stride : (int64_t)a.Nt * a.C * a.length;
    const int64_t total_w = (int64_t)a.B * a.C * ((a.L + 3) / 4);
    const uintptr_t al16 = reinterpret_cast<uintptr_t>(eps_c) | reinterpret_cast<uintptr_t>(eps_n);
    const uintptr_t al8 = reinterpret_cast<uintptr_t>(z_t) | reinterpret_cast<uintptr_t>(z_prev);
    const NoiseArgs nz{a.eta, a.step, a.seed, a.clip0};
    const bool w4 = a.length == 4 && a.stride == 4 && a.L % 2 == 0 && clip_stride % 4 == 0 && (al16 & 15) == 0 &&
                    (al8 & 7) == 0 && total_w < (int64_t(1) << 31) && !(a.eta > 0.0f);
    if (w4)
      AVD_CUDA(launch_pdl(cfg_ddim_audio_w4_kernel, dim3(cdiv(total_w, 256)), dim3(256), 0, s, eps_c, eps_n, z_t, z_prev, a.C,
                          a.L, a.Nt, a.guidance, a.t_now, a.t_prev, alpha_bar, (uint32_t)total_w, clip_stride));
    else
      AVD_CUDA(launch_pdl(cfg_ddim_audio_kernel, dim3(cdiv(total, 256)), dim3(256), 0, s, eps_c, eps_n, z_t, z_prev, a.C, a.L,
                          a.Nt, a.length, a.stride, a.guidance, a.t_now, a.t_prev, alpha_bar, total, clip_stride, nz));
    AVD_CUDA(cudaGetLastError());
  }
  return 0;
}

// =====================================================================================================
// operand preparation
// =====================================================================================================
// nseg > 0: bf16 segments.  nseg < 0: ONE fp16 value (AVD_PREC_F16), saturated to the finite fp16 range;
// NSEG_F16_X carries the residual-stream scale AVD_F16_XSCALE, NSEG_F16 is unscaled (weights, normalised values).
__device__ __forceinline__ void store_split(__nv_bfloat16* base, int64_t seg_stride, int64_t idx, int nseg, float v) {
  if (nseg < 0) {
    const float sv = fminf(fmaxf(nseg == NSEG_F16_X ? v * AVD_F16_XSCALE : v, -65504.0f), 65504.0f);
    reinterpret_cast<__half*>(base)[idx] = __float2half_rn(sv);
    return;
  }
  __nv_bfloat16 hi, mid, lo;
  split_bf16(v, hi, mid, lo);
  base[idx] = hi;
  if (nseg == 2) {
    // two segments: lo absorbs the whole remainder (round once)
    base[seg_stride + idx] = __float2bfloat16_rn(v - __bfloat162float(hi));
  } else if (nseg == 3) {
    base[seg_stride + idx] = mid;
    base[2 * seg_stride + idx] = lo;
  }
}

__global__ void pack_weight_kernel(const float* __restrict__ W, const float* __restrict__ col_scale,
                                   __nv_bfloat16* __restrict__ out, int N, int K, int nseg) {
  int64_t total = (int64_t)N * K;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    float v = W[i];
    if (col_scale != nullptr) v = __fmul_rn(v, col_scale[i % K]);
    store_split(out, total, i, nseg, v);
  }
}

int launch_pack_weight(const float* W, const float* col_scale, void* out, int N, int K, int nseg, cudaStream_t s) {
  int64_t total = (int64_t)N * K;
  if (total <= 0) return 0;
  LaunchScope ls(KIND_PACK, s);
  pack_weight_kernel<<<std::min(cdiv(total, 256), 4096), 256, 0, s>>>(W, col_scale, static_cast<__nv_bfloat16*>(out),
                                                                      N, K, nseg);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

__global__ void transpose_kernel(const float* __restrict__ W, float* __restrict__ Wt, int N, int K) {
  __shared__ float tile[32][33];
  int k0 = blockIdx.x * 32, n0 = blockIdx.y * 32;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    int n = n0 + r, k = k0 + threadIdx.x;
    tile[r][threadIdx.x] = (n < N && k < K) ? W[(int64_t)n * K + k] : 0.0f;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    int k = k0 + r, n = n0 + threadIdx.x;
    if (k < K && n < N) Wt[(int64_t)k * N + n] = tile[threadIdx.x][r];
  }
}

int launch_transpose(const float* W, float* Wt, int N, int K, cudaStream_t s) {
  if (N <= 0 || K <= 0) return 0;
  dim3 grid(cdiv(K, 32), cdiv(N, 32)), block(32, 8);
  LaunchScope ls(KIND_PACK, s);
  transpose_kernel<<<grid, block, 0, s>>>(W, Wt, N, K);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// one warp per row; each 64-column part: lane handles 2 consecutive columns
__global__ void rows_prepare_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ xa,
                                    float* __restrict__ ssq, int M, int d, int nseg) {
  int row = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  int lane = threadIdx.x % 32;
  if (row >= M) return;
  int parts = d / 64;
  int64_t seg_stride = (int64_t)M * d;
  for (int p = 0; p < parts; ++p) {
    int col = p * 64 + lane * 2;
    float2 v = *reinterpret_cast<const float2*>(x + (int64_t)row * d + col);
    store_split(xa, seg_stride, (int64_t)row * d + col, nseg, v.x);
    store_split(xa, seg_stride, (int64_t)row * d + col + 1, nseg, v.y);
    float s = warp_sum(v.x * v.x + v.y * v.y);
    if (lane == 0) ssq[(int64_t)row * parts + p] = s;
  }
}

int launch_rows_prepare(const float* x, void* xa, float* ssq, int M, int d, int nseg, cudaStream_t s) {
  if (M <= 0) return 0;
  LaunchScope ls(KIND_NORM, s);
  rows_prepare_kernel<<<cdiv(M, 8), 256, 0, s>>>(x, static_cast<__nv_bfloat16*>(xa), ssq, M, d, nseg);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

// RMSNorm of selected rows (mmdt.py:39-42): y = scale * x / (||x|| / sqrt(d) + eps); one warp per output row.
__global__ void rmsnorm_rows_kernel(const float* __restrict__ x, const float* __restrict__ scale, int n_seq, int N,
                                    int row_off, int n_rows, int d, float eps, float* __restrict__ out_f32,
                                    __nv_bfloat16* __restrict__ out_bf16, int nseg) {
  int orow = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  int lane = threadIdx.x % 32;
  int Mo = n_seq * n_rows;
  if (orow >= Mo) return;
  int sq = orow / n_rows, r = orow % n_rows;
  const float* xr = x + ((int64_t)sq * N + row_off + r) * d;
  float ss = 0.0f;
  for (int c = lane * 4; c < d; c += 128) {
    float4 v = *reinterpret_cast<const float4*>(xr + c);
    ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  ss = warp_sum(ss);
  float den = __fadd_rn(__fdiv_rn(__fsqrt_rn(ss), __fsqrt_rn((float)d)), eps);
  int64_t seg_stride = (int64_t)Mo * d;
  for (int c = lane * 4; c < d; c += 128) {
    float4 v = *reinterpret_cast<const float4*>(xr + c);
    float4 sc = *reinterpret_cast<const float4*>(scale + c);
    float y[4] = {__fdiv_rn(__fmul_rn(sc.x, v.x), den), __fdiv_rn(__fmul_rn(sc.y, v.y), den),
                  __fdiv_rn(__fmul_rn(sc.z, v.z), den), __fdiv_rn(__fmul_rn(sc.w, v.w), den)};
    int64_t o = (int64_t)orow * d + c;
    if (out_f32 != nullptr) *reinterpret_cast<float4*>(out_f32 + o) = make_float4(y[0], y[1], y[2], y[3]);
    if (out_bf16 != nullptr) {
#pragma unroll
      for (int k = 0; k < 4; ++k) store_split(out_bf16, seg_stride, o + k, nseg, y[k]);
    }
  }
}

int launch_rmsnorm_rows(const float* x, const float* scale, int n_seq, int N, int row_off, int n_rows, int d, float eps,
                        float* out_f32, void* out_bf16, int nseg, cudaStream_t s) {
  int Mo = n_seq * n_rows;
  if (Mo <= 0) return 0;
  LaunchScope ls(KIND_NORM, s);
  rmsnorm_rows_kernel<<<cdiv(Mo, 8), 256, 0, s>>>(x, scale, n_seq, N, row_off, n_rows, d, eps, out_f32,
                                                  static_cast<__nv_bfloat16*>(out_bf16), nseg);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

// LayerNorm (+ activation: 1 GELU erf, 2 LeakyReLU(0.1), 3 ReLU) rows; one warp per row held in registers (d <= 1024), two-pass
// statistics in fp32.
constexpr int LN_MAX_V4 = 8;   // float4 per lane
__global__ void __launch_bounds__(256) layernorm_act_kernel(const float* __restrict__ y, const float* __restrict__ w,
                                                            const float* __restrict__ b, int M, int d, float eps, int gelu,
                                                            float* __restrict__ out_f32,
                                                            __nv_bfloat16* __restrict__ out_bf16, int nseg) {
  pdl_wait();
  pdl_launch_dependents();
  int row = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
  int lane = threadIdx.x % 32;
  if (row >= M) return;
  const float* yr = y + (int64_t)row * d;
  const int nv = d / 128;                         // float4 per lane (d % 128 == 0 checked by the launcher)
  float4 v[LN_MAX_V4];
  float sum = 0.0f;
#pragma unroll
  for (int i = 0; i < LN_MAX_V4; ++i) {
    if (i < nv) {
      v[i] = *reinterpret_cast<const float4*>(yr + i * 128 + lane * 4);
      sum += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    }
  }
  float mu = warp_sum(sum) / (float)d;
  float sq = 0.0f;
#pragma unroll
  for (int i = 0; i < LN_MAX_V4; ++i) {
    if (i < nv) {
      float a0 = v[i].x - mu, a1 = v[i].y - mu, a2 = v[i].z - mu, a3 = v[i].w - mu;
      sq += a0 * a0 + a1 * a1 + a2 * a2 + a3 * a3;
    }
  }
  float var = warp_sum(sq) / (float)d;
  float rstd = 1.0f / sqrtf(var + eps);
  int64_t seg_stride = (int64_t)M * d;
#pragma unroll
  for (int i = 0; i < LN_MAX_V4; ++i) {
    if (i < nv) {
      const int c = i * 128 + lane * 4;
      float4 ww = *reinterpret_cast<const float4*>(w + c);
      float4 bb = *reinterpret_cast<const float4*>(b + c);
      float o[4] = {(v[i].x - mu) * rstd * ww.x + bb.x, (v[i].y - mu) * rstd * ww.y + bb.y,
                    (v[i].z - mu) * rstd * ww.z + bb.z, (v[i].w - mu) * rstd * ww.w + bb.w};
      if (gelu == 1 && out_f32 == nullptr && (nseg < 0 || nseg == 1)) {
        // one 16-bit value per element: the 1.5e-7 erf approximation disappears in its rounding (2^-9 / 2^-11), and erff
        // (~40 instructions per element) made this pass compute bound: 28.6 us for 52 MB at A2V batch 64
#pragma unroll
        for (int k = 0; k < 4; ++k) o[k] = gelu_fast(o[k]);
      } else if (gelu == 1) {
#pragma unroll
        for (int k = 0; k < 4; ++k) o[k] = gelu_erf(o[k]);
      } else if (gelu == 2) {                      // nn.LeakyReLU(0.1)
#pragma unroll
        for (int k = 0; k < 4; ++k) o[k] = o[k] >= 0.0f ? o[k] : __fmul_rn(o[k], 0.1f);
      } else if (gelu == 3) {                      // nn.ReLU
#pragma unroll
        for (int k = 0; k < 4; ++k) o[k] = fmaxf(o[k], 0.0f);
      }
      int64_t idx = (int64_t)row * d + c;
      if (out_f32 != nullptr) *reinterpret_cast<float4*>(out_f32 + idx) = make_float4(o[0], o[1], o[2], o[3]);
      if (out_bf16 != nullptr && nseg < 0) {       // one fp16 value (AVD_PREC_F16; normalised values: unscaled)
        *reinterpret_cast<uint2*>(out_bf16 + idx) = make_uint2(pack_f16x2(o[0], o[1]), pack_f16x2(o[2], o[3]));
      } else if (out_bf16 != nullptr) {
        for (int sg = 0; sg < nseg; ++sg) {        // hi / mid / lo bf16 terms, 4 columns = one 8-byte store
          uint32_t pk[2];
#pragma unroll
          for (int k = 0; k < 2; ++k) {
            __nv_bfloat16 h0 = __float2bfloat16_rn(o[2 * k]), h1 = __float2bfloat16_rn(o[2 * k + 1]);
            o[2 * k] -= __bfloat162float(h0);
            o[2 * k + 1] -= __bfloat162float(h1);
            pk[k] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
          }
          *reinterpret_cast<uint2*>(out_bf16 + sg * seg_stride + idx) = make_uint2(pk[0], pk[1]);
        }
      }
    }
  }
}

int launch_layernorm_act(const float* y, const float* w, const float* b, int M, int d, float eps, int gelu,
                         float* out_f32, void* out_bf16, int nseg, cudaStream_t s) {
  if (M <= 0) return 0;
  AVD_REQUIRE(d % 128 == 0 && d <= 128 * LN_MAX_V4, "LayerNorm rows must be a multiple of 128 wide, at most 1024");
  AVD_REQUIRE(gelu >= 0 && gelu <= 3, "activation must be 0 (none), 1 (GELU), 2 (LeakyReLU 0.1) or 3 (ReLU)");
  if (kind_skipped(KIND_LN_ACT)) return 0;
  LaunchScope ls(KIND_LN_ACT, s);
  AVD_CUDA(launch_pdl(layernorm_act_kernel, dim3(cdiv(M, 8)), dim3(256), 0, s, y, w, b, M, d, eps, gelu, out_f32,
                      static_cast<__nv_bfloat16*>(out_bf16), nseg));
  return 0;
}

// =====================================================================================================
// Fused input builder (sample_clip.py:322-330 / 363-371): for R consecutive tokens of one clip
//   tok = gather(latent)           (tube patch or 1-D chunk index map)
//   X[:, 0:d_adapt]   = tok @ W^T + b      (fp32 FMA; adapter weights pre-transposed [d_tok][d_adapt])
//   X[:, d_adapt:d]   = timestep embedding row (host table computed with the reference formula)
// and writes the joint-sequence row into x (fp32), xa (bf16 segments) and ssq (row sum of squares in
// part 0, zeros elsewhere) for the cond sequence b and/or the null sequence B+b.
// =====================================================================================================
constexpr int BX_R = 4;      // tokens per CTA
constexpr int BX_J = 4;      // adapter columns per thread (d_adapt <= 256 * BX_J)
constexpr int BX_THREADS = 256;

// KU = weight rows fetched per batch (KU x BX_J independent loads in flight per thread).  The first batch is issued
// BEFORE griddepcontrol.wait (weights do not depend on the previous kernel): with KU = 32 the whole audio adapter
// (d_tok = 32) is in registers when the latent arrives; later batches are double-buffered.
template <int KU>
__global__ void __launch_bounds__(BX_THREADS) build_x_kernel(BuildXArgs a, const float* __restrict__ z,
                                                            const float* __restrict__ wt,
                                                            const float* __restrict__ bias,
                                                            const float* __restrict__ temb, float* __restrict__ x,
                                                            __nv_bfloat16* __restrict__ xa, float* __restrict__ ssq,
                                                            const float* __restrict__ add_rows,
                                                            const float* __restrict__ tadd) {
  extern __shared__ float bx_smem[];
  float* tok = bx_smem;                       // [BX_R][d_tok]
  __shared__ float red[BX_R][BX_THREADS / 32];
  __shared__ float row_ss[BX_R];

  const int groups = (a.n_tok + BX_R - 1) / BX_R;
  const int b = blockIdx.x / groups;
  const int n0 = (blockIdx.x % groups) * BX_R;
  const int tid = threadIdx.x;
  const int ncol = (a.d_adapt + BX_THREADS - 1) / BX_THREADS;   // live adapter columns of this thread (<= BX_J)

  // adapter bias / first weight rows do not depend on the previous kernel
  float bv[BX_J];
#pragma unroll
  for (int j = 0; j < BX_J; ++j) {
    const int c = tid + j * BX_THREADS;
    bv[j] = (c < a.d_adapt) ? __ldg(bias + c) : 0.0f;
  }
  float wv[KU][BX_J];
  auto load_batch = [&](int k0) {
#pragma unroll
    for (int u = 0; u < KU; ++u)
#pragma unroll
      for (int j = 0; j < BX_J; ++j) {
        const int c = tid + j * BX_THREADS;
        wv[u][j] = (j < ncol && c < a.d_adapt && k0 + u < a.d_tok) ? __ldg(wt + (int64_t)(k0 + u) * a.d_adapt + c) : 0.0f;
      }
  };
  load_batch(0);
  pdl_wait();
  pdl_launch_dependents();

  // ---- gather R tokens
  for (int idx = tid; idx < BX_R * a.d_tok; idx += BX_THREADS) {
    int r = idx / a.d_tok, k = idx % a.d_tok;
    int n = n0 + r;
    float v = 0.0f;
    if (n < a.n_tok) {
      int64_t src;
      if (a.modality == 0) {
        int Hh = a.H / a.th, Ww = a.W / a.tw;
        int wk = n % Ww, hj = (n / Ww) % Hh, ti = n / (Ww * Hh);
        int dw = k % a.tw, dh = (k / a.tw) % a.th, dt = (k / (a.tw * a.th)) % a.tt, c = k / (a.tw * a.th * a.tt);
        src = ((((int64_t)b * a.C + c) * a.T + (ti * a.tt + dt)) * a.H + (hj * a.th + dh)) * a.W + (wk * a.tw + dw);
      } else {
        int j = k % a.length, c = k / a.length;
        src = ((int64_t)b * a.C + c) * a.L + (int64_t)n * a.stride + j;
      }
      v = z[src];
    }
    tok[idx] = v;
  }
  __syncthreads();

  // ---- adapter GEMV for R rows: BX_KU weight rows in flight per batch, k ascending (fixed summation order)
  float acc[BX_R][BX_J];
#pragma unroll
  for (int r = 0; r < BX_R; ++r)
#pragma unroll
    for (int j = 0; j < BX_J; ++j) acc[r][j] = 0.0f;
  if constexpr (KU >= 32) {          // single batch (the launcher guarantees d_tok <= KU): consume wv in place
#pragma unroll
    for (int u = 0; u < KU; ++u) {
      if (u < a.d_tok) {
#pragma unroll
        for (int r = 0; r < BX_R; ++r) {
          const float t = tok[r * a.d_tok + u];
#pragma unroll
          for (int j = 0; j < BX_J; ++j) acc[r][j] = fmaf(t, wv[u][j], acc[r][j]);
        }
      }
    }
  } else {
    for (int k0 = 0; k0 < a.d_tok; k0 += KU) {
      float wc[KU][BX_J];
#pragma unroll
      for (int u = 0; u < KU; ++u)
#pragma unroll
        for (int j = 0; j < BX_J; ++j) wc[u][j] = wv[u][j];
      if (k0 + KU < a.d_tok) load_batch(k0 + KU);        // next batch flies while this one is consumed
#pragma unroll
      for (int u = 0; u < KU; ++u) {
        if (k0 + u < a.d_tok) {
#pragma unroll
          for (int r = 0; r < BX_R; ++r) {
            const float t = tok[r * a.d_tok + k0 + u];
#pragma unroll
            for (int j = 0; j < BX_J; ++j) acc[r][j] = fmaf(t, wc[u][j], acc[r][j]);
          }
        }
      }
    }
  }

  // ---- write rows + sums of squares
  int64_t seg_stride = a.M * (int64_t)a.d;
#pragma unroll
  for (int r = 0; r < BX_R; ++r) {
    int n = n0 + r;
    bool live = n < a.n_tok;
    float part = 0.0f;
    int64_t row_c = ((int64_t)b * a.N + a.row_off + n) * a.d;
    int64_t row_n = ((int64_t)(a.B + b) * a.N + a.row_off + n) * a.d;
#pragma unroll
    for (int j = 0; j < BX_J; ++j) {
      int c = tid + j * BX_THREADS;
      if (live && c < a.d_adapt) {
        float v = acc[r][j] + bv[j];
        if (add_rows != nullptr) v += __ldg(add_rows + (int64_t)n * a.d_adapt + c);   // positional + modality embeddings
        if (tadd != nullptr) v += __ldg(tadd + c);                                    // additive time embedding
        part += v * v;
        if (a.write_cond) { x[row_c + c] = v; store_split(xa, seg_stride, row_c + c, a.nseg, v); }
        if (a.write_null) { x[row_n + c] = v; store_split(xa, seg_stride, row_n + c, a.nseg, v); }
      }
    }
    for (int c2 = tid; c2 < a.tstep_dim; c2 += BX_THREADS) {
      if (live) {
        float v = __ldg(temb + c2);
        int c = a.d_adapt + c2;
        part += v * v;
        if (a.write_cond) { x[row_c + c] = v; store_split(xa, seg_stride, row_c + c, a.nseg, v); }
        if (a.write_null) { x[row_n + c] = v; store_split(xa, seg_stride, row_n + c, a.nseg, v); }
      }
    }
    part = warp_sum(part);
    if ((tid & 31) == 0) red[r][tid >> 5] = part;
  }
  __syncthreads();
  if (tid < BX_R) {
    float s = 0.0f;
#pragma unroll
    for (int wv = 0; wv < BX_THREADS / 32; ++wv) s += red[tid][wv];
    row_ss[tid] = s;
  }
  __syncthreads();
  for (int idx = tid; idx < BX_R * a.ssq_parts; idx += BX_THREADS) {
    int r = idx / a.ssq_parts, p = idx % a.ssq_parts;
    int n = n0 + r;
    if (n < a.n_tok) {
      float v = (p == 0) ? row_ss[r] : 0.0f;
      if (a.write_cond) ssq[((int64_t)b * a.N + a.row_off + n) * a.ssq_parts + p] = v;
      if (a.write_null) ssq[((int64_t)(a.B + b) * a.N + a.row_off + n) * a.ssq_parts + p] = v;
    }
  }
}

// The same for many tokens (A2V: 96 video tokens x batch per step): 32 tokens per CTA so that each adapter weight, read
// once per CTA from L2, feeds 32 rows (the 4-token kernel re-reads the 256 KB video adapter 1536 times at batch 64:
// 393 MB of L2 traffic, 263 us).  Token values sit TRANSPOSED in smem ([k][32 rows]: one k = 8 broadcast 16-byte loads);
// thread = adapter column(s), k ascending -> the same fp32 summation order as build_x_kernel.
constexpr int BXW_R = 32;
constexpr int BXW_JMAX = 2;    // adapter columns per thread (d_adapt <= 512)
template <int BXW_J>
__global__ void __launch_bounds__(BX_THREADS) build_x_wide_kernel(BuildXArgs a, const float* __restrict__ z,
                                                                 const float* __restrict__ wt,
                                                                 const float* __restrict__ bias,
                                                                 const float* __restrict__ temb, float* __restrict__ x,
                                                                 __nv_bfloat16* __restrict__ xa, float* __restrict__ ssq,
                                                                 const float* __restrict__ add_rows,
                                                                 const float* __restrict__ tadd) {
  extern __shared__ float bx_smem[];
  float* tok = bx_smem;                       // [d_tok][BXW_R]
  __shared__ float red[BXW_R][BX_THREADS / 32];
  const int groups = (a.n_tok + BXW_R - 1) / BXW_R;
  const int b = blockIdx.x / groups;
  const int n0 = (blockIdx.x % groups) * BXW_R;
  const int tid = threadIdx.x;
  const int ncol = (a.d_adapt + BX_THREADS - 1) / BX_THREADS;
  float bv[BXW_J];
#pragma unroll
  for (int j = 0; j < BXW_J; ++j) {
    const int c = tid + j * BX_THREADS;
    bv[j] = (c < a.d_adapt) ? __ldg(bias + c) : 0.0f;
  }
  pdl_wait();
  pdl_launch_dependents();
  for (int idx = tid; idx < BXW_R * a.d_tok; idx += BX_THREADS) {
    const int r = idx % BXW_R, k = idx / BXW_R;           // consecutive threads = consecutive tokens: conflict-free smem stores
    const int n = n0 + r;
    float v = 0.0f;
    if (n < a.n_tok) {
      int64_t src;
      if (a.modality == 0) {
        const int Hh = a.H / a.th, Ww = a.W / a.tw;
        const int wk = n % Ww, hj = (n / Ww) % Hh, ti = n / (Ww * Hh);
        const int dw = k % a.tw, dh = (k / a.tw) % a.th, dt = (k / (a.tw * a.th)) % a.tt, c = k / (a.tw * a.th * a.tt);
        src = ((((int64_t)b * a.C + c) * a.T + (ti * a.tt + dt)) * a.H + (hj * a.th + dh)) * a.W + (wk * a.tw + dw);
      } else {
        const int j = k % a.length, c = k / a.length;
        src = ((int64_t)b * a.C + c) * a.L + (int64_t)n * a.stride + j;
      }
      v = z[src];
    }
    tok[k * BXW_R + r] = v;
  }
  __syncthreads();
  float acc[BXW_R][BXW_J];
#pragma unroll
  for (int r = 0; r < BXW_R; ++r)
#pragma unroll
    for (int j = 0; j < BXW_J; ++j) acc[r][j] = 0.0f;
  // adapter weights: KU rows per batch, the NEXT batch's loads in flight while this one is consumed (the loads come
  // from L2, ~600 cycles; without the double buffer every batch would wait them out)
  constexpr int KU = 8;
  float wn[KU][BXW_J];
  auto load_batch = [&](int k0) {
#pragma unroll
    for (int u = 0; u < KU; ++u)
#pragma unroll
      for (int j = 0; j < BXW_J; ++j) {
        const int c = tid + j * BX_THREADS;
        wn[u][j] = (j < ncol && c < a.d_adapt && k0 + u < a.d_tok) ? __ldg(wt + (int64_t)(k0 + u) * a.d_adapt + c) : 0.0f;
      }
  };
  load_batch(0);
  for (int k0 = 0; k0 < a.d_tok; k0 += KU) {
    float w[KU][BXW_J];
#pragma unroll
    for (int u = 0; u < KU; ++u)
#pragma unroll
      for (int j = 0; j < BXW_J; ++j) w[u][j] = wn[u][j];
    if (k0 + KU < a.d_tok) load_batch(k0 + KU);
#pragma unroll
    for (int u = 0; u < KU; ++u) {
      if (k0 + u < a.d_tok) {
        const float4* t4 = reinterpret_cast<const float4*>(tok + (k0 + u) * BXW_R);
#pragma unroll
        for (int q = 0; q < BXW_R / 4; ++q) {
          const float4 t = t4[q];
#pragma unroll
          for (int j = 0; j < BXW_J; ++j) {
            if (j < ncol) {
              acc[4 * q][j] = fmaf(t.x, w[u][j], acc[4 * q][j]);
              acc[4 * q + 1][j] = fmaf(t.y, w[u][j], acc[4 * q + 1][j]);
              acc[4 * q + 2][j] = fmaf(t.z, w[u][j], acc[4 * q + 2][j]);
              acc[4 * q + 3][j] = fmaf(t.w, w[u][j], acc[4 * q + 3][j]);
            }
          }
        }
      }
    }
  }
  const int64_t seg_stride = a.M * (int64_t)a.d;
#pragma unroll
  for (int r = 0; r < BXW_R; ++r) {
    const int n = n0 + r;
    const bool live = n < a.n_tok;
    float part = 0.0f;
    const int64_t row_c = ((int64_t)b * a.N + a.row_off + n) * a.d;
    const int64_t row_n = ((int64_t)(a.B + b) * a.N + a.row_off + n) * a.d;
#pragma unroll
    for (int j = 0; j < BXW_J; ++j) {
      const int c = tid + j * BX_THREADS;
      if (live && j < ncol && c < a.d_adapt) {
        float v = acc[r][j] + bv[j];
        if (add_rows != nullptr) v += __ldg(add_rows + (int64_t)n * a.d_adapt + c);
        if (tadd != nullptr) v += __ldg(tadd + c);
        part += v * v;
        if (a.write_cond) { x[row_c + c] = v; store_split(xa, seg_stride, row_c + c, a.nseg, v); }
        if (a.write_null) { x[row_n + c] = v; store_split(xa, seg_stride, row_n + c, a.nseg, v); }
      }
    }
    for (int c2 = tid; c2 < a.tstep_dim; c2 += BX_THREADS) {
      if (live) {
        const float v = __ldg(temb + c2);
        const int c = a.d_adapt + c2;
        part += v * v;
        if (a.write_cond) { x[row_c + c] = v; store_split(xa, seg_stride, row_c + c, a.nseg, v); }
        if (a.write_null) { x[row_n + c] = v; store_split(xa, seg_stride, row_n + c, a.nseg, v); }
      }
    }
    part = warp_sum(part);
    if ((tid & 31) == 0) red[r][tid >> 5] = part;
  }
  __syncthreads();
  for (int idx = tid; idx < BXW_R * a.ssq_parts; idx += BX_THREADS) {
    const int r = idx / a.ssq_parts, p = idx % a.ssq_parts;
    const int n = n0 + r;
    if (n < a.n_tok) {
      float v = 0.0f;
      if (p == 0) {
#pragma unroll
        for (int wv = 0; wv < BX_THREADS / 32; ++wv) v += red[r][wv];
      }
      if (a.write_cond) ssq[((int64_t)b * a.N + a.row_off + n) * a.ssq_parts + p] = v;
      if (a.write_null) ssq[((int64_t)(a.B + b) * a.N + a.row_off + n) * a.ssq_parts + p] = v;
    }
  }
}

// ---- tensor-core route of the input builder (many wide tokens: the A2V video rows at large batch) ----------------------
// (1) tokenize_split_kernel: gather the latent into the token matrix as bf16 segments (the GEMM's A operand);
// (2) launch_gemm: Y = tok W^T + b on tcgen05 (split-bf16 passes: the adapter keeps ~fp32 products);
// (3) assemble_x_kernel: Y (+ additive rows) and the timestep-embedding columns -> x / xa / ssq rows of the cond and null
//     sequences.  One warp per token.
__global__ void __launch_bounds__(256) tokenize_split_kernel(BuildXArgs a, const float* __restrict__ z,
                                                             __nv_bfloat16* __restrict__ tok, int nseg) {
  pdl_wait();
  pdl_launch_dependents();
  const int64_t total = (int64_t)a.B * a.n_tok * a.d_tok;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int k = (int)(i % a.d_tok);
    const int64_t r = i / a.d_tok;
    const int n = (int)(r % a.n_tok), b = (int)(r / a.n_tok);
    int64_t src;
    if (a.modality == 0) {
      const int Hh = a.H / a.th, Ww = a.W / a.tw;
      const int wk = n % Ww, hj = (n / Ww) % Hh, ti = n / (Ww * Hh);
      const int dw = k % a.tw, dh = (k / a.tw) % a.th, dt = (k / (a.tw * a.th)) % a.tt, c = k / (a.tw * a.th * a.tt);
      src = ((((int64_t)b * a.C + c) * a.T + (ti * a.tt + dt)) * a.H + (hj * a.th + dh)) * a.W + (wk * a.tw + dw);
    } else {
      const int j = k % a.length, c = k / a.length;
      src = ((int64_t)b * a.C + c) * a.L + (int64_t)n * a.stride + j;
    }
    store_split(tok, total, i, nseg, z[src]);
  }
}

__global__ void __launch_bounds__(256) assemble_x_kernel(BuildXArgs a, const float* __restrict__ y,
                                                         const float* __restrict__ temb, float* __restrict__ x,
                                                         __nv_bfloat16* __restrict__ xa, float* __restrict__ ssq,
                                                         const float* __restrict__ add_rows, const float* __restrict__ tadd) {
  pdl_wait();
  pdl_launch_dependents();
  const int row = blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);     // token index over the batch
  const int lane = threadIdx.x & 31;
  if (row >= a.B * a.n_tok) return;
  const int n = row % a.n_tok, b = row / a.n_tok;
  const int64_t seg_stride = a.M * (int64_t)a.d;
  const int64_t row_c = ((int64_t)b * a.N + a.row_off + n) * a.d;
  const int64_t row_n = ((int64_t)(a.B + b) * a.N + a.row_off + n) * a.d;
  float part = 0.0f;
  for (int c = lane; c < a.d; c += 32) {
    float v;
    if (c < a.d_adapt) {
      v = y[(int64_t)row * a.d_adapt + c];
      if (add_rows != nullptr) v += __ldg(add_rows + (int64_t)n * a.d_adapt + c);
      if (tadd != nullptr) v += __ldg(tadd + c);
    } else {
      v = __ldg(temb + c - a.d_adapt);
    }
    part = fmaf(v, v, part);
    if (a.write_cond) { x[row_c + c] = v; store_split(xa, seg_stride, row_c + c, a.nseg, v); }
    if (a.write_null) { x[row_n + c] = v; store_split(xa, seg_stride, row_n + c, a.nseg, v); }
  }
  part = warp_sum(part);
  for (int p = lane; p < a.ssq_parts; p += 32) {
    const float v = p == 0 ? part : 0.0f;
    if (a.write_cond) ssq[((int64_t)b * a.N + a.row_off + n) * a.ssq_parts + p] = v;
    if (a.write_null) ssq[((int64_t)(a.B + b) * a.N + a.row_off + n) * a.ssq_parts + p] = v;
  }
}

int launch_build_x_tc(const BuildXArgs& a, const float* z, const void* w_pack, int w_precision, const float* bias,
                      const float* temb, float* x, void* xa, float* ssq, void* tok_scratch, float* y_scratch,
                      cudaStream_t s, const float* add_rows, const float* tadd) {
  AVD_REQUIRE(a.d_tok % 64 == 0 && a.d_adapt % 32 == 0, "tensor-core input builder needs d_tok % 64 == 0 and d_adapt % 32 == 0");
  AVD_REQUIRE(a.tstep_dim == 0 || temb != nullptr, "missing timestep-embedding row");
  if (a.B <= 0 || a.n_tok <= 0) return 0;
  const int M = a.B * a.n_tok;
  const int nseg = nseg_of(w_precision);
  AVD_REQUIRE(nseg == 2 || nseg == 3, "adapter operands are split bf16 (bf16x3 or fp32 mode)");
  if (!kind_skipped(KIND_BUILD_X)) {
    LaunchScope ls(KIND_BUILD_X, s);
    const int64_t total = (int64_t)M * a.d_tok;
    AVD_CUDA(launch_pdl(tokenize_split_kernel, dim3((unsigned)std::min<int64_t>(cdiv(total, 256), 148 * 8)), dim3(256), 0, s, a,
                        z, static_cast<__nv_bfloat16*>(tok_scratch), nseg));
  }
  avd_gemm_args g{};
  g.M = M; g.N = a.d_adapt; g.K = a.d_tok; g.precision = w_precision;
  g.A = tok_scratch; g.W = w_pack; g.bias = bias; g.out_f32 = y_scratch;
  AVD_TRY(launch_gemm(g, s));
  if (kind_skipped(KIND_BUILD_X)) return 0;
  LaunchScope ls(KIND_BUILD_X, s);
  AVD_CUDA(launch_pdl(assemble_x_kernel, dim3(cdiv(M, 8)), dim3(256), 0, s, a, static_cast<const float*>(y_scratch), temb, x,
                      static_cast<__nv_bfloat16*>(xa), ssq, add_rows, tadd));
  return 0;
}

int launch_build_x(const BuildXArgs& a, const float* z, const float* wt, const float* bias, const float* temb,
                   float* x, void* xa, float* ssq, cudaStream_t s, const float* add_rows, const float* tadd) {
  AVD_REQUIRE(a.d_adapt <= BX_THREADS * BX_J, "adapter width too large for build_x");
  AVD_REQUIRE(a.d_tok * BX_R * 4 <= 160 * 1024, "token width too large for build_x");
  AVD_REQUIRE(a.tstep_dim == 0 || temb != nullptr, "missing timestep-embedding row");
  if (a.B <= 0 || a.n_tok <= 0) return 0;
  if (kind_skipped(KIND_BUILD_X)) return 0;
  // many wide tokens (A2V video rows at batch >= 32) when the tensor-core route is not available: 32 tokens per CTA
  const int wide_opt = options().bx_wide;
  const bool wide = a.d_adapt <= BX_THREADS * BXW_JMAX && a.d_tok * BXW_R * 4 <= 96 * 1024 &&
                    (wide_opt == 1 || (wide_opt != 0 && a.d_tok > 32 && (int64_t)a.B * a.n_tok >= 3072));   // measured: wins only at B >= 32
  if (wide) {
    const int groups = (a.n_tok + BXW_R - 1) / BXW_R;
    const size_t smem = (size_t)BXW_R * a.d_tok * sizeof(float);
    if (a.d_adapt <= BX_THREADS) {
      if (smem > 32 * 1024) AVD_TRY(ensure_dyn_smem(reinterpret_cast<const void*>(build_x_wide_kernel<1>), 96 * 1024));
      LaunchScope ls(KIND_BUILD_X, s);
      AVD_CUDA(launch_pdl(build_x_wide_kernel<1>, dim3(a.B * groups), dim3(BX_THREADS), smem, s, a, z, wt, bias, temb, x,
                          static_cast<__nv_bfloat16*>(xa), ssq, add_rows, tadd));
    } else {
      if (smem > 32 * 1024) AVD_TRY(ensure_dyn_smem(reinterpret_cast<const void*>(build_x_wide_kernel<2>), 96 * 1024));
      LaunchScope ls(KIND_BUILD_X, s);
      AVD_CUDA(launch_pdl(build_x_wide_kernel<2>, dim3(a.B * groups), dim3(BX_THREADS), smem, s, a, z, wt, bias, temb, x,
                          static_cast<__nv_bfloat16*>(xa), ssq, add_rows, tadd));
    }
    return 0;
  }
  int groups = (a.n_tok + BX_R - 1) / BX_R;
  size_t smem = (size_t)BX_R * a.d_tok * sizeof(float);
  if (smem > 48 * 1024) AVD_TRY(ensure_dyn_smem(reinterpret_cast<const void*>(build_x_kernel<8>), 160 * 1024));
  LaunchScope ls(KIND_BUILD_X, s);
  if (a.d_tok <= 32)
    AVD_CUDA(launch_pdl(build_x_kernel<32>, dim3(a.B * groups), dim3(BX_THREADS), smem, s, a, z, wt, bias, temb, x,
                        static_cast<__nv_bfloat16*>(xa), ssq, add_rows, tadd));
  else
    AVD_CUDA(launch_pdl(build_x_kernel<8>, dim3(a.B * groups), dim3(BX_THREADS), smem, s, a, z, wt, bias, temb, x,
                        static_cast<__nv_bfloat16*>(xa), ssq, add_rows, tadd));
  return 0;
}

}  // namespace avd
