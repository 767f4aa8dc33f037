// avd_vae.cu -- SURVEY §8f N1: VideoVAE.encode / decode (avdiff/models/encoders/vae_video3d.py:164-214) as CUDA kernels: the
// step either side of the sampling loop for the video modality.
//
//   encode  x [B,3,T,H,W] -> enc_net: blocks of Conv3d(3^3, pad 1) + GELU + GroupNorm(8) (vae_video3d.py:78-84, 92-96)
//                         -> AvgPool3d(t_down, s_down, s_down) (:98) -> to_lat / to_mu 1x1x1 (:102-106)  -> z [B,Cv,T',H',W']
//   decode  z -> from_lat 1x1x1 (:110) -> trilinear upsample, align_corners=False (:210) -> dec_net blocks (:111-114)
//             -> to_img 1x1x1 (:119) -> sigmoid | tanh (:122-127)                                         -> x_hat [B,3,T,H,W]
//
// Kernels (fp32, C = 64 trunk channels = 8 GroupNorm groups of 8):
//   conv3d_c64_kernel<MODE>   one CTA = 1 x 8 x 32 output voxels x 64 channels; the input tile (+halo) and the weights come
//                             through shared memory 8 input channels at a time; thread = (w column, group of 8 output
//                             channels) with 8 h x 8 co accumulators.  The LOADER is the fusion point:
//                               RAW   the video itself (first encoder block, C_in = 3)
//                               GN    the previous block's GELU output with its GroupNorm applied on the fly as a
//                                     per-(clip, channel) affine (zero padding stays zero: padding follows the norm)
//                               TRI   the trilinear upsample of from_lat(z), interpolated on the fly from the latent grid
//                             Epilogue: bias + GELU(erf) -> y, plus per-CTA (sum, sum of squares) of each GroupNorm group.
//   gn_finalize_kernel        per (clip, group): reduce the per-CTA partials in fp64 in a fixed order -> per-channel
//                             affine a = rstd * gamma, s = beta - mean * rstd * gamma.
//   vae_pool_lat_kernel       AvgPool3d + last GroupNorm (an affine commutes with the mean) + to_lat.
//   vae_from_lat_kernel, vae_to_img_kernel   the 1x1x1 convolutions (+ last GroupNorm + output activation).
// Everything is deterministic (no atomics).  The 64 -> 64 convolutions are CUDA-core fp32 direct convolutions; they run
// once per clip, outside the 50-step loop (DESIGN.md section 10 gives their measured cost next to a sampling call).
#include "avd_internal.h"
#include "avd_ptx.cuh"

namespace avd {

constexpr int V_C = 64;              // trunk channels
constexpr int V_TH = 8, V_TW = 32;   // output tile (one t plane)
constexpr int V_CI = 8;              // input channels per smem chunk
constexpr int V_IH = V_TH + 2, V_IW = V_TW + 2;
constexpr int V_IN_FLOATS = V_CI * 3 * V_IH * V_IW;      // 8160
constexpr int V_W_FLOATS = V_CI * 27 * V_C;              // 13824
constexpr int V_THREADS = 256;

enum { VL_RAW = 0, VL_GN = 1, VL_TRI = 2 };

struct Conv3dArgs {
  int B, Cin, T, H, W;               // output (= input) grid
  const float* in;                   // RAW / GN: [B,Cin,T,H,W]; TRI: latent-resolution [B,64,Tl,Hl,Wl]
  const float* gn_a; const float* gn_s;   // GN: per (b, channel) affine [B,64]
  int Tl, Hl, Wl;                    // TRI: latent grid
  float st, sh, sw;                  // TRI: in/out scales (ATen area_pixel_compute_scale, align_corners = False)
  const float* w; const float* bias; // [64,Cin,3,3,3], [64]
  float* out;                        // [B,64,T,H,W]
  float* partial;                    // [B][tiles][8][2] (sum, sumsq) per GroupNorm group
  int tiles_h, tiles_w;
};

__device__ __forceinline__ void tri_axis(int dst, float scale, int size, int& i0, int& i1, float& l1) {
  // ATen area_pixel_compute_source_index(scale, dst, align_corners=false, cubic=false) + guard_index_and_lambda
  float src = scale * ((float)dst + 0.5f) - 0.5f;
  if (src < 0.0f) src = 0.0f;
  i0 = min((int)src, size - 1);
  i1 = i0 + ((i0 < size - 1) ? 1 : 0);
  l1 = fminf(fmaxf(src - (float)i0, 0.0f), 1.0f);
}

template <int MODE>
__global__ void __launch_bounds__(V_THREADS) conv3d_c64_kernel(Conv3dArgs a) {
  extern __shared__ __align__(16) float v_smem[];
  float* in_s = v_smem;                      // [V_CI][3][V_IH][V_IW]
  float* w_s = v_smem + V_IN_FLOATS;         // [V_CI][27][64]   (co fastest)
  const int tid = threadIdx.x, vg = tid & 31, cog = tid >> 5, co0 = cog * 8;
  const int tile = blockIdx.x;
  const int tw = tile % a.tiles_w, th = (tile / a.tiles_w) % a.tiles_h, t = tile / (a.tiles_w * a.tiles_h);
  const int b = blockIdx.y;
  const int h0 = th * V_TH, w0 = tw * V_TW;
  const int64_t plane = (int64_t)a.H * a.W, vol = plane * a.T;
  float acc[V_TH][8];
#pragma unroll
  for (int h = 0; h < V_TH; ++h)
#pragma unroll
    for (int c = 0; c < 8; ++c) acc[h][c] = 0.0f;

  for (int c0 = 0; c0 < a.Cin; c0 += V_CI) {
    const int nci = min(V_CI, a.Cin - c0);
    __syncthreads();
    // ---- input tile + halo
    for (int idx = tid; idx < nci * 3 * V_IH * V_IW; idx += V_THREADS) {
      const int iw = idx % V_IW, ih = (idx / V_IW) % V_IH, it = (idx / (V_IW * V_IH)) % 3, ci = idx / (V_IW * V_IH * 3);
      const int tt = t + it - 1, hh = h0 + ih - 1, ww = w0 + iw - 1;
      float v = 0.0f;
      if (tt >= 0 && tt < a.T && hh >= 0 && hh < a.H && ww >= 0 && ww < a.W) {
        const int c = c0 + ci;
        if (MODE == VL_RAW) {
          v = __ldg(a.in + ((int64_t)b * a.Cin + c) * vol + tt * plane + (int64_t)hh * a.W + ww);
        } else if (MODE == VL_GN) {
          const float y = __ldg(a.in + ((int64_t)b * a.Cin + c) * vol + tt * plane + (int64_t)hh * a.W + ww);
          v = fmaf(y, __ldg(a.gn_a + b * V_C + c), __ldg(a.gn_s + b * V_C + c));
        } else {
          int t0i, t1i, h0i, h1i, w0i, w1i;
          float lt, lh, lw;
          tri_axis(tt, a.st, a.Tl, t0i, t1i, lt);
          tri_axis(hh, a.sh, a.Hl, h0i, h1i, lh);
          tri_axis(ww, a.sw, a.Wl, w0i, w1i, lw);
          const float* p = a.in + ((int64_t)b * V_C + c) * ((int64_t)a.Tl * a.Hl * a.Wl);
          const int64_t pl = (int64_t)a.Hl * a.Wl;
          auto at = [&](int ti, int hi, int wi) { return __ldg(p + ti * pl + (int64_t)hi * a.Wl + wi); };
          const float mt = 1.0f - lt, mh = 1.0f - lh, mw = 1.0f - lw;
          // the summation tree of ATen's upsample_trilinear3d CPU kernel
          v = mt * (mh * (mw * at(t0i, h0i, w0i) + lw * at(t0i, h0i, w1i)) + lh * (mw * at(t0i, h1i, w0i) + lw * at(t0i, h1i, w1i))) +
              lt * (mh * (mw * at(t1i, h0i, w0i) + lw * at(t1i, h0i, w1i)) + lh * (mw * at(t1i, h1i, w0i) + lw * at(t1i, h1i, w1i)));
        }
      }
      in_s[idx] = v;
    }
    // ---- weights of these input channels: global [co][Cin][27] -> smem [ci][tap][co]
    for (int idx = tid; idx < nci * 27 * V_C; idx += V_THREADS) {
      const int tap = idx % 27, ci = (idx / 27) % nci, co = idx / (27 * nci);
      w_s[(ci * 27 + tap) * V_C + co] = __ldg(a.w + ((int64_t)co * a.Cin + c0 + ci) * 27 + tap);
    }
    __syncthreads();
    for (int ci = 0; ci < nci; ++ci) {
#pragma unroll
      for (int dt = 0; dt < 3; ++dt) {
#pragma unroll
        for (int dw = 0; dw < 3; ++dw) {
          float x[V_IH];
          const float* col = in_s + ((ci * 3 + dt) * V_IH) * V_IW + vg + dw;
#pragma unroll
          for (int ih = 0; ih < V_IH; ++ih) x[ih] = col[ih * V_IW];
#pragma unroll
          for (int dh = 0; dh < 3; ++dh) {
            const float4* wp = reinterpret_cast<const float4*>(w_s + (ci * 27 + (dt * 3 + dh) * 3 + dw) * V_C + co0);
            const float4 wa = wp[0], wb = wp[1];
            const float wv[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
            for (int h = 0; h < V_TH; ++h)
#pragma unroll
              for (int c = 0; c < 8; ++c) acc[h][c] = fmaf(x[h + dh], wv[c], acc[h][c]);
          }
        }
      }
    }
  }
  // ---- epilogue: bias + GELU, store, GroupNorm partial statistics of this warp's channel group
  float s1 = 0.0f, s2 = 0.0f;
  const int ww = w0 + vg;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const float bias = __ldg(a.bias + co0 + c);
    float* dst = a.out + ((int64_t)b * V_C + co0 + c) * vol + t * plane;
#pragma unroll
    for (int h = 0; h < V_TH; ++h) {
      const int hh = h0 + h;
      if (hh < a.H && ww < a.W) {
        const float v = gelu_erf(acc[h][c] + bias);
        dst[(int64_t)hh * a.W + ww] = v;
        s1 += v;
        s2 = fmaf(v, v, s2);
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    s2 += __shfl_xor_sync(0xffffffffu, s2, o);
  }
  if (vg == 0) {
    float* pp = a.partial + (((int64_t)b * gridDim.x + tile) * 8 + cog) * 2;
    pp[0] = s1;
    pp[1] = s2;
  }
}

// per (clip, group): fixed-order fp64 reduction of the per-CTA partials -> per-channel affine of nn.GroupNorm(8, 64, eps)
__global__ void gn_finalize_kernel(const float* __restrict__ partial, int tiles, double count, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float eps, float* __restrict__ gn_a,
                                   float* __restrict__ gn_s) {
  const int g = blockIdx.x, b = blockIdx.y;
  __shared__ double r1[256], r2[256];
  double s1 = 0.0, s2 = 0.0;
  for (int i = threadIdx.x; i < tiles; i += blockDim.x) {
    const float* pp = partial + (((int64_t)b * tiles + i) * 8 + g) * 2;
    s1 += (double)pp[0];
    s2 += (double)pp[1];
  }
  r1[threadIdx.x] = s1;
  r2[threadIdx.x] = s2;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) { r1[threadIdx.x] += r1[threadIdx.x + o]; r2[threadIdx.x] += r2[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x < 8) {
    const double mean = r1[0] / count;
    const double var = fmax(r2[0] / count - mean * mean, 0.0);
    const float rstd = (float)(1.0 / sqrt(var + (double)eps));
    const int c = g * 8 + threadIdx.x;
    const float ga = gamma[c] * rstd;
    gn_a[b * V_C + c] = ga;
    gn_s[b * V_C + c] = beta[c] - (float)mean * ga;
  }
}

// AvgPool3d(td, sd, sd) of y, the last GroupNorm as an affine of the pooled value, to_lat 1x1x1 (64 -> Cv)
__global__ void __launch_bounds__(256) vae_pool_lat_kernel(const float* __restrict__ y, const float* __restrict__ gn_a,
                                                           const float* __restrict__ gn_s, const float* __restrict__ wl,
                                                           const float* __restrict__ bl, float* __restrict__ z, int T, int H,
                                                           int W, int td, int sd, int Cv) {
  const int Tl = T / td, Hl = H / sd, Wl = W / sd;
  const int v = blockIdx.x, b = blockIdx.y;
  const int wl_i = v % Wl, hl_i = (v / Wl) % Hl, tl_i = v / (Wl * Hl);
  const int win = td * sd * sd;
  __shared__ float red[V_C][8];
  __shared__ float pooled[V_C];
  const int64_t plane = (int64_t)H * W, vol = plane * T;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int c = 0; c < V_C; ++c) {
    float s = 0.0f;
    for (int i = threadIdx.x; i < win; i += blockDim.x) {
      const int dw = i % sd, dh = (i / sd) % sd, dt = i / (sd * sd);
      s += __ldg(y + ((int64_t)b * V_C + c) * vol + (int64_t)(tl_i * td + dt) * plane + (int64_t)(hl_i * sd + dh) * W + wl_i * sd + dw);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) red[c][warp] = s;
  }
  __syncthreads();
  if (threadIdx.x < V_C) {
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += red[threadIdx.x][k];
    pooled[threadIdx.x] = fmaf(s / (float)win, gn_a[b * V_C + threadIdx.x], gn_s[b * V_C + threadIdx.x]);
  }
  __syncthreads();
  if ((int)threadIdx.x < Cv) {
    float acc = __ldg(bl + threadIdx.x);
    for (int c = 0; c < V_C; ++c) acc = fmaf(__ldg(wl + threadIdx.x * V_C + c), pooled[c], acc);
    z[(((int64_t)b * Cv + threadIdx.x) * Tl + tl_i) * Hl * Wl + (int64_t)hl_i * Wl + wl_i] = acc;
  }
}

// from_lat: 1x1x1 conv Cv -> 64 on the latent grid
__global__ void vae_from_lat_kernel(const float* __restrict__ z, const float* __restrict__ wf, const float* __restrict__ bf,
                                    float* __restrict__ h, int Cv, int64_t nvox) {
  const int b = blockIdx.y;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvox * V_C; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t vx = i % nvox;
    const int c = (int)(i / nvox);
    float acc = __ldg(bf + c);
    for (int q = 0; q < Cv; ++q) acc = fmaf(__ldg(wf + c * Cv + q), __ldg(z + ((int64_t)b * Cv + q) * nvox + vx), acc);
    h[((int64_t)b * V_C + c) * nvox + vx] = acc;
  }
}

// last GroupNorm + to_img 1x1x1 (64 -> Cimg) + sigmoid (act = 0) / tanh (act = 1)
__global__ void __launch_bounds__(256) vae_to_img_kernel(const float* __restrict__ y, const float* __restrict__ gn_a,
                                                         const float* __restrict__ gn_s, const float* __restrict__ wi,
                                                         const float* __restrict__ bi, float* __restrict__ x, int Cimg,
                                                         int64_t nvox, int act) {
  __shared__ float ws[4 * V_C], ga[V_C], gs[V_C];
  const int b = blockIdx.y;
  for (int i = threadIdx.x; i < Cimg * V_C; i += blockDim.x) ws[i] = wi[i];
  for (int i = threadIdx.x; i < V_C; i += blockDim.x) { ga[i] = gn_a[b * V_C + i]; gs[i] = gn_s[b * V_C + i]; }
  __syncthreads();
  for (int64_t vx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; vx < nvox; vx += (int64_t)gridDim.x * blockDim.x) {
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int c = 0; c < V_C; ++c) {
      const float v = fmaf(__ldg(y + ((int64_t)b * V_C + c) * nvox + vx), ga[c], gs[c]);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (k < Cimg) acc[k] = fmaf(ws[k * V_C + c], v, acc[k]);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (k < Cimg) {
        const float o = acc[k] + __ldg(bi + k);
        x[((int64_t)b * Cimg + k) * nvox + vx] = act == 0 ? 1.0f / (1.0f + expf(-o)) : tanhf(o);
      }
  }
}

template <int MODE>
static int launch_conv3d(const Conv3dArgs& a, cudaStream_t st) {
  const int smem = (V_IN_FLOATS + V_W_FLOATS) * (int)sizeof(float);
  AVD_TRY(ensure_dyn_smem(reinterpret_cast<const void*>(conv3d_c64_kernel<MODE>), smem));
  LaunchScope ls(KIND_MISC, st);
  conv3d_c64_kernel<MODE><<<dim3(a.T * a.tiles_h * a.tiles_w, a.B), V_THREADS, smem, st>>>(a);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

static int conv_block(int mode, Conv3dArgs a, const float* gamma, const float* beta, float eps, float* gn_a, float* gn_s,
                      cudaStream_t st) {
  a.tiles_h = (a.H + V_TH - 1) / V_TH;
  a.tiles_w = (a.W + V_TW - 1) / V_TW;
  if (mode == VL_RAW) AVD_TRY(launch_conv3d<VL_RAW>(a, st));
  else if (mode == VL_GN) AVD_TRY(launch_conv3d<VL_GN>(a, st));
  else AVD_TRY(launch_conv3d<VL_TRI>(a, st));
  const int tiles = a.T * a.tiles_h * a.tiles_w;
  LaunchScope ls(KIND_MISC, st);
  gn_finalize_kernel<<<dim3(8, a.B), 256, 0, st>>>(a.partial, tiles, 8.0 * (double)a.T * a.H * a.W, gamma, beta, eps, gn_a, gn_s);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace avd

using namespace avd;

extern "C" {

int64_t avd_vae_scratch_bytes(int B, int T, int H, int W) {
  const int64_t vol = (int64_t)T * H * W;
  const int64_t tiles = (int64_t)T * ((H + V_TH - 1) / V_TH) * ((W + V_TW - 1) / V_TW);
  // two trunk activations [B,64,T,H,W], GroupNorm partials, 2 x per-channel affines, latent-resolution hidden (decode)
  return 2 * B * V_C * vol * 4 + B * tiles * 16 * 4 + 4 * B * V_C * 4 + B * V_C * vol * 4 / 16 + 4096;
}

int avd_video_encode(const avd_vae_weights* w, const float* x, float* z, void* scratch, int B, int T, int H, int W,
                     void* stream) {
  AVD_REQUIRE(w != nullptr && x != nullptr && z != nullptr && scratch != nullptr, "null argument");
  AVD_REQUIRE(w->base == V_C, "the VideoVAE kernels are built for 64 trunk channels (video.encoder.base / decoder.base)");
  AVD_REQUIRE(w->enc_blocks >= 1 && w->enc_blocks <= AVD_VAE_MAX_BLOCKS, "bad encoder block count");
  AVD_REQUIRE(B > 0 && B <= 65535 && T > 0 && H > 0 && W > 0, "bad shape");
  AVD_REQUIRE(w->t_down > 0 && w->s_down > 0 && T % w->t_down == 0 && H % w->s_down == 0 && W % w->s_down == 0,
              "T, H, W must be multiples of (t_down, s_down, s_down): crop first (vae_video3d.py:139-162)");
  AVD_REQUIRE(w->lat_ch > 0 && w->lat_ch <= 256 && w->in_ch > 0, "bad channel counts");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t vol = (int64_t)T * H * W;
  const int64_t tiles = (int64_t)T * ((H + V_TH - 1) / V_TH) * ((W + V_TW - 1) / V_TW);
  float* y[2] = {static_cast<float*>(scratch), static_cast<float*>(scratch) + (int64_t)B * V_C * vol};
  float* partial = y[1] + (int64_t)B * V_C * vol;
  float* gn_a = partial + B * tiles * 16;
  float* gn_s = gn_a + B * V_C;
  const float* cur = x;
  for (int i = 0; i < w->enc_blocks; ++i) {
    Conv3dArgs a{};
    a.B = B; a.Cin = i == 0 ? w->in_ch : V_C; a.T = T; a.H = H; a.W = W;
    a.in = cur; a.gn_a = gn_a; a.gn_s = gn_s;
    a.w = w->enc_w[i]; a.bias = w->enc_b[i]; a.out = y[i & 1]; a.partial = partial;
    // the affine of block i-1 is consumed by this launch's loader and then overwritten by this block's finalize
    AVD_TRY(conv_block(i == 0 ? VL_RAW : VL_GN, a, w->enc_gn_w[i], w->enc_gn_b[i], w->gn_eps, gn_a, gn_s, st));
    cur = y[i & 1];
  }
  const int Tl = T / w->t_down, Hl = H / w->s_down, Wl = W / w->s_down;
  LaunchScope ls(KIND_MISC, st);
  vae_pool_lat_kernel<<<dim3(Tl * Hl * Wl, B), 256, 0, st>>>(cur, gn_a, gn_s, w->to_lat_w, w->to_lat_b, z, T, H, W, w->t_down,
                                                            w->s_down, w->lat_ch);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

int avd_video_decode(const avd_vae_weights* w, const float* z, float* x, void* scratch, int B, int Tl, int Hl, int Wl, int T,
                     int H, int W, void* stream) {
  AVD_REQUIRE(w != nullptr && x != nullptr && z != nullptr && scratch != nullptr, "null argument");
  AVD_REQUIRE(w->base == V_C, "the VideoVAE kernels are built for 64 trunk channels (video.encoder.base / decoder.base)");
  AVD_REQUIRE(w->dec_blocks >= 1 && w->dec_blocks <= AVD_VAE_MAX_BLOCKS, "bad decoder block count");
  AVD_REQUIRE(B > 0 && B <= 65535 && T > 0 && H > 0 && W > 0 && Tl > 0 && Hl > 0 && Wl > 0, "bad shape");
  AVD_REQUIRE(w->in_ch > 0 && w->in_ch <= 4 && w->lat_ch > 0, "bad channel counts");
  AVD_REQUIRE((int64_t)Tl * Hl * Wl * 16 <= (int64_t)T * H * W, "latent grid too large for the scratch carving");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t vol = (int64_t)T * H * W, lvol = (int64_t)Tl * Hl * Wl;
  const int64_t tiles = (int64_t)T * ((H + V_TH - 1) / V_TH) * ((W + V_TW - 1) / V_TW);
  float* y[2] = {static_cast<float*>(scratch), static_cast<float*>(scratch) + (int64_t)B * V_C * vol};
  float* partial = y[1] + (int64_t)B * V_C * vol;
  float* gn_a = partial + B * tiles * 16;
  float* gn_s = gn_a + B * V_C;
  float* hlat = gn_s + 3 * B * V_C;
  {
    LaunchScope ls(KIND_MISC, st);
    vae_from_lat_kernel<<<dim3((unsigned)std::min<int64_t>((lvol * V_C + 255) / 256, 4096), B), 256, 0, st>>>(
        z, w->from_lat_w, w->from_lat_b, hlat, w->lat_ch, lvol);
    AVD_CUDA(cudaGetLastError());
  }
  const float* cur = hlat;
  for (int i = 0; i < w->dec_blocks; ++i) {
    Conv3dArgs a{};
    a.B = B; a.Cin = V_C; a.T = T; a.H = H; a.W = W;
    a.in = cur; a.gn_a = gn_a; a.gn_s = gn_s;
    a.Tl = Tl; a.Hl = Hl; a.Wl = Wl;
    a.st = (float)Tl / (float)T; a.sh = (float)Hl / (float)H; a.sw = (float)Wl / (float)W;
    a.w = w->dec_w[i]; a.bias = w->dec_b[i]; a.out = y[i & 1]; a.partial = partial;
    AVD_TRY(conv_block(i == 0 ? VL_TRI : VL_GN, a, w->dec_gn_w[i], w->dec_gn_b[i], w->gn_eps, gn_a, gn_s, st));
    cur = y[i & 1];
  }
  LaunchScope ls(KIND_MISC, st);
  vae_to_img_kernel<<<dim3((unsigned)std::min<int64_t>((vol + 255) / 256, 8192), B), 256, 0, st>>>(
      cur, gn_a, gn_s, w->to_img_w, w->to_img_b, x, w->in_ch, vol, w->out_tanh);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
