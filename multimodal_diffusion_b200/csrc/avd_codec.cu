// avd_codec.cu -- SURVEY §8f N1: AudioCodec.encode / decode (avdiff/models/encoders/audio_codec.py:184-214) as CUDA kernels:
// the step either side of the sampling loop for the audio modality.
//
//   encode   wav [B,1,L] -> pre: Conv1d(1->H,k9)+GELU, Conv1d(H->H,k9)+GELU (audio_codec.py:107-110)
//                        -> right-pad / crop to Fa*hop, avg_pool1d(hop) (:149-176) -> to_lat 1x1 (H->Ca)  -> z [B,Ca,Fa]
//            ONE kernel: a CTA owns one latent frame; it rebuilds the first conv's output for its window from the waveform
//            (9 MACs per value: cheaper than a round trip through HBM), runs the HxHx9 contraction on CUDA cores in fp32,
//            applies GELU, reduces the frame mean in a fixed order (deterministic), and applies the 1x1 projection.
//   decode   z [B,Ca,Fa] -> from_lat 1x1 (Ca->H) -> nearest upsample to Fa*hop (:204-206)
//                        -> smooth: Conv1d(H->H,k)+GELU, Conv1d(H->H,k)+GELU, Conv1d(H->1,k) (:117-123) -> tanh -> wav
//            TWO kernels: (A) upsampled from_lat rows built on the fly + first smoothing conv + GELU -> g1 [B,H,L];
//            (B) second smoothing conv + GELU into shared memory (tile + halo) + the H->1 conv + tanh.
//
// The contractions are fp32 FMA (H = 64: 3.5 GFLOP per 3 s clip -- < 0.5 % of a sampling call's FLOPs), parity
// tolerance 1e-5 against the reference modules (tests/golden/encoders.npz).
#include "avd_internal.h"
#include "avd_ptx.cuh"

#include <mutex>
#include <string>

namespace avd {

constexpr int AC_THREADS = 256;
constexpr int AC_TILE = 256;        // time steps per sub-tile: thread (lane tg, channel group cog) owns t = tg + 32 i, i < 8
constexpr int AC_MAXH = 64;         // hidden channels (8 per thread: cog < H / 8)

// acc[c][i] += sum_ci sum_j w[(co0 + c)][ci][j] * in[ci][tg + 32 i + j]   for the thread's 8 output channels and 8 times.
// in: smem [H][in_ld] (tile + K - 1 halo); w: global [H][H][K] (read through the read-only path: every lane of a warp reads
// the same address, the 8 channel groups of a CTA read different rows).
template <int K>
__device__ __forceinline__ void conv_tile_accumulate(float (&acc)[8][8], const float* __restrict__ in, int in_ld, int H,
                                                     const float* __restrict__ w, int co0, int tg) {
  for (int ci = 0; ci < H; ++ci) {
    const float* row = in + ci * in_ld + tg;
    float x[8 + 0][1];
    (void)x;
    float wv[8][K];
#pragma unroll
    for (int c = 0; c < 8; ++c)
#pragma unroll
      for (int j = 0; j < K; ++j) wv[c][j] = __ldg(w + ((int64_t)(co0 + c) * H + ci) * K + j);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float v[K];
#pragma unroll
      for (int j = 0; j < K; ++j) v[j] = row[32 * i + j];
#pragma unroll
      for (int c = 0; c < 8; ++c)
#pragma unroll
        for (int j = 0; j < K; ++j) acc[c][i] = fmaf(wv[c][j], v[j], acc[c][i]);
    }
  }
}

struct CodecEncArgs {
  int B, L, H, Ca, Fa, hop, total;     // total = Fa * hop (the padded / cropped length the pool sees)
  const float* wav;                    // [B,1,L]
  const float* w1; const float* b1;    // pre[0][0]: [H,1,9]
  const float* w2; const float* b2;    // pre[1][0]: [H,H,9]
  const float* wl; const float* bl;    // to_lat:    [Ca,H,1]
  float* z;                            // [B,Ca,Fa]
};

__global__ void __launch_bounds__(AC_THREADS) codec_encode_kernel(CodecEncArgs a) {
  constexpr int K = 9, PAD = 4;
  extern __shared__ float ac_smem[];
  const int in_ld = AC_TILE + K - 1;                 // 264
  float* h1 = ac_smem;                               // [H][in_ld]
  float* wav_s = h1 + a.H * in_ld;                   // [AC_TILE + 2 (K - 1)]
  float* mean_s = wav_s + AC_TILE + 2 * (K - 1);     // [H]
  const int f = blockIdx.x, b = blockIdx.y;
  const int tid = threadIdx.x, tg = tid & 31, cog = tid >> 5;
  const int co0 = cog * 8;
  const bool live = co0 < a.H;
  const float* wav = a.wav + (int64_t)b * a.L;
  float fsum[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int s0 = 0; s0 < a.hop; s0 += AC_TILE) {
    const int t0 = f * a.hop + s0;                   // first output position of this sub-tile
    const int nt = min(AC_TILE, a.hop - s0);
    // waveform window [t0 - 8, t0 + nt + 8): zero outside [0, L) (Conv1d zero padding; F.pad zeros beyond L)
    for (int i = tid; i < nt + 2 * (K - 1); i += AC_THREADS) {
      const int t = t0 - 2 * PAD + i;
      wav_s[i] = (t >= 0 && t < a.L) ? wav[t] : 0.0f;
    }
    __syncthreads();
    // first conv + GELU for positions [t0 - 4, t0 + nt + 4); positions outside [0, L) are the SECOND conv's zero padding
    for (int idx = tid; idx < a.H * (nt + K - 1); idx += AC_THREADS) {
      const int c = idx / (nt + K - 1), i = idx - c * (nt + K - 1);
      const int t = t0 - PAD + i;
      float v = 0.0f;
      if (t >= 0 && t < a.L) {
        float acc = __ldg(a.b1 + c);
#pragma unroll
        for (int j = 0; j < K; ++j) acc = fmaf(__ldg(a.w1 + c * K + j), wav_s[i + j], acc);
        v = gelu_erf(acc);
      }
      h1[c * in_ld + i] = v;
    }
    __syncthreads();
    if (live) {
      float acc[8][8];
#pragma unroll
      for (int c = 0; c < 8; ++c)
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[c][i] = 0.0f;
      conv_tile_accumulate<K>(acc, h1, in_ld, a.H, a.w2, co0, tg);
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const float bias = __ldg(a.b2 + co0 + c);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int tl = tg + 32 * i;
          // positions past the waveform (t >= L) are F.pad zeros of the pooled tensor (audio_codec.py:170-173)
          if (tl < nt && t0 + tl < a.L) fsum[c] += gelu_erf(acc[c][i] + bias);
        }
      }
    }
    __syncthreads();
  }
  if (live) {
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      float v = fsum[c];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (tg == 0) mean_s[co0 + c] = v / (float)a.hop;
    }
  }
  __syncthreads();
  if (tid < a.Ca) {                                   // to_lat: 1x1 conv over the H pooled channels
    float acc = __ldg(a.bl + tid);
    for (int c = 0; c < a.H; ++c) acc = fmaf(__ldg(a.wl + tid * a.H + c), mean_s[c], acc);
    a.z[((int64_t)b * a.Ca + tid) * a.Fa + f] = acc;
  }
}

struct CodecDecArgs {
  int B, Ca, Fa, H, hop, L, K;         // L = Fa * hop
  const float* z;                      // [B,Ca,Fa]
  const float* wf; const float* bf;    // from_lat: [H,Ca,1]
  const float* w0; const float* b0;    // smooth[0]: [H,H,K]
  const float* w2; const float* b2;    // smooth[2]: [H,H,K]
  const float* w4; const float* b4;    // smooth[4]: [1,H,K]
  float* g1;                           // scratch [B,H,L]
  float* wav;                          // [B,1,L]
};

// (A) g1 = gelu(conv(upsample(from_lat(z))))
template <int K>
__global__ void __launch_bounds__(AC_THREADS) codec_decode_a_kernel(CodecDecArgs a) {
  constexpr int PAD = K / 2;
  extern __shared__ float ac_smem[];
  const int in_ld = AC_TILE + K - 1;
  float* up = ac_smem;                               // [H][in_ld]
  const int t0 = blockIdx.x * AC_TILE, b = blockIdx.y;
  const int tid = threadIdx.x, tg = tid & 31, cog = tid >> 5, co0 = cog * 8;
  const int nt = min(AC_TILE, a.L - t0);
  // F.interpolate(mode="nearest", size=L): src = min(floor(dst * (float)Fa / L), Fa - 1) in fp32, exactly as ATen does
  const float scale = (float)a.Fa / (float)a.L;
  for (int idx = tid; idx < a.H * (nt + K - 1); idx += AC_THREADS) {
    const int c = idx / (nt + K - 1), i = idx - c * (nt + K - 1);
    const int t = t0 - PAD + i;
    float v = 0.0f;
    if (t >= 0 && t < a.L) {
      // ATen nearest_idx: identity / >> 1 shortcuts, else min(floor(dst * scale), in - 1) with scale in fp32
      const int fr = a.hop == 1 ? t : a.hop == 2 ? (t >> 1) : min((int)floorf(__fmul_rn((float)t, scale)), a.Fa - 1);
      float acc = __ldg(a.bf + c);
      for (int q = 0; q < a.Ca; ++q) acc = fmaf(__ldg(a.wf + c * a.Ca + q), __ldg(a.z + ((int64_t)b * a.Ca + q) * a.Fa + fr), acc);
      v = acc;
    }
    up[c * in_ld + i] = v;
  }
  __syncthreads();
  if (co0 >= a.H) return;
  float acc[8][8];
#pragma unroll
  for (int c = 0; c < 8; ++c)
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[c][i] = 0.0f;
  conv_tile_accumulate<K>(acc, up, in_ld, a.H, a.w0, co0, tg);
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const float bias = __ldg(a.b0 + co0 + c);
    float* dst = a.g1 + ((int64_t)b * a.H + co0 + c) * a.L + t0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int tl = tg + 32 * i;
      if (tl < nt) dst[tl] = gelu_erf(acc[c][i] + bias);
    }
  }
}

// (B) wav = tanh(conv_{H->1}(gelu(conv(g1))))   -- the middle conv is evaluated on the tile plus the last conv's halo
template <int K>
__global__ void __launch_bounds__(AC_THREADS) codec_decode_b_kernel(CodecDecArgs a) {
  constexpr int PAD = K / 2;
  constexpr int OUT_TILE = AC_TILE - (K - 1);        // output positions per CTA: the middle conv covers OUT_TILE + K - 1 = 256
  extern __shared__ float ac_smem[];
  const int in_ld = AC_TILE + K - 1;
  float* in = ac_smem;                               // [H][in_ld]   g1 window
  float* mid = in + a.H * in_ld;                     // [H][AC_TILE] gelu(conv(g1)) on [o0 - PAD, o0 - PAD + 256)
  const int o0 = blockIdx.x * OUT_TILE, b = blockIdx.y;
  const int tid = threadIdx.x, tg = tid & 31, cog = tid >> 5, co0 = cog * 8;
  const int m0 = o0 - PAD;                           // first position of the middle tile
  for (int idx = tid; idx < a.H * in_ld; idx += AC_THREADS) {
    const int c = idx / in_ld, i = idx - c * in_ld;
    const int t = m0 - PAD + i;
    in[idx] = (t >= 0 && t < a.L) ? __ldg(a.g1 + ((int64_t)b * a.H + c) * a.L + t) : 0.0f;
  }
  __syncthreads();
  if (co0 < a.H) {
    float acc[8][8];
#pragma unroll
    for (int c = 0; c < 8; ++c)
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[c][i] = 0.0f;
    conv_tile_accumulate<K>(acc, in, in_ld, a.H, a.w2, co0, tg);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const float bias = __ldg(a.b2 + co0 + c);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int tl = tg + 32 * i, t = m0 + tl;
        mid[(co0 + c) * AC_TILE + tl] = (t >= 0 && t < a.L) ? gelu_erf(acc[c][i] + bias) : 0.0f;   // zero padding of the last conv
      }
    }
  }
  __syncthreads();
  for (int tl = tid; tl < OUT_TILE; tl += AC_THREADS) {
    const int t = o0 + tl;
    if (t >= a.L) break;
    float acc = __ldg(a.b4);
    for (int c = 0; c < a.H; ++c)
#pragma unroll
      for (int j = 0; j < K; ++j) acc = fmaf(__ldg(a.w4 + c * K + j), mid[c * AC_TILE + tl + j], acc);
    a.wav[(int64_t)b * a.L + t] = tanhf(acc);
  }
}

static int check_hidden(int H) {
  AVD_REQUIRE(H > 0 && H <= AC_MAXH && H % 8 == 0, "codec hidden width must be a multiple of 8, at most 64");
  return 0;
}

}  // namespace avd

using namespace avd;

extern "C" {

int avd_audio_encode(const float* wav, float* z, int B, int L, int H, int Ca, int Fa, int hop, const float* w_pre0,
                     const float* b_pre0, const float* w_pre1, const float* b_pre1, const float* w_lat,
                     const float* b_lat, void* stream) {
  AVD_REQUIRE(wav != nullptr && z != nullptr, "null tensor");
  AVD_REQUIRE(B > 0 && L > 0 && Fa > 0 && hop > 0 && Ca > 0 && Ca <= AC_THREADS, "bad shape");
  AVD_REQUIRE(B <= 65535, "batch too large for one launch");
  AVD_TRY(check_hidden(H));
  CodecEncArgs a{B, L, H, Ca, Fa, hop, Fa * hop, wav, w_pre0, b_pre0, w_pre1, b_pre1, w_lat, b_lat, z};
  const int smem = (H * (AC_TILE + 8) + AC_TILE + 16 + H) * (int)sizeof(float);
  AVD_TRY(ensure_dyn_smem(reinterpret_cast<const void*>(codec_encode_kernel), smem));
  LaunchScope ls(KIND_MISC, static_cast<cudaStream_t>(stream));
  codec_encode_kernel<<<dim3(Fa, B), AC_THREADS, smem, static_cast<cudaStream_t>(stream)>>>(a);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

int64_t avd_audio_decode_scratch_bytes(int B, int H, int Fa, int hop) { return (int64_t)B * H * Fa * hop * 4; }

int avd_audio_decode(const float* z, float* wav, float* scratch, int B, int Ca, int Fa, int H, int hop, int K,
                     const float* w_from, const float* b_from, const float* w_s0, const float* b_s0, const float* w_s2,
                     const float* b_s2, const float* w_s4, const float* b_s4, void* stream) {
  AVD_REQUIRE(z != nullptr && wav != nullptr && scratch != nullptr, "null tensor");
  AVD_REQUIRE(B > 0 && Fa > 0 && hop > 0 && Ca > 0, "bad shape");
  AVD_REQUIRE(B <= 65535, "batch too large for one launch");
  AVD_REQUIRE(K == 3 || K == 5 || K == 7 || K == 9 || K == 11, "smoothing kernel size must be 3, 5, 7, 9 or 11");
  AVD_TRY(check_hidden(H));
  const int L = Fa * hop;
  CodecDecArgs a{B, Ca, Fa, H, hop, L, K, z, w_from, b_from, w_s0, b_s0, w_s2, b_s2, w_s4, b_s4, scratch, wav};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int smem_a = H * (AC_TILE + K - 1) * (int)sizeof(float);
  const int smem_b = (H * (AC_TILE + K - 1) + H * AC_TILE) * (int)sizeof(float);
  const int out_tile = AC_TILE - (K - 1);
  const dim3 ga((L + AC_TILE - 1) / AC_TILE, B), gb((L + out_tile - 1) / out_tile, B);
#define AVD_DEC(KK)                                                                                   \
  case KK: {                                                                                          \
    AVD_TRY(ensure_dyn_smem(reinterpret_cast<const void*>(codec_decode_a_kernel<KK>), smem_a));       \
    AVD_TRY(ensure_dyn_smem(reinterpret_cast<const void*>(codec_decode_b_kernel<KK>), smem_b));       \
    { LaunchScope ls(KIND_MISC, st); codec_decode_a_kernel<KK><<<ga, AC_THREADS, smem_a, st>>>(a); }  \
    { LaunchScope ls(KIND_MISC, st); codec_decode_b_kernel<KK><<<gb, AC_THREADS, smem_b, st>>>(a); }  \
    break;                                                                                            \
  }
  switch (K) {
    AVD_DEC(3) AVD_DEC(5) AVD_DEC(7) AVD_DEC(9) AVD_DEC(11)
    default: break;
  }
#undef AVD_DEC
  AVD_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"

// ---- allocation-free variant of the audio decoder -----------------------------------------------------------------
// The one intermediate of avd_audio_decode (the [B,H,Fa*hop] fp32 activation between its two kernels) lives in a
// file-scope device array instead of caller memory.  Every GPU that loads the library holds these 16 MiB until exit.
// In HOST code the name below is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress,
// looked up once per device (with that device current) and cached.
static __device__ __align__(256) float g_audio_decode_static_scratch[AVD_AUDIO_DECODE_STATIC_MAX_BYTES / 4];

static int audio_decode_static_ptr(float** out) {
  constexpr int MAX_DEV = 64;
  static std::mutex mu;
  static float* ptrs[MAX_DEV] = {};
  int dev = 0;
  AVD_CUDA(cudaGetDevice(&dev));
  AVD_REQUIRE(dev >= 0 && dev < MAX_DEV, "device ordinal beyond the static-scratch table");
  std::lock_guard<std::mutex> lk(mu);
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    AVD_CUDA(cudaGetSymbolAddress(&p, g_audio_decode_static_scratch));
    AVD_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null scratch address");
    ptrs[dev] = static_cast<float*>(p);
  }
  *out = ptrs[dev];
  return 0;
}

extern "C" int avd_audio_decode_static(const float* z, float* wav, int B, int Ca, int Fa, int H, int hop, int K,
                                       const float* w_from, const float* b_from, const float* w_s0, const float* b_s0,
                                       const float* w_s2, const float* b_s2, const float* w_s4, const float* b_s4,
                                       void* stream) {
  AVD_REQUIRE(B > 0 && Fa > 0 && hop > 0 && H > 0, "bad shape");
  const int64_t need = (int64_t)B * H * Fa * hop * 4;
  if (need > (int64_t)AVD_AUDIO_DECODE_STATIC_MAX_BYTES)
    AVD_FAIL("avd_audio_decode_static: B*H*Fa*hop*4 = " + std::to_string(need) + " bytes of scratch exceed the built-in " +
             std::to_string((int64_t)AVD_AUDIO_DECODE_STATIC_MAX_BYTES) + " (use avd_audio_decode with caller scratch)");
  float* scratch = nullptr;
  AVD_TRY(audio_decode_static_ptr(&scratch));
  // kernel A writes every element of [B,H,L] that kernel B reads, so nothing in the shared array has to start at zero
  return avd_audio_decode(z, wav, scratch, B, Ca, Fa, H, hop, K, w_from, b_from, w_s0, b_s0, w_s2, b_s2, w_s4, b_s4,
                          stream);
}
