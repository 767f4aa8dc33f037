// avd_stream.cu -- SURVEY.md §8(f) row N2: the stitching kernels of the sliding-window driver
// (reference: avdiff/models/infer/stream_infer.py crossfade_audio :82-113, crossfade_video :116-143).
//
// The reference scatters: for window i = 0..N-1 it adds chunk_i * w into out[i*hop : i*hop+L] and w into a norm buffer,
// then divides.  Here both are one GATHER pass: an output sample at position t looks up the (at most ceil(L/hop)) windows
// that cover it, in ascending i -- the order the reference's loop adds them -- with explicitly rounded fp32 multiplies,
// adds and the final divide, so the result is bit-identical to numpy's fp32 arithmetic and no norm buffer exists.
// HBM-bound: every window sample is read once, every output sample written once (audio 4 B + 4 B per window sample /
// output sample; video 1 B + 1 B).  The fade window w[L] comes from the host (numpy, the reference's formula), so the
// transcendental part is bit-exact by construction.
#include "avd_internal.h"

namespace avd {

// ---------------------------------------------------------------------------------------------------- audio
// chunks [rows, N, L] fp32, w [L] fp32, out [rows, (N-1)*hop + L] fp32; blockIdx.y = row.  rows = 1 and floor = 1e-6 is
// stream_infer.crossfade_audio; rows = B*C, w = hann and floor = 1e-8 is ops.overlap_add_1d(apply_hann=True)
// (avdiff/utils/ops.py:82-92), the same arithmetic in torch.
template <int VEC>
__global__ void __launch_bounds__(256) crossfade_audio_kernel(const float* __restrict__ chunks,
                                                              const float* __restrict__ w, float* __restrict__ out,
                                                              int N, int L, int hop, int64_t T_out, float floor_) {
  chunks += (int64_t)blockIdx.y * N * L;
  out += (int64_t)blockIdx.y * T_out;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * VEC;
  for (int64_t t0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * VEC; t0 < T_out; t0 += stride) {
    // windows covering t0 .. t0+VEC-1 (VEC > 1 only when hop % VEC == 0 and L % VEC == 0: the VEC samples then share
    // their set of windows and every access is VEC-aligned)
    int64_t i_lo = t0 >= L ? (t0 - L) / hop + 1 : 0;
    int64_t i_hi = t0 / hop;
    if (i_hi > N - 1) i_hi = N - 1;
    float acc[VEC], nrm[VEC];
#pragma unroll
    for (int v = 0; v < VEC; ++v) acc[v] = nrm[v] = 0.0f;
    for (int64_t i = i_lo; i <= i_hi; ++i) {
      const int64_t j = t0 - i * hop;            // position inside window i
      float c[VEC], ww[VEC];
      if (VEC == 4) {
        const float4 c4 = *reinterpret_cast<const float4*>(chunks + i * L + j);
        const float4 w4 = *reinterpret_cast<const float4*>(w + j);
        c[0] = c4.x; c[1 % VEC] = c4.y; c[2 % VEC] = c4.z; c[3 % VEC] = c4.w;
        ww[0] = w4.x; ww[1 % VEC] = w4.y; ww[2 % VEC] = w4.z; ww[3 % VEC] = w4.w;
      } else {
        c[0] = chunks[i * L + j];
        ww[0] = w[j];
      }
#pragma unroll
      for (int v = 0; v < VEC; ++v) {
        acc[v] = __fadd_rn(acc[v], __fmul_rn(c[v], ww[v]));   // y[a:a+L] += chunks[i] * w   (stream_infer.py:109)
        nrm[v] = __fadd_rn(nrm[v], ww[v]);                    // norm[a:a+L] += w            (:110)
      }
    }
    float r[VEC];
#pragma unroll
    for (int v = 0; v < VEC; ++v) r[v] = __fdiv_rn(acc[v], fmaxf(nrm[v], floor_));   // (:111-112)
    if (VEC == 4) *reinterpret_cast<float4*>(out + t0) = make_float4(r[0], r[1 % VEC], r[2 % VEC], r[3 % VEC]);
    else out[t0] = r[0];
  }
}

int launch_overlap_add_win(const float* chunks, const float* w, float* out, int rows, int N, int L, int hop,
                           float norm_floor, cudaStream_t s) {
  AVD_REQUIRE(chunks != nullptr && w != nullptr && out != nullptr, "null pointer");
  AVD_REQUIRE(rows > 0 && rows <= 65535 && N > 0 && L > 0 && hop > 0, "windowed overlap-add needs 0 < rows <= 65535 and N, L, hop > 0");
  const int64_t T_out = (int64_t)(N - 1) * hop + L;
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const bool vec4 = (hop % 4 == 0) && (L % 4 == 0) && (T_out % 4 == 0) && ((reinterpret_cast<uintptr_t>(chunks) | reinterpret_cast<uintptr_t>(w) |
                                                        reinterpret_cast<uintptr_t>(out)) % 16 == 0);
  const int64_t items = vec4 ? T_out / 4 : T_out;
  int64_t blocks = (items + 255) / 256;
  const int64_t cap = ((int64_t)sms * 16 + rows - 1) / rows;   // grid-stride beyond 16 resident blocks per SM
  if (blocks > cap) blocks = cap;
  if (kind_skipped(KIND_TOKENIZE)) return 0;
  LaunchScope ls(KIND_TOKENIZE, s);
  const dim3 grid((unsigned)blocks, (unsigned)rows);
  if (vec4) crossfade_audio_kernel<4><<<grid, 256, 0, s>>>(chunks, w, out, N, L, hop, T_out, norm_floor);
  else crossfade_audio_kernel<1><<<grid, 256, 0, s>>>(chunks, w, out, N, L, hop, T_out, norm_floor);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

int launch_crossfade_audio(const float* chunks, const float* w, float* out, int N, int L, int hop, cudaStream_t s) {
  return launch_overlap_add_win(chunks, w, out, 1, N, L, hop, 1e-6f, s);
}

// ---------------------------------------------------------------------------------------------------- video
// chunks [N, T, F] uint8 (F = H*W*3 bytes per frame), w [T] fp32 (one weight per frame), out [(N-1)*hop + T, F] uint8.
// A thread owns 16 consecutive bytes of one output frame; all of them share the frame's windows and weights.
// blockDim.x must be 256 (the block fills the 256-entry u8 -> u8/255 table).
template <int VEC>   // 16: uint4 accesses; 1: byte accesses (frame size or pointers not 16-byte aligned)
__global__ void __launch_bounds__(256) crossfade_video_kernel(const uint8_t* __restrict__ chunks,
                                                              const float* __restrict__ w, uint8_t* __restrict__ out,
                                                              int N, int T, int64_t F, int hop, int64_t T_out) {
  // u8 / 255.0 has 256 possible values: one correctly rounded division each, once per block, instead of one per byte
  __shared__ float lut[256];
  lut[threadIdx.x] = __fdiv_rn((float)threadIdx.x, 255.0f);
  __syncthreads();
  const int64_t per_frame = F / VEC;
  const int64_t total = T_out * per_frame;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
    const int64_t t = idx / per_frame;
    const int64_t x = (idx - t * per_frame) * VEC;
    int64_t i_lo = t >= T ? (t - T) / hop + 1 : 0;
    int64_t i_hi = t / hop;
    if (i_hi > N - 1) i_hi = N - 1;
    float acc[VEC];
#pragma unroll
    for (int v = 0; v < VEC; ++v) acc[v] = 0.0f;
    float nrm = 0.0f;
    for (int64_t i = i_lo; i <= i_hi; ++i) {
      const int64_t j = t - i * hop;
      const float wj = w[j];
      const uint8_t* src = chunks + (i * T + j) * F + x;
      uint8_t px[VEC];
      if (VEC == 16) *reinterpret_cast<uint4*>(px) = *reinterpret_cast<const uint4*>(src);
      else px[0] = src[0];
#pragma unroll
      for (int v = 0; v < VEC; ++v)   // chunk = u8 -> f32 / 255.0 ; out += chunk * w     (stream_infer.py:137-139)
        acc[v] = __fadd_rn(acc[v], __fmul_rn(lut[px[v]], wj));
      nrm = __fadd_rn(nrm, wj);       // norm += w                                         (:140)
    }
    const float den = fmaxf(nrm, 1e-6f);
    uint8_t res[VEC];
#pragma unroll
    for (int v = 0; v < VEC; ++v) {   // clip(out / norm, 0, 1) * 255 -> uint8 (truncation)  (:142-144)
      float q = den == 1.0f ? acc[v] : __fdiv_rn(acc[v], den);   // x / 1.0 == x exactly
      q = fminf(fmaxf(q, 0.0f), 1.0f);
      res[v] = (uint8_t)__fmul_rn(q, 255.0f);
    }
    uint8_t* dst = out + t * F + x;
    if (VEC == 16) *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(res);
    else dst[0] = res[0];
  }
}

int launch_crossfade_video(const uint8_t* chunks, const float* w, uint8_t* out, int N, int T, int64_t F, int hop,
                           cudaStream_t s) {
  AVD_REQUIRE(chunks != nullptr && w != nullptr && out != nullptr, "null pointer");
  AVD_REQUIRE(N > 0 && T > 0 && F > 0 && hop > 0, "crossfade_video needs N, T, frame bytes, hop > 0");
  const int64_t T_out = (int64_t)(N - 1) * hop + T;
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const bool vec = (F % 16 == 0) &&
                   ((reinterpret_cast<uintptr_t>(chunks) | reinterpret_cast<uintptr_t>(out)) % 16 == 0);
  const int64_t items = T_out * (vec ? F / 16 : F);
  int64_t blocks = (items + 255) / 256;
  const int64_t cap = (int64_t)sms * 16;
  if (blocks > cap) blocks = cap;
  if (kind_skipped(KIND_TOKENIZE)) return 0;
  LaunchScope ls(KIND_TOKENIZE, s);
  if (vec) crossfade_video_kernel<16><<<(unsigned)blocks, 256, 0, s>>>(chunks, w, out, N, T, F, hop, T_out);
  else crossfade_video_kernel<1><<<(unsigned)blocks, 256, 0, s>>>(chunks, w, out, N, T, F, hop, T_out);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace avd
