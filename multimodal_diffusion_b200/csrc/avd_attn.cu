// avd_attn.cu -- K5: joint video+audio self-attention (nn.MultiheadAttention core, mmdt.py:58-61).
//
// attn_tc_kernel (bf16 mode): flash-style, one CTA = one (sequence, head, 128-query tile), 128 threads,
// thread t owns query row t (TMEM lane t).
//   per 128-key block j:   S = Q K_j^T        tcgen05.mma  M=128, N=round16(keys), K=64   -> TMEM cols [0,128)
//                          online softmax     tcgen05.ld rows of S, fp32 max / exp2 / sum in registers,
//                                             P (bf16) written to 128B-swizzled smem
//                          O_j = P V_j        tcgen05.mma  M=128, N=64, K=round16(keys)   -> TMEM cols [128,192)
//                                             (V consumed MN-major straight from the [keys, 64] TMA tile)
//                          O = O*alpha + O_j  in registers
// K/V tiles are double-buffered: the TMA for block j+1 is issued as soon as block j's S-MMA is queued.
// The key count of the last block is rounded to 16 only (instruction-descriptor N / k-steps are runtime
// values), so N = 133 costs 128 + 16 key columns, not 256.  No mask: the reference's sampler passes none,
// and null-branch prompt rows stay in the sequence as keys/values.
//
// attn_f32_kernel (fp32-parity mode): CUDA-core fp32, one warp per query row, scores staged in smem.
#include "avd_internal.h"
#include "avd_ptx.cuh"

#include <cstdlib>

namespace avd {

constexpr int ATT_BQ = 128;
constexpr int ATT_BKV = 128;
constexpr int ATT_HD = 64;
constexpr int ATT_TILE_BYTES = ATT_BQ * ATT_HD * 2;  // 16 KB: [128 rows][64 bf16], 128B swizzle
constexpr int ATT_SMEM = ATT_TILE_BYTES * (1 + 2 + 2 + 2) + 128;  // Q, K[2], V[2], P (2 atoms), barriers (2 CTAs/SM fit)
constexpr uint32_t ATT_TMEM_COLS = 256;  // S: 128 cols, O: 64 cols

struct AttnParams {
  int N;            // tokens per sequence
  int d;            // model width (row stride of out)
  int n_heads;
  int64_t M;        // total rows (n_seq * N)
  __nv_bfloat16* out;
  int out_nseg;
  float scale_log2e;  // (1/sqrt(hd)) * log2(e)
  uint32_t v_lbo, v_sbo;  // MN-major V descriptor strides (bytes)
  int dbg_stop;           // debug: leave the main loop early at a given stage (0 = run normally)
};

__global__ void __launch_bounds__(128) attn_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, const AttnParams p) {
  extern __shared__ __align__(1024) uint8_t att_smem[];
  uint8_t* smem = att_smem;
  if ((smem_u32(smem) & 1023u) != 0) __trap();   // 128B-swizzle tiles need 1024-B aligned bases
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + ATT_TILE_BYTES;       // 2 stages
  uint8_t* sV = sK + 2 * ATT_TILE_BYTES;   // 2 stages
  uint8_t* sP = sV + 2 * ATT_TILE_BYTES;   // 2 swizzle atoms of [128][64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * ATT_TILE_BYTES);
  uint64_t* q_full = bars;
  uint64_t* k_full = bars + 1;   // [2]
  uint64_t* v_full = bars + 3;   // [2]
  uint64_t* s_done = bars + 5;
  uint64_t* o_done = bars + 6;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 7);

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const int q0 = blockIdx.x * ATT_BQ;
  const int head = blockIdx.y;
  const int seq = blockIdx.z;
  const int row_base = seq * p.N;           // first row of this sequence in the [M, 3d] matrix
  const int colQ = head * ATT_HD, colK = p.d + head * ATT_HD, colV = 2 * p.d + head * ATT_HD;
  const int nkv = (p.N + ATT_BKV - 1) / ATT_BKV;

  if (tid == 0) {
    tma_prefetch_desc(&tmQKV);
    for (int i = 0; i < 7; ++i) mbar_init(&bars[i], 1);
    fence_barrier_init();
  }
  __syncwarp();
  if (warp == 0) tmem_alloc<ATT_TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_S = tmem_base;
  const uint32_t tmem_O = tmem_base + 128;
  const uint32_t lane_sel = (uint32_t)(warp * 32) << 16;

  if (tid == 0) {
    mbar_expect_tx(q_full, ATT_TILE_BYTES);
    tma_load_3d(sQ, &tmQKV, q_full, colQ, row_base + q0, 0);
    mbar_expect_tx(&k_full[0], ATT_TILE_BYTES);
    tma_load_3d(sK, &tmQKV, &k_full[0], colK, row_base, 0);
    mbar_expect_tx(&v_full[0], ATT_TILE_BYTES);
    tma_load_3d(sV, &tmQKV, &v_full[0], colV, row_base, 0);
  }

  float o_acc[ATT_HD];
#pragma unroll
  for (int c = 0; c < ATT_HD; ++c) o_acc[c] = 0.0f;
  float m_run = -INFINITY, l_run = 0.0f;

  for (int j = 0; j < nkv && p.dbg_stop != 1; ++j) {
    const int st = j & 1;
    const uint32_t ph = (j >> 1) & 1;
    const int kv_valid = min(ATT_BKV, p.N - j * ATT_BKV);
    const int kv_n = (kv_valid + 15) & ~15;

    if (tid == 0) {
      if (j == 0) mbar_wait(q_full, 0);
      mbar_wait(&k_full[st], ph);
      tc_fence_after();
      const uint32_t idesc_s = make_idesc_bf16(ATT_BQ, kv_n, 0, 0);
      const uint32_t q_addr = smem_u32(sQ), k_addr = smem_u32(sK + st * ATT_TILE_BYTES);
#pragma unroll
      for (int k = 0; k < ATT_HD / 16; ++k) {
        umma_bf16(tmem_S, make_smem_desc_sw128(q_addr + k * 32, 16, 1024),
                  make_smem_desc_sw128(k_addr + k * 32, 16, 1024), idesc_s, k > 0 ? 1u : 0u);
      }
      umma_commit(s_done);
      if (j + 1 < nkv) {  // stage st^1 was last read by block j-1's MMAs, which completed before o_done(j-1)
        const int nst = st ^ 1;
        mbar_expect_tx(&k_full[nst], ATT_TILE_BYTES);
        tma_load_3d(sK + nst * ATT_TILE_BYTES, &tmQKV, &k_full[nst], colK, row_base + (j + 1) * ATT_BKV, 0);
        mbar_expect_tx(&v_full[nst], ATT_TILE_BYTES);
        tma_load_3d(sV + nst * ATT_TILE_BYTES, &tmQKV, &v_full[nst], colV, row_base + (j + 1) * ATT_BKV, 0);
      }
    }
    __syncwarp();
    mbar_wait(s_done, j & 1);
    tc_fence_after();
    if (p.dbg_stop == 2) break;

    // ---- pass 1: row max over the valid keys
    const int nch = (kv_n + 31) >> 5;
    float m_blk = -INFINITY;
    for (int ch = 0; ch < nch; ++ch) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_S + lane_sel + ch * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int c = 0; c < 32; ++c) {
        if (ch * 32 + c < kv_valid) m_blk = fmaxf(m_blk, __uint_as_float(r[c]));
      }
    }
    const float m_new = fmaxf(m_run, m_blk);
    const float alpha = exp2f((m_run - m_new) * p.scale_log2e);   // m_run = -inf on the first block -> 0
    const float m_sc = m_new * p.scale_log2e;

    // ---- pass 2: p = exp2(s*c - m*c), row sum, P -> smem (K-major, 128B swizzle, two 64-key atoms)
    float l_blk = 0.0f;
    for (int ch = 0; ch < nch; ++ch) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_S + lane_sel + ch * 32, r);
      tmem_ld_wait();
      uint32_t w[16];
#pragma unroll
      for (int c = 0; c < 32; c += 2) {
        float p0 = (ch * 32 + c < kv_valid) ? exp2f(fmaf(__uint_as_float(r[c]), p.scale_log2e, -m_sc)) : 0.0f;
        float p1 = (ch * 32 + c + 1 < kv_valid) ? exp2f(fmaf(__uint_as_float(r[c + 1]), p.scale_log2e, -m_sc)) : 0.0f;
        l_blk += p0 + p1;
        w[c >> 1] = pack_bf16x2(p0, p1);
      }
      uint8_t* atom = sP + (ch >> 1) * ATT_TILE_BYTES + tid * 128;
      const int q_base = (ch & 1) * 4;  // 16-byte chunk index of this 32-column group inside the 128-byte row
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int q = (q_base + i) ^ (tid & 7);
        *reinterpret_cast<uint4*>(atom + q * 16) = make_uint4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]);
      }
    }
    l_run = l_run * alpha + l_blk;
    m_run = m_new;

    fence_async_smem();   // P (generic-proxy stores) -> visible to the tensor core's async-proxy reads
    tc_fence_before();
    __syncthreads();
    if (p.dbg_stop == 3) break;

    if (tid == 0) {
      tc_fence_after();
      mbar_wait(&v_full[st], ph);
      tc_fence_after();
      const uint32_t idesc_o = make_idesc_bf16(ATT_BQ, ATT_HD, 0, 1);   // B = V is MN-major
      const uint32_t p_addr = smem_u32(sP), v_addr = smem_u32(sV + st * ATT_TILE_BYTES);
      const int ksteps = kv_n >> 4;
      for (int ks = 0; ks < ksteps; ++ks) {
        const uint32_t pa = p_addr + (ks >> 2) * ATT_TILE_BYTES + (ks & 3) * 32;
        const uint32_t va = v_addr + ks * 16 * 128;   // 16 keys = 16 rows of 128 B
        umma_bf16(tmem_O, make_smem_desc_sw128(pa, 16, 1024), make_smem_desc_sw128(va, p.v_lbo, p.v_sbo), idesc_o,
                  ks > 0 ? 1u : 0u);
      }
      umma_commit(o_done);
    }
    __syncwarp();
    mbar_wait(o_done, j & 1);
    tc_fence_after();

    // ---- O = O * alpha + O_j
#pragma unroll
    for (int ch = 0; ch < ATT_HD / 32; ++ch) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_O + lane_sel + ch * 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int c = 0; c < 32; ++c) o_acc[ch * 32 + c] = fmaf(o_acc[ch * 32 + c], alpha, __uint_as_float(r[c]));
    }
    tc_fence_before();
  }

  if (p.dbg_stop != 0 && tid == 0) {   // debug exits: drain the TMA loads that are still in flight
    mbar_wait(q_full, 0);
    mbar_wait(&k_full[0], 0);
    mbar_wait(&v_full[0], 0);
    if (nkv > 1 && p.dbg_stop >= 2) { mbar_wait(&k_full[1], 0); mbar_wait(&v_full[1], 0); }
  }

  // ---- normalise and store this thread's query row
  const int q = q0 + tid;
  if (q < p.N) {
    const float inv_l = 1.0f / l_run;
    const int64_t seg_stride = p.M * (int64_t)p.d;
    __nv_bfloat16* op = p.out + ((int64_t)row_base + q) * p.d + head * ATT_HD;
    float rem[ATT_HD];
#pragma unroll
    for (int c = 0; c < ATT_HD; ++c) rem[c] = o_acc[c] * inv_l;
    for (int sg = 0; sg < p.out_nseg; ++sg) {
      uint4* o4 = reinterpret_cast<uint4*>(op + sg * seg_stride);
#pragma unroll
      for (int v4 = 0; v4 < ATT_HD / 8; ++v4) {
        uint32_t w[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int c = v4 * 8 + k * 2;
          __nv_bfloat16 h0 = __float2bfloat16_rn(rem[c]);
          __nv_bfloat16 h1 = __float2bfloat16_rn(rem[c + 1]);
          rem[c] -= __bfloat162float(h0);
          rem[c + 1] -= __bfloat162float(h1);
          w[k] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
        }
        o4[v4] = make_uint4(w[0], w[1], w[2], w[3]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc<ATT_TMEM_COLS>(tmem_base);
  }
}

static uint32_t env_u32(const char* name, uint32_t dflt) {
  const char* v = std::getenv(name);
  return v ? (uint32_t)std::strtoul(v, nullptr, 0) : dflt;
}

int launch_attention_tc(const void* qkv_bf16, void* out, int out_nseg, int n_seq, int N, int n_heads, int hd,
                        cudaStream_t s) {
  AVD_REQUIRE(hd == ATT_HD, "tcgen05 attention kernel is built for head_dim 64");
  AVD_REQUIRE(n_seq > 0 && N > 0 && n_heads > 0, "empty attention");
  const int d = n_heads * hd;
  const int64_t M = (int64_t)n_seq * N;
  CUtensorMap tm;
  AVD_TRY(get_tensor_map_bf16_3d(qkv_bf16, (uint64_t)3 * d, (uint64_t)M, 1, (uint64_t)3 * d * 2,
                                 (uint64_t)M * 3 * d * 2, ATT_HD, ATT_BQ, &tm));
  static bool attr_set = false;
  if (!attr_set) {
    AVD_CUDA(cudaFuncSetAttribute(attn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_SMEM));
    attr_set = true;
  }
  AttnParams p{};
  p.N = N; p.d = d; p.n_heads = n_heads; p.M = M;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.out_nseg = out_nseg;
  p.scale_log2e = (1.0f / sqrtf((float)hd)) * 1.4426950408889634f;
  p.v_lbo = env_u32("AVD_DBG_V_LBO", 16);      // debug overrides for the MN-major descriptor
  p.v_sbo = env_u32("AVD_DBG_V_SBO", 1024);
  p.dbg_stop = (int)env_u32("AVD_DBG_ATT_STOP", 0);
  dim3 grid((N + ATT_BQ - 1) / ATT_BQ, n_heads, n_seq);
  LaunchScope ls(KIND_ATTN_TC, s);
  attn_tc_kernel<<<grid, 128, ATT_SMEM, s>>>(tm, p);
  AVD_CUDA(cudaGetLastError());
  return 0;
}

// =====================================================================================================
// fp32 CUDA-core attention (parity mode): one warp per query row.
// =====================================================================================================
constexpr int AF_WARPS = 4;
constexpr int AF_ROWS = 16;   // query rows per CTA

__global__ void __launch_bounds__(AF_WARPS * 32) attn_f32_kernel(const float* __restrict__ qkv,
                                                                 __nv_bfloat16* __restrict__ out, int out_nseg, int N,
                                                                 int d, int hd, int64_t M, float scale) {
  extern __shared__ float af_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* sc = af_smem + (size_t)warp * N;                 // scores of this warp's current row
  float* qs = af_smem + (size_t)AF_WARPS * N + warp * hd; // its query vector
  const int head = blockIdx.y, seq = blockIdx.z;
  const int64_t row0 = (int64_t)seq * N;
  const int ld = 3 * d;
  const float* Kp = qkv + row0 * ld + d + head * hd;
  const float* Vp = qkv + row0 * ld + 2 * d + head * hd;
  const int64_t seg_stride = M * (int64_t)d;

  for (int qi = warp; qi < AF_ROWS; qi += AF_WARPS) {
    const int q = blockIdx.x * AF_ROWS + qi;
    if (q >= N) break;
    const float* Qp = qkv + (row0 + q) * ld + head * hd;
    for (int c = lane; c < hd; c += 32) qs[c] = Qp[c];
    __syncwarp();
    float mx = -INFINITY;
    for (int j = lane; j < N; j += 32) {
      const float* kr = Kp + (int64_t)j * ld;
      float acc = 0.0f;
      for (int c = 0; c < hd; c += 4) {
        float4 k4 = *reinterpret_cast<const float4*>(kr + c);
        acc = fmaf(qs[c], k4.x, acc); acc = fmaf(qs[c + 1], k4.y, acc);
        acc = fmaf(qs[c + 2], k4.z, acc); acc = fmaf(qs[c + 3], k4.w, acc);
      }
      acc *= scale;
      sc[j] = acc;
      mx = fmaxf(mx, acc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.0f;
    for (int j = lane; j < N; j += 32) {
      float e = expf(sc[j] - mx);
      sc[j] = e;
      sum += e;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    __syncwarp();
    const float inv = 1.0f / sum;
    for (int c = lane; c < hd; c += 32) {
      float acc = 0.0f;
      for (int j = 0; j < N; ++j) acc = fmaf(sc[j], Vp[(int64_t)j * ld + c], acc);
      float v = acc * inv;
      int64_t idx = (row0 + q) * d + head * hd + c;
      __nv_bfloat16 h0 = __float2bfloat16_rn(v);
      out[idx] = h0;
      if (out_nseg >= 2) {
        float r1 = v - __bfloat162float(h0);
        __nv_bfloat16 h1 = __float2bfloat16_rn(r1);
        out[seg_stride + idx] = h1;
        if (out_nseg >= 3) out[2 * seg_stride + idx] = __float2bfloat16_rn(r1 - __bfloat162float(h1));
      }
    }
    __syncwarp();
  }
}

int launch_attention_f32(const float* qkv, void* out, int out_nseg, int n_seq, int N, int n_heads, int hd,
                         cudaStream_t s) {
  AVD_REQUIRE(hd % 4 == 0 && hd <= 256, "head_dim must be a multiple of 4, <= 256");
  AVD_REQUIRE(n_seq > 0 && N > 0 && n_heads > 0, "empty attention");
  const int d = n_heads * hd;
  size_t smem = ((size_t)AF_WARPS * N + AF_WARPS * hd) * sizeof(float);
  AVD_REQUIRE(smem <= 200 * 1024, "sequence too long for the fp32 parity attention kernel");
  if (smem > 48 * 1024) {
    AVD_CUDA(cudaFuncSetAttribute(attn_f32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  dim3 grid((N + AF_ROWS - 1) / AF_ROWS, n_heads, n_seq);
  LaunchScope ls(KIND_ATTN_F32, s);
  attn_f32_kernel<<<grid, AF_WARPS * 32, smem, s>>>(qkv, static_cast<__nv_bfloat16*>(out), out_nseg, N, d, hd,
                                                    (int64_t)n_seq * N, 1.0f / sqrtf((float)hd));
  AVD_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace avd
