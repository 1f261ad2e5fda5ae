// Global (un-windowed) multi-head attention that also RETURNS the attention map -- the core of the Depthformer decoder's
// SelfAttentionBlock (model/Depthformer/self_attention.py:70-80):
//
//     attn = softmax(q k^T * scale)   (B, nH, S, S), returned to the caller (fp32)
//     o    = attn v                   (B, S, nH * dv)
//
// S = (H/32)(W/32) tokens: 418 at 352x1216, 300 at 480x640; head dims 64.  The map has to be materialised, so the op is
// bound by writing it (B nH S^2 4 B: 22 MB at batch 8) -- the q / k / v reads are 50x smaller.
//
// CUDA-core kernel, fp32 math (this is also the fp32 mode: <= 1e-5 against the reference rules out bf16 / tf32 MMAs):
// one CTA = 32 query rows of one (image, head); its 32 x S block of logits lives in shared memory, so the softmax is
// exact and single-pass over smem, each map row is written ONCE by a warp (coalesced 128-byte segments), and K / V
// stream through a 64-key shared-memory tile.  Inputs are (B, S, nH * d) rows with an arbitrary row stride (slices of a
// packed projection work); fp32 or bf16.
#include "swin_common.cuh"

namespace swin {
namespace mha {

constexpr int kRows = 32;      // query rows per CTA
constexpr int kKeys = 64;      // keys per shared-memory tile
constexpr int kThreads = 256;
constexpr int kMaxHeadDim = 64;
constexpr int kTileStride = kMaxHeadDim + 4;   // words; 68 = 4 mod 32: 128-bit reads of consecutive rows are conflict-free

template <typename T>
__device__ __forceinline__ float ldf(const T* p);
template <>
__device__ __forceinline__ float ldf<float>(const float* p) { return __ldg(p); }
template <>
__device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }

template <typename T>
__device__ __forceinline__ void stf(T* p, float v);
template <>
__device__ __forceinline__ void stf<float>(float* p, float v) { *p = v; }
template <>
__device__ __forceinline__ void stf<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

struct Args {
  int B, S, Sk, nH, dq, dv, s_pad;   // S query rows, Sk keys (s_pad = Sk rounded up to the key tile)
  int64_t q_rs, k_rs, v_rs, o_rs;   // row strides in elements
  float scale;
  // optional ordered bias (model/ODA2/oda2_red_order_swin2_decoder.py:88-93): logit(i, j) += idx_bias[tok_idx[i] -
  // tok_idx[j] + n_emb - 1, head]; self-attention only (S == Sk); null = none
  const int* tok_idx;     // (B, S)
  const float* idx_bias;  // (2 n_emb - 1, nH)
  int n_emb;
};

template <typename T>
__global__ void __launch_bounds__(kThreads) mha_fwd_kernel(const Args a, const T* __restrict__ q, const T* __restrict__ k,
                                                           const T* __restrict__ v, T* __restrict__ o, float* __restrict__ attn) {
  extern __shared__ __align__(16) float sm[];
  float* sq = sm;                                   // [kRows][dq]
  float* skv0 = sq + kRows * a.dq;                  // 2 x [kKeys][kTileStride]: tile t in buffer t & 1
  float* ss = skv0 + 2 * kKeys * kTileStride;       // [kRows][s_pad]
  int* sidx = reinterpret_cast<int*>(ss + kRows * a.s_pad);   // [s_pad] token indices of this image (ordered bias only)
  float* sbias = reinterpret_cast<float*>(sidx + a.s_pad);    // [2 n_emb - 1] this head's bias column
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int row0 = blockIdx.x * kRows, h = blockIdx.y, b = blockIdx.z;
  const bool biased = a.tok_idx != nullptr;
  if (biased) {
    for (int i = tid; i < a.s_pad; i += kThreads) sidx[i] = i < a.Sk ? a.tok_idx[(int64_t)b * a.Sk + i] : 0;
    for (int i = tid; i < 2 * a.n_emb - 1; i += kThreads) sbias[i] = a.idx_bias[(int64_t)i * a.nH + h];
  }
  const T* qb = q + (int64_t)b * a.S * a.q_rs + h * a.dq;
  const T* kb = k + (int64_t)b * a.Sk * a.k_rs + h * a.dq;
  const T* vb = v + (int64_t)b * a.Sk * a.v_rs + h * a.dv;

  for (int i = tid; i < kRows * a.dq; i += kThreads) {
    const int r = i / a.dq, d = i - r * a.dq;
    sq[i] = row0 + r < a.S ? ldf(qb + (int64_t)(row0 + r) * a.q_rs + d) : 0.f;
  }

  // K / V tiles: the global loads of tile t + 1 are issued into registers BEFORE tile t is consumed and stored to the other
  // buffer after it, so their latency hides behind the FMAs (a tile is 64 x d <= 4096 values: <= 16 per thread).
  constexpr int kStage = kKeys * kMaxHeadDim / kThreads;
  float stage[kStage];
  auto tile_load = [&](const T* base, int64_t rs, int d_head, int c0) {
#pragma unroll
    for (int u = 0; u < kStage; ++u) {
      const int i = tid + u * kThreads, r = i / d_head, d = i - r * d_head;
      stage[u] = (i < kKeys * d_head && c0 + r < a.Sk) ? ldf(base + (int64_t)(c0 + r) * rs + d) : 0.f;
    }
  };
  auto tile_store = [&](float* buf, int d_head) {
#pragma unroll
    for (int u = 0; u < kStage; ++u) {
      const int i = tid + u * kThreads, r = i / d_head, d = i - r * d_head;
      if (i < kKeys * d_head) buf[r * kTileStride + d] = stage[u];
    }
  };
  const int n_tiles = a.s_pad / kKeys;

  // ---- logits: thread (j = tid % 64, ig = tid / 64) owns S[ig + 4 r][chunk + j], r = 0..7 ----
  const int j = tid & (kKeys - 1), ig = tid >> 6;
  tile_load(kb, a.k_rs, a.dq, 0);
  tile_store(skv0, a.dq);
  __syncthreads();                                  // sq and tile 0 ready
  for (int t = 0; t < n_tiles; ++t) {
    const int c0 = t * kKeys;
    const float* skv = skv0 + (t & 1) * kKeys * kTileStride;
    if (t + 1 < n_tiles) tile_load(kb, a.k_rs, a.dq, c0 + kKeys);
    float acc[8] = {};
    const float4* kr = reinterpret_cast<const float4*>(skv + j * kTileStride);
    for (int d4 = 0; d4 < a.dq / 4; ++d4) {      // qk_head_dim is a multiple of 4 (checked by the entry point)
      const float4 kv = kr[d4];
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const float4 qv = reinterpret_cast<const float4*>(sq + (ig + 4 * r) * a.dq)[d4];   // warp-wide broadcast
        acc[r] = fmaf(qv.x, kv.x, fmaf(qv.y, kv.y, fmaf(qv.z, kv.z, fmaf(qv.w, kv.w, acc[r]))));
      }
    }
    const bool live = c0 + j < a.Sk;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      float t = acc[r] * a.scale;
      if (biased) t += sbias[sidx[min(row0 + ig + 4 * r, a.S - 1)] - sidx[c0 + j] + a.n_emb - 1];
      ss[(ig + 4 * r) * a.s_pad + c0 + j] = live ? t : -INFINITY;
    }
    if (t + 1 < n_tiles) tile_store(skv0 + ((t + 1) & 1) * kKeys * kTileStride, a.dq);
    __syncthreads();                                // next tile stored, this one consumed, logits visible
  }
  tile_load(vb, a.v_rs, a.dv, 0);                   // first V tile: in flight during the softmax

  // ---- softmax: warp w owns rows w, w + 8, w + 16, w + 24; the normalised row goes to smem and to the map ----
  for (int r = warp; r < kRows; r += kThreads / 32) {
    float* srow = ss + r * a.s_pad;
    float m = -INFINITY;
    for (int c = lane; c < a.Sk; c += 32) m = fmaxf(m, srow[c]);
#pragma unroll
    for (int off = 16; off; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
    float sum = 0.f;
    for (int c = lane; c < a.s_pad; c += 32) {
      const float e = c < a.Sk ? expf(srow[c] - m) : 0.f;
      srow[c] = e;
      sum += e;
    }
#pragma unroll
    for (int off = 16; off; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
    const float inv = 1.f / sum;
    const bool row_live = row0 + r < a.S;
    float* arow = attn != nullptr && row_live ? attn + (((int64_t)b * a.nH + h) * a.S + row0 + r) * a.Sk : nullptr;
    for (int c = lane; c < a.s_pad; c += 32) {
      const float p = srow[c] * inv;
      srow[c] = p;
      if (arow != nullptr && c < a.Sk) arow[c] = p;
    }
  }

  // ---- o = P v: thread (d = tid % 64, ig) owns o[ig + 4 r][d] ----
  float acc[8] = {};
  tile_store(skv0, a.dv);
  __syncthreads();                                  // P rows and V tile 0 ready
  for (int t = 0; t < n_tiles; ++t) {
    const int c0 = t * kKeys;
    const float* skv = skv0 + (t & 1) * kKeys * kTileStride;
    if (t + 1 < n_tiles) tile_load(vb, a.v_rs, a.dv, c0 + kKeys);
    if (j < a.dv) {
      for (int kk = 0; kk < kKeys; kk += 4) {
        const float v0 = skv[kk * kTileStride + j], v1 = skv[(kk + 1) * kTileStride + j];
        const float v2 = skv[(kk + 2) * kTileStride + j], v3 = skv[(kk + 3) * kTileStride + j];
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          const float4 pv = *reinterpret_cast<const float4*>(ss + (ig + 4 * r) * a.s_pad + c0 + kk);   // broadcast
          acc[r] = fmaf(pv.w, v3, fmaf(pv.z, v2, fmaf(pv.y, v1, fmaf(pv.x, v0, acc[r]))));
        }
      }
    }
    if (t + 1 < n_tiles) tile_store(skv0 + ((t + 1) & 1) * kKeys * kTileStride, a.dv);
    __syncthreads();
  }
  if (j < a.dv) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int row = row0 + ig + 4 * r;
      if (row < a.S) stf(o + ((int64_t)b * a.S + row) * a.o_rs + h * a.dv + j, acc[r]);
    }
  }
}

}  // namespace mha

size_t mha_forward_smem_bytes(int kv_seq, int qk_head_dim, int n_emb) {
  const int s_pad = (kv_seq + mha::kKeys - 1) / mha::kKeys * mha::kKeys;
  return sizeof(float) * ((size_t)mha::kRows * qk_head_dim + 2 * (size_t)mha::kKeys * mha::kTileStride + (size_t)mha::kRows * s_pad +
                          (n_emb > 0 ? (size_t)s_pad + 2 * (size_t)n_emb : 0));
}

cudaError_t launch_mha_forward(int dtype, int batch, int seq, int kv_seq, int heads, int dq, int dv, int64_t q_rs, int64_t k_rs, int64_t v_rs,
                               int64_t o_rs, float scale, const void* q, const void* k, const void* v, void* o, float* attn,
                               const int* tok_idx, const float* idx_bias, int n_emb, cudaStream_t st) {
  mha::Args a;
  a.tok_idx = tok_idx; a.idx_bias = idx_bias; a.n_emb = tok_idx != nullptr ? n_emb : 0;
  a.B = batch; a.S = seq; a.Sk = kv_seq; a.nH = heads; a.dq = dq; a.dv = dv;
  a.s_pad = (kv_seq + mha::kKeys - 1) / mha::kKeys * mha::kKeys;
  a.q_rs = q_rs; a.k_rs = k_rs; a.v_rs = v_rs; a.o_rs = o_rs; a.scale = scale;
  const size_t smem = mha_forward_smem_bytes(kv_seq, dq, a.n_emb);
  const dim3 grid((seq + mha::kRows - 1) / mha::kRows, heads, batch);
  cudaError_t e;
  TimedRegion timed(st);
  if (dtype == SWIN_ATTN_F32) {
    e = cudaFuncSetAttribute(mha::mha_fwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    mha::mha_fwd_kernel<float><<<grid, mha::kThreads, smem, st>>>(a, static_cast<const float*>(q), static_cast<const float*>(k),
                                                                   static_cast<const float*>(v), static_cast<float*>(o), attn);
  } else {
    e = cudaFuncSetAttribute(mha::mha_fwd_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    mha::mha_fwd_kernel<__nv_bfloat16><<<grid, mha::kThreads, smem, st>>>(
        a, static_cast<const __nv_bfloat16*>(q), static_cast<const __nv_bfloat16*>(k), static_cast<const __nv_bfloat16*>(v),
        static_cast<__nv_bfloat16*>(o), attn);
  }
  count_launch();
  return cudaGetLastError();
}

}  // namespace swin
