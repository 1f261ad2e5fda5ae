// Block plumbing around the attention kernels: LayerNorm of the residual stream
// (fp32 in stage 0, bf16 after the first autocast Linear) written directly as the bf16 operand of the following Linear.
//
// Under torch.autocast the reference runs  x(fp32) -> LayerNorm (fp32 out) -> cast to bf16 -> Linear
// (newcrf_layers.py:207,253 / swin_transformer.py:201,243): 4+4 bytes per element for the norm and 4+2 for the
// cast.  This kernel reads x once and writes bf16 once (6 bytes per element), with the same fp32 arithmetic
// (mean, biased variance, rsqrt(var + eps), affine), so the Linear sees the same bf16 values up to the
// rounding of the fp32 statistics.
//
// One warp per row; every lane keeps its C/32 elements in registers (two-pass variance, no re-read).
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "swin_common.cuh"
#include "swin_gelu.cuh"

namespace swin {

__device__ __forceinline__ float2 load2(const float* p, int i) { return reinterpret_cast<const float2*>(p)[i]; }
__device__ __forceinline__ float2 load2(const __nv_bfloat16* p, int i) {
  return __bfloat1622float2(reinterpret_cast<const __nv_bfloat162*>(p)[i]);
}

template <int NC, typename TIn, typename TOut>   // NC pairs of elements per lane: C = 64 * NC
__global__ void __launch_bounds__(256) layernorm_rows_kernel(const TIn* __restrict__ x, const float* __restrict__ gamma,
                                                             const float* __restrict__ beta, float eps, int64_t rows,
                                                             TOut* __restrict__ out) {
  constexpr int C = 64 * NC;
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const TIn* xr = x + row * C;
  float2 v[NC];
  float sum = 0.f;
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    v[k] = load2(xr, lane + 32 * k);
    sum += v[k].x + v[k].y;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const float mean = sum * (1.0f / C);
  float sq = 0.f;
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    const float a = v[k].x - mean, b = v[k].y - mean;
    sq = fmaf(a, a, sq);
    sq = fmaf(b, b, sq);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  const float rstd = rsqrtf(sq * (1.0f / C) + eps);
  const float2* g2 = reinterpret_cast<const float2*>(gamma);
  const float2* b2 = reinterpret_cast<const float2*>(beta);
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    const float2 g = g2[lane + 32 * k], b = b2[lane + 32 * k];
    const float y0 = fmaf((v[k].x - mean) * rstd, g.x, b.x), y1 = fmaf((v[k].y - mean) * rstd, g.y, b.y);
    if constexpr (sizeof(TOut) == 2) {
      reinterpret_cast<__nv_bfloat162*>(out + row * C)[lane + 32 * k] = __floats2bfloat162_rn(y0, y1);
    } else {
      reinterpret_cast<float2*>(out + row * C)[lane + 32 * k] = make_float2(y0, y1);
    }
  }
}

// PatchMerging's gather fused with its LayerNorm (swin_transformer.py:271-287: pad H, W to even, concatenate the 2x2
// neighbourhood x[0::2,0::2] | x[1::2,0::2] | x[0::2,1::2] | x[1::2,1::2] along channels, then norm(4C)): output row
// (b, i, j) is read straight from the four pixels (2i + (s & 1), 2j + (s >> 1)), s = 0..3, of the (B, H, W, C) stage
// output (zeros beyond an odd H / W), so the concatenated tensor is never written.  Same lane <-> channel mapping and
// summation order as layernorm_rows_kernel on the concatenated row: bit-identical to cat + LayerNorm.
template <int NC, typename TIn>   // 4C = 64 * NC channels per output row, NC / 4 pairs per lane and segment
__global__ void __launch_bounds__(256) merge_layernorm_rows_kernel(const TIn* __restrict__ x, const float* __restrict__ gamma,
                                                                   const float* __restrict__ beta, float eps, int batch, int height,
                                                                   int width, __nv_bfloat16* __restrict__ out) {
  constexpr int C4 = 64 * NC, C = C4 / 4, KS = NC / 4;
  const int lane = threadIdx.x & 31;
  const int h2 = (height + 1) >> 1, w2 = (width + 1) >> 1;
  const int64_t rows = (int64_t)batch * h2 * w2;
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int j = (int)(row % w2), i = (int)((row / w2) % h2), b = (int)(row / ((int64_t)w2 * h2));
  float2 v[NC];
  float sum = 0.f;
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const int y = 2 * i + (s & 1), xx = 2 * j + (s >> 1);
    const bool inside = y < height && xx < width;
    const TIn* src = x + (((int64_t)b * height + y) * width + xx) * C;
#pragma unroll
    for (int k = 0; k < KS; ++k) {
      v[s * KS + k] = inside ? load2(src, lane + 32 * k) : make_float2(0.f, 0.f);
    }
  }
#pragma unroll
  for (int k = 0; k < NC; ++k) sum += v[k].x + v[k].y;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const float mean = sum * (1.0f / C4);
  float sq = 0.f;
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    const float a = v[k].x - mean, c = v[k].y - mean;
    sq = fmaf(a, a, sq);
    sq = fmaf(c, c, sq);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  const float rstd = rsqrtf(sq * (1.0f / C4) + eps);
  const float2* g2 = reinterpret_cast<const float2*>(gamma);
  const float2* b2 = reinterpret_cast<const float2*>(beta);
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    const float2 g = g2[lane + 32 * k], bb = b2[lane + 32 * k];
    const float y0 = fmaf((v[k].x - mean) * rstd, g.x, bb.x), y1 = fmaf((v[k].y - mean) * rstd, g.y, bb.y);
    reinterpret_cast<__nv_bfloat162*>(out + row * C4)[lane + 32 * k] = __floats2bfloat162_rn(y0, y1);
  }
}

// Residual add fused with the LayerNorm that follows it (block plumbing, swin_transformer.py:240-244 /
// newcrf_layers.py:251-255: `x = shortcut + attn; ... norm2(x)`, and `x + mlp` followed by the next block's norm1):
// s = x + delta rounded to the dtype of the residual stream (what torch's type promotion gives: fp32 + bf16 -> fp32,
// bf16 + bf16 -> bf16), y = LayerNorm(s) as bf16.  One read of x and delta, one write of s and y, instead of
// add (2 reads, 1 write) + norm (1 read, 1 write).
// `gate` (optional): the stochastic-depth scale of delta, one fp32 per sample (DropPath.gate: 0 or 1 / keep), applied as
// torch.addcmul does (one FMA): s = x + delta * gate[row / rows_per_sample].
template <int NC, typename TX>
__global__ void __launch_bounds__(256) add_layernorm_rows_kernel(const TX* __restrict__ x, const __nv_bfloat16* __restrict__ delta,
                                                                 const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                 const float* __restrict__ gate, int64_t rows_per_sample,
                                                                 float eps, int64_t rows, TX* __restrict__ sum_out,
                                                                 __nv_bfloat16* __restrict__ y_out) {
  constexpr int C = 64 * NC;
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const float sc = gate != nullptr ? gate[row / rows_per_sample] : 1.0f;
  float2 v[NC];
  float sum = 0.f;
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    const float2 a = load2(x + row * C, lane + 32 * k), d = load2(delta + row * C, lane + 32 * k);
    // without a gate sc == 1 and fmaf(d, 1, a) == a + d exactly
    const float s0 = fmaf(d.x, sc, a.x), s1 = fmaf(d.y, sc, a.y);
    if constexpr (sizeof(TX) == 2) {
      const __nv_bfloat162 r = __floats2bfloat162_rn(s0, s1);
      reinterpret_cast<__nv_bfloat162*>(sum_out + row * C)[lane + 32 * k] = r;
      v[k] = __bfloat1622float2(r);
    } else {
      v[k] = make_float2(s0, s1);
      reinterpret_cast<float2*>(sum_out + row * C)[lane + 32 * k] = v[k];
    }
    sum += v[k].x + v[k].y;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const float mean = sum * (1.0f / C);
  float sq = 0.f;
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    const float a = v[k].x - mean, b = v[k].y - mean;
    sq = fmaf(a, a, sq);
    sq = fmaf(b, b, sq);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  const float rstd = rsqrtf(sq * (1.0f / C) + eps);
  const float2* g2 = reinterpret_cast<const float2*>(gamma);
  const float2* b2 = reinterpret_cast<const float2*>(beta);
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    const float2 g = g2[lane + 32 * k], b = b2[lane + 32 * k];
    const float y0 = fmaf((v[k].x - mean) * rstd, g.x, b.x), y1 = fmaf((v[k].y - mean) * rstd, g.y, b.y);
    reinterpret_cast<__nv_bfloat162*>(y_out + row * C)[lane + 32 * k] = __floats2bfloat162_rn(y0, y1);
  }
}

// Backward of the kernel above, for training under autocast: dy is the bf16 gradient the Linear's dgrad produced,
// x the saved residual stream; mean / rstd are recomputed from the row (it is in registers anyway), so nothing but x
// is saved.  dx = rstd (g - mean(g) - xhat mean(g xhat)), g = dy gamma;  dgamma = sum_rows dy xhat;  dbeta = sum_rows dy.
// One warp per row, grid-stride over rows; every warp accumulates its dgamma / dbeta partials in a private shared-
// memory row (lane l owns the same channels for every row, so there are no conflicts and no atomics), the CTA
// reduces its 8 rows at the end and adds them to global memory with one atomic per channel.
// kResidual: backward of the fused residual add + LayerNorm (add_layernorm_rows_kernel).  `gsum` is the gradient that
// reaches the sum through the residual stream; dx = gsum + (LayerNorm backward of dy) is the gradient of the stream
// before the add, and ddelta = bf16(dx * gate) that of the bf16 branch output -- one pass instead of LayerNorm
// backward + autograd's gradient accumulation + the cast / gate multiply of the branch gradient.
template <int NC, typename TX, typename TDy, bool kResidual = false>
__global__ void __launch_bounds__(256) layernorm_bwd_kernel(const TX* __restrict__ x, const TDy* __restrict__ dy,
                                                            const float* __restrict__ gamma, float eps, int64_t rows,
                                                            TX* __restrict__ dx, float* __restrict__ dgamma,
                                                            float* __restrict__ dbeta, const TX* __restrict__ gsum = nullptr,
                                                            const float* __restrict__ gate = nullptr, int64_t rows_per_sample = 1,
                                                            __nv_bfloat16* __restrict__ ddelta = nullptr) {
  constexpr int C = 64 * NC;
  extern __shared__ __align__(16) float2 acc[];   // [8 warps][2 (dgamma, dbeta)][C / 2]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float2* ag = acc + (warp * 2) * (C / 2);
  float2* ab = ag + C / 2;
#pragma unroll
  for (int k = 0; k < NC; ++k) {
    ag[lane + 32 * k] = make_float2(0.f, 0.f);
    ab[lane + 32 * k] = make_float2(0.f, 0.f);
  }
  const float2* g2 = reinterpret_cast<const float2*>(gamma);
  for (int64_t row = (int64_t)blockIdx.x * 8 + warp; row < rows; row += (int64_t)gridDim.x * 8) {
    float2 v[NC], d[NC];
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < NC; ++k) {
      v[k] = load2(x + row * C, lane + 32 * k);
      d[k] = load2(dy + row * C, lane + 32 * k);
      sum += v[k].x + v[k].y;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float mean = sum * (1.0f / C);
    float sq = 0.f;
#pragma unroll
    for (int k = 0; k < NC; ++k) {
      v[k].x -= mean;
      v[k].y -= mean;
      sq = fmaf(v[k].x, v[k].x, sq);
      sq = fmaf(v[k].y, v[k].y, sq);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
    const float rstd = rsqrtf(sq * (1.0f / C) + eps);
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int k = 0; k < NC; ++k) {
      const float2 g = g2[lane + 32 * k];
      v[k].x *= rstd;   // xhat
      v[k].y *= rstd;
      float2 a = ag[lane + 32 * k], b = ab[lane + 32 * k];
      a.x = fmaf(d[k].x, v[k].x, a.x); a.y = fmaf(d[k].y, v[k].y, a.y);
      b.x += d[k].x; b.y += d[k].y;
      ag[lane + 32 * k] = a;
      ab[lane + 32 * k] = b;
      d[k].x *= g.x;    // g = dy gamma
      d[k].y *= g.y;
      s1 += d[k].x + d[k].y;
      s2 = fmaf(d[k].x, v[k].x, s2);
      s2 = fmaf(d[k].y, v[k].y, s2);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      s1 += __shfl_xor_sync(0xffffffffu, s1, o);
      s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    }
    s1 *= (1.0f / C);
    s2 *= (1.0f / C);
    float sc = 1.0f;
    if constexpr (kResidual) sc = gate != nullptr ? gate[row / rows_per_sample] : 1.0f;
#pragma unroll
    for (int k = 0; k < NC; ++k) {
      float y0 = rstd * (d[k].x - s1 - v[k].x * s2), y1 = rstd * (d[k].y - s1 - v[k].y * s2);
      if constexpr (kResidual) {
        const float2 g = load2(gsum + row * C, lane + 32 * k);
        y0 += g.x;
        y1 += g.y;
        reinterpret_cast<__nv_bfloat162*>(ddelta + row * C)[lane + 32 * k] = __floats2bfloat162_rn(y0 * sc, y1 * sc);
      }
      if constexpr (sizeof(TX) == 2) {
        reinterpret_cast<__nv_bfloat162*>(dx + row * C)[lane + 32 * k] = __floats2bfloat162_rn(y0, y1);
      } else {
        reinterpret_cast<float2*>(dx + row * C)[lane + 32 * k] = make_float2(y0, y1);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += 256) {
    float sg = 0.f, sb = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      sg += reinterpret_cast<const float*>(acc + (w * 2) * (C / 2))[i];
      sb += reinterpret_cast<const float*>(acc + (w * 2 + 1) * (C / 2))[i];
    }
    atomicAdd(dgamma + i, sg);
    atomicAdd(dbeta + i, sb);
  }
}

template <int NC>
static cudaError_t launch_ln_bwd(const void* x, int x_dtype, const void* dy, int dy_dtype, const float* g, float eps, int64_t rows, void* dx,
                                 float* dgamma, float* dbeta, int sms, cudaStream_t st) {
  const size_t smem = (size_t)8 * 2 * 64 * NC * sizeof(float);
  int64_t want = (rows + 7) / 8;
  const int64_t cap = (int64_t)sms * (smem > 96 * 1024 ? 1 : smem > 48 * 1024 ? 2 : 4);
  const unsigned grid = (unsigned)(want < cap ? want : cap);
  cudaError_t e = cudaMemsetAsync(dgamma, 0, sizeof(float) * 64 * NC, st);
  if (e == cudaSuccess) e = cudaMemsetAsync(dbeta, 0, sizeof(float) * 64 * NC, st);
  if (e != cudaSuccess) return e;
  using bf = __nv_bfloat16;
#define SWIN_LN_BWD(TX, TDY)                                                                                        \
  do {                                                                                                              \
    auto k = layernorm_bwd_kernel<NC, TX, TDY>;                                                                     \
    if ((e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e; \
    k<<<grid, 256, smem, st>>>(static_cast<const TX*>(x), static_cast<const TDY*>(dy), g, eps, rows, static_cast<TX*>(dx), dgamma, dbeta, \
                               nullptr, nullptr, 1, nullptr);                                                        \
  } while (0)
  if (x_dtype == SWIN_ATTN_F32 && dy_dtype == SWIN_ATTN_F32) SWIN_LN_BWD(float, float);
  else if (x_dtype == SWIN_ATTN_F32) SWIN_LN_BWD(float, bf);
  else if (dy_dtype == SWIN_ATTN_F32) SWIN_LN_BWD(bf, float);
  else SWIN_LN_BWD(bf, bf);
#undef SWIN_LN_BWD
  count_launch();
  return cudaGetLastError();
}

template <int NC>
static cudaError_t launch_add_ln_bwd(const void* x, int x_dtype, const void* dy, const void* gsum, const float* g, const float* gate,
                                     int64_t rows_per_sample, float eps, int64_t rows, void* dx, void* ddelta, float* dgamma, float* dbeta,
                                     int sms, cudaStream_t st) {
  const size_t smem = (size_t)8 * 2 * 64 * NC * sizeof(float);
  int64_t want = (rows + 7) / 8;
  const int64_t cap = (int64_t)sms * (smem > 96 * 1024 ? 1 : smem > 48 * 1024 ? 2 : 4);
  const unsigned grid = (unsigned)(want < cap ? want : cap);
  cudaError_t e = cudaMemsetAsync(dgamma, 0, sizeof(float) * 64 * NC, st);
  if (e == cudaSuccess) e = cudaMemsetAsync(dbeta, 0, sizeof(float) * 64 * NC, st);
  if (e != cudaSuccess) return e;
  using bf = __nv_bfloat16;
  if (x_dtype == SWIN_ATTN_F32) {
    auto k = layernorm_bwd_kernel<NC, float, bf, true>;
    if ((e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
    k<<<grid, 256, smem, st>>>(static_cast<const float*>(x), static_cast<const bf*>(dy), g, eps, rows, static_cast<float*>(dx), dgamma, dbeta,
                               static_cast<const float*>(gsum), gate, rows_per_sample, static_cast<bf*>(ddelta));
  } else {
    auto k = layernorm_bwd_kernel<NC, bf, bf, true>;
    if ((e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
    k<<<grid, 256, smem, st>>>(static_cast<const bf*>(x), static_cast<const bf*>(dy), g, eps, rows, static_cast<bf*>(dx), dgamma, dbeta,
                               static_cast<const bf*>(gsum), gate, rows_per_sample, static_cast<bf*>(ddelta));
  }
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_add_layernorm_backward(const void* x, int x_dtype, const void* dy, const void* gsum, const float* g, const float* gate,
                                          int64_t rows_per_sample, float eps, int64_t rows, int channels, void* dx, void* ddelta,
                                          float* dgamma, float* dbeta, int sms, cudaStream_t st) {
  switch (channels / 64) {
#define SWIN_LN_CASE(NC) \
  case NC: return launch_add_ln_bwd<NC>(x, x_dtype, dy, gsum, g, gate, rows_per_sample, eps, rows, dx, ddelta, dgamma, dbeta, sms, st)
    SWIN_LN_CASE(2); SWIN_LN_CASE(3); SWIN_LN_CASE(4); SWIN_LN_CASE(6); SWIN_LN_CASE(8);
    SWIN_LN_CASE(12); SWIN_LN_CASE(16); SWIN_LN_CASE(24);
#undef SWIN_LN_CASE
    default: return cudaErrorInvalidValue;
  }
}

cudaError_t launch_layernorm_backward(const void* x, int x_dtype, const void* dy, int dy_dtype, const float* g, float eps, int64_t rows,
                                      int channels, void* dx, float* dgamma, float* dbeta, int sms, cudaStream_t st) {
  switch (channels / 64) {
#define SWIN_LN_CASE(NC) case NC: return launch_ln_bwd<NC>(x, x_dtype, dy, dy_dtype, g, eps, rows, dx, dgamma, dbeta, sms, st)
    SWIN_LN_CASE(2); SWIN_LN_CASE(3); SWIN_LN_CASE(4); SWIN_LN_CASE(6); SWIN_LN_CASE(8);
    SWIN_LN_CASE(12); SWIN_LN_CASE(16); SWIN_LN_CASE(24);   // 3072 channels: the row does not fit in registers
#undef SWIN_LN_CASE
    default: return cudaErrorInvalidValue;
  }
}

template <int NC>
static cudaError_t launch_ln(const void* x, int x_dtype, const float* g, const float* b, float eps, int64_t rows, void* out,
                             int out_dtype, cudaStream_t st) {
  const unsigned grid = (unsigned)((rows + 7) / 8);
  using bf = __nv_bfloat16;
  if (x_dtype == SWIN_ATTN_F32 && out_dtype == SWIN_ATTN_BF16)
    layernorm_rows_kernel<NC, float, bf><<<grid, 256, 0, st>>>(static_cast<const float*>(x), g, b, eps, rows, static_cast<bf*>(out));
  else if (x_dtype == SWIN_ATTN_F32)
    layernorm_rows_kernel<NC, float, float><<<grid, 256, 0, st>>>(static_cast<const float*>(x), g, b, eps, rows, static_cast<float*>(out));
  else if (out_dtype == SWIN_ATTN_BF16)
    layernorm_rows_kernel<NC, bf, bf><<<grid, 256, 0, st>>>(static_cast<const bf*>(x), g, b, eps, rows, static_cast<bf*>(out));
  else
    layernorm_rows_kernel<NC, bf, float><<<grid, 256, 0, st>>>(static_cast<const bf*>(x), g, b, eps, rows, static_cast<float*>(out));
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_layernorm(const void* x, int x_dtype, const float* g, const float* b, float eps, int64_t rows, int channels,
                             void* out, int out_dtype, cudaStream_t st) {
  switch (channels / 64) {
#define SWIN_LN_CASE(NC) case NC: return launch_ln<NC>(x, x_dtype, g, b, eps, rows, out, out_dtype, st)
    SWIN_LN_CASE(2); SWIN_LN_CASE(3); SWIN_LN_CASE(4); SWIN_LN_CASE(6); SWIN_LN_CASE(8);
    SWIN_LN_CASE(12); SWIN_LN_CASE(16); SWIN_LN_CASE(24); SWIN_LN_CASE(48);
#undef SWIN_LN_CASE
    default: return cudaErrorInvalidValue;
  }
}

template <int NC>
static cudaError_t launch_merge_ln(const void* x, int x_dtype, const float* g, const float* b, float eps, int batch, int height, int width,
                                   void* out, cudaStream_t st) {
  const int64_t rows = (int64_t)batch * ((height + 1) / 2) * ((width + 1) / 2);
  const unsigned grid = (unsigned)((rows + 7) / 8);
  using bf = __nv_bfloat16;
  if (x_dtype == SWIN_ATTN_F32)
    merge_layernorm_rows_kernel<NC, float><<<grid, 256, 0, st>>>(static_cast<const float*>(x), g, b, eps, batch, height, width, static_cast<bf*>(out));
  else
    merge_layernorm_rows_kernel<NC, bf><<<grid, 256, 0, st>>>(static_cast<const bf*>(x), g, b, eps, batch, height, width, static_cast<bf*>(out));
  count_launch();
  return cudaGetLastError();
}

bool merge_layernorm_supported(int channels) {
  return channels == 128 || channels == 192 || channels == 256 || channels == 384 || channels == 512 || channels == 768;
}

cudaError_t launch_merge_layernorm(const void* x, int x_dtype, const float* g, const float* b, float eps, int batch, int height, int width,
                                   int channels, void* out, cudaStream_t st) {
  switch (channels / 16) {   // 4C / 64
    case 8: return launch_merge_ln<8>(x, x_dtype, g, b, eps, batch, height, width, out, st);
    case 12: return launch_merge_ln<12>(x, x_dtype, g, b, eps, batch, height, width, out, st);
    case 16: return launch_merge_ln<16>(x, x_dtype, g, b, eps, batch, height, width, out, st);
    case 32: return launch_merge_ln<32>(x, x_dtype, g, b, eps, batch, height, width, out, st);
    case 24: return launch_merge_ln<24>(x, x_dtype, g, b, eps, batch, height, width, out, st);
    case 48: return launch_merge_ln<48>(x, x_dtype, g, b, eps, batch, height, width, out, st);
    default: return cudaErrorInvalidValue;
  }
}

template <int NC>
static cudaError_t launch_add_ln(const void* x, int x_dtype, const void* delta, const float* g, const float* b, const float* gate,
                                 int64_t rows_per_sample, float eps, int64_t rows, void* sum_out, void* y_out, cudaStream_t st) {
  const unsigned grid = (unsigned)((rows + 7) / 8);
  using bf = __nv_bfloat16;
  if (x_dtype == SWIN_ATTN_F32)
    add_layernorm_rows_kernel<NC, float><<<grid, 256, 0, st>>>(static_cast<const float*>(x), static_cast<const bf*>(delta), g, b, gate, rows_per_sample,
                                                               eps, rows, static_cast<float*>(sum_out), static_cast<bf*>(y_out));
  else
    add_layernorm_rows_kernel<NC, bf><<<grid, 256, 0, st>>>(static_cast<const bf*>(x), static_cast<const bf*>(delta), g, b, gate, rows_per_sample, eps,
                                                            rows, static_cast<bf*>(sum_out), static_cast<bf*>(y_out));
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_add_layernorm(const void* x, int x_dtype, const void* delta, const float* g, const float* b, const float* gate,
                                 int64_t rows_per_sample, float eps, int64_t rows, int channels, void* sum_out, void* y_out,
                                 cudaStream_t st) {
  switch (channels / 64) {
#define SWIN_LN_CASE(NC) case NC: return launch_add_ln<NC>(x, x_dtype, delta, g, b, gate, rows_per_sample, eps, rows, sum_out, y_out, st)
    SWIN_LN_CASE(2); SWIN_LN_CASE(3); SWIN_LN_CASE(4); SWIN_LN_CASE(6); SWIN_LN_CASE(8);
    SWIN_LN_CASE(12); SWIN_LN_CASE(16); SWIN_LN_CASE(24);
#undef SWIN_LN_CASE
    default: return cudaErrorInvalidValue;
  }
}

// ---- exact (erf) GELU of the MLP, bf16 -> bf16 ------------------------------------------------------------------
// nn.GELU() between fc1 and fc2 (newcrf_layers.py:9-27, swin_transformer.py Mlp).  ATen's kernel evaluates
// x/2 (1 + erff(x/sqrt 2)) and is bound by erff's instruction count (~0.6 of copy speed); here erfc is the
// exp(-z^2 + P(t)) t form with t = 1/(1 + z/2) (fractional error < 1.2e-7 everywhere, Numerical Recipes erfcc), i.e.
// two MUFU ops and ten FMAs, and the negative tail x/2 erfc(|x|/sqrt 2) has no cancellation -- closer to the fp64 value
// than 1 + erff is.  16-byte vectors, two per thread.
template <bool kBackward>
__global__ void __launch_bounds__(256) gelu_kernel(const uint4* __restrict__ x, const uint4* __restrict__ dy, uint4* __restrict__ out,
                                                   int64_t nvec) {
  const int64_t half = (nvec + 1) / 2;
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= half) return;
  const int64_t j = i + half;
  const bool two = j < nvec;
  uint4 a[2], g[2];
  a[0] = x[i];
  a[1] = two ? x[j] : make_uint4(0, 0, 0, 0);
  if (kBackward) {
    g[0] = dy[i];
    g[1] = two ? dy[j] : make_uint4(0, 0, 0, 0);
  }
#pragma unroll
  for (int v = 0; v < 2; ++v) {
    uint32_t* w = reinterpret_cast<uint32_t*>(&a[v]);
    const uint32_t* gw = reinterpret_cast<const uint32_t*>(&g[v]);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[k]));
      float2 d = make_float2(0.f, 0.f);
      if (kBackward) d = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gw[k]));
      const float2 o = gelu_value2<kBackward>(f, d);
      const __nv_bfloat162 r = __floats2bfloat162_rn(o.x, o.y);
      w[k] = *reinterpret_cast<const uint32_t*>(&r);
    }
  }
  out[i] = a[0];
  if (two) out[j] = a[1];
}

cudaError_t launch_gelu(const void* x, const void* dy, void* out, int64_t n, cudaStream_t st) {
  const int64_t nvec = n / 8, half = (nvec + 1) / 2;
  const unsigned grid = (unsigned)((half + 255) / 256);
  if (dy == nullptr)
    gelu_kernel<false><<<grid, 256, 0, st>>>(static_cast<const uint4*>(x), nullptr, static_cast<uint4*>(out), nvec);
  else
    gelu_kernel<true><<<grid, 256, 0, st>>>(static_cast<const uint4*>(x), static_cast<const uint4*>(dy), static_cast<uint4*>(out), nvec);
  count_launch();
  return cudaGetLastError();
}

// ---- bias gradient of the per-token Linears: column sums of the bf16 output gradient -----------------------------
// autograd of `x @ W^T + b` (qkv / qk / proj / fc1 / fc2, newcrf_layers.py:102-147, swin_transformer.py:104-142) needs
// db = sum over tokens of dy.  ATen runs a generic column reduction at ~2 TB/s, 137 times per training step (5.5 ms of a
// 48 ms step); this kernel streams dy once with 16-byte loads: a warp covers 256 columns of a row, a CTA 8 rows at a
// time over its row chunk, partials are combined across the 8 warps in shared memory and added with one fp32 atomic
// per column per CTA (the output is zeroed by the launcher).
__global__ void __launch_bounds__(256) colsum_kernel(const __nv_bfloat16* __restrict__ dy, int64_t rows, int cols, int64_t rows_per_cta,
                                                     float* __restrict__ out) {
  __shared__ float part[8][256];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col = blockIdx.x * 256 + lane * 8;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_cta, r1 = min(r0 + rows_per_cta, rows);
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (col < cols) {
    for (int64_t r = r0 + warp; r < r1; r += 8) {
      const uint4 v = *reinterpret_cast<const uint4*>(dy + r * cols + col);
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[k]));
        acc[2 * k] += f.x;
        acc[2 * k + 1] += f.y;
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) part[warp][lane * 8 + k] = acc[k];
  __syncthreads();
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c < cols) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += part[w][threadIdx.x];
    atomicAdd(out + c, s);
  }
}

cudaError_t launch_colsum(const void* dy, int64_t rows, int cols, float* out, int sms, cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(out, 0, sizeof(float) * cols, st);
  if (e != cudaSuccess) return e;
  const int col_tiles = (cols + 255) / 256;
  int64_t chunks = ((int64_t)sms * 8 + col_tiles - 1) / col_tiles;          // ~8 CTAs per SM in total
  const int64_t max_chunks = (rows + 63) / 64;                                // at least 64 rows per CTA
  if (chunks > max_chunks) chunks = max_chunks;
  if (chunks < 1) chunks = 1;
  if (chunks > 65535) chunks = 65535;
  const int64_t per = (rows + chunks - 1) / chunks;
  colsum_kernel<<<dim3(col_tiles, (unsigned)((rows + per - 1) / per)), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(dy), rows, cols, per, out);
  count_launch();
  return cudaGetLastError();
}

bool layernorm_supported(int channels) {
  if (channels % 64) return false;
  switch (channels / 64) {
    case 2: case 3: case 4: case 6: case 8: case 12: case 16: case 24: case 48: return true;
    default: return false;
  }
}

}  // namespace swin
