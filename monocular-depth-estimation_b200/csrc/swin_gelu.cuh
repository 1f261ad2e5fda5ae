// Exact (erf) GELU arithmetic shared by the stand-alone GELU kernels (swin_block_ops.cu) and the epilogue of the
// fused Linear + GELU kernel (swin_linear_tc.cu): both produce the same bits for the same bf16 input.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

namespace swin {

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// two elements at a time: packed fp32x2 FMAs halve the instruction count of the polynomial (the kernel is issue-bound
// otherwise: ~25 scalar instructions per element against ~4 bytes of traffic)
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ float2 splat(float v) { return make_float2(v, v); }

// 0.5 erfc(|x| / sqrt 2) = Phi(-|x|) from nx = -|x| (two elements)
__device__ __forceinline__ float2 half_erfc2(float2 nx) {
  const float2 z = __fmul2_rn(nx, splat(-0.70710678118654752440f));
  const float2 zn = __fmul2_rn(nx, splat(0.70710678118654752440f));
  const float2 d = __ffma2_rn(splat(0.5f), z, splat(1.0f));
  const float2 t = make_float2(rcp_approx(d.x), rcp_approx(d.y));
  float2 p = splat(0.17087277f);
  p = __ffma2_rn(p, t, splat(-0.82215223f));
  p = __ffma2_rn(p, t, splat(1.48851587f));
  p = __ffma2_rn(p, t, splat(-1.13520398f));
  p = __ffma2_rn(p, t, splat(0.27886807f));
  p = __ffma2_rn(p, t, splat(-0.18628806f));
  p = __ffma2_rn(p, t, splat(0.09678418f));
  p = __ffma2_rn(p, t, splat(0.37409196f));
  p = __ffma2_rn(p, t, splat(1.00002368f));
  p = __ffma2_rn(p, t, splat(-1.26551223f));
  // exp(-z^2 + P(t)) / 2 as one ex2: the halving is the -1 in the exponent
  const float2 a = __ffma2_rn(__ffma2_rn(z, zn, p), splat(1.4426950408889634f), splat(-1.0f));
  return __fmul2_rn(t, make_float2(ex2_approx(a.x), ex2_approx(a.y)));
}

// forward: x Phi(x) = relu(x) - |x| Phi(-|x|) (one FMA, no cancellation in the negative tail, no select);
// backward: dy (Phi(x) + x phi(x))
template <bool kBackward>
__device__ __forceinline__ float2 gelu_value2(float2 x, float2 dy) {
  // the clamp keeps -|x| finite so that an infinite input gives inf / NaN as x Phi(x) does
  const float2 nx = make_float2(fmaxf(-fabsf(x.x), -3.0e38f), fmaxf(-fabsf(x.y), -3.0e38f));
  const float2 h = half_erfc2(nx);
  if (!kBackward) return __ffma2_rn(nx, h, make_float2(fmaxf(x.x, 0.f), fmaxf(x.y, 0.f)));
  const float2 cdf = make_float2(x.x >= 0.f ? 1.0f - h.x : h.x, x.y >= 0.f ? 1.0f - h.y : h.y);   // Phi(x)
  const float2 q = __fmul2_rn(__fmul2_rn(x, x), splat(-0.72134752044448170f));
  const float2 pdf = __fmul2_rn(splat(0.3989422804014327f), make_float2(ex2_approx(q.x), ex2_approx(q.y)));   // exp(-x^2/2)/sqrt(2 pi)
  return __fmul2_rn(dy, __ffma2_rn(x, pdf, cdf));
}

}  // namespace swin
