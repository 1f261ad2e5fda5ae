// Shared definitions for the shifted-window attention kernels (sm_100a).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "swin_attn.h"

namespace swin {

constexpr int kHeadDim = 32;

// Launch-time description of one attention call; all strides in elements.
struct Problem {
  const void* q;
  const void* k;
  const void* v;
  const void* q_pad;   // nullable, C elements
  const void* k_pad;
  const void* v_pad;
  const float* table;  // ((2ws-1)^2, nH)
  const float* mask;   // nullable explicit additive mask (mask_nw, N, N)
  int mask_nw;
  void* o;             // (B,H,W,*)
  float* lse;          // (B,Hp,Wp,nH) or null
  // backward only
  const void* d_o;     // (B,H,W,C) dense
  const void* o_in;    // (B,H,W,C) dense
  void* dq;
  void* dk;
  void* dv;
  float* d_table;
  float* d_kpad;       // (C) fp32 or null: gradient of k_pad / v_pad (sum of dK / dV over the padding tokens) when
  float* d_vpad;       //   the backward runs on un-padded tensors (tcgen05 kernel only)
  long long* trace;    // debug: per-pair clock64 stamps of CTA 0 (tcgen05 kernel), or null
  int B, H, W, Hp, Wp, nH, C, ws, shift;
  int nWy, nWx;        // windows per column / row of the padded image
  int64_t q_ps, k_ps, v_ps, o_ps;
  float scale;
  float dropout_p;
  uint64_t seed;
};

// Token t of window (wy, wx) -> un-shifted coordinates in the padded image.
// Restates the composition roll(-s) . window_partition of the reference
// (newcrf_layers.py:220-227 with :39-41): shifted[ys] = x[(ys + s) mod Hp].
__device__ __forceinline__ void token_coord(const Problem& p, int wy, int wx, int t, int& y, int& x) {
  int i = t / p.ws, j = t - i * p.ws;
  y = wy * p.ws + i + p.shift;
  x = wx * p.ws + j + p.shift;
  if (y >= p.Hp) y -= p.Hp;
  if (x >= p.Wp) x -= p.Wp;
}

// SW-MSA region id of a token in SHIFTED coordinates (newcrf_layers.py:335-345):
// slices [0,Hp-ws) [Hp-ws,Hp-s) [Hp-s,Hp) along each axis, id = 3*rh + rw.
__device__ __forceinline__ int region_id(const Problem& p, int wy, int wx, int t) {
  int i = t / p.ws, j = t - i * p.ws;
  int ys = wy * p.ws + i, xs = wx * p.ws + j;
  int rh = ys < p.Hp - p.ws ? 0 : (ys < p.Hp - p.shift ? 1 : 2);
  int rw = xs < p.Wp - p.ws ? 0 : (xs < p.Wp - p.shift ? 1 : 2);
  return 3 * rh + rw;
}

// Attention dropout (nn.Dropout on the softmax output, newcrf_layers.py:137 / swin_transformer.py:139): element
// (window, head, query row, key column) is kept iff a counter-based hash of (seed, element index) is >= p 2^32,
// and kept probabilities are scaled by 1 / (1 - p).  The same function regenerates the mask in the backward.
// (Bitwise parity with torch's Philox stream is out of scope; every NewCRFs configuration uses p = 0.)
__host__ __device__ __forceinline__ uint32_t dropout_hash(uint64_t seed, uint64_t index) {
  uint64_t z = seed + 0x9E3779B97F4A7C15ull * (index + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (uint32_t)(z >> 32);
}
struct Dropout {
  uint64_t seed, base;   // base = ((window * nH + head) * N) * N
  uint32_t threshold;    // keep iff hash >= threshold
  float keep_scale;      // 1 / (1 - p); 1 when p == 0
  bool on;
  __device__ __forceinline__ float scale_of(int n, int row, int col) const {
    if (!on) return 1.0f;
    return dropout_hash(seed, base + (uint64_t)row * n + col) >= threshold ? keep_scale : 0.0f;
  }
};
__device__ __forceinline__ Dropout make_dropout(const struct Problem& p, int win, int head, int n);

// relative_position_index[i, j] (newcrf_layers.py:89-100).
__device__ __forceinline__ int rel_index(int ws, int ti, int tj) {
  int yi = ti / ws, xi = ti - yi * ws;
  int yj = tj / ws, xj = tj - yj * ws;
  return (yi - yj + ws - 1) * (2 * ws - 1) + (xi - xj + ws - 1);
}

__device__ __forceinline__ Dropout make_dropout(const Problem& p, int win, int head, int n) {
  Dropout d;
  d.on = p.dropout_p > 0.0f;
  d.seed = p.seed;
  d.base = ((uint64_t)win * p.nH + head) * (uint64_t)n * n;
  const double t = (double)p.dropout_p * 4294967296.0;
  d.threshold = t >= 4294967295.0 ? 0xffffffffu : (uint32_t)t;
  d.keep_scale = d.on ? 1.0f / (1.0f - p.dropout_p) : 1.0f;
  return d;
}

template <typename T>
struct Elem;
template <>
struct Elem<float> {
  static __device__ __forceinline__ float4 load4(const float* p) { return *reinterpret_cast<const float4*>(p); }
  static __device__ __forceinline__ void store4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
  static __device__ __forceinline__ void store1(float* p, float v) { *p = v; }
};
template <>
struct Elem<__nv_bfloat16> {
  static __device__ __forceinline__ float4 load4(const __nv_bfloat16* p) {
    uint2 raw = *reinterpret_cast<const uint2*>(p);
    __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&raw.x);
    __nv_bfloat162 b = *reinterpret_cast<__nv_bfloat162*>(&raw.y);
    float2 fa = __bfloat1622float2(a), fb = __bfloat1622float2(b);
    return make_float4(fa.x, fa.y, fb.x, fb.y);
  }
  static __device__ __forceinline__ void store4(__nv_bfloat16* p, float4 v) {
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
    uint2 raw;
    raw.x = *reinterpret_cast<uint32_t*>(&a);
    raw.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(p) = raw;
  }
  static __device__ __forceinline__ void store1(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
};

// Launchers implemented by the kernel translation units.
cudaError_t launch_simt_forward(const Problem& p, int dtype, cudaStream_t st);
cudaError_t launch_simt_backward(const Problem& p, int dtype, cudaStream_t st);
bool tc_forward_supported(const Problem& p, int dtype);
cudaError_t launch_tc_forward(const Problem& p, cudaStream_t st, const char** err);
bool tc12_forward_supported(const Problem& p, int dtype);
cudaError_t launch_tc12_forward(const Problem& p, cudaStream_t st, const char** err);
bool tc_backward_supported(const Problem& p, int dtype);
bool mma12_backward_supported(const Problem& p, int dtype);                 // window 12, bf16, aligned: warp-MMA kernel
cudaError_t launch_mma12_backward(const Problem& p, cudaStream_t st);
bool mma12_forward_supported(const Problem& p, int dtype);                  // window 12, bf16, padded or not: warp-MMA kernel
cudaError_t launch_mma12_forward(const Problem& p, cudaStream_t st);
cudaError_t launch_tc_backward(const Problem& p, cudaStream_t st, const char** err);

cudaError_t launch_layernorm(const void* x, int x_dtype, const float* gamma, const float* beta, float eps, int64_t rows, int channels,
                             void* out, int out_dtype, cudaStream_t st);
cudaError_t launch_layernorm_backward(const void* x, int x_dtype, const void* dy, int dy_dtype, const float* gamma, float eps, int64_t rows,
                                      int channels, void* dx, float* dgamma, float* dbeta, int sms, cudaStream_t st);
cudaError_t launch_add_layernorm(const void* x, int x_dtype, const void* delta, const float* gamma, const float* beta, const float* gate,
                                 int64_t rows_per_sample, float eps, int64_t rows, int channels, void* sum_out, void* y_out,
                                 cudaStream_t st);
cudaError_t launch_add_layernorm_backward(const void* x, int x_dtype, const void* dy, const void* gsum, const float* gamma, const float* gate,
                                          int64_t rows_per_sample, float eps, int64_t rows, int channels, void* dx, void* ddelta,
                                          float* dgamma, float* dbeta, int sms, cudaStream_t st);
cudaError_t launch_gelu(const void* x, const void* dy, void* out, int64_t n, cudaStream_t st);
cudaError_t launch_colsum(const void* dy, int64_t rows, int cols, float* out, int sms, cudaStream_t st);
bool layernorm_supported(int channels);
bool merge_layernorm_supported(int channels);
cudaError_t launch_merge_layernorm(const void* x, int x_dtype, const float* g, const float* b, float eps, int batch, int height, int width,
                                   int channels, void* out, cudaStream_t st);
bool tc_linear_supported(int64_t rows, int in_features, int out_features);
cudaError_t launch_tc_linear(const void* x, const void* weight, const void* bias, const void* residual, void* pre_act, void* y, int64_t rows,
                             int in_features, int out_features, int activation, cudaStream_t st, const char** err);

// Global multi-head attention that returns its map (Depthformer SelfAttentionBlock core; mha_global.cu)
size_t mha_forward_smem_bytes(int kv_seq, int qk_head_dim, int n_emb);
cudaError_t launch_mha_forward(int dtype, int batch, int seq, int kv_seq, int heads, int dq, int dv, int64_t q_rs, int64_t k_rs, int64_t v_rs,
                               int64_t o_rs, float scale, const void* q, const void* k, const void* v, void* o, float* attn,
                               const int* tok_idx, const float* idx_bias, int n_emb, cudaStream_t st);

bool mha_tc_supported(int dtype, int seq, int kv_seq, int dq, int dv, int64_t q_rs, int64_t k_rs, int64_t v_rs, int64_t o_rs, const void* q,
                      const void* k, const void* v, const void* o);
cudaError_t launch_mha_tc_forward(int batch, int seq, int kv_seq, int heads, int64_t q_rs, int64_t k_rs, int64_t v_rs, int64_t o_rs, float scale,
                                  const void* q, const void* k, const void* v, void* o, float* attn, const int* tok_idx,
                                  const float* idx_bias, int n_emb, cudaStream_t st, const char** err);

void set_tc_linear_pair_mode(int mode);
void count_launch(int n = 1);

// Launch-adjacent timing (swin_attn_time_next_launch): a pair of events armed by the calling thread is recorded
// immediately before and after the op's device work, i.e. after the host-side preparation (tensor-map encoding,
// attribute calls), so that a host-bound caller does not count its own launch path as kernel time.
struct LaunchEvents {
  cudaEvent_t start = nullptr, stop = nullptr;
};
LaunchEvents take_launch_events();   // returns and disarms the calling thread's events (null when none are armed)
struct TimedRegion {
  LaunchEvents ev;
  cudaStream_t st;
  explicit TimedRegion(cudaStream_t s) : ev(take_launch_events()), st(s) {
    if (ev.start) record(ev.start);
  }
  ~TimedRegion() {
    if (ev.stop) record(ev.stop);
  }
  // On a capturing stream the pair becomes two EXTERNAL event-record nodes of the graph: every replay re-records them,
  // so the launch can be timed inside the CUDA-graph replay (cudaEventElapsedTime after the replay has finished).
  void record(cudaEvent_t e) const {
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cs) == cudaSuccess && cs == cudaStreamCaptureStatusActive) cudaEventRecordWithFlags(e, st, cudaEventRecordExternal);
    else cudaEventRecord(e, st);
  }
  TimedRegion(const TimedRegion&) = delete;
  TimedRegion& operator=(const TimedRegion&) = delete;
};

}  // namespace swin
