// C ABI of libswin_attn.so (see include/swin_attn.h): argument checking,
// dispatch between the tcgen05 and the CUDA-core kernels, host-buffer entry.
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "swin_common.cuh"

namespace swin {
static std::atomic<uint64_t> g_launches{0};
void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }
static thread_local std::string g_error;
static thread_local LaunchEvents g_launch_events;
LaunchEvents take_launch_events() {
  const LaunchEvents ev = g_launch_events;
  g_launch_events = LaunchEvents();
  return ev;
}
static long long* g_trace = nullptr;

static int fail(const std::string& msg) {
  g_error = msg;
  return 1;
}
static int fail_cuda(const char* what, cudaError_t e) {
  g_error = std::string(what) + ": " + cudaGetErrorString(e);
  return 2;
}

static size_t elem_size(int dtype) { return dtype == SWIN_ATTN_F32 ? 4 : 2; }

// Validate the shape (mirrors the reference's asserts: shift in [0, ws),
// newcrf_layers.py:180; C divisible by heads) and fill the launch description.
static int make_problem(const swin_attn_shape_t* s, Problem& p) {
  if (s == nullptr) return fail("shape is null");
  if (s->dtype != SWIN_ATTN_F32 && s->dtype != SWIN_ATTN_BF16) return fail("dtype must be SWIN_ATTN_F32 or SWIN_ATTN_BF16");
  if (s->window != 7 && s->window != 12) return fail("window must be 7 or 12 (got " + std::to_string(s->window) + ")");
  if (s->shift < 0 || s->shift >= s->window) return fail("shift_size must in 0-window_size");
  if (s->batch <= 0 || s->height <= 0 || s->width <= 0 || s->num_heads <= 0) return fail("batch/height/width/num_heads must be positive");
  if (!(s->dropout_p >= 0.0f && s->dropout_p < 1.0f)) return fail("dropout_p must be in [0, 1)");
  memset(&p, 0, sizeof(p));
  p.B = s->batch; p.H = s->height; p.W = s->width; p.nH = s->num_heads;
  p.C = s->num_heads * kHeadDim;
  p.ws = s->window; p.shift = s->shift;
  p.nWy = (p.H + p.ws - 1) / p.ws; p.nWx = (p.W + p.ws - 1) / p.ws;
  p.Hp = p.nWy * p.ws; p.Wp = p.nWx * p.ws;
  p.q_ps = s->q_pix_stride; p.k_ps = s->k_pix_stride; p.v_ps = s->v_pix_stride; p.o_ps = s->o_pix_stride;
  if (p.q_ps < p.C || p.k_ps < p.C || p.v_ps < p.C || p.o_ps < p.C) return fail("pixel strides must be >= num_heads*32");
  if ((p.q_ps | p.k_ps | p.v_ps | p.o_ps) & 3) return fail("pixel strides must be multiples of 4 elements");
  if ((int64_t)p.B * p.Hp * p.Wp * p.nH >= (1ll << 31)) return fail("problem too large for 32-bit tile indices");
  p.scale = s->scale; p.dropout_p = s->dropout_p; p.seed = s->seed;
  return 0;
}

static int set_mask(const swin_attn_shape_t* s, const float* mask, Problem& p) {
  if (mask == nullptr) { p.mask = nullptr; p.mask_nw = 0; return 0; }
  if (s->mask_windows <= 0) return fail("attn_mask given but mask_windows <= 0");
  if ((p.B * p.nWy * p.nWx) % s->mask_windows != 0) return fail("number of windows is not a multiple of mask_windows");
  p.mask = mask; p.mask_nw = s->mask_windows;
  return 0;
}

static bool misaligned(const void* ptr, int dtype) {
  return (reinterpret_cast<uintptr_t>(ptr) & (4 * elem_size(dtype) - 1)) != 0;
}

}  // namespace swin

using namespace swin;

extern "C" {

int swin_attn_forward(const swin_attn_shape_t* shape, const void* q, const void* k, const void* v,
                      const void* q_pad, const void* k_pad, const void* v_pad, const float* bias_table,
                      const float* attn_mask, void* o, float* lse, void* stream) {
  Problem p;
  if (int rc = make_problem(shape, p)) return rc;
  if (!q || !k || !v || !bias_table || !o) return fail("q, k, v, bias_table and o must be non-null");
  if (misaligned(q, shape->dtype) || misaligned(k, shape->dtype) || misaligned(v, shape->dtype) || misaligned(o, shape->dtype))
    return fail("q/k/v/o must be aligned to 4 elements");
  p.q = q; p.k = k; p.v = v; p.q_pad = q_pad; p.k_pad = k_pad; p.v_pad = v_pad;
  p.table = bias_table; p.o = o; p.lse = lse; p.trace = g_trace;
  if (int rc = set_mask(shape, attn_mask, p)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // window 12, impl AUTO: the warp-MMA forward -- every lane carries a row, two CTAs per SM, padding is free -- beats the
  // tcgen05 window-12 kernel by 1.03x (aligned 96 x 96 grids) to 1.5x (the un-aligned stage grids of the window-12
  // encoder on a 352 x 1216 image): profiles/r2_micro_fwd_ws12_mma.jsonl.  impl = TCGEN05 forces the tcgen05 kernel.
  const bool mma_ok = attn_mask == nullptr && mma12_forward_supported(p, shape->dtype);
  if (shape->impl == SWIN_ATTN_IMPL_WARP_MMA && !mma_ok) return fail("warp-MMA forward does not support this shape/dtype (window 12, bf16, no explicit mask)");
  if (mma_ok && (shape->impl == SWIN_ATTN_IMPL_WARP_MMA || shape->impl == SWIN_ATTN_IMPL_AUTO)) {
    cudaError_t e = launch_mma12_forward(p, st);
    if (e != cudaSuccess) return fail_cuda("window-12 warp-MMA forward launch", e);
    return 0;
  }
  const bool want_tc12 = shape->impl != SWIN_ATTN_IMPL_SIMT && attn_mask == nullptr && tc12_forward_supported(p, shape->dtype);
  bool want_tc = want_tc12 || (shape->impl != SWIN_ATTN_IMPL_SIMT && attn_mask == nullptr && tc_forward_supported(p, shape->dtype));
  if (shape->impl == SWIN_ATTN_IMPL_TCGEN05 && !want_tc) return fail("tcgen05 forward does not support this shape/dtype");
  if (want_tc) {
    const char* why = nullptr;
    cudaError_t e = want_tc12 ? launch_tc12_forward(p, st, &why) : launch_tc_forward(p, st, &why);
    if (e != cudaSuccess) return why ? fail(std::string("tcgen05 forward: ") + why) : fail_cuda("tcgen05 forward launch", e);
    return 0;
  }
  cudaError_t e = launch_simt_forward(p, shape->dtype, st);
  if (e != cudaSuccess) return fail_cuda("forward launch", e);
  return 0;
}

static int backward_impl(const swin_attn_shape_t* shape, const void* q, const void* k, const void* v, const void* k_pad,
                         const void* v_pad, const void* o, const void* d_o, const float* lse, const float* bias_table,
                         const float* attn_mask, void* dq, void* dk, void* dv, float* d_bias_table, float* d_k_pad, float* d_v_pad,
                         bool unpadded, void* stream) {
  Problem p;
  if (int rc = make_problem(shape, p)) return rc;
  const bool ragged = p.H != p.Hp || p.W != p.Wp;
  if (ragged && !unpadded) return fail("backward needs height and width to be multiples of the window (pad explicitly)");
  if (!q || !k || !v || !o || !d_o || !lse || !bias_table || !dq || !dk || !dv || !d_bias_table)
    return fail("backward: all pointers must be non-null");
  if (p.o_ps != p.C) return fail("backward: o and d_o must be dense (pixel stride == C)");
  if (misaligned(q, shape->dtype) || misaligned(k, shape->dtype) || misaligned(v, shape->dtype) || misaligned(o, shape->dtype) ||
      misaligned(d_o, shape->dtype) || misaligned(dq, shape->dtype) || misaligned(dk, shape->dtype) || misaligned(dv, shape->dtype))
    return fail("backward: tensors must be aligned to 4 elements");
  p.q = q; p.k = k; p.v = v; p.o_in = o; p.d_o = d_o; p.lse = const_cast<float*>(lse);
  p.table = bias_table; p.dq = dq; p.dk = dk; p.dv = dv; p.d_table = d_bias_table;
  p.k_pad = k_pad; p.v_pad = v_pad; p.d_kpad = d_k_pad; p.d_vpad = d_v_pad; p.trace = g_trace;
  if (int rc = set_mask(shape, attn_mask, p)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  bool want_tc = shape->impl != SWIN_ATTN_IMPL_SIMT && attn_mask == nullptr && tc_backward_supported(p, shape->dtype);
  if (shape->impl == SWIN_ATTN_IMPL_TCGEN05 && !want_tc) return fail("tcgen05 backward does not support this shape/dtype");
  const bool want_mma12 = !want_tc && (shape->impl == SWIN_ATTN_IMPL_AUTO || shape->impl == SWIN_ATTN_IMPL_WARP_MMA) && attn_mask == nullptr &&
                          mma12_backward_supported(p, shape->dtype);
  if (ragged && !want_tc && !want_mma12)
    return fail("backward on un-padded tensors needs the tcgen05 kernel (bf16, window 7, shift 0 or 3) or the warp-MMA kernel (bf16, "
                "window 12, shift 0 or 6, height and width >= shift), no mask, 16-byte aligned tensors; pad explicitly otherwise");
  if (want_tc) {
    const char* why = nullptr;
    cudaError_t e = launch_tc_backward(p, st, &why);
    if (e != cudaSuccess) return why ? fail(std::string("tcgen05 backward: ") + why) : fail_cuda("tcgen05 backward launch", e);
    return 0;
  }
  if (shape->impl == SWIN_ATTN_IMPL_WARP_MMA && !mma12_backward_supported(p, shape->dtype))
    return fail("warp-MMA backward does not support this shape/dtype (window 12, bf16, window-aligned, no explicit mask / dropout)");
  if (want_mma12) {   // window 12: warp-MMA kernel
    cudaError_t e = launch_mma12_backward(p, st);
    if (e != cudaSuccess) return fail_cuda("window-12 backward launch", e);
    return 0;
  }
  cudaError_t e = launch_simt_backward(p, shape->dtype, st);
  if (e != cudaSuccess) return fail_cuda("backward launch", e);
  return 0;
}

int swin_attn_backward(const swin_attn_shape_t* shape, const void* q, const void* k, const void* v,
                       const void* o, const void* d_o, const float* lse, const float* bias_table,
                       const float* attn_mask, void* dq, void* dk, void* dv, float* d_bias_table, void* stream) {
  return backward_impl(shape, q, k, v, nullptr, nullptr, o, d_o, lse, bias_table, attn_mask, dq, dk, dv, d_bias_table, nullptr, nullptr,
                       false, stream);
}

int swin_attn_backward_unpadded(const swin_attn_shape_t* shape, const void* q, const void* k, const void* v, const void* k_pad,
                                const void* v_pad, const void* o, const void* d_o, const float* lse, const float* bias_table,
                                void* dq, void* dk, void* dv, float* d_bias_table, float* d_k_pad, float* d_v_pad, void* stream) {
  if ((k_pad != nullptr) != (d_k_pad != nullptr) || (v_pad != nullptr) != (d_v_pad != nullptr))
    return fail("backward (un-padded): d_k_pad / d_v_pad must be given exactly when k_pad / v_pad are");
  return backward_impl(shape, q, k, v, k_pad, v_pad, o, d_o, lse, bias_table, nullptr, dq, dk, dv, d_bias_table, d_k_pad, d_v_pad, true,
                       stream);
}

// Device scratch for the host-buffer entry point; grown on demand, reused.
static std::mutex g_host_mu;
static void* g_scratch = nullptr;
static size_t g_scratch_bytes = 0;

int swin_attn_forward_host(const swin_attn_shape_t* shape, const void* q, const void* k, const void* v,
                           const void* q_pad, const void* k_pad, const void* v_pad, const float* bias_table, void* o) {
  Problem p;
  if (int rc = make_problem(shape, p)) return rc;
  if (!q || !k || !v || !bias_table || !o) return fail("q, k, v, bias_table and o must be non-null");
  if (p.q_ps != p.C || p.k_ps != p.C || p.v_ps != p.C || p.o_ps != p.C) return fail("host entry takes dense (B,H,W,C) buffers");
  std::lock_guard<std::mutex> lock(g_host_mu);
  const size_t es = elem_size(shape->dtype);
  const size_t tens = (size_t)p.B * p.H * p.W * p.C * es;
  const size_t tens_al = (tens + 255) & ~size_t(255);
  const size_t vec = ((size_t)p.C * es + 255) & ~size_t(255);
  const size_t tab = (((size_t)(2 * p.ws - 1) * (2 * p.ws - 1) * p.nH * 4) + 255) & ~size_t(255);
  const size_t need = 4 * tens_al + 3 * vec + tab;
  cudaError_t e;
  if (need > g_scratch_bytes) {
    if (g_scratch) cudaFree(g_scratch);
    g_scratch = nullptr; g_scratch_bytes = 0;
    if ((e = cudaMalloc(&g_scratch, need)) != cudaSuccess) return fail_cuda("cudaMalloc scratch", e);
    g_scratch_bytes = need;
  }
  char* base = static_cast<char*>(g_scratch);
  void* dq = base; void* dk = base + tens_al; void* dv = base + 2 * tens_al; void* dout = base + 3 * tens_al;
  char* vecs = base + 4 * tens_al;
  void* dqp = q_pad ? vecs : nullptr; void* dkp = k_pad ? vecs + vec : nullptr; void* dvp = v_pad ? vecs + 2 * vec : nullptr;
  float* dtab = reinterpret_cast<float*>(vecs + 3 * vec);
  cudaStream_t st = nullptr;
  if ((e = cudaMemcpyAsync(dq, q, tens, cudaMemcpyHostToDevice, st)) != cudaSuccess) return fail_cuda("H2D q", e);
  if ((e = cudaMemcpyAsync(dk, k, tens, cudaMemcpyHostToDevice, st)) != cudaSuccess) return fail_cuda("H2D k", e);
  if ((e = cudaMemcpyAsync(dv, v, tens, cudaMemcpyHostToDevice, st)) != cudaSuccess) return fail_cuda("H2D v", e);
  if (q_pad) cudaMemcpyAsync(dqp, q_pad, p.C * es, cudaMemcpyHostToDevice, st);
  if (k_pad) cudaMemcpyAsync(dkp, k_pad, p.C * es, cudaMemcpyHostToDevice, st);
  if (v_pad) cudaMemcpyAsync(dvp, v_pad, p.C * es, cudaMemcpyHostToDevice, st);
  cudaMemcpyAsync(dtab, bias_table, (size_t)(2 * p.ws - 1) * (2 * p.ws - 1) * p.nH * 4, cudaMemcpyHostToDevice, st);
  int rc = swin_attn_forward(shape, dq, dk, dv, dqp, dkp, dvp, dtab, nullptr, dout, nullptr, st);
  if (rc) return rc;
  if ((e = cudaMemcpyAsync(o, dout, tens, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return fail_cuda("D2H o", e);
  if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return fail_cuda("host forward", e);
  return 0;
}

// ---- host-buffer forward that allocates nothing (swin_attn_forward_host_static) -------------------------------------
// The staging scratch of swin_attn_forward_host, but in a file-scope device array instead of a cudaMalloc'd buffer:
// for callers that may not allocate (CUDA graph capture around them, tight serving loops, harnesses that forbid
// cudaMalloc).  Every GPU whose context loads this library holds SWIN_ATTN_STATIC_SCRATCH_BYTES of it until exit.
// In HOST code `g_static_scratch` is only a host-side stand-in for the symbol, never a device address: the address is
// resolved with cudaGetSymbolAddress, once per device, with that device current, and that pointer is what the copies
// and the kernel get.
static __device__ __align__(256) unsigned char g_static_scratch[SWIN_ATTN_STATIC_SCRATCH_BYTES];

static constexpr int kMaxStaticDevices = 64;
static std::mutex g_static_mu;
static void* g_static_ptr[kMaxStaticDevices] = {};   // device address of g_static_scratch on device i (null = not resolved yet)

// Device address of the static scratch on the CURRENT device (resolved on first use per device).  g_static_mu held.
static int static_scratch_address(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail_cuda("swin_attn_forward_host_static: cudaGetDevice", e);
  if (dev < 0 || dev >= kMaxStaticDevices) return fail("swin_attn_forward_host_static: device ordinal out of range");
  if (g_static_ptr[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_static_scratch);
    if (e != cudaSuccess) return fail_cuda("swin_attn_forward_host_static: cudaGetSymbolAddress", e);
    if (ptr == nullptr) return fail("swin_attn_forward_host_static: cudaGetSymbolAddress returned a null address");
    g_static_ptr[dev] = ptr;
  }
  *out = g_static_ptr[dev];
  return 0;
}

int swin_attn_forward_host_static(const swin_attn_shape_t* shape, const void* q, const void* k, const void* v,
                                  const void* q_pad, const void* k_pad, const void* v_pad, const float* bias_table, void* o) {
  Problem p;
  if (int rc = make_problem(shape, p)) return rc;
  if (!q || !k || !v || !bias_table || !o) return fail("q, k, v, bias_table and o must be non-null");
  if (p.q_ps != p.C || p.k_ps != p.C || p.v_ps != p.C || p.o_ps != p.C) return fail("host entry takes dense (B,H,W,C) buffers");
  const size_t es = elem_size(shape->dtype);
  const size_t tens = (size_t)p.B * p.H * p.W * p.C * es;
  const size_t tens_al = (tens + 255) & ~size_t(255);
  const size_t vec = ((size_t)p.C * es + 255) & ~size_t(255);
  const size_t tab_bytes = (size_t)(2 * p.ws - 1) * (2 * p.ws - 1) * p.nH * 4;
  const size_t tab = (tab_bytes + 255) & ~size_t(255);
  const size_t need = 4 * tens_al + 3 * vec + tab;
  if (need > (size_t)SWIN_ATTN_STATIC_SCRATCH_BYTES)   // checked before any CUDA call
    return fail("swin_attn_forward_host_static: this shape needs " + std::to_string(need) + " bytes of device staging, the static "
                "scratch holds SWIN_ATTN_STATIC_SCRATCH_BYTES = " + std::to_string((size_t)SWIN_ATTN_STATIC_SCRATCH_BYTES) +
                " (use swin_attn_forward_host, which allocates)");
  // One array per device, shared by every call: calls are serialised here and the call synchronises before it returns,
  // so two calls of this entry point never overlap on the GPU.
  std::lock_guard<std::mutex> lock(g_static_mu);
  void* scratch = nullptr;
  if (int rc = static_scratch_address(&scratch)) return rc;
  char* base = static_cast<char*>(scratch);
  void* dq = base; void* dk = base + tens_al; void* dv = base + 2 * tens_al; void* dout = base + 3 * tens_al;
  char* vecs = base + 4 * tens_al;
  void* dqp = q_pad ? vecs : nullptr; void* dkp = k_pad ? vecs + vec : nullptr; void* dvp = v_pad ? vecs + 2 * vec : nullptr;
  float* dtab = reinterpret_cast<float*>(vecs + 3 * vec);
  cudaStream_t st = nullptr;
  cudaError_t e;
  // The array keeps the previous call's contents.  q / k / v / pads / table are overwritten in full below; the output
  // staging is cleared, so that no byte of an earlier call's result can pass for this call's.
  if ((e = cudaMemsetAsync(dout, 0, tens_al, st)) != cudaSuccess) return fail_cuda("clear output staging", e);
  if ((e = cudaMemcpyAsync(dq, q, tens, cudaMemcpyHostToDevice, st)) != cudaSuccess) return fail_cuda("H2D q", e);
  if ((e = cudaMemcpyAsync(dk, k, tens, cudaMemcpyHostToDevice, st)) != cudaSuccess) return fail_cuda("H2D k", e);
  if ((e = cudaMemcpyAsync(dv, v, tens, cudaMemcpyHostToDevice, st)) != cudaSuccess) return fail_cuda("H2D v", e);
  if (q_pad && (e = cudaMemcpyAsync(dqp, q_pad, p.C * es, cudaMemcpyHostToDevice, st)) != cudaSuccess) return fail_cuda("H2D q_pad", e);
  if (k_pad && (e = cudaMemcpyAsync(dkp, k_pad, p.C * es, cudaMemcpyHostToDevice, st)) != cudaSuccess) return fail_cuda("H2D k_pad", e);
  if (v_pad && (e = cudaMemcpyAsync(dvp, v_pad, p.C * es, cudaMemcpyHostToDevice, st)) != cudaSuccess) return fail_cuda("H2D v_pad", e);
  if ((e = cudaMemcpyAsync(dtab, bias_table, tab_bytes, cudaMemcpyHostToDevice, st)) != cudaSuccess) return fail_cuda("H2D bias_table", e);
  int rc = swin_attn_forward(shape, dq, dk, dv, dqp, dkp, dvp, dtab, nullptr, dout, nullptr, st);
  if (rc) return rc;
  if ((e = cudaMemcpyAsync(o, dout, tens, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return fail_cuda("D2H o", e);
  if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return fail_cuda("host forward (static scratch)", e);
  return 0;
}

int swin_layernorm_forward(const void* x, int32_t x_dtype, const float* gamma, const float* beta, float eps, int64_t rows, int32_t channels,
                           void* out, int32_t out_dtype, void* stream) {
  if (!x || !gamma || !beta || !out) return fail("layernorm: x, gamma, beta and out must be non-null");
  if (rows < 0 || rows > (int64_t)8 * 0x7fffffff) return fail("layernorm: bad row count");
  if ((out_dtype != SWIN_ATTN_F32 && out_dtype != SWIN_ATTN_BF16) || (x_dtype != SWIN_ATTN_F32 && x_dtype != SWIN_ATTN_BF16))
    return fail("layernorm: x_dtype / out_dtype must be SWIN_ATTN_F32 or SWIN_ATTN_BF16");
  if (!layernorm_supported(channels))
    return fail("layernorm: channels must be one of 128, 192, 256, 384, 512, 768, 1024, 1536, 3072 (got " + std::to_string(channels) + ")");
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(gamma) | reinterpret_cast<uintptr_t>(beta) |
       reinterpret_cast<uintptr_t>(out)) & 7)
    return fail("layernorm: pointers must be 8-byte aligned");
  if (rows == 0) return 0;
  cudaError_t e = launch_layernorm(x, x_dtype, gamma, beta, eps, rows, channels, out, out_dtype, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return fail_cuda("layernorm launch", e);
  return 0;
}

int swin_patch_merge_layernorm_forward(const void* x, int32_t x_dtype, const float* gamma, const float* beta, float eps, int32_t batch,
                                       int32_t height, int32_t width, int32_t channels, void* out, void* stream) {
  if (!x || !gamma || !beta || !out) return fail("patch merge + layernorm: x, gamma, beta and out must be non-null");
  if (batch < 0 || height < 0 || width < 0) return fail("patch merge + layernorm: bad batch / height / width");
  if (x_dtype != SWIN_ATTN_F32 && x_dtype != SWIN_ATTN_BF16) return fail("patch merge + layernorm: x_dtype must be SWIN_ATTN_F32 or SWIN_ATTN_BF16");
  if (!merge_layernorm_supported(channels))
    return fail("patch merge + layernorm: channels must be one of 128, 192, 256, 384, 512, 768 (got " + std::to_string(channels) + ")");
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(gamma) | reinterpret_cast<uintptr_t>(beta) |
       reinterpret_cast<uintptr_t>(out)) & 7)
    return fail("patch merge + layernorm: pointers must be 8-byte aligned");
  if ((int64_t)batch * height * width == 0) return 0;
  cudaError_t e = launch_merge_layernorm(x, x_dtype, gamma, beta, eps, batch, height, width, channels, out, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return fail_cuda("patch merge + layernorm launch", e);
  return 0;
}

static int add_ln_checks(const char* what, int32_t x_dtype, int64_t rows, int32_t channels, const float* gate, int64_t rows_per_sample) {
  if (rows < 0) return fail(std::string(what) + ": bad row count");
  if (x_dtype != SWIN_ATTN_F32 && x_dtype != SWIN_ATTN_BF16) return fail(std::string(what) + ": x_dtype must be SWIN_ATTN_F32 or SWIN_ATTN_BF16");
  if (!layernorm_supported(channels) || channels > 1536)
    return fail(std::string(what) + ": channels must be one of 128, 192, 256, 384, 512, 768, 1024, 1536 (got " + std::to_string(channels) + ")");
  if (gate != nullptr && (rows_per_sample <= 0 || rows % rows_per_sample))
    return fail(std::string(what) + ": with a gate, rows must be a multiple of rows_per_sample > 0");
  return 0;
}

int swin_add_layernorm_forward_gated(const void* x, int32_t x_dtype, const void* delta, const float* gamma, const float* beta,
                                     const float* gate, int64_t rows_per_sample, float eps, int64_t rows, int32_t channels, void* sum_out,
                                     void* y_out, void* stream) {
  if (!x || !delta || !gamma || !beta || !sum_out || !y_out) return fail("add+layernorm: all pointers must be non-null");
  if (int rc = add_ln_checks("add+layernorm", x_dtype, rows, channels, gate, rows_per_sample)) return rc;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(sum_out) | reinterpret_cast<uintptr_t>(gamma) |
       reinterpret_cast<uintptr_t>(beta)) & 7 || (reinterpret_cast<uintptr_t>(delta) | reinterpret_cast<uintptr_t>(y_out)) & 3)
    return fail("add+layernorm: pointers must be 8-byte aligned (bf16 delta / y: 4)");
  if (rows == 0) return 0;
  cudaError_t e = launch_add_layernorm(x, x_dtype, delta, gamma, beta, gate, rows_per_sample, eps, rows, channels, sum_out, y_out,
                                       static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return fail_cuda("add+layernorm launch", e);
  return 0;
}

int swin_add_layernorm_forward(const void* x, int32_t x_dtype, const void* delta, const float* gamma, const float* beta, float eps,
                               int64_t rows, int32_t channels, void* sum_out, void* y_out, void* stream) {
  return swin_add_layernorm_forward_gated(x, x_dtype, delta, gamma, beta, nullptr, 1, eps, rows, channels, sum_out, y_out, stream);
}

int swin_add_layernorm_backward(const void* sum, int32_t x_dtype, const void* dy, const void* d_sum, const float* gamma, const float* gate,
                                int64_t rows_per_sample, float eps, int64_t rows, int32_t channels, void* dx, void* d_delta, float* dgamma,
                                float* dbeta, void* stream) {
  if (!sum || !dy || !d_sum || !gamma || !dx || !d_delta || !dgamma || !dbeta)
    return fail("add+layernorm backward: all pointers (except gate) must be non-null");
  if (int rc = add_ln_checks("add+layernorm backward", x_dtype, rows, channels, gate, rows_per_sample)) return rc;
  if ((reinterpret_cast<uintptr_t>(sum) | reinterpret_cast<uintptr_t>(d_sum) | reinterpret_cast<uintptr_t>(dx) |
       reinterpret_cast<uintptr_t>(gamma)) & 7 || (reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(d_delta)) & 3)
    return fail("add+layernorm backward: pointers must be 8-byte aligned (bf16 dy / d_delta: 4)");
  if (rows == 0) {
    cudaMemsetAsync(dgamma, 0, sizeof(float) * channels, static_cast<cudaStream_t>(stream));
    cudaMemsetAsync(dbeta, 0, sizeof(float) * channels, static_cast<cudaStream_t>(stream));
    return 0;
  }
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
  }
  cudaError_t e = launch_add_layernorm_backward(sum, x_dtype, dy, d_sum, gamma, gate, rows_per_sample, eps, rows, channels, dx, d_delta, dgamma,
                                                dbeta, sms, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return fail_cuda("add+layernorm backward launch", e);
  return 0;
}

static int gelu_entry(const char* what, const void* x, const void* dy, void* out, int64_t n, void* stream) {
  if (!x || !out) return fail(std::string(what) + ": all pointers must be non-null");
  if (n < 0 || n % 8) return fail(std::string(what) + ": the element count must be a non-negative multiple of 8");
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(dy)) & 15)
    return fail(std::string(what) + ": pointers must be 16-byte aligned");
  if (n == 0) return 0;
  cudaError_t e = launch_gelu(x, dy, out, n, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return fail_cuda(what, e);
  return 0;
}

int swin_gelu_forward(const void* x, void* y, int64_t n, void* stream) { return gelu_entry("gelu forward", x, nullptr, y, n, stream); }

int swin_gelu_backward(const void* x, const void* dy, void* dx, int64_t n, void* stream) {
  if (!dy) return fail("gelu backward: all pointers must be non-null");
  return gelu_entry("gelu backward", x, dy, dx, n, stream);
}

int swin_linear_forward(const void* x, const void* weight, const void* bias, void* y, int64_t rows, int32_t in_features,
                        int32_t out_features, int32_t activation, void* stream) {
  if (!x || !weight || !bias || !y) return fail("linear forward: all pointers must be non-null");
  if (activation != SWIN_ACT_NONE && activation != SWIN_ACT_GELU) return fail("linear forward: activation must be SWIN_ACT_NONE or SWIN_ACT_GELU");
  if (rows < 0 || rows >= (1ll << 31) - 128) return fail("linear forward: bad row count");
  if (in_features < 64 || in_features % 64) return fail("linear forward: in_features must be a positive multiple of 64");
  if (out_features < 256 || out_features % 256) return fail("linear forward: out_features must be a positive multiple of 256");
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(weight) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(bias)) & 15)
    return fail("linear forward: x, weight, bias and y must be 16-byte aligned");
  if (rows == 0) return 0;
  if (!tc_linear_supported(rows, in_features, out_features))
    return fail("linear forward: the tcgen05 kernel is unavailable (cuTensorMapEncodeTiled not found in the driver)");
  const char* why = nullptr;
  cudaError_t e = launch_tc_linear(x, weight, bias, nullptr, nullptr, y, rows, in_features, out_features, activation, static_cast<cudaStream_t>(stream), &why);
  if (e != cudaSuccess) return fail_cuda(why ? why : "linear forward launch", e);
  return 0;
}

int swin_linear_gelu_forward_train(const void* x, const void* weight, const void* bias, void* h, void* y, int64_t rows, int32_t in_features,
                                   int32_t out_features, void* stream) {
  if (!x || !weight || !bias || !h || !y) return fail("linear + GELU (training): all pointers must be non-null");
  if (rows < 0 || rows >= (1ll << 31) - 128) return fail("linear + GELU (training): bad row count");
  if (in_features < 64 || in_features % 64) return fail("linear + GELU (training): in_features must be a positive multiple of 64");
  if (out_features < 256 || out_features % 256) return fail("linear + GELU (training): out_features must be a positive multiple of 256");
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(weight) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(bias) |
       reinterpret_cast<uintptr_t>(h)) & 15)
    return fail("linear + GELU (training): x, weight, bias, h and y must be 16-byte aligned");
  if (rows == 0) return 0;
  if (!tc_linear_supported(rows, in_features, out_features))
    return fail("linear + GELU (training): the tcgen05 kernel is unavailable (cuTensorMapEncodeTiled not found in the driver)");
  const char* why = nullptr;
  cudaError_t e = launch_tc_linear(x, weight, bias, nullptr, h, y, rows, in_features, out_features, SWIN_ACT_GELU, static_cast<cudaStream_t>(stream), &why);
  if (e != cudaSuccess) return fail_cuda(why ? why : "linear + GELU (training) launch", e);
  return 0;
}

int swin_linear_residual_forward(const void* x, const void* weight, const void* bias, const void* residual, void* y, int64_t rows,
                                 int32_t in_features, int32_t out_features, void* stream) {
  if (!x || !weight || !bias || !residual || !y) return fail("linear + residual: all pointers must be non-null");
  if (rows < 0 || rows >= (1ll << 31) - 128) return fail("linear + residual: bad row count");
  if (in_features < 64 || in_features % 64) return fail("linear + residual: in_features must be a positive multiple of 64");
  if (out_features < 256 || out_features % 256) return fail("linear + residual: out_features must be a positive multiple of 256");
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(weight) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(bias) |
       reinterpret_cast<uintptr_t>(residual)) & 15)
    return fail("linear + residual: x, weight, bias, residual and y must be 16-byte aligned");
  if (rows == 0) return 0;
  if (!tc_linear_supported(rows, in_features, out_features))
    return fail("linear + residual: the tcgen05 kernel is unavailable (cuTensorMapEncodeTiled not found in the driver)");
  const char* why = nullptr;
  cudaError_t e = launch_tc_linear(x, weight, bias, residual, nullptr, y, rows, in_features, out_features, SWIN_ACT_NONE, static_cast<cudaStream_t>(stream), &why);
  if (e != cudaSuccess) return fail_cuda(why ? why : "linear + residual launch", e);
  return 0;
}

int swin_linear_bias_grad(const void* dy, int64_t rows, int32_t cols, float* dbias, void* stream) {
  if (!dy || !dbias) return fail("linear bias grad: all pointers must be non-null");
  if (rows < 0 || cols <= 0 || cols % 8) return fail("linear bias grad: rows >= 0 and cols a positive multiple of 8");
  if (reinterpret_cast<uintptr_t>(dy) & 15 || reinterpret_cast<uintptr_t>(dbias) & 3)
    return fail("linear bias grad: dy must be 16-byte aligned, dbias 4-byte aligned");
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
  }
  cudaError_t e = launch_colsum(dy, rows, cols, dbias, sms, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return fail_cuda("linear bias grad launch", e);
  return 0;
}

int swin_layernorm_backward(const void* x, int32_t x_dtype, const void* dy, int32_t dy_dtype, const float* gamma, float eps, int64_t rows,
                            int32_t channels, void* dx, float* dgamma, float* dbeta, void* stream) {
  if (!x || !dy || !gamma || !dx || !dgamma || !dbeta) return fail("layernorm backward: all pointers must be non-null");
  if (rows < 0) return fail("layernorm backward: bad row count");
  if (x_dtype != SWIN_ATTN_F32 && x_dtype != SWIN_ATTN_BF16) return fail("layernorm backward: x_dtype must be SWIN_ATTN_F32 or SWIN_ATTN_BF16");
  if (dy_dtype != SWIN_ATTN_F32 && dy_dtype != SWIN_ATTN_BF16) return fail("layernorm backward: dy_dtype must be SWIN_ATTN_F32 or SWIN_ATTN_BF16");
  if (!layernorm_supported(channels) || channels > 1536)
    return fail("layernorm backward: channels must be one of 128, 192, 256, 384, 512, 768, 1024, 1536 (got " + std::to_string(channels) + ")");
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dx) | reinterpret_cast<uintptr_t>(gamma)) & 7 ||
      (reinterpret_cast<uintptr_t>(dy) & (dy_dtype == SWIN_ATTN_F32 ? 7 : 3)))
    return fail("layernorm backward: pointers must be 8-byte aligned (bf16 dy: 4)");
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
  }
  cudaError_t e = launch_layernorm_backward(x, x_dtype, dy, dy_dtype, gamma, eps, rows, channels, dx, dgamma, dbeta, sms,
                                            static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return fail_cuda("layernorm backward launch", e);
  return 0;
}

// Debug hook (not part of the public header): device buffer of >= 16 * pairs-per-CTA int64 that receives
// clock64 stamps of CTA 0 of the tcgen05 kernel; pass NULL to switch tracing off.
static int mha_impl(const mha_shape_t* s, const void* q, const void* k, const void* v, void* o, float* attn_map, const int32_t* tok_idx,
                    const float* idx_bias, int n_emb, void* stream) {
  if (s == nullptr) return fail("mha forward: shape is null");
  if (!q || !k || !v || !o) return fail("mha forward: q, k, v and o must be non-null");
  if (s->dtype != SWIN_ATTN_F32 && s->dtype != SWIN_ATTN_BF16) return fail("mha forward: dtype must be SWIN_ATTN_F32 or SWIN_ATTN_BF16");
  if (s->batch < 0 || s->seq < 0 || s->kv_seq < 0 || s->num_heads <= 0) return fail("mha forward: batch / seq / kv_seq must be >= 0 and num_heads positive");
  if (s->qk_head_dim <= 0 || s->qk_head_dim > 64 || s->v_head_dim <= 0 || s->v_head_dim > 64)
    return fail("mha forward: head dims must be in 1..64 (got " + std::to_string(s->qk_head_dim) + ", " + std::to_string(s->v_head_dim) + ")");
  if (s->q_row_stride < (int64_t)s->num_heads * s->qk_head_dim || s->k_row_stride < (int64_t)s->num_heads * s->qk_head_dim ||
      s->v_row_stride < (int64_t)s->num_heads * s->v_head_dim || s->o_row_stride < (int64_t)s->num_heads * s->v_head_dim)
    return fail("mha forward: row strides must cover num_heads * head_dim elements");
  const int kv_seq = s->kv_seq > 0 ? s->kv_seq : s->seq;       // 0: self-attention, keys = queries' tokens
  if (s->batch == 0 || s->seq == 0) return 0;
  if (s->qk_head_dim % 4) return fail("mha forward: qk_head_dim must be a multiple of 4");
  if (s->batch > 65535 || s->num_heads > 65535) return fail("mha forward: batch and num_heads must be <= 65535");
  if (s->impl != SWIN_ATTN_IMPL_AUTO && s->impl != SWIN_ATTN_IMPL_SIMT) return fail("mha forward: impl must be SWIN_ATTN_IMPL_AUTO or SWIN_ATTN_IMPL_SIMT");
  if (tok_idx != nullptr) {
    if (idx_bias == nullptr || n_emb < 1) return fail("mha forward (ordered bias): index_bias must be non-null and num_emb positive");
    if (kv_seq != s->seq) return fail("mha forward (ordered bias): self-attention only (kv_seq must equal seq)");
  }
  if (s->impl == SWIN_ATTN_IMPL_AUTO && (tok_idx == nullptr || n_emb <= 256) &&
      mha_tc_supported(s->dtype, s->seq, kv_seq, s->qk_head_dim, s->v_head_dim, s->q_row_stride, s->k_row_stride, s->v_row_stride,
                       s->o_row_stride, q, k, v, o)) {
    const char* why = nullptr;
    cudaError_t e = launch_mha_tc_forward(s->batch, s->seq, kv_seq, s->num_heads, s->q_row_stride, s->k_row_stride, s->v_row_stride,
                                          s->o_row_stride, s->scale, q, k, v, o, attn_map, tok_idx, idx_bias, n_emb,
                                          static_cast<cudaStream_t>(stream), &why);
    if (e != cudaSuccess) return fail_cuda(why ? why : "mha forward (tcgen05) launch", e);
    return 0;
  }
  const size_t smem = mha_forward_smem_bytes(kv_seq, s->qk_head_dim, tok_idx != nullptr ? n_emb : 0);
  if (smem > 227u * 1024u)
    return fail("mha forward: " + std::to_string(kv_seq) + " keys need " + std::to_string(smem) + " B of shared memory per CTA (limit 232448; kv_seq <= 1472)");
  cudaError_t e = launch_mha_forward(s->dtype, s->batch, s->seq, kv_seq, s->num_heads, s->qk_head_dim, s->v_head_dim, s->q_row_stride,
                                     s->k_row_stride, s->v_row_stride, s->o_row_stride, s->scale, q, k, v, o, attn_map, tok_idx, idx_bias,
                                     n_emb, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return fail_cuda("mha forward launch", e);
  return 0;
}

int mha_attention_forward(const mha_shape_t* s, const void* q, const void* k, const void* v, void* o, float* attn_map, void* stream) {
  return mha_impl(s, q, k, v, o, attn_map, nullptr, nullptr, 0, stream);
}

int mha_attention_forward_ordered(const mha_shape_t* s, const void* q, const void* k, const void* v, const int32_t* token_index,
                                  const float* index_bias, int32_t num_emb, void* o, float* attn_map, void* stream) {
  if (token_index == nullptr) return fail("mha forward (ordered bias): token_index must be non-null");
  return mha_impl(s, q, k, v, o, attn_map, token_index, index_bias, num_emb, stream);
}

// Debug hook (not part of the public header): CTA-pair mode of the tcgen05 linear kernel, -1 automatic, 0 off, 1 on.
void swin_linear_debug_set_pair_mode(int mode) { set_tc_linear_pair_mode(mode); }

void swin_attn_debug_set_trace(void* device_buffer) { g_trace = static_cast<long long*>(device_buffer); }

// Measurement probe (bench.py): what the event bracket of swin_attn_time_next_launch costs around an EMPTY kernel that
// follows other work in the stream -- event record, launch latency of a 148 x 384 grid, completion, event record --
// optionally with the backward's memset in front.  Median of `reps` brackets in microseconds, < 0 on a CUDA error.
// The attention launches' event-timed durations carry this fixed cost; it is reported next to the roofline.
namespace {
__global__ void probe_empty_kernel(int* sink) {
  if (sink != nullptr && threadIdx.x == 1u << 20) sink[0] = 1;
}
__global__ void probe_busy_kernel(float* p, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = 1.f;
}
}  // namespace

double swin_attn_probe_empty_bracket_us(void* stream, int with_memset, int reps) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (reps < 1) reps = 1;
  float* busy = nullptr;
  const size_t n_busy = (size_t)8 << 20;
  cudaEvent_t a = nullptr, b = nullptr;
  if (cudaMalloc(&busy, n_busy * sizeof(float)) != cudaSuccess || cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) {
    cudaGetLastError();
    cudaFree(busy);
    return -1.0;
  }
  std::vector<float> us;
  bool ok = true;
  for (int i = 0; i < reps + 2 && ok; ++i) {
    probe_busy_kernel<<<148 * 4, 256, 0, st>>>(busy, n_busy);   // the bracket follows a running kernel, as in a model step
    cudaEventRecord(a, st);
    if (with_memset) cudaMemsetAsync(busy, 0, 20480, st);
    probe_empty_kernel<<<148, 384, 0, st>>>(nullptr);
    cudaEventRecord(b, st);
    ok = cudaStreamSynchronize(st) == cudaSuccess;
    float ms = 0.f;
    if (ok && cudaEventElapsedTime(&ms, a, b) == cudaSuccess && i >= 2) us.push_back(ms * 1e3f);
  }
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  cudaFree(busy);
  if (!ok || us.empty()) {
    cudaGetLastError();
    return -1.0;
  }
  std::sort(us.begin(), us.end());
  return us[us.size() / 2];
}

void swin_attn_time_next_launch(void* start_event, void* stop_event) {
  g_launch_events.start = static_cast<cudaEvent_t>(start_event);
  g_launch_events.stop = static_cast<cudaEvent_t>(stop_event);
}

uint64_t swin_attn_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int swin_attn_uses_tcgen05(const swin_attn_shape_t* shape, int backward) {
  Problem p;
  if (make_problem(shape, p)) return 0;
  if (shape->impl == SWIN_ATTN_IMPL_SIMT || shape->impl == SWIN_ATTN_IMPL_WARP_MMA) return 0;
  if (backward) return tc_backward_supported(p, shape->dtype) ? 1 : 0;
  if (shape->impl == SWIN_ATTN_IMPL_AUTO && mma12_forward_supported(p, shape->dtype)) return 0;   // window 12: warp-MMA forward
  return tc_forward_supported(p, shape->dtype) || tc12_forward_supported(p, shape->dtype) ? 1 : 0;
}

const char* swin_attn_last_error(void) { return g_error.c_str(); }
const char* swin_attn_version(void) { return "swin_attn 0.1 (sm_100a)"; }

}  // extern "C"
