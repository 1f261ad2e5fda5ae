// torch.library registration of the C ABI (include/swin_attn.h) as dispatcher ops with C++ kernels:
//   swin_attn::fwd / swin_attn::bwd / swin_attn::layer_norm / swin_attn::layer_norm_bwd
// CUDA implementations (argument checks, layout normalisation, output allocation, one C-ABI call on the current
// stream), Meta implementations (shape inference for FakeTensor / torch.compile / DDP), and Autograd-key
// implementations (torch::autograd::Function) so that a training step pays a few microseconds of host time per op
// instead of the ~60-300 us of Python-defined custom ops -- the step was host-bound with those.
//
// Nothing here computes: every number comes from libswin_attn.so.  No CPU kernels are registered on purpose.
#include <ATen/ATen.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <cuda_runtime.h>
#include <torch/csrc/autograd/autograd_not_implemented_fallback.h>
#include <torch/csrc/autograd/custom_function.h>
#include <torch/library.h>

#include <mutex>
#include <vector>

#include "../../include/swin_attn.h"

namespace {

using at::Tensor;
using c10::optional;
using torch::autograd::AutogradContext;
using torch::autograd::variable_list;

// ---- optional per-launch timing (bench.py's live roofline) --------------------------------------------------
struct TimingRecord {
  cudaEvent_t start, stop;
  int64_t bytes, flops;
  int64_t tag[6];   // kind (0 fwd, 1 bwd), window, shift, heads, height, width
};
std::mutex g_timing_mutex;
// 0 off; 1 eager: every launch is bracketed, records are consumed by swin_torch_take_timing; 2 graph: only launches that
// are being CAPTURED are bracketed (the library records the events as external event-record nodes of the graph), the
// records persist with the graph and are read after every replay by swin_torch_read_graph_timing.
int g_timing_on = 0;
std::vector<TimingRecord> g_records, g_graph_records;

inline bool stream_is_capturing(cudaStream_t st) {
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  return cudaStreamIsCapturing(st, &cs) == cudaSuccess && cs == cudaStreamCaptureStatusActive;
}

struct TimedLaunch {
  bool on, graph;
  TimingRecord rec;
  cudaStream_t stream;
  explicit TimedLaunch(cudaStream_t st) : on(g_timing_on != 0), graph(g_timing_on == 2), stream(st) {
    if (graph && !stream_is_capturing(st)) on = false;
    if (on) {
      cudaEventCreate(&rec.start);
      cudaEventCreate(&rec.stop);
      // recorded by the library right around its device work (after the tensor maps are encoded)
      swin_attn_time_next_launch(rec.start, rec.stop);
    }
  }
  void finish(int64_t bytes, int64_t flops, int64_t kind, int64_t window, int64_t shift, int64_t heads, int64_t h, int64_t w) {
    if (!on) return;
    swin_attn_time_next_launch(nullptr, nullptr);
    rec.bytes = bytes;
    rec.flops = flops;
    const int64_t tag[6] = {kind, window, shift, heads, h, w};
    for (int i = 0; i < 6; ++i) rec.tag[i] = tag[i];
    std::lock_guard<std::mutex> lock(g_timing_mutex);
    (graph ? g_graph_records : g_records).push_back(rec);
  }
};

inline void check_rc(int rc) { TORCH_CHECK(rc == 0, "swin_attn: ", swin_attn_last_error()); }

inline int32_t dtype_code(at::ScalarType t) {
  TORCH_CHECK(t == at::kFloat || t == at::kBFloat16, "swin_attn: dtype ", t, " is not supported (float32 or bfloat16)");
  return t == at::kFloat ? SWIN_ATTN_F32 : SWIN_ATTN_BF16;
}

// Dense rows / images with unit channel stride (the pixel stride may exceed C: slices of a packed projection).
inline Tensor image_layout(const Tensor& t) {
  const auto s = t.strides();
  const auto d = t.sizes();
  if (s[3] == 1 && s[1] == d[2] * s[2] && s[0] == d[1] * s[1] && s[2] % 4 == 0 && s[2] >= d[3]) return t;
  return t.contiguous();
}

inline int64_t split_channels(const Tensor& packed, const optional<Tensor>& v, int64_t num_heads) {
  const int64_t c = num_heads * 32;
  const int64_t want = v.has_value() ? 2 * c : 3 * c;
  TORCH_CHECK(packed.dim() == 4 && packed.size(3) == want, "swin_attn: packed projection must be (B,H,W,", want, ") for ",
              num_heads, " heads of dim 32, got ", packed.sizes());
  if (v.has_value())   // reference assert, newcrf_layers.py:143
    TORCH_CHECK(v->dim() == 4 && v->size(0) == packed.size(0) && v->size(1) == packed.size(1) && v->size(2) == packed.size(2) &&
                    v->size(3) == c, "self.dim != v.shape[-1]");
  return c;
}

inline Tensor mask_arg(const optional<Tensor>& attn_mask, int64_t window, int32_t* n_windows) {
  *n_windows = 0;
  if (!attn_mask.has_value()) return Tensor();
  const int64_t n = window * window;
  TORCH_CHECK(attn_mask->dim() == 3 && attn_mask->size(1) == n && attn_mask->size(2) == n, "swin_attn: attn_mask must be (nW, ", n,
              ", ", n, "), got ", attn_mask->sizes());
  *n_windows = (int32_t)attn_mask->size(0);
  return attn_mask->detach().to(at::kFloat).contiguous();
}

inline swin_attn_shape_t make_shape(int64_t b, int64_t h, int64_t w, int64_t heads, int64_t window, int64_t shift, at::ScalarType dt,
                                    int64_t impl, double scale, int64_t q_ps, int64_t k_ps, int64_t v_ps, int64_t o_ps,
                                    int32_t mask_windows, double dropout_p, int64_t seed) {
  swin_attn_shape_t s;
  memset(&s, 0, sizeof(s));
  s.batch = (int32_t)b; s.height = (int32_t)h; s.width = (int32_t)w; s.num_heads = (int32_t)heads;
  s.window = (int32_t)window; s.shift = (int32_t)shift; s.dtype = dtype_code(dt); s.impl = (int32_t)impl;
  s.scale = (float)scale; s.dropout_p = (float)dropout_p; s.seed = (uint64_t)seed;
  s.q_pix_stride = q_ps; s.k_pix_stride = k_ps; s.v_pix_stride = v_ps; s.o_pix_stride = o_ps;
  s.mask_windows = mask_windows;
  return s;
}

inline const char* byte_ptr(const Tensor& t) { return static_cast<const char*>(t.data_ptr()); }

// ---- swin_attn::fwd ------------------------------------------------------------------------------------------
std::tuple<Tensor, Tensor> fwd_cuda(const Tensor& packed_in, const optional<Tensor>& v_in, const Tensor& bias_table,
                                    const optional<Tensor>& packed_pad_in, const optional<Tensor>& v_pad_in, int64_t window,
                                    int64_t shift, int64_t num_heads, double scale, bool need_lse, int64_t impl,
                                    const optional<Tensor>& attn_mask, double dropout_p, int64_t seed) {
  TORCH_CHECK(window > 0, "swin_attn: window must be positive");
  const int64_t c = split_channels(packed_in, v_in, num_heads);
  int32_t mask_nw = 0;
  const Tensor mask = mask_arg(attn_mask, window, &mask_nw);
  const Tensor packed = image_layout(packed_in);
  const auto dt = packed.scalar_type();
  dtype_code(dt);
  const int64_t b = packed.size(0), h = packed.size(1), w = packed.size(2);
  const int64_t ps = packed.stride(2), es = packed.element_size();
  const char* base = byte_ptr(packed);
  Tensor v, packed_pad, v_pad;
  const char *v_ptr, *qpad = nullptr, *kpad = nullptr, *vpad = nullptr;
  int64_t v_ps;
  if (packed_pad_in.has_value()) {
    packed_pad = packed_pad_in->to(dt).contiguous();
    qpad = byte_ptr(packed_pad);
    kpad = qpad + c * es;
  }
  if (!v_in.has_value()) {
    v_ptr = base + 2 * c * es;
    v_ps = ps;
    if (qpad) vpad = qpad + 2 * c * es;
  } else {
    v = image_layout(v_in->to(dt));
    v_ptr = byte_ptr(v);
    v_ps = v.stride(2);
    if (v_pad_in.has_value()) {
      v_pad = v_pad_in->to(dt).contiguous();
      vpad = byte_ptr(v_pad);
    }
  }
  const Tensor table = bias_table.detach().to(at::kFloat).contiguous();
  const int64_t hp = (h + window - 1) / window * window, wp = (w + window - 1) / window * window;
  Tensor o = at::empty({b, h, w, c}, packed.options());
  Tensor lse = need_lse ? at::empty({b, hp, wp, num_heads}, packed.options().dtype(at::kFloat))
                        : at::empty({0}, packed.options().dtype(at::kFloat));
  if (o.numel() == 0) return {o, lse};   // empty batch: nothing to launch (the reference returns empty tensors too)
  const swin_attn_shape_t shp = make_shape(b, h, w, num_heads, window, shift, dt, impl, scale, ps, ps, v_ps, c, mask_nw, dropout_p, seed);
  c10::cuda::CUDAGuard guard(packed.device());
  cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  TimedLaunch timed(stream);
  check_rc(swin_attn_forward(&shp, base, base + c * es, v_ptr, qpad, kpad, vpad, table.data_ptr<float>(),
                             mask.defined() ? mask.data_ptr<float>() : nullptr, o.data_ptr(),
                             need_lse ? lse.data_ptr<float>() : nullptr, stream));
  // algorithmic traffic: read q, k, v and write o once, valid tokens only
  timed.finish(4 * b * h * w * c * es + (need_lse ? lse.numel() * 4 : 0), 4 * b * hp * wp * window * window * c, 0, window, shift,
               num_heads, h, w);
  return {o, lse};
}

std::tuple<Tensor, Tensor> fwd_meta(const Tensor& packed, const optional<Tensor>& v, const Tensor&, const optional<Tensor>&,
                                    const optional<Tensor>&, int64_t window, int64_t, int64_t num_heads, double, bool need_lse,
                                    int64_t, const optional<Tensor>&, double, int64_t) {
  const auto b = packed.sym_size(0), h = packed.sym_size(1), w = packed.sym_size(2);
  const c10::SymInt c = num_heads * 32;
  const auto hp = (h + (window - 1)) / window * window, wp = (w + (window - 1)) / window * window;
  Tensor o = at::empty_symint({b, h, w, c}, packed.options());
  Tensor lse = need_lse ? at::empty_symint({b, hp, wp, c10::SymInt(num_heads)}, packed.options().dtype(at::kFloat))
                        : at::empty({0}, packed.options().dtype(at::kFloat));
  return {o, lse};
}

// ---- swin_attn::bwd ------------------------------------------------------------------------------------------
std::tuple<Tensor, Tensor, Tensor> bwd_cuda(const Tensor& d_o_in, const Tensor& packed_in, const optional<Tensor>& v_in,
                                            const Tensor& o_in, const Tensor& lse, const Tensor& bias_table, int64_t window,
                                            int64_t shift, int64_t num_heads, double scale, const optional<Tensor>& attn_mask,
                                            int64_t impl, double dropout_p, int64_t seed) {
  const int64_t c = split_channels(packed_in, v_in, num_heads);
  int32_t mask_nw = 0;
  const Tensor mask = mask_arg(attn_mask, window, &mask_nw);
  Tensor packed = image_layout(packed_in);
  const auto dt = packed.scalar_type();
  dtype_code(dt);
  const int64_t b = packed.size(0), h = packed.size(1), w = packed.size(2), pc = packed.size(3);
  const int64_t es = packed.element_size();
  Tensor v;
  if (v_in.has_value()) v = image_layout(v_in->to(dt));
  // Gradients are written with the strides of d_packed / d_v (dense): the inputs must be dense as well, because one
  // shape struct carries one stride per tensor role.
  if (packed.stride(2) != pc) packed = packed.contiguous();
  if (v.defined() && v.stride(2) != c) v = v.contiguous();
  const int64_t ps = packed.stride(2);
  const Tensor d_o = d_o_in.to(dt).contiguous();
  const Tensor o = o_in.contiguous();
  Tensor d_packed = at::empty({b, h, w, pc}, packed.options());
  Tensor d_v;
  const char* v_ptr;
  char* dv_ptr;
  int64_t v_ps;
  char* dbase = static_cast<char*>(d_packed.data_ptr());
  const char* base = byte_ptr(packed);
  if (!v.defined()) {
    v_ptr = base + 2 * c * es;
    v_ps = ps;
    d_v = at::empty({0}, packed.options());
    dv_ptr = dbase + 2 * c * es;
  } else {
    v_ptr = byte_ptr(v);
    v_ps = v.stride(2);
    d_v = at::empty({b, h, w, c}, packed.options());
    dv_ptr = static_cast<char*>(d_v.data_ptr());
  }
  const Tensor table = bias_table.detach().to(at::kFloat).contiguous();
  Tensor d_table = at::empty_like(table);
  if (d_packed.numel() == 0) return {d_packed, d_v, at::zeros_like(bias_table)};   // empty batch
  const swin_attn_shape_t shp = make_shape(b, h, w, num_heads, window, shift, dt, impl, scale, ps, ps, v_ps, c, mask_nw, dropout_p, seed);
  c10::cuda::CUDAGuard guard(packed.device());
  cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  TimedLaunch timed(stream);
  check_rc(swin_attn_backward(&shp, base, base + c * es, v_ptr, o.data_ptr(), d_o.data_ptr(), lse.data_ptr<float>(),
                              table.data_ptr<float>(), mask.defined() ? mask.data_ptr<float>() : nullptr, dbase, dbase + c * es,
                              dv_ptr, d_table.data_ptr<float>(), stream));
  // algorithmic traffic of the recompute backward: read q, k, v, dO + lse, write dq, dk, dv
  timed.finish(7 * b * h * w * c * es + lse.numel() * 4, 10 * b * h * w * window * window * c, 1, window, shift, num_heads, h, w);
  return {d_packed, d_v, d_table.to(bias_table.scalar_type())};
}

std::tuple<Tensor, Tensor, Tensor> bwd_meta(const Tensor&, const Tensor& packed, const optional<Tensor>& v, const Tensor&,
                                            const Tensor&, const Tensor& bias_table, int64_t, int64_t, int64_t, double,
                                            const optional<Tensor>&, int64_t, double, int64_t) {
  Tensor d_packed = at::empty_symint(packed.sym_sizes(), packed.options());
  Tensor d_v = v.has_value() ? at::empty_symint(v->sym_sizes(), packed.options()) : at::empty({0}, packed.options());
  return {d_packed, d_v, at::empty_like(bias_table)};
}

// ---- swin_attn::bwd_unpadded: backward on un-padded tensors, pad vectors as differentiable inputs ------------------
std::tuple<Tensor, Tensor, Tensor, Tensor, Tensor> bwd_unpadded_cuda(const Tensor& d_o_in, const Tensor& packed_in, const optional<Tensor>& v_in,
                                                                     const Tensor& o_in, const Tensor& lse, const Tensor& bias_table,
                                                                     const optional<Tensor>& packed_pad_in, const optional<Tensor>& v_pad_in,
                                                                     int64_t window, int64_t shift, int64_t num_heads, double scale) {
  TORCH_CHECK(window > 0, "swin_attn: window must be positive");
  const int64_t c = split_channels(packed_in, v_in, num_heads);
  Tensor packed = image_layout(packed_in);
  const auto dt = packed.scalar_type();
  dtype_code(dt);
  const int64_t b = packed.size(0), h = packed.size(1), w = packed.size(2), pc = packed.size(3);
  const int64_t es = packed.element_size();
  Tensor v;
  if (v_in.has_value()) v = image_layout(v_in->to(dt));
  if (packed.stride(2) != pc) packed = packed.contiguous();
  if (v.defined() && v.stride(2) != c) v = v.contiguous();
  const Tensor d_o = d_o_in.to(dt).contiguous();
  const Tensor o = o_in.contiguous();
  Tensor d_packed = at::empty({b, h, w, pc}, packed.options());
  Tensor d_v = v.defined() ? at::empty({b, h, w, c}, packed.options()) : at::empty({0}, packed.options());
  const Tensor table = bias_table.detach().to(at::kFloat).contiguous();
  // The small fp32 gradients live in ONE buffer, q_pad | k_pad | v_pad | bias table: the library clears buffers that
  // touch with a single memset (k_pad | v_pad | table; swin_attn_tc_bwd.cu), and the q_pad part -- no gradient flows into
  // the padding QUERIES, their output rows are cropped -- is cleared here.
  // pad vectors: k_pad = packed_pad[C:2C]; v_pad = packed_pad[2C:3C] (fused qkv) or the separate v_pad (external v)
  Tensor packed_pad, v_pad, d_packed_pad, d_v_pad;
  const char *kpad = nullptr, *vpad = nullptr;
  float *d_kpad = nullptr, *d_vpad = nullptr;
  const auto fopt = packed.options().dtype(at::kFloat);
  const bool has_pp = packed_pad_in.has_value(), has_vp = v.defined() && v_pad_in.has_value();
  if (has_pp) TORCH_CHECK(packed_pad_in->numel() == pc, "swin_attn: packed_pad must have ", pc, " elements");
  if (has_vp) TORCH_CHECK(v_pad_in->numel() == c, "swin_attn: v_pad must have ", c, " elements");
  const int64_t n_pp = has_pp ? pc : 0, n_vp = has_vp ? c : 0;
  Tensor small = at::empty({n_pp + n_vp + table.numel()}, fopt);
  Tensor d_table = small.narrow(0, n_pp + n_vp, table.numel()).view(table.sizes());
  if (has_pp) {
    packed_pad = packed_pad_in->detach().to(dt).contiguous();
    d_packed_pad = small.narrow(0, 0, pc);
    d_packed_pad.narrow(0, 0, c).zero_();
    kpad = byte_ptr(packed_pad) + c * es;
    d_kpad = d_packed_pad.data_ptr<float>() + c;
    if (!v.defined()) {
      vpad = byte_ptr(packed_pad) + 2 * c * es;
      d_vpad = d_packed_pad.data_ptr<float>() + 2 * c;
    }
  } else {
    d_packed_pad = at::empty({0}, fopt);
  }
  if (has_vp) {
    v_pad = v_pad_in->detach().to(dt).contiguous();
    d_v_pad = small.narrow(0, n_pp, c);
    vpad = byte_ptr(v_pad);
    d_vpad = d_v_pad.data_ptr<float>();
  } else {
    d_v_pad = at::empty({0}, fopt);
  }
  if (d_packed.numel() == 0) return {d_packed, d_v, at::zeros_like(bias_table), d_packed_pad, d_v_pad};
  const int64_t ps = packed.stride(2);
  const char* base = byte_ptr(packed);
  char* dbase = static_cast<char*>(d_packed.data_ptr());
  const char* v_ptr = v.defined() ? byte_ptr(v) : base + 2 * c * es;
  char* dv_ptr = v.defined() ? static_cast<char*>(d_v.data_ptr()) : dbase + 2 * c * es;
  const int64_t v_ps = v.defined() ? v.stride(2) : ps;
  const swin_attn_shape_t shp = make_shape(b, h, w, num_heads, window, shift, dt, SWIN_ATTN_IMPL_AUTO, scale, ps, ps, v_ps, c, 0, 0.0, 0);
  c10::cuda::CUDAGuard guard(packed.device());
  cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  TimedLaunch timed(stream);
  check_rc(swin_attn_backward_unpadded(&shp, base, base + c * es, v_ptr, kpad, vpad, o.data_ptr(), d_o.data_ptr(), lse.data_ptr<float>(),
                                       table.data_ptr<float>(), dbase, dbase + c * es, dv_ptr, d_table.data_ptr<float>(), d_kpad, d_vpad,
                                       stream));
  timed.finish(7 * b * h * w * c * es + lse.numel() * 4, 10 * b * lse.size(1) * lse.size(2) * window * window * c, 1, window, shift, num_heads, h, w);
  return {d_packed, d_v, d_table.to(bias_table.scalar_type()), d_packed_pad, d_v_pad};
}

std::tuple<Tensor, Tensor, Tensor, Tensor, Tensor> bwd_unpadded_meta(const Tensor&, const Tensor& packed, const optional<Tensor>& v, const Tensor&,
                                                                     const Tensor&, const Tensor& bias_table,
                                                                     const optional<Tensor>& packed_pad, const optional<Tensor>& v_pad, int64_t,
                                                                     int64_t, int64_t, double) {
  const auto fopt = packed.options().dtype(at::kFloat);
  Tensor d_packed = at::empty_symint(packed.sym_sizes(), packed.options());
  Tensor d_v = v.has_value() ? at::empty_symint(v->sym_sizes(), packed.options()) : at::empty({0}, packed.options());
  Tensor dpp = packed_pad.has_value() ? at::empty_symint({packed.sym_size(3)}, fopt) : at::empty({0}, fopt);
  Tensor dvp = v.has_value() && v_pad.has_value() ? at::empty_symint({v->sym_size(3)}, fopt) : at::empty({0}, fopt);
  return {d_packed, d_v, at::empty_like(bias_table), dpp, dvp};
}

// ---- swin_attn::layer_norm / layer_norm_bwd ------------------------------------------------------------------
Tensor layer_norm_cuda(const Tensor& x_in, const Tensor& weight, const Tensor& bias, double eps, bool out_bf16) {
  const int32_t xd = dtype_code(x_in.scalar_type());
  const Tensor x = x_in.contiguous();
  const int64_t c = x.size(-1);
  Tensor out = at::empty(x.sizes(), x.options().dtype(out_bf16 ? at::kBFloat16 : at::kFloat));
  const Tensor w = weight.detach().to(at::kFloat).contiguous(), b = bias.detach().to(at::kFloat).contiguous();
  if (x.numel() == 0) return out;
  c10::cuda::CUDAGuard guard(x.device());
  check_rc(swin_layernorm_forward(x.data_ptr(), xd, w.data_ptr<float>(), b.data_ptr<float>(), (float)eps, c ? x.numel() / c : 0,
                                  (int32_t)c, out.data_ptr(), out_bf16 ? SWIN_ATTN_BF16 : SWIN_ATTN_F32,
                                  at::cuda::getCurrentCUDAStream()));
  return out;
}

Tensor layer_norm_meta(const Tensor& x, const Tensor&, const Tensor&, double, bool out_bf16) {
  return at::empty_symint(x.sym_sizes(), x.options().dtype(out_bf16 ? at::kBFloat16 : at::kFloat));
}

// ---- swin_attn::patch_merge_norm (inference) ------------------------------------------------------------------
// x: (B, H*W, C) or (B, H, W, C) stage output; returns bf16 (B, ceil(H/2)*ceil(W/2), 4C) = norm(cat of the 2x2 neighbourhood)
Tensor patch_merge_norm_cuda(const Tensor& x_in, int64_t height, int64_t width, const Tensor& weight, const Tensor& bias, double eps) {
  const int32_t xd = dtype_code(x_in.scalar_type());
  const Tensor x = x_in.contiguous();
  const int64_t c = x.size(-1), b = x.size(0);
  TORCH_CHECK(height >= 0 && width >= 0 && x.numel() == b * height * width * c, "swin_attn::patch_merge_norm: input feature has wrong size");
  TORCH_CHECK(weight.numel() == 4 * c && bias.numel() == 4 * c, "swin_attn::patch_merge_norm: the norm must have 4C channels");
  const int64_t h2 = (height + 1) / 2, w2 = (width + 1) / 2;
  Tensor out = at::empty({b, h2 * w2, 4 * c}, x.options().dtype(at::kBFloat16));
  const Tensor w = weight.detach().to(at::kFloat).contiguous(), bb = bias.detach().to(at::kFloat).contiguous();
  if (out.numel() == 0) return out;
  c10::cuda::CUDAGuard guard(x.device());
  check_rc(swin_patch_merge_layernorm_forward(x.data_ptr(), xd, w.data_ptr<float>(), bb.data_ptr<float>(), (float)eps, (int32_t)b,
                                              (int32_t)height, (int32_t)width, (int32_t)c, out.data_ptr(),
                                              at::cuda::getCurrentCUDAStream()));
  return out;
}

Tensor patch_merge_norm_meta(const Tensor& x, int64_t height, int64_t width, const Tensor&, const Tensor&, double) {
  const auto b = x.sym_size(0), c = x.sym_size(-1);
  return at::empty_symint({b, c10::SymInt(((height + 1) / 2) * ((width + 1) / 2)), c * 4}, x.options().dtype(at::kBFloat16));
}

// gate: per-sample stochastic-depth scale (B, 1, .., 1) fp32 or None; rows_per_sample = rows / B
static const float* gate_ptr(const optional<Tensor>& scale, const Tensor& x, Tensor& holder, int64_t& rows_per_sample) {
  rows_per_sample = 1;
  if (!scale.has_value() || !scale->defined()) return nullptr;
  TORCH_CHECK(x.dim() >= 2 && scale->numel() == x.size(0), "swin_attn::add_layer_norm: scale must hold one value per sample");
  holder = scale->detach().to(at::kFloat).contiguous();
  rows_per_sample = x.numel() / x.size(-1) / x.size(0);
  return holder.data_ptr<float>();
}

std::tuple<Tensor, Tensor> add_layer_norm_cuda(const Tensor& x_in, const Tensor& delta_in, const Tensor& weight, const Tensor& bias,
                                               double eps, const optional<Tensor>& scale) {
  const int32_t xd = dtype_code(x_in.scalar_type());
  TORCH_CHECK(delta_in.scalar_type() == at::kBFloat16, "swin_attn::add_layer_norm: delta must be bfloat16");
  TORCH_CHECK(delta_in.sizes() == x_in.sizes(), "swin_attn::add_layer_norm: x and delta must have the same shape");
  const Tensor x = x_in.contiguous(), delta = delta_in.contiguous();
  const int64_t c = x.size(-1);
  Tensor sum = at::empty_like(x);
  Tensor y = at::empty(x.sizes(), x.options().dtype(at::kBFloat16));
  const Tensor w = weight.detach().to(at::kFloat).contiguous(), b = bias.detach().to(at::kFloat).contiguous();
  if (x.numel() == 0) return {sum, y};
  Tensor gate_holder;
  int64_t rps = 1;
  const float* gate = gate_ptr(scale, x, gate_holder, rps);
  c10::cuda::CUDAGuard guard(x.device());
  check_rc(swin_add_layernorm_forward_gated(x.data_ptr(), xd, delta.data_ptr(), w.data_ptr<float>(), b.data_ptr<float>(), gate, rps,
                                            (float)eps, c ? x.numel() / c : 0, (int32_t)c, sum.data_ptr(), y.data_ptr(),
                                            at::cuda::getCurrentCUDAStream()));
  return {sum, y};
}

std::tuple<Tensor, Tensor> add_layer_norm_meta(const Tensor& x, const Tensor&, const Tensor&, const Tensor&, double,
                                               const optional<Tensor>&) {
  return {at::empty_symint(x.sym_sizes(), x.options()), at::empty_symint(x.sym_sizes(), x.options().dtype(at::kBFloat16))};
}

// backward of the pair: (dx, d_delta, dgamma, dbeta)
std::tuple<Tensor, Tensor, Tensor, Tensor> add_layer_norm_bwd_cuda(const Tensor& d_sum_in, const Tensor& dy_in, const Tensor& sum_in,
                                                                   const Tensor& weight, double eps, const optional<Tensor>& scale) {
  const int32_t xd = dtype_code(sum_in.scalar_type());
  TORCH_CHECK(dy_in.scalar_type() == at::kBFloat16 && d_sum_in.scalar_type() == sum_in.scalar_type(),
              "swin_attn::add_layer_norm_bwd: dy must be bfloat16 and d_sum of the stream dtype");
  TORCH_CHECK(dy_in.sizes() == sum_in.sizes() && d_sum_in.sizes() == sum_in.sizes(), "swin_attn::add_layer_norm_bwd: shape mismatch");
  const Tensor sum = sum_in.contiguous(), dy = dy_in.contiguous(), d_sum = d_sum_in.contiguous();
  const int64_t c = sum.size(-1);
  Tensor dx = at::empty_like(sum);
  Tensor d_delta = at::empty(sum.sizes(), sum.options().dtype(at::kBFloat16));
  Tensor dgamma = at::zeros({c}, sum.options().dtype(at::kFloat)), dbeta = at::zeros({c}, sum.options().dtype(at::kFloat));
  if (sum.numel() == 0) return {dx, d_delta, dgamma, dbeta};
  const Tensor w = weight.detach().to(at::kFloat).contiguous();
  Tensor gate_holder;
  int64_t rps = 1;
  const float* gate = gate_ptr(scale, sum, gate_holder, rps);
  c10::cuda::CUDAGuard guard(sum.device());
  check_rc(swin_add_layernorm_backward(sum.data_ptr(), xd, dy.data_ptr(), d_sum.data_ptr(), w.data_ptr<float>(), gate, rps, (float)eps,
                                       sum.numel() / c, (int32_t)c, dx.data_ptr(), d_delta.data_ptr(), dgamma.data_ptr<float>(),
                                       dbeta.data_ptr<float>(), at::cuda::getCurrentCUDAStream()));
  return {dx, d_delta, dgamma, dbeta};
}

std::tuple<Tensor, Tensor, Tensor, Tensor> add_layer_norm_bwd_meta(const Tensor&, const Tensor&, const Tensor& sum, const Tensor& weight, double,
                                                                   const optional<Tensor>&) {
  return {at::empty_symint(sum.sym_sizes(), sum.options()), at::empty_symint(sum.sym_sizes(), sum.options().dtype(at::kBFloat16)),
          at::empty_symint({sum.sym_size(-1)}, sum.options().dtype(at::kFloat)),
          at::empty_symint({sum.sym_size(-1)}, sum.options().dtype(at::kFloat))};
}

Tensor gelu_cuda(const Tensor& x_in) {
  TORCH_CHECK(x_in.scalar_type() == at::kBFloat16 && x_in.numel() % 8 == 0, "swin_attn::gelu: bfloat16 with a multiple of 8 elements");
  const Tensor x = x_in.contiguous();
  Tensor y = at::empty_like(x);
  if (x.numel() == 0) return y;
  c10::cuda::CUDAGuard guard(x.device());
  check_rc(swin_gelu_forward(x.data_ptr(), y.data_ptr(), x.numel(), at::cuda::getCurrentCUDAStream()));
  return y;
}

// act(x W^T + b) with the tcgen05 GEMM (inference; activation: 0 none, 1 exact GELU)
Tensor linear_act_cuda(const Tensor& x_in, const Tensor& weight, const Tensor& bias, int64_t activation) {
  TORCH_CHECK(x_in.scalar_type() == at::kBFloat16 && weight.scalar_type() == at::kBFloat16 && bias.scalar_type() == at::kBFloat16,
              "swin_attn::linear_act: x, weight and bias must be bfloat16");
  TORCH_CHECK(weight.dim() == 2 && bias.dim() == 1 && x_in.dim() >= 1 && x_in.size(-1) == weight.size(1) && bias.size(0) == weight.size(0),
              "swin_attn::linear_act: x (..., in), weight (out, in), bias (out)");
  const Tensor x = x_in.contiguous(), w = weight.contiguous(), b = bias.contiguous();
  std::vector<int64_t> shape(x.sizes().begin(), x.sizes().end());
  shape.back() = w.size(0);
  Tensor y = at::empty(shape, x.options());
  if (y.numel() == 0) return y;
  c10::cuda::CUDAGuard guard(x.device());
  cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  const int64_t rows = x.numel() / x.size(-1), k = w.size(1), n = w.size(0);
  TimedLaunch timed(stream);
  check_rc(swin_linear_forward(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), rows, (int32_t)k, (int32_t)n, (int32_t)activation,
                               stream));
  // algorithmic traffic: x, W and y once; flops of the product (the GELU is not counted)
  timed.finish(2 * (rows * k + n * k + rows * n), 2 * rows * n * k, 2, k, activation, n, rows, 0);
  return y;
}

// (h, GELU(h)) = linear_gelu_both(x, W, b): the training form of fc1 + GELU, one tcgen05 GEMM writing both tensors
std::tuple<Tensor, Tensor> linear_gelu_both_cuda(const Tensor& x_in, const Tensor& weight, const Tensor& bias) {
  TORCH_CHECK(x_in.scalar_type() == at::kBFloat16 && weight.scalar_type() == at::kBFloat16 && bias.scalar_type() == at::kBFloat16,
              "swin_attn::linear_gelu_both: x, weight and bias must be bfloat16");
  TORCH_CHECK(weight.dim() == 2 && bias.dim() == 1 && x_in.dim() >= 1 && x_in.size(-1) == weight.size(1) && bias.size(0) == weight.size(0),
              "swin_attn::linear_gelu_both: x (..., in), weight (out, in), bias (out)");
  const Tensor x = x_in.contiguous(), w = weight.contiguous(), b = bias.contiguous();
  std::vector<int64_t> shape(x.sizes().begin(), x.sizes().end());
  shape.back() = w.size(0);
  Tensor h = at::empty(shape, x.options()), y = at::empty(shape, x.options());
  if (y.numel() == 0) return {h, y};
  c10::cuda::CUDAGuard guard(x.device());
  cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  const int64_t rows = x.numel() / x.size(-1), k = w.size(1), n = w.size(0);
  TimedLaunch timed(stream);
  check_rc(swin_linear_gelu_forward_train(x.data_ptr(), w.data_ptr(), b.data_ptr(), h.data_ptr(), y.data_ptr(), rows, (int32_t)k, (int32_t)n,
                                          stream));
  timed.finish(2 * (rows * k + n * k + 2 * rows * n), 2 * rows * n * k, 2, k, 3, n, rows, 0);
  return {h, y};
}

std::tuple<Tensor, Tensor> linear_gelu_both_meta(const Tensor& x, const Tensor& weight, const Tensor&) {
  std::vector<c10::SymInt> shape(x.sym_sizes().begin(), x.sym_sizes().end());
  shape.back() = weight.sym_size(0);
  return {at::empty_symint(shape, x.options()), at::empty_symint(shape, x.options())};
}

// residual + (x W^T + b) with the tcgen05 GEMM (inference; bf16 residual stream)
Tensor linear_residual_cuda(const Tensor& x_in, const Tensor& weight, const Tensor& bias, const Tensor& residual_in) {
  TORCH_CHECK(x_in.scalar_type() == at::kBFloat16 && weight.scalar_type() == at::kBFloat16 && bias.scalar_type() == at::kBFloat16 &&
                  residual_in.scalar_type() == at::kBFloat16,
              "swin_attn::linear_residual: x, weight, bias and residual must be bfloat16");
  TORCH_CHECK(weight.dim() == 2 && bias.dim() == 1 && x_in.dim() >= 1 && x_in.size(-1) == weight.size(1) && bias.size(0) == weight.size(0),
              "swin_attn::linear_residual: x (..., in), weight (out, in), bias (out)");
  TORCH_CHECK(residual_in.dim() >= 1 && residual_in.size(-1) == weight.size(0) &&
                  residual_in.numel() / weight.size(0) == x_in.numel() / weight.size(1),
              "swin_attn::linear_residual: residual must hold one row of out_features per row of x");
  const Tensor x = x_in.contiguous(), w = weight.contiguous(), b = bias.contiguous(), r = residual_in.contiguous();
  Tensor y = at::empty(r.sizes(), r.options());
  if (y.numel() == 0) return y;
  c10::cuda::CUDAGuard guard(x.device());
  cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  const int64_t rows = x.numel() / x.size(-1), k = w.size(1), n = w.size(0);
  TimedLaunch timed(stream);
  check_rc(swin_linear_residual_forward(x.data_ptr(), w.data_ptr(), b.data_ptr(), r.data_ptr(), y.data_ptr(), rows, (int32_t)k, (int32_t)n,
                                        stream));
  // algorithmic traffic: x, W, the residual and y once; flops of the product
  timed.finish(2 * (rows * k + n * k + 2 * rows * n), 2 * rows * n * k, 2, k, 2, n, rows, 0);
  return y;
}

Tensor linear_residual_meta(const Tensor&, const Tensor&, const Tensor&, const Tensor& residual) {
  return at::empty_symint(residual.sym_sizes(), residual.options());
}

Tensor linear_act_meta(const Tensor& x, const Tensor& weight, const Tensor&, int64_t) {
  std::vector<c10::SymInt> shape(x.sym_sizes().begin(), x.sym_sizes().end());
  shape.back() = weight.sym_size(0);
  return at::empty_symint(shape, x.options());
}

Tensor gelu_bwd_cuda(const Tensor& dy_in, const Tensor& x_in) {
  TORCH_CHECK(x_in.scalar_type() == at::kBFloat16 && x_in.numel() % 8 == 0, "swin_attn::gelu_bwd: bfloat16 with a multiple of 8 elements");
  TORCH_CHECK(dy_in.sizes() == x_in.sizes(), "swin_attn::gelu_bwd: dy and x must have the same shape");
  const Tensor x = x_in.contiguous(), dy = dy_in.to(at::kBFloat16).contiguous();
  Tensor dx = at::empty_like(x);
  if (x.numel() == 0) return dx;
  c10::cuda::CUDAGuard guard(x.device());
  check_rc(swin_gelu_backward(x.data_ptr(), dy.data_ptr(), dx.data_ptr(), x.numel(), at::cuda::getCurrentCUDAStream()));
  return dx;
}

Tensor gelu_meta(const Tensor& x) { return at::empty_symint(x.sym_sizes(), x.options()); }
Tensor gelu_bwd_meta(const Tensor&, const Tensor& x) { return at::empty_symint(x.sym_sizes(), x.options()); }

// ---- swin_attn::linear: the autocast Linear with a fused bias gradient ---------------------------------------------
Tensor bias_grad_cuda(const Tensor& dy_in) {
  TORCH_CHECK(dy_in.dim() == 2 && dy_in.scalar_type() == at::kBFloat16 && dy_in.size(1) % 8 == 0,
              "swin_attn::bias_grad: (rows, cols) bfloat16 with cols a multiple of 8");
  const Tensor dy = dy_in.contiguous();
  Tensor out = at::empty({dy.size(1)}, dy.options().dtype(at::kFloat));
  if (dy.numel() == 0) return out.zero_();
  c10::cuda::CUDAGuard guard(dy.device());
  check_rc(swin_linear_bias_grad(dy.data_ptr(), dy.size(0), (int32_t)dy.size(1), out.data_ptr<float>(), at::cuda::getCurrentCUDAStream()));
  return out;
}

Tensor bias_grad_meta(const Tensor& dy) { return at::empty_symint({dy.sym_size(1)}, dy.options().dtype(at::kFloat)); }

// what torch.autocast(bf16) computes for nn.Linear: operands rounded to bf16, cuBLAS addmm with fp32 accumulation
Tensor linear_any(const Tensor& x, const Tensor& weight, const optional<Tensor>& bias) {
  return at::linear(x.to(at::kBFloat16), weight.to(at::kBFloat16),
                    bias.has_value() ? optional<Tensor>(bias->to(at::kBFloat16)) : c10::nullopt);
}

std::tuple<Tensor, Tensor, Tensor> layer_norm_bwd_cuda(const Tensor& dy_in, const Tensor& x_in, const Tensor& weight, double eps) {
  const int32_t xd = dtype_code(x_in.scalar_type());
  const Tensor x = x_in.contiguous();
  Tensor dy = dy_in;
  if (dy.scalar_type() != at::kFloat && dy.scalar_type() != at::kBFloat16) dy = dy.to(at::kFloat);
  dy = dy.contiguous();
  const int64_t c = x.size(-1);
  Tensor dx = at::empty_like(x);
  const Tensor w = weight.detach().to(at::kFloat).contiguous();
  Tensor dw = at::empty_like(w), db = at::empty_like(w);
  if (x.numel() == 0) return {dx, at::zeros_like(w), at::zeros_like(w)};
  c10::cuda::CUDAGuard guard(x.device());
  check_rc(swin_layernorm_backward(x.data_ptr(), xd, dy.data_ptr(), dtype_code(dy.scalar_type()), w.data_ptr<float>(), (float)eps,
                                   c ? x.numel() / c : 0, (int32_t)c, dx.data_ptr(), dw.data_ptr<float>(), db.data_ptr<float>(),
                                   at::cuda::getCurrentCUDAStream()));
  return {dx, dw, db};
}

std::tuple<Tensor, Tensor, Tensor> layer_norm_bwd_meta(const Tensor&, const Tensor& x, const Tensor& weight, double) {
  Tensor w = at::empty_symint(weight.sym_sizes(), weight.options().dtype(at::kFloat));
  return {at::empty_symint(x.sym_sizes(), x.options()), w, at::empty_like(w)};
}

// ---- dispatcher handles ----------------------------------------------------------------------------------------
using FwdSig = std::tuple<Tensor, Tensor>(const Tensor&, const optional<Tensor>&, const Tensor&, const optional<Tensor>&,
                                          const optional<Tensor>&, int64_t, int64_t, int64_t, double, bool, int64_t,
                                          const optional<Tensor>&, double, int64_t);
using BwdSig = std::tuple<Tensor, Tensor, Tensor>(const Tensor&, const Tensor&, const optional<Tensor>&, const Tensor&, const Tensor&,
                                                  const Tensor&, int64_t, int64_t, int64_t, double, const optional<Tensor>&, int64_t,
                                                  double, int64_t);
using LnSig = Tensor(const Tensor&, const Tensor&, const Tensor&, double, bool);
using LnBwdSig = std::tuple<Tensor, Tensor, Tensor>(const Tensor&, const Tensor&, const Tensor&, double);

inline auto fwd_op() {
  static auto op = c10::Dispatcher::singleton().findSchemaOrThrow("swin_attn::fwd", "").typed<FwdSig>();
  return op;
}
inline auto bwd_op() {
  static auto op = c10::Dispatcher::singleton().findSchemaOrThrow("swin_attn::bwd", "").typed<BwdSig>();
  return op;
}
using BwdUnpaddedSig = std::tuple<Tensor, Tensor, Tensor, Tensor, Tensor>(const Tensor&, const Tensor&, const optional<Tensor>&, const Tensor&,
                                                                          const Tensor&, const Tensor&, const optional<Tensor>&,
                                                                          const optional<Tensor>&, int64_t, int64_t, int64_t, double);
inline auto bwd_unpadded_op() {
  static auto op = c10::Dispatcher::singleton().findSchemaOrThrow("swin_attn::bwd_unpadded", "").typed<BwdUnpaddedSig>();
  return op;
}
inline auto ln_op() {
  static auto op = c10::Dispatcher::singleton().findSchemaOrThrow("swin_attn::layer_norm", "").typed<LnSig>();
  return op;
}
inline auto ln_bwd_op() {
  static auto op = c10::Dispatcher::singleton().findSchemaOrThrow("swin_attn::layer_norm_bwd", "").typed<LnBwdSig>();
  return op;
}

// ---- autograd ------------------------------------------------------------------------------------------------
struct AttnFunction : public torch::autograd::Function<AttnFunction> {
  static variable_list forward(AutogradContext* ctx, const Tensor& packed, const optional<Tensor>& v, const Tensor& bias_table,
                               const optional<Tensor>& packed_pad, const optional<Tensor>& v_pad, int64_t window, int64_t shift,
                               int64_t num_heads, double scale, bool need_lse, int64_t impl, const optional<Tensor>& attn_mask,
                               double dropout_p, int64_t seed) {
    at::AutoDispatchBelowADInplaceOrView guard;
    auto out = fwd_op().call(packed, v, bias_table, packed_pad, v_pad, window, shift, num_heads, scale, need_lse, impl, attn_mask,
                             dropout_p, seed);
    const Tensor& o = std::get<0>(out);
    const Tensor& lse = std::get<1>(out);
    ctx->save_for_backward({packed, v.has_value() ? *v : Tensor(), o, lse, bias_table, attn_mask.has_value() ? *attn_mask : Tensor(),
                            packed_pad.has_value() ? *packed_pad : Tensor(), v_pad.has_value() ? *v_pad : Tensor()});
    ctx->saved_data["window"] = window;
    ctx->saved_data["shift"] = shift;
    ctx->saved_data["heads"] = num_heads;
    ctx->saved_data["scale"] = scale;
    ctx->saved_data["dropout_p"] = dropout_p;
    ctx->saved_data["seed"] = seed;
    const bool aligned = packed.size(1) % window == 0 && packed.size(2) % window == 0;
    // un-padded tensors: the tcgen05 backward takes them with the pad vectors (swin_attn_backward_unpadded)
    // (window 7: tcgen05 kernel; window 12: warp-MMA kernel, which needs the pad rows not to wrap around the rolled image)
    const bool unpadded_ok = !aligned && packed.scalar_type() == at::kBFloat16 &&
                             ((window == 7 && (shift == 0 || shift == 3)) ||
                              (window == 12 && (shift == 0 || shift == 6) && packed.size(1) >= shift && packed.size(2) >= shift)) &&
                             !attn_mask.has_value() && dropout_p == 0.0;
    ctx->saved_data["ok"] = need_lse && (aligned || unpadded_ok);
    ctx->saved_data["unpadded"] = !aligned;
    ctx->mark_non_differentiable({lse});
    return {o, lse};
  }

  static variable_list backward(AutogradContext* ctx, variable_list grads) {
    TORCH_CHECK(ctx->saved_data["ok"].toBool(),
                "swin_attn: backward needs the forward to have run with need_lse=True on a tensor whose height and width are "
                "multiples of the window (the layers pad explicitly whenever gradients are enabled)");
    const auto saved = ctx->get_saved_variables();
    const Tensor &packed = saved[0], &v = saved[1], &o = saved[2], &lse = saved[3], &table = saved[4], &mask = saved[5];
    if (ctx->saved_data["unpadded"].toBool()) {
      const Tensor &packed_pad = saved[6], &v_pad = saved[7];
      auto out = bwd_unpadded_op().call(grads[0], packed, v.defined() ? optional<Tensor>(v) : c10::nullopt, o, lse, table,
                                        packed_pad.defined() ? optional<Tensor>(packed_pad) : c10::nullopt,
                                        v_pad.defined() ? optional<Tensor>(v_pad) : c10::nullopt, ctx->saved_data["window"].toInt(),
                                        ctx->saved_data["shift"].toInt(), ctx->saved_data["heads"].toInt(),
                                        ctx->saved_data["scale"].toDouble());
      Tensor d_v, d_pp, d_vp;
      if (v.defined()) d_v = std::get<1>(out).to(v.scalar_type());
      if (packed_pad.defined()) d_pp = std::get<3>(out).to(packed_pad.scalar_type()).view(packed_pad.sizes());
      if (v.defined() && v_pad.defined()) d_vp = std::get<4>(out).to(v_pad.scalar_type()).view(v_pad.sizes());
      return {std::get<0>(out), d_v, std::get<2>(out), d_pp, d_vp, Tensor(), Tensor(), Tensor(), Tensor(), Tensor(), Tensor(),
              Tensor(), Tensor(), Tensor()};
    }
    auto out = bwd_op().call(grads[0], packed, v.defined() ? optional<Tensor>(v) : c10::nullopt, o, lse, table,
                             ctx->saved_data["window"].toInt(), ctx->saved_data["shift"].toInt(), ctx->saved_data["heads"].toInt(),
                             ctx->saved_data["scale"].toDouble(), mask.defined() ? optional<Tensor>(mask) : c10::nullopt,
                             SWIN_ATTN_IMPL_AUTO, ctx->saved_data["dropout_p"].toDouble(), ctx->saved_data["seed"].toInt());
    Tensor d_v;
    if (v.defined()) {
      d_v = std::get<1>(out);
      if (d_v.scalar_type() != v.scalar_type()) d_v = d_v.to(v.scalar_type());
    }
    return {std::get<0>(out), d_v, std::get<2>(out), Tensor(), Tensor(), Tensor(), Tensor(), Tensor(), Tensor(), Tensor(), Tensor(),
            Tensor(), Tensor(), Tensor()};
  }
};

std::tuple<Tensor, Tensor> fwd_autograd(const Tensor& packed, const optional<Tensor>& v, const Tensor& bias_table,
                                        const optional<Tensor>& packed_pad, const optional<Tensor>& v_pad, int64_t window,
                                        int64_t shift, int64_t num_heads, double scale, bool need_lse, int64_t impl,
                                        const optional<Tensor>& attn_mask, double dropout_p, int64_t seed) {
  auto out = AttnFunction::apply(packed, v, bias_table, packed_pad, v_pad, window, shift, num_heads, scale, need_lse, impl, attn_mask,
                                 dropout_p, seed);
  return {out[0], out[1]};
}

struct LayerNormFunction : public torch::autograd::Function<LayerNormFunction> {
  static Tensor forward(AutogradContext* ctx, const Tensor& x, const Tensor& weight, const Tensor& bias, double eps, bool out_bf16) {
    at::AutoDispatchBelowADInplaceOrView guard;
    ctx->save_for_backward({x, weight});
    ctx->saved_data["eps"] = eps;
    ctx->saved_data["out_bf16"] = out_bf16;
    return ln_op().call(x, weight, bias, eps, out_bf16);
  }

  static variable_list backward(AutogradContext* ctx, variable_list grads) {
    const auto saved = ctx->get_saved_variables();
    const Tensor &x = saved[0], &weight = saved[1];
    TORCH_CHECK(x.size(-1) <= 1536, "swin_attn::layer_norm: the backward covers <= 1536 channels");
    const Tensor dy = grads[0].to(ctx->saved_data["out_bf16"].toBool() ? at::kBFloat16 : at::kFloat);
    auto out = ln_bwd_op().call(dy, x, weight, ctx->saved_data["eps"].toDouble());
    return {std::get<0>(out), std::get<1>(out).to(weight.scalar_type()), std::get<2>(out).to(weight.scalar_type()), Tensor(), Tensor()};
  }
};

// (sum, y) = add_layer_norm(x, delta, weight, bias, eps, scale): residual add (+ DropPath gate) + LayerNorm as one node
struct AddLayerNormFunction : public torch::autograd::Function<AddLayerNormFunction> {
  static variable_list forward(AutogradContext* ctx, const Tensor& x, const Tensor& delta, const Tensor& weight, const Tensor& bias,
                               double eps, const optional<Tensor>& scale) {
    at::AutoDispatchBelowADInplaceOrView guard;
    static auto op = c10::Dispatcher::singleton()
                         .findSchemaOrThrow("swin_attn::add_layer_norm", "")
                         .typed<std::tuple<Tensor, Tensor>(const Tensor&, const Tensor&, const Tensor&, const Tensor&, double,
                                                           const optional<Tensor>&)>();
    auto out = op.call(x, delta, weight, bias, eps, scale);
    ctx->save_for_backward({std::get<0>(out), weight, scale.has_value() ? *scale : Tensor()});
    ctx->saved_data["eps"] = eps;
    return {std::get<0>(out), std::get<1>(out)};
  }

  static variable_list backward(AutogradContext* ctx, variable_list grads) {
    static auto op = c10::Dispatcher::singleton()
                         .findSchemaOrThrow("swin_attn::add_layer_norm_bwd", "")
                         .typed<std::tuple<Tensor, Tensor, Tensor, Tensor>(const Tensor&, const Tensor&, const Tensor&, const Tensor&, double,
                                                                           const optional<Tensor>&)>();
    const auto saved = ctx->get_saved_variables();
    const Tensor &sum = saved[0], &weight = saved[1];
    const optional<Tensor> scale = saved[2].defined() ? optional<Tensor>(saved[2]) : c10::nullopt;
    const Tensor d_sum = grads[0].defined() ? grads[0].to(sum.scalar_type()) : at::zeros_like(sum);
    const Tensor dy = grads[1].defined() ? grads[1].to(at::kBFloat16) : at::zeros(sum.sizes(), sum.options().dtype(at::kBFloat16));
    auto out = op.call(d_sum, dy, sum, weight, ctx->saved_data["eps"].toDouble(), scale);
    return {std::get<0>(out), std::get<1>(out), std::get<2>(out).to(weight.scalar_type()), std::get<3>(out).to(weight.scalar_type()),
            Tensor(), Tensor()};
  }
};

std::tuple<Tensor, Tensor> add_layer_norm_autograd(const Tensor& x, const Tensor& delta, const Tensor& weight, const Tensor& bias, double eps,
                                                   const optional<Tensor>& scale) {
  auto out = AddLayerNormFunction::apply(x, delta, weight, bias, eps, scale);
  return {out[0], out[1]};
}

struct GeluFunction : public torch::autograd::Function<GeluFunction> {
  static Tensor forward(AutogradContext* ctx, const Tensor& x) {
    at::AutoDispatchBelowADInplaceOrView guard;
    ctx->save_for_backward({x});
    static auto op = c10::Dispatcher::singleton().findSchemaOrThrow("swin_attn::gelu", "").typed<Tensor(const Tensor&)>();
    return op.call(x);
  }
  static variable_list backward(AutogradContext* ctx, variable_list grads) {
    static auto op = c10::Dispatcher::singleton().findSchemaOrThrow("swin_attn::gelu_bwd", "").typed<Tensor(const Tensor&, const Tensor&)>();
    return {op.call(grads[0], ctx->get_saved_variables()[0])};
  }
};

Tensor gelu_autograd(const Tensor& x) { return GeluFunction::apply(x); }

struct LinearFunction : public torch::autograd::Function<LinearFunction> {
  static Tensor forward(AutogradContext* ctx, const Tensor& x, const Tensor& weight, const optional<Tensor>& bias) {
    at::AutoDispatchBelowADInplaceOrView guard;
    const Tensor xb = x.to(at::kBFloat16), wb = weight.to(at::kBFloat16);
    ctx->save_for_backward({xb, wb});
    ctx->saved_data["x_dtype"] = (int64_t)x.scalar_type();
    ctx->saved_data["w_dtype"] = (int64_t)weight.scalar_type();
    ctx->saved_data["b_dtype"] = bias.has_value() ? (int64_t)bias->scalar_type() : (int64_t)-1;
    return at::linear(xb, wb, bias.has_value() ? optional<Tensor>(bias->to(at::kBFloat16)) : c10::nullopt);
  }
  static variable_list backward(AutogradContext* ctx, variable_list grads) {
    static auto op = c10::Dispatcher::singleton().findSchemaOrThrow("swin_attn::bias_grad", "").typed<Tensor(const Tensor&)>();
    const auto saved = ctx->get_saved_variables();
    const Tensor &xb = saved[0], &wb = saved[1];
    const int64_t n = wb.size(0), k = wb.size(1);
    const Tensor dy = grads[0].to(at::kBFloat16).reshape({-1, n}).contiguous();
    Tensor dx, dw, db;
    if (ctx->needs_input_grad(0))
      dx = at::mm(dy, wb).view(xb.sizes()).to((at::ScalarType)ctx->saved_data["x_dtype"].toInt());
    if (ctx->needs_input_grad(1)) {
      // fp32 master weights: cuBLAS writes the fp32 gradient directly (autocast rounds it to bf16 and casts it back
      // in a second kernel)
      const auto wd = (at::ScalarType)ctx->saved_data["w_dtype"].toInt();
      dw = wd == at::kFloat ? at::mm(dy.t(), xb.reshape({-1, k}), at::kFloat) : at::mm(dy.t(), xb.reshape({-1, k})).to(wd);
    }
    const int64_t bd = ctx->saved_data["b_dtype"].toInt();
    if (bd >= 0 && ctx->needs_input_grad(2)) db = (n % 8 == 0 ? op.call(dy) : dy.sum(0)).to((at::ScalarType)bd);
    return {dx, dw, db};
  }
};

// GELU(x W^T + b) in training under bf16 autocast: ONE forward kernel (the tcgen05 GEMM writes h and GELU(h)); backward =
// the fused GELU backward on the saved h, then the Linear backward of LinearFunction (cuBLAS dgrad / wgrad in fp32 for
// fp32 master weights, fused column sum for the bias).
struct LinearGeluFunction : public torch::autograd::Function<LinearGeluFunction> {
  static Tensor forward(AutogradContext* ctx, const Tensor& x, const Tensor& weight, const Tensor& bias) {
    at::AutoDispatchBelowADInplaceOrView guard;
    static auto op = c10::Dispatcher::singleton()
                         .findSchemaOrThrow("swin_attn::linear_gelu_both", "")
                         .typed<std::tuple<Tensor, Tensor>(const Tensor&, const Tensor&, const Tensor&)>();
    const Tensor xb = x.to(at::kBFloat16), wb = weight.to(at::kBFloat16);
    auto out = op.call(xb, wb, bias.to(at::kBFloat16));
    ctx->save_for_backward({xb, wb, std::get<0>(out)});
    ctx->saved_data["x_dtype"] = (int64_t)x.scalar_type();
    ctx->saved_data["w_dtype"] = (int64_t)weight.scalar_type();
    ctx->saved_data["b_dtype"] = (int64_t)bias.scalar_type();
    return std::get<1>(out);
  }
  static variable_list backward(AutogradContext* ctx, variable_list grads) {
    static auto colsum = c10::Dispatcher::singleton().findSchemaOrThrow("swin_attn::bias_grad", "").typed<Tensor(const Tensor&)>();
    static auto gelu_bwd = c10::Dispatcher::singleton().findSchemaOrThrow("swin_attn::gelu_bwd", "").typed<Tensor(const Tensor&, const Tensor&)>();
    const auto saved = ctx->get_saved_variables();
    const Tensor &xb = saved[0], &wb = saved[1], &h = saved[2];
    const int64_t n = wb.size(0), k = wb.size(1);
    const Tensor dy = gelu_bwd.call(grads[0], h).reshape({-1, n});
    Tensor dx, dw, db;
    if (ctx->needs_input_grad(0))
      dx = at::mm(dy, wb).view(xb.sizes()).to((at::ScalarType)ctx->saved_data["x_dtype"].toInt());
    if (ctx->needs_input_grad(1)) {
      const auto wd = (at::ScalarType)ctx->saved_data["w_dtype"].toInt();
      dw = wd == at::kFloat ? at::mm(dy.t(), xb.reshape({-1, k}), at::kFloat) : at::mm(dy.t(), xb.reshape({-1, k})).to(wd);
    }
    if (ctx->needs_input_grad(2)) db = colsum.call(dy).to((at::ScalarType)ctx->saved_data["b_dtype"].toInt());
    return {dx, dw, db};
  }
};

Tensor linear_gelu_autograd(const Tensor& x, const Tensor& weight, const Tensor& bias) {
  return LinearGeluFunction::apply(x, weight, bias);
}

// below autograd (no_grad callers): the fused forward alone
Tensor linear_gelu_plain(const Tensor& x, const Tensor& weight, const Tensor& bias) {
  static auto op = c10::Dispatcher::singleton()
                       .findSchemaOrThrow("swin_attn::linear_act", "")
                       .typed<Tensor(const Tensor&, const Tensor&, const Tensor&, int64_t)>();
  return op.call(x.to(at::kBFloat16), weight.to(at::kBFloat16), bias.to(at::kBFloat16), 1);
}

Tensor linear_autograd(const Tensor& x, const Tensor& weight, const optional<Tensor>& bias) {
  return LinearFunction::apply(x, weight, bias);
}

Tensor layer_norm_autograd(const Tensor& x, const Tensor& weight, const Tensor& bias, double eps, bool out_bf16) {
  return LayerNormFunction::apply(x, weight, bias, eps, out_bf16);
}

}  // namespace

TORCH_LIBRARY(swin_attn, m) {
  m.def("fwd(Tensor packed, Tensor? v, Tensor bias_table, Tensor? packed_pad, Tensor? v_pad, int window, int shift, "
        "int num_heads, float scale, bool need_lse, int impl, Tensor? attn_mask=None, float dropout_p=0.0, int seed=0) "
        "-> (Tensor, Tensor)");
  m.def("bwd(Tensor d_o, Tensor packed, Tensor? v, Tensor o, Tensor lse, Tensor bias_table, int window, int shift, "
        "int num_heads, float scale, Tensor? attn_mask=None, int impl=0, float dropout_p=0.0, int seed=0) "
        "-> (Tensor, Tensor, Tensor)");
  m.def("bwd_unpadded(Tensor d_o, Tensor packed, Tensor? v, Tensor o, Tensor lse, Tensor bias_table, Tensor? packed_pad, "
        "Tensor? v_pad, int window, int shift, int num_heads, float scale) -> (Tensor, Tensor, Tensor, Tensor, Tensor)");
  m.def("layer_norm(Tensor x, Tensor weight, Tensor bias, float eps, bool out_bf16) -> Tensor");
  m.def("layer_norm_bwd(Tensor dy, Tensor x, Tensor weight, float eps) -> (Tensor, Tensor, Tensor)");
  m.def("add_layer_norm(Tensor x, Tensor delta, Tensor weight, Tensor bias, float eps, Tensor? scale=None) -> (Tensor, Tensor)");
  m.def("add_layer_norm_bwd(Tensor d_sum, Tensor dy, Tensor sum, Tensor weight, float eps, Tensor? scale=None) "
        "-> (Tensor, Tensor, Tensor, Tensor)");
  m.def("patch_merge_norm(Tensor x, int height, int width, Tensor weight, Tensor bias, float eps) -> Tensor");
  m.def("linear(Tensor x, Tensor weight, Tensor? bias) -> Tensor");
  m.def("bias_grad(Tensor dy) -> Tensor");
  m.def("gelu(Tensor x) -> Tensor");
  m.def("gelu_bwd(Tensor dy, Tensor x) -> Tensor");
  m.def("linear_act(Tensor x, Tensor weight, Tensor bias, int activation) -> Tensor");
  m.def("linear_residual(Tensor x, Tensor weight, Tensor bias, Tensor residual) -> Tensor");
  m.def("linear_gelu_both(Tensor x, Tensor weight, Tensor bias) -> (Tensor, Tensor)");
  m.def("linear_gelu(Tensor x, Tensor weight, Tensor bias) -> Tensor");
}

TORCH_LIBRARY_IMPL(swin_attn, CUDA, m) {
  m.impl("fwd", fwd_cuda);
  m.impl("bwd", bwd_cuda);
  m.impl("bwd_unpadded", bwd_unpadded_cuda);
  m.impl("layer_norm", layer_norm_cuda);
  m.impl("layer_norm_bwd", layer_norm_bwd_cuda);
  m.impl("add_layer_norm", add_layer_norm_cuda);
  m.impl("add_layer_norm_bwd", add_layer_norm_bwd_cuda);
  m.impl("patch_merge_norm", patch_merge_norm_cuda);
  m.impl("linear", linear_any);
  m.impl("bias_grad", bias_grad_cuda);
  m.impl("gelu", gelu_cuda);
  m.impl("gelu_bwd", gelu_bwd_cuda);
  m.impl("linear_act", linear_act_cuda);
  m.impl("linear_residual", linear_residual_cuda);
  m.impl("linear_gelu_both", linear_gelu_both_cuda);
  m.impl("linear_gelu", linear_gelu_plain);
}

TORCH_LIBRARY_IMPL(swin_attn, Meta, m) {
  m.impl("fwd", fwd_meta);
  m.impl("bwd", bwd_meta);
  m.impl("bwd_unpadded", bwd_unpadded_meta);
  m.impl("layer_norm", layer_norm_meta);
  m.impl("layer_norm_bwd", layer_norm_bwd_meta);
  m.impl("add_layer_norm", add_layer_norm_meta);
  m.impl("add_layer_norm_bwd", add_layer_norm_bwd_meta);
  m.impl("patch_merge_norm", patch_merge_norm_meta);
  m.impl("linear", linear_any);
  m.impl("bias_grad", bias_grad_meta);
  m.impl("gelu", gelu_meta);
  m.impl("gelu_bwd", gelu_bwd_meta);
  m.impl("linear_act", linear_act_meta);
  m.impl("linear_residual", linear_residual_meta);
  m.impl("linear_gelu_both", linear_gelu_both_meta);
  m.impl("linear_gelu", linear_gelu_plain);
}

TORCH_LIBRARY_IMPL(swin_attn, Autograd, m) {
  m.impl("fwd", fwd_autograd);
  m.impl("layer_norm", layer_norm_autograd);
  m.impl("add_layer_norm", add_layer_norm_autograd);
  m.impl("gelu", gelu_autograd);
  m.impl("linear", linear_autograd);
  m.impl("linear_act", torch::autograd::autogradNotImplementedFallback());   // inference only
  m.impl("patch_merge_norm", torch::autograd::autogradNotImplementedFallback());   // inference only
  m.impl("linear_residual", torch::autograd::autogradNotImplementedFallback());    // inference only
  m.impl("linear_gelu_both", torch::autograd::autogradNotImplementedFallback());   // differentiated through linear_gelu
  m.impl("linear_gelu", linear_gelu_autograd);
}

// ---- swin_attn::mha_fwd: global multi-head attention that returns its map (Depthformer SelfAttentionBlock core) -------
// q, k: (B, S, nH * dq), v: (B, S, nH * dv), unit channel stride (row strides may exceed the width: slices of a packed
// projection).  Returns o (B, S, nH * dv) in the input dtype and the fp32 map (B, nH, S, S).
std::tuple<Tensor, Tensor> mha_fwd_impl(const Tensor& q_in, const Tensor& k_in, const Tensor& v_in, const optional<Tensor>& token_index,
                                        const optional<Tensor>& index_bias, int64_t num_heads, double scale) {
  TORCH_CHECK(q_in.dim() == 3 && k_in.dim() == 3 && v_in.dim() == 3, "mha_fwd: q, k, v must be (B, S, D)");
  TORCH_CHECK(q_in.size(0) == k_in.size(0) && q_in.size(2) == k_in.size(2) && v_in.size(0) == k_in.size(0) && v_in.size(1) == k_in.size(1),
              "mha_fwd: q (B, S, D), k (B, Sk, D), v (B, Sk, Dv) shapes disagree");
  TORCH_CHECK(num_heads > 0 && q_in.size(2) % num_heads == 0 && v_in.size(2) % num_heads == 0, "Hidden dim not multiple of num heads.");
  const auto dt = q_in.scalar_type();
  const int32_t code = dtype_code(dt);
  auto rows = [&](const Tensor& t) {
    Tensor u = t.to(dt);
    return (u.stride(2) == 1 && u.stride(0) == u.size(1) * u.stride(1) && u.stride(1) >= u.size(2)) ? u : u.contiguous();
  };
  const Tensor q = rows(q_in), k = rows(k_in), v = rows(v_in);
  const int64_t b = q.size(0), s = q.size(1), sk = k.size(1);
  Tensor o = at::empty({b, s, v.size(2)}, q.options());
  Tensor attn = at::empty({b, num_heads, s, sk}, q.options().dtype(at::kFloat));
  mha_shape_t shp;
  memset(&shp, 0, sizeof(shp));
  shp.batch = (int32_t)b; shp.seq = (int32_t)s; shp.kv_seq = (int32_t)sk; shp.num_heads = (int32_t)num_heads;
  shp.qk_head_dim = (int32_t)(q.size(2) / num_heads); shp.v_head_dim = (int32_t)(v.size(2) / num_heads);
  shp.dtype = code; shp.scale = (float)scale;
  shp.q_row_stride = q.stride(1); shp.k_row_stride = k.stride(1); shp.v_row_stride = v.stride(1); shp.o_row_stride = o.stride(1);
  c10::cuda::CUDAGuard guard(q.device());
  cudaStream_t stream = at::cuda::getCurrentCUDAStream();
  if (token_index.has_value()) {
    TORCH_CHECK(index_bias.has_value() && index_bias->dim() == 2 && index_bias->size(1) == num_heads && index_bias->size(0) % 2 == 1,
                "mha_fwd_ordered: index_bias must be (2 * num_emb - 1, num_heads)");
    TORCH_CHECK(token_index->dim() == 2 && token_index->size(0) == b && token_index->size(1) == s && s == sk,
                "mha_fwd_ordered: token_index must be (B, S) and the attention self-attention");
    const Tensor idx = token_index->to(at::kInt).contiguous();
    const Tensor bias = index_bias->detach().to(at::kFloat).contiguous();
    check_rc(mha_attention_forward_ordered(&shp, q.data_ptr(), k.data_ptr(), v.data_ptr(), idx.data_ptr<int32_t>(), bias.data_ptr<float>(),
                                           (int32_t)((bias.size(0) + 1) / 2), o.data_ptr(), attn.data_ptr<float>(), stream));
  } else {
    check_rc(mha_attention_forward(&shp, q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), attn.data_ptr<float>(), stream));
  }
  return {o, attn};
}

std::tuple<Tensor, Tensor> mha_fwd_cuda(const Tensor& q, const Tensor& k, const Tensor& v, int64_t num_heads, double scale) {
  return mha_fwd_impl(q, k, v, c10::nullopt, c10::nullopt, num_heads, scale);
}

std::tuple<Tensor, Tensor> mha_fwd_ordered_cuda(const Tensor& q, const Tensor& k, const Tensor& v, const Tensor& token_index,
                                                const Tensor& index_bias, int64_t num_heads, double scale) {
  return mha_fwd_impl(q, k, v, token_index, index_bias, num_heads, scale);
}

std::tuple<Tensor, Tensor> mha_fwd_ordered_meta(const Tensor& q, const Tensor& k, const Tensor& v, const Tensor&, const Tensor&,
                                                int64_t num_heads, double) {
  return {at::empty_symint({q.sym_size(0), q.sym_size(1), v.sym_size(2)}, q.options()),
          at::empty_symint({q.sym_size(0), c10::SymInt(num_heads), q.sym_size(1), k.sym_size(1)}, q.options().dtype(at::kFloat))};
}

std::tuple<Tensor, Tensor> mha_fwd_meta(const Tensor& q, const Tensor& k, const Tensor& v, int64_t num_heads, double) {
  return {at::empty_symint({q.sym_size(0), q.sym_size(1), v.sym_size(2)}, q.options()),
          at::empty_symint({q.sym_size(0), c10::SymInt(num_heads), q.sym_size(1), k.sym_size(1)}, q.options().dtype(at::kFloat))};
}

TORCH_LIBRARY_FRAGMENT(swin_attn, m) {
  m.def("mha_fwd(Tensor q, Tensor k, Tensor v, int num_heads, float scale) -> (Tensor, Tensor)");
  m.def("mha_fwd_ordered(Tensor q, Tensor k, Tensor v, Tensor token_index, Tensor index_bias, int num_heads, float scale) -> (Tensor, Tensor)");
}
TORCH_LIBRARY_IMPL(swin_attn, CUDA, m) {
  m.impl("mha_fwd", mha_fwd_cuda);
  m.impl("mha_fwd_ordered", mha_fwd_ordered_cuda);
}
TORCH_LIBRARY_IMPL(swin_attn, Meta, m) {
  m.impl("mha_fwd", mha_fwd_meta);
  m.impl("mha_fwd_ordered", mha_fwd_ordered_meta);
}

// ---- timing control (ctypes; not tensor ops) -------------------------------------------------------------------
extern "C" {

void swin_torch_set_timing(int mode) {
  std::lock_guard<std::mutex> lock(g_timing_mutex);
  g_timing_on = mode;
}

// Graph mode: elapsed times of the event pairs captured into CUDA graphs, as recorded by the LAST replay (the caller
// has synchronised).  Same layout as swin_torch_take_timing; the records stay.  Returns the number written.
int swin_torch_read_graph_timing(double* ms, int64_t* meta, int cap) {
  std::lock_guard<std::mutex> lock(g_timing_mutex);
  int n = 0;
  for (auto& r : g_graph_records) {
    if (n >= cap) break;
    float t = 0.f;
    if (cudaEventElapsedTime(&t, r.start, r.stop) != cudaSuccess) {
      cudaGetLastError();
      t = -1.f;
    }
    ms[n] = t;
    meta[8 * n] = r.bytes;
    meta[8 * n + 1] = r.flops;
    for (int i = 0; i < 6; ++i) meta[8 * n + 2 + i] = r.tag[i];
    ++n;
  }
  return n;
}

int swin_torch_graph_timing_count(void) {
  std::lock_guard<std::mutex> lock(g_timing_mutex);
  return (int)g_graph_records.size();
}

// Forget the graph-mode records (their graphs must have been destroyed, or never be replayed again).
void swin_torch_clear_graph_timing(void) {
  std::lock_guard<std::mutex> lock(g_timing_mutex);
  for (auto& r : g_graph_records) {
    cudaEventDestroy(r.start);
    cudaEventDestroy(r.stop);
  }
  g_graph_records.clear();
}

// Synchronises on the recorded events and moves up to `cap` records out: ms[i], meta[8 i .. 8 i + 7] =
// {bytes, flops, kind, window, shift, heads, height, width}.  Returns the number written.
int swin_torch_take_timing(double* ms, int64_t* meta, int cap) {
  std::vector<TimingRecord> recs;
  {
    std::lock_guard<std::mutex> lock(g_timing_mutex);
    recs.swap(g_records);
  }
  int n = 0;
  for (auto& r : recs) {
    cudaEventSynchronize(r.stop);
    float t = 0.f;
    cudaEventElapsedTime(&t, r.start, r.stop);
    if (n < cap) {
      ms[n] = t;
      meta[8 * n] = r.bytes;
      meta[8 * n + 1] = r.flops;
      for (int i = 0; i < 6; ++i) meta[8 * n + 2 + i] = r.tag[i];
      ++n;
    }
    cudaEventDestroy(r.start);
    cudaEventDestroy(r.stop);
  }
  return n;
}

int swin_torch_pending_timing(void) {
  std::lock_guard<std::mutex> lock(g_timing_mutex);
  return (int)g_records.size();
}

}  // extern "C"
