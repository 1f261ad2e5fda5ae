/*
 * swin_attn.h -- C ABI of the B200 (sm_100a) shifted-window attention library.
 *
 * This is the drop-in boundary for ONE hot path of
 * pitlover/Monocular-Depth-Estimation: everything the reference computes
 * between the q/k(/v) Linear and the output Linear of its two WindowAttention
 * modules, together with the pad / torch.roll / window_partition /
 * window_reverse / crop plumbing of the blocks that call them.
 *
 *   reference interface replaced                         (paths under /root/reference)
 *   ---------------------------------------------------------------------------------
 *   WindowAttention.forward (decoder, x & external v)    model/NewCRFs/newcrf_layers.py:110-149
 *   WindowAttention.forward (encoder, fused qkv)         model/NewCRFs/swin_transformer.py:112-144
 *   window_partition / window_reverse                    newcrf_layers.py:30-59, swin_transformer.py:32-61
 *   F.pad / torch.roll / crop in the block forwards      newcrf_layers.py:210-249, swin_transformer.py:204-238
 *   SW-MSA mask construction                             newcrf_layers.py:331-350, swin_transformer.py:361-380
 *   relative_position_bias_table[index] gather           newcrf_layers.py:124-127, swin_transformer.py:126-129
 *   autograd backward of all of the above                (implicit in the reference)
 *
 * The reference has no FFI; its boundary is the Python nn.Module API, which
 * monocular-depth-estimation_b200/layers.py mirrors on top of these entry
 * points (bound with ctypes).
 *
 * Conventions
 *   - All tensors are "image order": (B, H, W, channels), un-shifted,
 *     un-partitioned.  The cyclic shift, the window tiling, the zero padding
 *     up to (Hp, Wp) = (ceil(H/ws)*ws, ceil(W/ws)*ws) and the final crop are
 *     address arithmetic inside the kernels; nothing is materialised.
 *   - q/k/v may be channel slices of a wider tensor (e.g. the (B,H,W,3C)
 *     output of the qkv Linear): `*_pix_stride` is the distance in ELEMENTS
 *     between consecutive pixels, the pointer already includes the channel
 *     offset.  Rows and images are dense: row stride = W * pix_stride.
 *   - head_dim must be 32 (every NewCRFs / Swin configuration); window 7 or 12.
 *   - dtype: SWIN_ATTN_F32 (fp32 storage, fp32 FFMA math, for <=1e-5 parity)
 *            SWIN_ATTN_BF16 (bf16 storage, tcgen05 bf16 MMA, fp32 accumulate/softmax)
 *   - `*_pad` (nullable): C-element vectors giving the value of q/k/v at the
 *     padding tokens (y >= H or x >= W).  The reference pads BEFORE the
 *     Linear, so pad tokens carry the Linear bias; NULL means zeros (decoder
 *     v, newcrf_layers.py:216).
 *   - `stream` is a cudaStream_t (NULL = default stream).  Device entry points
 *     never synchronise.
 *   - Return value: 0 on success, non-zero on error; swin_attn_last_error()
 *     then returns a message (thread-local).
 */
#ifndef SWIN_ATTN_H_
#define SWIN_ATTN_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { SWIN_ATTN_F32 = 0, SWIN_ATTN_BF16 = 1 };

/* Which kernel family serves the bf16 forward/backward.  AUTO picks the
 * tcgen05/TMEM kernel whenever the shape allows it, except for window 12,
 * where the warp-MMA kernels are the faster ones (forward and backward).  WARP_MMA forces the warp-level tensor-core kernels
 * (window 12 only; an error elsewhere). */
enum { SWIN_ATTN_IMPL_AUTO = 0, SWIN_ATTN_IMPL_SIMT = 1, SWIN_ATTN_IMPL_TCGEN05 = 2, SWIN_ATTN_IMPL_WARP_MMA = 3 };

typedef struct {
  int32_t batch;       /* B */
  int32_t height;      /* H: valid rows of q/k/v/o */
  int32_t width;       /* W: valid columns */
  int32_t num_heads;   /* nH; channels C = nH * 32 */
  int32_t window;      /* ws: 7 or 12 */
  int32_t shift;       /* 0 (W-MSA) or 0 < shift < ws (SW-MSA; reference uses ws/2) */
  int32_t dtype;       /* SWIN_ATTN_F32 | SWIN_ATTN_BF16 */
  int32_t impl;        /* SWIN_ATTN_IMPL_* */
  float scale;         /* q scale, head_dim^-0.5 unless qk_scale was given */
  float dropout_p;     /* attention dropout probability (0 in every NewCRFs config) */
  uint64_t seed;       /* Philox seed for dropout */
  int64_t q_pix_stride;
  int64_t k_pix_stride;
  int64_t v_pix_stride;
  int64_t o_pix_stride; /* output (B,H,W,*) pixel stride, >= C */
  int32_t mask_windows; /* nW of the explicit mask (0 when attn_mask is NULL) */
  int32_t reserved;
} swin_attn_shape_t;

/* Forward.  o: (B,H,W,C) un-shifted & cropped.  lse (nullable): fp32
 * (B,Hp,Wp,nH) log-sum-exp of every query row, kept for the backward; entries of
 * padding tokens (y >= H or x >= W) are unspecified. */
int swin_attn_forward(const swin_attn_shape_t* shape,
                      const void* q, const void* k, const void* v,
                      const void* q_pad, const void* k_pad, const void* v_pad,
                      const float* bias_table, /* ((2ws-1)^2, nH) fp32 */
                      const float* attn_mask,  /* nullable: explicit additive mask (mask_windows, N, N) fp32,
                                                  window w uses slice w % mask_windows, exactly as
                                                  WindowAttention.forward(x, mask) does (newcrf_layers.py:130-133).
                                                  NULL: the SW-MSA mask is derived analytically from `shift`. */
                      void* o, float* lse, void* stream);

/* Backward (recomputes the probabilities from q, k, lse).  Requires H, W to be
 * multiples of ws (the caller pads explicitly when it needs gradients, so that
 * pad-token gradients reach the Linear bias through autograd).
 * dq/dk/dv use the q/k/v pixel strides of `shape`; d_bias_table is fp32
 * ((2ws-1)^2, nH) and is OVERWRITTEN (zeroed, then accumulated). */
int swin_attn_backward(const swin_attn_shape_t* shape,
                       const void* q, const void* k, const void* v,
                       const void* o, const void* d_o, const float* lse,
                       const float* bias_table, const float* attn_mask,
                       void* dq, void* dk, void* dv, float* d_bias_table,
                       void* stream);

/* Host-buffer forward: q/k/v/o are HOST pointers (pinned or pageable) laid out
 * as dense (B,H,W,C) tensors; the call stages them through device scratch that
 * the library owns, runs swin_attn_forward and copies o back before returning.
 * This is the end-to-end entry a non-PyTorch caller would bind. */
int swin_attn_forward_host(const swin_attn_shape_t* shape,
                           const void* q, const void* k, const void* v,
                           const void* q_pad, const void* k_pad, const void* v_pad,
                           const float* bias_table, void* o);

/* Host-buffer forward that ALLOCATES NOTHING.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid allocation).  Same arguments, results and error behaviour as swin_attn_forward_host, but the device
 * staging of q / k / v / o, the pad vectors and the bias table lives in a file-scope `static __device__` array of
 * SWIN_ATTN_STATIC_SCRATCH_BYTES inside the library instead of a cudaMalloc'd buffer; the call makes no cudaMalloc /
 * cudaFree and calls no allocating library routine.
 *   - Limit: 4 * align256(B*H*W*C*elem) + 3 * align256(C*elem) + align256((2ws-1)^2 * nH * 4) must not exceed
 *     SWIN_ATTN_STATIC_SCRATCH_BYTES (8 MiB: e.g. bf16, B*H*W*C <= ~1.0 M elements).  Larger shapes are rejected with
 *     an error naming both sizes, before any CUDA call; use swin_attn_forward_host for them.
 *   - Cost: every GPU whose context loads the library holds those 8 MiB until the process exits.
 *   - Sharing: all calls on a device use the SAME array, which keeps the previous call's contents (the entry point
 *     overwrites its inputs in full and clears its output staging first).  Calls of this entry point are serialised by a
 *     mutex and each synchronises before returning, so they never overlap; anything else that were to use the array
 *     concurrently on the GPU would race with them. */
#define SWIN_ATTN_STATIC_SCRATCH_BYTES (8u << 20)
int swin_attn_forward_host_static(const swin_attn_shape_t* shape,
                                  const void* q, const void* k, const void* v,
                                  const void* q_pad, const void* k_pad, const void* v_pad,
                                  const float* bias_table, void* o);

/* Backward on UN-PADDED tensors (height / width need not be multiples of the window): the counterpart of calling
 * swin_attn_forward with pad vectors, i.e. the autograd of pad -> Linear -> attention -> crop of the reference
 * (newcrf_layers.py:212-217,248-249) without materialising the padded tensors.  TMA zero-fills the padding tokens, the
 * kernel substitutes k_pad / v_pad for their K / V rows (q_pad is irrelevant: padded query rows are cropped, their
 * gradient is zero), drops their dq / dk / dv rows and returns instead d_k_pad / d_v_pad (C fp32 each, OVERWRITTEN): the
 * sums of dK / dV over all padding tokens, which the caller adds to the gradient of the projection bias.  k_pad / v_pad may
 * be NULL (zero padding; then the matching d_*_pad must be NULL).  lse: (B, Hp, Wp, nH) as written by the forward.
 * bf16, no mask; window 7 with shift 0 or 3 (tcgen05 kernel) or window 12 with shift 0 or 6 on a grid at least `shift` tall
 * and wide (warp-MMA kernel); returns an error otherwise. */
int swin_attn_backward_unpadded(const swin_attn_shape_t* shape, const void* q, const void* k, const void* v, const void* k_pad,
                                const void* v_pad, const void* o, const void* d_o, const float* lse, const float* bias_table,
                                void* dq, void* dk, void* dv, float* d_bias_table, float* d_k_pad, float* d_v_pad, void* stream);

/* LayerNorm of the residual stream (fp32, or bf16 once an autocast Linear has produced it) over the last
 * dimension, in fp32 arithmetic, written as the (bf16 or fp32) operand of the following Linear: the norm1 / norm2 of the block forwards (newcrf_layers.py:207,253, swin_transformer.py:201,243)
 * fused with the casts torch.autocast inserts around it.  x: (rows, channels) of x_dtype, dense;
 * gamma, beta: (channels) fp32; out: (rows, channels) of out_dtype.  channels in {128, 192, 256, 384, 512, 768,
 * 1024, 1536, 3072}.  Biased variance, eps inside the square root, as nn.LayerNorm. */
int swin_layernorm_forward(const void* x, int32_t x_dtype, const float* gamma, const float* beta, float eps, int64_t rows,
                           int32_t channels, void* out, int32_t out_dtype, void* stream);

/* Residual add fused with the LayerNorm that follows it in the block forwards (`x = shortcut + attn` then `norm2(x)`,
 * newcrf_layers.py:251-253, swin_transformer.py:240-243; `x + mlp(...)` then the next block's `norm1`):
 * sum_out = x + delta in x_dtype (torch's type promotion: fp32 + bf16 -> fp32, bf16 + bf16 -> bf16, rounded once),
 * y_out = LayerNorm(sum_out) as bf16.  delta, y_out: (rows, channels) bf16; x, sum_out: (rows, channels) of x_dtype
 * (may alias).  channels in {128, 192, 256, 384, 512, 768, 1024, 1536}. */
int swin_add_layernorm_forward(const void* x, int32_t x_dtype, const void* delta, const float* gamma, const float* beta, float eps,
                               int64_t rows, int32_t channels, void* sum_out, void* y_out, void* stream);

/* The same with the stochastic-depth gate of the branch folded in (training: `x + drop_path(branch)`,
 * newcrf_layers.py:251,255; timm DropPath): sum_out = x + delta * gate[row / rows_per_sample] as torch.addcmul
 * computes it (one FMA), gate: one fp32 per sample (0 or 1 / keep) or NULL (no gate); x may be fp32 (stage 0) or bf16 (later stages).
 * Backward of the pair (autograd of the add, the DropPath scaling and nn.LayerNorm in one pass):
 *   dx = d_sum + LayerNormBackward(dy; sum, gamma)    (x_dtype)   -- gradient of the residual stream before the add
 *   d_delta = bf16(dx * gate)                                      -- gradient of the bf16 branch output
 *   dgamma, dbeta (channels) fp32, OVERWRITTEN.
 * sum: the forward's sum_out (mean / rstd are recomputed); dy: (rows, channels) bf16 gradient of y_out;
 * d_sum: (rows, channels) x_dtype gradient reaching sum_out through the residual stream. */
int swin_add_layernorm_forward_gated(const void* x, int32_t x_dtype, const void* delta, const float* gamma, const float* beta,
                                     const float* gate, int64_t rows_per_sample, float eps, int64_t rows, int32_t channels, void* sum_out,
                                     void* y_out, void* stream);
int swin_add_layernorm_backward(const void* sum, int32_t x_dtype, const void* dy, const void* d_sum, const float* gamma, const float* gate,
                                int64_t rows_per_sample, float eps, int64_t rows, int32_t channels, void* dx, void* d_delta, float* dgamma,
                                float* dbeta, void* stream);

/* Exact (erf) GELU of the block MLPs, `nn.GELU()` between fc1 and fc2 (newcrf_layers.py:9-27, swin_transformer.py Mlp),
 * on the bf16 output of the autocast fc1: y = x Phi(x), elementwise, x and y: n bf16 elements, n a multiple of 8,
 * 16-byte aligned (may alias).  Backward: dx = dy (Phi(x) + x phi(x)). */
int swin_gelu_forward(const void* x, void* y, int64_t n, void* stream);
int swin_gelu_backward(const void* x, const void* dy, void* dx, int64_t n, void* stream);

/* Per-token Linear with the activation folded into its epilogue, tcgen05 GEMM:
 *   y = act(x W^T + b),  x: (rows, in_features) bf16 dense, weight: (out_features, in_features) bf16 (nn.Linear.weight),
 *   bias: (out_features) bf16, y: (rows, out_features) bf16.
 * activation = SWIN_ACT_GELU replaces `self.act(self.fc1(x))` of the block MLP (newcrf_layers.py:21-23; Mlp of
 * swin_transformer.py) under bf16 autocast: fp32 accumulation, + bias, rounded to bf16 (the value the autocast Linear
 * returns), exact GELU in fp32 on that value, rounded to bf16 -- bit-identical to swin_gelu_forward applied to the
 * bf16 Linear output of the same accumulation.  SWIN_ACT_NONE is the plain Linear.
 * in_features a multiple of 64, out_features a multiple of 256, all pointers 16-byte aligned.  Inference only. */
#define SWIN_ACT_NONE 0
#define SWIN_ACT_GELU 1
int swin_linear_forward(const void* x, const void* weight, const void* bias, void* y, int64_t rows, int32_t in_features,
                        int32_t out_features, int32_t activation, void* stream);

/* Training form of fc1 + GELU: the same GEMM writes BOTH h = x W^T + b (bf16, what the autocast Linear returns and what
 * the GELU backward needs) and y = GELU(h) (bf16), so `self.act(self.fc1(x))` (newcrf_layers.py:21-23) costs one kernel
 * and one read of x / W in training as well; h, y: (rows, out_features) bf16.  y is bit-identical to
 * swin_linear_forward(.., SWIN_ACT_GELU), h to swin_linear_forward(.., SWIN_ACT_NONE). */
int swin_linear_gelu_forward_train(const void* x, const void* weight, const void* bias, void* h, void* y, int64_t rows, int32_t in_features,
                                   int32_t out_features, void* stream);

/* Per-token Linear with the block's residual add folded into its epilogue (same tcgen05 GEMM):
 *   y = residual + (x W^T + b),  residual, y: (rows, out_features) bf16 (may not alias), other operands as above.
 * Replaces `shortcut + drop_path(self.attn(...))` behind `attn.proj` and `x + drop_path(self.mlp(...))` behind `Mlp.fc2`
 * (swin_transformer.py:240-244, newcrf_layers.py:251-255) at inference under bf16 autocast where the residual stream is
 * bf16: the Linear output is rounded to bf16 (what the autocast Linear returns), added to the residual in fp32 and the
 * sum rounded to bf16 (torch's bf16 + bf16) -- bit-identical to the two separate steps.  Inference only. */
int swin_linear_residual_forward(const void* x, const void* weight, const void* bias, const void* residual, void* y, int64_t rows,
                                 int32_t in_features, int32_t out_features, void* stream);

/* Bias gradient of the per-token Linears either side of the attention (qkv / qk / proj, fc1 / fc2;
 * newcrf_layers.py:102,118,147, swin_transformer.py:104,120,142): dbias[c] = sum over rows of dy[row, c].
 * dy: (rows, cols) bf16, dense, 16-byte aligned, cols a multiple of 8; dbias: (cols) fp32, OVERWRITTEN. */
int swin_linear_bias_grad(const void* dy, int64_t rows, int32_t cols, float* dbias, void* stream);

/* Backward of swin_layernorm_forward for training under autocast (autograd of nn.LayerNorm in the reference).
 * dy: (rows, channels) of dy_dtype (bf16 for the bf16-output form, fp32 for the fp32-output form); x: the forward's
 * input (mean / rstd are recomputed).
 * dx: (rows, channels) of x_dtype; dgamma, dbeta: (channels) fp32, OVERWRITTEN (zeroed, then accumulated).
 * channels in {128, 192, 256, 384, 512, 768, 1024, 1536}. */
int swin_layernorm_backward(const void* x, int32_t x_dtype, const void* dy, int32_t dy_dtype, const float* gamma, float eps, int64_t rows,
                            int32_t channels, void* dx, float* dgamma, float* dbeta, void* stream);

/* PatchMerging's 2x2 gather fused with its LayerNorm (swin_transformer.py:271-287: `F.pad` to even H / W, `torch.cat` of
 * x[:, 0::2, 0::2], x[:, 1::2, 0::2], x[:, 0::2, 1::2], x[:, 1::2, 1::2] along channels, `self.norm`): x is the stage
 * output (batch, height, width, channels) in x_dtype, out is bf16 (batch, ceil(height/2) * ceil(width/2), 4 * channels)
 * -- the operand of the `reduction` Linear under bf16 autocast.  channels: 128, 192, 256, 384, 512 or 768 (Swin-B / Swin-L stages).  Bit-identical to
 * the concatenation followed by swin_layernorm_forward. */
int swin_patch_merge_layernorm_forward(const void* x, int32_t x_dtype, const float* gamma, const float* beta, float eps, int32_t batch,
                                       int32_t height, int32_t width, int32_t channels, void* out, void* stream);

/* Measurement hook: arms a pair of cudaEvent_t for the calling thread.  The next attention / linear entry point this
 * thread calls records `start_event` immediately before and `stop_event` immediately after its device work (kernel
 * plus its memsets) on the op's stream -- after the host-side preparation (tensor-map encoding), so a host-bound
 * caller does not count its own launch path as kernel time -- and disarms.  NULL, NULL disarms.  (No counterpart in
 * the reference, which has no timing of its own; used by bench.py's live roofline.) */
void swin_attn_time_next_launch(void* start_event, void* stop_event);

/* Global (un-windowed) multi-head attention that returns its attention map: the core of the Depthformer decoder's
 * SelfAttentionBlock / PostNormSelfAttentionBlock (model/Depthformer/self_attention.py:70-80, :148-158) and of both halves
 * of its Luna blocks (pack: model/Depthformer/luna_layer.py:85-94, unpack: :118-127):
 *     attn = softmax(q k^T * scale) over the last axis, (batch, num_heads, seq, kv_seq) fp32, written to attn_map when non-NULL;
 *     o    = attn v, (batch, seq, num_heads * v_head_dim) in `dtype`.
 * q: (batch, seq, num_heads * qk_head_dim), k: (batch, kv_seq, num_heads * qk_head_dim), v: (batch, kv_seq, num_heads *
 * v_head_dim), head h at channel offset h * head_dim (what `_reshape_qkv` / `_reshape_3d` view), rows `*_row_stride` elements
 * apart (slices of a packed projection are fine), batches seq (kv_seq) rows apart.  kv_seq = 0 means kv_seq = seq.
 * Two kernels: tcgen05 (bf16, both head dims 64, 16 <= kv_seq <= 448, 16-byte aligned pointers, row strides multiples of 8
 * -- the Depthformer shapes; S and the bf16 P live in tensor memory, fp32 softmax) and CUDA cores (fp32 math for both
 * dtypes -- the fp32 mode must hold 1e-5 --, head dims <= 64 with qk_head_dim a multiple of 4, kv_seq <= 1472).  Attention
 * dropout is not applied (the reference returns the un-dropped map; callers that train with attn_drop_prob > 0 apply it to
 * attn_map themselves). */
typedef struct {
  int32_t batch, seq, kv_seq, num_heads, qk_head_dim, v_head_dim, dtype;
  int32_t impl; /* SWIN_ATTN_IMPL_AUTO: tcgen05 kernel when it applies, else CUDA cores; SWIN_ATTN_IMPL_SIMT: CUDA cores */
  float scale;
  int32_t reserved;
  int64_t q_row_stride, k_row_stride, v_row_stride, o_row_stride;
} mha_shape_t;
int mha_attention_forward(const mha_shape_t* shape, const void* q, const void* k, const void* v, void* o, float* attn_map, void* stream);

/* The same attention with an ORDERED bias -- the "pluggable bias source" of the ODA2 depth-order decoder windows
 * (`PreNormOrderedSwinSA.forward`, model/ODA2/oda2_red_order_swin2_decoder.py:88-93, :112-117): every
 * token carries an integer index (its depth bin), and
 *     logit(i, j) = q_i . k_j * scale + index_bias[token_index[i] - token_index[j] + num_emb - 1, head]
 * (the `F.embedding` gather of `depth_embedding` by the pairwise index difference, added after the scaling, before the
 * softmax; no shift mask).  A window of the reference is one batch element here: q, k, v are the projections of the
 * window-partitioned tokens, (num_windows * B, r * r, C).  token_index: (batch, seq) int32 in [0, num_emb); index_bias:
 * (2 * num_emb - 1, num_heads) fp32.  Self-attention only (kv_seq == seq).  The tcgen05 kernel needs num_emb <= 256 on top of
 * its other conditions; otherwise the CUDA-core kernel runs. */
int mha_attention_forward_ordered(const mha_shape_t* shape, const void* q, const void* k, const void* v, const int32_t* token_index,
                                  const float* index_bias, int32_t num_emb, void* o, float* attn_map, void* stream);

/* Measurement probe: median, in microseconds, of `reps` event brackets (as armed by swin_attn_time_next_launch) around an
 * EMPTY 148 x 384 kernel that follows a running kernel in `stream`, with (with_memset != 0) or without the backward's
 * memset in front of it: the fixed cost every event-timed attention launch carries (about 5 us / 8 us on a B200).
 * Allocates and frees a 32 MB scratch buffer; < 0 on a CUDA error.  (No counterpart in the reference; used by bench.py,
 * which reports it next to the roofline.) */
double swin_attn_probe_empty_bracket_us(void* stream, int32_t with_memset, int32_t reps);

/* Number of kernel launches issued by this library since load (all threads). */
uint64_t swin_attn_launch_count(void);

/* 1 if a tcgen05 kernel serves this shape under shape->impl (AUTO or TCGEN05), else 0 (CUDA cores or warp-level MMAs). */
int swin_attn_uses_tcgen05(const swin_attn_shape_t* shape, int backward);

const char* swin_attn_last_error(void);
const char* swin_attn_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SWIN_ATTN_H_ */
