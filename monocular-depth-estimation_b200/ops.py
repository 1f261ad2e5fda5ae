"""``torch.library`` custom ops over the C ABI of ``libswin_attn.so``.

``swin_attn::fwd`` / ``swin_attn::bwd`` take the *packed* output of the
reference's projection Linear -- ``qkv`` (B,H,W,3C) for the encoder
(swin_transformer.py:120) or ``qk`` (B,H,W,2C) plus an external ``v``
(B,H,W,C) for the NeWCRF decoder (newcrf_layers.py:118,143-144) -- in image
order, and return the attention output in image order, cropped to (H, W).
Everything in between (pad, roll, partition, bias gather, mask, softmax, PV,
reverse, un-roll, crop) happens inside one kernel launch.

CUDA only.  There is deliberately no CPU implementation registered: calling
the ops with CPU tensors raises.
"""
from __future__ import annotations

import ctypes
import math
from typing import Optional, Tuple

import torch
from torch import Tensor

from . import _lib

_DTYPES = {torch.float32: _lib.F32, torch.bfloat16: _lib.BF16}

# Optional per-launch timing (bench.py): a list that receives
# (start_event, end_event, algorithmic_bytes, algorithmic_flops, tag) for every
# forward launch, the events recorded on the launching (current) stream.
_kernel_timing = None


def set_kernel_timing(records):
    global _kernel_timing
    _kernel_timing = records


def _ptr(t: Optional[Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _image_layout(t: Tensor) -> Tensor:
    """Dense rows/images with unit channel stride (pixel stride may exceed C)."""
    b, h, w, c = t.shape
    s = t.stride()
    if s[3] == 1 and s[1] == w * s[2] and s[0] == h * s[1] and s[2] % 4 == 0 and s[2] >= c:
        return t
    return t.contiguous()


def _shape(batch, height, width, num_heads, window, shift, dtype, impl, scale, q_ps, k_ps, v_ps, o_ps,
           mask_windows=0, dropout_p=0.0, seed=0) -> _lib.Shape:
    if dtype not in _DTYPES:
        raise RuntimeError(f"swin_attn: dtype {dtype} is not supported (float32 or bfloat16)")
    return _lib.Shape(batch, height, width, num_heads, window, shift, _DTYPES[dtype], impl,
                      float(scale), float(dropout_p), int(seed), q_ps, k_ps, v_ps, o_ps, mask_windows, 0)


def _mask_arg(attn_mask: Optional[Tensor], window: int):
    if attn_mask is None:
        return None, 0
    n = window * window
    if attn_mask.dim() != 3 or attn_mask.shape[1:] != (n, n):
        raise RuntimeError(f"swin_attn: attn_mask must be (nW, {n}, {n}), got {tuple(attn_mask.shape)}")
    return attn_mask.detach().to(torch.float32).contiguous(), attn_mask.shape[0]


def _split(packed: Tensor, v: Optional[Tensor], num_heads: int):
    c = num_heads * 32
    want = 3 * c if v is None else 2 * c
    if packed.dim() != 4 or packed.shape[-1] != want:
        raise RuntimeError(f"swin_attn: packed projection must be (B,H,W,{want}) for {num_heads} heads of dim 32, "
                           f"got {tuple(packed.shape)}")
    if v is not None and (v.shape[:3] != packed.shape[:3] or v.shape[-1] != c):
        raise RuntimeError("self.dim != v.shape[-1]")  # reference assert, newcrf_layers.py:143
    return c


@torch.library.custom_op("swin_attn::fwd", mutates_args=(), device_types="cuda")
def _fwd(packed: Tensor, v: Optional[Tensor], bias_table: Tensor, packed_pad: Optional[Tensor],
         v_pad: Optional[Tensor], window: int, shift: int, num_heads: int, scale: float,
         need_lse: bool, impl: int, attn_mask: Optional[Tensor] = None, dropout_p: float = 0.0,
         seed: int = 0) -> Tuple[Tensor, Tensor]:
    c = _split(packed, v, num_heads)
    mask, mask_nw = _mask_arg(attn_mask, window)
    lib = _lib.load()
    packed = _image_layout(packed)
    b, h, w, _ = packed.shape
    dt = packed.dtype
    ps = packed.stride(2)
    es = packed.element_size()
    base = packed.data_ptr()
    if v is None:
        v_ptr, v_ps = base + 2 * c * es, ps
        vpad_ptr = None if packed_pad is None else packed_pad.data_ptr() + 2 * c * es
    else:
        v = _image_layout(v.to(dt))
        v_ptr, v_ps = v.data_ptr(), v.stride(2)
        if v_pad is not None:
            v_pad = v_pad.to(dt).contiguous()
        vpad_ptr = None if v_pad is None else v_pad.data_ptr()
    if packed_pad is not None:
        packed_pad = packed_pad.to(dt).contiguous()
        if v is None:
            vpad_ptr = packed_pad.data_ptr() + 2 * c * es
    table = bias_table.detach().to(torch.float32).contiguous()
    hp, wp = -(-h // window) * window, -(-w // window) * window
    o = torch.empty((b, h, w, c), dtype=dt, device=packed.device)
    lse = torch.empty((b, hp, wp, num_heads) if need_lse else (0,), dtype=torch.float32, device=packed.device)
    shp = _shape(b, h, w, num_heads, window, shift, dt, impl, scale, ps, ps, v_ps, c, mask_nw, dropout_p, seed)
    qpad_ptr = None if packed_pad is None else packed_pad.data_ptr()
    kpad_ptr = None if packed_pad is None else packed_pad.data_ptr() + c * es
    with torch.cuda.device(packed.device):
        stream = torch.cuda.current_stream().cuda_stream
        if _kernel_timing is not None:
            ev0 = torch.cuda.Event(enable_timing=True)
            ev0.record()
        _lib.check(lib.swin_attn_forward(
            ctypes.byref(shp), ctypes.c_void_p(base), ctypes.c_void_p(base + c * es), ctypes.c_void_p(v_ptr),
            ctypes.c_void_p(qpad_ptr), ctypes.c_void_p(kpad_ptr), ctypes.c_void_p(vpad_ptr),
            _ptr(table), _ptr(mask), _ptr(o), _ptr(lse) if need_lse else None, ctypes.c_void_p(stream)))
        if _kernel_timing is not None:
            ev1 = torch.cuda.Event(enable_timing=True)
            ev1.record()
            tokens = b * h * w
            # algorithmic traffic: read q, k, v and write o once, valid tokens only
            _kernel_timing.append((ev0, ev1, 4 * tokens * c * es + (lse.numel() * 4 if need_lse else 0),
                                   4 * b * hp * wp * window * window * c, ("fwd", window, shift, num_heads, h, w)))
    return o, lse


@_fwd.register_fake
def _(packed, v, bias_table, packed_pad, v_pad, window, shift, num_heads, scale, need_lse, impl, attn_mask=None,
      dropout_p=0.0, seed=0):
    b, h, w, _ = packed.shape
    c = num_heads * 32
    hp, wp = -(-h // window) * window, -(-w // window) * window
    o = packed.new_empty((b, h, w, c))
    lse = packed.new_empty((b, hp, wp, num_heads) if need_lse else (0,), dtype=torch.float32)
    return o, lse


@torch.library.custom_op("swin_attn::bwd", mutates_args=(), device_types="cuda")
def _bwd(d_o: Tensor, packed: Tensor, v: Optional[Tensor], o: Tensor, lse: Tensor, bias_table: Tensor,
         window: int, shift: int, num_heads: int, scale: float,
         attn_mask: Optional[Tensor] = None, impl: int = 0, dropout_p: float = 0.0,
         seed: int = 0) -> Tuple[Tensor, Tensor, Tensor]:
    c = _split(packed, v, num_heads)
    mask, mask_nw = _mask_arg(attn_mask, window)
    lib = _lib.load()
    packed = _image_layout(packed)
    b, h, w, pc = packed.shape
    dt = packed.dtype
    es = packed.element_size()
    ps = packed.stride(2)
    d_o = d_o.to(dt).contiguous()
    o = o.contiguous()
    d_packed = torch.empty((b, h, w, pc), dtype=dt, device=packed.device)
    if v is None:
        v_ptr, v_ps = packed.data_ptr() + 2 * c * es, ps
        d_v = torch.empty((0,), dtype=dt, device=packed.device)
        dv_ptr = d_packed.data_ptr() + 2 * c * es
    else:
        v = _image_layout(v.to(dt))
        v_ptr, v_ps = v.data_ptr(), v.stride(2)
        d_v = torch.empty((b, h, w, c), dtype=dt, device=packed.device)
        dv_ptr = d_v.data_ptr()
    table = bias_table.detach().to(torch.float32).contiguous()
    d_table = torch.empty_like(table)
    base = packed.data_ptr()
    with torch.cuda.device(packed.device):
        stream = torch.cuda.current_stream().cuda_stream
        # Gradients are written with the strides of d_packed / d_v, which can
        # differ from those of the inputs: run q/k and v through two shapes
        # only when they do.
        if ps == pc and (v is None or v_ps == c):
            shp = _shape(b, h, w, num_heads, window, shift, dt, impl, scale, ps, ps, v_ps, c, mask_nw, dropout_p, seed)
            if _kernel_timing is not None:
                ev0 = torch.cuda.Event(enable_timing=True)
                ev0.record()
            _lib.check(lib.swin_attn_backward(
                ctypes.byref(shp), ctypes.c_void_p(base), ctypes.c_void_p(base + c * es), ctypes.c_void_p(v_ptr),
                _ptr(o), _ptr(d_o), _ptr(lse), _ptr(table), _ptr(mask),
                ctypes.c_void_p(d_packed.data_ptr()), ctypes.c_void_p(d_packed.data_ptr() + c * es),
                ctypes.c_void_p(dv_ptr), _ptr(d_table), ctypes.c_void_p(stream)))
            if _kernel_timing is not None:
                ev1 = torch.cuda.Event(enable_timing=True)
                ev1.record()
                # algorithmic traffic of the recompute backward: read q, k, v, dO + lse, write dq, dk, dv
                _kernel_timing.append((ev0, ev1, 7 * b * h * w * c * es + lse.numel() * 4,
                                       10 * b * h * w * window * window * c, ("bwd", window, shift, num_heads, h, w)))
        else:
            packed_c = packed.contiguous()
            v_c = None if v is None else v.contiguous()
            return _bwd(d_o, packed_c, v_c, o, lse, bias_table, window, shift, num_heads, scale, attn_mask, impl,
                        dropout_p, seed)
    return d_packed, d_v, d_table.to(bias_table.dtype)


@_bwd.register_fake
def _(d_o, packed, v, o, lse, bias_table, window, shift, num_heads, scale, attn_mask=None, impl=0, dropout_p=0.0, seed=0):
    d_packed = packed.new_empty(packed.shape)
    d_v = packed.new_empty((0,)) if v is None else packed.new_empty(v.shape)
    return d_packed, d_v, torch.empty_like(bias_table)


def _setup_context(ctx, inputs, output):
    (packed, v, bias_table, packed_pad, v_pad, window, shift, num_heads, scale, need_lse, impl, attn_mask,
     dropout_p, seed) = inputs
    o, lse = output
    ctx.save_for_backward(packed, v, o, lse, bias_table, attn_mask)
    ctx.attrs = (window, shift, num_heads, scale)
    ctx.dropout = (dropout_p, seed)
    ctx.padded = packed.shape[1] % window != 0 or packed.shape[2] % window != 0
    ctx.has_lse = need_lse
    ctx.v_dtype = None if v is None else v.dtype


def _backward(ctx, d_o, d_lse):
    packed, v, o, lse, bias_table, attn_mask = ctx.saved_tensors
    window, shift, num_heads, scale = ctx.attrs
    if ctx.padded or not ctx.has_lse:
        raise RuntimeError(
            "swin_attn: backward needs the forward to have run with need_lse=True on a tensor whose height and "
            "width are multiples of the window (the layers pad explicitly whenever gradients are enabled)")
    d_packed, d_v, d_table = torch.ops.swin_attn.bwd(d_o, packed, v, o, lse, bias_table, window, shift, num_heads, scale,
                                                     attn_mask, _lib.IMPL_AUTO, ctx.dropout[0], ctx.dropout[1])
    if v is None:
        d_v = None
    elif d_v.dtype != ctx.v_dtype:
        d_v = d_v.to(ctx.v_dtype)
    return d_packed, d_v, d_table, None, None, None, None, None, None, None, None, None, None, None


_fwd.register_autograd(_backward, setup_context=_setup_context)


_LN_CHANNELS = (128, 192, 256, 384, 512, 768, 1024, 1536, 3072)


@torch.library.custom_op("swin_attn::layer_norm", mutates_args=(), device_types="cuda")
def _layer_norm(x: Tensor, weight: Tensor, bias: Tensor, eps: float, out_bf16: bool) -> Tensor:
    """LayerNorm over the last dim of an fp32 / bf16 tensor in fp32 arithmetic, written as bf16 (or fp32):
    ``swin_layernorm_forward``."""
    if x.dtype not in _DTYPES:
        raise RuntimeError(f"swin_attn::layer_norm: dtype {x.dtype} is not supported (float32 or bfloat16)")
    lib = _lib.load()
    x = x.contiguous()
    c = x.shape[-1]
    out = torch.empty(x.shape, dtype=torch.bfloat16 if out_bf16 else torch.float32, device=x.device)
    w, b = weight.detach().float().contiguous(), bias.detach().float().contiguous()
    with torch.cuda.device(x.device):
        _lib.check(lib.swin_layernorm_forward(_ptr(x), _DTYPES[x.dtype], _ptr(w), _ptr(b), float(eps), x.numel() // c, c, _ptr(out),
                                              _lib.BF16 if out_bf16 else _lib.F32,
                                              ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return out


@_layer_norm.register_fake
def _(x, weight, bias, eps, out_bf16):
    return x.new_empty(x.shape, dtype=torch.bfloat16 if out_bf16 else torch.float32)


@torch.library.custom_op("swin_attn::layer_norm_bwd", mutates_args=(), device_types="cuda")
def _layer_norm_bwd(dy: Tensor, x: Tensor, weight: Tensor, eps: float) -> Tuple[Tensor, Tensor, Tensor]:
    """``swin_layernorm_backward``: (dx, dweight, dbias) from the output gradient (bf16 or fp32, the dtype of the
    forward's output) and the saved input."""
    lib = _lib.load()
    x = x.contiguous()
    if dy.dtype not in _DTYPES:
        dy = dy.to(torch.float32)
    dy = dy.contiguous()
    c = x.shape[-1]
    dx = torch.empty_like(x)
    w = weight.detach().float().contiguous()
    dw, db = torch.empty_like(w), torch.empty_like(w)
    with torch.cuda.device(x.device):
        _lib.check(lib.swin_layernorm_backward(_ptr(x), _DTYPES[x.dtype], _ptr(dy), _DTYPES[dy.dtype], _ptr(w), float(eps), x.numel() // c, c,
                                               _ptr(dx), _ptr(dw), _ptr(db),
                                               ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return dx, dw, db


@_layer_norm_bwd.register_fake
def _(dy, x, weight, eps):
    w = weight.new_empty(weight.shape, dtype=torch.float32)
    return torch.empty_like(x), w, torch.empty_like(w)


def _ln_setup_context(ctx, inputs, output):
    x, weight, bias, eps, out_bf16 = inputs
    ctx.save_for_backward(x, weight)
    ctx.eps, ctx.out_bf16 = eps, out_bf16
    ctx.param_dtype = weight.dtype


def _ln_backward(ctx, dy):
    x, weight = ctx.saved_tensors
    if x.shape[-1] > 1536:
        raise RuntimeError("swin_attn::layer_norm: the backward covers <= 1536 channels")
    dy = dy.to(torch.bfloat16 if ctx.out_bf16 else torch.float32)
    dx, dw, db = torch.ops.swin_attn.layer_norm_bwd(dy, x, weight, ctx.eps)
    return dx, dw.to(ctx.param_dtype), db.to(ctx.param_dtype), None, None


_layer_norm.register_autograd(_ln_backward, setup_context=_ln_setup_context)


def norm_for_linear(norm: torch.nn.Module, x: Tensor) -> Tensor:
    """``norm(x)`` for a LayerNorm whose output only feeds a Linear.  Under bf16 autocast this is the fused kernel
    (fp32 / bf16 residual in, bf16 out: the values autocast would hand the Linear, and in training the matching
    fused backward); every other case -- fp32 mode, unsupported widths, non-LayerNorm ``norm_layer`` -- is the
    module itself."""
    if (type(norm) is torch.nn.LayerNorm and norm.elementwise_affine and norm.bias is not None and x.is_cuda
            and x.dtype in _DTYPES and torch.is_autocast_enabled()
            and torch.get_autocast_dtype("cuda") == torch.bfloat16 and len(norm.normalized_shape) == 1
            and x.shape[-1] in _LN_CHANNELS):
        needs_grad = torch.is_grad_enabled() and (x.requires_grad or norm.weight.requires_grad or norm.bias.requires_grad)
        if not needs_grad or x.shape[-1] <= 1536:
            return torch.ops.swin_attn.layer_norm(x, norm.weight, norm.bias, norm.eps, True)
    return norm(x)


def norm_fp32(norm: torch.nn.Module, x: Tensor) -> Tensor:
    """``norm(x)`` for a LayerNorm whose output is kept (stage outputs, ``norm_crf``, the patch-embedding norm).
    Under bf16 autocast this is the fused kernel with an fp32 result -- the dtype autocast gives ``nn.LayerNorm`` --
    every other case (fp32 mode, unsupported widths, training above 1536 channels) is the module itself."""
    if (type(norm) is torch.nn.LayerNorm and norm.elementwise_affine and norm.bias is not None and x.is_cuda
            and x.dtype in _DTYPES and torch.is_autocast_enabled()
            and torch.get_autocast_dtype("cuda") == torch.bfloat16 and len(norm.normalized_shape) == 1
            and x.shape[-1] in _LN_CHANNELS):
        needs_grad = torch.is_grad_enabled() and (x.requires_grad or norm.weight.requires_grad or norm.bias.requires_grad)
        if not needs_grad or x.shape[-1] <= 1536:
            return torch.ops.swin_attn.layer_norm(x, norm.weight, norm.bias, norm.eps, False)
    return norm(x)


def window_attention(packed: Tensor, v: Optional[Tensor], bias_table: Tensor, window: int, shift: int,
                     num_heads: int, scale: Optional[float] = None, packed_pad: Optional[Tensor] = None,
                     v_pad: Optional[Tensor] = None, impl: int = _lib.IMPL_AUTO,
                     attn_mask: Optional[Tensor] = None, dropout_p: float = 0.0, seed: Optional[int] = None) -> Tensor:
    """Fused shifted-window attention on image-order tensors (see module docstring).

    ``packed``: (B,H,W,3C) qkv or (B,H,W,2C) qk with ``v`` (B,H,W,C) separate.
    ``packed_pad`` / ``v_pad``: value of the projections at padding tokens
    (the Linear bias; None = zeros).  ``dropout_p`` > 0 drops attention probabilities (nn.Dropout on the softmax
    output in the reference); ``seed`` defaults to a draw from torch's CPU generator.  Returns (B,H,W,C).
    """
    if not 0 <= shift < window:
        raise AssertionError("shift_size must in 0-window_size")  # newcrf_layers.py:180
    if scale is None:
        scale = 32 ** -0.5
    need_grad = torch.is_grad_enabled() and (
        packed.requires_grad or bias_table.requires_grad or (v is not None and v.requires_grad))
    if dropout_p > 0.0 and seed is None:
        seed = int(torch.randint(0, 2 ** 62, (1,)).item())   # CPU generator: reproducible under torch.manual_seed, no sync
    o, _ = torch.ops.swin_attn.fwd(packed, v, bias_table, packed_pad, v_pad, window, shift, num_heads,
                                   float(scale), need_grad, impl, attn_mask, float(dropout_p), int(seed or 0))
    return o


def window_attention_host(q, k, v, bias_table, window: int, shift: int, num_heads: int,
                          scale: Optional[float] = None, q_pad=None, k_pad=None, v_pad=None,
                          impl: int = _lib.IMPL_AUTO) -> Tensor:
    """Host-buffer entry (``swin_attn_forward_host``): CPU tensors in, CPU tensor
    out, host<->device copies inside the call.  q, k, v: dense (B,H,W,C)."""
    lib = _lib.load()
    q, k, v = q.contiguous(), k.contiguous(), v.contiguous()
    b, h, w, c = q.shape
    if scale is None:
        scale = 32 ** -0.5
    o = torch.empty_like(q)
    table = bias_table.detach().to(torch.float32).contiguous()
    pads = [None if t is None else t.to(q.dtype).contiguous() for t in (q_pad, k_pad, v_pad)]
    shp = _shape(b, h, w, num_heads, window, shift, q.dtype, impl, scale, c, c, c, c)
    _lib.check(lib.swin_attn_forward_host(ctypes.byref(shp), _ptr(q), _ptr(k), _ptr(v), _ptr(pads[0]),
                                          _ptr(pads[1]), _ptr(pads[2]), _ptr(table), _ptr(o)))
    return o
