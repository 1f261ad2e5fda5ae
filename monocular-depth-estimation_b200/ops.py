"""``torch.library`` ops over the C ABI of ``libswin_attn.so`` (C++ dispatcher kernels, csrc/torch_binding.cpp).

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
import os
from typing import Optional, Tuple

import torch
from torch import Tensor

from . import _lib

_DTYPES = {torch.float32: _lib.F32, torch.bfloat16: _lib.BF16}

TORCH_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libswin_attn_torch.so")


def _load_ops():
    """Register swin_attn::{fwd,bwd,layer_norm,layer_norm_bwd}: C++ dispatcher kernels (CUDA, Meta, Autograd keys) in
    libswin_attn_torch.so, built from csrc/torch_binding.cpp over the C ABI.  No fallback: a missing build raises."""
    _lib.load()
    if not os.path.exists(TORCH_LIB_PATH):
        raise RuntimeError(
            f"{TORCH_LIB_PATH} is missing: the torch binding has not been built (run __graft_entry__.build()); "
            "this package has no CPU or PyTorch fallback")
    torch.ops.load_library(TORCH_LIB_PATH)
    lib = ctypes.CDLL(TORCH_LIB_PATH)
    lib.swin_torch_set_timing.argtypes = [ctypes.c_int]
    lib.swin_torch_take_timing.argtypes = [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int64), ctypes.c_int]
    lib.swin_torch_take_timing.restype = ctypes.c_int
    lib.swin_torch_pending_timing.restype = ctypes.c_int
    lib.swin_torch_read_graph_timing.argtypes = [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int64), ctypes.c_int]
    lib.swin_torch_read_graph_timing.restype = ctypes.c_int
    lib.swin_torch_graph_timing_count.restype = ctypes.c_int
    lib.swin_torch_clear_graph_timing.restype = None
    return lib


_torch_lib = _load_ops()

# Optional per-launch timing (bench.py): while a list is installed, every attention launch is bracketed by CUDA
# events on the launching stream (inside the C++ op); ``set_kernel_timing(None)`` synchronises on them and appends
# (milliseconds, algorithmic_bytes, algorithmic_flops, tag) per launch, tag = ("fwd"|"bwd", window, shift, heads, H, W).
_kernel_timing = None


def set_kernel_timing(records):
    global _kernel_timing
    if records is None:
        _torch_lib.swin_torch_set_timing(0)
        if _kernel_timing is not None:
            n = _torch_lib.swin_torch_pending_timing()
            ms = (ctypes.c_double * max(n, 1))()
            meta = (ctypes.c_int64 * (8 * max(n, 1)))()
            n = _torch_lib.swin_torch_take_timing(ms, meta, n)
            for i in range(n):
                m = meta[8 * i:8 * i + 8]
                _kernel_timing.append((ms[i], m[0], m[1], (("fwd", "bwd", "lin")[m[2]],) + tuple(m[3:8])))
        _kernel_timing = None
    else:
        _kernel_timing = records
        _torch_lib.swin_torch_set_timing(1)


def capture_graph_timing(on: bool):
    """Graph mode of the per-launch timing (bench.py): while on, every attention / linear launch that is being CAPTURED
    into a CUDA graph is bracketed by two external event-record nodes of that graph (eager launches are not timed).
    After each replay has finished, ``read_graph_timing()`` returns the launches' times inside that replay."""
    _torch_lib.swin_torch_set_timing(2 if on else 0)


def read_graph_timing():
    """Records of the instrumented graphs, as timed by their LAST replay (synchronise first): same tuples as
    ``set_kernel_timing``; a pair whose events were not recorded yet has ms = -1."""
    n = _torch_lib.swin_torch_graph_timing_count()
    ms = (ctypes.c_double * max(n, 1))()
    meta = (ctypes.c_int64 * (8 * max(n, 1)))()
    n = _torch_lib.swin_torch_read_graph_timing(ms, meta, n)
    out = []
    for i in range(n):
        m = meta[8 * i:8 * i + 8]
        out.append((ms[i], m[0], m[1], (("fwd", "bwd", "lin")[m[2]],) + tuple(m[3:8])))
    return out


def clear_graph_timing():
    """Drop the graph-mode records (destroy their instrumented graphs first, or never replay them again)."""
    _torch_lib.swin_torch_clear_graph_timing()


def _ptr(t: Optional[Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _shape(batch, height, width, num_heads, window, shift, dtype, impl, scale, q_ps, k_ps, v_ps, o_ps,
           mask_windows=0, dropout_p=0.0, seed=0) -> _lib.Shape:
    if dtype not in _DTYPES:
        raise RuntimeError(f"swin_attn: dtype {dtype} is not supported (float32 or bfloat16)")
    return _lib.Shape(batch, height, width, num_heads, window, shift, _DTYPES[dtype], impl,
                      float(scale), float(dropout_p), int(seed), q_ps, k_ps, v_ps, o_ps, mask_windows, 0)


_LN_CHANNELS = (128, 192, 256, 384, 512, 768, 1024, 1536, 3072)


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


def patch_merge_norm(norm: torch.nn.Module, x: Tensor, height: int, width: int) -> Optional[Tensor]:
    """``norm(cat(2x2 neighbourhood of x))`` of PatchMerging (swin_transformer.py:271-287) as one kernel that gathers
    the four pixels while normalising -- the concatenated (B, H/2*W/2, 4C) tensor is never written.  Inference under
    bf16 autocast only (bit-identical to the concatenation followed by ``norm_for_linear``); returns None when the
    fused kernel does not apply and the caller takes the reference's steps."""
    if (type(norm) is torch.nn.LayerNorm and norm.elementwise_affine and norm.bias is not None and x.is_cuda
            and x.dtype in _DTYPES and not torch.is_grad_enabled() and torch.is_autocast_enabled()
            and torch.get_autocast_dtype("cuda") == torch.bfloat16 and len(norm.normalized_shape) == 1
            and x.shape[-1] in (128, 192, 256, 384, 512, 768) and norm.normalized_shape[0] == 4 * x.shape[-1]):
        return torch.ops.swin_attn.patch_merge_norm(x, height, width, norm.weight, norm.bias, norm.eps)
    return None


def linear(lin: torch.nn.Module, x: Tensor) -> Tensor:
    """``lin(x)`` for the per-token Linears of the blocks.  In training under bf16 autocast this is one C++ autograd
    node computing what autocast computes (bf16 operands, cuBLAS addmm) whose backward sums the bias gradient with the
    fused column-sum kernel instead of ATen's generic reduction; every other case is the module itself."""
    if (type(lin) is torch.nn.Linear and lin.bias is not None and x.is_cuda and torch.is_grad_enabled()
            and torch.is_autocast_enabled() and torch.get_autocast_dtype("cuda") == torch.bfloat16
            and lin.out_features % 8 == 0 and x.dtype in _DTYPES and lin.weight.dtype in _DTYPES
            and (x.requires_grad or lin.weight.requires_grad)):
        return torch.ops.swin_attn.linear(x, lin.weight, lin.bias)
    return lin(x)


def linear_gelu(lin: torch.nn.Module, act: torch.nn.Module, x: Tensor) -> Tensor:
    """``act(lin(x))`` of the block MLP (fc1 -> GELU, newcrf_layers.py:21-23).  Inference under bf16 autocast (or with
    bf16 parameters after ``harness.prepare_inference``) runs ONE tcgen05 GEMM whose epilogue adds the bias, rounds to
    bf16 as the autocast Linear does, applies the exact GELU and rounds again: the hidden tensor is written once.
    Training under bf16 autocast runs the same GEMM with both h and GELU(h) written (``swin_attn::linear_gelu``, one
    autograd node: GELU backward on the saved h, then the Linear backward).  Every other case is
    ``gelu(act, linear(lin, x))``."""
    if (type(lin) is torch.nn.Linear and lin.bias is not None and type(act) is torch.nn.GELU and act.approximate == "none"
            and x.is_cuda and not torch.is_grad_enabled() and lin.in_features % 64 == 0 and lin.out_features % 256 == 0
            and (x.dtype == torch.bfloat16 or x.dtype == torch.float32)
            and ((torch.is_autocast_enabled() and torch.get_autocast_dtype("cuda") == torch.bfloat16)
                 or (x.dtype == torch.bfloat16 and lin.weight.dtype == torch.bfloat16))):
        bf = torch.bfloat16
        return torch.ops.swin_attn.linear_act(x.to(bf), lin.weight.to(bf), lin.bias.to(bf), 1)
    if (type(lin) is torch.nn.Linear and lin.bias is not None and type(act) is torch.nn.GELU and act.approximate == "none"
            and x.is_cuda and torch.is_grad_enabled() and lin.in_features % 64 == 0 and lin.out_features % 256 == 0
            and x.dtype in _DTYPES and lin.weight.dtype in _DTYPES and (x.requires_grad or lin.weight.requires_grad)
            and torch.is_autocast_enabled() and torch.get_autocast_dtype("cuda") == torch.bfloat16):
        # training: one autograd node; its forward is the same GEMM writing h (for the GELU backward) and GELU(h)
        return torch.ops.swin_attn.linear_gelu(x, lin.weight, lin.bias)
    return gelu(act, linear(lin, x))


def linear_residual(lin: torch.nn.Module, x: Tensor, residual: Tensor) -> Optional[Tensor]:
    """``residual + lin(x)`` -- the residual add behind ``attn.proj`` / ``Mlp.fc2`` (swin_transformer.py:240-244,
    newcrf_layers.py:251-255) -- as ONE tcgen05 GEMM whose epilogue rounds the Linear output to bf16 (as the autocast
    Linear does), adds the bf16 residual stream in fp32 and rounds the sum to bf16 (torch's bf16 + bf16): the branch
    output is never written.  Inference with a bf16 stream only; returns None when the fused kernel does not apply
    (fp32 stream of stage 0, widths that are not a multiple of 256, training) and the caller takes the separate steps."""
    if (type(lin) is torch.nn.Linear and lin.bias is not None and x.is_cuda and not torch.is_grad_enabled()
            and residual.dtype == torch.bfloat16 and lin.in_features % 64 == 0 and lin.out_features % 256 == 0
            and (x.dtype == torch.bfloat16 or x.dtype == torch.float32)
            and residual.shape[-1] == lin.out_features and residual.numel() // lin.out_features == x.numel() // lin.in_features
            and ((torch.is_autocast_enabled() and torch.get_autocast_dtype("cuda") == torch.bfloat16)
                 or (x.dtype == torch.bfloat16 and lin.weight.dtype == torch.bfloat16))):
        bf = torch.bfloat16
        return torch.ops.swin_attn.linear_residual(x.to(bf), lin.weight.to(bf), lin.bias.to(bf), residual)
    return None


def gelu(act: torch.nn.Module, x: Tensor) -> Tensor:
    """``act(x)`` for the MLP activation: the fused exact-GELU kernel when ``act`` is ``nn.GELU()`` (erf form) and ``x``
    is the bf16 output of an autocast Linear; otherwise the module itself."""
    if (type(act) is torch.nn.GELU and act.approximate == "none" and x.is_cuda and x.dtype == torch.bfloat16
            and x.numel() % 8 == 0):
        return torch.ops.swin_attn.gelu(x)
    return act(x)


def residual(x: Tensor, delta: Tensor, scale: Optional[Tensor] = None) -> Tensor:
    """``x + delta`` or, with a stochastic-depth gate, ``x + delta * scale`` in one kernel (DropPath.gate)."""
    if scale is None:
        return x + delta
    if scale.dtype != delta.dtype:   # the model-wide draw is fp32 (layers.draw_drop_path_gates); keep the stream's dtype
        scale = scale.to(delta.dtype)
    return torch.addcmul(x, delta, scale)


def add_norm(norm: torch.nn.Module, x: Tensor, delta: Tensor, scale: Optional[Tensor] = None) -> Tuple[Tensor, Tensor]:
    """``x + delta`` and ``norm(x + delta)`` as the operand of the following Linear: the residual add of a block
    with the LayerNorm behind it (``norm2`` of the same block, ``norm1`` of the next one).  One fused kernel under bf16
    autocast (bf16 ``delta``, <= 1536 channels), in training one autograd node whose backward is one kernel as well
    (LayerNorm backward + the gradient arriving through the residual stream + the bf16 / gated gradient of the branch);
    otherwise the two separate steps.  ``scale`` is the stochastic-depth gate of ``delta`` (training only, one value
    per sample; fp32 stream of stage 0 or the bf16 stream behind the first PatchMerging -- the fused add is the single
    fp32 FMA + rounding ``torch.addcmul`` performs, bit for bit)."""
    if (type(norm) is torch.nn.LayerNorm and norm.elementwise_affine and norm.bias is not None and x.is_cuda
            and x.dtype in _DTYPES and delta.dtype == torch.bfloat16 and delta.shape == x.shape
            and torch.is_autocast_enabled() and torch.get_autocast_dtype("cuda") == torch.bfloat16
            and len(norm.normalized_shape) == 1 and x.shape[-1] in _LN_CHANNELS and x.shape[-1] <= 1536
            and (scale is None or (x.dim() >= 2 and scale.numel() == x.shape[0]))):
        out = torch.ops.swin_attn.add_layer_norm(x, delta, norm.weight, norm.bias, norm.eps, scale)
        return out[0], out[1]
    x = residual(x, delta, scale)
    return x, norm_for_linear(norm, x)


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
        packed.requires_grad or bias_table.requires_grad or (v is not None and v.requires_grad)
        or (packed_pad is not None and packed_pad.requires_grad) or (v_pad is not None and v_pad.requires_grad))
    if dropout_p > 0.0 and seed is None:
        # CPU generator: reproducible under torch.manual_seed, no device sync.  The rank is folded in so that identically
        # seeded data-parallel replicas do not share one mask.  The seed is a host scalar: a CUDA graph would bake it in
        # (harness.GraphedTrainStep refuses models with attn_drop > 0 for that reason).
        seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            seed = (seed + 0x9E3779B97F4A7C15 * (torch.distributed.get_rank() + 1)) % (2 ** 63)
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
