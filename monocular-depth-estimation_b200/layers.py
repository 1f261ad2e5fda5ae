"""Drop-in ``nn.Module``s for the reference's window-attention stack.

Constructor arguments, ``forward`` signatures, return values and
``state_dict`` keys equal those of

    model/NewCRFs/newcrf_layers.py   WindowAttention :62, CRFBlock :152,
                                     BasicCRFLayer :260, NewCRF :366, Mlp :9
    model/NewCRFs/swin_transformer.py WindowAttention :64 (here
                                     ``EncoderWindowAttention``),
                                     SwinTransformerBlock :147, PatchMerging :249,
                                     BasicLayer :292, PatchEmbed :396

so a reference checkpoint loads with ``strict=True`` and the callers
(``NewCRFDepth``) run unchanged.  What differs is the execution: between the
qk/qkv Linear and the proj Linear everything -- pad, cyclic shift, window
partition, bias gather, SW-MSA mask, softmax, PV, window reverse, un-shift,
crop -- is ONE launch of the sm_100a kernel behind ``ops.window_attention``;
no rolled / partitioned / padded copy and no (nW, N, N) mask is materialised.
The kernels are the only implementation: there is no PyTorch fallback.
"""
from __future__ import annotations

from typing import Optional

import os

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops


def _pair(v):
    return tuple(v) if isinstance(v, (tuple, list)) else (v, v)


class DropPath(nn.Module):
    """Per-sample stochastic depth (the reference takes it from timm 0.5.4)."""

    def __init__(self, drop_prob: float = 0.0):
        super().__init__()
        self.drop_prob = float(drop_prob)
        self._banked = None   # this forward's gates (a block uses its DropPath twice), drawn with every other DropPath's in one go

    def forward(self, x):
        if self.drop_prob == 0.0 or not self.training:
            return x
        return x * self.gate(x).to(x.dtype)   # one pass over x (timm: x.div(keep) * gate)

    def gate(self, x):
        """Per-sample scale 0 | 1/keep of shape (B, 1, ..), or None when the path is never dropped (eval, rate 0):
        lets the caller fold the scaling into the residual add (``ops.residual``).  fp32 when it comes from the
        model-wide draw (what the fused add + LayerNorm kernels take), else the dtype of ``x``; the value 1/keep is
        rounded to the dtype of ``x`` either way."""
        if self.drop_prob == 0.0 or not self.training:
            return None
        if self._banked:
            g = self._banked.pop()
            if g.shape[0] == x.shape[0] and g.device == x.device:
                return g.view((x.shape[0],) + (1,) * (x.ndim - 1))
        keep = 1.0 - self.drop_prob
        return x.new_empty((x.shape[0],) + (1,) * (x.ndim - 1)).bernoulli_(keep) / keep


def draw_drop_path_gates(root: nn.Module, batch: int, device, dtype) -> None:
    """The stochastic-depth gates of EVERY DropPath under ``root`` for one training forward, drawn by three kernels
    (uniform numbers, compare, scale) instead of two or three per module and two more per module for the fp32 copies the
    fused add + LayerNorm nodes take -- ~250 launches of a Swin-L training step.  Each module's ``gate`` then returns its
    rows (two per module and forward); the per-sample statistics are those of ``bernoulli_(keep) / keep`` per module (timm 0.5.4 ``drop_path``), the
    value 1/keep rounded to ``dtype`` as the per-module draw on a ``dtype`` tensor rounds it."""
    cache = root.__dict__.get("_drop_path_cache")
    if cache is None or cache[1].device != torch.device(device) or cache[3] != dtype:
        mods = [m for m in root.modules() if isinstance(m, DropPath) and m.drop_prob > 0.0]
        keep = torch.tensor([1.0 - m.drop_prob for m in mods], dtype=torch.float32)
        inv = (1.0 / keep).to(dtype).float()
        cache = (mods, keep.to(device).unsqueeze(1), inv.to(device).unsqueeze(1), dtype)
        root.__dict__["_drop_path_cache"] = cache
    mods, keep, inv, _ = cache
    if not mods:
        return
    # two independent gates per module: a block calls its DropPath on the attention branch and on the MLP branch
    bank = (torch.rand(2, len(mods), batch, device=keep.device) < keep).float() * inv
    for i, m in enumerate(mods):
        m._banked = [bank[1, i], bank[0, i]]


def _gate(drop_path, x):
    return drop_path.gate(x) if isinstance(drop_path, DropPath) else None


# Residual adds in the proj / fc2 GEMM epilogues (ops.linear_residual, csrc/swin_linear_tc.cu).  Off by default: measured
# on the B200 the 128 x 256-tile, one-CTA tcgen05 GEMM loses more against cuBLAS on these shapes
# (N = C: three n-blocks, 2.1 waves; 0.90 vs 1.52 PFLOP/s at 13376 x 3072 x 768) than the saved pass over the stream
# gains, so the default keeps cuBLAS + the add fused into the following LayerNorm (ops.add_norm).
FUSE_RESIDUAL_EPILOGUE = os.environ.get("SWIN_FUSE_RESIDUAL", "0") == "1"


def _linear_into_stream(lin, drop, drop_path, x, t):
    """``x + drop_path(drop(lin(t)))`` with the add in the GEMM epilogue (``ops.linear_residual``) when nothing random
    sits between the Linear and the add (inference; bf16 residual stream); None otherwise."""
    if not FUSE_RESIDUAL_EPILOGUE:
        return None
    if torch.is_grad_enabled() or (drop.training and drop.p > 0.0) or \
            (isinstance(drop_path, DropPath) and drop_path.training and drop_path.drop_prob > 0.0):
        return None
    return ops.linear_residual(lin, t, x)


class Mlp(nn.Module):
    """fc1 -> act -> drop -> fc2 -> drop (newcrf_layers.py:9-27)."""

    def __init__(self, in_features, hidden_features=None, out_features=None, act_layer=nn.GELU, drop=0.0):
        super().__init__()
        self.fc1 = nn.Linear(in_features, hidden_features or in_features)
        self.act = act_layer()
        self.fc2 = nn.Linear(hidden_features or in_features, out_features or in_features)
        self.drop = nn.Dropout(drop)

    def hidden(self, x):
        return self.drop(ops.linear_gelu(self.fc1, self.act, x))

    def forward(self, x):
        return self.drop(ops.linear(self.fc2, self.hidden(x)))


def _relative_position_index(wh: int, ww: int) -> torch.Tensor:
    ys, xs = torch.meshgrid(torch.arange(wh), torch.arange(ww), indexing="ij")
    ys, xs = ys.flatten(), xs.flatten()
    return (ys[:, None] - ys[None, :] + wh - 1) * (2 * ww - 1) + (xs[:, None] - xs[None, :] + ww - 1)


class _WindowAttentionBase(nn.Module):
    """Parameters and buffers shared by both reference flavours."""

    def __init__(self, dim, window_size, num_heads, out_dim, qk_scale, attn_drop, proj_drop, in_proj, qkv_bias):
        super().__init__()
        self.dim = dim
        self.window_size = _pair(window_size)
        if self.window_size[0] != self.window_size[1]:
            raise ValueError("only square windows are supported (the reference only builds square ones)")
        self.num_heads = num_heads
        head_dim = dim // num_heads
        if head_dim * num_heads != dim or head_dim != 32:
            raise ValueError(f"the sm_100a kernels need head_dim == 32 (dim={dim}, num_heads={num_heads})")
        self.scale = qk_scale or head_dim ** -0.5
        wh, ww = self.window_size
        self.relative_position_bias_table = nn.Parameter(torch.zeros((2 * wh - 1) * (2 * ww - 1), num_heads))
        self.register_buffer("relative_position_index", _relative_position_index(wh, ww))
        # registered in the reference's order (table, qk | qkv, proj: newcrf_layers.py:86-105, swin_transformer.py:88-107)
        # so that model.parameters() -- and with it an index-keyed optimizer state_dict -- lines up with the reference's
        name, factor = in_proj
        setattr(self, name, nn.Linear(dim, dim * factor, bias=qkv_bias))
        # applied to the softmax output inside the kernel while training (0 in every NewCRFs configuration)
        self.attn_drop = nn.Dropout(attn_drop)
        self.proj = nn.Linear(out_dim, out_dim)
        self.proj_drop = nn.Dropout(proj_drop)
        nn.init.trunc_normal_(self.relative_position_bias_table, std=0.02)
        self.softmax = nn.Softmax(dim=-1)  # kept for module-tree parity; the kernel does the softmax

    def _fused_pad_backward(self, shift, height, width):
        """True when the backward kernel takes un-padded tensors (bf16 autocast, no attention dropout; window 7, shift
        0 / 3: tcgen05 kernel; window 12, shift 0 / 6 on a grid at least ``shift`` tall and wide: warp-MMA kernel): training
        then runs the same call as inference -- no F.pad before the Linear, no crop after the attention -- and the
        gradient of the pad tokens reaches the Linear bias through the op's ``packed_pad`` input."""
        ws = self.window_size[0]
        shape_ok = (ws == 7 and shift in (0, 3)) or (ws == 12 and shift in (0, 6) and height >= shift and width >= shift)
        return (shape_ok and self._drop_p() == 0.0 and torch.is_autocast_enabled()
                and torch.get_autocast_dtype("cuda") == torch.bfloat16)

    def _drop_p(self):
        return self.attn_drop.p if self.training else 0.0

    def _windows_as_images(self, t):
        ws = self.window_size[0]
        return t.reshape(t.shape[0], ws, ws, t.shape[-1])


class WindowAttention(_WindowAttentionBase):
    """NeWCRF decoder attention: q, k from ``x``; ``v`` external (newcrf_layers.py:62-149)."""

    def __init__(self, dim, window_size, num_heads, v_dim, qkv_bias=True, qk_scale=None, attn_drop=0., proj_drop=0.):
        super().__init__(dim, window_size, num_heads, v_dim, qk_scale, attn_drop, proj_drop, ("qk", 2), qkv_bias)

    def forward(self, x, v, mask=None):
        """x, v: (num_windows*B, N, C) window-order tokens; mask: (nW, N, N) or None."""
        b_, n, c = x.shape
        assert self.dim == v.shape[-1], "self.dim != v.shape[-1]"
        qk = self._windows_as_images(self.qk(x))
        o = ops.window_attention(qk, self._windows_as_images(v), self.relative_position_bias_table,
                                 self.window_size[0], 0, self.num_heads, self.scale, attn_mask=mask,
                                 dropout_p=self._drop_p())
        return self.proj_drop(self.proj(o.reshape(b_, n, c)))

    def forward_image(self, x_img, v_img, shift):
        return self.proj_drop(ops.linear(self.proj, self.attend_image(x_img, v_img, shift)))

    def attend_image(self, x_img, v_img, shift):
        """x_img, v_img: (B, H, W, C) image-order (x already normalised) -> attention output before ``proj``.  Pad
        tokens take the Linear bias for q/k and zero for v, as in the reference
        where ``F.pad`` precedes the Linear (newcrf_layers.py:212-216)."""
        ws = self.window_size[0]
        if torch.is_grad_enabled() and not self._fused_pad_backward(shift, x_img.shape[1], x_img.shape[2]) and (
                x_img.requires_grad or v_img.requires_grad or self.relative_position_bias_table.requires_grad):
            h, w = x_img.shape[1:3]
            pb, pr = (-h) % ws, (-w) % ws
            if pb or pr:
                x_img = F.pad(x_img, (0, 0, 0, pr, 0, pb))
                v_img = F.pad(v_img, (0, 0, 0, pr, 0, pb))
            o = ops.window_attention(ops.linear(self.qk, x_img), v_img, self.relative_position_bias_table, ws, shift,
                                     self.num_heads, self.scale, dropout_p=self._drop_p())
            o = o[:, :h, :w, :]
        else:
            o = ops.window_attention(ops.linear(self.qk, x_img), v_img, self.relative_position_bias_table, ws, shift,
                                     self.num_heads, self.scale, packed_pad=self.qk.bias, v_pad=None,
                                     dropout_p=self._drop_p())
        return o


class EncoderWindowAttention(_WindowAttentionBase):
    """Swin encoder attention with fused qkv (swin_transformer.py:64-144)."""

    def __init__(self, dim, window_size, num_heads, qkv_bias=True, qk_scale=None, attn_drop=0., proj_drop=0.):
        super().__init__(dim, window_size, num_heads, dim, qk_scale, attn_drop, proj_drop, ("qkv", 3), qkv_bias)

    def forward(self, x, mask=None):
        b_, n, c = x.shape
        qkv = self._windows_as_images(self.qkv(x))
        o = ops.window_attention(qkv, None, self.relative_position_bias_table, self.window_size[0], 0,
                                 self.num_heads, self.scale, attn_mask=mask, dropout_p=self._drop_p())
        return self.proj_drop(self.proj(o.reshape(b_, n, c)))

    def forward_image(self, x_img, shift):
        return self.proj_drop(ops.linear(self.proj, self.attend_image(x_img, shift)))

    def attend_image(self, x_img, shift):
        """(B, H, W, C) normalised image-order tokens -> attention output before ``proj``."""
        ws = self.window_size[0]
        if torch.is_grad_enabled() and not self._fused_pad_backward(shift, x_img.shape[1], x_img.shape[2]) and (
                x_img.requires_grad or self.relative_position_bias_table.requires_grad):
            h, w = x_img.shape[1:3]
            pb, pr = (-h) % ws, (-w) % ws
            if pb or pr:
                x_img = F.pad(x_img, (0, 0, 0, pr, 0, pb))
            o = ops.window_attention(ops.linear(self.qkv, x_img), None, self.relative_position_bias_table, ws, shift,
                                     self.num_heads, self.scale, dropout_p=self._drop_p())
            o = o[:, :h, :w, :]
        else:
            o = ops.window_attention(ops.linear(self.qkv, x_img), None, self.relative_position_bias_table, ws, shift,
                                     self.num_heads, self.scale, packed_pad=self.qkv.bias, dropout_p=self._drop_p())
        return o


def _block_tail(blk, x, o):
    """Everything of a block behind the attention core (swin_transformer.py:228-246, newcrf_layers.py:239-257):
    ``x = x + drop_path(proj(o))``, ``norm2``, MLP.  Returns ``(stream, pending, scale)``: the MLP branch still to be
    added (with its stochastic-depth gate), or ``pending = None`` when the add already happened in the fc2 GEMM.
    At inference on a bf16 stream both residual adds ride in the epilogues of the proj / fc2 GEMMs
    (``ops.linear_residual``); otherwise the add is fused with the LayerNorm behind it (``ops.add_norm``)."""
    attn = blk.attn
    s = _linear_into_stream(attn.proj, attn.proj_drop, blk.drop_path, x, o)
    if s is not None:
        x, y2 = s, ops.norm_for_linear(blk.norm2, s)
    else:
        a = attn.proj_drop(ops.linear(attn.proj, o))
        x, y2 = ops.add_norm(blk.norm2, x, a, _gate(blk.drop_path, a))
    hidden = blk.mlp.hidden(y2)
    s = _linear_into_stream(blk.mlp.fc2, blk.mlp.drop, blk.drop_path, x, hidden)
    if s is not None:
        return s, None, None
    m = blk.mlp.drop(ops.linear(blk.mlp.fc2, hidden))
    return x, m, _gate(blk.drop_path, m)


class CRFBlock(nn.Module):
    """norm1 -> (S)W-MSA(x, v) -> +shortcut -> norm2 -> MLP -> + (newcrf_layers.py:152-257).

    ``mask_matrix`` is accepted for signature parity; the kernel derives the
    identical SW-MSA mask from ``shift_size`` analytically.
    """

    def __init__(self, dim, num_heads, v_dim, window_size=7, shift_size=0, mlp_ratio=4., qkv_bias=True,
                 qk_scale=None, drop=0., attn_drop=0., drop_path=0., act_layer=nn.GELU, norm_layer=nn.LayerNorm):
        super().__init__()
        self.dim, self.num_heads, self.v_dim = dim, num_heads, v_dim
        self.window_size, self.shift_size, self.mlp_ratio = window_size, shift_size, mlp_ratio
        assert 0 <= self.shift_size < self.window_size, "shift_size must in 0-window_size"
        self.norm1 = norm_layer(dim)
        self.attn = WindowAttention(dim, window_size=_pair(window_size), num_heads=num_heads, v_dim=v_dim,
                                    qkv_bias=qkv_bias, qk_scale=qk_scale, attn_drop=attn_drop, proj_drop=drop)
        self.drop_path = DropPath(drop_path) if drop_path > 0. else nn.Identity()
        self.norm2 = norm_layer(v_dim)
        self.mlp = Mlp(in_features=v_dim, hidden_features=int(v_dim * mlp_ratio), act_layer=act_layer, drop=drop)
        self.H = None
        self.W = None

    def forward(self, x, v, mask_matrix=None):
        x, pending, scale = self.forward_parts(x, v, ops.norm_for_linear(self.norm1, x))
        return x if pending is None else ops.residual(x, pending, scale)

    def forward_parts(self, x, v, y):
        """The block without its last residual add: ``y`` is ``norm1(x)``; returns the stream after the attention
        residual and the (drop-path'ed) MLP branch still to be added, so that a stage can fuse that add with the
        next block's ``norm1`` (BasicCRFLayer.forward)."""
        b, l, c = x.shape
        h, w = self.H, self.W
        assert l == h * w, "input feature has wrong size"
        o = self.attn.attend_image(y.view(b, h, w, c), v, self.shift_size).reshape(b, h * w, self.v_dim)
        return _block_tail(self, x, o)


class BasicCRFLayer(nn.Module):
    """One NeWCRF stage: blocks alternating shift 0 / ws//2 (newcrf_layers.py:260-363)."""

    def __init__(self, dim, depth, num_heads, v_dim, window_size=7, mlp_ratio=4., qkv_bias=True, qk_scale=None,
                 drop=0., attn_drop=0., drop_path=0., norm_layer=nn.LayerNorm, downsample=None,
                 use_checkpoint=False):
        super().__init__()
        self.window_size, self.shift_size = window_size, window_size // 2
        self.depth, self.use_checkpoint = depth, use_checkpoint
        self.blocks = nn.ModuleList([
            CRFBlock(dim=dim, num_heads=num_heads, v_dim=v_dim, window_size=window_size,
                     shift_size=0 if i % 2 == 0 else window_size // 2, mlp_ratio=mlp_ratio, qkv_bias=qkv_bias,
                     qk_scale=qk_scale, drop=drop, attn_drop=attn_drop,
                     drop_path=drop_path[i] if isinstance(drop_path, list) else drop_path, norm_layer=norm_layer)
            for i in range(depth)])
        self.downsample = downsample(dim=dim, norm_layer=norm_layer) if downsample is not None else None

    def forward(self, x, v, H, W):
        if v.is_cuda and v.dtype == torch.float32 and torch.is_autocast_enabled() and \
                torch.get_autocast_dtype("cuda") == torch.bfloat16:
            v = v.to(torch.bfloat16)   # the value every block's kernel reads; cast once per level instead of per block
        pending = scale = None
        for blk in self.blocks:
            blk.H, blk.W = H, W
            if self.use_checkpoint and torch.is_grad_enabled():
                # The reference's branch (newcrf_layers.py:353-355) calls ``checkpoint(blk, x, attn_mask)`` without ``v``
                # and raises a TypeError; NewCRF hard-codes use_checkpoint=False (:404), so it never runs there.  Here
                # the flag does what it says: the block is recomputed in the backward (same outputs and gradients).
                if pending is not None:
                    x, pending = ops.residual(x, pending, scale), None
                x = torch.utils.checkpoint.checkpoint(blk, x, v, None, use_reentrant=False)
                continue
            if pending is None:
                y = ops.norm_for_linear(blk.norm1, x)
            else:
                x, y = ops.add_norm(blk.norm1, x, pending, scale)   # previous block's last residual add + this norm1
            x, pending, scale = blk.forward_parts(x, v, y)
        if pending is not None:
            x = ops.residual(x, pending, scale)
        if self.downsample is not None:
            return x, H, W, self.downsample(x, H, W), (H + 1) // 2, (W + 1) // 2
        return x, H, W, x, H, W


class NewCRF(nn.Module):
    """3x3 conv projections -> BasicCRFLayer -> LayerNorm -> NCHW (newcrf_layers.py:366-432)."""

    def __init__(self, input_dim=96, embed_dim=96, v_dim=64, window_size=7, num_heads=4, depth=2, patch_size=4,
                 in_chans=3, norm_layer=nn.LayerNorm, patch_norm=True):
        super().__init__()
        self.embed_dim, self.patch_norm = embed_dim, patch_norm
        self.proj_x = nn.Conv2d(input_dim, embed_dim, 3, padding=1) if input_dim != embed_dim else None
        self.proj_v = nn.Conv2d(v_dim, embed_dim, 3, padding=1) if v_dim != embed_dim else None
        self.crf_layer = BasicCRFLayer(dim=embed_dim, depth=depth, num_heads=num_heads, v_dim=embed_dim,
                                       window_size=window_size, mlp_ratio=4., qkv_bias=True, qk_scale=None,
                                       drop=0., attn_drop=0., drop_path=0., norm_layer=norm_layer,
                                       downsample=None, use_checkpoint=False)
        self.norm_crf = norm_layer(embed_dim)

    def forward(self, x, v):
        # channels-last strides through the level: the token views below are then free, and cuDNN's NHWC kernels
        # take the conv inputs as they are (no layout conversion kernels around every conv)
        if x.is_cuda:
            x = x.contiguous(memory_format=torch.channels_last)
            v = v.contiguous(memory_format=torch.channels_last)
        if self.proj_x is not None:
            x = self.proj_x(x)
        if self.proj_v is not None:
            v = self.proj_v(v)
        h, w = x.shape[2], x.shape[3]
        x = x.flatten(2).transpose(1, 2)
        v = v.permute(0, 2, 3, 1)
        x_out, h, w, _, _, _ = self.crf_layer(x, v, h, w)
        x_out = ops.norm_fp32(self.norm_crf, x_out)
        x_out = x_out.view(-1, h, w, self.embed_dim).permute(0, 3, 1, 2)
        return x_out if x_out.is_cuda else x_out.contiguous()


class SwinTransformerBlock(nn.Module):
    """Swin block (swin_transformer.py:147-246); see CRFBlock about ``mask_matrix``."""

    def __init__(self, dim, num_heads, window_size=7, shift_size=0, mlp_ratio=4., qkv_bias=True, qk_scale=None,
                 drop=0., attn_drop=0., drop_path=0., act_layer=nn.GELU, norm_layer=nn.LayerNorm):
        super().__init__()
        self.dim, self.num_heads = dim, num_heads
        self.window_size, self.shift_size, self.mlp_ratio = window_size, shift_size, mlp_ratio
        assert 0 <= self.shift_size < self.window_size, "shift_size must in 0-window_size"
        self.norm1 = norm_layer(dim)
        self.attn = EncoderWindowAttention(dim, window_size=_pair(window_size), num_heads=num_heads,
                                           qkv_bias=qkv_bias, qk_scale=qk_scale, attn_drop=attn_drop, proj_drop=drop)
        self.drop_path = DropPath(drop_path) if drop_path > 0. else nn.Identity()
        self.norm2 = norm_layer(dim)
        self.mlp = Mlp(in_features=dim, hidden_features=int(dim * mlp_ratio), act_layer=act_layer, drop=drop)
        self.H = None
        self.W = None

    def forward(self, x, mask_matrix=None):
        x, pending, scale = self.forward_parts(x, ops.norm_for_linear(self.norm1, x))
        return x if pending is None else ops.residual(x, pending, scale)

    def forward_parts(self, x, y):
        """The block without its last residual add (see CRFBlock.forward_parts)."""
        b, l, c = x.shape
        h, w = self.H, self.W
        assert l == h * w, "input feature has wrong size"
        o = self.attn.attend_image(y.view(b, h, w, c), self.shift_size).reshape(b, h * w, c)
        return _block_tail(self, x, o)


class PatchMerging(nn.Module):
    """2x2 neighbourhood concat -> LN -> Linear(4C, 2C) (swin_transformer.py:249-289)."""

    def __init__(self, dim, norm_layer=nn.LayerNorm):
        super().__init__()
        self.dim = dim
        self.reduction = nn.Linear(4 * dim, 2 * dim, bias=False)
        self.norm = norm_layer(4 * dim)

    def forward(self, x, H, W):
        b, l, c = x.shape
        assert l == H * W, "input feature has wrong size"
        y = ops.patch_merge_norm(self.norm, x, H, W)   # inference: gather + norm in one kernel
        if y is not None:
            return self.reduction(y)
        x = x.view(b, H, W, c)
        if H % 2 or W % 2:
            x = F.pad(x, (0, 0, 0, W % 2, 0, H % 2))
        x = torch.cat([x[:, 0::2, 0::2], x[:, 1::2, 0::2], x[:, 0::2, 1::2], x[:, 1::2, 1::2]], -1)
        return self.reduction(ops.norm_for_linear(self.norm, x.view(b, -1, 4 * c)))


class BasicLayer(nn.Module):
    """One Swin stage (swin_transformer.py:292-393)."""

    def __init__(self, dim, depth, num_heads, window_size=7, mlp_ratio=4., qkv_bias=True, qk_scale=None, drop=0.,
                 attn_drop=0., drop_path=0., norm_layer=nn.LayerNorm, downsample=None, use_checkpoint=False):
        super().__init__()
        self.window_size, self.shift_size = window_size, window_size // 2
        self.depth, self.use_checkpoint = depth, use_checkpoint
        self.blocks = nn.ModuleList([
            SwinTransformerBlock(dim=dim, num_heads=num_heads, window_size=window_size,
                                 shift_size=0 if i % 2 == 0 else window_size // 2, mlp_ratio=mlp_ratio,
                                 qkv_bias=qkv_bias, qk_scale=qk_scale, drop=drop, attn_drop=attn_drop,
                                 drop_path=drop_path[i] if isinstance(drop_path, list) else drop_path,
                                 norm_layer=norm_layer)
            for i in range(depth)])
        self.downsample = downsample(dim=dim, norm_layer=norm_layer) if downsample is not None else None

    def forward(self, x, H, W):
        pending = scale = None
        for blk in self.blocks:
            blk.H, blk.W = H, W
            if self.use_checkpoint and torch.is_grad_enabled():
                if pending is not None:
                    x, pending = ops.residual(x, pending, scale), None
                x = torch.utils.checkpoint.checkpoint(blk, x, None, use_reentrant=False)
                continue
            if pending is None:
                y = ops.norm_for_linear(blk.norm1, x)
            else:
                x, y = ops.add_norm(blk.norm1, x, pending, scale)   # previous block's last residual add + this norm1
            x, pending, scale = blk.forward_parts(x, y)
        if pending is not None:
            x = ops.residual(x, pending, scale)
        if self.downsample is not None:
            return x, H, W, self.downsample(x, H, W), (H + 1) // 2, (W + 1) // 2
        return x, H, W, x, H, W


class PatchEmbed(nn.Module):
    """4x4/4 conv patchify (+LN) (swin_transformer.py:396-436)."""

    def __init__(self, patch_size=4, in_chans=3, embed_dim=96, norm_layer=None):
        super().__init__()
        self.patch_size = _pair(patch_size)
        self.in_chans, self.embed_dim = in_chans, embed_dim
        self.proj = nn.Conv2d(in_chans, embed_dim, kernel_size=self.patch_size, stride=self.patch_size)
        self.norm = norm_layer(embed_dim) if norm_layer is not None else None

    def forward(self, x):
        _, _, h, w = x.shape
        ph, pw = self.patch_size
        if w % pw:
            x = F.pad(x, (0, pw - w % pw))
        if h % ph:
            x = F.pad(x, (0, 0, 0, ph - h % ph))
        if (x.is_cuda and not torch.is_grad_enabled() and torch.is_autocast_enabled()
                and torch.get_autocast_dtype("cuda") == torch.bfloat16 and (self.in_chans * ph * pw) % 8 == 0):
            # A stride == kernel conv is a per-patch Linear: gather the patches as bf16 rows (one strided copy) and run
            # one GEMM with the bias in its epilogue, instead of cuDNN's small-channel conv (10x off the memory
            # roofline here) + layout copy + separate bias add.  Same bf16 operands and fp32 accumulation as autocast's conv.
            b, c, hh, ww = x.shape[0], x.shape[1], x.shape[2] // ph, x.shape[3] // pw
            cols = torch.empty(b, hh, ww, c, ph, pw, dtype=torch.bfloat16, device=x.device)
            cols.copy_(x.contiguous().view(b, c, hh, ph, ww, pw).permute(0, 2, 4, 1, 3, 5))
            bias = None if self.proj.bias is None else self.proj.bias.to(torch.bfloat16)
            y = F.linear(cols.view(b, hh * ww, c * ph * pw), self.proj.weight.reshape(self.embed_dim, -1).to(torch.bfloat16), bias)
            if self.norm is not None:
                y = ops.norm_fp32(self.norm, y)
            return y.transpose(1, 2).view(b, self.embed_dim, hh, ww)
        if x.is_cuda:
            x = x.contiguous(memory_format=torch.channels_last)   # conv output in NHWC: the token view below is free
        x = self.proj(x)
        if self.norm is not None:
            hh, ww = x.shape[2], x.shape[3]
            x = ops.norm_fp32(self.norm, x.flatten(2).transpose(1, 2)).transpose(1, 2).view(-1, self.embed_dim, hh, ww)
        return x
