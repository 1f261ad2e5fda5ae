"""Sibling consumers of the window-attention kernels: drop-in blocks for the reference's OTHER Swin
flavours, over the same fused op (``ops.window_attention``) as the NewCRFs modules.

* ``Oda2WindowAttention`` / ``Oda2SwinTransformerBlock`` / ``Oda2SwinTransformerStage`` /
  ``Oda2PatchMerging`` -- ``model/ODA2/oda2_swin_transformer.py:95-183, 186-295, 341-452, 298-338``: identical attention
  math, but tokens are padded to a multiple of the window with ``F.pad(mode="replicate")`` (:12, :254-258) and ``q`` is
  scaled in place (:160).  A replicated token is a copy of the edge token AFTER ``norm1``, and the qkv Linear is
  per-token, so replicating the projected ``qkv`` rows is the same thing: the block projects the H x W tokens only, pads the
  packed projection with ``replicate`` (one small copy, only when H or W is not a multiple of the window) and runs the
  window-ALIGNED kernel -- the pad-vector trick of the NewCRFs modules does not apply, replicated tokens differ per
  row / column.  Autograd sends the pad tokens' gradient back to the edge tokens through ``F.pad``'s backward.
* ``OdaSwinTransformerBlock`` -- ``model/ODA/oda_swin_transformer.py:28-160``: the timm (0.5.4) Swin block at a FIXED
  ``input_resolution`` with a non-persistent ``attn_mask`` buffer and ``reset_resolution``; window 12 for Swin-L/384.  The
  buffer is kept (same construction, same name) so that checkpoints and callers that touch it keep working; when it still
  equals the standard SW-MSA mask of the resolution -- always, unless a caller overwrote it -- the kernel's analytic
  region test is used (tcgen05 forward for window 7 and 12), otherwise the explicit-mask path of the op.

Constructor signatures, attribute names and ``state_dict`` keys follow the reference; head_dim must be 32 (every Swin
the reference builds).
"""
from __future__ import annotations

import math
from typing import Optional, Tuple, Union

import torch
import torch.nn as nn
import torch.nn.functional as F
import torch.utils.checkpoint as checkpoint

from . import ops
from .layers import DropPath, EncoderWindowAttention, Mlp, _relative_position_index

_SWIN_PADDING_MODE = "replicate"   # oda2_swin_transformer.py:12


def sw_msa_mask(hp: int, wp: int, window: int, shift: int, device=None, dtype=torch.float32) -> torch.Tensor:
    """The SW-MSA mask every Swin flavour of the reference builds for a (hp, wp) grid (oda_swin_transformer.py:72-94,
    oda2_swin_transformer.py:409-432, swin_transformer.py:361-380): (nW, N, N), 0 where two tokens of a shifted window come
    from the same region, -100 otherwise."""
    img = torch.zeros(1, hp, wp, 1, dtype=dtype, device=device)
    cnt = 0
    for hs in (slice(0, -window), slice(-window, -shift), slice(-shift, None)):
        for ws_ in (slice(0, -window), slice(-window, -shift), slice(-shift, None)):
            img[:, hs, ws_, :] = cnt
            cnt += 1
    win = img.view(1, hp // window, window, wp // window, window, 1).permute(0, 1, 3, 2, 4, 5).reshape(-1, window * window)
    m = win.unsqueeze(1) - win.unsqueeze(2)
    return m.masked_fill(m != 0, -100.0).masked_fill(m == 0, 0.0)


class _MaskCheck:
    """Remembers which mask tensors were verified to be the standard SW-MSA mask (by identity and version)."""

    def __init__(self):
        self._seen = {}

    def is_standard(self, mask: Optional[torch.Tensor], hp: int, wp: int, window: int, shift: int) -> bool:
        if mask is None:
            return False
        key = (mask.data_ptr(), mask._version, tuple(mask.shape), hp, wp, window, shift)
        hit = self._seen.get(key)
        if hit is None:
            n_win = (hp // window) * (wp // window)
            hit = tuple(mask.shape) == (n_win, window * window, window * window) and bool(
                torch.equal(mask.float(), sw_msa_mask(hp, wp, window, shift, device=mask.device)))
            if len(self._seen) > 64:
                self._seen.clear()
            self._seen[key] = hit
        return hit


def _attend_windows(mod, qkv_windows: torch.Tensor, mask: Optional[torch.Tensor], scale: float) -> torch.Tensor:
    """(num_windows*B, N, 3C) projected windows -> (num_windows*B, N, C): each window is a one-window image."""
    ws = mod.window_size[0]
    b_, n, c3 = qkv_windows.shape
    o = ops.window_attention(qkv_windows.reshape(b_, ws, ws, c3), None, mod.relative_position_bias_table, ws, 0, mod.num_heads,
                             scale, attn_mask=mask, dropout_p=mod.attn_drop.p if mod.training else 0.0)
    return o.reshape(b_, n, c3 // 3)


class Oda2WindowAttention(nn.Module):
    """``model/ODA2/oda2_swin_transformer.py:95-183``."""

    def __init__(self, dim: int, window_size: Union[int, Tuple[int, int]], num_heads: int, qkv_bias: bool = True,
                 attn_drop_prob: float = 0.0, drop_prob: float = 0.0) -> None:
        super().__init__()
        self.dim = dim
        self.window_size = (window_size, window_size) if isinstance(window_size, int) else tuple(window_size)
        self.num_heads = num_heads
        if dim % num_heads != 0:
            raise ValueError(f"Dim {dim} is not divisible by num_heads {num_heads}.")   # :118
        head_dim = dim // num_heads
        if head_dim != 32 or self.window_size[0] != self.window_size[1]:
            raise ValueError(f"the sm_100a kernels need head_dim == 32 and square windows (dim={dim}, num_heads={num_heads})")
        self.attn_scale = math.sqrt(1 / head_dim)
        wh, ww = self.window_size
        self.relative_position_bias_table = nn.Parameter(torch.zeros((2 * wh - 1) * (2 * ww - 1), num_heads))
        self.register_buffer("relative_position_index", _relative_position_index(wh, ww))
        self.qkv = nn.Linear(dim, dim * 3, bias=qkv_bias)
        self.proj = nn.Linear(dim, dim, bias=True)
        self.attn_drop = nn.Dropout(attn_drop_prob)
        self.proj_drop = nn.Dropout(drop_prob)
        nn.init.trunc_normal_(self.relative_position_bias_table, std=0.02)
        self.softmax = nn.Softmax(dim=-1)   # module-tree parity; the kernel does the softmax

    def forward(self, x: torch.Tensor, mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        """x: (num_windows*B, N, C); mask: (num_windows, N, N) or None (:150-183)."""
        return self.proj_drop(self.proj(self.attend_windows(self.qkv(x), mask)))

    def attend_windows(self, qkv_windows: torch.Tensor, mask: Optional[torch.Tensor]) -> torch.Tensor:
        return _attend_windows(self, qkv_windows, mask, self.attn_scale)

    def attend_image(self, y_img: torch.Tensor, shift: int) -> torch.Tensor:
        """Normalised image-order tokens (B, H, W, C) -> attention output before ``proj``, replicate-padded windows with
        the standard SW-MSA mask (analytic in the kernel)."""
        ws = self.window_size[0]
        h, w = y_img.shape[1:3]
        qkv = ops.linear(self.qkv, y_img)
        pb, pr = (-h) % ws, (-w) % ws
        if pb or pr:
            qkv = F.pad(qkv, (0, 0, 0, pr, 0, pb), mode=_SWIN_PADDING_MODE)   # the reference pads x; qkv is per-token
        o = ops.window_attention(qkv, None, self.relative_position_bias_table, ws, shift, self.num_heads, self.attn_scale,
                                 dropout_p=self.attn_drop.p if self.training else 0.0)
        return o[:, :h, :w, :] if (pb or pr) else o


class OdaWindowAttention(EncoderWindowAttention):
    """timm 0.5.4 ``WindowAttention`` as ``model/ODA/oda_swin_transformer.py:25`` imports it: the fused-qkv Swin attention
    (same parameters as the reference's own ``model/NewCRFs/swin_transformer.py:64-144``)."""

    def attend_windows(self, qkv_windows: torch.Tensor, mask: Optional[torch.Tensor]) -> torch.Tensor:
        return _attend_windows(self, qkv_windows, mask, self.scale)


class SwinMLP(Mlp):
    """``oda2_swin_transformer.py:19-38`` (same parameters as ``Mlp``: fc1, act, fc2, drop)."""

    def __init__(self, in_features: int, hidden_features: int, drop_prob: float = 0.0, act_layer=nn.GELU) -> None:
        super().__init__(in_features, hidden_features, None, act_layer, drop_prob)


class SwinWindowing(nn.Module):
    """Parameter-free placeholder kept for module-tree parity (``oda2_swin_transformer.py:41-92``); the kernels address
    windows through tensor maps, nothing is partitioned."""

    def __init__(self, window_size: int):
        super().__init__()
        self.window_size = window_size
        self.H: Optional[int] = None
        self.W: Optional[int] = None

    def window_partition(self, x: torch.Tensor) -> torch.Tensor:
        r = self.window_size
        b, h, w, c = x.shape
        self.H, self.W = h, w
        return x.view(b, h // r, r, w // r, r, c).permute(0, 1, 3, 2, 4, 5).reshape(-1, r, r, c)

    def window_reverse(self, windows: torch.Tensor, orig_h: Optional[int] = None, orig_w: Optional[int] = None) -> torch.Tensor:
        orig_h = orig_h if orig_h is not None else self.H
        orig_w = orig_w if orig_w is not None else self.W
        r = self.window_size
        sh, sw = orig_h // r, orig_w // r
        b = windows.shape[0] // (sh * sw)
        return windows.view(b, sh, sw, r, r, -1).permute(0, 1, 3, 2, 4, 5).reshape(b, orig_h, orig_w, -1)

    def extra_repr(self):
        return f"window_size={self.window_size}"


def _finish(blk, x, o):
    from .layers import _block_tail
    x, pending, scale = _block_tail(blk, x, o)
    return x if pending is None else ops.residual(x, pending, scale)


class Oda2SwinTransformerBlock(nn.Module):
    """``model/ODA2/oda2_swin_transformer.py:186-295``; ``forward(x, mask_matrix)`` with ``blk.H, blk.W`` set by the stage."""

    def __init__(self, dim: int, num_heads: int, window_size: int = 7, shift_size: int = 0, mlp_ratio: float = 4.,
                 qkv_bias: bool = True, attn_drop_prob: float = 0.0, drop_prob: float = 0.0, path_drop_prob: float = 0.0,
                 act_layer=nn.GELU) -> None:
        super().__init__()
        self.dim, self.num_heads = dim, num_heads
        self.window_size, self.shift_size, self.mlp_ratio = window_size, shift_size, mlp_ratio
        if not (0 <= self.shift_size < self.window_size):
            raise ValueError(f"shift_size {shift_size} must in [0, window_size {window_size})")   # :219
        self.norm1 = nn.LayerNorm(dim)
        self.attn = Oda2WindowAttention(dim, window_size=self.window_size, num_heads=num_heads, qkv_bias=qkv_bias,
                                        attn_drop_prob=attn_drop_prob, drop_prob=drop_prob)
        self.drop_path = DropPath(path_drop_prob) if (path_drop_prob > 0) else nn.Identity()
        self.norm2 = nn.LayerNorm(dim)
        self.mlp = SwinMLP(in_features=dim, hidden_features=int(dim * mlp_ratio), act_layer=act_layer, drop_prob=drop_prob)
        self.windowing = SwinWindowing(window_size=window_size)
        self.H: Optional[int] = None
        self.W: Optional[int] = None
        self._mask_check = _MaskCheck()

    def forward(self, x: torch.Tensor, mask_matrix: Optional[torch.Tensor]) -> torch.Tensor:
        b, n, c = x.shape
        h, w = self.H, self.W
        if n != h * w:
            raise ValueError(f"Input shape {tuple(x.shape)} does not match with size ({h}, {w}).")   # :244
        ws, shift = self.window_size, self.shift_size
        y = ops.norm_for_linear(self.norm1, x).view(b, h, w, c)
        hp, wp = h + (-h) % ws, w + (-w) % ws
        if shift == 0 or self._mask_check.is_standard(mask_matrix, hp, wp, ws, shift):
            o = self.attn.attend_image(y, shift)
        else:
            # a caller-made mask: the reference's op order with the explicit-mask form of the kernel
            yp = F.pad(y, (0, 0, 0, wp - w, 0, hp - h), mode=_SWIN_PADDING_MODE) if (hp != h or wp != w) else y
            yp = torch.roll(yp, shifts=(-shift, -shift), dims=(1, 2))
            win = self.windowing.window_partition(yp).view(-1, ws * ws, c)
            ow = self.attn.attend_windows(self.attn.qkv(win), mask_matrix)
            o = torch.roll(self.windowing.window_reverse(ow.view(-1, ws, ws, c)), shifts=(shift, shift), dims=(1, 2))[:, :h, :w, :]
        return _finish(self, x, o.reshape(b, h * w, c))


class Oda2PatchMerging(nn.Module):
    """``oda2_swin_transformer.py:298-338``.  (The reference pads H with ``w % 2`` and W with ``h % 2`` -- its pad tuple is
    ``(0, 0, 0, h % 2, 0, w % 2)`` -- which only works when both are odd or both even; reproduced as written.)"""

    def __init__(self, dim: int):
        super().__init__()
        self.dim = dim
        self.reduction = nn.Linear(4 * dim, 2 * dim, bias=False)
        self.norm = nn.LayerNorm(4 * dim)

    def forward(self, x: torch.Tensor, h: int, w: int) -> torch.Tensor:
        b, n, c = x.shape
        if n != h * w:
            raise ValueError(f"Input {tuple(x.shape)} does not match with size ({h}, {w})")
        x = x.view(b, h, w, c)
        if (h % 2 == 1) or (w % 2 == 1):
            x = F.pad(x, (0, 0, 0, h % 2, 0, w % 2), mode=_SWIN_PADDING_MODE)
        x = torch.cat([x[:, 0::2, 0::2, :], x[:, 1::2, 0::2, :], x[:, 0::2, 1::2, :], x[:, 1::2, 1::2, :]], -1)
        x = x.view(b, -1, 4 * c)
        return ops.linear(self.reduction, ops.norm_for_linear(self.norm, x))


class Oda2SwinTransformerStage(nn.Module):
    """``oda2_swin_transformer.py:341-452``: returns ``(x, h, w, x_down, h_down, w_down)``."""

    def __init__(self, dim: int, depth: int, num_heads: int, window_size: int = 7, mlp_ratio: float = 4., qkv_bias: bool = True,
                 drop_prob: float = 0.0, attn_drop_prob: float = 0.0, path_drop_prob: Union[float, Tuple[float, ...]] = 0.0,
                 downsample=None, use_checkpoint=False):
        super().__init__()
        self.window_size = window_size
        self.shift_size = window_size // 2
        self.depth = depth
        self.use_checkpoint = use_checkpoint
        self.blocks = nn.ModuleList([
            Oda2SwinTransformerBlock(
                dim=dim, num_heads=num_heads, window_size=window_size, shift_size=0 if (i % 2 == 0) else window_size // 2,
                mlp_ratio=mlp_ratio, qkv_bias=qkv_bias, drop_prob=drop_prob, attn_drop_prob=attn_drop_prob,
                path_drop_prob=path_drop_prob[i] if isinstance(path_drop_prob, (tuple, list)) else path_drop_prob)
            for i in range(depth)])
        self.windowing = SwinWindowing(window_size=window_size)
        self.downsample = downsample(dim=dim) if downsample is not None else None

    def forward(self, x: torch.Tensor, h: int, w: int):
        # The reference builds the (nW, N, N) SW-MSA mask here on every forward (:409-432); the kernels derive it from the
        # window position, so no mask tensor is built: the blocks get None and use the analytic test.
        for blk in self.blocks:
            blk.H, blk.W = h, w
            if self.use_checkpoint and self.training:
                x = checkpoint.checkpoint(_Oda2BlockNoMask(blk), x, use_reentrant=False)
            else:
                x = blk._forward_standard(x)
        if self.downsample is not None:
            return x, h, w, self.downsample(x, h, w), (h + 1) // 2, (w + 1) // 2
        return x, h, w, x, h, w


class _Oda2BlockNoMask(nn.Module):
    def __init__(self, blk):
        super().__init__()
        self.blk = blk

    def forward(self, x):
        return self.blk._forward_standard(x)


def _oda2_forward_standard(self, x: torch.Tensor) -> torch.Tensor:
    b, n, c = x.shape
    h, w = self.H, self.W
    if n != h * w:
        raise ValueError(f"Input shape {tuple(x.shape)} does not match with size ({h}, {w}).")
    y = ops.norm_for_linear(self.norm1, x).view(b, h, w, c)
    return _finish(self, x, self.attn.attend_image(y, self.shift_size).reshape(b, h * w, c))


Oda2SwinTransformerBlock._forward_standard = _oda2_forward_standard


class OdaSwinTransformerBlock(nn.Module):
    """``model/ODA/oda_swin_transformer.py:28-160`` (timm 0.5.4 Swin block, fixed ``input_resolution``)."""

    def __init__(self, dim, input_resolution, num_heads, window_size=7, shift_size=0, mlp_ratio=4., qkv_bias=True, drop=0.,
                 attn_drop=0., drop_path=0., act_layer=nn.GELU, norm_layer=nn.LayerNorm):
        super().__init__()
        self.dim = dim
        self.input_resolution = tuple(input_resolution)
        self.num_heads = num_heads
        self.window_size = window_size
        self.shift_size = shift_size
        self.mlp_ratio = mlp_ratio
        if min(self.input_resolution) <= self.window_size:
            # if window size is larger than input resolution, we don't partition windows (:55-58)
            self.shift_size = 0
            self.window_size = min(self.input_resolution)
        assert 0 <= self.shift_size < self.window_size, "shift_size must in 0-window_size"
        self.norm1 = norm_layer(dim)
        self.attn = OdaWindowAttention(dim, window_size=(self.window_size, self.window_size), num_heads=num_heads,
                                       qkv_bias=qkv_bias, attn_drop=attn_drop, proj_drop=drop)
        self.drop_path = DropPath(drop_path) if drop_path > 0. else nn.Identity()
        self.norm2 = norm_layer(dim)
        self.mlp = Mlp(in_features=dim, hidden_features=int(dim * mlp_ratio), act_layer=act_layer, drop=drop)
        attn_mask = None
        if self.shift_size > 0:
            h, w = self.input_resolution
            attn_mask = sw_msa_mask(h, w, self.window_size, self.shift_size)
        self.register_buffer("attn_mask", attn_mask, persistent=False)
        self._mask_check = _MaskCheck()

    def reset_resolution(self, size):
        """As written in the reference (:96-121): ``input_resolution`` is assigned BEFORE it is compared with ``size``, so
        the early return always fires and the mask buffer is never rebuilt; reproduced, not repaired."""
        self.input_resolution = size
        if (self.attn_mask is None) or (self.input_resolution == size):
            return

    def forward(self, x):
        h, w = self.input_resolution
        b, l, c = x.shape
        assert l == h * w, f"input feature has wrong size, {self.input_resolution} vs {tuple(x.shape)}"
        ws, shift = self.window_size, self.shift_size
        if h % ws or w % ws:
            raise RuntimeError(f"shape '{(b, h // ws, ws, w // ws, ws, c)}' is invalid for input of size {x.numel()}")  # timm window_partition's view
        y = ops.norm_for_linear(self.norm1, x).view(b, h, w, c)
        if shift == 0 or self._mask_check.is_standard(self.attn_mask, h, w, ws, shift):
            o = self.attn.attend_image(y, shift)   # window-aligned: no padding of any kind
        else:
            yr = torch.roll(y, shifts=(-shift, -shift), dims=(1, 2))
            win = yr.view(b, h // ws, ws, w // ws, ws, c).permute(0, 1, 3, 2, 4, 5).reshape(-1, ws * ws, c)
            ow = self.attn.attend_windows(self.attn.qkv(win), self.attn_mask)
            o = ow.view(b, h // ws, w // ws, ws, ws, c).permute(0, 1, 3, 2, 4, 5).reshape(b, h, w, c)
            o = torch.roll(o, shifts=(shift, shift), dims=(1, 2))
        return _finish(self, x, o.reshape(b, h * w, c))


class PreNormOrderedSwinSA(nn.Module):
    """``model/ODA2/oda2_red_order_swin2_decoder.py:13-132``: pre-norm window self-attention of the ODA2 depth-order decoder --
    windows of 4 / 8 / 16, NO shift mask, and a data-dependent bias: every pixel carries an integer depth index and the logit
    of (query i, key j) gets ``depth_embedding[idx_i - idx_j + num_emb - 1, head]``.  Returns ``(x + attention, attn)`` with
    the attention map (num_windows * B, heads, r*r, r*r).

    The windows are the batch of the global-attention kernels (``mha_attention_forward_ordered``: one launch computes the
    scaled dot product, the index-difference bias gather, the softmax, the returned map and the value product); roll,
    window partition / reverse and the projections stay torch ops, as in the reference.  ``attn_drop_prob > 0`` in
    training (the reference drops the scaled LOGITS, before the bias and the softmax, :110) takes the reference's op sequence
    in torch."""

    def __init__(self, in_dims: int, num_heads: int, num_emb: int, window_size: int = 8, shift_size: int = 0,
                 attn_drop_prob: float = 0.0, drop_prob: float = 0.0, bias_type: str = "depth", bias_init: str = "linear"):
        super().__init__()
        self.in_dims, self.num_heads = in_dims, num_heads
        if in_dims % num_heads != 0:
            raise ValueError(f"Input dim {in_dims} is not divisible by num_heads {num_heads}.")
        self.head_dim = in_dims // num_heads
        if self.head_dim > 64 or self.head_dim % 4:
            raise ValueError("the CUDA kernels serve head dims up to 64 (multiples of 4)")
        self.norm = nn.LayerNorm(in_dims)
        self.q_proj, self.k_proj = nn.Linear(in_dims, in_dims), nn.Linear(in_dims, in_dims)
        self.v_proj, self.o_proj = nn.Linear(in_dims, in_dims), nn.Linear(in_dims, in_dims)
        self.attn_scale = math.sqrt(1 / self.head_dim)
        self.drop = nn.Dropout(drop_prob)
        self.attn_drop = nn.Dropout(attn_drop_prob)
        self.window_size, self.shift_size = window_size, shift_size
        self.windowing = SwinWindowing(window_size=window_size)
        assert (self.window_size == 16) or (self.window_size == 8) or (self.window_size == 4)
        self.num_emb, self.bias_type = num_emb, bias_type
        if bias_type == "depth":
            if bias_init == "linear":      # :50-59
                with torch.no_grad():
                    de = torch.linspace(1, 2 * num_emb - 1, 2 * num_emb - 1) - num_emb
                    de = de.unsqueeze(-1).expand(2 * num_emb - 1, num_heads).contiguous()
                    init = torch.ones(num_heads, dtype=torch.float32).uniform_(0.01, 0.04)
                    de[:num_emb] *= init
                    de[-num_emb:] *= (-init)
            elif bias_init == "random":
                de = torch.zeros(2 * num_emb - 1, num_heads).uniform_(-0.05, 0.05)
            else:
                raise ValueError(f"Unsupported bias init {bias_init}.")
            self.depth_embedding = nn.Parameter(de, requires_grad=True)
        elif bias_type == "none":
            pass
        elif bias_type == "pos":
            raise NotImplementedError
        else:
            raise ValueError(f"Unsupported bias type {bias_type}.")

    def forward(self, x: torch.Tensor, indices: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        from .depthformer import global_attention
        b, h, w, d = x.shape
        assert d == self.in_dims
        r, s = self.window_size, self.shift_size
        identity = x
        if s > 0:
            x = torch.roll(x, shifts=(-s, -s), dims=(1, 2))
            indices = torch.roll(indices, shifts=(-s, -s), dims=(1, 2))
        xw = self.windowing.window_partition(x).reshape(-1, r * r, d)          # (nW*B, rr, d)
        x_norm = self.norm(xw)
        q, k, v = self.q_proj(x_norm), self.k_proj(x_norm), self.v_proj(x_norm)
        depth = self.bias_type == "depth"
        idx = self.windowing.window_partition(indices.unsqueeze(-1)).reshape(-1, r * r) if depth else None
        if self.training and self.attn_drop.p > 0.0:
            out, attn = self._reference_core(q, k, v, idx)
        else:
            out, attn = global_attention(q, k, v, self.num_heads, self.attn_scale, idx, self.depth_embedding if depth else None)
        out = self.drop(self.o_proj(out)).reshape(-1, r, r, d)
        out = self.windowing.window_reverse(out, h, w)
        if s > 0:
            out = torch.roll(out, shifts=(s, s), dims=(1, 2))
        return out + identity, attn

    def _reference_core(self, q, k, v, idx):
        """:104-122 as written (training with dropout on the scaled logits)."""
        n, rr, d = q.shape

        def heads(t):
            return t.view(n, rr, self.num_heads, d // self.num_heads).transpose(1, 2).contiguous()

        attn = torch.matmul(heads(q), heads(k).transpose(-1, -2)) * self.attn_scale
        attn = self.attn_drop(attn)
        if idx is not None:
            rel = idx.view(-1, rr, 1) - idx.view(-1, 1, rr) + (self.num_emb - 1)
            attn = attn + F.embedding(rel, weight=self.depth_embedding).permute(0, 3, 1, 2)
        attn = torch.softmax(attn, dim=-1)
        return torch.matmul(attn.to(v.dtype), heads(v)).transpose(1, 2).reshape(n, rr, d), attn
