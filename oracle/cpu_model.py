"""CPU execution of the NewCRFs model with the oracle attention.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  The product modules have
no CPU path; for the parity tests and for the timed CPU baseline this file
derives, from a product model, a copy whose attention modules run the
plain-PyTorch restatement in ``oracle/window_attention.py`` -- i.e. the
reference's op sequence (pad, roll, partition, bmm, gather, mask, softmax,
bmm, reverse, roll, crop).  The dependency points one way only: the oracle
imports the product's module definitions, never the reverse.
"""
from __future__ import annotations

import copy

import torch

import mde_b200
from mde_b200 import layers, siblings

from .window_attention import partition, unpartition, window_attention


def _table_bias(attn):
    idx = attn.relative_position_index.reshape(-1)
    n = attn.window_size[0] * attn.window_size[1]
    return attn.relative_position_bias_table[idx].reshape(n, n, -1).permute(2, 0, 1)


def _windows_forward(attn, q, k, v, mask):
    """WindowAttention.forward on window-order heads (newcrf_layers.py:120-146)."""
    b_, n, c = q.shape
    nh = attn.num_heads

    def heads(t):
        return t.reshape(b_, n, nh, c // nh).transpose(1, 2)

    logits = (heads(q) * attn.scale) @ heads(k).transpose(-2, -1) + _table_bias(attn)[None]
    if mask is not None:
        nw = mask.shape[0]
        logits = (logits.view(b_ // nw, nw, nh, n, n) + mask[None, :, None]).view(-1, nh, n, n)
    p = torch.softmax(logits, dim=-1)
    return (p @ heads(v)).transpose(1, 2).reshape(b_, n, c)


class OracleWindowAttention(layers.WindowAttention):
    def forward(self, x, v, mask=None):
        c = self.dim
        qk = self.qk(x)
        o = _windows_forward(self, qk[..., :c], qk[..., c:], v.to(qk.dtype), mask)
        return self.proj_drop(self.proj(o))

    def attend_image(self, x_img, v_img, shift):   # the attention core before ``proj`` (the block applies proj)
        c = self.dim
        qk = self.qk(x_img)
        bias = self.qk.bias
        o, _ = window_attention(qk[..., :c], qk[..., c:], v_img.to(qk.dtype), self.relative_position_bias_table,
                                self.window_size[0], shift, self.num_heads, self.scale,
                                q_pad=None if bias is None else bias[:c].to(qk.dtype),
                                k_pad=None if bias is None else bias[c:].to(qk.dtype), v_pad=None)
        return o


class OracleEncoderWindowAttention(layers.EncoderWindowAttention):
    def forward(self, x, mask=None):
        c = self.dim
        qkv = self.qkv(x)
        o = _windows_forward(self, qkv[..., :c], qkv[..., c:2 * c], qkv[..., 2 * c:], mask)
        return self.proj_drop(self.proj(o))

    def attend_image(self, x_img, shift):
        c = self.dim
        qkv = self.qkv(x_img)
        bias = self.qkv.bias
        pads = [None] * 3 if bias is None else [bias[i * c:(i + 1) * c].to(qkv.dtype) for i in range(3)]
        o, _ = window_attention(qkv[..., :c], qkv[..., c:2 * c], qkv[..., 2 * c:], self.relative_position_bias_table,
                                self.window_size[0], shift, self.num_heads, self.scale,
                                q_pad=pads[0], k_pad=pads[1], v_pad=pads[2])
        return o


def _split3(qkv, c):
    return qkv[..., :c], qkv[..., c:2 * c], qkv[..., 2 * c:]


class OracleOda2WindowAttention(siblings.Oda2WindowAttention):
    """model/ODA2/oda2_swin_transformer.py:150-183 (windows) and :254-288 (replicate pad, roll, partition ... crop)."""

    scale = property(lambda self: self.attn_scale)

    def attend_windows(self, qkv_windows, mask):
        return _windows_forward(self, *_split3(qkv_windows, self.dim), mask)

    def attend_image(self, y_img, shift):
        o, _ = window_attention(*_split3(self.qkv(y_img), self.dim), self.relative_position_bias_table, self.window_size[0], shift,
                                self.num_heads, self.attn_scale, pad_mode="replicate")
        return o


class OracleOdaWindowAttention(OracleEncoderWindowAttention, siblings.OdaWindowAttention):
    """timm WindowAttention inside model/ODA/oda_swin_transformer.py:123-160 (window-aligned grids only)."""

    def attend_windows(self, qkv_windows, mask):
        return _windows_forward(self, *_split3(qkv_windows, self.dim), mask)


_SWAP = {layers.WindowAttention: OracleWindowAttention, layers.EncoderWindowAttention: OracleEncoderWindowAttention,
         siblings.Oda2WindowAttention: OracleOda2WindowAttention, siblings.OdaWindowAttention: OracleOdaWindowAttention}


def to_oracle(module: torch.nn.Module, device="cpu") -> torch.nn.Module:
    """Deep copy of ``module`` (on the CPU by default) whose attention runs the oracle.

    With ``device="cuda"`` the same restatement runs as PyTorch-eager ATen kernels on the GPU:
    that is "the reference on the B200" (used to measure the reference's own bf16-autocast
    error and the eager baseline), still never the product path."""
    from .global_attention import SWAP as mha_swap
    from .ordered_window_attention import SWAP as ordered_swap
    mha_swap = {**mha_swap, **ordered_swap}
    clone = copy.deepcopy(module).to(device)
    for m in clone.modules():
        if type(m) in _SWAP:
            m.__class__ = _SWAP[type(m)]
        elif type(m) in mha_swap:
            m.__class__ = mha_swap[type(m)]
    return clone


def build_newcrfs(version: str = "large07", seed: int = 0, max_depth: float = 80.0) -> torch.nn.Module:
    torch.manual_seed(seed)
    return to_oracle(mde_b200.NewCRFDepth(version=version, max_depth=max_depth)).eval()
