"""Depthformer decoder attention: drop-in ``SelfAttentionBlock`` / ``PostNormSelfAttentionBlock`` /
``FeedForwardBlock`` / ``PostNormFeedForwardBlock`` / ``ViTLayer`` for ``model/Depthformer/{self_attention,feed_forward,
vit_layer}.py`` -- same constructors, attribute names, ``state_dict`` keys and return values (the block returns the
attention map, so the map is materialised).

The attention core -- ``softmax(q k^T * scale)`` returned as a (B, nH, S, S) map, times ``v`` -- runs in ONE CUDA kernel of
``libswin_attn.so`` (``mha_attention_forward``, csrc/mha_global.cu) instead of the reference's matmul / mul_ / softmax /
matmul / transpose chain; the projections stay ``nn.Linear``.  The backward of the core is computed from the saved map
with library matmuls (no kernel of ours: the op is 0.2 GFLOP per image); training with ``attn_drop_prob > 0`` takes the
map from the kernel and applies ``nn.Dropout`` + the value product with torch ops, exactly as the reference orders them.
There is no CPU path: the kernel needs CUDA tensors.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
import torch.nn as nn

from . import ops  # noqa: F401  (loads libswin_attn_torch.so and registers swin_attn::mha_fwd)


class _GlobalAttention(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, num_heads, scale):
        o, attn = torch.ops.swin_attn.mha_fwd(q, k, v, num_heads, scale)
        ctx.save_for_backward(q, k, v, attn)
        ctx.num_heads, ctx.scale = num_heads, scale
        return o, attn

    @staticmethod
    def backward(ctx, d_o, d_attn):
        q, k, v, attn = ctx.saved_tensors
        nh, scale = ctx.num_heads, ctx.scale
        b, s, _ = q.shape

        def heads(t):
            return t.reshape(b, s, nh, -1).transpose(1, 2).float()

        qh, kh, vh, doh = heads(q), heads(k), heads(v), heads(d_o)
        d_p = doh @ vh.transpose(-2, -1)
        if d_attn is not None:
            d_p = d_p + d_attn
        d_s = attn * (d_p - (attn * d_p).sum(-1, keepdim=True))
        d_v = attn.transpose(-2, -1) @ doh
        d_q = (d_s @ kh) * scale
        d_k = (d_s.transpose(-2, -1) @ qh) * scale

        def back(t, like):
            return t.transpose(1, 2).reshape(b, s, -1).to(like.dtype)

        return back(d_q, q), back(d_k, k), back(d_v, v), None, None


def global_attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, num_heads: int, scale: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """``attn = softmax(q k^T * scale)`` (B, nH, S, S) fp32 and ``o = attn v`` (B, S, D) for (B, S, D) projections whose
    heads are consecutive channel groups (self_attention.py:36-40, :70-80)."""
    if not q.is_cuda:
        raise RuntimeError("global_attention needs CUDA tensors: this package has no CPU path (the oracle under oracle/ is the CPU restatement)")
    if torch.is_autocast_enabled():   # the Linears produced bf16 under autocast; the reference's matmuls run in it too
        dt = torch.get_autocast_dtype("cuda")
        q, k, v = q.to(dt), k.to(dt), v.to(dt)
    return _GlobalAttention.apply(q, k, v, num_heads, float(scale))


class _SelfAttentionBase(nn.Module):
    def __init__(self, hidden_dim: int, key_query_dim: int, num_heads: int, attn_drop_prob: float, drop_prob: float):
        super().__init__()
        self.hidden_dim = hidden_dim
        self.key_query_dim = key_query_dim
        self.num_heads = num_heads
        if (hidden_dim % num_heads != 0) or (key_query_dim % num_heads != 0):
            raise ValueError("Hidden dim not multiple of num heads.")   # self_attention.py:21-22
        self.head_dim = key_query_dim // num_heads
        if self.head_dim > 64 or hidden_dim // num_heads > 64:
            raise ValueError("the CUDA kernel serves head dims up to 64")

    def _core(self, hidden: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        q, k, v = self.query_proj(hidden), self.key_proj(hidden), self.value_proj(hidden)
        if self.training and self.attn_drop.p > 0.0:
            # the reference drops the probabilities that enter the value product and returns the un-dropped map (:76-80)
            _, attn = global_attention(q, k, v, self.num_heads, self.attn_scale)
            b, s, _ = v.shape
            vh = v.view(b, s, self.num_heads, -1).transpose(1, 2)
            out = torch.matmul(self.attn_drop(attn).to(vh.dtype), vh).transpose(1, 2).reshape(b, s, -1)
            return out, attn
        return global_attention(q, k, v, self.num_heads, self.attn_scale)


class SelfAttentionBlock(_SelfAttentionBase):
    """Pre-norm self-attention sub-block (``model/Depthformer/self_attention.py:8-93``): returns
    ``(hidden + drop(out_proj(attention)), attn)``."""

    def __init__(self, hidden_dim: int, key_query_dim: int, num_heads: int, attn_drop_prob: float = 0.0, drop_prob: float = 0.1):
        super().__init__(hidden_dim, key_query_dim, num_heads, attn_drop_prob, drop_prob)
        self.norm = nn.LayerNorm(hidden_dim, eps=1e-5)
        self.query_proj = nn.Linear(hidden_dim, key_query_dim)
        self.key_proj = nn.Linear(hidden_dim, key_query_dim)
        self.value_proj = nn.Linear(hidden_dim, hidden_dim)
        self.out_proj = nn.Linear(hidden_dim, hidden_dim)
        self.attn_scale = math.sqrt(1.0 / self.head_dim)
        self.attn_drop = nn.Dropout(attn_drop_prob, inplace=False)
        self.drop = nn.Dropout(drop_prob, inplace=True)

    def forward(self, hidden: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        residual = hidden
        output, attn = self._core(self.norm(hidden))
        output = self.drop(self.out_proj(output))
        return output + residual, attn


class PostNormSelfAttentionBlock(_SelfAttentionBase):
    """Post-norm variant (``self_attention.py:96-173``): ``norm(hidden + drop(out_proj(attention)))``."""

    def __init__(self, hidden_dim: int, key_query_dim: int, num_heads: int, attn_drop_prob: float = 0.0, drop_prob: float = 0.1):
        super().__init__(hidden_dim, key_query_dim, num_heads, attn_drop_prob, drop_prob)
        self.query_proj = nn.Linear(hidden_dim, key_query_dim)
        self.key_proj = nn.Linear(hidden_dim, key_query_dim)
        self.value_proj = nn.Linear(hidden_dim, hidden_dim)
        self.out_proj = nn.Linear(hidden_dim, hidden_dim)
        self.attn_scale = math.sqrt(1.0 / self.head_dim)
        self.attn_drop = nn.Dropout(attn_drop_prob, inplace=False)
        self.drop = nn.Dropout(drop_prob, inplace=True)
        self.norm = nn.LayerNorm(hidden_dim, eps=1e-5)

    def forward(self, hidden: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        identity = hidden
        output, attn = self._core(hidden)
        output = self.drop(self.out_proj(output))
        return self.norm(output + identity), attn


class FeedForwardBlock(nn.Module):
    """``model/Depthformer/feed_forward.py:6-46``: ``hidden + add_weight * drop(fc2(drop(act(fc1(norm(hidden))))))``."""

    def __init__(self, hidden_dim: int, feedforward_dim: Optional[int] = None, drop_prob: float = 0.1, act_layer=nn.GELU,
                 add_weight: float = 1.0):
        super().__init__()
        self.hidden_dim = hidden_dim
        self.feedforward_dim = hidden_dim * 4 if feedforward_dim is None else feedforward_dim
        self.norm = nn.LayerNorm(hidden_dim, eps=1e-5)
        self.fc1 = nn.Linear(hidden_dim, self.feedforward_dim)
        self.act = act_layer()
        self.fc2 = nn.Linear(self.feedforward_dim, hidden_dim)
        self.drop = nn.Dropout(drop_prob, inplace=True)
        self.add_weight = add_weight

    def forward(self, hidden: torch.Tensor) -> torch.Tensor:
        x = self.drop(self.act(self.fc1(self.norm(hidden))))
        return hidden + self.drop(self.fc2(x)) * self.add_weight


class PostNormFeedForwardBlock(nn.Module):
    """``feed_forward.py:49-89``."""

    def __init__(self, hidden_dim: int, feedforward_dim: Optional[int] = None, drop_prob: float = 0.1, act_layer=nn.GELU,
                 add_weight: float = 1.0):
        super().__init__()
        self.hidden_dim = hidden_dim
        self.feedforward_dim = hidden_dim * 4 if feedforward_dim is None else feedforward_dim
        self.fc1 = nn.Linear(hidden_dim, self.feedforward_dim)
        self.act = act_layer()
        self.fc2 = nn.Linear(self.feedforward_dim, hidden_dim)
        self.drop = nn.Dropout(drop_prob, inplace=True)
        self.norm = nn.LayerNorm(hidden_dim, eps=1e-5)
        self.add_weight = add_weight

    def forward(self, hidden: torch.Tensor) -> torch.Tensor:
        x = self.drop(self.act(self.fc1(hidden)))
        return self.norm(hidden + self.drop(self.fc2(x)) * self.add_weight)


class ViTLayer(nn.Module):
    """``model/Depthformer/vit_layer.py:9-45``: ``num_repeat`` x (self-attention, feed-forward) with SHARED weights; returns the
    last repeat's attention map."""

    def __init__(self, hidden_dim: int, key_query_dim: int, num_heads: int, *, num_repeat: int = 1,
                 feedforward_dim: Optional[int] = None, attn_drop_prob: float = 0.0, drop_prob: float = 0.1, act_layer=nn.GELU):
        super().__init__()
        if num_repeat < 1:
            raise ValueError("num_repeat is less than 1.")
        self.num_repeat = num_repeat
        self.self_attn = SelfAttentionBlock(hidden_dim, key_query_dim, num_heads, attn_drop_prob, drop_prob)
        self.feed_forward = FeedForwardBlock(hidden_dim, feedforward_dim, drop_prob, act_layer)

    def forward(self, hidden: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        attn_weight = None
        for _ in range(self.num_repeat):
            hidden, attn_weight = self.self_attn(hidden)
            hidden = self.feed_forward(hidden)
        return hidden, attn_weight
