"""Depthformer decoder attention: drop-in ``SelfAttentionBlock`` / ``PostNormSelfAttentionBlock`` /
``FeedForwardBlock`` / ``PostNormFeedForwardBlock`` / ``ViTLayer`` / ``LunaBlock`` / ``PreNormLunaBlock`` / ``LunaLayer`` /
``PreNormLunaLayer`` for ``model/Depthformer/{self_attention,feed_forward,vit_layer,luna_layer}.py`` -- same constructors,
attribute names, ``state_dict`` keys and return values (the blocks return their attention maps, so the maps are
materialised).

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
    def forward(ctx, q, k, v, num_heads, scale, token_index=None, index_bias=None):
        if token_index is None:
            o, attn = torch.ops.swin_attn.mha_fwd(q, k, v, num_heads, scale)
        else:
            o, attn = torch.ops.swin_attn.mha_fwd_ordered(q, k, v, token_index, index_bias, num_heads, scale)
        ctx.save_for_backward(q, k, v, attn)
        ctx.num_heads, ctx.scale = num_heads, scale
        ctx.token_index, ctx.bias_shape = token_index, None if index_bias is None else (index_bias.shape, index_bias.dtype)
        return o, attn

    @staticmethod
    def backward(ctx, d_o, d_attn):
        q, k, v, attn = ctx.saved_tensors
        nh, scale = ctx.num_heads, ctx.scale
        b = q.shape[0]

        def heads(t):
            return t.reshape(b, t.shape[1], nh, -1).transpose(1, 2).float()

        qh, kh, vh, doh = heads(q), heads(k), heads(v), heads(d_o)
        d_p = doh @ vh.transpose(-2, -1)
        if d_attn is not None:
            d_p = d_p + d_attn
        d_s = attn * (d_p - (attn * d_p).sum(-1, keepdim=True))
        d_v = attn.transpose(-2, -1) @ doh
        d_q = (d_s @ kh) * scale
        d_k = (d_s.transpose(-2, -1) @ qh) * scale

        def back(t, like):
            return t.transpose(1, 2).reshape(b, like.shape[1], -1).to(like.dtype)

        d_bias = None
        if ctx.token_index is not None and ctx.needs_input_grad[6]:
            # the bias enters the logits as index_bias[idx_i - idx_j + n - 1, head]: scatter dS back by that index
            shape, dtype = ctx.bias_shape
            idx = ctx.token_index.long()
            rel = (idx[:, :, None] - idx[:, None, :] + (shape[0] + 1) // 2 - 1).reshape(-1)
            d_bias = torch.zeros(shape, dtype=torch.float32, device=q.device).index_add_(0, rel, d_s.permute(0, 2, 3, 1).reshape(-1, nh))
            d_bias = d_bias.to(dtype)
        return back(d_q, q), back(d_k, k), back(d_v, v), None, None, None, d_bias


def global_attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, num_heads: int, scale: float,
                     token_index: Optional[torch.Tensor] = None, index_bias: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """``attn = softmax(q k^T * scale)`` (B, nH, S, Sk) fp32 and ``o = attn v`` (B, S, Dv) for projections q (B, S, D),
    k (B, Sk, D), v (B, Sk, Dv) whose heads are consecutive channel groups (self_attention.py:36-40, :70-80;
    luna_layer.py:50-54, :85-94, :118-127).  ``token_index`` (B, S) integers with ``index_bias`` (2 n - 1, nH) add the ordered
    bias ``index_bias[idx_i - idx_j + n - 1, head]`` to the scaled logits (oda2_red_order_swin2_decoder.py:88-93, :112-117)."""
    if not q.is_cuda:
        raise RuntimeError("global_attention needs CUDA tensors: this package has no CPU path (the oracle under oracle/ is the CPU restatement)")
    if torch.is_autocast_enabled():   # the Linears produced bf16 under autocast; the reference's matmuls run in it too
        dt = torch.get_autocast_dtype("cuda")
        q, k, v = q.to(dt), k.to(dt), v.to(dt)
    return _GlobalAttention.apply(q, k, v, num_heads, float(scale), token_index, index_bias)


def _attend(mod: nn.Module, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """The attention core of a block ``mod`` (``num_heads``, ``attn_scale``, ``attn_drop``): the kernel, or -- training with
    attention dropout -- the kernel's map, ``nn.Dropout`` and the value product in torch, as the reference orders them: it
    drops the probabilities that enter the value product and returns the un-dropped map (self_attention.py:74-80)."""
    if mod.training and mod.attn_drop.p > 0.0:
        _, attn = global_attention(q, k, v, mod.num_heads, mod.attn_scale)
        b, sk, _ = v.shape
        vh = v.view(b, sk, mod.num_heads, -1).transpose(1, 2)
        out = torch.matmul(mod.attn_drop(attn).to(vh.dtype), vh).transpose(1, 2).reshape(b, q.shape[1], -1)
        return out, attn
    return global_attention(q, k, v, mod.num_heads, mod.attn_scale)


class _AttentionCore(nn.Module):
    """Blocks reach the attention core through this one method (the oracle's CPU copies override it)."""

    def _attend(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        return _attend(self, q, k, v)


class _SelfAttentionBase(_AttentionCore):
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
        return self._attend(self.query_proj(hidden), self.key_proj(hidden), self.value_proj(hidden))


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


class _LunaBase(_AttentionCore):
    """Luna: linear unified nested attention (luna_layer.py:10-131, :134-259): PACK -- K auxiliary tokens attend over the HW
    image tokens, map (B, nH, K, HW) -- then UNPACK -- the image tokens attend over the K packed tokens, map (B, nH, HW, K).
    Both are ``global_attention`` calls with different query and key counts."""

    def __init__(self, hidden_dim: int, aux_dim: int, qk_proj_dim: int, num_heads: int, attn_drop_prob: float, drop_prob: float):
        super().__init__()
        self.hidden_dim, self.aux_dim, self.qk_proj_dim, self.num_heads = hidden_dim, aux_dim, qk_proj_dim, num_heads
        if hidden_dim % num_heads != 0:
            raise ValueError("Hidden dim not multiple of num heads.")
        self.head_dim = hidden_dim // num_heads
        if self.head_dim > 64 or qk_proj_dim % num_heads != 0 or qk_proj_dim // num_heads > 64:
            raise ValueError("the CUDA kernel serves head dims up to 64 (qk_proj_dim must split over the heads)")

    def _projections(self):
        h, a, d = self.hidden_dim, self.aux_dim, self.qk_proj_dim
        self.q1_proj, self.k1_proj, self.v1_proj, self.o1_proj = nn.Linear(a, d), nn.Linear(h, d), nn.Linear(h, h), nn.Linear(h, a)
        self.q2_proj, self.k2_proj, self.v2_proj, self.o2_proj = nn.Linear(h, d), nn.Linear(a, d), nn.Linear(a, h), nn.Linear(h, h)

    def _tail(self, attn_drop_prob, drop_prob):
        self.attn_scale = math.sqrt(1.0 / self.head_dim)
        self.attn_drop = nn.Dropout(attn_drop_prob, inplace=False)
        self.drop = nn.Dropout(drop_prob, inplace=False)


class LunaBlock(_LunaBase):
    """Post-norm Luna sub-block (``luna_layer.py:10-131``): returns ``(out, aux_out, attn1, attn2)``."""

    def __init__(self, hidden_dim: int, aux_dim: int, qk_proj_dim: int, num_heads: int, attn_drop_prob: float = 0.0, drop_prob: float = 0.1):
        super().__init__(hidden_dim, aux_dim, qk_proj_dim, num_heads, attn_drop_prob, drop_prob)
        self._projections()
        self._tail(attn_drop_prob, drop_prob)
        self.aux_norm = nn.LayerNorm(aux_dim, eps=1e-5)
        self.norm = nn.LayerNorm(hidden_dim, eps=1e-5)

    def forward(self, hidden: torch.Tensor, aux: torch.Tensor):
        assert hidden.shape[2] == self.hidden_dim
        out1, attn1 = self._attend(self.q1_proj(aux), self.k1_proj(hidden), self.v1_proj(hidden))
        out1 = self.drop(self.o1_proj(out1))
        aux_out = self.aux_norm(aux + out1)
        out2, attn2 = self._attend(self.q2_proj(hidden), self.k2_proj(out1), self.v2_proj(out1))
        out2 = self.drop(self.o2_proj(out2))
        return self.norm(hidden + out2), aux_out, attn1, attn2


class PreNormLunaBlock(_LunaBase):
    """Pre-norm Luna sub-block (``luna_layer.py:134-259``)."""

    def __init__(self, hidden_dim: int, aux_dim: int, qk_proj_dim: int, num_heads: int, attn_drop_prob: float = 0.0, drop_prob: float = 0.1):
        super().__init__(hidden_dim, aux_dim, qk_proj_dim, num_heads, attn_drop_prob, drop_prob)
        self.aux_norm = nn.LayerNorm(aux_dim, eps=1e-5)
        self.inter_norm = nn.LayerNorm(aux_dim, eps=1e-5)
        self.norm = nn.LayerNorm(hidden_dim, eps=1e-5)
        self._projections()
        self._tail(attn_drop_prob, drop_prob)

    def forward(self, hidden: torch.Tensor, aux: torch.Tensor):
        assert hidden.shape[2] == self.hidden_dim
        aux_n, hidden_n = self.aux_norm(aux), self.norm(hidden)
        out1, attn1 = self._attend(self.q1_proj(aux_n), self.k1_proj(hidden_n), self.v1_proj(hidden_n))
        out1 = self.drop(self.o1_proj(out1))
        aux_out = aux + out1
        out_n = self.inter_norm(out1)
        out2, attn2 = self._attend(self.q2_proj(hidden_n), self.k2_proj(out_n), self.v2_proj(out_n))
        out2 = self.drop(self.o2_proj(out2))
        return hidden + out2, aux_out, attn1, attn2


class _LunaLayerBase(nn.Module):
    def forward(self, hidden: torch.Tensor, aux: torch.Tensor):
        """hidden (B, D, H, W), aux (B, K, aux_dim) -> (hidden (B, D, H, W), aux, attn1 (B, nH, K, HW), attn2 (B, nH, HW, K))."""
        b, d, h, w = hidden.shape
        hidden = hidden.view(b, d, h * w).transpose(1, 2).contiguous()
        hidden, aux, attn1, attn2 = self.luna_attn(hidden, aux)
        hidden = self.feed_forward(hidden)
        return hidden.transpose(1, 2).reshape(b, d, h, w), aux, attn1, attn2


class LunaLayer(_LunaLayerBase):
    """``luna_layer.py:262-302``: post-norm Luna block + post-norm feed-forward."""

    def __init__(self, hidden_dim: int, aux_dim: int, qk_proj_dim: int, num_heads: int, *, feedforward_dim: Optional[int] = None,
                 attn_drop_prob: float = 0.0, drop_prob: float = 0.1, act_layer=nn.GELU):
        super().__init__()
        self.luna_attn = LunaBlock(hidden_dim, aux_dim, qk_proj_dim, num_heads, attn_drop_prob, drop_prob)
        self.feed_forward = PostNormFeedForwardBlock(hidden_dim, feedforward_dim, drop_prob, act_layer, add_weight=1.0)


class PreNormLunaLayer(_LunaLayerBase):
    """``luna_layer.py:305-345``: pre-norm Luna block + pre-norm feed-forward."""

    def __init__(self, hidden_dim: int, aux_dim: int, qk_proj_dim: int, num_heads: int, *, feedforward_dim: Optional[int] = None,
                 attn_drop_prob: float = 0.0, drop_prob: float = 0.1, act_layer=nn.GELU):
        super().__init__()
        self.luna_attn = PreNormLunaBlock(hidden_dim, aux_dim, qk_proj_dim, num_heads, attn_drop_prob, drop_prob)
        self.feed_forward = FeedForwardBlock(hidden_dim, feedforward_dim, drop_prob, act_layer, add_weight=1.0)
