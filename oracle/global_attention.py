"""CPU restatement of the Depthformer decoder's global self-attention (TEST INFRASTRUCTURE ONLY, see oracle/__init__.py).

``global_attention`` follows ``model/Depthformer/self_attention.py:36-46`` (head split / merge) and ``:70-84`` (scaled
dot product, softmax, value product): plain PyTorch, any dtype / device.  The product's kernel
(``mha_attention_forward``) is checked against it; the fixtures under tests/golden/ (``depthformer_*``) come from the
reference modules themselves (oracle/make_golden.py) and pin this file.
"""
from __future__ import annotations

from typing import Tuple

import torch

from mde_b200 import depthformer


def global_attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, num_heads: int, scale: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """q: (B, S, nH*dq), k: (B, Sk, nH*dq), v: (B, Sk, nH*dv) -> o (B, S, nH*dv), attn (B, nH, S, Sk).  Self-attention
    (self_attention.py:70-80) has Sk = S; the two halves of a Luna block (luna_layer.py:85-94, :118-127) do not."""
    b, s, _ = q.shape

    def heads(t):                                            # _reshape_qkv, self_attention.py:36-40; _reshape_3d, luna_layer.py:50-54
        return t.reshape(b, t.shape[1], num_heads, -1).transpose(1, 2).contiguous()

    attn = torch.matmul(heads(q), heads(k).transpose(-2, -1))   # :72
    attn = attn * scale                                      # :73 (in place in the reference)
    attn = torch.softmax(attn, dim=-1)                       # :74
    out = torch.matmul(attn, heads(v))                       # :80 (attn_drop is the identity in eval / at p = 0)
    return out.transpose(1, 2).reshape(b, s, -1), attn       # _reshape_output, :42-46


class _OracleCore:
    def _attend(self, q, k, v):
        out, attn = global_attention(q, k, v, self.num_heads, self.attn_scale)
        if self.training and self.attn_drop.p > 0.0:
            b, sk, _ = v.shape
            vh = v.view(b, sk, self.num_heads, -1).transpose(1, 2)
            out = torch.matmul(self.attn_drop(attn), vh).transpose(1, 2).reshape(b, q.shape[1], -1)
        return out, attn


class OracleSelfAttentionBlock(_OracleCore, depthformer.SelfAttentionBlock):
    pass


class OraclePostNormSelfAttentionBlock(_OracleCore, depthformer.PostNormSelfAttentionBlock):
    pass


class OracleLunaBlock(_OracleCore, depthformer.LunaBlock):
    pass


class OraclePreNormLunaBlock(_OracleCore, depthformer.PreNormLunaBlock):
    pass


SWAP = {depthformer.SelfAttentionBlock: OracleSelfAttentionBlock,
        depthformer.PostNormSelfAttentionBlock: OraclePostNormSelfAttentionBlock,
        depthformer.LunaBlock: OracleLunaBlock, depthformer.PreNormLunaBlock: OraclePreNormLunaBlock}
