"""CPU restatement of the ODA2 depth-order window attention core (TEST INFRASTRUCTURE ONLY, see oracle/__init__.py).

``ordered_attention`` follows ``model/ODA2/oda2_red_order_swin2_decoder.py:88-93`` (bias from the pairwise difference of
the tokens' depth indices through ``F.embedding``) and ``:104-117`` (head split, scaled dot product, bias add, softmax,
value product) on window-order tensors; plain PyTorch, any dtype / device.  The fixtures ``oda2_ordered_sa_*`` come from
the reference module itself (oracle/make_golden.py) and pin this file; the product's kernel
(``mha_attention_forward_ordered``) is checked against it.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn.functional as F

from mde_b200 import siblings


def ordered_attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, num_heads: int, scale: float,
                      token_index: Optional[torch.Tensor], index_bias: Optional[torch.Tensor]) -> Tuple[torch.Tensor, torch.Tensor]:
    """q, k, v: (nW*B, rr, D); token_index: (nW*B, rr) ints in [0, n); index_bias: (2n-1, nH) -> o (nW*B, rr, D), attn."""
    n_win, rr, d = q.shape

    def heads(t):                                                            # _reshape_4d, :72-75
        return t.reshape(n_win, rr, num_heads, d // num_heads).transpose(1, 2).contiguous()

    attn = torch.matmul(heads(q), heads(k).transpose(-1, -2))               # :107
    attn = attn * scale                                                      # :108
    if token_index is not None:
        n_emb = (index_bias.shape[0] + 1) // 2
        rel = token_index.view(-1, rr, 1) - token_index.view(-1, 1, rr) + (n_emb - 1)        # :89-90
        attn = attn + F.embedding(rel.long(), weight=index_bias.to(attn.dtype)).permute(0, 3, 1, 2)   # :91-92, :110
    attn = torch.softmax(attn, dim=-1)                                       # :111
    out = torch.matmul(attn, heads(v))                                       # :113
    return out.transpose(1, 2).reshape(n_win, rr, d), attn                   # :114


class OraclePreNormOrderedSwinSA(siblings.PreNormOrderedSwinSA):
    def forward(self, x, indices):
        import mde_b200.depthformer as dfm
        real = dfm.global_attention
        dfm.global_attention = lambda q, k, v, nh, sc, idx=None, bias=None: ordered_attention(q, k, v, nh, sc, idx, bias)
        try:
            return super().forward(x, indices)
        finally:
            dfm.global_attention = real


SWAP = {siblings.PreNormOrderedSwinSA: OraclePreNormOrderedSwinSA}
