"""Plain-PyTorch (CPU, fp32/fp64) restatement of the reference's shifted-window
attention path, on image-order tensors.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Every function cites the reference lines it restates (paths relative to
/root/reference).  The reference computes the path as

    pad -> roll(-s) -> window_partition -> [Linear] -> scale, QK^T, +bias,
    +mask, softmax, PV -> [Linear] -> window_reverse -> roll(+s) -> crop

and because the two Linears are per-token they commute with pad/roll/partition;
this file restates everything between them as ONE function of image-order
q/k/v tensors, which is exactly the contract of the CUDA op
(include/swin_attn.h).
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
import torch.nn.functional as F


def relative_position_index(ws: int) -> torch.Tensor:
    """(N, N) int64 index into the (2ws-1)^2-row bias table.

    model/NewCRFs/newcrf_layers.py:89-100 == swin_transformer.py:91-102:
    index[i, j] = (yi - yj + ws - 1) * (2 ws - 1) + (xi - xj + ws - 1).
    """
    ys, xs = torch.meshgrid(torch.arange(ws), torch.arange(ws), indexing="ij")
    ys, xs = ys.reshape(-1), xs.reshape(-1)
    dy = ys[:, None] - ys[None, :] + (ws - 1)
    dx = xs[:, None] - xs[None, :] + (ws - 1)
    return dy * (2 * ws - 1) + dx


def partition(x: torch.Tensor, ws: int) -> torch.Tensor:
    """(B, Hp, Wp, C) -> (B * nW, ws*ws, C); newcrf_layers.py:30-42."""
    b, hp, wp, c = x.shape
    x = x.reshape(b, hp // ws, ws, wp // ws, ws, c)
    return x.permute(0, 1, 3, 2, 4, 5).reshape(-1, ws * ws, c)


def unpartition(win: torch.Tensor, ws: int, b: int, hp: int, wp: int) -> torch.Tensor:
    """(B * nW, ws*ws, C) -> (B, Hp, Wp, C); newcrf_layers.py:45-59."""
    c = win.shape[-1]
    x = win.reshape(b, hp // ws, wp // ws, ws, ws, c)
    return x.permute(0, 1, 3, 2, 4, 5).reshape(b, hp, wp, c)


def shift_mask(hp: int, wp: int, ws: int, shift: int, dtype=torch.float32, device="cpu") -> torch.Tensor:
    """(nW, N, N) additive SW-MSA mask with values {0, -100}.

    newcrf_layers.py:331-350 == swin_transformer.py:361-380: nine regions
    from the slices [0, -ws), [-ws, -shift), [-shift, end) along each axis of
    the *shifted, padded* image; pairs in different regions get -100.
    """
    region = torch.zeros(1, hp, wp, 1, dtype=dtype, device=device)
    cuts = (slice(0, -ws), slice(-ws, -shift), slice(-shift, None))
    n = 0
    for hs in cuts:
        for wsl in cuts:
            region[:, hs, wsl, :] = n
            n += 1
    ids = partition(region, ws).squeeze(-1)              # (nW, N)
    diff = ids[:, None, :] - ids[:, :, None]
    return torch.where(diff != 0, torch.full_like(diff, -100.0), torch.zeros_like(diff))


def _pad_to_window(x: torch.Tensor, ws: int, fill: Optional[torch.Tensor], mode: str = "constant") -> torch.Tensor:
    """Right/bottom pad to a multiple of ws (newcrf_layers.py:212-217).

    The reference pads the *pre-Linear* tensor with zeros, so a pad token's
    q/k (and the encoder's v) equal the Linear bias; ``fill`` carries that
    per-channel vector.  ``fill=None`` is a true zero pad (decoder v,
    newcrf_layers.py:216).
    """
    b, h, w, c = x.shape
    pb, pr = (-h) % ws, (-w) % ws
    if pb == 0 and pr == 0:
        return x
    if mode == "replicate":   # model/ODA2/oda2_swin_transformer.py:12,254-258: pad tokens copy the edge tokens
        return F.pad(x, (0, 0, 0, pr, 0, pb), mode="replicate")
    out = F.pad(x, (0, 0, 0, pr, 0, pb))
    if fill is not None:
        out[:, h:, :, :] = fill
        out[:, :, w:, :] = fill
    return out


def window_attention(
    q: torch.Tensor,
    k: torch.Tensor,
    v: torch.Tensor,
    bias_table: torch.Tensor,
    window: int,
    shift: int,
    num_heads: int,
    scale: Optional[float] = None,
    q_pad: Optional[torch.Tensor] = None,
    k_pad: Optional[torch.Tensor] = None,
    v_pad: Optional[torch.Tensor] = None,
    drop_scale: Optional[torch.Tensor] = None,
    pad_mode: str = "constant",
) -> Tuple[torch.Tensor, torch.Tensor]:
    """Shifted-window multi-head attention on image-order tensors.

    q, k, v: (B, H, W, C); bias_table: ((2ws-1)^2, nH).  Returns
    ``o`` (B, H, W, C) in un-shifted image coordinates, cropped to (H, W), and
    ``lse`` (B, Hp, Wp, nH): log-sum-exp of each query row's logits (natural
    log), also in un-shifted coordinates of the *padded* image.

    ``drop_scale``: optional multiplicative dropout mask on the probabilities.

    Restates WindowAttention.forward (newcrf_layers.py:117-146,
    swin_transformer.py:120-141) plus the pad/roll/partition/reverse/crop of
    CRFBlock.forward (newcrf_layers.py:210-249) and SwinTransformerBlock.forward
    (swin_transformer.py:204-238).
    """
    b, h, w, c = q.shape
    ws, nh = window, num_heads
    hd = c // nh
    if scale is None:
        scale = hd ** -0.5
    n = ws * ws
    qp = _pad_to_window(q, ws, q_pad, pad_mode)   # pad_mode="replicate": the ODA2 Swin (the pad vectors are unused)
    kp = _pad_to_window(k, ws, k_pad, pad_mode)
    vp = _pad_to_window(v, ws, v_pad, pad_mode)
    hp, wp = qp.shape[1], qp.shape[2]
    if shift > 0:
        qp, kp, vp = (torch.roll(t, shifts=(-shift, -shift), dims=(1, 2)) for t in (qp, kp, vp))

    def heads(t):                                         # (B_, N, C) -> (B_, nH, N, hd)
        return partition(t, ws).reshape(-1, n, nh, hd).transpose(1, 2)

    qw, kw, vw = heads(qp), heads(kp), heads(vp)
    logits = (qw * scale) @ kw.transpose(-2, -1)          # (B_, nH, N, N)
    idx = relative_position_index(ws).reshape(-1).to(q.device)
    bias = bias_table.to(logits.dtype)[idx].reshape(n, n, nh).permute(2, 0, 1)
    logits = logits + bias[None]
    if shift > 0:
        m = shift_mask(hp, wp, ws, shift, torch.float32, q.device)   # (nW, N, N), fp32 as in the reference
        nw = m.shape[0]
        logits = (logits.reshape(b, nw, nh, n, n) + m[None, :, None]).reshape(-1, nh, n, n)
    lse = torch.logsumexp(logits, dim=-1)                 # (B_, nH, N)
    p = torch.softmax(logits, dim=-1)
    if drop_scale is not None:                            # nn.Dropout on the softmax output (newcrf_layers.py:137):
        p = p * drop_scale.to(p.dtype)                    # (B_, nH, N, N) of {0, 1/(1-p)}, supplied by the test
    ow = (p @ vw).transpose(1, 2).reshape(-1, n, c)
    o = unpartition(ow, ws, b, hp, wp)
    lse_img = unpartition(lse.transpose(1, 2), ws, b, hp, wp)
    if shift > 0:
        o = torch.roll(o, shifts=(shift, shift), dims=(1, 2))
        lse_img = torch.roll(lse_img, shifts=(shift, shift), dims=(1, 2))
    return o[:, :h, :w, :].contiguous(), lse_img.contiguous()


def attention_core_bytes_flops(b: int, hp: int, wp: int, c: int, ws: int, elem: int, backward: bool = False):
    """Algorithmic bytes / flops of one launch."""
    tokens = b * hp * wp
    n = ws * ws
    if backward:
        return 8 * tokens * c * elem, 10 * tokens * n * c
    return 4 * tokens * c * elem, 4 * tokens * n * c
