"""Deterministic synthetic weights and inputs keyed by NAME, not by creation
order, so the reference modules (here) and the product modules (on the GPU
box) get bit-identical parameters without shipping multi-hundred-MB fixtures.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
"""
from __future__ import annotations

import math
import zlib

import numpy as np
import torch


def _gen(name: str, salt: int = 0) -> torch.Generator:
    return torch.Generator().manual_seed(zlib.crc32(name.encode()) + 7919 * salt)


def synthetic_state(module: torch.nn.Module, salt: int = 0) -> dict:
    """State dict for ``module`` with O(1)-effect biases, bias tables and norms
    (so pad-token, bias-gather and mask errors cannot hide under small weights,
    SURVEY.md section 8d)."""
    out = {}
    for key, ref in module.state_dict().items():
        g = _gen(key, salt)
        if not ref.dtype.is_floating_point:
            out[key] = ref.clone()
            continue
        shape = tuple(ref.shape)
        if key.endswith("running_var"):
            t = 0.5 + torch.rand(shape, generator=g)
        elif key.endswith("running_mean"):
            t = 0.1 * torch.randn(shape, generator=g)
        elif key.endswith("relative_position_bias_table"):
            t = 0.5 * torch.randn(shape, generator=g)
        elif ref.dim() == 1 and key.endswith("weight"):
            t = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif ref.dim() == 1:
            scale = 0.3 if (".qk." in key or ".qkv." in key) else 0.1
            t = scale * torch.randn(shape, generator=g)
        else:
            fan_in = int(np.prod(shape[1:]))
            t = torch.randn(shape, generator=g) / math.sqrt(fan_in)
        out[key] = t.to(ref.dtype)
    return out


def randn(name: str, *shape, salt: int = 0) -> torch.Tensor:
    return torch.randn(*shape, generator=_gen(name, salt))


def synthetic_depth_gt(name: str, b: int, h: int, w: int, max_depth: float = 80.0, sparsity: float = 0.7,
                       min_depth: float = 1e-3):
    """KITTI-like GT: uniform(min_depth, max_depth) with ``sparsity`` of pixels invalid (=0)
    (dataset/depth_dataset.py:68-69).  ``min_depth`` = 1e-3 is the config's validity floor
    (``min_depth_eval``): abs_rel divides by gt, so a handful of near-zero pixels carry its mean and the metric is
    ill-conditioned there.  ``min_depth`` = 1.0 is the realistic range (no KITTI / NYU pixel is closer than that order)
    on which the absolute 1e-3 budget of north_star is held."""
    g = _gen(name)
    gt = min_depth + (max_depth - min_depth) * torch.rand(b, 1, h, w, generator=g)
    keep = torch.rand(b, 1, h, w, generator=g) > sparsity
    return gt * keep


def silog_loss(pred: torch.Tensor, gt: torch.Tensor, alpha: float = 10.0, beta: float = 0.15, min_depth: float = 1e-3):
    """Scale-invariant log loss per image, averaged (json/kitti/newcrfs/newcrfs_github_eval.json: alpha 10, beta 0.15,
    per_image): alpha * sqrt(mean(d^2) - beta * mean(d)^2), d = log(pred) - log(gt) over the valid pixels.  The
    reference's loss module is absent from the snapshot; this restatement is what the golden gradient
    fixture differentiates, independently of the product's ``harness.SILogLoss``."""
    losses = []
    for p, g in zip(pred, gt):
        valid = g > min_depth
        d = torch.log(p[valid].clamp_min(min_depth)) - torch.log(g[valid])
        losses.append(alpha * torch.sqrt(((d * d).mean() - beta * d.mean() ** 2).clamp_min(1e-12)))
    return torch.stack(losses).mean()


def depth_errors(gt: np.ndarray, pred: np.ndarray) -> dict:
    """abs_rel and rmse exactly as utils/depth_utils.py:32-54 computes them."""
    abs_rel = float(np.mean(np.abs(gt - pred) / gt))
    rmse = float(np.sqrt(((gt - pred) ** 2).mean()))
    return dict(abs_rel=abs_rel, rmse=rmse)
