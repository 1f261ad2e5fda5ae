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


def synthetic_depth_gt(name: str, b: int, h: int, w: int, max_depth: float = 80.0, sparsity: float = 0.7):
    """KITTI-like GT: uniform(1e-3, max_depth) with ``sparsity`` of pixels
    invalid (=0) (dataset/depth_dataset.py:68-69)."""
    g = _gen(name)
    gt = 1e-3 + (max_depth - 1e-3) * torch.rand(b, 1, h, w, generator=g)
    keep = torch.rand(b, 1, h, w, generator=g) > sparsity
    return gt * keep


def depth_errors(gt: np.ndarray, pred: np.ndarray) -> dict:
    """abs_rel and rmse exactly as utils/depth_utils.py:32-54 computes them."""
    abs_rel = float(np.mean(np.abs(gt - pred) / gt))
    rmse = float(np.sqrt(((gt - pred) ** 2).mean()))
    return dict(abs_rel=abs_rel, rmse=rmse)
