"""The reference's OWN model, built from the staged / upstream sources (oracle/reference_shim.py).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Imports nothing from the product package, so the reference arm
of ``bench.py`` never loads ``libswin_attn*.so``.  ``NewCRFDepth`` here is ``model/NewCRFs/NewCRFDepth.py:11-148``
itself, executed unmodified (the only stand-ins are the four timm / mmcv symbols of the shim, none of which carries
arithmetic of the hot path).
"""
from __future__ import annotations

import contextlib
import io
import time

import torch

from . import reference_shim as rs
from .synthetic import synthetic_state


def available() -> bool:
    return rs.available()


def images(batch: int, height: int, width: int, seed: int = 0) -> torch.Tensor:
    """ImageNet-normalised-like N(0,1) images (dataset/depth_dataset.py:290) -- the same generator and seed rule as
    ``mde_b200.harness.synthetic_images`` so both arms of the bench see the same pixels."""
    g = torch.Generator().manual_seed(seed)
    return torch.randn(batch, 3, height, width, generator=g)


def build_newcrfs(version: str = "large07", max_depth: float = 80.0, seed: int = 0, synthetic: bool = False,
                  device="cpu") -> torch.nn.Module:
    """Reference ``NewCRFDepth(version)`` in eval mode: default init under ``torch.manual_seed(seed)`` (what the
    product's ``harness.build_model`` does) or, with ``synthetic=True``, the name-keyed weights of
    ``oracle/synthetic.py`` that the parity fixtures use."""
    mod = rs.newcrf_depth()
    torch.manual_seed(seed)
    with contextlib.redirect_stdout(io.StringIO()):      # the constructor prints "== Load encoder backbone from: None"
        model = mod.NewCRFDepth(version=version, max_depth=max_depth, pretrained=None)
    if synthetic:
        model.load_state_dict(synthetic_state(model))
    return model.eval().to(device)


def decoder_only_forward(model: torch.nn.Module, feats, ppm_out):
    """BASELINE configs[0]: the NeWCRF decoder block crf3 .. crf0 on given encoder + PSP features
    (model/NewCRFs/NewCRFDepth.py:131-137)."""
    shuffle = torch.nn.PixelShuffle(2)
    e3 = shuffle(model.crf3(feats[3], ppm_out))
    e2 = shuffle(model.crf2(feats[2], e3))
    e1 = shuffle(model.crf1(feats[1], e2))
    return model.crf0(feats[0], e1)


def time_forward(fn, steps: int, warmup: int):
    """Wall-clock seconds of each of ``steps`` calls after ``warmup`` untimed ones (CPU arm)."""
    with torch.no_grad():
        for _ in range(warmup):
            fn()
        out = []
        for _ in range(steps):
            t0 = time.perf_counter()
            fn()
            out.append(time.perf_counter() - t0)
    return out


def time_forward_cuda(fn, steps: int, warmup: int, flush: torch.Tensor = None):
    """Milliseconds of each of ``steps`` calls on the current CUDA stream (CUDA events, L2 flushed before each)."""
    with torch.no_grad():
        for _ in range(warmup):
            fn()
        evs = []
        for _ in range(steps):
            if flush is not None:
                flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            evs.append((a, b))
        torch.cuda.synchronize()
    return [a.elapsed_time(b) for a, b in evs]
