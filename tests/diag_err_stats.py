#!/usr/bin/env python
"""(diagnostic script, not collected by pytest; lives under tests/ because it uses the oracle)
Error statistics of the bf16 forward kernels against the fp64 oracle: max-normalised error, RMS error and
mean signed error (bias), tcgen05 vs CUDA-core kernel, padded / shifted cases."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from mde_b200 import _lib  # noqa: E402
from oracle import window_attention as wa  # noqa: E402

dev = "cuda"
for (b, h, w, nh, shift, sigma) in [(2, 22, 76, 6, 0, 1.0), (2, 22, 76, 6, 3, 1.0), (2, 28, 77, 6, 3, 1.0), (2, 22, 76, 6, 3, 3.0)]:
    c = nh * 32
    g = torch.Generator().manual_seed(7)
    qkv = (sigma * torch.randn(b, h, w, 3 * c, generator=g)).bfloat16()
    table = 0.5 * torch.randn(169, nh, generator=g)
    pad = (0.3 * torch.randn(3 * c, generator=g)).bfloat16()
    q, k, v = qkv.double()[..., :c], qkv.double()[..., c:2 * c], qkv.double()[..., 2 * c:]
    pd = pad.double()
    want, _ = wa.window_attention(q, k, v, table.double(), 7, shift, nh, None, pd[:c], pd[c:2 * c], pd[2 * c:])
    for name, impl in (("tcgen05", _lib.IMPL_AUTO), ("simt", _lib.IMPL_SIMT)):
        o, _ = torch.ops.swin_attn.fwd(qkv.to(dev), None, table.to(dev), pad.to(dev), None, 7, shift, nh, 32 ** -0.5, False, impl, None)
        e = o.double().cpu() - want
        print(f"B{b} {h}x{w} nH{nh} shift{shift} sigma{sigma} {name:8s} max-norm {float(e.abs().max() / want.abs().max()):.3e} "
              f"rms {float(e.pow(2).mean().sqrt()):.3e} (signal rms {float(want.pow(2).mean().sqrt()):.3e}) "
              f"mean {float(e.mean()):+.3e}")
