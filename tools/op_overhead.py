#!/usr/bin/env python
"""Host-side cost per call of the custom ops (tiny inputs, launches only, no sync in the loop)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import mde_b200  # noqa: E402,F401
from mde_b200 import _lib  # noqa: E402

dev = torch.device("cuda")
x = torch.randn(64, 192, device=dev)
w, b = torch.ones(192, device=dev), torch.zeros(192, device=dev)
qkv = torch.randn(1, 14, 14, 3 * 64, device=dev).bfloat16()
table = torch.zeros(169, 2, device=dev)
ln = torch.nn.LayerNorm(192).to(dev)


def bench(name, fn, n=2000):
    for _ in range(50):
        fn()
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(n):
        fn()
    dt = (time.perf_counter() - t) / n * 1e6
    torch.cuda.synchronize()
    print(f"{name:40s} {dt:7.1f} us/call")


with torch.no_grad():
    bench("aten layer_norm (no grad)", lambda: ln(x))
    bench("swin_attn::layer_norm (no grad)", lambda: torch.ops.swin_attn.layer_norm(x, w, b, 1e-5, True))
    bench("swin_attn::fwd (no grad)", lambda: torch.ops.swin_attn.fwd(qkv, None, table, None, None, 7, 3, 2, 0.17, False, 0, None))
    bench("aten add", lambda: x + x)
xg = x.clone().requires_grad_()
wg, bg = w.clone().requires_grad_(), b.clone().requires_grad_()
bench("swin_attn::layer_norm fwd (grad)", lambda: torch.ops.swin_attn.layer_norm(xg, wg, bg, 1e-5, True))
bench("aten layer_norm fwd (grad)", lambda: torch.nn.functional.layer_norm(xg, (192,), wg, bg))
y = torch.ops.swin_attn.layer_norm(xg, wg, bg, 1e-5, True)
dy = torch.ones_like(y)
bench("swin_attn::layer_norm fwd+bwd", lambda: torch.autograd.grad(torch.ops.swin_attn.layer_norm(xg, wg, bg, 1e-5, True), [xg, wg, bg], dy))
bench("aten layer_norm fwd+bwd", lambda: torch.autograd.grad(torch.nn.functional.layer_norm(xg, (192,), wg, bg), [xg, wg, bg], torch.ones_like(x)))
qg, tg = qkv.clone().requires_grad_(), table.clone().requires_grad_()
do = torch.ones(1, 14, 14, 64, device=dev).bfloat16()
bench("swin_attn::fwd+bwd", lambda: torch.autograd.grad(torch.ops.swin_attn.fwd(qg, None, tg, None, None, 7, 3, 2, 0.17, True, 0, None)[0], [qg, tg], do))
