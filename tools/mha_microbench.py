#!/usr/bin/env python
"""Global attention that returns its map (Depthformer SelfAttentionBlock core): kernel time at the model's shapes against
the bytes of the map it has to write, and against the same math in eager PyTorch (cuBLAS bmm + softmax)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from mde_b200 import ops  # noqa: E402,F401


def timed(fn, iters=30):
    for _ in range(3):
        fn()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6500.0
    for dtype in (torch.bfloat16, torch.float32):
        for b, s in ((8, 418), (8, 300), (1, 418)):
            q, k, v = (torch.randn(b, s, 256, device="cuda").to(dtype) for _ in range(3))
            us = timed(lambda: torch.ops.swin_attn.mha_fwd(q, k, v, 4, 0.125))

            def eager():
                qh, kh, vh = (t.view(b, s, 4, 64).transpose(1, 2).contiguous() for t in (q, k, v))
                at = torch.matmul(qh, kh.transpose(-2, -1))
                at *= 0.125
                at = torch.softmax(at.float(), dim=-1)
                return torch.matmul(at.to(vh.dtype), vh).transpose(1, 2).reshape(b, s, 256), at
            us_e = timed(eager)
            nbytes = b * 4 * s * s * 4 + 4 * b * s * 256 * q.element_size()
            print(json.dumps({"dtype": str(dtype).split(".")[1], "batch": b, "seq": s, "us": round(us, 2), "algorithmic_MB": round(nbytes / 1e6, 2),
                              "GBps": round(nbytes / us / 1e3, 1), "hbm_frac": round(nbytes / us / 1e3 / peak, 4),
                              "eager_torch_us": round(us_e, 2), "speedup_vs_eager": round(us_e / us, 2)}))


if __name__ == "__main__":
    main()
