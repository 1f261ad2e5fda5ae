"""Micro-benchmark of the fused fc1 + GELU tcgen05 GEMM (swin_linear_forward) against the two-kernel path it replaces
(cuBLAS bf16 Linear with bias epilogue, then swin_gelu_forward) on the MLP shapes of NewCRFs Swin-L at 352x1216, batch 8.
Prints one JSON line per shape: times in us (CUDA events, cold L2 between iterations), algorithmic GB/s
(x + W + y once) and TFLOP/s of the fused kernel."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import mde_b200  # noqa: F401
from mde_b200 import _lib, ops  # noqa: F401

SHAPES = [  # name, rows, in, out
    ("enc.s0", 8 * 88 * 304, 192, 768), ("enc.s1", 8 * 44 * 152, 384, 1536), ("enc.s2", 8 * 22 * 76, 768, 3072),
    ("enc.s3", 8 * 11 * 38, 1536, 6144), ("crf0", 8 * 88 * 304, 128, 512), ("crf1", 8 * 44 * 152, 256, 1024),
    ("crf2", 8 * 22 * 76, 512, 2048), ("crf3", 8 * 11 * 38, 1024, 4096),
]


def timed(fn, flush, iters=10):
    for _ in range(3):
        fn()
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
    dev = "cuda"
    only = sys.argv[1] if len(sys.argv) > 1 else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for name, rows, k, n in SHAPES:
        if only and name != only:
            continue
        x = torch.randn(rows, k, device=dev).bfloat16()
        w = (torch.randn(n, k, device=dev) * k ** -0.5).bfloat16()
        b = torch.randn(n, device=dev).bfloat16()
        lib = _lib.load()
        lib.swin_linear_debug_set_pair_mode(0)                       # one CTA per 128 x 256 tile
        t_fused = timed(lambda: torch.ops.swin_attn.linear_act(x, w, b, 1), flush)
        t_plain = timed(lambda: torch.ops.swin_attn.linear_act(x, w, b, 0), flush)
        ref1 = torch.ops.swin_attn.linear_act(x, w, b, 1)
        lib.swin_linear_debug_set_pair_mode(1)                       # CTA pairs (cta_group::2), 256 x 256 tile over two SMs
        t_fused2 = timed(lambda: torch.ops.swin_attn.linear_act(x, w, b, 1), flush)
        t_plain2 = timed(lambda: torch.ops.swin_attn.linear_act(x, w, b, 0), flush)
        same = bool(torch.equal(ref1, torch.ops.swin_attn.linear_act(x, w, b, 1)))
        lib.swin_linear_debug_set_pair_mode(-1)
        t_cublas = timed(lambda: torch.nn.functional.linear(x, w, b), flush)
        h = torch.nn.functional.linear(x, w, b)
        t_gelu = timed(lambda: torch.ops.swin_attn.gelu(h), flush)
        byt = 2 * (rows * k + n * k + rows * n)
        print(json.dumps({"shape": name, "rows": rows, "in": k, "out": n, "fused_us": round(t_fused, 1), "plain_us": round(t_plain, 1),
                          "pair_fused_us": round(t_fused2, 1), "pair_plain_us": round(t_plain2, 1), "pair_equals_single": same,
                          "cublas_us": round(t_cublas, 1), "gelu_us": round(t_gelu, 1),
                          "speedup_vs_two_kernels": round((t_cublas + t_gelu) / t_fused, 2),
                          "fused_GBps": round(byt / t_fused / 1e3, 1), "fused_TFLOPs": round(2 * rows * k * n / t_fused / 1e6, 1)}))
        sys.stdout.flush()


if __name__ == "__main__":
    main()
