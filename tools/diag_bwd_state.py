#!/usr/bin/env python
"""Does the tcgen05 BACKWARD depend on what the previously running kernel left on the SMs (shared memory / tensor memory)?
Same inputs, same lse; the backward is launched right after (a) a tcgen05 forward, (b) a CUDA-core forward, (c) a kernel
that fills all shared memory with NaN patterns, and compared with the fp64 oracle's autograd and with each other."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from mde_b200 import _lib  # noqa: E402
from oracle import window_attention as wa  # noqa: E402
from oracle.synthetic import randn  # noqa: E402


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def main():
    for b, h, w, nh, shift in [(1, 15, 20, 48, 0), (1, 15, 20, 48, 3), (1, 30, 40, 24, 0), (8, 15, 20, 48, 0)]:
        c = nh * 32
        name = f"bstate{(b, h, w, nh, shift)}"
        qkv = randn(name + ".p", b, h, w, 3 * c).bfloat16().cuda()
        table = (0.5 * randn(name + ".t", 169, nh)).cuda()
        pads = (0.3 * randn(name + ".pad", 3 * c)).bfloat16().cuda()
        d_o = randn(name + ".c", b, h, w, c).bfloat16().cuda()
        sc = 32 ** -0.5
        o, lse = torch.ops.swin_attn.fwd(qkv, None, table, pads, None, 7, shift, nh, sc, True, _lib.IMPL_AUTO, None, 0.0, 0)
        # fp64 oracle gradients
        pr, tr, dr = qkv.double().cpu().requires_grad_(), table.double().cpu().requires_grad_(), pads.double().cpu().requires_grad_()
        o_ref, _ = wa.window_attention(pr[..., :c], pr[..., c:2 * c], pr[..., 2 * c:], tr, 7, shift, nh, None, dr[:c], dr[c:2 * c], dr[2 * c:])
        g_ref = torch.autograd.grad((o_ref * d_o.double().cpu()).sum(), [pr, tr, dr])

        def bwd():
            return torch.ops.swin_attn.bwd_unpadded(d_o, qkv, None, o, lse, table, pads.float(), None, 7, shift, nh, sc)

        results = {}
        for label, impl in (("after_tc_fwd", _lib.IMPL_AUTO), ("after_simt_fwd", _lib.IMPL_SIMT)):
            torch.ops.swin_attn.fwd(qkv, None, table, pads, None, 7, shift, nh, sc, True, impl, None, 0.0, 0)
            results[label] = bwd()
            torch.cuda.synchronize()
        big = torch.full((64 << 20,), float("nan"), device="cuda")   # touch L2 / nothing on smem, but a different predecessor
        results["after_fill"] = bwd()
        torch.cuda.synchronize()
        del big
        line = f"{(b, h, w, nh, shift)}: "
        for label, r in results.items():
            line += f"{label}: d_qkv {rel(r[0], g_ref[0]):.3e} d_table {rel(r[2], g_ref[1]):.3e} d_pad {rel(r[3], g_ref[2]):.3e} | "
        a, bb = results["after_tc_fwd"], results["after_simt_fwd"]
        line += f"tc-vs-simt predecessor: d_qkv {rel(a[0], bb[0]):.2e} d_table {rel(a[2], bb[2]):.2e} d_pad {rel(a[3], bb[3]):.2e}"
        print(line, flush=True)


if __name__ == "__main__":
    main()
