#!/usr/bin/env python
"""Inside the real model in training-like mode (grad on, eval modules, bf16 autocast, 480x640 fixture): for selected
attention calls (a) are the inputs of the op untouched by the forward kernel?  (b) does the gradient autograd produced
for `packed` equal a backward recomputed from a DIRECT forward call's lse on the same inputs?  (c) both against the
fp64 oracle's autograd."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import mde_b200  # noqa: E402
from mde_b200 import _lib, harness, ops  # noqa: E402
from oracle import window_attention as wa  # noqa: E402
from oracle.synthetic import randn, synthetic_depth_gt, synthetic_state  # noqa: E402

WATCH = {0, 4, 22, 23, 24}


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def main():
    name = "newcrfdepth_large07_nyu_grads"
    m = mde_b200.NewCRFDepth(version="large07", max_depth=10.0)
    m.load_state_dict(synthetic_state(m))
    m = m.cuda().eval()
    img = randn(name + ".img", 1, 3, 480, 640).cuda()
    gt = synthetic_depth_gt(name + ".gt", 1, 480, 640, 10.0, sparsity=0.3, min_depth=0.5).cuda()
    orig = ops.window_attention
    n = [0]
    rec = {}

    def wrapped(packed, v, table, window, shift, nh, scale=None, packed_pad=None, v_pad=None, **kw):
        i = n[0]
        n[0] += 1
        if i not in WATCH:
            return orig(packed, v, table, window, shift, nh, scale, packed_pad=packed_pad, v_pad=v_pad, **kw)
        before = packed.detach().clone()
        vb = None if v is None else v.detach().clone()
        out = orig(packed, v, table, window, shift, nh, scale, packed_pad=packed_pad, v_pad=v_pad, **kw)
        torch.cuda.synchronize()
        same = torch.equal(before, packed.detach()) and (v is None or torch.equal(vb, v.detach()))
        r = rec[i] = dict(packed=before, v=vb, table=table.detach(), pad=None if packed_pad is None else packed_pad.detach(),
                          shift=shift, nh=nh, scale=float(scale if scale is not None else 32 ** -0.5), inputs_untouched=same)
        out.register_hook(lambda g, r=r: r.__setitem__("d_o", g.detach().clone()))
        packed.register_hook(lambda g, r=r: r.__setitem__("d_packed", g.detach().clone()))
        return out

    ops.window_attention = wrapped
    with torch.autocast("cuda", dtype=torch.bfloat16):
        pred = m(img)
    harness.SILogLoss()(pred, gt).backward()
    torch.cuda.synchronize()
    for i, r in sorted(rec.items()):
        packed, v, table, pad = r["packed"], r["v"], r["table"], r["pad"]
        c = r["nh"] * 32
        padb = None if pad is None else pad.to(packed.dtype)
        o1, l1 = torch.ops.swin_attn.fwd(packed, v, table, padb, None, 7, r["shift"], r["nh"], r["scale"], True, _lib.IMPL_AUTO, None, 0.0, 0)
        g = torch.ops.swin_attn.bwd_unpadded(r["d_o"].contiguous(), packed, v, o1, l1, table, pad, None, 7, r["shift"], r["nh"], r["scale"])
        line = (f"call {i} {tuple(packed.shape)} shift{r['shift']} ext_v={v is not None}: inputs untouched by the forward: {r['inputs_untouched']}; "
                f"autograd d_packed vs recomputed-from-direct-forward: {rel(r['d_packed'], g[0]):.2e}")
        if packed.shape[1] <= 30:   # fp64 oracle on the small stages
            pr = packed.double().cpu().requires_grad_()
            tr = table.double().cpu()
            dd = None if pad is None else pad.double().cpu()
            if v is None:
                o_ref, _ = wa.window_attention(pr[..., :c], pr[..., c:2 * c], pr[..., 2 * c:], tr, 7, r["shift"], r["nh"], None,
                                               dd[:c], dd[c:2 * c], dd[2 * c:])
            else:
                o_ref, _ = wa.window_attention(pr[..., :c], pr[..., c:], v.double().cpu(), tr, 7, r["shift"], r["nh"], None, dd[:c], dd[c:], None)
            (gp,) = torch.autograd.grad((o_ref * r["d_o"].double().cpu()).sum(), [pr])
            line += f"; vs fp64 oracle: autograd {rel(r['d_packed'], gp):.2e}, recomputed {rel(g[0], gp):.2e}"
        print(line, flush=True)


if __name__ == "__main__":
    main()
