#!/usr/bin/env python
"""Inside the real model (480x640 gradient fixture, bf16 autocast): at every window-attention call, run the tcgen05
forward and the independent CUDA-core forward on the SAME inputs and report how o and lse differ (mean / max, and where)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import mde_b200  # noqa: E402
from mde_b200 import _lib, ops  # noqa: E402
from oracle.synthetic import randn, synthetic_state  # noqa: E402


def main():
    name = "newcrfdepth_large07_nyu_grads"
    m = mde_b200.NewCRFDepth(version="large07", max_depth=10.0)
    m.load_state_dict(synthetic_state(m))
    m = m.cuda().eval()
    img = randn(name + ".img", 1, 3, 480, 640).cuda()
    orig = ops.window_attention
    n = [0]

    def wrapped(packed, v, table, window, shift, nh, scale=None, packed_pad=None, v_pad=None, **kw):
        out = orig(packed, v, table, window, shift, nh, scale, packed_pad=packed_pad, v_pad=v_pad, **kw)
        sc = float(scale if scale is not None else 32 ** -0.5)
        pp = None if packed_pad is None else packed_pad.to(packed.dtype)
        o1, l1 = torch.ops.swin_attn.fwd(packed, v, table, pp, v_pad, window, shift, nh, sc, True, _lib.IMPL_AUTO, None, 0.0, 0)
        o2, l2 = torch.ops.swin_attn.fwd(packed, v, table, pp, v_pad, window, shift, nh, sc, True, _lib.IMPL_SIMT, None, 0.0, 0)
        h, w = packed.shape[1:3]
        eo = (o1.float() - o2.float())
        e_real = (out.detach().float() - o2.float())      # the output of the REAL call path (autograd op when grads are on)
        el = (l1 - l2)[:, :h, :w]
        bad = el.abs() > 0.02
        where = ""
        if bad.any():
            idx = bad.nonzero()[:3].tolist()
            where = f" lse>0.02 at {int(bad.sum())} entries, e.g. (b,y,x,head) {idx}"
        print(f"call {n[0]:2d} {'x'.join(map(str, packed.shape))} nH{nh} shift{shift} v_ext={v is not None}: "
              f"o mean|d| {float(eo.abs().mean()):.2e} max {float(eo.abs().max()):.2e} (|o| max {float(o2.float().abs().max()):.2f}) | "
              f"REAL-path o mean|d| {float(e_real.abs().mean()):.2e} max {float(e_real.abs().max()):.2e} | "
              f"lse mean|d| {float(el.abs().mean()):.2e} max {float(el.abs().max()):.2e} |logit-ish| lse max {float(l2[:, :h, :w].abs().max()):.1f}{where}", flush=True)
        n[0] += 1
        return out

    ops.window_attention = wrapped
    grad_mode = len(sys.argv) > 1 and sys.argv[1] == "grad"
    with torch.set_grad_enabled(grad_mode), torch.autocast("cuda", dtype=torch.bfloat16):
        m(img)


if __name__ == "__main__":
    main()
