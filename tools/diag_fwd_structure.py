#!/usr/bin/env python
"""On REAL activations of the 480x640 fixture: error of the tcgen05 and of the CUDA-core forward against the fp64 oracle
for selected attention calls, with the structure of the tcgen05 error: by window, by head, by token row, signed."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import mde_b200  # noqa: E402
from mde_b200 import _lib, ops  # noqa: E402
from oracle import window_attention as wa  # noqa: E402
from oracle.synthetic import randn, synthetic_state  # noqa: E402

WATCH = {4, 22, 23}


def main():
    name = "newcrfdepth_large07_nyu_grads"
    m = mde_b200.NewCRFDepth(version="large07", max_depth=10.0)
    m.load_state_dict(synthetic_state(m))
    m = m.cuda().eval()
    img = randn(name + ".img", 1, 3, 480, 640).cuda()
    orig = ops.window_attention
    n = [0]

    def wrapped(packed, v, table, window, shift, nh, scale=None, packed_pad=None, v_pad=None, **kw):
        i = n[0]
        n[0] += 1
        out = orig(packed, v, table, window, shift, nh, scale, packed_pad=packed_pad, v_pad=v_pad, **kw)
        if i not in WATCH:
            return out
        sc = float(scale if scale is not None else 32 ** -0.5)
        c = nh * 32
        pp = packed_pad.to(packed.dtype)
        o1, _ = torch.ops.swin_attn.fwd(packed, v, table, pp, None, window, shift, nh, sc, True, _lib.IMPL_AUTO, None, 0.0, 0)
        o2, _ = torch.ops.swin_attn.fwd(packed, v, table, pp, None, window, shift, nh, sc, True, _lib.IMPL_SIMT, None, 0.0, 0)
        pk, pd = packed.double().cpu(), pp.double().cpu()
        want, _ = wa.window_attention(pk[..., :c], pk[..., c:2 * c], pk[..., 2 * c:], table.double().cpu(), window, shift, nh, None,
                                      pd[:c], pd[c:2 * c], pd[2 * c:])
        e1, e2 = o1.double().cpu() - want, o2.double().cpu() - want
        b, h, w, _ = packed.shape
        print(f"call {i} {tuple(packed.shape)} shift{shift}: |packed| max {float(packed.abs().max()):.1f} rms {float(packed.float().pow(2).mean().sqrt()):.2f}; "
              f"|o| rms {float(want.pow(2).mean().sqrt()):.3f}")
        for label, e in (("tcgen05", e1), ("cuda-core", e2)):
            print(f"   {label:9s} vs fp64: mean|e| {float(e.abs().mean()):.3e} rms {float(e.pow(2).mean().sqrt()):.3e} max {float(e.abs().max()):.3e} "
                  f"mean signed {float(e.mean()):+.3e}; mean of e*sign(o) {float((e * want.sign()).mean()):+.3e}")
        # coherent (row-scale) errors: alpha = <e, o> / <o, o> per (token, head) row of 32 channels
        for label, e in (("tcgen05", e1), ("cuda-core", e2)):
            er, orow = e.view(b, h, w, nh, 32), want.view(b, h, w, nh, 32)
            alpha = (er * orow).sum(-1) / orow.pow(2).sum(-1).clamp_min(1e-30)
            resid = er - alpha.unsqueeze(-1) * orow
            print(f"   {label:9s} row-scale error alpha: mean {float(alpha.mean()):+.3e} std {float(alpha.std()):.3e} max|a| {float(alpha.abs().max()):.3e}; "
                  f"share of squared error that is a row scale: {float((alpha.unsqueeze(-1) * orow).pow(2).sum() / er.pow(2).sum()):.3f}; residual rms {float(resid.pow(2).mean().sqrt()):.3e}")
            # windows with padding (last window row: image rows >= 14 at 15x20; last window col: cols >= 14..)
            big = alpha.abs() > 4 * alpha.std()
            if big.any():
                idx = big.nonzero()
                print(f"      |alpha| > 4 sigma at {int(big.sum())} rows; image rows {sorted(set(idx[:, 1].tolist()))[:20]} cols {sorted(set(idx[:, 2].tolist()))[:20]}")
        # structure of the tcgen05 error: per head, per window (un-shifted grid), per position inside the window
        eh = e1.view(b, h, w, nh, 32)
        per_head = eh.abs().mean(dim=(0, 1, 2, 4))
        print("   tcgen05 mean|e| per head: min %.2e median %.2e max %.2e (head %d)" % (float(per_head.min()), float(per_head.median()), float(per_head.max()), int(per_head.argmax())))
        eh2 = e2.view(b, h, w, nh, 32).abs().mean(dim=(0, 1, 2, 4))
        print("   cuda-core mean|e| per head: min %.2e median %.2e max %.2e" % (float(eh2.min()), float(eh2.median()), float(eh2.max())))
        rows = e1.abs().mean(dim=(0, 2, 3))
        cols = e1.abs().mean(dim=(0, 1, 3))
        print("   tcgen05 mean|e| per image row   : " + " ".join(f"{float(x) * 1e4:.1f}" for x in rows))
        print("   cuda-core mean|e| per image row : " + " ".join(f"{float(x) * 1e4:.1f}" for x in e2.abs().mean(dim=(0, 2, 3))))
        print("   tcgen05 mean|e| per image col   : " + " ".join(f"{float(x) * 1e4:.1f}" for x in cols))
        return out

    ops.window_attention = wrapped
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        m(img)


if __name__ == "__main__":
    main()
