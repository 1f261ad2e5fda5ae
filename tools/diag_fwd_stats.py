#!/usr/bin/env python
"""Error STATISTICS (not just the maximum) of the tcgen05 forward against the fp64 oracle on the stage shapes of the
480x640 training fixture: mean |err|, max |err| and mean signed error of o and lse, per region (rows of windows with /
without padding).  A max-normalised parity bound cannot see a 2x loss of average accuracy; this can."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from mde_b200 import _lib  # noqa: E402
from oracle import window_attention as wa  # noqa: E402
from oracle.synthetic import randn  # noqa: E402

CASES = [(1, 15, 20, 48, 0), (1, 15, 20, 48, 3), (1, 30, 40, 24, 0), (1, 30, 40, 24, 3), (1, 60, 80, 12, 0), (2, 21, 28, 4, 0)]


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "build"
    for b, h, w, nh, shift in CASES:
        c = nh * 32
        name = f"fstat{(b, h, w, nh, shift)}"
        qkv = randn(name + ".p", b, h, w, 3 * c).bfloat16()
        table = 0.5 * randn(name + ".t", 169, nh)
        pads = (0.3 * randn(name + ".pad", 3 * c)).bfloat16()
        q, k, v = (qkv[..., i * c:(i + 1) * c].double() for i in range(3))
        pd = pads.double()
        want, want_lse = wa.window_attention(q, k, v, table.double(), 7, shift, nh, None, pd[:c], pd[c:2 * c], pd[2 * c:])
        for impl, iname in ((_lib.IMPL_AUTO, "tc"), (_lib.IMPL_SIMT, "simt")):
            o, lse = torch.ops.swin_attn.fwd(qkv.cuda(), None, table.cuda(), pads.cuda(), None, 7, shift, nh, 32 ** -0.5, True, impl, None)
            eo = (o.double().cpu() - want)
            el = (lse.double().cpu()[:, :h, :w] - want_lse[:, :h, :w])
            # rows that belong to windows containing padding (last window row / column, before the shift)
            print(json.dumps({"build": tag, "case": [b, h, w, nh, shift], "impl": iname,
                              "o_mean_abs": float(eo.abs().mean()), "o_max_abs": float(eo.abs().max()), "o_mean_signed": float(eo.mean()),
                              "o_rms": float(eo.pow(2).mean().sqrt()),
                              "lse_mean_abs": float(el.abs().mean()), "lse_max_abs": float(el.abs().max()), "lse_mean_signed": float(el.mean())}), flush=True)


if __name__ == "__main__":
    main()
