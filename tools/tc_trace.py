#!/usr/bin/env python
"""Timeline of CTA 0 of the tcgen05 forward kernel (debug hook swin_attn_debug_set_trace).
Prints, per local pair, SM-clock stamps (relative, in cycles) of the producer / MMA / softmax events."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from mde_b200 import _lib  # noqa: E402

EV = ["prod_start", "prod_issued", "mma_full", "mma_slotfree", "S_issued", "mma_Pready", "PV_issued",
      "sm_wait", "sm_Sready", "sm_ld", "sm_math", "sm_Parrive", "sm_Oready", "sm_Rarrive"]


def main():
    b, h, w, nh, ws = 8, 22, 76, 24, 7
    shift = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    dev = torch.device("cuda")
    c = nh * 32
    qkv = torch.randn(b, h, w, 3 * c, device=dev).bfloat16()
    table = 0.02 * torch.randn(169, nh, device=dev)
    pad = (0.1 * torch.randn(3 * c, device=dev)).bfloat16()
    lib = _lib.load()
    lib.swin_attn_debug_set_trace.argtypes = [ctypes.c_void_p]
    trace = torch.zeros(64 * 16, dtype=torch.int64, device=dev)

    def run():
        return torch.ops.swin_attn.fwd(qkv, None, table, pad, None, ws, shift, nh, 32 ** -0.5, False, _lib.IMPL_AUTO, None)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    torch.empty(256 << 20, dtype=torch.uint8, device=dev).zero_()   # cold L2, as between the blocks of the model
    lib.swin_attn_debug_set_trace(ctypes.c_void_p(trace.data_ptr()))
    run()
    torch.cuda.synchronize()
    lib.swin_attn_debug_set_trace(None)
    t = trace.cpu().view(64, 16)
    t0 = int(t[:, :14][t[:, :14] > 0].min())
    print(f"kernel entry {int(t[0][14]) - t0}, exit {int(t[0][15]) - t0} (cycles relative to the first stamp)")
    print("pair " + " ".join(f"{e:>12s}" for e in EV))
    for n in range(30):
        if int(t[n].max()) == 0:
            break
        print(f"{n:4d} " + " ".join(f"{(int(v) - t0) if v > 0 else -1:12d}" for v in t[n][:len(EV)]))


if __name__ == "__main__":
    main()
