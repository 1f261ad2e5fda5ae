#!/usr/bin/env python
"""Timeline of CTA 0 of the tcgen05 BACKWARD kernel (trace build: SWIN_TC_TRACE=1 bash .../csrc/build.sh).
Per local pair, SM-clock stamps (cycles, relative) of the producer / phase-1 issuer / phase-2 issuer / softmax events."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from mde_b200 import _lib  # noqa: E402

EV = ["prod_issue", "p1_top", "p1_landed", "p1_fixed", "p1_issued", "p2_Pseen", "p2_issued",
      "sm_wait", "sm_Sready", "sm_loaded", "sm_math", "sm_Pdone", "sm_Oready", "sm_released", "sm_stored"]


def main():
    stage = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    shift = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    b = 8
    h, w, nh = {0: (120, 160, 6), 1: (60, 80, 12), 2: (30, 40, 24), 3: (15, 20, 48)}[stage]   # 480x640 training grids (un-padded)
    c = nh * 32
    dev = "cuda"
    qkv = torch.randn(b, h, w, 3 * c, device=dev).bfloat16()
    table = 0.02 * torch.randn(169, nh, device=dev)
    pad = (0.1 * torch.randn(3 * c, device=dev)).bfloat16()
    d_o = torch.randn(b, h, w, c, device=dev).bfloat16()
    sc = 32 ** -0.5
    o, lse = torch.ops.swin_attn.fwd(qkv, None, table, pad, None, 7, shift, nh, sc, True, _lib.IMPL_AUTO, None, 0.0, 0)
    lib = _lib.load()
    lib.swin_attn_debug_set_trace.argtypes = [ctypes.c_void_p]
    trace = torch.zeros(16384, dtype=torch.int64, device=dev)

    def run():
        return torch.ops.swin_attn.bwd_unpadded(d_o, qkv, None, o, lse, table, pad.float(), None, 7, shift, nh, sc)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    torch.empty(256 << 20, dtype=torch.uint8, device=dev).zero_()
    lib.swin_attn_debug_set_trace(ctypes.c_void_p(trace.data_ptr()))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run()
    e1.record()
    torch.cuda.synchronize()
    lib.swin_attn_debug_set_trace(None)
    t = trace.cpu()[12288:12288 + 240 * 16].view(240, 16)
    used = t[:, :15][t[:, :15] > 0]
    if used.numel() == 0:
        print("no stamps: not a trace build (SWIN_TC_TRACE=1 bash monocular-depth-estimation_b200/csrc/build.sh)")
        return
    t0 = int(used.min())
    print(f"backward stage {stage} ({b}x{h}x{w}, {nh} heads) shift {shift}: CUDA events around the launch {e0.elapsed_time(e1) * 1e3:.1f} us")
    print("pair " + " ".join(f"{e:>11s}" for e in EV))
    once = ["entry", "bars_init", "tmem_alloc", "fills_done", "setup_barrier", "wg0_last_stored", "both_wgs_done", "grads_added", "exit"]
    print("once: " + "  ".join(f"{name}={int(t[239][i]) - t0 if t[239][i] > 0 else -1}" for i, name in enumerate(once)))
    for n in range(239):
        if int(t[n].max()) == 0:
            break
        print(f"{n:4d} " + " ".join(f"{(int(v) - t0) if v > 0 else -1:11d}" for v in t[n][:len(EV)]))


if __name__ == "__main__":
    main()
