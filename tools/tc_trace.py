#!/usr/bin/env python
"""Timeline of CTA 0 of the tcgen05 forward kernel (debug hook swin_attn_debug_set_trace).
Prints, per local pair, SM-clock stamps (relative, in cycles) of the producer / MMA / softmax events.
Needs a trace build of the library:  SWIN_TC_TRACE=1 bash monocular-depth-estimation_b200/csrc/build.sh
(the default build compiles the stamps out)."""
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
    shift = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    stage = int(sys.argv[2]) if len(sys.argv) > 2 else 2          # Swin-L stage grid at 352x1216, batch 8
    b, ws = 8, 7
    h, w, nh = {0: (88, 304, 6), 1: (44, 152, 12), 2: (22, 76, 24), 3: (11, 38, 48)}[stage]
    dev = torch.device("cuda")
    c = nh * 32
    qkv = torch.randn(b, h, w, 3 * c, device=dev).bfloat16()
    table = 0.02 * torch.randn(169, nh, device=dev)
    pad = (0.1 * torch.randn(3 * c, device=dev)).bfloat16()
    lib = _lib.load()
    lib.swin_attn_debug_set_trace.argtypes = [ctypes.c_void_p]
    trace = torch.zeros(16384, dtype=torch.int64, device=dev)   # 256 x 16 pair stamps of CTA 0, then (entry, exit) ns per CTA at 4096

    def run():
        return torch.ops.swin_attn.fwd(qkv, None, table, pad, None, ws, shift, nh, 32 ** -0.5, False, _lib.IMPL_AUTO, None)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    torch.empty(256 << 20, dtype=torch.uint8, device=dev).zero_()   # cold L2, as between the blocks of the model
    lib.swin_attn_debug_set_trace(ctypes.c_void_p(trace.data_ptr()))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run()
    e1.record()
    torch.cuda.synchronize()
    lib.swin_attn_debug_set_trace(None)
    cta = trace.cpu()[4096:4096 + 2 * 148].view(148, 2)
    cta = cta[cta[:, 1] > 0]
    g0 = int(cta[:, 0].min())
    ent, ext = (cta[:, 0] - g0).float() / 1e3, (cta[:, 1] - g0).float() / 1e3
    print(f"stage {stage} shift {shift}: {cta.shape[0]} CTAs; CUDA events around the launch {e0.elapsed_time(e1) * 1e3:.1f} us; "
          f"globaltimer: CTA entry {ent.min():.1f}..{ent.max():.1f} us, exit min/median/max "
          f"{ext.min():.1f}/{ext.median():.1f}/{ext.max():.1f} us, CTA lifetime min/median/max "
          f"{(ext - ent).min():.1f}/{(ext - ent).median():.1f}/{(ext - ent).max():.1f} us")
    t = trace.cpu()[:4096].view(256, 16)
    t0 = int(t[:, :14][t[:, :14] > 0].min())
    print(f"kernel entry {int(t[0][14]) - t0}, exit {int(t[0][15]) - t0} (cycles relative to the first stamp)")
    print("pair " + " ".join(f"{e:>12s}" for e in EV))
    for n in range(0, 256, 1 if stage >= 2 else 8):
        if int(t[n].max()) == 0:
            break
        print(f"{n:4d} " + " ".join(f"{(int(v) - t0) if v > 0 else -1:12d}" for v in t[n][:len(EV)]))


    t2 = trace.cpu()[8192:8192 + 4096].view(256, 16)
    if int(t2.max()) > 0:
        ev2 = ["pass1", "pass2", "bar1", "-", "o_staged", "bar2", "loop_end"]
        # PV issuer (absolute, relative to t0): loop top, V landed, before the P wait, MMAs issued
        print("PV issuer: pair   loop_top   v_landed  before_P   (Pready)  mmas_out  (PV_issued)")
        for n in range(0, 256, 1 if stage >= 2 else 8):
            if int(t[n].max()) == 0:
                break
            f = lambda v: (int(v) - t0) if int(v) > 0 else -1
            print(f"           {n:4d} {f(t2[n][8]):10d} {f(t2[n][9]):10d} {f(t2[n][10]):10d} {f(t[n][5]):10d} {f(t2[n][11]):10d} {f(t[n][6]):10d}")
        print("softmax sub-phases of tile 0 row 0 (cycles since sm_ld; then since sm_Rarrive)")
        print("pair " + " ".join(f"{e:>10s}" for e in ev2))
        for n in range(0, 256, 1 if stage >= 2 else 8):
            if int(t[n].max()) == 0:
                break
            base1, base2 = int(t[n][9]), int(t[n][13])
            vals = [int(t2[n][i]) - (base1 if i < 4 else base2) if int(t2[n][i]) > 0 else -1 for i in range(7)]
            print(f"{n:4d} " + " ".join(f"{v:10d}" for v in vals))


if __name__ == "__main__":
    main()
