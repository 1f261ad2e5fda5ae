#!/usr/bin/env python
"""(benchmark script, not collected by pytest; lives under tests/ because it executes the oracle)

BASELINE.json configs[4], "vs the reference PyTorch module": times the reference's eager op sequence for the
scoped path on the GPU -- pad, roll, window_partition, scale, QK^T bmm, bias gather, mask add, softmax, PV bmm,
head merge, window_reverse, roll, crop (oracle/window_attention.py, the restatement pinned against the
reference modules) -- next to the fused kernel, on the stage grids of NewCRFs Swin-L at batch 8.

    python tests/bench_reference_eager_gpu.py [--iters 10] [--only NAME] [--bwd]
    python tests/bench_reference_eager_gpu.py --model     # whole NewCRFs Swin-L forward, reference op sequence on the GPU

``--model`` is the "stronger bar" of SURVEY.md 8d / BASELINE.md 4: the UNMODIFIED reference ``NewCRFDepth("large07")`` (staged
copy ``oracle/_ref``, built by ``oracle/ref_model.py``; no product module or kernel in it) on the same B200 in fp32 and under
torch.autocast(bf16), batch 8 at 352x1216, beside this repo's model (harness.infer on a prepare_inference'd model).  The same
numbers are the ``reference_gpu_eager`` key of the default ``bench.py`` line.  (The per-op mode above times the oracle's
restatement of the scoped path, i.e. the reference's eager OP SEQUENCE, not its modules.)
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from mde_b200 import _lib  # noqa: E402
from oracle import window_attention as wa  # noqa: E402

SHAPES = {  # name: (B, H, W, nH, ws)
    "enc_s0_352x1216": (8, 88, 304, 6, 7), "enc_s1_352x1216": (8, 44, 152, 12, 7), "enc_s2_352x1216": (8, 22, 76, 24, 7),
    "enc_s3_352x1216": (8, 11, 38, 48, 7), "oda_ws12_s0": (8, 96, 96, 6, 12),
}


def timed(fn, iters, flush):
    for _ in range(2):
        fn()
    evs = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in evs) / iters * 1e3


def model_mode(iters):
    import mde_b200
    from mde_b200.harness import build_model, infer, prepare_inference, synthetic_images
    from oracle import ref_model
    if not ref_model.available():
        raise SystemExit("the staged reference (oracle/_ref) is missing: run __graft_entry__.build() where /root/reference exists")
    dev = torch.device("cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    model = build_model("large07", 80.0, seed=0, device=dev).eval()
    ref = ref_model.build_newcrfs("large07", 80.0, seed=0, synthetic=True, device=dev).eval()   # the reference's own modules
    images = synthetic_images(8, 352, 1216, seed=0).to(dev)
    out = {}
    with torch.no_grad():
        out["reference_eager_fp32_ms"] = timed(lambda: ref(images), iters, flush) / 1e3
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out["reference_eager_bf16_autocast_ms"] = timed(lambda: ref(images), iters, flush) / 1e3
        out["ours_fp32_ms"] = timed(lambda: infer(model, images, bf16=False), iters, flush) / 1e3
    prepare_inference(model)
    out["ours_bf16_autocast_ms"] = timed(lambda: infer(model, images), iters, flush) / 1e3
    out = {k: round(v, 3) for k, v in out.items()}
    out.update({"workload": "NewCRFs large07 forward, batch 8, 3x352x1216, random-init weights, L2 flushed between steps",
                "images_per_s": {k[:-3]: round(8e3 / v, 1) for k, v in out.items()},
                "speedup_vs_reference_bf16": round(out["reference_eager_bf16_autocast_ms"] / out["ours_bf16_autocast_ms"], 2),
                "speedup_vs_reference_fp32": round(out["reference_eager_fp32_ms"] / out["ours_fp32_ms"], 2)})
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", action="store_true")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--only", default=None)
    ap.add_argument("--bwd", action="store_true")
    args = ap.parse_args()
    if args.model:
        return model_mode(args.iters)
    dev = torch.device("cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for name, (b, h, w, nh, ws) in SHAPES.items():
        if args.only and args.only not in name:
            continue
        c = nh * 32
        hp, wp = -(-h // ws) * ws, -(-w // ws) * ws
        if args.bwd:
            h, w = hp, wp
        g = torch.Generator(device=dev).manual_seed(0)
        qkv = torch.randn(b, h, w, 3 * c, device=dev, generator=g).bfloat16().requires_grad_(args.bwd)
        table = (0.02 * torch.randn((2 * ws - 1) ** 2, nh, device=dev, generator=g)).requires_grad_(args.bwd)
        pad = (0.1 * torch.randn(3 * c, device=dev, generator=g)).bfloat16()
        d_o = torch.randn(b, h, w, c, device=dev, generator=g).bfloat16()
        for shift in (0, ws // 2):
            def eager():
                # the reference under autocast: bf16 bmm, fp32 bias/mask/softmax
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    o, _ = wa.window_attention(qkv[..., :c], qkv[..., c:2 * c], qkv[..., 2 * c:], table, ws, shift, nh, None,
                                               pad[:c], pad[c:2 * c], pad[2 * c:])
                if args.bwd:
                    torch.autograd.grad((o * d_o).sum(), [qkv, table])
                return o

            def fused():
                if args.bwd:
                    o, lse = torch.ops.swin_attn.fwd(qkv.detach(), None, table.detach(), None, None, ws, shift, nh, 32 ** -0.5, True,
                                                     _lib.IMPL_AUTO, None)
                    return torch.ops.swin_attn.bwd(d_o, qkv.detach(), None, o, lse, table.detach(), ws, shift, nh, 32 ** -0.5, None,
                                                   _lib.IMPL_AUTO)
                return torch.ops.swin_attn.fwd(qkv.detach(), None, table.detach(), pad, None, ws, shift, nh, 32 ** -0.5, False,
                                               _lib.IMPL_AUTO, None)
            t_ref, t_new = timed(eager, args.iters, flush), timed(fused, args.iters, flush)
            print(json.dumps({"case": name, "pass": "fwd+bwd" if args.bwd else "fwd", "shift": shift,
                              "reference_eager_us": round(t_ref, 1), "fused_us": round(t_new, 1),
                              "speedup": round(t_ref / t_new, 1)}), flush=True)


if __name__ == "__main__":
    main()
