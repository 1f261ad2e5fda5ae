#!/usr/bin/env python
"""Per-parameter bf16 gradient errors at the 480x640 training fixture (tests/golden/newcrfdepth_large07_nyu_grads.pt):
product vs the reference's fp32 gradients, next to the UNMODIFIED reference's own autocast run, twice (run-to-run
spread of the atomics).  Evidence for the bounds in tests/test_reference_gpu.py."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import mde_b200  # noqa: E402
from mde_b200 import harness  # noqa: E402
from oracle import ref_model  # noqa: E402
from oracle.synthetic import randn, synthetic_depth_gt, synthetic_state  # noqa: E402

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def grads(model, img, gt):
    for p in model.parameters():
        p.grad = None
    with torch.autocast("cuda", dtype=torch.bfloat16):
        pred = model(img)
    harness.SILogLoss()(pred, gt).backward()
    return {k: p.grad.detach().float().cpu() for k, p in model.named_parameters()}


def main():
    name = "newcrfdepth_large07_nyu_grads"
    g = torch.load(os.path.join(ROOT, "tests", "golden", name + ".pt"), weights_only=False)
    cfg = g["cfg"]
    img = randn(name + ".img", cfg["b"], 3, cfg["h"], cfg["w"]).cuda()
    gt = synthetic_depth_gt(name + ".gt", cfg["b"], cfg["h"], cfg["w"], cfg["max_depth"], sparsity=0.3, min_depth=0.5).cuda()
    runs = {}
    m = mde_b200.NewCRFDepth(version=cfg["version"], max_depth=cfg["max_depth"])
    m.load_state_dict(synthetic_state(m))
    m = m.cuda().eval()
    runs["product_a"] = grads(m, img, gt)
    runs["product_b"] = grads(m, img, gt)
    del m
    torch.cuda.empty_cache()
    if ref_model.available():
        r = ref_model.build_newcrfs(cfg["version"], cfg["max_depth"], synthetic=True, device="cuda")
        runs["reference_autocast"] = grads(r, img, gt)
        del r
    # every parameter: relative deviation of the gradient NORM from the reference's fp32 value, median per group
    def group(k):
        if k.startswith("backbone.layers."):
            kind = "attn" if ".attn." in k else "mlp" if ".mlp." in k else "norm" if ".norm" in k else "other"
            return "enc.stage" + k.split(".")[2] + "." + kind
        return k.split(".")[0]
    groups = {}
    for k, (_s, nrm) in g["grad_sums"].items():
        if nrm > 0:
            groups.setdefault(group(k), []).append(k)
    print("group (median |norm - ref| / ref)".ljust(40) + "".join(n.rjust(20) for n in runs))
    for gname in sorted(groups):
        row = gname.ljust(40)
        for n in runs:
            errs = sorted(abs(float(runs[n][k].double().norm()) - g["grad_sums"][k][1]) / g["grad_sums"][k][1] for k in groups[gname])
            row += f"{errs[len(errs) // 2]:20.3e}"
        print(row)
    keys = [k for k in g["grads"] if k.endswith("relative_position_bias_table")]
    print("table".ljust(62) + "".join(n.rjust(20) for n in runs))
    for k in keys:
        print(k.ljust(62) + "".join(f"{rel(runs[n][k], g['grads'][k]):20.3e}" for n in runs))
    for kind, sel in (("tables", lambda k: k.endswith("relative_position_bias_table")), ("biases", lambda k: not k.endswith("relative_position_bias_table"))):
        for n in runs:
            errs = sorted(rel(runs[n][k], g["grads"][k]) for k in g["grads"] if sel(k))
            print(json.dumps({"kind": kind, "run": n, "max": errs[-1], "median": errs[len(errs) // 2], "mean": sum(errs) / len(errs)}))


if __name__ == "__main__":
    main()
