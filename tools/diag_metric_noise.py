#!/usr/bin/env python
"""How far do AbsRel / RMSE on the synthetic GTs move under bf16 autocast -- for the product and for the UNMODIFIED
reference (oracle/_ref) on the same GPU -- relative to the reference's fp32 golden values?  Evidence for the bounds in
tests/test_reference_gpu.py.  (Round 2 used it to A/B a single-pass upper-bound softmax against the exact two-pass
form: profiles/r2_softmax_bound_vs_exact.md.)"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import mde_b200  # noqa: E402
from oracle import ref_model  # noqa: E402
from oracle.synthetic import depth_errors, randn, synthetic_depth_gt, synthetic_state  # noqa: E402

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def main():
    tag = os.environ.get("DIAG_TAG", "bf16")
    for name, mname, gname in (("newcrfdepth_tiny07", "newcrfdepth_tiny07_metrics", "newcrfdepth_tiny07"),
                               ("newcrfdepth_large07_kitti", "newcrfdepth_large07_kitti", "newcrfdepth_large07_kitti")):
        g = torch.load(os.path.join(ROOT, "tests", "golden", gname + ".pt"), weights_only=False)
        gm = torch.load(os.path.join(ROOT, "tests", "golden", mname + ".pt"), weights_only=False)
        cfg = g["cfg"]
        img = (g["img"] if "img" in g else randn(name + ".img", cfg["b"], 3, cfg["h"], cfg["w"])).cuda()
        gts = {lo: synthetic_depth_gt(name + ".gt", cfg["b"], cfg["h"], cfg["w"], cfg["max_depth"], min_depth=lo) for lo in (1.0, 1e-3)}

        def metrics(depth, lo):
            gt = gts[lo]
            valid = (gt > 1e-3).numpy()
            return depth_errors(gt.numpy()[valid], depth.clamp(1e-3, cfg["max_depth"]).numpy()[valid])

        models = {}
        m = mde_b200.NewCRFDepth(version=cfg["version"], max_depth=cfg["max_depth"])
        m.load_state_dict(synthetic_state(m))
        models["product_" + tag] = m.cuda().eval()
        if ref_model.available():
            models["reference_autocast"] = ref_model.build_newcrfs(cfg["version"], cfg["max_depth"], synthetic=True, device="cuda")
        for label, model in models.items():
            with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
                d = model(img).float().cpu()
            real, ill = metrics(d, 1.0), metrics(d, 1e-3)
            diff = (d - g["depth"]).double()
            print(json.dumps({"fixture": name, "model": label,
                              "d_abs_rel_realgt": abs(real["abs_rel"] - gm["abs_rel_realgt"]), "d_rmse_realgt_rel": abs(real["rmse"] - gm["rmse_realgt"]) / gm["rmse_realgt"],
                              "d_abs_rel_illgt_rel": abs(ill["abs_rel"] - gm["abs_rel"]) / gm["abs_rel"],
                              "abs_rel_realgt_golden": gm["abs_rel_realgt"],
                              "depth_max_norm_err": float(diff.abs().max() / g["depth"].abs().max()),
                              "depth_mean_signed_err_m": float(diff.mean()), "depth_rms_err_m": float(diff.pow(2).mean().sqrt())}), flush=True)
        del models
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
