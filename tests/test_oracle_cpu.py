"""The oracle, pinned against outputs of the reference modules (tests/golden/,
made by oracle/make_golden.py from /root/reference).  CPU only."""
import pytest
import torch

import mde_b200
from mde_b200 import layers
from oracle import window_attention as wa
from oracle.cpu_model import to_oracle
from oracle.synthetic import depth_errors, synthetic_depth_gt, synthetic_state
from conftest import rel_err

TOL = 1e-5  # fp32, max-normalised (north_star)


def _grad(out, cot, inputs):
    return torch.autograd.grad((out * cot).sum(), inputs)


@pytest.mark.parametrize("name", ["crf_block_ws7_s0", "crf_block_ws7_s3", "crf_block_ws7_s3_exact"])
def test_crf_block_matches_reference(golden, name):
    g = golden(name)
    cfg = g["cfg"]
    blk = layers.CRFBlock(cfg["c"], cfg["nh"], cfg["c"], window_size=cfg["ws"], shift_size=cfg["shift"])
    blk.load_state_dict(synthetic_state(blk), strict=True)
    blk = to_oracle(blk).eval()
    blk.H, blk.W = cfg["h"], cfg["w"]
    x, v = g["x"].clone().requires_grad_(), g["v"].clone().requires_grad_()
    out = blk(x, v, None)
    assert rel_err(out, g["out"]) < TOL
    a = blk.attn
    gs = _grad(out, g["cot"], [x, v, a.relative_position_bias_table, a.qk.weight, a.qk.bias, a.proj.weight])
    for got, key in zip(gs, ["dx", "dv", "dtable", "dqk_w", "dqk_b", "dproj_w"]):
        assert rel_err(got, g[key]) < TOL, key


@pytest.mark.parametrize("name", ["swin_block_ws7_s0", "swin_block_ws7_s3", "swin_block_ws12_s6", "swin_block_ws12_s0"])
def test_swin_block_matches_reference(golden, name):
    g = golden(name)
    cfg = g["cfg"]
    blk = layers.SwinTransformerBlock(cfg["c"], cfg["nh"], window_size=cfg["ws"], shift_size=cfg["shift"])
    blk.load_state_dict(synthetic_state(blk), strict=True)
    blk = to_oracle(blk).eval()
    blk.H, blk.W = cfg["h"], cfg["w"]
    x = g["x"].clone().requires_grad_()
    out = blk(x, None)
    assert rel_err(out, g["out"]) < TOL
    a = blk.attn
    gs = _grad(out, g["cot"], [x, a.relative_position_bias_table, a.qkv.weight, a.qkv.bias])
    for got, key in zip(gs, ["dx", "dtable", "dqkv_w", "dqkv_b"]):
        assert rel_err(got, g[key]) < TOL, key


def test_masked_window_attention_matches_reference(golden):
    g = golden("masked_windows_ws7")
    cfg = g["cfg"]
    att = layers.WindowAttention(cfg["c"], (cfg["ws"],) * 2, cfg["nh"], cfg["c"])
    att.load_state_dict(synthetic_state(att), strict=True)
    att = to_oracle(att).eval()
    with torch.no_grad():
        assert rel_err(att(g["x"], g["v"], g["mask"]), g["out"]) < TOL
        assert rel_err(att(g["x"], g["v"], None), g["out_nomask"]) < TOL


def test_newcrf_level_matches_reference(golden):
    g = golden("newcrf_level")
    cfg = g["cfg"]
    m = layers.NewCRF(input_dim=cfg["cin"], embed_dim=cfg["embed"], v_dim=cfg["vdim"], window_size=7,
                      num_heads=cfg["nh"])
    m.load_state_dict(synthetic_state(m), strict=True)
    m = to_oracle(m).eval()
    x, v = g["x"].clone().requires_grad_(), g["v"].clone().requires_grad_()
    out = m(x, v)
    assert rel_err(out, g["out"]) < TOL
    gs = _grad(out, g["cot"], [x, v, m.crf_layer.blocks[1].attn.relative_position_bias_table])
    for got, key in zip(gs, ["dx", "dv", "dtable1"]):
        assert rel_err(got, g[key]) < TOL, key


@pytest.mark.parametrize("fixture", ["newcrfdepth_tiny07", "newcrfdepth_tiny12"])
def test_full_model_matches_reference(golden, fixture):
    g = golden(fixture)
    cfg = g["cfg"]
    m = mde_b200.NewCRFDepth(version=cfg["version"], max_depth=cfg["max_depth"])
    m.load_state_dict(synthetic_state(m), strict=True)
    m = to_oracle(m).eval()
    with torch.no_grad():
        depth = m(g["img"])
    assert rel_err(depth, g["depth"]) < TOL
    gt = synthetic_depth_gt(fixture + ".gt", cfg["b"], cfg["h"], cfg["w"], cfg["max_depth"])
    valid = (gt > 1e-3).numpy()
    err = depth_errors(gt.numpy()[valid], depth.clamp(1e-3, cfg["max_depth"]).numpy()[valid])
    assert abs(err["abs_rel"] - g["abs_rel"]) < 1e-3 and abs(err["rmse"] - g["rmse"]) < 1e-3


def test_shift_mask_regions_and_index():
    # negative controls: the restated mask must actually mask, the index must be the reference's
    m = wa.shift_mask(14, 21, 7, 3)
    assert m.shape == (6, 49, 49) and set(m.unique().tolist()) == {-100.0, 0.0}
    assert (m[0] == 0).all() and (m[-1] != 0).any()
    idx = wa.relative_position_index(7)
    assert idx.shape == (49, 49) and idx.min() == 0 and idx.max() == 168 and (idx.diagonal() == 84).all()


def test_empty_and_ragged_inputs():
    t = torch.zeros(169, 2)
    for h, w in [(1, 1), (7, 7), (8, 6), (13, 15)]:
        q = torch.randn(1, h, w, 64)
        o, lse = wa.window_attention(q, q, q, t, 7, 3 if min(h, w) > 1 else 0, 2)
        assert o.shape == q.shape and torch.isfinite(o).all()
        assert lse.shape == (1, -(-h // 7) * 7, -(-w // 7) * 7, 2)


def test_full_model_matches_reference_at_the_benchmarked_config(golden):
    """BASELINE configs[1]: the oracle's CPU model of ``large07`` on one 352x1216 image against the reference's depth map
    (tests/golden/newcrfdepth_large07_kitti.pt), AbsRel / RMSE on both synthetic GTs within 1e-3 absolute."""
    from oracle.synthetic import randn
    name = "newcrfdepth_large07_kitti"
    g = golden(name)
    cfg = g["cfg"]
    m = mde_b200.NewCRFDepth(version=cfg["version"], max_depth=cfg["max_depth"])
    m.load_state_dict(synthetic_state(m), strict=True)
    m = to_oracle(m).eval()
    with torch.no_grad():
        depth = m(randn(name + ".img", cfg["b"], 3, cfg["h"], cfg["w"]))
    assert rel_err(depth, g["depth"]) < TOL
    for tag, lo in (("", 1e-3), ("_realgt", 1.0)):
        gt = synthetic_depth_gt(name + ".gt", cfg["b"], cfg["h"], cfg["w"], cfg["max_depth"], min_depth=lo)
        valid = (gt > 1e-3).numpy()
        err = depth_errors(gt.numpy()[valid], depth.clamp(1e-3, cfg["max_depth"]).numpy()[valid])
        assert abs(err["abs_rel"] - g["abs_rel" + tag]) < 1e-3 and abs(err["rmse"] - g["rmse" + tag]) < 1e-3


def test_training_gradient_fixture_is_well_formed(golden):
    """tests/golden/newcrfdepth_large07_nyu_grads.pt: loss.backward() through the reference at 480x640 -- every
    attention parameter the fused path differentiates is present in full, every parameter has a norm entry."""
    g = golden("newcrfdepth_large07_nyu_grads")
    tables = [k for k in g["grads"] if k.endswith("relative_position_bias_table")]
    biases = [k for k in g["grads"] if k.endswith("qkv.bias") or k.endswith("qk.bias")]
    assert len(tables) == 24 + 8 and len(biases) == 24 + 8
    assert g["grads"]["backbone.layers.2.blocks.7.attn.relative_position_bias_table"].shape == (169, 24)
    assert g["grads"]["crf3.crf_layer.blocks.1.attn.qk.bias"].shape == (2048,)
    m = mde_b200.NewCRFDepth(version="large07", max_depth=10.0)
    assert set(g["grad_sums"]) == {k for k, _ in m.named_parameters()}
    assert all(torch.isfinite(v).all() and float(v.abs().max()) > 0 for v in g["grads"].values())
    assert g["loss"] > 0


def test_silog_restatement_equals_the_harness_loss():
    """oracle.synthetic.silog_loss (what the gradient fixture differentiates) == mde_b200.harness.SILogLoss."""
    from mde_b200.harness import SILogLoss
    from oracle.synthetic import silog_loss
    gen = torch.Generator().manual_seed(3)
    pred = torch.rand(3, 1, 20, 30, generator=gen) * 9 + 0.2
    gt = synthetic_depth_gt("silog.gt", 3, 20, 30, 10.0, sparsity=0.3, min_depth=0.5)
    assert abs(float(silog_loss(pred, gt)) - float(SILogLoss()(pred, gt))) < 1e-5


def test_staged_reference_is_a_byte_copy_and_imports():
    """oracle/_ref (when staged): every file matches MANIFEST.json's sha256, NewCRFDepth builds through the shim, and the
    reference's tcompute_errors equals the restated depth_errors."""
    import hashlib
    import json
    import os
    import numpy as np
    from oracle import ref_model, reference_shim as rs, stage_reference
    if not stage_reference.staged():
        pytest.skip("oracle/_ref not staged")
    man = json.load(open(os.path.join(stage_reference.DEST, "MANIFEST.json")))
    assert len(man["files"]) > 50
    for rel, sha in man["files"].items():
        assert hashlib.sha256(open(os.path.join(stage_reference.DEST, rel), "rb").read()).hexdigest() == sha, rel
    assert ref_model.available()
    m = ref_model.build_newcrfs("tiny07", 80.0, synthetic=True)
    ours = mde_b200.NewCRFDepth(version="tiny07", max_depth=80.0)
    assert [k for k, _ in m.named_parameters()] and set(m.state_dict()) == set(ours.state_dict())
    gt, pred = np.random.rand(100) * 70 + 1, np.random.rand(100) * 70 + 1
    ref_err = rs.module("utils.depth_utils").tcompute_errors(gt, pred)
    mine = depth_errors(gt, pred)
    assert abs(mine["abs_rel"] - ref_err["abs_rel"]) < 1e-12 and abs(mine["rmse"] - ref_err["rmse"]) < 1e-12


def test_reference_arm_of_the_bench_loads_no_product_library():
    """`bench.py --impl reference` times the unmodified reference and must not dlopen libswin_attn*.so."""
    import json
    import os
    import subprocess
    import sys
    from oracle import ref_model
    if not ref_model.available():
        pytest.skip("no reference tree (neither oracle/_ref nor /root/reference)")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = ("import sys, runpy\nsys.argv = ['bench.py', '--impl', 'reference', '--steps', '1', '--warmup', '0', '--version', 'tiny07']\n"
            "import bench\nbench.cpu_reference_configs0 = lambda threads: None\n"
            "try:\n    bench.main()\nexcept SystemExit:\n    pass\n"
            "print('MAPS', [l.split()[-1] for l in open('/proc/self/maps') if 'swin_attn' in l], file=sys.stderr)\n")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=root, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["cpu_baseline"]["kind"] == "reference" and line["value"] > 0
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["gpu_launches"] == 0
    assert "MAPS []" in r.stderr, r.stderr[-500:]
