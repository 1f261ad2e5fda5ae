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


def test_full_model_matches_reference(golden):
    g = golden("newcrfdepth_tiny07")
    cfg = g["cfg"]
    m = mde_b200.NewCRFDepth(version=cfg["version"], max_depth=cfg["max_depth"])
    m.load_state_dict(synthetic_state(m), strict=True)
    m = to_oracle(m).eval()
    with torch.no_grad():
        depth = m(g["img"])
    assert rel_err(depth, g["depth"]) < TOL
    gt = synthetic_depth_gt("newcrfdepth_tiny07.gt", cfg["b"], cfg["h"], cfg["w"], cfg["max_depth"])
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
