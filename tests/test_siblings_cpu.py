"""SURVEY §8f row 3, CPU side: the oracle's restatement of the ODA2 (replicate padding) and ODA (timm block, fixed
attn_mask buffer) Swin blocks pinned against fixtures generated from the reference modules (oracle/make_golden.py:
model/ODA2/oda2_swin_transformer.py SwinTransformerStage, model/ODA/oda_swin_transformer.py SwinTransformerBlock)."""
import pytest
import torch

from mde_b200 import siblings
from oracle.cpu_model import to_oracle
from oracle.synthetic import synthetic_state
from oracle import window_attention as wa
from conftest import rel_err

TOL = 1e-5  # fp32, max-normalised


def oda2_stage_from_golden(g):
    cfg = g["cfg"]
    st = siblings.Oda2SwinTransformerStage(dim=cfg["c"], depth=2, num_heads=cfg["nh"], window_size=cfg["ws"])
    st.load_state_dict(synthetic_state(st), strict=True)
    return st.eval()


def oda_block_from_golden(g):
    cfg = g["cfg"]
    blk = siblings.OdaSwinTransformerBlock(cfg["c"], (cfg["h"], cfg["w"]), cfg["nh"], window_size=cfg["ws"], shift_size=cfg["shift"])
    blk.load_state_dict(synthetic_state(blk), strict=True)
    return blk.eval()


def test_oda2_stage_oracle_matches_reference(golden):
    g = golden("oda2_stage_ws7_replicate")
    cfg = g["cfg"]
    st = to_oracle(oda2_stage_from_golden(g)).eval()
    x = g["x"].clone().requires_grad_()
    out = st(x, cfg["h"], cfg["w"])[0]
    assert rel_err(out, g["out"]) < TOL
    a0, a1 = st.blocks[0].attn, st.blocks[1].attn
    gs = torch.autograd.grad((out * g["cot"]).sum(), [x, a0.relative_position_bias_table, a1.relative_position_bias_table,
                                                     a1.qkv.weight, a1.qkv.bias])
    for got, key in zip(gs, ["dx", "dtable0", "dtable1", "dqkv_w1", "dqkv_b1"]):
        assert rel_err(got, g[key]) < TOL, key


def test_oda2_block_with_the_callers_mask_matrix(golden):
    """The public block signature forward(x, mask_matrix): the reference's mask (standard) and a caller-made one."""
    g = golden("oda2_stage_ws7_replicate")
    cfg = g["cfg"]
    st = to_oracle(oda2_stage_from_golden(g)).eval()
    h, w, ws = cfg["h"], cfg["w"], cfg["ws"]
    hp, wp = -(-h // ws) * ws, -(-w // ws) * ws
    blk = st.blocks[1]
    blk.H, blk.W = h, w
    std = siblings.sw_msa_mask(hp, wp, ws, ws // 2)
    assert torch.equal(std, wa.shift_mask(hp, wp, ws, ws // 2))
    with torch.no_grad():
        x = g["x"]
        a = blk(x, std)                                  # verified standard -> analytic form
        b = blk._forward_standard(x)
        assert torch.equal(a, b)
        other = std.clone()
        other[0, 0, 1] = -100.0                          # a caller-made mask goes the explicit way and changes the result
        c = blk(x, other)
        assert rel_err(c, a) > 1e-7
        d = blk(x, std.clone())                          # the explicit way with the standard mask equals the analytic one
        blk._mask_check._seen.clear()
        assert rel_err(d, a) < TOL


@pytest.mark.parametrize("name", ["oda_block_ws12_s6", "oda_block_ws12_s0", "oda_block_ws7_s3"])
def test_oda_block_oracle_matches_reference(golden, name):
    g = golden(name)
    blk = to_oracle(oda_block_from_golden(g)).eval()
    x = g["x"].clone().requires_grad_()
    out = blk(x)
    assert rel_err(out, g["out"]) < TOL
    a = blk.attn
    gs = torch.autograd.grad((out * g["cot"]).sum(), [x, a.relative_position_bias_table, a.qkv.weight, a.qkv.bias])
    for got, key in zip(gs, ["dx", "dtable", "dqkv_w", "dqkv_b"]):
        assert rel_err(got, g[key]) < TOL, key


def test_oda_block_keeps_the_reference_interface(golden):
    g = golden("oda_block_ws12_s6")
    blk = oda_block_from_golden(g)
    cfg = g["cfg"]
    assert "attn_mask" not in blk.state_dict() and blk.attn_mask.shape == ((cfg["h"] // 12) * (cfg["w"] // 12), 144, 144)
    assert torch.equal(blk.attn_mask, wa.shift_mask(cfg["h"], cfg["w"], 12, 6))
    keys = set(blk.state_dict())
    assert {"norm1.weight", "attn.relative_position_bias_table", "attn.relative_position_index", "attn.qkv.weight", "attn.qkv.bias",
            "attn.proj.weight", "attn.proj.bias", "norm2.bias", "mlp.fc1.weight", "mlp.fc2.bias"} <= keys
    blk.reset_resolution((24, 24))                       # as written in the reference: records the size, keeps the mask
    assert blk.input_resolution == (24, 24) and blk.attn_mask.shape[0] == (cfg["h"] // 12) * (cfg["w"] // 12)
    small = siblings.OdaSwinTransformerBlock(64, (6, 6), 2, window_size=7, shift_size=3)
    assert small.window_size == 6 and small.shift_size == 0 and small.attn_mask is None   # oda_swin_transformer.py:55-58
    with pytest.raises(AssertionError):
        to_oracle(blk)(torch.zeros(1, 5, cfg["c"]))


def test_integration_recipe_names_exist_in_the_reference():
    """INTEGRATION.md §1 'Sibling model families': every attribute the recipe assigns exists in the reference modules
    (staged copy or /root/reference) and in the drop-in modules."""
    from oracle import reference_shim as rs
    if not rs.available():
        pytest.skip("reference not staged")
    from mde_b200 import depthformer as dfm
    oda2, oda = rs.oda2_swin_transformer(), rs.oda_swin_transformer()
    for name, ours in [("WindowAttention", siblings.Oda2WindowAttention), ("SwinTransformerBlock", siblings.Oda2SwinTransformerBlock),
                       ("SwinTransformerStage", siblings.Oda2SwinTransformerStage), ("PatchMerging", siblings.Oda2PatchMerging)]:
        assert hasattr(oda2, name) and isinstance(ours, type)
    assert hasattr(oda, "SwinTransformerBlock")
    sa, luna = rs.depthformer_self_attention(), rs.depthformer_luna_layer()
    for mod, names in [(sa, ["SelfAttentionBlock", "PostNormSelfAttentionBlock"]), (luna, ["LunaBlock", "PreNormLunaBlock", "LunaLayer", "PreNormLunaLayer"]),
                       (rs.depthformer_vit_layer(), ["ViTLayer", "SelfAttentionBlock", "FeedForwardBlock"])]:
        for n in names:
            assert hasattr(mod, n) and hasattr(dfm, n), n
    # same state_dict keys as the reference modules
    ref = luna.PreNormLunaBlock(64, 32, 64, 2)
    assert set(ref.state_dict()) == set(dfm.PreNormLunaBlock(64, 32, 64, 2).state_dict())
    ref = oda2.SwinTransformerStage(dim=64, depth=2, num_heads=2, window_size=7, downsample=oda2.PatchMerging)
    ours = siblings.Oda2SwinTransformerStage(dim=64, depth=2, num_heads=2, window_size=7, downsample=siblings.Oda2PatchMerging)
    assert set(ref.state_dict()) == set(ours.state_dict())
    ref = oda.SwinTransformerBlock(64, (14, 14), 2, window_size=7, shift_size=3)
    assert set(ref.state_dict()) == set(siblings.OdaSwinTransformerBlock(64, (14, 14), 2, window_size=7, shift_size=3).state_dict())


def ordered_sa_from_golden(g):
    cfg = g["cfg"]
    torch.manual_seed(0)
    m = siblings.PreNormOrderedSwinSA(cfg["dims"], cfg["nh"], cfg["n_emb"], window_size=cfg["r"], shift_size=cfg["shift"])
    m.load_state_dict(synthetic_state(m), strict=True)
    return m.eval()


@pytest.mark.parametrize("name", ["oda2_ordered_sa_r8_s4", "oda2_ordered_sa_r4_s0"])
def test_oda2_ordered_window_attention_oracle_matches_reference(golden, name):
    """model/ODA2/oda2_red_order_swin2_decoder.py PreNormOrderedSwinSA: depth-order bias, windows of 8 / 4, no shift mask."""
    g = golden(name)
    m = to_oracle(ordered_sa_from_golden(g)).eval()
    x = g["x"].clone().requires_grad_()
    out, attn = m(x, g["indices"])
    assert rel_err(out, g["out"]) < TOL and rel_err(attn, g["attn"]) < TOL
    gs = torch.autograd.grad((out * g["cot_o"]).sum() + (attn * g["cot_a"]).sum(), [x, m.depth_embedding, m.k_proj.weight, m.v_proj.bias])
    for got, key in zip(gs, ["dx", "dbias", "dk_w", "dv_b"]):
        assert rel_err(got, g[key]) < TOL, key
    cfg = g["cfg"]
    n_win = cfg["b"] * (cfg["h"] // cfg["r"]) * (cfg["w"] // cfg["r"])
    assert attn.shape == (n_win, cfg["nh"], cfg["r"] ** 2, cfg["r"] ** 2)
    assert set(m.state_dict()) == {"depth_embedding"} | {f"{p}.{w}" for p in ("norm", "q_proj", "k_proj", "v_proj", "o_proj") for w in ("weight", "bias")}
    with pytest.raises(AssertionError):
        siblings.PreNormOrderedSwinSA(64, 2, 8, window_size=7)      # the reference accepts windows of 4, 8, 16 only (:45)
