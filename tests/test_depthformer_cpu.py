"""SURVEY §8f row 4, CPU side: the oracle's restatement of the Depthformer decoder attention (ViTLayer = pre-norm
SelfAttentionBlock + FeedForwardBlock; PostNormSelfAttentionBlock) pinned against fixtures generated from the reference
modules (oracle/make_golden.py: model/Depthformer/{vit_layer,self_attention,feed_forward}.py)."""
import ctypes

import pytest
import torch

from mde_b200 import _lib, depthformer
from oracle.cpu_model import to_oracle
from oracle.synthetic import synthetic_state
from conftest import rel_err

TOL = 1e-5  # fp32, max-normalised


def vit_from_golden(g):
    cfg = g["cfg"]
    m = depthformer.ViTLayer(cfg["hidden"], cfg["kq"], cfg["nh"], num_repeat=cfg["repeat"])
    m.load_state_dict(synthetic_state(m), strict=True)
    return m.eval()


def postnorm_from_golden(g):
    cfg = g["cfg"]
    m = depthformer.PostNormSelfAttentionBlock(cfg["hidden"], cfg["kq"], cfg["nh"])
    m.load_state_dict(synthetic_state(m), strict=True)
    return m.eval()


@pytest.mark.parametrize("name", ["depthformer_vit_hd64", "depthformer_vit_kq32"])
def test_vit_layer_oracle_matches_reference(golden, name):
    g = golden(name)
    m = to_oracle(vit_from_golden(g)).eval()
    x = g["x"].clone().requires_grad_()
    out, attn = m(x)
    assert rel_err(out, g["out"]) < TOL and rel_err(attn, g["attn"]) < TOL
    sa = m.self_attn
    gs = torch.autograd.grad((out * g["cot_o"]).sum() + (attn * g["cot_a"]).sum(),
                             [x, sa.query_proj.weight, sa.key_proj.weight, sa.value_proj.weight, sa.out_proj.bias])
    for got, key in zip(gs, ["dx", "dq_w", "dk_w", "dv_w", "dout_b"]):
        assert rel_err(got, g[key]) < TOL, key


def test_postnorm_block_oracle_matches_reference(golden):
    g = golden("depthformer_postnorm_sa")
    m = to_oracle(postnorm_from_golden(g)).eval()
    x = g["x"].clone().requires_grad_()
    out, attn = m(x)
    assert rel_err(out, g["out"]) < TOL and rel_err(attn, g["attn"]) < TOL
    dx, dk = torch.autograd.grad((out * g["cot_o"]).sum(), [x, m.key_proj.weight])
    assert rel_err(dx, g["dx"]) < TOL and rel_err(dk, g["dk_w"]) < TOL


def luna_from_golden(g):
    cfg = g["cfg"]
    m = getattr(depthformer, cfg["cls"])(cfg["hidden"], cfg["aux_dim"], cfg["qk"], cfg["nh"])
    m.load_state_dict(synthetic_state(m), strict=True)
    return m.eval()


def check_luna(m, g, x, aux, bound, check_fn):
    outs = m(x, aux)
    for i, (o, want) in enumerate(zip(outs, g["outs"])):
        check_fn(rel_err(o.float(), want), bound, f"{g['cfg']['cls']} output {i}")
    la = m.luna_attn
    gs = torch.autograd.grad(sum((o.float() * c.to(o.device)).sum() for o, c in zip(outs, g["cots"])),
                             [x, aux, la.k1_proj.weight, la.q2_proj.weight, la.v2_proj.weight])
    for got, key in zip(gs, ["dx", "daux", "dk1_w", "dq2_w", "dv2_w"]):
        check_fn(rel_err(got, g[key]), bound, f"{g['cfg']['cls']} {key}")
    return outs


@pytest.mark.parametrize("name", ["depthformer_luna_postnorm", "depthformer_luna_prenorm"])
def test_luna_layer_oracle_matches_reference(golden, name):
    g = golden(name)
    m = to_oracle(luna_from_golden(g)).eval()

    def plain(err, bound, what):
        assert err < bound, what
    outs = check_luna(m, g, g["x"].clone().requires_grad_(), g["aux"].clone().requires_grad_(), TOL, plain)
    cfg = g["cfg"]
    assert outs[2].shape == (cfg["b"], cfg["nh"], cfg["k"], cfg["h"] * cfg["w"])      # pack map (B, nH, K, HW)
    assert outs[3].shape == (cfg["b"], cfg["nh"], cfg["h"] * cfg["w"], cfg["k"])      # unpack map (B, nH, HW, K)


def test_reference_interface_and_errors():
    m = depthformer.ViTLayer(128, 64, 2, num_repeat=2)
    assert set(m.state_dict()) == {f"self_attn.{p}.{w}" for p in ("norm", "query_proj", "key_proj", "value_proj", "out_proj")
                                   for w in ("weight", "bias")} | {f"feed_forward.{p}.{w}" for p in ("norm", "fc1", "fc2")
                                                                   for w in ("weight", "bias")}
    assert m.self_attn.head_dim == 32 and m.self_attn.attn_scale == (1.0 / 32) ** 0.5
    with pytest.raises(ValueError, match="Hidden dim not multiple of num heads"):
        depthformer.SelfAttentionBlock(100, 64, 3)
    with pytest.raises(ValueError, match="num_repeat"):
        depthformer.ViTLayer(128, 64, 2, num_repeat=0)
    with pytest.raises(RuntimeError, match="no CPU path"):      # the product fails loudly without CUDA tensors
        m(torch.zeros(1, 4, 128))


def test_c_abi_argument_checks():
    lib = _lib.load()
    shp = _lib.MhaShape(1, 8, 0, 2, 96, 64, _lib.F32, 0, 0.125, 0, 192, 192, 128, 128)
    dummy = ctypes.c_void_p(ctypes.addressof(ctypes.create_string_buffer(64)))
    assert lib.mha_attention_forward(ctypes.byref(shp), dummy, dummy, dummy, dummy, None, None) != 0
    assert "head dims must be in 1..64" in lib.swin_attn_last_error().decode()
    shp = _lib.MhaShape(1, 8, 4000, 2, 64, 64, _lib.F32, 0, 0.125, 0, 128, 128, 128, 128)
    assert lib.mha_attention_forward(ctypes.byref(shp), dummy, dummy, dummy, dummy, None, None) != 0
    assert "shared memory" in lib.swin_attn_last_error().decode()
    shp = _lib.MhaShape(1, 8, 0, 2, 64, 64, _lib.F32, 0, 0.125, 0, 100, 128, 128, 128)
    assert lib.mha_attention_forward(ctypes.byref(shp), dummy, dummy, dummy, dummy, None, None) != 0
    assert "row strides" in lib.swin_attn_last_error().decode()
    assert lib.mha_attention_forward(ctypes.byref(shp), None, dummy, dummy, dummy, None, None) != 0
    shp = _lib.MhaShape(0, 8, 0, 2, 64, 64, _lib.F32, 0, 0.125, 0, 128, 128, 128, 128)      # empty batch: nothing to do
    assert lib.mha_attention_forward(ctypes.byref(shp), dummy, dummy, dummy, dummy, None, None) == 0
