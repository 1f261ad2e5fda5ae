"""Generate tests/golden/*.pt by RUNNING THE REFERENCE MODULES (fp32, CPU).

Run in the build container only (needs /root/reference):

    python -m oracle.make_golden

TEST INFRASTRUCTURE ONLY.  Each fixture holds seeded inputs, the reference's
outputs and the gradients autograd computes through the reference code; the
weights are regenerated from parameter names by oracle/synthetic.py, so the
fixtures stay a few hundred KB.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

from . import reference_shim as rs
from .synthetic import depth_errors, randn, silog_loss, synthetic_depth_gt, synthetic_state
from .window_attention import shift_mask

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _grads(out, inputs, seed_name):
    w = randn(seed_name + ".cotangent", *out.shape)
    gs = torch.autograd.grad((out * w).sum(), inputs, allow_unused=True)
    return w, gs


def crf_block(name, ws, shift, b, h, w, c, nh):
    L = rs.newcrf_layers()
    blk = L.CRFBlock(c, nh, c, window_size=ws, shift_size=shift).eval()
    blk.load_state_dict(synthetic_state(blk))
    x = randn(name + ".x", b, h * w, c).requires_grad_()
    v = randn(name + ".v", b, h, w, c).requires_grad_()
    hp, wp = -(-h // ws) * ws, -(-w // ws) * ws
    blk.H, blk.W = h, w
    out = blk(x, v, shift_mask(hp, wp, ws, ws // 2) if shift else None)
    params = [blk.attn.relative_position_bias_table, blk.attn.qk.weight, blk.attn.qk.bias, blk.attn.proj.weight]
    cot, gs = _grads(out, [x, v] + params, name)
    return dict(kind="crf_block", cfg=dict(ws=ws, shift=shift, b=b, h=h, w=w, c=c, nh=nh), x=x.detach(), v=v.detach(),
                out=out.detach(), cot=cot, dx=gs[0], dv=gs[1], dtable=gs[2], dqk_w=gs[3], dqk_b=gs[4], dproj_w=gs[5])


def swin_block(name, ws, shift, b, h, w, c, nh):
    S = rs.swin_transformer()
    blk = S.SwinTransformerBlock(c, nh, window_size=ws, shift_size=shift).eval()
    blk.load_state_dict(synthetic_state(blk))
    x = randn(name + ".x", b, h * w, c).requires_grad_()
    hp, wp = -(-h // ws) * ws, -(-w // ws) * ws
    blk.H, blk.W = h, w
    out = blk(x, shift_mask(hp, wp, ws, ws // 2) if shift else None)
    params = [blk.attn.relative_position_bias_table, blk.attn.qkv.weight, blk.attn.qkv.bias]
    cot, gs = _grads(out, [x] + params, name)
    return dict(kind="swin_block", cfg=dict(ws=ws, shift=shift, b=b, h=h, w=w, c=c, nh=nh), x=x.detach(),
                out=out.detach(), cot=cot, dx=gs[0], dtable=gs[1], dqkv_w=gs[2], dqkv_b=gs[3])


def oda2_stage(name, ws, b, h, w, c, nh):
    """model/ODA2 SwinTransformerStage (depth 2: un-shifted + shifted block, replicate padding, mask built per forward)."""
    S = rs.oda2_swin_transformer()
    st = S.SwinTransformerStage(dim=c, depth=2, num_heads=nh, window_size=ws).eval()
    st.load_state_dict(synthetic_state(st))
    x = randn(name + ".x", b, h * w, c).requires_grad_()
    out = st(x, h, w)[0]
    a0, a1 = st.blocks[0].attn, st.blocks[1].attn
    cot, gs = _grads(out, [x, a0.relative_position_bias_table, a1.relative_position_bias_table, a1.qkv.weight, a1.qkv.bias], name)
    return dict(kind="oda2_stage", cfg=dict(ws=ws, b=b, h=h, w=w, c=c, nh=nh), x=x.detach(), out=out.detach(), cot=cot,
                dx=gs[0], dtable0=gs[1], dtable1=gs[2], dqkv_w1=gs[3], dqkv_b1=gs[4])


def oda_block(name, ws, shift, b, h, w, c, nh):
    """model/ODA SwinTransformerBlock (timm block, fixed input_resolution, attn_mask buffer)."""
    S = rs.oda_swin_transformer()
    blk = S.SwinTransformerBlock(c, (h, w), nh, window_size=ws, shift_size=shift).eval()
    blk.load_state_dict(synthetic_state(blk))
    x = randn(name + ".x", b, h * w, c).requires_grad_()
    out = blk(x)
    cot, gs = _grads(out, [x, blk.attn.relative_position_bias_table, blk.attn.qkv.weight, blk.attn.qkv.bias], name)
    return dict(kind="oda_block", cfg=dict(ws=ws, shift=shift, b=b, h=h, w=w, c=c, nh=nh), x=x.detach(), out=out.detach(),
                cot=cot, dx=gs[0], dtable=gs[1], dqkv_w=gs[2], dqkv_b=gs[3])


def depthformer_vit(name, b, s, hidden, kq, nh, repeat):
    """model/Depthformer ViTLayer (pre-norm SelfAttentionBlock + FeedForwardBlock, shared weights over `repeat`), eval mode;
    cotangents on BOTH outputs (hidden and the returned attention map)."""
    V = rs.depthformer_vit_layer()
    m = V.ViTLayer(hidden, kq, nh, num_repeat=repeat).eval()
    m.load_state_dict(synthetic_state(m))
    x = randn(name + ".x", b, s, hidden).requires_grad_()
    out, attn = m(x)
    cot_o, cot_a = randn(name + ".cot_o", *out.shape), randn(name + ".cot_a", *attn.shape)
    sa = m.self_attn
    gs = torch.autograd.grad((out * cot_o).sum() + (attn * cot_a).sum(),
                             [x, sa.query_proj.weight, sa.key_proj.weight, sa.value_proj.weight, sa.out_proj.bias])
    return dict(kind="depthformer_vit", cfg=dict(b=b, s=s, hidden=hidden, kq=kq, nh=nh, repeat=repeat), x=x.detach(),
                out=out.detach(), attn=attn.detach(), cot_o=cot_o, cot_a=cot_a, dx=gs[0], dq_w=gs[1], dk_w=gs[2], dv_w=gs[3],
                dout_b=gs[4])


def depthformer_postnorm_sa(name, b, s, hidden, kq, nh):
    S = rs.depthformer_self_attention()
    m = S.PostNormSelfAttentionBlock(hidden, kq, nh).eval()
    m.load_state_dict(synthetic_state(m))
    x = randn(name + ".x", b, s, hidden).requires_grad_()
    out, attn = m(x)
    cot_o = randn(name + ".cot_o", *out.shape)
    gs = torch.autograd.grad((out * cot_o).sum(), [x, m.key_proj.weight])
    return dict(kind="depthformer_postnorm_sa", cfg=dict(b=b, s=s, hidden=hidden, kq=kq, nh=nh), x=x.detach(), out=out.detach(),
                attn=attn.detach(), cot_o=cot_o, dx=gs[0], dk_w=gs[1])


def depthformer_luna(name, cls, b, h, w, hidden, aux_dim, qk, nh, k_tokens):
    """model/Depthformer LunaLayer (post-norm) / PreNormLunaLayer: pack + unpack attention and the feed-forward, eval mode;
    cotangents on all four outputs."""
    Lm = rs.depthformer_luna_layer()
    m = getattr(Lm, cls)(hidden, aux_dim, qk, nh).eval()
    m.load_state_dict(synthetic_state(m))
    x = randn(name + ".x", b, hidden, h, w).requires_grad_()
    aux = randn(name + ".aux", b, k_tokens, aux_dim).requires_grad_()
    outs = m(x, aux)
    cots = [randn(f"{name}.cot{i}", *o.shape) for i, o in enumerate(outs)]
    la = m.luna_attn
    gs = torch.autograd.grad(sum((o * c).sum() for o, c in zip(outs, cots)), [x, aux, la.k1_proj.weight, la.q2_proj.weight, la.v2_proj.weight])
    return dict(kind="depthformer_luna", cfg=dict(cls=cls, b=b, h=h, w=w, hidden=hidden, aux_dim=aux_dim, qk=qk, nh=nh, k=k_tokens),
                x=x.detach(), aux=aux.detach(), outs=[o.detach() for o in outs], cots=cots, dx=gs[0], daux=gs[1], dk1_w=gs[2],
                dq2_w=gs[3], dv2_w=gs[4])


def oda2_ordered_sa(name, b, h, w, dims, nh, n_emb, r, shift):
    """model/ODA2 PreNormOrderedSwinSA: depth-order bias from per-pixel integer indices, window r, no shift mask; eval mode;
    cotangents on both outputs; the bias table is perturbed so that its gradient is not symmetric by construction."""
    D = rs.oda2_ordered_swin_decoder()
    torch.manual_seed(0)
    m = D.PreNormOrderedSwinSA(dims, nh, n_emb, window_size=r, shift_size=shift).eval()
    m.load_state_dict(synthetic_state(m))
    x = randn(name + ".x", b, h, w, dims).requires_grad_()
    g = torch.Generator().manual_seed(1234)
    indices = torch.randint(0, n_emb, (b, h, w), generator=g)
    out, attn = m(x, indices)
    cot_o, cot_a = randn(name + ".cot_o", *out.shape), randn(name + ".cot_a", *attn.shape)
    gs = torch.autograd.grad((out * cot_o).sum() + (attn * cot_a).sum(), [x, m.depth_embedding, m.k_proj.weight, m.v_proj.bias])
    return dict(kind="oda2_ordered_sa", cfg=dict(b=b, h=h, w=w, dims=dims, nh=nh, n_emb=n_emb, r=r, shift=shift), x=x.detach(),
                indices=indices, out=out.detach(), attn=attn.detach(), cot_o=cot_o, cot_a=cot_a, dx=gs[0], dbias=gs[1], dk_w=gs[2],
                dv_b=gs[3])


def masked_windows(name, ws, nw, groups, c, nh):
    """WindowAttention.forward(x, v, mask) on window-order tokens with an arbitrary mask."""
    L = rs.newcrf_layers()
    att = L.WindowAttention(c, (ws, ws), nh, c).eval()
    att.load_state_dict(synthetic_state(att))
    n = ws * ws
    x = randn(name + ".x", nw * groups, n, c)
    v = randn(name + ".v", nw * groups, n, c)
    mask = torch.where(randn(name + ".m", nw, n, n) > 0.8, -100.0, 0.0)
    with torch.no_grad():
        out, out_nomask = att(x, v, mask), att(x, v, None)
    return dict(kind="masked_windows", cfg=dict(ws=ws, nw=nw, groups=groups, c=c, nh=nh), x=x, v=v, mask=mask,
                out=out, out_nomask=out_nomask)


def newcrf_level(name, b, h, w, cin, embed, vdim, nh):
    L = rs.newcrf_layers()
    m = L.NewCRF(input_dim=cin, embed_dim=embed, v_dim=vdim, window_size=7, num_heads=nh).eval()
    m.load_state_dict(synthetic_state(m))
    x = randn(name + ".x", b, cin, h, w).requires_grad_()
    v = randn(name + ".v", b, vdim, h, w).requires_grad_()
    out = m(x, v)
    cot, gs = _grads(out, [x, v, m.crf_layer.blocks[1].attn.relative_position_bias_table], name)
    return dict(kind="newcrf_level", cfg=dict(b=b, h=h, w=w, cin=cin, embed=embed, vdim=vdim, nh=nh), x=x.detach(),
                v=v.detach(), out=out.detach(), cot=cot, dx=gs[0], dv=gs[1], dtable1=gs[2])


def full_model(name, version, b, h, w, max_depth=80.0):
    D = rs.newcrf_depth()
    m = D.NewCRFDepth(version=version, max_depth=max_depth).eval()
    m.load_state_dict(synthetic_state(m))
    img = randn(name + ".img", b, 3, h, w)
    with torch.no_grad():
        depth = m(img)
    gt = synthetic_depth_gt(name + ".gt", b, h, w, max_depth)
    valid = (gt > 1e-3).numpy()
    pred = depth.clamp(1e-3, max_depth).numpy()
    # the reference's own metric code (utils/depth_utils.py:32-54)
    sys.path.insert(0, rs.REFERENCE_ROOT)
    from utils.depth_utils import tcompute_errors
    ref_err = tcompute_errors(gt.numpy()[valid], pred[valid])
    mine = depth_errors(gt.numpy()[valid], pred[valid])
    assert abs(mine["abs_rel"] - ref_err["abs_rel"]) < 1e-12 and abs(mine["rmse"] - ref_err["rmse"]) < 1e-12
    return dict(kind="full_model", cfg=dict(version=version, b=b, h=h, w=w, max_depth=max_depth), img=img, depth=depth,
                abs_rel=float(ref_err["abs_rel"]), rmse=float(ref_err["rmse"]))


def _metrics(depth, gt, max_depth):
    """AbsRel / RMSE through the reference's own utils/depth_utils.py:32-54 (and the restatement agrees)."""
    tcompute_errors = rs.module("utils.depth_utils").tcompute_errors
    valid = (gt > 1e-3).numpy()
    pred = depth.clamp(1e-3, max_depth).numpy()
    ref_err = tcompute_errors(gt.numpy()[valid], pred[valid])
    mine = depth_errors(gt.numpy()[valid], pred[valid])
    assert abs(mine["abs_rel"] - ref_err["abs_rel"]) < 1e-12 and abs(mine["rmse"] - ref_err["rmse"]) < 1e-12
    return float(ref_err["abs_rel"]), float(ref_err["rmse"])


def full_model_v2(name, version, b, h, w, max_depth=80.0, keep_depth=True):
    """The benchmarked configuration itself: reference NewCRFDepth(version) on one synthetic image at the BASELINE shape;
    depth map + AbsRel / RMSE on the ill-conditioned GT (uniform(1e-3, max)) and on the realistic one (uniform(1, max))."""
    D = rs.newcrf_depth()
    m = D.NewCRFDepth(version=version, max_depth=max_depth).eval()
    m.load_state_dict(synthetic_state(m))
    img = randn(name + ".img", b, 3, h, w)
    with torch.no_grad():
        depth = m(img)
    out = dict(kind="full_model_v2", cfg=dict(version=version, b=b, h=h, w=w, max_depth=max_depth))
    for tag, lo in (("", 1e-3), ("_realgt", 1.0)):
        gt = synthetic_depth_gt(name + ".gt", b, h, w, max_depth, min_depth=lo)
        out["abs_rel" + tag], out["rmse" + tag] = _metrics(depth, gt, max_depth)
    if keep_depth:
        out["depth"] = depth
    return out


def full_model_grads(name, version, b, h, w, max_depth=10.0):
    """A training-shape gradient fixture: reference NewCRFDepth(version) in eval mode (DropPath off, BatchNorm on its
    running statistics: the deterministic half of SURVEY 8c "training parity"), SILog loss on a synthetic GT, backward.
    Kept: the loss, the FULL gradient of every relative_position_bias_table and of every qkv / qk bias (the parameters
    the fused attention path differentiates itself), and fp64 sum / L2 norm of every other parameter's gradient."""
    D = rs.newcrf_depth()
    m = D.NewCRFDepth(version=version, max_depth=max_depth).eval()
    m.load_state_dict(synthetic_state(m))
    img = randn(name + ".img", b, 3, h, w)
    gt = synthetic_depth_gt(name + ".gt", b, h, w, max_depth, sparsity=0.3, min_depth=0.5)
    depth = m(img)
    loss = silog_loss(depth, gt)
    loss.backward()
    full, sums = {}, {}
    for k, p in m.named_parameters():
        g = p.grad
        if k.endswith("relative_position_bias_table") or k.endswith("qkv.bias") or k.endswith("qk.bias"):
            full[k] = g.detach().clone()
        sums[k] = (float(g.double().sum()), float(g.double().norm()))
    return dict(kind="full_model_grads", cfg=dict(version=version, b=b, h=h, w=w, max_depth=max_depth), loss=float(loss),
                depth_mean=float(depth.mean()), grads=full, grad_sums=sums)


CASES = {
    "crf_block_ws7_s0": lambda n: crf_block(n, 7, 0, 2, 9, 11, 64, 2),
    "crf_block_ws7_s3": lambda n: crf_block(n, 7, 3, 2, 9, 11, 64, 2),
    "crf_block_ws7_s3_exact": lambda n: crf_block(n, 7, 3, 1, 14, 21, 128, 4),
    "swin_block_ws7_s0": lambda n: swin_block(n, 7, 0, 1, 10, 16, 96, 3),
    "swin_block_ws7_s3": lambda n: swin_block(n, 7, 3, 1, 10, 16, 96, 3),
    "swin_block_ws12_s6": lambda n: swin_block(n, 12, 6, 1, 13, 25, 64, 2),
    "swin_block_ws12_s0": lambda n: swin_block(n, 12, 0, 1, 24, 12, 64, 2),
    "masked_windows_ws7": lambda n: masked_windows(n, 7, 4, 2, 64, 2),
    "oda2_stage_ws7_replicate": lambda n: oda2_stage(n, 7, 2, 16, 23, 96, 3),
    "oda_block_ws12_s6": lambda n: oda_block(n, 12, 6, 1, 24, 36, 64, 2),
    "oda_block_ws12_s0": lambda n: oda_block(n, 12, 0, 1, 24, 24, 64, 2),
    "oda_block_ws7_s3": lambda n: oda_block(n, 7, 3, 2, 14, 21, 96, 3),
    # ODA2 depth-order decoder windows: window 8 shifted by 4 with head dim 64, window 4 un-shifted with head dim 32
    "oda2_ordered_sa_r8_s4": lambda n: oda2_ordered_sa(n, 1, 16, 24, 128, 2, 16, 8, 4),
    "oda2_ordered_sa_r4_s0": lambda n: oda2_ordered_sa(n, 2, 8, 12, 64, 2, 9, 4, 0),
    # Depthformer decoder attention: S = 11 x 10 tokens, head dim 64 (the model's) / 32 with a narrower key-query width
    "depthformer_vit_hd64": lambda n: depthformer_vit(n, 2, 110, 128, 128, 2, 2),
    "depthformer_vit_kq32": lambda n: depthformer_vit(n, 1, 75, 128, 64, 2, 1),
    "depthformer_postnorm_sa": lambda n: depthformer_postnorm_sa(n, 1, 97, 192, 192, 3),
    # Luna pack / unpack: K = 24 auxiliary tokens against 9 x 11 image tokens, head dim 64
    "depthformer_luna_postnorm": lambda n: depthformer_luna(n, "LunaLayer", 2, 9, 11, 128, 96, 128, 2, 24),
    "depthformer_luna_prenorm": lambda n: depthformer_luna(n, "PreNormLunaLayer", 1, 7, 9, 128, 128, 64, 2, 17),
    "newcrf_level": lambda n: newcrf_level(n, 1, 10, 13, 48, 64, 32, 2),
    "newcrfdepth_tiny07": lambda n: full_model(n, "tiny07", 1, 64, 96),
    # the window-12 encoder (version "..12", NewCRFDepth.py:26) on a 96 x 160 image: stage grids 24x40, 12x20, 6x10, 3x5 --
    # padding along W at every stage, along H from stage 2 on, and a last stage smaller than the shift
    "newcrfdepth_tiny12": lambda n: full_model(n, "tiny12", 1, 96, 160),
    # metrics only (same image / weights as newcrfdepth_tiny07; adds the realistic-range GT)
    "newcrfdepth_tiny07_metrics": lambda n: full_model_v2("newcrfdepth_tiny07", "tiny07", 1, 64, 96, keep_depth=False),
    # BASELINE configs[1]: the benchmarked model at the benchmarked shape, one image
    "newcrfdepth_large07_kitti": lambda n: full_model_v2(n, "large07", 1, 352, 1216),
    # BASELINE configs[2]: the training shape, loss.backward() through the reference
    "newcrfdepth_large07_nyu_grads": lambda n: full_model_grads(n, "large07", 1, 480, 640),
}


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(8)
    only = sys.argv[1:]
    for name, fn in CASES.items():
        if only and name not in only:
            continue
        blob = fn(name)
        blob = {k: (v.to(torch.float32).contiguous() if torch.is_tensor(v) else v) for k, v in blob.items()}
        path = os.path.join(OUT, name + ".pt")
        torch.save(blob, path)
        print(f"{name}: {os.path.getsize(path) / 1024:.0f} KB")


if __name__ == "__main__":
    main()
