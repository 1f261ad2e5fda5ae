#!/usr/bin/env python
"""Which output of the tcgen05 forward degrades the model-level bf16 gradients?  Every window-attention call of the
480x640 gradient fixture goes through a custom autograd function whose forward takes o and lse from the tcgen05 kernel
or from the independent CUDA-core kernel; the backward is always swin_attn::bwd_unpadded.  The sweep adds ONE bf16 ulp
to ONE element of the output of encoder call k.  Optional third family: an older build of the library at
tools/_v9/libswin_attn_v9.so (git show <commit>:.../csrc/*.cu compiled as in csrc/build.sh), called through the C ABI.
Result of round 2: profiles/r2_grad_metric_bimodal.md."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import mde_b200  # noqa: E402
from mde_b200 import _lib, harness, ops  # noqa: E402
from oracle.synthetic import randn, synthetic_depth_gt, synthetic_state  # noqa: E402

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
MODE = {"o": "tc", "lse": "tc", "flip": None}
CALL = [0]
_V9 = None


def v9_forward(packed, v, table, padb, shift, nh, scale):
    """o from the committed (v9) kernel: tools/_v9/libswin_attn_v9.so through the C ABI."""
    import ctypes
    global _V9
    if _V9 is None:
        _V9 = ctypes.CDLL(os.path.join(ROOT, "tools", "_v9", "libswin_attn_v9.so"))
        _V9.swin_attn_forward.argtypes = [ctypes.POINTER(_lib.Shape)] + [ctypes.c_void_p] * 11
        _V9.swin_attn_forward.restype = ctypes.c_int
    pk = packed.contiguous()
    b, h, w, pc = pk.shape
    c, es = nh * 32, pk.element_size()
    o = torch.empty(b, h, w, c, dtype=pk.dtype, device=pk.device)
    lse = torch.empty(b, -(-h // 7) * 7, -(-w // 7) * 7, nh, device=pk.device)
    tb = table.float().contiguous()
    shp = _lib.Shape(b, h, w, nh, 7, shift, _lib.BF16, _lib.IMPL_AUTO, scale, 0.0, 0, pc, pc, pc if v is None else c, c, 0, 0)
    base, pb = pk.data_ptr(), padb.data_ptr()
    vptr = base + 2 * c * es if v is None else v.contiguous().data_ptr()
    rc = _V9.swin_attn_forward(ctypes.byref(shp), base, base + c * es, vptr, pb, pb + c * es, pb + 2 * c * es if v is None else None,
                               tb.data_ptr(), None, o.data_ptr(), lse.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    return o, lse


class Mixed(torch.autograd.Function):
    @staticmethod
    def forward(ctx, packed, v, table, pad, shift, nh, scale):
        padb = None if pad is None else pad.to(packed.dtype)
        outs = {}
        for name, impl in (("tc", _lib.IMPL_AUTO), ("simt", _lib.IMPL_SIMT)):
            if name in MODE.values():
                outs[name] = torch.ops.swin_attn.fwd(packed, v, table, padb, None, 7, shift, nh, scale, True, impl, None, 0.0, 0)
        if "v9" in MODE.values():
            outs["v9"] = v9_forward(packed, v, table, padb.contiguous(), shift, nh, scale)
        if MODE.get("check") and "tc" in outs:
            ref = v9_forward(packed, v, table, padb.contiguous(), shift, nh, scale)
            d = (outs["tc"][0].float() - ref[0].float())
            ne = d != 0
            if int(ne.sum()) > 1:
                px = ne.any(dim=3)[0]
                ys, xs = px.nonzero(as_tuple=True)
                hd = ne.view(*ne.shape[:3], nh, 32).any(dim=4).any(dim=0).any(dim=0).any(dim=0).nonzero().flatten().tolist()
                print(f"   in-flow call {CALL[0]} {tuple(packed.shape)} shift{shift}: tc(v10) vs v9 differ in {int(ne.sum())} elements, max|d| {float(d.abs().max()):.3e}; "
                      f"{int(px.sum())} pixels y {int(ys.min())}..{int(ys.max())} x {int(xs.min())}..{int(xs.max())}; heads {hd[:12]}{'...' if len(hd) > 12 else ''} ({len(hd)}); "
                      f"packed contiguous {packed.is_contiguous()} strides {packed.stride()}", flush=True)
        o, lse = outs[MODE["o"]][0], outs[MODE["lse"]][1]
        if MODE.get("clone"):
            o = o.clone()
        if MODE.get("sync"):
            torch.cuda.synchronize()
        if MODE["flip"] is not None and CALL[0] == MODE["flip"]:     # one element, one bf16 ulp
            o = o.clone()
            flat = o.view(torch.int16).view(-1)
            flat[flat.numel() // 3] += 1
        CALL[0] += 1
        ctx.save_for_backward(packed, v if v is not None else torch.empty(0), table, pad if pad is not None else torch.empty(0), o, lse)
        ctx.meta = (shift, nh, scale, v is not None, pad is not None)
        return o

    @staticmethod
    def backward(ctx, d_o):
        packed, v, table, pad, o, lse = ctx.saved_tensors
        shift, nh, scale, has_v, has_pad = ctx.meta
        g = torch.ops.swin_attn.bwd_unpadded(d_o.contiguous(), packed, v if has_v else None, o, lse, table, pad if has_pad else None, None,
                                             7, shift, nh, scale)
        d_pad = g[3].to(pad.dtype).view(pad.shape) if has_pad else None
        return g[0], (g[1] if has_v else None), g[2], d_pad, None, None, None


def patched(packed, v, table, window, shift, nh, scale=None, packed_pad=None, v_pad=None, **kw):
    assert window == 7 and v_pad is None
    return Mixed.apply(packed, v, table, packed_pad, shift, nh, float(scale if scale is not None else 32 ** -0.5))


def rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def main():
    name = "newcrfdepth_large07_nyu_grads"
    g = torch.load(os.path.join(ROOT, "tests", "golden", name + ".pt"), weights_only=False)
    m = mde_b200.NewCRFDepth(version="large07", max_depth=10.0)
    m.load_state_dict(synthetic_state(m))
    m = m.cuda().eval()
    img = randn(name + ".img", 1, 3, 480, 640).cuda()
    gt = synthetic_depth_gt(name + ".gt", 1, 480, 640, 10.0, sparsity=0.3, min_depth=0.5).cuda()
    stock = ops.window_attention
    # PSP branches: pooled (1x1 .. 6x6) -> 1x1 conv -> GroupNorm -> ReLU: units that act on the whole image
    acts = {}
    for bi, branch in enumerate(m.decoder.psp_modules):
        branch[1].activate.register_forward_hook(lambda mod, inp, out, bi=bi: acts.__setitem__(bi, (inp[0].detach().float().cpu().clone(), out.detach().float().cpu().clone())))
    saved = {}
    have_v9 = os.path.exists(os.path.join(ROOT, "tools", "_v9", "libswin_attn_v9.so"))   # optional second build (see the docstring)
    sweep = [("o=tc lse=tc", "tc", "tc", True, None, {}), ("o=simt lse=simt", "simt", "simt", True, None, {})]
    if have_v9:
        sweep.append(("o=v9lib lse=v9lib", "v9", "v9", True, None, {}))
    for k in range(0, 24, 2):
        sweep.append((f"o=tc flip@{k}", "tc", "tc", True, k, {}))
        if have_v9:
            sweep.append((f"o=v9lib flip@{k}", "v9", "v9", True, k, {}))
        sweep.append((f"o=simt flip@{k}", "simt", "simt", True, k, {}))
    for label, o_src, l_src, patch, flip, extra in sweep:
        MODE["o"], MODE["lse"], MODE["flip"] = o_src, l_src, flip
        MODE["clone"], MODE["sync"], MODE["check"] = extra.get("clone"), extra.get("sync"), False
        CALL[0] = 0
        ops.window_attention = patched if patch else stock
        for p in m.parameters():
            p.grad = None
        with torch.autocast("cuda", dtype=torch.bfloat16):
            pred = m(img)
        loss = harness.SILogLoss()(pred, gt)
        loss.backward()
        saved[label] = dict(loss=float(loss))
        gr = {k: p.grad.detach().float().cpu() for k, p in m.named_parameters()}
        enc = sorted(rel(gr[k], g["grads"][k]) for k in g["grads"] if k.endswith("relative_position_bias_table") and k.startswith("backbone"))
        dec = sorted(rel(gr[k], g["grads"][k]) for k in g["grads"] if k.endswith("relative_position_bias_table") and not k.startswith("backbone"))
        print(f"{label:22s} enc median {enc[len(enc) // 2]:.4f} max {enc[-1]:.4f} | dec median {dec[len(dec) // 2]:.4f}", flush=True)
        continue
        print(json.dumps({"mode": label, "enc_tables_median": enc[len(enc) // 2], "enc_tables_max": enc[-1],
                          "dec_tables_median": dec[len(dec) // 2], "dec_tables_max": dec[-1],
                          "s3b0": rel(gr["backbone.layers.3.blocks.0.attn.relative_position_bias_table"],
                                      g["grads"]["backbone.layers.3.blocks.0.attn.relative_position_bias_table"])}), flush=True)


    base = "o=v9lib lse=v9lib"
    for label, sv in list(saved.items())[:0]:
        line = f"{label}: loss {sv['loss']:.7f}; depth vs [{base}]: max|d| {float((sv['pred'] - saved[base]['pred']).abs().max()):.3e} mean|d| {float((sv['pred'] - saved[base]['pred']).abs().mean()):.3e}"
        for bi in sorted(sv["acts"]):
            on_a, on_b = sv["acts"][bi][1] > 0, saved[base]["acts"][bi][1] > 0
            pre = sv["acts"][bi][0]
            line += f"; PSP branch {bi}: {int((on_a != on_b).sum())} of {on_a.numel()} ReLU units differ"
        print(line)
    _MAIN_DONE = True


if __name__ == "__main__":
    main()
