#!/usr/bin/env python
"""SURVEY §8f rows 3 and 4 on the GPU: the sibling consumers of the kernels, product module vs the UNMODIFIED reference module
(oracle/_ref, eager PyTorch) on the same B200, same weights, bf16 autocast and fp32 -- one JSON line per case with both
times and the max-normalised difference of the outputs.  The stand-in for BASELINE configs[3] ("Depthformer ..."), whose
whole model cannot be built offline (hub encoder): the ODA2 Swin-L stage that carries 18 of the 24 blocks, an ODA
(timm) Swin-L/384 stage-0 block pair at window 12, and the Depthformer decoder's ViT / Luna layers at 352x1216.

    python tools/siblings_bench.py > profiles/r2_siblings_bench.jsonl
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from mde_b200 import depthformer, siblings  # noqa: E402
from oracle import reference_shim as rs  # noqa: E402
from oracle.synthetic import synthetic_state  # noqa: E402

DEV = "cuda"


def timed(fn, iters=10):
    for _ in range(3):
        fn()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def rel(a, b):
    return float((a.float() - b.float()).abs().max() / b.float().abs().max())


def first(x):
    return x[0] if isinstance(x, (tuple, list)) else x


def case(name, ours, ref, args, shape_note):
    ours, ref = ours.to(DEV).eval(), ref.to(DEV).eval()
    ours.load_state_dict(synthetic_state(ours))
    ref.load_state_dict(synthetic_state(ref))
    out = {"case": name, "shape": shape_note}
    with torch.no_grad():
        for mode in ("bf16_autocast", "fp32"):
            ctx = torch.autocast("cuda", dtype=torch.bfloat16, enabled=mode == "bf16_autocast")
            with ctx:
                o, r = first(ours(*args)), first(ref(*args))
                out[mode] = {"product_ms": round(timed(lambda: ours(*args)), 4), "reference_eager_ms": round(timed(lambda: ref(*args)), 4),
                             "max_normalised_diff": rel(o, r)}
            out[mode]["speedup"] = round(out[mode]["reference_eager_ms"] / out[mode]["product_ms"], 2)
    print(json.dumps(out), flush=True)


class _Stage(torch.nn.Module):
    """(x) -> stage(x, h, w)[0] so that both stages take one argument."""

    def __init__(self, stage, h, w):
        super().__init__()
        self.stage, self.h, self.w = stage, h, w

    def forward(self, x):
        return self.stage(x, self.h, self.w)[0]

    def load_state_dict(self, sd):
        return self.stage.load_state_dict({k[len("stage."):]: v for k, v in sd.items()})


def main():
    if not rs.available():
        raise SystemExit("the staged reference (oracle/_ref) is missing: run __graft_entry__.build() where /root/reference exists")
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    b = 8
    oda2, oda = rs.oda2_swin_transformer(), rs.oda_swin_transformer()
    # ODA2 Swin-L stage 2 at 352x1216 (22 x 76 tokens per image: replicate padding to 28 x 77), 18 blocks, C = 768, 24 heads
    h, w, c, nh, depth = 22, 76, 768, 24, 18
    x = torch.randn(b, h * w, c, device=DEV)
    case("oda2_swin_large_stage2", _Stage(siblings.Oda2SwinTransformerStage(dim=c, depth=depth, num_heads=nh, window_size=7), h, w),
         _Stage(oda2.SwinTransformerStage(dim=c, depth=depth, num_heads=nh, window_size=7), h, w), (x,),
         f"batch {b}, {h}x{w} tokens, C {c}, {nh} heads, {depth} blocks, window 7, replicate padding")
    # ODA2 stage 0 (2 blocks, C = 192, 6 heads) at 88 x 304 tokens (padding to 91 x 308)
    h, w, c, nh, depth = 88, 304, 192, 6, 2
    x = torch.randn(b, h * w, c, device=DEV)
    case("oda2_swin_large_stage0", _Stage(siblings.Oda2SwinTransformerStage(dim=c, depth=depth, num_heads=nh, window_size=7), h, w),
         _Stage(oda2.SwinTransformerStage(dim=c, depth=depth, num_heads=nh, window_size=7), h, w), (x,),
         f"batch {b}, {h}x{w} tokens, C {c}, {nh} heads, {depth} blocks, window 7, replicate padding")
    # ODA / timm Swin-L at 384 x 384: stage 0, 96 x 96 tokens, window 12, shifted block (fixed attn_mask buffer)
    h = w = 96
    x = torch.randn(b, h * w, 192, device=DEV)
    case("oda_timm_swin_large_384_block_ws12_shift6", siblings.OdaSwinTransformerBlock(192, (h, w), 6, window_size=12, shift_size=6),
         oda.SwinTransformerBlock(192, (h, w), 6, window_size=12, shift_size=6), (x,),
         f"batch {b}, {h}x{w} tokens, C 192, 6 heads, window 12, shift 6")
    # Depthformer decoder at 352 x 1216: 11 x 38 = 418 tokens, hidden 256, 4 heads of 64
    vit, luna = rs.depthformer_vit_layer(), rs.depthformer_luna_layer()
    x = torch.randn(b, 418, 256, device=DEV)
    case("depthformer_vit_layer", depthformer.ViTLayer(256, 256, 4, num_repeat=1), vit.ViTLayer(256, 256, 4, num_repeat=1), (x,),
         f"batch {b}, 418 tokens, hidden 256, 4 heads of 64")
    xi, aux = torch.randn(b, 256, 11, 38, device=DEV), torch.randn(b, 32, 256, device=DEV)
    case("depthformer_prenorm_luna_layer", depthformer.PreNormLunaLayer(256, 256, 256, 4), luna.PreNormLunaLayer(256, 256, 256, 4), (xi, aux),
         f"batch {b}, 418 image tokens x 32 auxiliary tokens, hidden 256, 4 heads of 64")


if __name__ == "__main__":
    main()
