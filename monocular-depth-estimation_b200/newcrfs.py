"""NewCRFs depth network assembled from the drop-in layers.

The callers of the hot path (model/NewCRFs/NewCRFDepth.py:11-164,
swin_transformer.py:439-620, uper_crf_head.py:9-58,318-364 of the reference)
restated so the headline benchmark -- full-model images/s -- can be run
without the reference tree (``/root/reference`` does not exist on the GPU box,
and its ``timm`` / ``mmcv`` imports are not installed anywhere).  Module names
and ``state_dict`` keys equal the reference's, so its checkpoints load.
Only the window attention is accelerated; convs, Linears, norms and pooling
are stock PyTorch (cuDNN / cuBLAS), as in the reference.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops
from .layers import BasicLayer, NewCRF, PatchEmbed, PatchMerging, draw_drop_path_gates

_VARIANTS = {
    "tiny": dict(embed_dim=96, depths=[2, 2, 6, 2], num_heads=[3, 6, 12, 24]),
    "base": dict(embed_dim=128, depths=[2, 2, 18, 2], num_heads=[4, 8, 16, 32]),
    "large": dict(embed_dim=192, depths=[2, 2, 18, 2], num_heads=[6, 12, 24, 48]),
}


class SwinTransformer(nn.Module):
    """Swin backbone returning the four stage outputs as NCHW (swin_transformer.py:439-620)."""

    def __init__(self, pretrain_img_size=224, patch_size=4, in_chans=3, embed_dim=96, depths=(2, 2, 6, 2),
                 num_heads=(3, 6, 12, 24), window_size=7, mlp_ratio=4., qkv_bias=True, qk_scale=None,
                 drop_rate=0., attn_drop_rate=0., drop_path_rate=0.2, norm_layer=nn.LayerNorm, ape=False,
                 patch_norm=True, out_indices=(0, 1, 2, 3), frozen_stages=-1, use_checkpoint=False):
        super().__init__()
        if ape:
            raise NotImplementedError("absolute position embedding is unused by NewCRFs (ape=False)")
        self.pretrain_img_size = pretrain_img_size
        self.num_layers = len(depths)
        self.embed_dim, self.ape, self.patch_norm = embed_dim, ape, patch_norm
        self.out_indices, self.frozen_stages = out_indices, frozen_stages
        self.patch_embed = PatchEmbed(patch_size=patch_size, in_chans=in_chans, embed_dim=embed_dim,
                                      norm_layer=norm_layer if patch_norm else None)
        self.pos_drop = nn.Dropout(p=drop_rate)
        total = sum(depths)
        dpr = [drop_path_rate * i / max(total - 1, 1) for i in range(total)]
        self.layers = nn.ModuleList()
        for i in range(self.num_layers):
            self.layers.append(BasicLayer(
                dim=int(embed_dim * 2 ** i), depth=depths[i], num_heads=num_heads[i], window_size=window_size,
                mlp_ratio=mlp_ratio, qkv_bias=qkv_bias, qk_scale=qk_scale, drop=drop_rate, attn_drop=attn_drop_rate,
                drop_path=dpr[sum(depths[:i]):sum(depths[:i + 1])], norm_layer=norm_layer,
                downsample=PatchMerging if i < self.num_layers - 1 else None, use_checkpoint=use_checkpoint))
        self.num_features = [int(embed_dim * 2 ** i) for i in range(self.num_layers)]
        for i in out_indices:
            self.add_module(f"norm{i}", norm_layer(self.num_features[i]))
        self._freeze_stages()

    def _freeze_stages(self):
        if self.frozen_stages >= 0:
            self.patch_embed.eval()
            for p in self.patch_embed.parameters():
                p.requires_grad = False
        if self.frozen_stages >= 2:
            self.pos_drop.eval()
            for i in range(self.frozen_stages - 1):
                self.layers[i].eval()
                for p in self.layers[i].parameters():
                    p.requires_grad = False

    def init_weights(self, pretrained=None):
        def init(m):
            if isinstance(m, nn.Linear):
                nn.init.trunc_normal_(m.weight, std=.02)
                if m.bias is not None:
                    nn.init.zeros_(m.bias)
            elif isinstance(m, nn.LayerNorm):
                nn.init.zeros_(m.bias)
                nn.init.ones_(m.weight)
        self.apply(init)
        if isinstance(pretrained, str):
            self.load_pretrained_state(torch.load(pretrained, map_location="cpu"))
        elif pretrained is not None:
            raise TypeError("pretrained must be a str or None")

    def load_pretrained_state(self, checkpoint):
        """What the reference's ``load_checkpoint(self, pretrained, strict=False)`` does with a checkpoint dict
        (newcrf_utils.py:194-263): unwrap ``state_dict`` / ``model``, strip a ``module.`` prefix, keep the online
        (``encoder.``) branch of a MoBY checkpoint, resize every ``relative_position_bias_table`` whose window size
        differs from this model's by bicubic interpolation (a window-7 checkpoint into a window-12 model; tables whose
        head count differs are skipped, as there), then load non-strictly.  Returns the (missing, unexpected) keys."""
        if not isinstance(checkpoint, dict):
            raise RuntimeError("No state_dict found in checkpoint")
        state = checkpoint.get("state_dict", checkpoint.get("model", checkpoint))
        if state and next(iter(state)).startswith("module."):
            state = {k[7:]: v for k, v in state.items()}
        if state and sorted(state)[0].startswith("encoder"):
            state = {k.replace("encoder.", ""): v for k, v in state.items() if k.startswith("encoder.")}
        state = dict(state)
        state.pop("absolute_pos_embed", None)            # ape=False in every NewCRFs configuration
        own = self.state_dict()
        for key in [k for k in state if "relative_position_bias_table" in k]:
            if key not in own:
                continue
            (l1, nh1), (l2, nh2) = state[key].shape, own[key].shape
            if nh1 != nh2:
                del state[key]                           # the reference logs "Error in loading ..., pass"
            elif l1 != l2:
                s1, s2 = int(l1 ** 0.5), int(l2 ** 0.5)
                resized = F.interpolate(state[key].float().permute(1, 0).reshape(1, nh1, s1, s1), size=(s2, s2), mode="bicubic")
                state[key] = resized.reshape(nh2, l2).permute(1, 0)
        # relative_position_index depends on the window size only; a checkpoint of another window size carries a stale one
        for key in [k for k in state if k.endswith("relative_position_index") and k in own and state[k].shape != own[k].shape]:
            del state[key]
        return self.load_state_dict(state, strict=False)

    def forward(self, x):
        x = self.patch_embed(x)
        wh, ww = x.shape[2], x.shape[3]
        x = self.pos_drop(x.flatten(2).transpose(1, 2))
        outs = []
        for i, layer in enumerate(self.layers):
            x_out, h, w, x, wh, ww = layer(x, wh, ww)
            if i in self.out_indices:
                x_out = ops.norm_fp32(getattr(self, f"norm{i}"), x_out)
                # (B, C, H, W) in channels-last strides: the values and shape of the reference's contiguous NCHW
                # copy (swin_transformer.py:607-608) without the copy; the decoder convs run NHWC in cuDNN anyway
                outs.append(x_out.view(-1, h, w, self.num_features[i]).permute(0, 3, 1, 2))
        return tuple(outs)

    def train(self, mode=True):
        super().train(mode)
        self._freeze_stages()


class ConvModule(nn.Module):
    """conv -> BN|GN -> ReLU with mmcv's child names (``conv``, ``bn`` / ``gn``, ``activate``)."""

    def __init__(self, in_ch, out_ch, kernel_size, padding=0, norm="bn", num_groups=None):
        super().__init__()
        self.conv = nn.Conv2d(in_ch, out_ch, kernel_size, padding=padding, bias=False)
        self.norm_name = norm
        self.add_module(norm, nn.GroupNorm(num_groups, out_ch) if norm == "gn" else nn.BatchNorm2d(out_ch))
        self.activate = nn.ReLU(inplace=True)

    def forward(self, x):
        return self.activate(getattr(self, self.norm_name)(self.conv(x)))


_RESIZE_CACHE = {}


def _bilinear_resize_small(x, size):
    """``F.interpolate(x, size, mode="bilinear", align_corners=False)`` for a tiny source grid (the pooled PSP
    branches, 1x1 .. 6x6) as ONE fp32 matmul with the interpolation matrix: ATen's NCHW bilinear kernel
    parallelises over the output pixels only and loops over N*C inside each thread (0.8 ms per branch at
    batch 8).  The matrix is the response of the same ATen op to the identity, so the weights are identical."""
    n, c, h, w = x.shape
    key = (h, w, size[0], size[1], x.device)
    k = _RESIZE_CACHE.get(key)
    if k is None:
        eye = torch.eye(h * w, device=x.device, dtype=torch.float32).view(1, h * w, h, w)
        k = F.interpolate(eye, size=size, mode="bilinear", align_corners=False).reshape(h * w, size[0] * size[1])
        _RESIZE_CACHE[key] = k
    with torch.autocast(x.device.type, enabled=False):
        return torch.matmul(x.float().reshape(n * c, h * w), k).view(n, c, size[0], size[1]).to(x.dtype)


def _adaptive_avg_pool_small(x, out_size):
    """``F.adaptive_avg_pool2d(x, out_size)`` of the last encoder feature (11x38 .. 15x20 pixels) as one fp32 matmul
    with the pooling matrix (ATen's NHWC kernel takes ~100 us per pool scale at batch 8 for 0.4 MFLOP of work).  The
    matrix is the response of the same ATen op to the identity, so the bin boundaries and weights are identical."""
    n, c, h, w = x.shape
    s = out_size if isinstance(out_size, (tuple, list)) else (out_size, out_size)
    key = ("pool", h, w, s[0], s[1], x.device)
    k = _RESIZE_CACHE.get(key)
    if k is None:
        eye = torch.eye(h * w, device=x.device, dtype=torch.float32).view(1, h * w, h, w)
        k = F.adaptive_avg_pool2d(eye, s).reshape(h * w, s[0] * s[1])
        _RESIZE_CACHE[key] = k
    with torch.autocast(x.device.type, enabled=False):
        # (n, hw, c) is a free view of a channels-last feature; one batched (s*s, hw) x (hw, c) product
        tokens = x.permute(0, 2, 3, 1).reshape(n, h * w, c).float()
        pooled = torch.matmul(k.t(), tokens)                       # (n, s*s, c)
        # dense NCHW: the 1x1 conv behind it then returns a dense weight gradient (DDP bucket views want that)
        return pooled.transpose(1, 2).reshape(n, c, s[0], s[1]).to(x.dtype).contiguous()


class PSP(nn.Module):
    """Pyramid pooling over the last encoder feature (uper_crf_head.py:9-58, 318-364).

    The reference rebinds ``norm_cfg`` to GroupNorm(256) at the first pool scale
    (uper_crf_head.py:35) and never restores it, so all four pooled branches
    use GroupNorm; only the bottleneck keeps BatchNorm.
    """

    def __init__(self, pool_scales=(1, 2, 3, 6), in_channels=(96, 192, 384, 768), channels=512, **_unused):
        super().__init__()
        self.in_channels, self.channels, self.align_corners = list(in_channels), channels, False
        self.psp_modules = nn.ModuleList([
            nn.Sequential(nn.AdaptiveAvgPool2d(s),
                          ConvModule(self.in_channels[-1], channels, 1,
                                     norm="gn", num_groups=256))
            for s in pool_scales])
        self.bottleneck = ConvModule(self.in_channels[-1] + len(pool_scales) * channels, channels, 3, padding=1)

    def init_weights(self):
        pass

    def forward(self, inputs):
        x = inputs[-1]
        outs = [x]
        for m in self.psp_modules:   # Sequential(AdaptiveAvgPool2d(s), ConvModule): same modules, the pool as a matmul
            pooled = _adaptive_avg_pool_small(x, m[0].output_size) if x.is_cuda else m[0](x)
            outs.append(_bilinear_resize_small(m[1](pooled), x.shape[2:]))
        return self.bottleneck(torch.cat(outs, dim=1))


class DispHead(nn.Module):
    """3x3 conv -> sigmoid -> bilinear upsample (NewCRFDepth.py:151-164)."""

    def __init__(self, input_dim=100):
        super().__init__()
        self.conv1 = nn.Conv2d(input_dim, 1, 3, padding=1)
        self.sigmoid = nn.Sigmoid()

    def forward(self, x, scale):
        x = self.sigmoid(self.conv1(x))
        if scale > 1:
            x = F.interpolate(x, scale_factor=scale, mode="bilinear", align_corners=False)
        return x


class NewCRFDepth(nn.Module):
    """Swin encoder -> PSP -> 4 NeWCRF levels with PixelShuffle -> DispHead (NewCRFDepth.py:11-148)."""

    def __init__(self, version=None, inv_depth=False, pretrained=None, frozen_stages=-1, min_depth=0.1,
                 max_depth=100.0, **kwargs):
        super().__init__()
        self.inv_depth = inv_depth
        self.with_auxiliary_head = False
        self.with_neck = False
        cfg = _VARIANTS[version[:-2]]
        window_size = int(version[-2:])
        in_channels = [cfg["embed_dim"] * 2 ** i for i in range(4)]
        self.backbone = SwinTransformer(embed_dim=cfg["embed_dim"], depths=cfg["depths"], num_heads=cfg["num_heads"],
                                        window_size=window_size, ape=False, drop_path_rate=0.3, patch_norm=True,
                                        use_checkpoint=False, frozen_stages=frozen_stages)
        embed_dim = 512
        crf_dims = [128, 256, 512, 1024]
        v_dims = [64, 128, 256, embed_dim]
        heads = [4, 8, 16, 32]
        for i in (3, 2, 1, 0):
            setattr(self, f"crf{i}", NewCRF(input_dim=in_channels[i], embed_dim=crf_dims[i], window_size=7,
                                            v_dim=v_dims[i], num_heads=heads[i]))
        self.decoder = PSP(pool_scales=(1, 2, 3, 6), in_channels=in_channels, channels=embed_dim)
        self.disp_head1 = DispHead(input_dim=crf_dims[0])
        self.up_mode = "bilinear"
        self.min_depth, self.max_depth = min_depth, max_depth
        self.init_weights(pretrained=pretrained)

    def init_weights(self, pretrained=None):
        self.backbone.init_weights(pretrained=pretrained)
        self.decoder.init_weights()

    def forward(self, imgs):
        if self.training and torch.is_grad_enabled():   # every stochastic-depth gate of this forward in one draw
            bf16 = torch.is_autocast_enabled() and torch.get_autocast_dtype(imgs.device.type) == torch.bfloat16
            draw_drop_path_gates(self, imgs.shape[0], imgs.device, torch.bfloat16 if bf16 else imgs.dtype)
        feats = self.backbone(imgs)
        e = self.decoder(feats)
        e = F.pixel_shuffle(self.crf3(feats[3], e), 2)
        e = F.pixel_shuffle(self.crf2(feats[2], e), 2)
        e = F.pixel_shuffle(self.crf1(feats[1], e), 2)
        e = self.crf0(feats[0], e)
        return self.disp_head1(e, 4) * self.max_depth
