"""Import the reference repository (``/root/reference``) in the build container.

TEST INFRASTRUCTURE ONLY.  ``/root/reference`` does not exist on the GPU box,
so nothing that runs there may import this module; it is used by
``oracle/make_golden.py`` (fixture generation) and by CPU tests that skip when
the reference is absent.

The reference needs two packages that are not installed:
``timm`` (pin 0.5.4: ``DropPath``, ``to_2tuple``, ``trunc_normal_``) and
``mmcv`` (``ConvModule``).  None of their arithmetic is on the hot path; the
stand-ins below restate their documented behaviour.
"""
from __future__ import annotations

import os
import sys
import types

import torch
import torch.nn as nn

REFERENCE_ROOT = os.environ.get("MDE_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "model", "NewCRFs"))


class _DropPath(nn.Module):
    """Stochastic depth per sample (timm 0.5.4 ``DropPath``)."""

    def __init__(self, drop_prob=0.0):
        super().__init__()
        self.drop_prob = drop_prob

    def forward(self, x):
        if self.drop_prob == 0.0 or not self.training:
            return x
        keep = 1.0 - self.drop_prob
        shape = (x.shape[0],) + (1,) * (x.ndim - 1)
        gate = x.new_empty(shape).bernoulli_(keep)
        return x * gate / keep


def _to_2tuple(v):
    return tuple(v) if isinstance(v, (tuple, list)) else (v, v)


def _trunc_normal_(t, mean=0.0, std=1.0, a=-2.0, b=2.0):
    return nn.init.trunc_normal_(t, mean=mean, std=std, a=a, b=b)


class _ConvModule(nn.Module):
    """conv (bias only without norm) -> BN | GN -> ReLU, as ``mmcv.cnn.ConvModule``
    does for the configurations uper_crf_head.py uses (:35-45, :333-340)."""

    def __init__(self, in_ch, out_ch, kernel_size, padding=0, conv_cfg=None, norm_cfg=None,
                 act_cfg=dict(type="ReLU"), **kw):
        super().__init__()
        self.conv = nn.Conv2d(in_ch, out_ch, kernel_size, padding=padding, bias=norm_cfg is None)
        self.norm_name = None
        if norm_cfg is not None:
            if norm_cfg["type"] == "GN":
                self.norm_name = "gn"
                self.add_module("gn", nn.GroupNorm(norm_cfg["num_groups"], out_ch))
            else:
                self.norm_name = "bn"
                self.add_module("bn", nn.BatchNorm2d(out_ch))
        self.activate = nn.ReLU(inplace=True) if act_cfg is not None else None

    def forward(self, x):
        x = self.conv(x)
        if self.norm_name is not None:
            x = getattr(self, self.norm_name)(x)
        if self.activate is not None:
            x = self.activate(x)
        return x


def install() -> None:
    """Put the reference on sys.path with the timm/mmcv stand-ins registered."""
    if not available():
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT}")
    if "timm" not in sys.modules:
        timm = types.ModuleType("timm")
        models = types.ModuleType("timm.models")
        layers = types.ModuleType("timm.models.layers")
        layers.DropPath, layers.to_2tuple, layers.trunc_normal_ = _DropPath, _to_2tuple, _trunc_normal_
        timm.models, models.layers = models, layers
        sys.modules.update({"timm": timm, "timm.models": models, "timm.models.layers": layers})
    if "mmcv" not in sys.modules:
        mmcv = types.ModuleType("mmcv")
        cnn = types.ModuleType("mmcv.cnn")
        cnn.ConvModule = _ConvModule
        mmcv.cnn = cnn
        utils = types.ModuleType("mmcv.utils")
        mmcv.utils = utils
        sys.modules.update({"mmcv": mmcv, "mmcv.cnn": cnn, "mmcv.utils": utils})
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)


def newcrf_layers():
    install()
    import importlib
    return importlib.import_module("model.NewCRFs.newcrf_layers")


def swin_transformer():
    install()
    import importlib
    return importlib.import_module("model.NewCRFs.swin_transformer")


def newcrf_depth():
    install()
    import importlib
    return importlib.import_module("model.NewCRFs.NewCRFDepth")
