"""Import the UNMODIFIED reference modules: from ``oracle/_ref`` (the copy staged by
``oracle/stage_reference.py`` during ``__graft_entry__.build()``; it travels to the GPU box)
or, in the build container, straight from ``/root/reference``.

TEST INFRASTRUCTURE ONLY.  Used by ``oracle/make_golden.py`` (fixture generation), by the
reference arm / ``cpu_baseline`` / ``reference_gpu_eager`` legs of ``bench.py`` and by the
tests that compare the product with the reference's own code; every user checks
``available()`` first.  ``/root/reference`` itself is never read on the GPU box.

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

_STAGED = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def _find_root() -> str:
    env = os.environ.get("MDE_REFERENCE_ROOT")
    if env:
        return env
    for root in (_STAGED, "/root/reference"):
        if os.path.isfile(os.path.join(root, "model", "NewCRFs", "NewCRFDepth.py")):
            return root
    return _STAGED


REFERENCE_ROOT = _find_root()


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "model", "NewCRFs", "NewCRFDepth.py"))


def source() -> str:
    """Where the reference modules come from: "oracle/_ref" (staged copy) or the upstream path."""
    return "oracle/_ref" if os.path.abspath(REFERENCE_ROOT) == os.path.abspath(_STAGED) else REFERENCE_ROOT


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
        _install_timm_swin(models, layers)
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


def _install_timm_swin(models, layers) -> None:
    """What ``model/ODA/oda_swin_transformer.py:22-26`` imports from timm 0.5.4.  ``window_partition``, ``window_reverse``,
    ``WindowAttention`` and ``Mlp`` of timm's ``swin_transformer.py`` are the published Swin algorithm (Liu et al. 2021,
    microsoft/Swin-Transformer) that the reference ALSO carries in its own ``model/NewCRFs/swin_transformer.py:11-144``:
    the stand-ins are those reference classes (same constructor keywords, same parameter names), resolved lazily because
    the NewCRFs module imports ``timm.models.layers`` itself.  The other names are only touched by code paths no fixture
    runs (model builders, checkpoint filters) and are inert placeholders."""
    fx = types.ModuleType("timm.models.fx_features")
    fx.register_notrace_function = lambda f: f
    vit = types.ModuleType("timm.models.vision_transformer")
    vit.checkpoint_filter_fn = lambda state_dict, model: state_dict
    vit._init_vit_weights = lambda *a, **k: None
    swin = types.ModuleType("timm.models.swin_transformer")

    def _lazy(name):
        def get(*a, **k):
            import importlib
            return getattr(importlib.import_module("model.NewCRFs.swin_transformer"), name)(*a, **k)
        return get

    class _LazyClass:
        """Class stand-in whose construction is forwarded to the reference's NewCRFs Swin class of the same name."""

        def __init__(self, name):
            self._name = name

        def __call__(self, *a, **k):
            return _lazy(self._name)(*a, **k)

    swin.window_partition, swin.window_reverse = _lazy("window_partition"), _lazy("window_reverse")
    swin.WindowAttention, swin.PatchMerging = _LazyClass("WindowAttention"), _LazyClass("PatchMerging")
    layers.Mlp, layers.PatchEmbed = _LazyClass("Mlp"), _LazyClass("PatchEmbed")
    models.fx_features, models.vision_transformer, models.swin_transformer = fx, vit, swin
    sys.modules.update({"timm.models.fx_features": fx, "timm.models.vision_transformer": vit, "timm.models.swin_transformer": swin})


def oda_swin_transformer():
    install()
    import importlib
    return importlib.import_module("model.ODA.oda_swin_transformer")


def oda2_swin_transformer():
    install()
    import importlib
    return importlib.import_module("model.ODA2.oda2_swin_transformer")


def depthformer_vit_layer():
    install()
    import importlib
    return importlib.import_module("model.Depthformer.vit_layer")


def oda2_ordered_swin_decoder():
    install()
    import importlib
    return importlib.import_module("model.ODA2.oda2_red_order_swin2_decoder")


def depthformer_luna_layer():
    install()
    import importlib
    return importlib.import_module("model.Depthformer.luna_layer")


def depthformer_self_attention():
    install()
    import importlib
    return importlib.import_module("model.Depthformer.self_attention")


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


def module(name: str):
    """Any other reference module by dotted name, e.g. ``model.ODA2.oda2_swin_transformer``,
    ``model.Depthformer.self_attention``, ``utils.depth_utils``."""
    install()
    import importlib
    return importlib.import_module(name)
