"""B200-native (sm_100a) shifted-window attention for NewCRFs.

Drop-in replacements for the reference's ``WindowAttention`` / ``CRFBlock`` /
``BasicCRFLayer`` / ``NewCRF`` / ``SwinTransformerBlock`` / ``BasicLayer``
modules (model/NewCRFs/{newcrf_layers,swin_transformer}.py of
pitlover/Monocular-Depth-Estimation) whose attention path runs in the
hand-written CUDA kernels of ``libswin_attn.so``.
"""
from . import _lib  # noqa: F401
from .ops import window_attention, window_attention_host  # noqa: F401
from .layers import (  # noqa: F401
    BasicCRFLayer, BasicLayer, CRFBlock, Mlp, NewCRF, PatchEmbed, PatchMerging,
    SwinTransformerBlock, WindowAttention, EncoderWindowAttention,
)
from .newcrfs import NewCRFDepth, SwinTransformer, PSP, DispHead  # noqa: F401
from . import depthformer, siblings  # noqa: F401  (Depthformer attention, ODA / ODA2 Swin blocks)
