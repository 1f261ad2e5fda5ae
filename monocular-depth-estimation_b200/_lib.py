"""ctypes binding of libswin_attn.so (C ABI in include/swin_attn.h).

There is no fallback: if the library has not been built (``python -c "import
__graft_entry__ as g; g.build()"`` or ``csrc/build.sh``) importing the ops
raises, and every entry point raises ``RuntimeError`` with the library's own
message on a non-zero return code.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_float, c_int, c_int32, c_int64, c_uint64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libswin_attn.so")

F32, BF16 = 0, 1
STATIC_SCRATCH_BYTES = 8 << 20   # SWIN_ATTN_STATIC_SCRATCH_BYTES (include/swin_attn.h)
IMPL_AUTO, IMPL_SIMT, IMPL_TCGEN05, IMPL_WARP_MMA = 0, 1, 2, 3

EXPORTS = (
    "swin_attn_forward",
    "swin_attn_backward",
    "swin_attn_backward_unpadded",
    "swin_attn_forward_host",
    "swin_attn_forward_host_static",
    "swin_layernorm_forward",
    "swin_layernorm_backward",
    "swin_patch_merge_layernorm_forward",
    "swin_add_layernorm_forward",
    "swin_add_layernorm_forward_gated",
    "swin_add_layernorm_backward",
    "swin_gelu_forward",
    "swin_gelu_backward",
    "swin_linear_forward",
    "swin_linear_residual_forward",
    "swin_linear_gelu_forward_train",
    "swin_linear_bias_grad",
    "mha_attention_forward",
    "mha_attention_forward_ordered",
    "swin_attn_time_next_launch",
    "swin_attn_probe_empty_bracket_us",
    "swin_attn_launch_count",
    "swin_attn_uses_tcgen05",
    "swin_attn_last_error",
    "swin_attn_version",
)


class Shape(Structure):
    """Mirror of ``swin_attn_shape_t``."""

    _fields_ = [
        ("batch", c_int32), ("height", c_int32), ("width", c_int32), ("num_heads", c_int32),
        ("window", c_int32), ("shift", c_int32), ("dtype", c_int32), ("impl", c_int32),
        ("scale", c_float), ("dropout_p", c_float), ("seed", c_uint64),
        ("q_pix_stride", c_int64), ("k_pix_stride", c_int64), ("v_pix_stride", c_int64), ("o_pix_stride", c_int64),
        ("mask_windows", c_int32), ("reserved", c_int32),
    ]


class MhaShape(Structure):
    """Mirror of ``mha_shape_t``."""

    _fields_ = [
        ("batch", c_int32), ("seq", c_int32), ("kv_seq", c_int32), ("num_heads", c_int32), ("qk_head_dim", c_int32),
        ("v_head_dim", c_int32), ("dtype", c_int32), ("impl", c_int32), ("scale", c_float), ("reserved", c_int32),
        ("q_row_stride", c_int64), ("k_row_stride", c_int64), ("v_row_stride", c_int64), ("o_row_stride", c_int64),
    ]


_lib = None


def load() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: the CUDA extension has not been built (run __graft_entry__.build()); "
            "this package has no CPU or PyTorch fallback")
    lib = ctypes.CDLL(LIB_PATH)
    vp = c_void_p
    lib.swin_attn_forward.argtypes = [POINTER(Shape)] + [vp] * 11
    lib.swin_attn_forward.restype = c_int
    lib.swin_attn_backward.argtypes = [POINTER(Shape)] + [vp] * 13
    lib.swin_attn_backward.restype = c_int
    lib.swin_attn_backward_unpadded.argtypes = [POINTER(Shape)] + [vp] * 16
    lib.swin_attn_backward_unpadded.restype = c_int
    lib.swin_attn_forward_host.argtypes = [POINTER(Shape)] + [vp] * 8
    lib.swin_attn_forward_host.restype = c_int
    lib.swin_attn_forward_host_static.argtypes = [POINTER(Shape)] + [vp] * 8
    lib.swin_attn_forward_host_static.restype = c_int
    lib.mha_attention_forward.argtypes = [POINTER(MhaShape)] + [vp] * 6
    lib.mha_attention_forward.restype = c_int
    lib.mha_attention_forward_ordered.argtypes = [POINTER(MhaShape), vp, vp, vp, vp, vp, c_int32, vp, vp, vp]
    lib.mha_attention_forward_ordered.restype = c_int
    lib.swin_layernorm_forward.argtypes = [vp, c_int32, vp, vp, c_float, c_int64, c_int32, vp, c_int32, vp]
    lib.swin_layernorm_forward.restype = c_int
    lib.swin_layernorm_backward.argtypes = [vp, c_int32, vp, c_int32, vp, c_float, c_int64, c_int32, vp, vp, vp, vp]
    lib.swin_layernorm_backward.restype = c_int
    lib.swin_add_layernorm_forward.argtypes = [vp, c_int32, vp, vp, vp, c_float, c_int64, c_int32, vp, vp, vp]
    lib.swin_add_layernorm_forward.restype = c_int
    lib.swin_add_layernorm_forward_gated.argtypes = [vp, c_int32, vp, vp, vp, vp, c_int64, c_float, c_int64, c_int32, vp, vp, vp]
    lib.swin_add_layernorm_forward_gated.restype = c_int
    lib.swin_add_layernorm_backward.argtypes = [vp, c_int32, vp, vp, vp, vp, c_int64, c_float, c_int64, c_int32, vp, vp, vp, vp, vp]
    lib.swin_add_layernorm_backward.restype = c_int
    lib.swin_gelu_forward.argtypes = [vp, vp, c_int64, vp]
    lib.swin_gelu_forward.restype = c_int
    lib.swin_gelu_backward.argtypes = [vp, vp, vp, c_int64, vp]
    lib.swin_gelu_backward.restype = c_int
    lib.swin_linear_forward.argtypes = [vp, vp, vp, vp, c_int64, c_int32, c_int32, c_int32, vp]
    lib.swin_linear_forward.restype = c_int
    lib.swin_linear_bias_grad.argtypes = [vp, c_int64, c_int32, vp, vp]
    lib.swin_linear_bias_grad.restype = c_int
    lib.swin_linear_gelu_forward_train.argtypes = [vp, vp, vp, vp, vp, c_int64, c_int32, c_int32, vp]
    lib.swin_linear_gelu_forward_train.restype = c_int
    lib.swin_linear_residual_forward.argtypes = [vp, vp, vp, vp, vp, c_int64, c_int32, c_int32, vp]
    lib.swin_linear_residual_forward.restype = c_int
    lib.swin_patch_merge_layernorm_forward.argtypes = [vp, c_int32, vp, vp, c_float, c_int32, c_int32, c_int32, c_int32, vp, vp]
    lib.swin_patch_merge_layernorm_forward.restype = c_int
    lib.swin_attn_time_next_launch.argtypes = [vp, vp]
    lib.swin_attn_time_next_launch.restype = None
    lib.swin_attn_probe_empty_bracket_us.argtypes = [vp, c_int32, c_int32]
    lib.swin_attn_probe_empty_bracket_us.restype = ctypes.c_double
    lib.swin_attn_launch_count.argtypes = []
    lib.swin_attn_launch_count.restype = c_uint64
    lib.swin_attn_uses_tcgen05.argtypes = [POINTER(Shape), c_int]
    lib.swin_attn_uses_tcgen05.restype = c_int
    lib.swin_attn_last_error.restype = c_char_p
    lib.swin_attn_version.restype = c_char_p
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        raise RuntimeError("swin_attn: " + load().swin_attn_last_error().decode())


def launch_count() -> int:
    return int(load().swin_attn_launch_count())
