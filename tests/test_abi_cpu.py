"""The C-ABI library loads and exports every symbol include/swin_attn.h declares;
argument validation answers without touching a GPU.  CPU only (no compute calls)."""
import ctypes
import os
import re

import pytest

from mde_b200 import _lib
from conftest import ROOT


def _declared():
    text = open(os.path.join(ROOT, "include", "swin_attn.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b((?:swin_(?:attn|layernorm|add_layernorm|patch_merge_layernorm|gelu|linear)|mha_attention)_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = _declared()
    assert set(names) == set(_lib.EXPORTS), names
    for n in names:
        assert getattr(lib, n) is not None


def test_shape_struct_matches_header_layout():
    # 8 x int32, 2 x float, uint64, 4 x int64, 2 x int32
    assert ctypes.sizeof(_lib.Shape) == 8 * 4 + 2 * 4 + 8 + 4 * 8 + 2 * 4
    # mha_shape_t: 8 x int32, float, int32, 4 x int64
    assert ctypes.sizeof(_lib.MhaShape) == 8 * 4 + 4 + 4 + 4 * 8


@pytest.mark.parametrize("field,value,msg", [
    ("window", 8, "window must be 7 or 12"),
    ("shift", 7, "shift_size must in 0-window_size"),
    ("dtype", 5, "dtype must be"),
    ("num_heads", 0, "must be positive"),
    ("q_pix_stride", 16, "pixel strides must be >="),
    ("dropout_p", 1.5, "dropout_p must be in"),
])
def test_validation_errors_without_gpu(field, value, msg):
    lib = _lib.load()
    shp = _lib.Shape(1, 7, 7, 2, 7, 0, _lib.F32, 0, 0.1767, 0.0, 0, 64, 64, 64, 64, 0, 0)
    setattr(shp, field, value)
    rc = lib.swin_attn_forward(ctypes.byref(shp), *([None] * 11))
    assert rc != 0
    assert msg in lib.swin_attn_last_error().decode()
    with pytest.raises(RuntimeError, match=msg):
        _lib.check(rc)


def test_null_pointers_are_rejected():
    lib = _lib.load()
    shp = _lib.Shape(1, 7, 7, 2, 7, 0, _lib.F32, 0, 0.1767, 0.0, 0, 64, 64, 64, 64, 0, 0)
    assert lib.swin_attn_forward(ctypes.byref(shp), *([None] * 11)) != 0
    assert "non-null" in lib.swin_attn_last_error().decode()
    shp.height = 8
    assert lib.swin_attn_backward(ctypes.byref(shp), *([None] * 13)) != 0
    assert "multiples of the window" in lib.swin_attn_last_error().decode()
    assert lib.swin_attn_version().decode().startswith("swin_attn")


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libswin_attn.so")
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        _lib.load()


def test_ops_have_no_cpu_kernel():
    import torch
    from mde_b200 import ops
    with pytest.raises((RuntimeError, NotImplementedError)):
        ops.window_attention(torch.zeros(1, 7, 7, 96), None, torch.zeros(169, 1), 7, 0, 1)


def test_layernorm_validation_without_gpu():
    lib = _lib.load()
    buf = (ctypes.c_float * 256)()
    ptr = ctypes.cast(buf, ctypes.c_void_p)
    assert lib.swin_layernorm_forward(ptr, _lib.F32, ptr, ptr, 1e-5, 1, 100, ptr, _lib.BF16, None) != 0
    assert b"channels must be one of" in lib.swin_attn_last_error()
    assert lib.swin_layernorm_forward(None, _lib.F32, ptr, ptr, 1e-5, 1, 128, ptr, _lib.BF16, None) != 0
    assert lib.swin_layernorm_forward(ptr, _lib.BF16, ptr, ptr, 1e-5, 0, 128, ptr, _lib.F32, None) == 0   # empty input: no launch


def test_patch_merge_layernorm_validation_without_gpu():
    """swin_patch_merge_layernorm_forward (PatchMerging gather + norm, swin_transformer.py:271-287)."""
    lib = _lib.load()
    buf = (ctypes.c_float * 256)()
    ptr = ctypes.cast(buf, ctypes.c_void_p)
    err = lib.swin_attn_last_error
    assert lib.swin_patch_merge_layernorm_forward(ptr, _lib.F32, ptr, ptr, 1e-5, 1, 4, 4, 96, ptr, None) != 0
    assert b"channels must be one of" in err()
    assert lib.swin_patch_merge_layernorm_forward(ptr, 5, ptr, ptr, 1e-5, 1, 4, 4, 192, ptr, None) != 0 and b"x_dtype" in err()
    assert lib.swin_patch_merge_layernorm_forward(None, _lib.F32, ptr, ptr, 1e-5, 1, 4, 4, 192, ptr, None) != 0
    assert lib.swin_patch_merge_layernorm_forward(ptr, _lib.F32, ptr, ptr, 1e-5, 1, -1, 4, 192, ptr, None) != 0
    assert lib.swin_patch_merge_layernorm_forward(ptr, _lib.BF16, ptr, ptr, 1e-5, 0, 4, 4, 192, ptr, None) == 0   # empty: no launch


def test_block_plumbing_entries_validate_without_gpu():
    """swin_add_layernorm_forward / swin_gelu_* / swin_linear_bias_grad / swin_layernorm_backward reject bad arguments
    before touching the device and treat empty inputs as a no-op."""
    lib = _lib.load()
    buf = (ctypes.c_float * 256)()
    base = ctypes.addressof(buf)
    base += (-base) % 16
    ptr, odd = ctypes.c_void_p(base), ctypes.c_void_p(base + 2)
    err = lib.swin_attn_last_error
    assert lib.swin_add_layernorm_forward(ptr, _lib.F32, ptr, ptr, ptr, 1e-5, 1, 3072, ptr, ptr, None) != 0
    assert b"channels must be one of" in err()
    assert lib.swin_add_layernorm_forward(ptr, 7, ptr, ptr, ptr, 1e-5, 1, 128, ptr, ptr, None) != 0 and b"x_dtype" in err()
    assert lib.swin_add_layernorm_forward(ptr, _lib.F32, None, ptr, ptr, 1e-5, 1, 128, ptr, ptr, None) != 0
    assert lib.swin_add_layernorm_forward(ptr, _lib.BF16, ptr, ptr, ptr, 1e-5, 0, 128, ptr, ptr, None) == 0
    assert lib.swin_gelu_forward(ptr, ptr, 12, None) != 0 and b"multiple of 8" in err()
    assert lib.swin_gelu_forward(odd, ptr, 16, None) != 0 and b"16-byte aligned" in err()
    assert lib.swin_gelu_backward(ptr, None, ptr, 16, None) != 0
    assert lib.swin_gelu_forward(ptr, ptr, 0, None) == 0
    assert lib.swin_linear_bias_grad(ptr, 4, 12, ptr, None) != 0 and b"multiple of 8" in err()
    assert lib.swin_linear_bias_grad(None, 4, 16, ptr, None) != 0
    assert lib.swin_layernorm_backward(ptr, _lib.F32, ptr, 9, ptr, 1e-5, 1, 128, ptr, ptr, ptr, None) != 0 and b"dy_dtype" in err()
    assert lib.swin_layernorm_backward(ptr, _lib.F32, ptr, _lib.BF16, ptr, 1e-5, 1, 3072, ptr, ptr, ptr, None) != 0


def test_linear_forward_entry_validates_without_gpu():
    """swin_linear_forward (tcgen05 GEMM with bias / GELU epilogue): argument checks before any device work."""
    lib = _lib.load()
    buf = (ctypes.c_float * 256)()
    base = ctypes.addressof(buf)
    base += (-base) % 16
    ptr, odd = ctypes.c_void_p(base), ctypes.c_void_p(base + 2)
    err = lib.swin_attn_last_error
    assert lib.swin_linear_forward(ptr, ptr, ptr, ptr, 4, 192, 768, 2, None) != 0 and b"activation" in err()
    assert lib.swin_linear_forward(ptr, ptr, ptr, ptr, 4, 100, 768, 1, None) != 0 and b"in_features" in err()
    assert lib.swin_linear_forward(ptr, ptr, ptr, ptr, 4, 192, 576, 1, None) != 0 and b"out_features" in err()
    assert lib.swin_linear_forward(ptr, ptr, ptr, ptr, -1, 192, 768, 1, None) != 0 and b"row count" in err()
    assert lib.swin_linear_forward(ptr, None, ptr, ptr, 4, 192, 768, 1, None) != 0 and b"non-null" in err()
    assert lib.swin_linear_forward(ptr, ptr, odd, ptr, 4, 192, 768, 0, None) != 0 and b"16-byte aligned" in err()
    assert lib.swin_linear_forward(ptr, ptr, ptr, ptr, 0, 192, 768, 1, None) == 0          # empty batch: no-op
    # the residual and the training (h and GELU(h)) forms share the checks
    assert lib.swin_linear_residual_forward(ptr, ptr, ptr, None, ptr, 4, 192, 768, None) != 0 and b"non-null" in err()
    assert lib.swin_linear_residual_forward(ptr, ptr, ptr, odd, ptr, 4, 192, 768, None) != 0 and b"16-byte aligned" in err()
    assert lib.swin_linear_residual_forward(ptr, ptr, ptr, ptr, ptr, 4, 192, 192, None) != 0 and b"out_features" in err()
    assert lib.swin_linear_residual_forward(ptr, ptr, ptr, ptr, ptr, 0, 192, 768, None) == 0
    assert lib.swin_linear_gelu_forward_train(ptr, ptr, ptr, None, ptr, 4, 192, 768, None) != 0 and b"non-null" in err()
    assert lib.swin_linear_gelu_forward_train(ptr, ptr, ptr, ptr, ptr, 4, 100, 768, None) != 0 and b"in_features" in err()
    assert lib.swin_linear_gelu_forward_train(ptr, ptr, ptr, ptr, ptr, 0, 192, 768, None) == 0


def test_backward_unpadded_entry_validates_without_gpu():
    """swin_attn_backward_unpadded: pad-gradient outputs must mirror the pad inputs; ragged shapes are accepted by the
    argument checks of this entry (and only this one); fp32 cannot be served (CUDA-core kernels need explicit padding)."""
    lib = _lib.load()
    buf = (ctypes.c_float * 4096)()
    base = ctypes.addressof(buf)
    base += (-base) % 16
    ptr = ctypes.c_void_p(base)
    c = 64
    shp = _lib.Shape(1, 9, 11, 2, 7, 3, _lib.BF16, _lib.IMPL_AUTO, 32 ** -0.5, 0.0, 0, 3 * c, 3 * c, 3 * c, c, 0, 0)
    err = lib.swin_attn_last_error
    # k_pad given without d_k_pad
    assert lib.swin_attn_backward_unpadded(ctypes.byref(shp), ptr, ptr, ptr, ptr, None, ptr, ptr, ptr, ptr, ptr, ptr, ptr, ptr, None, None,
                                           None) != 0
    assert b"d_k_pad / d_v_pad must be given exactly when" in err()
    # the aligned-only entry rejects the same ragged shape
    assert lib.swin_attn_backward(ctypes.byref(shp), ptr, ptr, ptr, ptr, ptr, ptr, ptr, None, ptr, ptr, ptr, ptr, None) != 0
    assert b"multiples of the window" in err()
    # fp32 on ragged tensors: no tensor-core kernel, explicit padding required
    shp32 = _lib.Shape(1, 9, 11, 2, 7, 3, _lib.F32, _lib.IMPL_AUTO, 32 ** -0.5, 0.0, 0, 3 * c, 3 * c, 3 * c, c, 0, 0)
    assert lib.swin_attn_backward_unpadded(ctypes.byref(shp32), ptr, ptr, ptr, None, None, ptr, ptr, ptr, ptr, ptr, ptr, ptr, ptr, None,
                                           None, None) != 0
    assert b"needs the tcgen05 kernel" in err()
    assert lib.swin_attn_backward_unpadded(ctypes.byref(shp), ptr, ptr, None, None, None, ptr, ptr, ptr, ptr, ptr, ptr, ptr, ptr, None, None,
                                           None) != 0
    assert b"non-null" in err()


def test_torch_binding_registers_ops_without_a_gpu():
    """libswin_attn_torch.so (csrc/torch_binding.cpp) loads on a CPU-only host, registers the four dispatcher ops with
    the documented schemas, has Meta kernels for shape inference and no CPU kernels (CPU tensors raise)."""
    import torch
    from mde_b200 import ops
    assert os.path.exists(ops.TORCH_LIB_PATH)
    for name in ("fwd", "bwd", "layer_norm", "layer_norm_bwd", "add_layer_norm", "gelu", "gelu_bwd", "linear", "bias_grad",
                 "bwd_unpadded"):
        assert hasattr(torch.ops.swin_attn, name)
    s = str(torch.ops.swin_attn.fwd.default._schema)
    assert "Tensor? attn_mask=None, float dropout_p=0., int seed=0" in s and s.endswith("-> (Tensor, Tensor)")
    qkv = torch.empty(2, 14, 21, 192, device="meta", dtype=torch.bfloat16)
    o, lse = torch.ops.swin_attn.fwd(qkv, None, torch.empty(169, 2, device="meta"), None, None, 7, 3, 2, 0.17, True, 0)
    assert o.shape == (2, 14, 21, 64) and lse.shape == (2, 14, 21, 2) and lse.dtype == torch.float32
    y = torch.ops.swin_attn.layer_norm(torch.empty(5, 192, device="meta"), torch.empty(192, device="meta"),
                                       torch.empty(192, device="meta"), 1e-5, True)
    assert y.shape == (5, 192) and y.dtype == torch.bfloat16
    with pytest.raises(NotImplementedError):
        torch.ops.swin_attn.fwd(torch.zeros(1, 7, 7, 192), None, torch.zeros(169, 2), None, None, 7, 0, 2, 0.17, False, 0)
    assert ops._torch_lib.swin_torch_pending_timing() == 0
    # the block-plumbing / GEMM ops of the callers either side of the path: registered, shape inference on meta tensors
    for name in ("patch_merge_norm", "linear_act", "linear_residual", "linear_gelu", "linear_gelu_both"):
        assert hasattr(torch.ops.swin_attn, name)
    bf = torch.bfloat16
    x = torch.empty(2, 11 * 38, 192, device="meta", dtype=bf)            # odd H: padded to 12 x 38 -> 6 x 19 outputs of 4C
    y = torch.ops.swin_attn.patch_merge_norm(x, 11, 38, torch.empty(768, device="meta"), torch.empty(768, device="meta"), 1e-5)
    assert y.shape == (2, 6 * 19, 768) and y.dtype == bf
    t, w, b = torch.empty(3, 50, 768, device="meta", dtype=bf), torch.empty(256, 768, device="meta", dtype=bf), \
        torch.empty(256, device="meta", dtype=bf)
    assert torch.ops.swin_attn.linear_act(t, w, b, 1).shape == (3, 50, 256)
    assert torch.ops.swin_attn.linear_residual(t, w, b, torch.empty(3, 50, 256, device="meta", dtype=bf)).shape == (3, 50, 256)
    h, g = torch.ops.swin_attn.linear_gelu_both(t, w, b)
    assert h.shape == g.shape == (3, 50, 256) and h.dtype == bf
    with pytest.raises(NotImplementedError):
        torch.ops.swin_attn.patch_merge_norm(torch.zeros(1, 4, 192), 2, 2, torch.ones(768), torch.zeros(768), 1e-5)


def test_swin_checkpoint_loading_follows_the_reference_loader():
    """SwinTransformer.load_pretrained_state == newcrf_utils.load_checkpoint (:194-263): `module.` prefix, MoBY `encoder.`
    branch, bicubic resize of a window-7 relative_position_bias_table into a window-12 model."""
    import torch
    import torch.nn.functional as F
    from mde_b200.newcrfs import SwinTransformer
    kw = dict(embed_dim=96, depths=(2, 2), num_heads=(3, 6), out_indices=(0, 1))
    src = SwinTransformer(window_size=7, **kw)
    with torch.no_grad():
        for p in src.parameters():
            p.normal_(0, 0.5)
    dst = SwinTransformer(window_size=12, **kw)
    ckpt = {"model": {"module." + k: v for k, v in src.state_dict().items()}}
    missing, unexpected = dst.load_pretrained_state(ckpt)
    assert not unexpected and all(k.endswith("relative_position_index") for k in missing)
    key = "layers.1.blocks.0.attn.relative_position_bias_table"
    want = F.interpolate(src.state_dict()[key].permute(1, 0).reshape(1, 6, 13, 13), size=(23, 23), mode="bicubic")
    assert torch.equal(dst.state_dict()[key], want.reshape(6, 529).permute(1, 0))
    assert torch.equal(dst.state_dict()["layers.0.blocks.1.attn.qkv.weight"], src.state_dict()["layers.0.blocks.1.attn.qkv.weight"])
    assert dst.state_dict()["layers.0.blocks.0.attn.relative_position_index"].shape == (144, 144)
    moby = {"encoder." + k: v for k, v in src.state_dict().items()}
    moby.update({"encoder_k." + k: torch.zeros_like(v) for k, v in src.state_dict().items()})
    same = SwinTransformer(window_size=7, **kw)
    missing, unexpected = same.load_pretrained_state({"state_dict": moby})
    assert not missing and not unexpected
    assert all(torch.equal(a, b) for a, b in zip(same.state_dict().values(), src.state_dict().values()))


def test_impl_selectors_match_the_header_and_are_validated_without_gpu():
    """SWIN_ATTN_IMPL_*: the ctypes mirror carries the header's values; impl = WARP_MMA on a shape that has no warp-MMA kernel
    (window 7) is an error, not a silent downgrade; swin_attn_uses_tcgen05 says which family serves a shape (host logic)."""
    import re
    hdr = open(os.path.join(ROOT, "include", "swin_attn.h")).read()
    enum = dict((k, int(v)) for k, v in re.findall(r"SWIN_ATTN_(IMPL_[A-Z0-9_]+) = (\d+)", hdr))
    assert enum == {"IMPL_AUTO": _lib.IMPL_AUTO, "IMPL_SIMT": _lib.IMPL_SIMT, "IMPL_TCGEN05": _lib.IMPL_TCGEN05,
                    "IMPL_WARP_MMA": _lib.IMPL_WARP_MMA}
    lib = _lib.load()
    buf = (ctypes.c_float * 4096)()
    base = ctypes.addressof(buf)
    base += (-base) % 16
    ptr = ctypes.c_void_p(base)
    c = 64
    shp7 = _lib.Shape(1, 14, 14, 2, 7, 3, _lib.BF16, _lib.IMPL_WARP_MMA, 32 ** -0.5, 0.0, 0, 3 * c, 3 * c, 3 * c, c, 0, 0)
    assert lib.swin_attn_forward(ctypes.byref(shp7), ptr, ptr, ptr, None, None, None, ptr, None, ptr, None, None) != 0
    assert b"warp-MMA forward does not support" in lib.swin_attn_last_error()
    assert lib.swin_attn_backward(ctypes.byref(shp7), ptr, ptr, ptr, ptr, ptr, ptr, ptr, None, ptr, ptr, ptr, ptr, None) != 0
    assert b"warp-MMA backward does not support" in lib.swin_attn_last_error()

    def uses_tc(ws, shift, h, w, impl, backward):
        s = _lib.Shape(1, h, w, 2, ws, shift, _lib.BF16, impl, 32 ** -0.5, 0.0, 0, 3 * c, 3 * c, 3 * c, c, 0, 0)
        return lib.swin_attn_uses_tcgen05(ctypes.byref(s), backward)

    if lib.swin_attn_uses_tcgen05(ctypes.byref(_lib.Shape(1, 14, 14, 2, 7, 3, _lib.BF16, _lib.IMPL_AUTO, 32 ** -0.5, 0.0, 0, 3 * c,
                                                          3 * c, 3 * c, c, 0, 0)), 0) == 1:   # the driver API is present (TMA encode)
        assert uses_tc(7, 3, 15, 23, _lib.IMPL_AUTO, 0) == 1 and uses_tc(7, 0, 14, 14, _lib.IMPL_AUTO, 1) == 1
        assert uses_tc(12, 6, 24, 36, _lib.IMPL_TCGEN05, 0) == 1          # the tcgen05 window-12 forward, on request
    assert uses_tc(12, 6, 24, 36, _lib.IMPL_AUTO, 0) == 0                 # window 12: warp-MMA kernels by default
    assert uses_tc(12, 6, 13, 25, _lib.IMPL_AUTO, 0) == 0 and uses_tc(12, 0, 24, 36, _lib.IMPL_AUTO, 1) == 0
    assert uses_tc(7, 3, 14, 14, _lib.IMPL_SIMT, 0) == 0 and uses_tc(7, 3, 14, 14, _lib.IMPL_WARP_MMA, 0) == 0
