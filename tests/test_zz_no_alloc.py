"""swin_attn_forward_host_static: the host-buffer forward that allocates nothing (static __device__ staging array).

Collected LAST on purpose (``zz``): a failure here must not hide other tests.
Calls the C ABI directly through ctypes (``_lib``), compares with the existing entry point ``swin_attn_forward_host``
(which stages through a cudaMalloc'd buffer) on two calls in a row, and checks the size limit.  Tolerances are the
suite's: max-normalised <= 1e-5 fp32, <= 1e-2 bf16 (tests/test_parity_gpu.py); the two entry points run the same
kernel on the same bytes, so they are also expected to agree exactly.
"""
import ctypes

import pytest
import torch

from mde_b200 import _lib
from oracle import window_attention as wa
from oracle.synthetic import randn
from conftest import check, rel_err

TOL = {torch.float32: 1e-5, torch.bfloat16: 1e-2}


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _call(entry, q, k, v, table, ws, shift, nh, pads):
    """One call of a host-buffer entry point of the C ABI; returns the output tensor (CPU)."""
    lib = _lib.load()
    b, h, w, c = q.shape
    dtype = _lib.F32 if q.dtype == torch.float32 else _lib.BF16
    shp = _lib.Shape(b, h, w, nh, ws, shift, dtype, _lib.IMPL_AUTO, 32 ** -0.5, 0.0, 0, c, c, c, c, 0, 0)
    o = torch.full_like(q, float("nan"))
    rc = getattr(lib, entry)(ctypes.byref(shp), _ptr(q), _ptr(k), _ptr(v), _ptr(pads[0]), _ptr(pads[1]), _ptr(pads[2]),
                             _ptr(table), _ptr(o))
    assert rc == 0, lib.swin_attn_last_error().decode()
    return o


def _inputs(name, b, h, w, nh, ws, dtype):
    c = nh * 32
    q, k, v = (randn(f"{name}.{n}", b, h, w, c).to(dtype).contiguous() for n in "qkv")
    table = (0.5 * randn(name + ".t", (2 * ws - 1) ** 2, nh)).contiguous()
    pads = [(0.3 * randn(f"{name}.pad{n}", c)).to(dtype).contiguous() for n in "qkv"]
    return q, k, v, table, pads


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_static_scratch_entry_matches_the_allocating_entry_on_two_calls_in_a_row(dtype):
    """Two calls in a row with DIFFERENT shapes and data (the second is smaller, so stale bytes of the first call lie
    behind and inside its staging regions): each must equal swin_attn_forward_host on the same inputs, and the oracle."""
    cases = [("noalloc.a", 2, 16, 23, 4, 7, 3), ("noalloc.b", 1, 9, 11, 2, 7, 0)]
    for name, b, h, w, nh, ws, shift in cases:
        q, k, v, table, pads = _inputs(name, b, h, w, nh, ws, dtype)
        got = _call("swin_attn_forward_host_static", q, k, v, table, ws, shift, nh, pads)
        want = _call("swin_attn_forward_host", q, k, v, table, ws, shift, nh, pads)
        assert got.device.type == "cpu" and got.shape == q.shape and torch.isfinite(got.float()).all()
        check(rel_err(got.float(), want.float()), TOL[dtype], f"static-scratch entry vs allocating entry {name} {dtype}")
        assert torch.equal(got, want)   # same kernel, same bytes
        ref, _ = wa.window_attention(q.double(), k.double(), v.double(), table.double(), ws, shift, nh, None,
                                     *[p.double() for p in pads])
        check(rel_err(got.float(), ref), TOL[dtype], f"static-scratch entry vs oracle {name} {dtype}")
    # and once more with the first case: the result does not depend on what the array held before
    name, b, h, w, nh, ws, shift = cases[0]
    q, k, v, table, pads = _inputs(name, b, h, w, nh, ws, dtype)
    again = _call("swin_attn_forward_host_static", q, k, v, table, ws, shift, nh, pads)
    assert torch.equal(again, _call("swin_attn_forward_host", q, k, v, table, ws, shift, nh, pads))


@pytest.mark.gpu
def test_static_scratch_entry_without_pad_vectors_and_with_ws12():
    """NULL pad vectors (zero padding) and the window-12 kernel through the same static staging."""
    q, k, v, table, _ = _inputs("noalloc.c", 1, 24, 36, 3, 12, torch.bfloat16)
    none = [None, None, None]
    got = _call("swin_attn_forward_host_static", q, k, v, table, 12, 6, 3, none)
    want = _call("swin_attn_forward_host", q, k, v, table, 12, 6, 3, none)
    check(rel_err(got.float(), want.float()), TOL[torch.bfloat16], "static-scratch entry ws12 vs allocating entry")


def test_static_scratch_entry_rejects_sizes_beyond_the_limit():
    """4 * B*H*W*C*elem beyond SWIN_ATTN_STATIC_SCRATCH_BYTES (8 MiB) -> a clear error before any CUDA call (no GPU
    needed: runs in the CPU suite too); a shape just inside the limit is not rejected for its size."""
    lib = _lib.load()
    assert _lib.STATIC_SCRATCH_BYTES == 8 << 20
    nh, c = 4, 128
    b, h, w = 8, 88, 304                      # 8*88*304*128 bf16 = 54.8 MB per tensor
    shp = _lib.Shape(b, h, w, nh, 7, 0, _lib.BF16, _lib.IMPL_AUTO, 32 ** -0.5, 0.0, 0, c, c, c, c, 0, 0)
    dummy = ctypes.c_void_p(ctypes.addressof(ctypes.create_string_buffer(64)))
    rc = lib.swin_attn_forward_host_static(ctypes.byref(shp), dummy, dummy, dummy, None, None, None, dummy, dummy)
    msg = lib.swin_attn_last_error().decode()
    assert rc != 0 and "SWIN_ATTN_STATIC_SCRATCH_BYTES" in msg and str(8 << 20) in msg, msg
    # the other argument checks come first, as in swin_attn_forward_host
    rc = lib.swin_attn_forward_host_static(ctypes.byref(shp), None, dummy, dummy, None, None, None, dummy, dummy)
    assert rc != 0 and "non-null" in lib.swin_attn_last_error().decode()
