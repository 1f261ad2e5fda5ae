"""Out-of-bounds canaries for the kernels behind the C ABI (compute-sanitizer is not available on the GPU pool).

Every output of a call (o, lse, dq, dk, dv, d_table) lives INSIDE a larger caller-owned allocation whose bytes before and
after it hold a sentinel; every input sits at the very end of its allocation's used part, followed by NaN bytes.  After the
call the guard bands must be bit-identical to the sentinel (no write outside an output) and the outputs must be finite (no
read of the NaN bytes behind an input reached a result) and equal to the same call on plain tensors."""
import ctypes

import pytest
import torch

import mde_b200  # noqa: F401
from mde_b200 import _lib
from oracle.synthetic import randn

pytestmark = pytest.mark.gpu
DEV = "cuda"
GUARD = 4096          # elements on either side (a multiple of 16 bytes for every dtype used)
SENTINEL = -12345.0


def _guarded_out(shape, dtype):
    n = 1
    for s in shape:
        n *= s
    full = torch.full((n + 2 * GUARD,), SENTINEL, dtype=dtype, device=DEV)
    return full[GUARD:GUARD + n].view(shape), full


def _guarded_in(t):
    """t on the device, followed by NaNs (and preceded by NaNs)."""
    full = torch.full((t.numel() + 2 * GUARD,), float("nan"), dtype=t.dtype, device=DEV)
    view = full[GUARD:GUARD + t.numel()].view(t.shape)
    view.copy_(t)
    return view, full


def _bands_intact(full, n):
    ref = torch.full((GUARD,), SENTINEL, dtype=full.dtype, device=full.device)
    return torch.equal(full[:GUARD], ref) and torch.equal(full[GUARD + n:], ref)


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


# ws, shift, B, H, W, nH, impl, with_pads: every tensor-core forward family, aligned and not
FWD = [
    (7, 3, 2, 15, 23, 3, _lib.IMPL_AUTO, True), (7, 0, 1, 14, 21, 2, _lib.IMPL_AUTO, False),
    (12, 6, 1, 13, 25, 2, _lib.IMPL_AUTO, True), (12, 0, 2, 22, 31, 3, _lib.IMPL_WARP_MMA, True),
    (12, 6, 1, 24, 36, 2, _lib.IMPL_WARP_MMA, False), (12, 6, 1, 13, 25, 2, _lib.IMPL_TCGEN05, True),
    (12, 0, 1, 35, 24, 2, _lib.IMPL_TCGEN05, True), (12, 6, 2, 24, 12, 2, _lib.IMPL_TCGEN05, False),
    (7, 3, 1, 9, 11, 2, _lib.IMPL_SIMT, True), (12, 6, 1, 13, 14, 1, _lib.IMPL_SIMT, True),
]


@pytest.mark.parametrize("case", FWD)
def test_forward_writes_nothing_outside_its_outputs(case):
    ws, shift, b, h, w, nh, impl, with_pads = case
    c = nh * 32
    name = f"guardf{case}"
    hp, wp = -(-h // ws) * ws, -(-w // ws) * ws
    packed, packed_full = _guarded_in(randn(name + ".p", b, h, w, 3 * c).bfloat16())
    table, _t = _guarded_in(0.5 * randn(name + ".t", (2 * ws - 1) ** 2, nh))
    pads, _p = _guarded_in(randn(name + ".pad", 3 * c).bfloat16()) if with_pads else (None, None)
    o, o_full = _guarded_out((b, h, w, c), torch.bfloat16)
    lse, lse_full = _guarded_out((b, hp, wp, nh), torch.float32)
    es = 2
    base = packed.data_ptr()
    pad_ptr = (lambda i: None) if pads is None else (lambda i: ctypes.c_void_p(pads.data_ptr() + i * c * es))
    shp = _lib.Shape(b, h, w, nh, ws, shift, _lib.BF16, impl, 32 ** -0.5, 0.0, 0, 3 * c, 3 * c, 3 * c, c, 0, 0)
    lib = _lib.load()
    rc = lib.swin_attn_forward(ctypes.byref(shp), ctypes.c_void_p(base), ctypes.c_void_p(base + c * es), ctypes.c_void_p(base + 2 * c * es),
                               pad_ptr(0), pad_ptr(1), pad_ptr(2), _ptr(table), None, _ptr(o), _ptr(lse), None)
    assert rc == 0, lib.swin_attn_last_error().decode()
    torch.cuda.synchronize()
    assert _bands_intact(o_full, o.numel()), "write outside o"
    assert _bands_intact(lse_full, lse.numel()), "write outside lse"
    assert torch.isfinite(o.float()).all() and torch.isfinite(lse[:, :h, :w]).all()
    # the same call through the torch op on plain tensors
    want, want_lse = torch.ops.swin_attn.fwd(packed.clone(), None, table.clone(), None if pads is None else pads.clone(), None, ws, shift,
                                             nh, 32 ** -0.5, True, impl, None)
    assert torch.equal(o, want)
    assert torch.equal(lse[:, :h, :w], want_lse[:, :h, :w])


# ws, shift, B, H, W, nH: window-aligned (what swin_attn_backward takes); tcgen05 (7) and warp-MMA (12) kernels, CUDA cores
BWD = [(7, 3, 2, 14, 21, 3, _lib.IMPL_AUTO), (7, 0, 1, 7, 7, 2, _lib.IMPL_AUTO), (12, 6, 2, 12, 24, 3, _lib.IMPL_AUTO),
       (12, 0, 1, 36, 24, 2, _lib.IMPL_WARP_MMA), (12, 6, 1, 12, 12, 1, _lib.IMPL_SIMT), (7, 3, 1, 14, 7, 2, _lib.IMPL_SIMT)]


@pytest.mark.parametrize("case", BWD)
def test_backward_writes_nothing_outside_its_outputs(case):
    ws, shift, b, h, w, nh, impl = case
    c = nh * 32
    name = f"guardb{case}"
    packed, _pf = _guarded_in(randn(name + ".p", b, h, w, 3 * c).bfloat16())
    table, _t = _guarded_in(0.5 * randn(name + ".t", (2 * ws - 1) ** 2, nh))
    o_t, lse_t = torch.ops.swin_attn.fwd(packed.clone(), None, table.clone(), None, None, ws, shift, nh, 32 ** -0.5, True, impl, None)
    o, _o = _guarded_in(o_t)
    lse, _l = _guarded_in(lse_t)
    d_o, _d = _guarded_in(randn(name + ".c", b, h, w, c).bfloat16())
    d_packed, dp_full = _guarded_out((b, h, w, 3 * c), torch.bfloat16)
    d_table, dt_full = _guarded_out(((2 * ws - 1) ** 2, nh), torch.float32)
    es = 2
    base, dbase = packed.data_ptr(), d_packed.data_ptr()
    shp = _lib.Shape(b, h, w, nh, ws, shift, _lib.BF16, impl, 32 ** -0.5, 0.0, 0, 3 * c, 3 * c, 3 * c, c, 0, 0)
    lib = _lib.load()
    vp = ctypes.c_void_p
    rc = lib.swin_attn_backward(ctypes.byref(shp), vp(base), vp(base + c * es), vp(base + 2 * c * es), _ptr(o), _ptr(d_o), _ptr(lse),
                                _ptr(table), None, vp(dbase), vp(dbase + c * es), vp(dbase + 2 * c * es), _ptr(d_table), None)
    assert rc == 0, lib.swin_attn_last_error().decode()
    torch.cuda.synchronize()
    assert _bands_intact(dp_full, d_packed.numel()), "write outside dq / dk / dv"
    assert _bands_intact(dt_full, d_table.numel()), "write outside d_table"
    assert torch.isfinite(d_packed.float()).all() and torch.isfinite(d_table).all()
    want = torch.ops.swin_attn.bwd(d_o.clone(), packed.clone(), None, o_t, lse_t, table.clone(), ws, shift, nh, 32 ** -0.5, None, impl)
    if impl == _lib.IMPL_SIMT:   # the CUDA-core backward accumulates dk / dv with atomics: the order is not fixed
        assert float((d_packed.float() - want[0].float()).abs().max()) <= 1e-2 * float(want[0].float().abs().max())
    else:
        assert torch.equal(d_packed, want[0])
    # the bins are summed with atomics across CTAs: the order is not fixed
    assert torch.allclose(d_table, want[2], rtol=1e-4, atol=1e-5 * float(want[2].abs().max()))


@pytest.mark.parametrize("case", [(7, 3, 2, 15, 23, 3), (12, 6, 1, 13, 25, 2), (12, 0, 2, 22, 31, 3), (12, 6, 1, 29, 17, 5)])
def test_unpadded_backward_writes_nothing_outside_its_outputs(case):
    """swin_attn_backward_unpadded (tcgen05 window 7, warp-MMA window 12): guard bands around dq | dk | dv, d_table and the
    pad-vector gradients; finite results."""
    ws, shift, b, h, w, nh = case
    c = nh * 32
    name = f"guardu{case}"
    hp, wp = -(-h // ws) * ws, -(-w // ws) * ws
    packed, _pf = _guarded_in(randn(name + ".p", b, h, w, 3 * c).bfloat16())
    table, _t = _guarded_in(0.5 * randn(name + ".t", (2 * ws - 1) ** 2, nh))
    pads, _p = _guarded_in(randn(name + ".pad", 3 * c).bfloat16())
    o_t, lse_t = torch.ops.swin_attn.fwd(packed.clone(), None, table.clone(), pads.clone(), None, ws, shift, nh, 32 ** -0.5, True,
                                         _lib.IMPL_AUTO, None)
    assert lse_t.shape == (b, hp, wp, nh)
    o, _o = _guarded_in(o_t)
    lse, _l = _guarded_in(lse_t)
    d_o, _d = _guarded_in(randn(name + ".c", b, h, w, c).bfloat16())
    d_packed, dp_full = _guarded_out((b, h, w, 3 * c), torch.bfloat16)
    d_table, dt_full = _guarded_out(((2 * ws - 1) ** 2, nh), torch.float32)
    d_kv_pad, dkv_full = _guarded_out((2 * c,), torch.float32)
    es = 2
    base, dbase = packed.data_ptr(), d_packed.data_ptr()
    shp = _lib.Shape(b, h, w, nh, ws, shift, _lib.BF16, _lib.IMPL_AUTO, 32 ** -0.5, 0.0, 0, 3 * c, 3 * c, 3 * c, c, 0, 0)
    lib = _lib.load()
    vp = ctypes.c_void_p
    rc = lib.swin_attn_backward_unpadded(ctypes.byref(shp), vp(base), vp(base + c * es), vp(base + 2 * c * es),
                                         vp(pads.data_ptr() + c * es), vp(pads.data_ptr() + 2 * c * es), _ptr(o), _ptr(d_o), _ptr(lse),
                                         _ptr(table), vp(dbase), vp(dbase + c * es), vp(dbase + 2 * c * es), _ptr(d_table),
                                         vp(d_kv_pad.data_ptr()), vp(d_kv_pad.data_ptr() + c * 4), None)
    assert rc == 0, lib.swin_attn_last_error().decode()
    torch.cuda.synchronize()
    assert _bands_intact(dp_full, d_packed.numel()), "write outside dq / dk / dv"
    assert _bands_intact(dt_full, d_table.numel()), "write outside d_table"
    assert _bands_intact(dkv_full, d_kv_pad.numel()), "write outside d_k_pad / d_v_pad"
    assert torch.isfinite(d_packed.float()).all() and torch.isfinite(d_table).all() and torch.isfinite(d_kv_pad).all()
    assert float(d_kv_pad.abs().max()) > 0.0


@pytest.mark.parametrize("case", [(2, 418, 4, 64, torch.bfloat16), (1, 300, 2, 64, torch.bfloat16), (2, 77, 3, 32, torch.float32),
                                  (1, 449, 2, 64, torch.bfloat16)])
def test_global_attention_writes_nothing_outside_its_outputs(case):
    b, s, nh, hd, dtype = case
    c = nh * hd
    name = f"guardm{case}"
    q, _q = _guarded_in(randn(name + ".q", b, s, c).to(dtype))
    k, _k = _guarded_in(randn(name + ".k", b, s, c).to(dtype))
    v, _v = _guarded_in(randn(name + ".v", b, s, c).to(dtype))
    o, o_full = _guarded_out((b, s, c), dtype)
    amap, m_full = _guarded_out((b, nh, s, s), torch.float32)
    shp = _lib.MhaShape(b, s, s, nh, hd, hd, _lib.F32 if dtype == torch.float32 else _lib.BF16, _lib.IMPL_AUTO, hd ** -0.5, 0, c, c, c, c)
    lib = _lib.load()
    rc = lib.mha_attention_forward(ctypes.byref(shp), _ptr(q), _ptr(k), _ptr(v), _ptr(o), _ptr(amap), None)
    assert rc == 0, lib.swin_attn_last_error().decode()
    torch.cuda.synchronize()
    assert _bands_intact(o_full, o.numel()), "write outside o"
    assert _bands_intact(m_full, amap.numel()), "write outside the attention map"
    assert torch.isfinite(o.float()).all() and torch.isfinite(amap).all()
    rows = amap.sum(-1)
    assert torch.allclose(rows, torch.ones_like(rows), atol=2e-2 if dtype == torch.bfloat16 else 1e-4)
