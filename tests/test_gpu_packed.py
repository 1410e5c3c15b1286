"""GPU parity of the packed-slab feed / sink passes against the oracle restatement, which is pinned
against the reference's own ScaleValue / data_caster templates in test_oracle_vs_ref.py.  Bar: bit-exact."""
import os

import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import fp

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def P():
    import fimex_b200
    assert fimex_b200.lib.fimexcu_device_count() >= 1
    from fimex_b200 import packed
    return packed


def _torch_dtype(dtype):
    import torch
    return {0: torch.int8, 1: torch.int16, 2: torch.int32, 3: torch.uint8, 4: getattr(torch, "uint16", None), 5: getattr(torch, "uint32", None)}[dtype]


@pytest.mark.parametrize("dtype", sorted(ol.PACKED_DTYPES))
def test_unpack_pack_vs_oracle(P, dtype):
    import torch
    td = _torch_dtype(dtype)
    if td is None:
        pytest.skip("torch build without this unsigned type")
    rng = np.random.default_rng(50 + dtype)
    T = ol.PACKED_DTYPES[dtype]
    info = np.iinfo(T)
    n = 300007
    raw = rng.integers(info.min, int(info.max) + 1, n, dtype=np.int64).astype(T)
    bad = float(info.max) if dtype != 1 else -32767.0
    raw[::23] = T(bad)
    o = ol.oracle()
    want = np.empty(n, np.float32)
    assert o.orc_unpack_f(dtype, raw.ctypes.data, n, bad, fp(want)) == 1
    d_raw = torch.from_numpy(raw.view(np.int16) if T == np.uint16 else raw.view(np.int32) if T == np.uint32 else raw)
    d_raw = d_raw.cuda().view(td)
    got = P.unpack(d_raw, bad).cpu().numpy()
    assert np.array_equal(got, want, equal_nan=True) and np.isnan(want).any()
    want_raw = want.copy()
    assert o.orc_unpack_scaled_f(dtype, raw.ctypes.data, n, bad, 0.01, 273.15, fp(want)) == 1
    got = P.unpackScaled(d_raw, bad, 0.01, 273.15).cpu().numpy()
    assert np.array_equal(got, want, equal_nan=True)
    span = float(info.max) - float(info.min)
    v = rng.uniform(info.min - 0.05 * span, info.max + 0.05 * span, n).astype(np.float32)
    v[:5000] = np.round(v[:5000]) + np.float32(0.5)
    v[5000:6000] = np.nan
    v[6000:6100] = np.float32(0.49999997) * rng.choice([-1, 1], 100)
    wantp = np.empty(n, T)
    assert o.orc_pack_f(dtype, fp(v), n, bad, wantp.ctypes.data) == 1
    gotp = P.pack(torch.from_numpy(v).cuda(), td, bad)
    gotp = gotp.view(torch.int16 if T == np.uint16 else torch.int32 if T == np.uint32 else td).cpu().numpy().view(T)
    assert np.array_equal(gotp, wantp)
    # bases off the 16-byte boundary (a slice of a larger array): the element-by-element path, same answers
    got1 = P.unpack(d_raw[1:], bad).cpu().numpy()
    assert np.array_equal(got1, want_raw[1:], equal_nan=True)
    gotp1 = P.pack(torch.from_numpy(v).cuda()[3:], td, bad)
    gotp1 = gotp1.view(torch.int16 if T == np.uint16 else torch.int32 if T == np.uint32 else td).cpu().numpy().view(T)
    assert np.array_equal(gotp1, wantp[3:])


def test_packed_vertical_chain_on_device(P):
    """CDMVerticalInterpolator::getDataSlice for a CDM_SHORT variable (CDMVerticalInterpolator.cc:478-480, 563): raw values -> float,
    fill -> NaN, hybrid-sigma levels to pressure levels, back to short; ps arrives packed too and is read with
    getScaledDataSliceInUnit.  All passes on the device; compared with the oracle chain, bit-exact."""
    import torch
    import test_gpu_vertical_forward as T
    from fimex_b200.vertical import verticalHybridPressure
    rng = np.random.default_rng(5)
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_vertical.npz"))
    ap, b = g["en_ap_pa"] * 0.01, g["en_b"]
    nz, nI = 65, 3001
    levels = np.array([1000, 925, 850, 700, 500, 400, 300, 250, 200, 150, 100, 50], dtype=np.float64)
    raw = rng.integers(-20000, 20000, (1, nz, nI)).astype(np.int16)
    raw[rng.random(raw.shape) < 0.02] = -32767
    ps_raw = rng.integers(-25000, 10000, (1, nI)).astype(np.int16)  # Pa = 90000 + 1.5 * raw
    ps_raw[0, ::97] = -32767
    o = ol.oracle()
    f = np.empty(raw.shape, np.float32)
    o.orc_unpack_f(1, raw.ctypes.data, raw.size, -32767.0, fp(f))
    ps = np.empty(ps_raw.shape, np.float32)
    o.orc_unpack_scaled_f(1, ps_raw.ctypes.data, ps_raw.size, -32767.0, 1.5 * 0.01, 90000.0 * 0.01, fp(ps))  # scaled to hPa like getScaledDataSliceInUnit
    with np.errstate(all="ignore"):
        p = (ap[:, None] + b[:, None] * ps.astype(np.float64)).astype(np.float32)[None]
    want_f = T._oracle_vertical(ol.VINT_LIN, f, np.ascontiguousarray(p), levels=levels)
    want = np.empty(want_f.shape, np.int16)
    o.orc_pack_f(1, fp(np.ascontiguousarray(want_f)), want_f.size, -32767.0, want.ctypes.data)
    d_f = P.unpack(raw, -32767.0)
    d_ps = P.unpackScaled(ps_raw, -32767.0, 1.5 * 0.01, 90000.0 * 0.01)
    d_out = verticalHybridPressure(ol.VINT_LIN, d_f, ap, b, d_ps, levels, ps_scale=1.0)
    got = P.pack(d_out, torch.int16, -32767.0).cpu().numpy()
    assert np.array_equal(got, want)
    assert (want == -32767).any() and (want != -32767).any()
