"""Full-size parity on every BASELINE config, against the reference itself.

`oracle/_ref/libfimexref_omp.so` is fimex's own CachedInterpolation / CachedForwardInterpolation / vector-rotation code compiled
from /root/reference (OpenMP build); it regrids a whole 65-level MEPS slab in well under a second, so every level of
S->T1 and S->T2 is compared bit for bit (values, undefined masks, fill values) -- no sampling, no tolerance:
  configs[1]  cached bilinear, staged (padded) kernel, dense device entry and the host-pointer ABI
  configs[2]  bicubic and nearest with 2 % NaN blobs and the _FillValue variant; x_wind / y_wind through the fused rotation
  configs[3]  forward_mean / forward_max / forward_median on the 2030x1354 swath
  configs[4]  hybrid-sigma -> 12 pressure levels (log) chained behind the bilinear regridding, every column of one member
              (checker: the oracle restatement; CDMVerticalInterpolator itself cannot be linked without NetCDF/udunits)
"""
import os

import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import dp, fp

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FILL = np.float32(9.96921e36)  # CDM's default float fill value (include/fimex/CDMconstants.h:151-160)


def _eq(a, b):
    """same undefined mask, same bits everywhere else"""
    a, b = np.asarray(a), np.asarray(b)
    if a.shape != b.shape:
        return False
    na, nb = np.isnan(a), np.isnan(b)
    return np.array_equal(na, nb) and np.array_equal(a[~na], b[~nb])


@pytest.fixture(scope="module")
def fx():
    import fimex_b200
    assert fimex_b200.lib.fimexcu_device_count() >= 1, "no sm_100 device: the product has no CPU fallback"
    return fimex_b200


@pytest.fixture(scope="module")
def refomp():
    r = ol.ref(omp=True)
    if r is None:
        pytest.skip("oracle/_ref not built")
    return r


def _blobs(fld, rng, n=40, size=20):
    """~2 % of the source cells undefined in 40 random 20x20 blobs (the same cells on every level)"""
    _, ny, nx = fld.shape
    for _ in range(n):
        y0, x0 = rng.integers(0, ny - size), rng.integers(0, nx - size)
        fld[:, y0:y0 + size, x0:x0 + size] = np.nan
    return fld


@pytest.fixture(scope="module")
def slab():
    from fimex_b200 import synthetic as S
    fld = S.meps_slab_numpy(t=7)  # all 65 levels
    return _blobs(fld, np.random.default_rng(5))


def _ref_apply(method, px, py, ox, oy, fld):
    from fimex_b200 import synthetic as S
    out, info = ol.r_cached_apply(method, px, py, S.MEPS_NX, S.MEPS_NY, ox, oy, fld, omp=True)
    return out, info


# ---------------------------------------------------------------------------------- configs[1] and [2]: all 65 levels
@pytest.mark.parametrize("target", ["T1", "T2"])
@pytest.mark.parametrize("method", [ol.BILINEAR, ol.NEAREST, ol.BICUBIC])
def test_full_slab_equals_reference(fx, refomp, slab, target, method):
    import torch
    from fimex_b200 import synthetic as S
    px, py, ox, oy = S.meps_index_tables(target)
    want, info = _ref_apply(method, px, py, ox, oy, slab)
    assert want.shape == (S.MEPS_NZ, oy, ox)
    ci = fx.createCachedInterpolation("x", "y", method, px, py, S.MEPS_NX, S.MEPS_NY, ox, oy)
    assert (ci.getInX(), ci.getInY()) == (info["inX"], info["inY"])
    sub_h = ci.reduceInput(slab, S.MEPS_NX, S.MEPS_NY)
    # (a) the host-pointer ABI (what CachedInterpolationInterface::interpolateValues forwards to)
    got, n = ci.interpolateValues(sub_h)
    assert n == want.size and _eq(got.reshape(want.shape), want)
    # (b) device-resident, the reference's dense layout
    sub_d = torch.from_numpy(sub_h).cuda()
    dense, _ = ci.interpolateValues(sub_d)
    torch.cuda.synchronize()
    assert _eq(dense.cpu().numpy().reshape(want.shape), want)
    del dense
    # (c) device-resident, padded layout (the TMA-staged kernels for bilinear and nearest)
    out = ci.interpolateValuesPadded(ci.toPadded(sub_d))
    torch.cuda.synchronize()
    assert _eq(out[:, :, :ox].cpu().numpy(), want)
    del out
    # (d) the _FillValue variant: undefined input arrives as the fill value and the result carries it (CDMInterpolator.cc:140-151)
    raw = sub_h.copy()
    raw[np.isnan(raw)] = FILL
    wfill = want.copy()
    wfill[np.isnan(wfill)] = FILL
    gotf, _ = ci.interpolateValues(raw, badValue=float(FILL), fillValue=float(FILL))
    assert _eq(gotf.reshape(want.shape), wfill)
    outf = ci.interpolateValuesPadded(ci.toPadded(torch.from_numpy(raw).cuda()), badValue=float(FILL), fillValue=float(FILL))
    torch.cuda.synchronize()
    assert _eq(outf[:, :, :ox].cpu().numpy(), wfill)
    assert np.isnan(want).any() and (~np.isnan(want)).mean() > 0.5


@pytest.mark.parametrize("method", [ol.BILINEAR, ol.BICUBIC, ol.NEAREST])
def test_full_slab_wind_rotation_equals_reference_steps(fx, refomp, slab, method):
    """configs[2]: x_wind / y_wind S->T1, all 65 levels: interpolate u, interpolate v, rotate by the matrix -- the reference's
    three steps with its own code (CDMInterpolator.cc:232-259, reproject.cc:488-507) against the fused kernels.
    The wind is a constant geostrophic-like flow (10, 5) m/s plus noise, so the rotated components are sane numbers."""
    import torch
    from fimex_b200 import synthetic as S
    from fimex_b200.interpolation import interpolateVectorValues, interpolateVectorValuesPadded
    px, py, ox, oy = S.meps_index_tables("T1")
    lon, lat = S.target_axes("T1")
    # rotation angle of the Lambert grid against true north at every target cell: n * (lon - lon0)
    n_cone = np.sin(np.radians(63.3))
    ang = np.radians(np.broadcast_to(n_cone * (lon - 15.0)[None, :], (oy, ox))).ravel()
    m3 = np.stack([np.cos(ang), np.sin(ang), np.degrees(ang)], axis=1).copy()
    rng = np.random.default_rng(17)
    u = (10.0 + slab - 280.0).astype(np.float32)
    v = (5.0 + 0.5 * rng.standard_normal(slab.shape)).astype(np.float32)
    v[np.isnan(slab[:, ::-1, :])] = np.nan  # v's undefined blobs sit elsewhere than u's
    wu, _ = _ref_apply(method, px, py, ox, oy, u)
    wv, _ = _ref_apply(method, px, py, ox, oy, v)
    refomp.ref_vector_reproject_values_by_matrix_f(dp(m3), ox, oy, fp(wu), fp(wv), S.MEPS_NZ)
    ci = fx.createCachedInterpolation("x", "y", method, px, py, S.MEPS_NX, S.MEPS_NY, ox, oy)
    cvr = fx.CachedVectorReprojection(m3, ox, oy)
    su, sv = ci.reduceInput(u, S.MEPS_NX, S.MEPS_NY), ci.reduceInput(v, S.MEPS_NX, S.MEPS_NY)
    gu, gv = interpolateVectorValues(ci, cvr, su, sv)
    assert _eq(np.asarray(gu).reshape(wu.shape), wu) and _eq(np.asarray(gv).reshape(wv.shape), wv)
    if method == ol.BILINEAR:  # the staged fused kernel
        pu, pv = interpolateVectorValuesPadded(ci, cvr, ci.toPadded(torch.from_numpy(su).cuda()), ci.toPadded(torch.from_numpy(sv).cuda()))
        torch.cuda.synchronize()
        assert _eq(pu[:, :, :ox].cpu().numpy(), wu) and _eq(pv[:, :, :ox].cpu().numpy(), wv)
    assert (~np.isnan(wu)).mean() > 0.5


# ---------------------------------------------------------------------------------- configs[3]: forward aggregation on the swath
@pytest.mark.parametrize("method", [ol.FWD_MEAN, ol.FWD_MAX, ol.FWD_MEDIAN, ol.FWD_SUM, ol.FWD_UNDEF_MEAN])
def test_swath_forward_equals_reference(fx, refomp, method):
    """2030x1354 swath (30 % masked) -> 0.025 deg grid, 4 levels, against CachedForwardInterpolation compiled from the reference
    (so with its level-0 quirk, level_compat = 1), every cell"""
    from test_coordinate_search import swath
    rng = np.random.default_rng(33)
    nx, ny = 1354, 2030
    lon, lat = swath(rng, nx, ny, nan_frac=0.005)
    ox, oy, nz = 800, 480, 4
    pox = ((lon - 0.0) / 0.025).ravel()   # positions of the source cells on the target axes (lon 0..20, lat 58..70)
    poy = ((lat - 58.0) / 0.025).ravel()
    fld = rng.normal(280, 5, size=(nz, ny, nx)).astype(np.float32)
    fld[:, rng.random((ny, nx)) < 0.3] = np.nan
    want, info = ol.r_cached_apply(method, pox, poy, nx, ny, ox, oy, fld, omp=True)
    ci = fx.CachedForwardInterpolation("x", "y", method, pox, poy, nx, ny, ox, oy)
    assert (ci.getInX(), ci.getInY()) == (info["inX"], info["inY"])
    got, _ = ci.interpolateValues(ci.reduceInput(fld, nx, ny))
    assert _eq(np.asarray(got).reshape(want.shape), want)
    assert (~np.isnan(want)).mean() > 0.2


# ---------------------------------------------------------------------------------- configs[4]: bilinear, then hybrid -> pressure
def test_config5_chain_every_column_of_a_member(fx):
    """One ensemble member at full size: 65 hybrid levels + ps regridded S->T2 (padded, staged kernel), then 65 -> 12 pressure
    levels (log) with the pressure formed from the REGRIDDED ps -- all 449 361 columns against the oracle chain
    (undefined masks identical; values within the 1e-5 relative of BASELINE.json, the device's log may differ from glibc's
    in the last bit of the weight)."""
    import torch
    from fimex_b200 import synthetic as S
    from fimex_b200.vertical import verticalHybridPressure
    g = np.load(os.path.join(GOLD, "reference_vertical.npz"))
    ap, b = g["en_ap_pa"] * 0.01, g["en_b"]
    px, py, ox, oy = S.meps_index_tables("T2")
    rng = np.random.default_rng(23)
    fld = S.meps_slab_numpy(t=2)
    yy, xx = np.meshgrid(np.arange(S.MEPS_NY), np.arange(S.MEPS_NX), indexing="ij")
    h = 1250.0 * (1 + np.sin(xx / 97.0) * np.cos(yy / 131.0))
    ps = (101325.0 - 12.0 * h + rng.normal(0, 300, h.shape)).astype(np.float32)[None]
    levels = np.array([1000, 925, 850, 700, 500, 400, 300, 250, 200, 150, 100, 50], dtype=np.float64)
    # reference chain: fimex's own bilinear (oracle/_ref), then the oracle's vertical restatement
    h_f, _ = ol.r_cached_apply(ol.BILINEAR, px, py, S.MEPS_NX, S.MEPS_NY, ox, oy, fld, omp=True)
    h_ps, _ = ol.r_cached_apply(ol.BILINEAR, px, py, S.MEPS_NX, S.MEPS_NY, ox, oy, ps, omp=True)
    ncol = ox * oy
    p = (ap[:, None] + b[:, None] * (h_ps.reshape(1, ncol).astype(np.float64) * 0.01)).astype(np.float32)
    o = ol.oracle()
    want = np.empty((12, ncol), np.float32)
    hf2 = np.ascontiguousarray(h_f.reshape(65, ncol))
    assert o.orc_vertical_interpolate_f(ol.VINT_LOG, fp(hf2), fp(p), 65, ncol, dp(levels), None, 12, None, None, float("nan"), float("nan"), fp(want)) == 1
    # device chain
    ci = fx.CachedInterpolation("x", "y", ol.BILINEAR, px, py, S.MEPS_NX, S.MEPS_NY, ox, oy)
    r_f = ci.interpolateValuesPadded(ci.toPadded(torch.from_numpy(ci.reduceInput(fld, S.MEPS_NX, S.MEPS_NY)).cuda()))
    r_ps = ci.interpolateValuesPadded(ci.toPadded(torch.from_numpy(ci.reduceInput(ps, S.MEPS_NX, S.MEPS_NY)).cuda()))
    pitch = r_f.shape[2]
    n_inner = oy * pitch
    out = verticalHybridPressure(ol.VINT_LOG, r_f.reshape(1, 65, n_inner), ap, b, r_ps.reshape(1, n_inner), levels, ps_scale=0.01)
    torch.cuda.synchronize()
    got = out.reshape(12, oy, pitch)[..., :ox].cpu().numpy().reshape(12, ncol)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    ok = ~np.isnan(want)
    rel = np.abs(got[ok] - want[ok]) / np.abs(want[ok])
    assert rel.max() <= 1e-5, rel.max()
    assert (got[ok] == want[ok]).mean() > 0.999  # in fact all but the rare 1-ulp weights are bit-equal
    assert ok.mean() > 0.9
