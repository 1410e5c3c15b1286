"""Degenerate inputs through the C ABI: empty slabs, single cells, targets wholly outside the source, single-level columns,
all-NaN fields.  The reference's loops simply run zero times for these (CachedInterpolation.cc:173-209 computes newSize =
size / (inX*inY) * outX*outY and loops over it; CDMVerticalInterpolator.cc:488-561 loops over the output levels); the drop-in must
return the same (empty or undefined) results, never an error or a launch with a zero-sized grid."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import dp, fp
from test_gpu_vertical_forward import _close, _oracle_vertical

pytestmark = pytest.mark.gpu


def _eq(a, b):
    a, b = np.asarray(a), np.asarray(b)
    na, nb = np.isnan(a), np.isnan(b)
    return a.shape == b.shape and np.array_equal(na, nb) and np.array_equal(a[~na], b[~nb])


@pytest.fixture(scope="module")
def fx():
    import fimex_b200
    assert fimex_b200.lib.fimexcu_device_count() >= 1
    return fimex_b200


@pytest.mark.parametrize("method", [ol.BILINEAR, ol.BICUBIC, ol.NEAREST])
def test_empty_slab_gives_an_empty_result(fx, method):
    """size = 0: newSize = 0 and nothing is touched (host, device and packed entry points)"""
    import torch
    rng = np.random.default_rng(1)
    in_x, in_y, out_x, out_y = 12, 9, 20, 15
    px, py = rng.uniform(0, in_x - 1, out_x * out_y), rng.uniform(0, in_y - 1, out_x * out_y)
    ci = fx.createCachedInterpolation("x", "y", method, px, py, in_x, in_y, out_x, out_y)
    got, new_size = ci.interpolateValues(np.empty((0, ci.getInY(), ci.getInX()), np.float32))
    assert new_size == 0 and got.size == 0
    got_dev, n = ci.interpolateValues(torch.empty((0, ci.getInY(), ci.getInX()), dtype=torch.float32, device="cuda"))
    assert n == 0 and got_dev.numel() == 0
    packed = ci.interpolateValuesPacked(np.empty((0, ci.getInY(), ci.getInX()), np.int16), -32767.0)
    assert packed.size == 0
    # fewer values than one level: the reference's integer division gives zero levels
    new = C.c_size_t(123)
    few = np.zeros(ci.getInX() * ci.getInY() - 1, np.float32)
    out = np.full(out_x * out_y, 7.0, np.float32)
    nan = float("nan")
    fx._lib.check(fx.lib.fimexcu_cache_apply_f(ci._h, few.ctypes.data, few.size, out.ctypes.data, C.byref(new), nan, nan), "apply")
    assert new.value == 0 and (out == 7.0).all()


@pytest.mark.parametrize("method", [ol.BILINEAR, ol.BICUBIC, ol.NEAREST])
def test_target_wholly_outside_the_source(fx, method):
    """no target cell has a source cell: every value undefined (NaN, or the fill value), no row band at all"""
    in_x, in_y, out_x, out_y, nz = 10, 8, 40, 30, 3
    gx, gy = np.meshgrid(np.arange(out_x, dtype=np.float64), np.arange(out_y, dtype=np.float64))
    px, py = 100.0 + gx, -50.0 - gy
    fld = np.random.default_rng(2).normal(size=(nz, in_y, in_x)).astype(np.float32)
    want, _ = ol.o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, fld)
    assert np.isnan(want).all()
    ci = fx.createCachedInterpolation("x", "y", method, px, py, in_x, in_y, out_x, out_y)
    bands, frac = ci.rowBands()
    assert frac == 0.0 and all(x0 >= x1 for _, _, x0, x1 in bands)
    sub = ci.reduceInput(fld, in_x, in_y)
    got, _ = ci.interpolateValues(sub)
    assert _eq(got.reshape(want.shape), want)
    got, _ = ci.interpolateValues(sub, fillValue=-1.5)
    assert (got == np.float32(-1.5)).all()


def test_single_cell_grids(fx):
    """1 x 1 target on a 2 x 2 source, one level; and a 1 x 1 source (only an exact hit is defined for bilinear)"""
    fld = np.array([[[1.0, 2.0], [3.0, 5.0]]], np.float32)
    for method in (ol.BILINEAR, ol.NEAREST, ol.BICUBIC):
        for x, y in ((0.25, 0.5), (0.0, 0.0), (1.0, 1.0), (0.999, 0.001), (1.2, 0.3)):
            px, py = np.array([x]), np.array([y])
            want, _ = ol.o_cached_apply(method, px, py, 2, 2, 1, 1, fld)
            ci = fx.createCachedInterpolation("x", "y", method, px, py, 2, 2, 1, 1)
            got, _ = ci.interpolateValues(ci.reduceInput(fld, 2, 2))
            assert _eq(got.reshape(want.shape), want), (method, x, y)
    one = np.array([[[4.0]]], np.float32)
    for method in (ol.BILINEAR, ol.NEAREST):
        for x, y in ((0.0, 0.0), (0.2, 0.0), (-0.4, 0.3)):
            px, py = np.array([x]), np.array([y])
            want, _ = ol.o_cached_apply(method, px, py, 1, 1, 1, 1, one)
            ci = fx.createCachedInterpolation("x", "y", method, px, py, 1, 1, 1, 1)
            got, _ = ci.interpolateValues(ci.reduceInput(one, 1, 1))
            assert _eq(got.reshape(want.shape), want), (method, x, y)


def test_all_nan_field_and_nan_positions(fx):
    """a field without a single number, and index tables with NaN / infinite entries (CachedInterpolation casts them: undefined)"""
    rng = np.random.default_rng(3)
    in_x, in_y, out_x, out_y, nz = 15, 11, 23, 17, 2
    px, py = rng.uniform(0, in_x - 1, out_x * out_y), rng.uniform(0, in_y - 1, out_x * out_y)
    px[::7] = np.nan
    py[3::11] = np.inf
    px[5::13] = -np.inf
    good = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    for method in (ol.BILINEAR, ol.NEAREST, ol.BICUBIC):
        ci = fx.createCachedInterpolation("x", "y", method, px, py, in_x, in_y, out_x, out_y)
        for fld in (np.full((nz, in_y, in_x), np.nan, np.float32), good):
            want, _ = ol.o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, fld)
            got, _ = ci.interpolateValues(ci.reduceInput(fld, in_x, in_y))
            assert _eq(got.reshape(want.shape), want), method


@pytest.mark.parametrize("method", [ol.FWD_MEAN, ol.FWD_MEDIAN, ol.FWD_MAX, ol.FWD_SUM])
def test_forward_with_empty_cells_and_empty_slab(fx, method):
    """most target cells receive no source point (NaN), some sources fall outside the target; zero levels"""
    import torch
    rng = np.random.default_rng(4)
    in_x, in_y, out_x, out_y, nz = 9, 7, 40, 30, 3
    px, py = rng.uniform(-3, out_x + 3, in_x * in_y), rng.uniform(-3, out_y + 3, in_x * in_y)
    fld = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    fld[0, 2, 3] = np.nan
    want, _ = ol.o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, fld, level_bug=0)
    fwd = fx.CachedForwardInterpolation("x", "y", method, px, py, in_x, in_y, out_x, out_y, level_compat=False)
    got, _ = fwd.interpolateValues(fld)
    assert _eq(got.reshape(want.shape), want) and np.isnan(want).mean() > 0.9
    got0, n0 = fwd.interpolateValues(np.empty((0, in_y, in_x), np.float32))
    assert n0 == 0 and got0.size == 0
    got0d, n0d = fwd.interpolateValues(torch.empty((0, in_y, in_x), dtype=torch.float32, device="cuda"))
    assert n0d == 0 and got0d.numel() == 0


@pytest.mark.parametrize("method", [ol.VINT_LIN, ol.VINT_LOG, ol.VINT_NN])
def test_vertical_degenerate_columns(fx, method):
    """two-level columns, a column of equal coordinates, target levels on, between and beyond the column, a NaN coordinate, one
    target level, no columns at all"""
    from fimex_b200 import vertical as V
    rng = np.random.default_rng(5)
    ncol = 97
    for nzi in (2, 3):
        p = np.sort(rng.uniform(100.0, 1000.0, (nzi, ncol)), axis=0)[::-1].astype(np.float32).copy()   # decreasing like pressure
        p[:, 5] = 500.0                       # no direction at all
        p[0, 9] = np.nan
        p[:, 11] = p[::-1, 11]                # increasing
        fld = rng.normal(size=(1, nzi, ncol)).astype(np.float32)
        fld[0, 0, 13] = np.nan
        for levels in (np.array([1000.0, 850.0, 500.0, 101.0, 50.0]), np.array([float(p[0, 0])]), np.array([2000.0])):
            want = _oracle_vertical(method, fld, np.ascontiguousarray(p[None]), levels=levels)
            got = V.verticalInterpolate(method, fld, np.ascontiguousarray(p[None]), level1=levels)
            ok = _close if method == ol.VINT_LOG else _eq   # the LOG weights may differ by one f32 ulp where the device log is used
            assert ok(np.asarray(got).reshape(want.shape), want), (nzi, levels)
    empty = V.verticalInterpolate(method, np.empty((1, 3, 0), np.float32), np.empty((1, 3, 0), np.float32), level1=np.array([500.0]))
    assert np.asarray(empty).size == 0


def test_hybrid_vertical_two_levels_and_constant_surface(fx):
    """hybrid-sigma source with the minimum of two levels, every column the same surface pressure (one histogram bin), levels beyond
    both ends"""
    from fimex_b200 import vertical as V
    rng = np.random.default_rng(6)
    ncol = 5000
    ap, b = np.array([0.0, 0.0]), np.array([0.3, 1.0])
    ps = np.full((1, ncol), 1000.0, np.float32)
    ps[0, ::97] = np.nan
    fld = rng.normal(size=(1, 2, ncol)).astype(np.float32)
    levels = np.array([1100.0, 1000.0, 650.0, 300.0, 200.0])
    p = (ap[:, None] + b[:, None] * ps.astype(np.float64)).astype(np.float32)[None]
    for method in (ol.VINT_LIN, ol.VINT_LOG):
        want = _oracle_vertical(method, fld, np.ascontiguousarray(p), levels=levels)
        got = V.verticalHybridPressure(method, fld, ap, b, ps, levels)
        ok = _close if method == ol.VINT_LOG else _eq
        assert ok(np.asarray(got).reshape(want.shape), want), method


def test_coordinate_search_without_targets_or_without_sources(fx):
    from fimex_b200 import coordinates as K
    rng = np.random.default_rng(7)
    lon, lat = rng.uniform(0, 20, 50 * 40), rng.uniform(50, 65, 50 * 40)
    # no targets
    ix, iy = K.flannTranslatePointsToClosestInputCell(0.1, np.empty(0), np.empty(0), lon.reshape(40, 50), lat.reshape(40, 50))
    assert ix.size == 0 and iy.size == 0
    ix, iy = K.fastTranslatePointsToClosestInputCell(np.empty(0), np.empty(0), lon.reshape(40, 50), lat.reshape(40, 50))
    assert ix.size == 0 and iy.size == 0
    # a source cloud without a single valid position: every target -1 (kd-tree) as in the reference
    nanlon = np.full((40, 50), np.nan)
    tx, ty = rng.uniform(0, 20, 30), rng.uniform(50, 65, 30)
    ix, iy = K.flannTranslatePointsToClosestInputCell(50000.0, tx.copy(), ty.copy(), nanlon, nanlon)
    wx, wy = tx.copy(), ty.copy()
    assert ol.ref().ref_kdtree_translate(50000.0, dp(wx), dp(wy), wx.size, dp(nanlon.ravel().copy()), dp(nanlon.ravel().copy()), 50, 40) == 0
    assert np.array_equal(np.asarray(ix).ravel(), wx) and np.array_equal(np.asarray(iy).ravel(), wy)


def test_reproject_nothing_and_fill_nothing(fx):
    from fimex_b200 import reproject as RP
    from fimex_b200 import interpolation as I
    x, y = RP.reproject_values("+proj=longlat +R=6371000", "+proj=stere +lat_0=90 +lon_0=0 +lat_ts=60 +R=6371000", np.empty(0), np.empty(0))
    assert x.size == 0 and y.size == 0
    allnan = np.full((6, 7), np.nan, np.float32)
    filled, n = I.mifi_fill2d_f(allnan.copy())
    wanted = allnan.copy()
    cnt = C.c_size_t()
    assert ol.oracle().orc_fill2d_f(7, 6, fp(wanted), 0.01, 1.6, 100, C.byref(cnt)) == 1
    assert _eq(filled, wanted)
    full = np.arange(42, dtype=np.float32).reshape(6, 7)
    filled, n = I.mifi_fill2d_f(full.copy())
    assert np.array_equal(filled, full)


@pytest.mark.parametrize("method", [ol.VINT_LIN, ol.VINT_LOG, ol.VINT_LIN_NO])
def test_vertical_explicit_coordinate_unsorted_and_repeated_levels(fx, method):
    """target levels in no order, one of them twice, one NaN, through the explicit-coordinate path (scan + bisection: the level
    list need not be sorted), on increasing, decreasing and a few non-monotone columns"""
    from fimex_b200 import vertical as V
    rng = np.random.default_rng(8)
    nzi, ncol = 33, 4000
    eta = np.linspace(0.02, 1.0, nzi)[:, None]
    ps = rng.uniform(900.0, 1040.0, (1, ncol))
    p = (ps * eta + 3.0 * (1 - eta)).astype(np.float32)
    p[:, ::5] = p[::-1, ::5]                 # every fifth column decreasing
    p[7, 11] = p[6, 11]                      # a repeated value: not strictly monotone
    p[3, 13] = np.nan
    fld = rng.normal(280, 20, (1, nzi, ncol)).astype(np.float32)
    fld[rng.random(fld.shape) < 0.02] = np.nan
    levels = np.array([500.0, 1000.0, 30.0, 850.0, 500.0, np.nan, 925.0, 1100.0, 1.0, float(p[5, 0])])
    want = _oracle_vertical(method, fld, np.ascontiguousarray(p[None]), levels=levels)
    got = V.verticalInterpolate(method, fld, np.ascontiguousarray(p[None]), level1=levels)
    ok = _close if method == ol.VINT_LOG else _eq
    assert ok(np.asarray(got).reshape(want.shape), want)


@pytest.mark.parametrize("nzi", [93, 137, 300])
def test_vertical_explicit_coordinate_long_columns(fx, nzi):
    """columns whose staged copy needs more than the default 48 KB of shared memory per block (137 levels: opt-in limit), and more
    than the kernel takes at all (300 levels: the sweep kernel serves them)"""
    from fimex_b200 import vertical as V
    rng = np.random.default_rng(9 + nzi)
    ncol = 3000
    eta = np.linspace(0.005, 1.0, nzi)[:, None]
    ps = rng.uniform(900.0, 1040.0, (1, ncol))
    p = (ps * eta + 2.0 * (1 - eta)).astype(np.float32)
    fld = rng.normal(280, 20, (1, nzi, ncol)).astype(np.float32)
    levels = np.array([1000.0, 925.0, 850.0, 700.0, 500.0, 300.0, 100.0, 10.0, 1.0])
    for method in (ol.VINT_LIN, ol.VINT_LOG):
        want = _oracle_vertical(method, fld, np.ascontiguousarray(p[None]), levels=levels)
        got = V.verticalInterpolate(method, fld, np.ascontiguousarray(p[None]), level1=levels)
        ok = _close if method == ol.VINT_LOG else _eq
        assert ok(np.asarray(got).reshape(want.shape), want), (nzi, method)
