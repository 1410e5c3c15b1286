"""Projection arithmetic pinned against REAL PROJ output.

The reference's NetCDF fixtures carry, for six projected grids, the double-precision longitude/latitude fields fimex generated
with PROJ (reproject::reproject_values, src/reproject.cc:351-415, proj4 string -> "+proj=longlat").  tests/golden/make_proj_pins.py
extracted them (axes, proj4 string, lon, lat) into tests/golden/reference_proj_pins.npz.  Checked here, for both the oracle
(oracle/miniproj.c) and the device projections (fimexcu_reproject_values, fxcu_proj.cu):

  inverse  proj(x, y) -> lon, lat  against the stored fields          tolerance 1e-11 degree  (1 um on the ground)
  forward  proj(lon, lat) -> x, y  against the stored axes            tolerance 1e-6 m / 1e-11 degree (ob_tran axes)

Observed: 3e-14 degree for all four fixtures with floating-point axes -- the last bits of asin/atan2 -- i.e. the same formulas
as PROJ for ob_tran (rotated lon/lat), lcc (1 standard parallel, sphere) and polar stere (sphere): the three projection families
the bench workloads and BASELINE.json's configs use.  Two fixtures (coordTest.nc, merge_target_top.nc) store their axes as
integers rounded to the metre after lon/lat had been generated; they pin to that rounding only (1 m, 1e-5 degree).
Ellipsoidal forms and the other projections have no PROJ output anywhere in the reference tree; they are pinned against Snyder's
published worked examples instead (tests/test_more_projections.py)."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import dp

PINS = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_proj_pins.npz"))
CASES = [str(c) for c in PINS["cases"]]
LL = "+proj=longlat +R=6.371e+06 +no_defs"


def _case(name):
    g = lambda k: PINS[name + "/" + k]  # noqa: E731
    return str(g("proj")), g("x"), g("y"), g("lon"), g("lat"), float(g("axis_rounding"))


def _oracle_tr(src, dst, x, y):
    xv, yv = np.array(x, dtype=np.float64).ravel().copy(), np.array(y, dtype=np.float64).ravel().copy()
    err = C.create_string_buffer(200)
    rc = ol.oracle().miniproj_transform_str(src.encode(), dst.encode(), xv.size, dp(xv), dp(yv), err, 200)
    assert rc >= 0, err.value
    return xv, yv


def _device_tr(src, dst, x, y):
    from fimex_b200 import reproject
    return reproject.reproject_values(src, dst, np.asarray(x, dtype=np.float64), np.asarray(y, dtype=np.float64))


def _check(tr, name):
    proj, x, y, lon, lat, rounding = _case(name)
    degrees_axes = "ob_tran" in proj
    # inverse: axes -> lon/lat
    glon, glat = tr(proj, LL, x, y)
    dlon = (glon - lon + 180.0) % 360.0 - 180.0
    away_from_pole = np.abs(lat) < 89.9999
    tol_deg = 1e-11 if rounding == 0.0 else 1e-5 * rounding
    assert np.abs(glat - lat).max() < tol_deg, (name, np.abs(glat - lat).max())
    assert np.abs(dlon[away_from_pole] * np.cos(np.radians(lat[away_from_pole]))).max() < tol_deg, name
    # forward: lon/lat -> axes
    fx, fy = tr(LL, proj, lon, lat)
    if degrees_axes:
        dx = (fx - x + 180.0) % 360.0 - 180.0
        assert np.abs(dx * np.cos(np.radians(y))).max() < 1e-11 and np.abs(fy - y).max() < 1e-11, name
    else:
        tol_m = 1e-6 if rounding == 0.0 else rounding
        assert np.abs(fx - x).max() < tol_m and np.abs(fy - y).max() < tol_m, (name, np.abs(fx - x).max(), np.abs(fy - y).max())


def test_fixture_is_what_the_script_writes():
    assert CASES == ["c11", "testdata_vertical_ensemble_in", "verticalPressure", "verticalSigmaHybrid", "coordTest", "merge_target_top"]
    kinds = {c: _case(c)[0].split()[0] for c in CASES}
    assert sorted(set(kinds.values())) == ["+proj=lcc", "+proj=ob_tran", "+proj=stere"]
    assert sum(_case(c)[1].size for c in CASES) > 2000


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_proj_output(name):
    _check(_oracle_tr, name)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_device_matches_proj_output(name):
    import fimex_b200
    assert fimex_b200.lib.fimexcu_device_count() >= 1
    _check(_device_tr, name)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_device_axes_reprojection_matches_proj_output(name):
    """the cache-build form of the same arithmetic: reproject_axes over the fixture's full lattice (K1, fxcu_proj.cu)"""
    from fimex_b200 import reproject
    proj, x, y, lon, lat, rounding = _case(name)
    xs, ys = np.unique(x), np.unique(y)
    px, py = reproject.reproject_axes(proj, LL, xs, ys)
    px, py = np.asarray(px).reshape(ys.size, xs.size), np.asarray(py).reshape(ys.size, xs.size)
    ix, iy = np.searchsorted(xs, x), np.searchsorted(ys, y)
    tol_deg = 1e-11 if rounding == 0.0 else 1e-5 * rounding
    dlon = (px[iy, ix] - lon + 180.0) % 360.0 - 180.0
    assert np.abs(py[iy, ix] - lat).max() < tol_deg
    assert np.abs(dlon * np.cos(np.radians(lat))).max() < tol_deg
