"""Pins oracle/miniproj.c, the CPU restatement of the projections (PROJ itself is not in the reference tree and not in this
image; its stored output for ob_tran / lcc / stere on the sphere is pinned in tests/test_proj_pins.py).  Pins here: round trips, closed-form / published coordinates, and the
reference's own PROJ-dependent unit tests replayed through its unmodified src/reproject.cc compiled over the proj.h shim
(test/testInterpolation.cc:274-470)."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import dp, fp
from test_gpu_projection import LL, LL_WGS, PROJS


def _tr(src, dst, x, y):
    xv, yv = np.array(x, dtype=np.float64).ravel().copy(), np.array(y, dtype=np.float64).ravel().copy()
    err = C.create_string_buffer(200)
    rc = ol.oracle().miniproj_transform_str(src.encode(), dst.encode(), xv.size, dp(xv), dp(yv), err, 200)
    assert rc >= 0, err.value
    return xv, yv


@pytest.mark.parametrize("name", sorted(PROJS))
def test_round_trip(name):
    rng = np.random.default_rng(1)
    lon, lat = rng.uniform(-20, 50, 5000), rng.uniform(45, 85, 5000)
    x, y = _tr(LL, PROJS[name], lon, lat)
    lon2, lat2 = _tr(PROJS[name], LL, x, y)
    tol = 2e-8 if name == "laea_eu" else 1e-9
    assert np.abs((lon2 - lon + 180) % 360 - 180).max() < tol and np.abs(lat2 - lat).max() < tol


def test_published_coordinates():
    x, y = _tr(LL_WGS, PROJS["utm33"], [15.0], [60.0])
    assert abs(x[0] - 500000.0) < 1e-6 and abs(y[0] - 0.9996 * 6654072.819) < 2e-3
    # polar stereographic on the sphere, true at 60N: rho = R (1 + sin 60) tan(45 - lat/2)
    R = 6371000.0
    x, y = _tr(LL, "+proj=stere +lat_0=90 +lon_0=0 +lat_ts=60 +R=6371000", [0.0, 90.0], [60.0, 60.0])
    rho = R * (1 + np.sin(np.radians(60))) * np.tan(np.radians(45 - 30))
    assert abs(x[0]) < 1e-6 and abs(y[0] + rho) < 1e-6 and abs(x[1] - rho) < 1e-6 and abs(y[1]) < 1e-6
    # spherical Mercator: y = R ln tan(45 + lat/2)
    x, y = _tr(LL, "+proj=merc +R=6371000", [10.0], [45.0])
    assert abs(x[0] - R * np.radians(10)) < 1e-6 and abs(y[0] - R * np.log(np.tan(np.radians(67.5)))) < 1e-6
    # rotated pole: the rotated north pole maps to rotated latitude 90
    x, y = _tr(LL, PROJS["hirlam"], [-40.0 + 180.0], [22.0])
    assert abs(y[0] - 90.0) < 1e-9
    # unprojectable point -> +inf (PROJ's HUGE_VAL), invalid definition -> error
    x, y = _tr(LL, "+proj=merc +R=6371000", [0.0], [90.0])
    assert np.isinf(x[0]) and np.isinf(y[0])
    err = C.create_string_buffer(200)
    xv = np.zeros(1)
    assert ol.oracle().miniproj_transform_str(b"+proj=fork +lat_1000=123", LL.encode(), 1, dp(xv), dp(xv.copy()), err, 200) == -1


def test_reference_emep_regrid_cell(ref):
    """test/testInterpolation.cc:274-385 through the reference's own reproject_f: EMEP 170x150 polar stereographic -> 0.5 deg
    lat/lon, cell (9, 77) must be country 33 for nearest, bilinear and bicubic"""
    path = "/root/reference/test/inData.txt"
    if not os.path.exists(path):
        pytest.skip("reference test data not present")
    iS, jS, lonS, latS = 170, 150, 180, 90
    arr = np.full((1, jS, iS), np.nan, np.float32)
    d = np.loadtxt(path)
    arr[0, d[:, 1].astype(int) - 1, d[:, 0].astype(int) - 1] = d[:, 2]
    assert arr[0, 50, 93] == 4
    emep = b"+proj=stere +ellps=WGS84 +lat_0=90 +lon_0=-32 +lat_ts=60 +x_0=7 +y_0=109"
    latlong = b"+proj=latlong +ellps=WGS84"
    iax = (np.arange(iS) + 1 - iS / 2.0) * 50000.0
    jax = (np.arange(jS) + 1 - jS / 2.0) * 50000.0
    lon = (np.arange(lonS) + 1) / 2.0 - 30
    lat = (np.arange(latS) + 1) / 2.0 + 30
    for method in (ol.NEAREST, ol.BILINEAR, ol.BICUBIC):
        out = np.full((1, latS, lonS), np.nan, np.float32)
        assert ref.ref_reproject_f(method, emep, fp(arr), dp(iax), dp(jax), 0, 0, iS, jS, 1, latlong, fp(out), dp(lon), dp(lat), 1, 2, lonS, latS) == 0
        assert abs(out[0, 77, 9] - 33) < 1e-5, method


def test_reference_vector_rotation_90_and_180(ref):
    """test/testInterpolation.cc:387-470: rotating the polar stereographic grid by 90 / 180 degrees turns (u,v) into (v,-u) / (-u,-v)"""
    a = b"+proj=stere +lat_0=90 +lon_0=0 +lat_ts=60 +ellps=sphere +a=6371000 +e=0"
    ax = np.arange(-5, 6) * 50000.0
    ay = np.arange(-4, 5) * 50000.0 + 1e6
    for lon0, want in ((90, lambda u, v: (v, -u)), (180, lambda u, v: (-u, -v))):
        b = f"+proj=stere +lat_0=90 +lon_0={lon0} +lat_ts=60 +ellps=sphere +a=6371000 +e=0".encode()
        m = np.empty(ax.size * ay.size * 3)
        assert ref.ref_get_vector_reproject_matrix(a, b, dp(ax), dp(ay), ax.size, ay.size, dp(m)) == 0
        u = np.full((1, ay.size, ax.size), 3.0, np.float32)
        v = np.full((1, ay.size, ax.size), -2.0, np.float32)
        ref.ref_vector_reproject_values_by_matrix_f(dp(m), ax.size, ay.size, fp(u), fp(v), 1)
        wu, wv = want(3.0, -2.0)
        assert np.abs(u - wu).max() < 1e-4 and np.abs(v - wv).max() < 1e-4
