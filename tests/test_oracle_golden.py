"""Pins oracle/fimex_oracle.c on the golden vectors of the reference's own unit tests
(test/testInterpolation.cc, test/testUtils.cc of fimex 2.4.1)."""
import ctypes as C
import math

import numpy as np
import pytest

from oracle_lib import PROJ_AXIS, LONGITUDE, fp, dp, o_points2position

NAN = float("nan")


def _point(fn, field, x, y):
    f = np.ascontiguousarray(field, dtype=np.float32)
    iz, iy, ix = f.shape
    out = np.zeros(iz, dtype=np.float32)
    fn(fp(f), fp(out), x, y, ix, iy, iz)
    return out


def _two(fn, a, b, xa, xb, x):
    a = np.atleast_1d(np.asarray(a, dtype=np.float32))
    b = np.atleast_1d(np.asarray(b, dtype=np.float32))
    out = np.full(a.shape, 12345, dtype=np.float32)
    rc = fn(fp(a), fp(b), fp(out), a.size, xa, xb, x)
    return out, rc


def test_points2position_goldens():
    # testInterpolation.cc:55-75
    ax = np.array([1., 2., 3., 4., 5.])
    np.testing.assert_allclose(o_points2position([-3., 5., 1.3, 2., 6.], ax, PROJ_AXIS), [-4., 4., 0.3, 1., 5.], atol=1e-10)
    np.testing.assert_allclose(o_points2position([-3., 5., 1.3, 2., 6.], ax[::-1].copy(), PROJ_AXIS), [8., 0., 3.7, 3., -1.], atol=1e-10)


def test_points2position_circular_longitude():
    # 0..359 axis: -0.2 and 359.8 both -> -0.2, 400 -> 40; one-sided normalisation on -10..30
    ax = np.arange(0., 360., 1.)
    r = o_points2position([-0.2, 359.8, 400., float("inf"), NAN], ax, LONGITUDE)
    np.testing.assert_allclose(r[:3], [-0.2, -0.2, 40.], atol=1e-9)
    assert r[3] == -999. and r[4] == -999.
    ax2 = np.arange(-10., 31., 1.)
    r2 = o_points2position([370., -350.], ax2, LONGITUDE)
    np.testing.assert_allclose(r2, [20., -340.], atol=1e-9)


def test_bsearch(oracle):
    # testInterpolation.cc:698-704
    v = np.array([1, 3, 5, 7, 9, 13, 17, 21, 24, 28], dtype=np.float64)
    assert oracle.orc_bsearch_double_index(7., dp(v), 10, 1) == 3
    assert oracle.orc_bsearch_double_index(8., dp(v), 10, 1) == -5


def test_bilinear_goldens(oracle):
    # testInterpolation.cc:77-110
    f = np.array([[[1., 2.], [2., 1 + np.sqrt(np.float32(2.0))]]], dtype=np.float32)
    fn = oracle.orc_get_values_bilinear_f
    assert abs(_point(fn, f, 0.3, 0.)[0] - 1.3) < 1e-6
    assert abs(_point(fn, f, 0.3, 0.0001)[0] - 1.3) < 1e-4
    assert abs(_point(fn, f, 0., 0.3)[0] - 1.3) < 1e-6
    assert abs(_point(fn, f, 0.0001, 0.3)[0] - 1.3) < 1e-4
    assert not math.isnan(_point(fn, f, 0, 0)[0])
    assert not math.isnan(_point(fn, f, 1, 1)[0])
    for x, y in [(1.5, 0.5), (0.5, 1.5), (0.5, -0.5), (-0.5, 0.5)]:
        assert math.isnan(_point(fn, f, x, y)[0])


def test_bilinear_edge_rules(oracle):
    # SURVEY 8a, verified there against the compiled reference on g[y][x] = 3y + x
    g = (3 * np.arange(3)[:, None] + np.arange(3)[None, :]).astype(np.float32)[None]
    fn = oracle.orc_get_values_bilinear_f
    assert _point(fn, g, -0.4, 1)[0] == 3
    assert math.isnan(_point(fn, g, -0.6, 1)[0])
    assert _point(fn, g, 1, 2.4)[0] == 7
    assert math.isnan(_point(fn, g, 1, 2.6)[0])
    assert _point(fn, g, 2.4, 2.4)[0] == 8
    assert _point(fn, g, 2, 1.5)[0] == 6.5


def test_bicubic_goldens(oracle):
    # testInterpolation.cc:112-152
    f = np.array([[1, 1, 1, 1], [2, 2, 2, 2], [2, 2, 2, 2], [1, 1, 1, 1]], dtype=np.float32)[None]
    ft = np.ascontiguousarray(f.transpose(0, 2, 1))
    fn = oracle.orc_get_values_bicubic_f
    for fld, pts in ((f, [((1, 1), 2.), ((1, 1.99999), 2.), ((1, 1.5), 2.125), ((1.5, 1), 2.)]),
                     (ft, [((1, 1), 2.), ((1.99999, 1), 2.), ((1.5, 1), 2.125), ((1, 1.5), 2.)])):
        for (x, y), want in pts:
            assert _point(fn, fld, x, y)[0] == pytest.approx(want, rel=1e-3)
    for x, y in [(.5, 1), (1, .5), (2.5, 1), (1, 2.5)]:
        assert math.isnan(_point(fn, f, x, y)[0])


def test_linear_goldens(oracle):
    # testInterpolation.cc:154-177
    a, b = [0, 1, -1, 1], [1, -1, 0, 1]
    out, _ = _two(oracle.orc_get_values_linear_f, a, b, 1., 1., .5)
    np.testing.assert_allclose(out, a, rtol=1e-5)
    out, _ = _two(oracle.orc_get_values_linear_f, a, b, 1., 2., 1.5)
    np.testing.assert_allclose(out, 0.5 * (np.array(a) + np.array(b)), rtol=1e-5)
    out, _ = _two(oracle.orc_get_values_linear_f, a, b, 0., 1., 2.)
    np.testing.assert_allclose(out, np.array(a) + 2. * (np.array(b) - np.array(a)), rtol=1e-5)


@pytest.mark.parametrize("name,want", [
    # testInterpolation.cc:630-696; a=2, b=3, in={200,300}, x = 0.5, 1.5, 2.5, 3.5, 4.5
    ("linear_no_extrapol", [NAN, NAN, 250, NAN, NAN]),
    ("linear_const_extrapol", [200, 200, 250, 300, 300]),
    ("linear_weak_extrapol", [NAN, 150, 250, 350, NAN]),
    ("linear", [50, 150, 250, 350, 450]),
])
def test_extrapolation_flavours(oracle, name, want):
    fn = getattr(oracle, f"orc_get_values_{name}_f")
    for x, w in zip([0.5, 1.5, 2.5, 3.5, 4.5], want):
        out, _ = _two(fn, 200, 300, 2., 3., x)
        if math.isnan(w):
            assert math.isnan(out[0]), (name, x)
        else:
            assert abs(out[0] - w) < 0.01, (name, x)


def test_log_goldens(oracle):
    # testInterpolation.cc:205-254 (NCL vintp2p_ecmwf numbers), tol 1e-3 relative
    for x, w in [(100., 100.), (1000., 1000.), (500., 729.073), (1500., 1158.482), (200., 370.927), (800., 912.781)]:
        out, rc = _two(oracle.orc_get_values_log_f, 1000., 100., 1000., 100., x)
        assert rc == 1 and out[0] == pytest.approx(w, rel=1e-3)
    for x, w in [(500., 763.1873), (200., 408.0904), (800., 926.384)]:
        out, rc = _two(oracle.orc_get_values_log_log_f, 1000., 100., 1000., 100., x)
        assert rc == 1 and out[0] == pytest.approx(w, rel=1e-3)
    # non-positive coordinate: MIFI_ERROR and the output untouched (interpolation.c:524-526)
    out, rc = _two(oracle.orc_get_values_log_f, 1., 2., -1., 100., 5.)
    assert rc == -1 and out[0] == 12345


def _pair(fn, ary, x):
    a = np.asarray(ary, dtype=np.float32)
    p0, p1 = C.c_size_t(), C.c_size_t()
    fn(fp(a), a.size, 1, x, C.byref(p0), C.byref(p1))
    return p0.value, p1.value


def test_neighbor_search_goldens(oracle):
    # testUtils.cc:130-173
    ary = [1, 2, 3, 4, 4, -1, -2, 5]
    f = oracle.orc_find_closest_distinct_elements
    assert _pair(f, ary, 1.5) == (0, 1)
    assert _pair(f, ary, 4) == (3, 7)
    assert _pair(f, ary, -1) == (5, 6)
    assert _pair(f, ary, -3) == (6, 5)
    assert _pair(f, ary, 5) == (7, 3)
    g = oracle.orc_find_closest_neighbor_distinct_elements
    assert _pair(g, ary, 1.5) == (0, 1)
    assert _pair(g, ary, 4) == (3, 7)
    assert _pair(g, ary, -1) == (5, 0)
    assert _pair(g, ary, -3) == (6, 5)
    assert _pair(g, ary, 5) == (7, 3)


def test_neighbor_search_first_element_quirk(oracle):
    # SURVEY quirk 6, verified against the compiled header: column {10,30,50,70,100}
    col = [10, 30, 50, 70, 100]
    g = oracle.orc_find_closest_neighbor_distinct_elements
    assert _pair(g, col, 5) == (0, 0)
    assert _pair(g, col, 10) == (0, 0)
    assert _pair(g, col, 120) == (4, 3)


def test_method_names(oracle):
    # interpolation.c:72-107 incl. the forward_undef_min quirk (:103-104)
    assert oracle.orc_string_to_interpolation_method(b"bilinear") == 1
    assert oracle.orc_string_to_interpolation_method(b"coord_kdtree") == 4
    assert oracle.orc_string_to_interpolation_method(b"forward_undef_max") == 13
    assert oracle.orc_string_to_interpolation_method(b"forward_undef_min") == 9
    assert oracle.orc_string_to_interpolation_method(b"nope") == -1
    from fimex_b200.constants import mifi_string_to_interpolation_method as host
    for n in (b"bilinear", b"nearestneighbor", b"bicubic", b"coord_nearestneighbor", b"coord_kdtree", b"forward_sum", b"forward_mean",
              b"forward_median", b"forward_max", b"forward_min", b"forward_undef_sum", b"forward_undef_mean", b"forward_undef_median",
              b"forward_undef_max", b"forward_undef_min", b"bogus"):
        assert host(n.decode()) == oracle.orc_string_to_interpolation_method(n)


def test_hybrid_sigma_pressure_formula(oracle):
    # vertical_coordinate_transformations.c:62-68
    ap = np.array([0., 2000., 5000.])
    b = np.array([1., 0.5, 0.])
    p = np.zeros(3)
    oracle.orc_atmosphere_hybrid_sigma_ap_pressure(3, 101325., dp(ap), dp(b), dp(p))
    np.testing.assert_array_equal(p, ap + b * 101325.)


# ---------------------------------------------------------------------------------- the reference's own output files (vertical)
def _vertical_golden():
    import os
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vertical.npz"))


def cdmcmp_close(a, b, tol=1e-3):
    """test/cdmcmp.cc:87-164: |a-b| <= tol*|a| and <= tol*|b| (or equal)"""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    d = np.abs(a - b)
    return (a == b) | ((d <= tol * np.abs(a)) & (d <= tol * np.abs(b)))


def test_vertical_oracle_reproduces_fimex_output_sigma_hybrid(oracle):
    """verticalSigmaHybrid.nc -> verticalSigmaHybridPressure.nc (testVerticalInterpolatorSigmaHybridToPressure.sh)"""
    g = _vertical_golden()
    ap, b, lev = g["sh_ap_pa"] * 0.01, g["sh_b"], g["sh_levels"]
    ps = g["sh_ps_raw"].astype(np.float64) * float(g["sh_ps_scale"])  # hPa
    bad = np.float32(9.96921e36)
    for name in ("air_potential_temperature", "specific_humidity"):
        v, e = g["sh_in_" + name], g["sh_ex_" + name]
        nt, nk, ny, nx = v.shape
        for t in range(nt):
            fld = v[t].reshape(nk, ny * nx).copy()
            oracle.orc_bad2nanf(fp(fld), fld.size, bad)
            p = np.ascontiguousarray((ap[:, None] + b[:, None] * ps[t].reshape(1, -1)).astype(np.float32))
            out = np.empty((lev.size, ny * nx), np.float32)
            assert oracle.orc_vertical_interpolate_f(1, fp(fld), fp(p), nk, ny * nx, dp(lev), None, lev.size, None, None, NAN, NAN, fp(out)) == 1
            oracle.orc_nanf2bad(fp(out), out.size, bad)
            assert cdmcmp_close(out, e[t].reshape(lev.size, -1)).all(), (name, t)


def test_vertical_oracle_reproduces_fimex_output_ensemble(oracle):
    """testdata_vertical_ensemble_in.nc -> _ex.nc: (time, hybrid=65, ensemble=3, y, x), packed int32 variables whose raw
    numbers are interpolated as f32 and written back with lround (DataUtils.h:55-58)."""
    g = _vertical_golden()
    ap, b, lev = g["en_ap_pa"] * 0.01, g["en_b"], g["en_levels"]
    ps = g["en_ps_pa"].astype(np.float64) * 0.01  # Pa -> hPa
    for name in ("specific_humidity_ml", "x_wind_ml", "y_wind_ml", "upward_air_velocity_ml"):
        v, e = g["en_in_" + name], g["en_ex_" + name]
        nt, nk = v.shape[:2]
        ncol = int(np.prod(v.shape[2:]))
        for t in range(nt):
            fld = v[t].reshape(nk, ncol).astype(np.float32)
            oracle.orc_bad2nanf(fp(fld), fld.size, np.float32(-2147483647))
            p = np.ascontiguousarray((ap[:, None] + b[:, None] * ps[t].reshape(1, ncol)).astype(np.float32))
            out = np.empty((lev.size, ncol), np.float32)
            oracle.orc_vertical_interpolate_f(1, fp(fld), fp(p), nk, ncol, dp(lev), None, lev.size, None, None, NAN, NAN, fp(out))
            got = np.where(np.isnan(out), -2147483647, np.sign(out) * np.floor(np.abs(out.astype(np.float64)) + 0.5)).astype(np.int64)
            assert np.array_equal(got, e[t].reshape(lev.size, ncol).astype(np.int64)), (name, t)
