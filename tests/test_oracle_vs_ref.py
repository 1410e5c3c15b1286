"""Pins the C restatement (oracle/liboracle.so) against the reference's OWN code compiled from
/root/reference into oracle/_ref (seeded random inputs, bit-exact).  Skipped when oracle/_ref is absent."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import dp, fp


def _eq(a, b):
    """bit-level equality that treats any NaN as equal to any NaN"""
    a, b = np.asarray(a), np.asarray(b)
    na, nb = np.isnan(a), np.isnan(b)
    return np.array_equal(na, nb) and np.array_equal(a[~na], b[~nb])


def test_points2position_matches_reference(oracle, ref):
    rng = np.random.default_rng(11)
    axes = [np.linspace(-1060084., 1309916., 949), np.linspace(75.25, 49.75, 1021), np.arange(0., 360., 0.5), np.arange(-180., 180., 0.25),
            np.cumsum(rng.uniform(0.5, 2.0, 200)), np.arange(-10., 31., 1.)]
    for k, ax in enumerate(axes):
        for t in (ol.PROJ_AXIS, ol.LONGITUDE, ol.LATITUDE):
            lo, hi = ax.min(), ax.max()
            pts = rng.uniform(lo - 0.3 * (hi - lo), hi + 0.3 * (hi - lo), 5000)
            pts[:200] = rng.choice(ax, 200)  # exact hits
            pts[200:210] = [np.nan, np.inf, -np.inf, 360., -360., 180., -180., 0., 720., -0.0]
            a, b = pts.copy(), pts.copy()
            oracle.orc_points2position(dp(a), a.size, dp(ax), ax.size, t)
            ref.mifi_points2position(dp(b), b.size, dp(ax), ax.size, t)
            assert _eq(a, b), (k, t)


@pytest.mark.parametrize("name", ["f", "bilinear_f", "bicubic_f"])
def test_point_kernels_match_reference(oracle, ref, name):
    rng = np.random.default_rng(5)
    ix, iy, iz = 23, 17, 4
    fld = rng.normal(size=(iz, iy, ix)).astype(np.float32)
    fld[rng.random(fld.shape) < 0.05] = np.nan
    fo, fr = getattr(oracle, "orc_get_values_" + name), getattr(ref, "mifi_get_values_" + name)
    xs = rng.uniform(-2, ix + 1, 20000)
    ys = rng.uniform(-2, iy + 1, 20000)
    xs[:4000] = np.round(xs[:4000] * 2) / 2  # half-integers and integers: the rounding edge cases
    ys[:4000] = np.round(ys[:4000] * 2) / 2
    a, b = np.zeros(iz, np.float32), np.zeros(iz, np.float32)
    for x, y in zip(xs, ys):
        if name == "bilinear_f" and (iy - 0.5 <= y < iy + 0.5) and not (0 <= np.floor(x) and np.floor(x) + 1 < ix):
            continue  # the reference reads out of bounds there (interpolation.c:322); defined as NaN in oracle and CUDA
        fo(fp(fld), fp(a), x, y, ix, iy, iz)
        fr(fp(fld), fp(b), x, y, ix, iy, iz)
        assert _eq(a, b), (x, y)


@pytest.mark.parametrize("name", ["nearest", "linear", "linear_weak_extrapol", "linear_no_extrapol", "linear_const_extrapol", "log", "log_log"])
def test_two_level_kernels_match_reference(oracle, ref, name):
    rng = np.random.default_rng(7)
    fo, fr = getattr(oracle, f"orc_get_values_{name}_f"), getattr(ref, f"mifi_get_values_{name}_f")
    n = 64
    for _ in range(300):
        A = rng.normal(size=n).astype(np.float32)
        B = rng.normal(size=n).astype(np.float32)
        A[rng.random(n) < 0.1] = np.nan
        xa, xb = rng.uniform(1, 1000, 2)
        x = rng.choice([xa, xb, rng.uniform(-100, 1500), (xa + xb) / 2])
        if rng.random() < 0.05:
            xb = xa
        o1 = np.full(n, 777, np.float32)
        o2 = np.full(n, 777, np.float32)
        r1 = fo(fp(A), fp(B), fp(o1), n, xa, xb, x)
        r2 = fr(fp(A), fp(B), fp(o2), n, xa, xb, x)
        assert r1 == r2 and _eq(o1, o2)


def _tables(rng, in_x, in_y, out_x, out_y, margin=3.0):
    """index tables with points inside, on the borders and outside"""
    px = rng.uniform(-margin, in_x - 1 + margin, out_x * out_y)
    py = rng.uniform(-margin, in_y - 1 + margin, out_x * out_y)
    k = out_x * out_y // 10
    px[:k] = np.round(px[:k] * 2) / 2
    py[:k] = np.round(py[:k] * 2) / 2
    px[k:k + 5] = -999.
    py[k:k + 5] = -999.
    # keep clear of the reference's out-of-bounds corner strip (interpolation.c:322)
    strip = (py >= in_y - 0.5) & (py < in_y + 0.5) & ~((np.floor(px) >= 0) & (np.floor(px) + 1 < in_x))
    py[strip] += 1.0
    return px, py


@pytest.mark.parametrize("method", [ol.BILINEAR, ol.BICUBIC, ol.NEAREST])
@pytest.mark.parametrize("window", ["full", "reduced"])
def test_cached_apply_matches_reference(oracle, ref, method, window):
    rng = np.random.default_rng(100 + method)
    in_x, in_y, out_x, out_y, nz = 61, 47, 53, 39, 3
    px, py = _tables(rng, in_x, in_y, out_x, out_y)
    if window == "reduced":  # targets cover only part of the source -> reduced domain kicks in
        px = 10 + px * 0.4
        py = 8 + py * 0.5
    fld = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    fld[rng.random(fld.shape) < 0.03] = np.nan
    o, oi = ol.o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, fld)
    r, ri = ol.r_cached_apply(method, px, py, in_x, in_y, out_x, out_y, fld)
    assert oi == ri
    if window == "reduced":
        assert oi["reduced"] and (oi["inX"] < in_x and oi["inY"] < in_y)
    # (the bilinear/bicubic cache always records a "reduced" domain, even one that spans the whole grid,
    #  CachedInterpolation.cc:236-252)
    assert _eq(o, r)


@pytest.mark.parametrize("method", [ol.FWD_SUM, ol.FWD_MEAN, ol.FWD_MEDIAN, ol.FWD_MAX, ol.FWD_MIN, ol.FWD_UNDEF_SUM, ol.FWD_UNDEF_MEAN,
                                    ol.FWD_UNDEF_MAX, ol.FWD_UNDEF_MIN])
def test_forward_matches_reference(oracle, ref, method):
    rng = np.random.default_rng(200 + method)
    in_x, in_y, out_x, out_y, nz = 83, 59, 21, 17, 2
    # every source cell gets a target position; several sources per target cell, some outside
    pox = rng.uniform(-2, out_x + 1, in_x * in_y)
    poy = rng.uniform(-2, out_y + 1, in_x * in_y)
    fld = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    fld[rng.random(fld.shape) < 0.1] = np.nan
    o, oi = ol.o_cached_apply(method, pox, poy, in_x, in_y, out_x, out_y, fld, level_bug=1)
    r, ri = ol.r_cached_apply(method, pox, poy, in_x, in_y, out_x, out_y, fld)
    assert oi == ri
    assert _eq(o, r)
    # the reference aggregates input level 0 into every output level
    assert _eq(r[0], r[1])


def test_forward_reduced_window_matches_reference(oracle, ref):
    rng = np.random.default_rng(31)
    in_x, in_y, out_x, out_y = 90, 70, 12, 10
    pox = rng.uniform(-40, out_x + 40, in_x * in_y)  # most sources fall outside the target
    poy = rng.uniform(-40, out_y + 40, in_x * in_y)
    gx, gy = np.meshgrid(np.arange(in_x), np.arange(in_y))
    keep = ((gx > 20) & (gx < 60) & (gy > 15) & (gy < 50)).ravel()
    pox[~keep] = -500.
    fld = rng.normal(size=(1, in_y, in_x)).astype(np.float32)
    o, oi = ol.o_cached_apply(ol.FWD_MEAN, pox, poy, in_x, in_y, out_x, out_y, fld)
    r, ri = ol.r_cached_apply(ol.FWD_MEAN, pox, poy, in_x, in_y, out_x, out_y, fld)
    assert oi == ri and oi["reduced"]
    assert _eq(o, r)


def test_round_and_clamp_matches_reference(oracle, ref):
    vals = [-1.0, -0.5, -0.49999, -0.0, 0.0, 0.49999, 0.5, 1.5, 2.5, 9.4999, 9.5, 10.0, 1e12, -1e12]  # NaN: undefined behaviour in the reference (double -> size_t)
    for v in vals:
        assert oracle.orc_round_and_clamp(v, 0, 9, ol.INVALID_IDX) == ref.ref_round_and_clamp(v, 0, 9, ol.INVALID_IDX), v


def test_neighbor_search_matches_reference(oracle, ref):
    rng = np.random.default_rng(3)
    for _ in range(2000):
        n = int(rng.integers(1, 12))
        col = rng.choice([1., 2., 3., 5., 8., 13., -4., 0.5], n).astype(np.float32) if rng.random() < 0.5 else np.sort(
            rng.uniform(0, 1000, n)).astype(np.float32)
        if rng.random() < 0.5:
            col = col[::-1].copy()
        x = float(rng.choice(np.concatenate([col, rng.uniform(-10, 1100, 3).astype(np.float32)])))
        for fo, fr in ((oracle.orc_find_closest_neighbor_distinct_elements, ref.ref_find_closest_neighbor_distinct_elements),
                       (oracle.orc_find_closest_distinct_elements, ref.ref_find_closest_distinct_elements)):
            a0, a1 = C.c_size_t(), C.c_size_t()
            b0, b1 = C.c_long(), C.c_long()
            fo(fp(col), n, 1, x, C.byref(a0), C.byref(a1))
            fr(fp(col), n, 1, x, C.byref(b0), C.byref(b1))
            assert (a0.value, a1.value) == (b0.value, b1.value), (col, x)


def test_vector_rotation_matches_reference(oracle, ref):
    rng = np.random.default_rng(9)
    ox, oy, nz = 31, 17, 3
    ang = rng.uniform(-np.pi, np.pi, ox * oy)
    m = np.stack([np.cos(ang), np.sin(ang), np.degrees(ang)], axis=1).copy()
    u = rng.normal(size=(nz, oy, ox)).astype(np.float32)
    v = rng.normal(size=(nz, oy, ox)).astype(np.float32)
    u[0, 0, :3] = np.nan
    u1, v1, u2, v2 = u.copy(), v.copy(), u.copy(), v.copy()
    oracle.orc_vector_reproject_values_by_matrix_f(dp(m), ox * oy, fp(u1), fp(v1), nz)
    ref.ref_vector_reproject_values_by_matrix_f(dp(m), ox, oy, fp(u2), fp(v2), nz)
    assert _eq(u1, u2) and _eq(v1, v2)
    d = rng.uniform(0, 360, size=(nz, oy, ox)).astype(np.float32)
    d1, d2 = d.copy(), d.copy()
    oracle.orc_vector_reproject_direction_by_matrix_f(dp(m), ox * oy, fp(d1), nz)
    ref.ref_vector_reproject_direction_by_matrix_f(dp(m), ox, oy, fp(d2), nz)
    assert _eq(d1, d2)


def test_reproject_matrix_restatement_matches_reference_logic(oracle, ref):
    """reproject.cc's matrix logic (compiled unmodified over the proj.h shim) against its restatement."""
    lcc = b"+proj=lcc +lat_0=63.3 +lon_0=15 +lat_1=63.3 +lat_2=63.3 +no_defs +R=6.371e+06"
    stere = b"+proj=stere +lat_0=90 +lon_0=0 +lat_ts=60 +R=6.371e+06 +no_defs"
    ll = b"+proj=latlong +R=6.371e+06 +no_defs"
    lon = np.arange(-10., 40., 2.5)
    lat = np.arange(52., 74., 2.)
    xs = np.arange(-8e5, 8e5, 1e5)
    ys = np.arange(-3.2e6, -1.6e6, 1e5)
    for src, dst, ax, ay in ((lcc, ll, lon, lat), (lcc, stere, xs, ys), (ll, stere, xs, ys)):
        a = np.zeros(ax.size * ay.size * 3)
        b = np.zeros_like(a)
        assert oracle.orc_get_vector_reproject_matrix(src, dst, dp(ax), dp(ay), ax.size, ay.size, dp(a)) == 0
        assert ref.ref_get_vector_reproject_matrix(src, dst, dp(ax), dp(ay), ax.size, ay.size, dp(b)) == 0
        assert _eq(a, b)
    ox_, oy_ = np.zeros(lon.size * lat.size), np.zeros(lon.size * lat.size)
    rx_, ry_ = np.zeros_like(ox_), np.zeros_like(oy_)
    oracle.orc_reproject_axes(ll, lcc, dp(lon), dp(lat), lon.size, lat.size, dp(ox_), dp(oy_))
    ref.ref_reproject_axes(ll, lcc, dp(lon), dp(lat), lon.size, lat.size, dp(rx_), dp(ry_))
    assert _eq(ox_, rx_) and _eq(oy_, ry_)


def test_packed_short_conversions_match_reference(oracle, ref):
    """the conversions CDMInterpolator::getDataSlice wraps around the interpolation for a CDM_SHORT variable, and the scaled read of
    the vertical converters: the restatement against the reference's own header templates (ScaleValue, data_caster)"""
    rng = np.random.default_rng(5)
    n = 200000
    for bad in (-32767.0, 9999.0, 1.0e10, -32768.0):
        raw = rng.integers(-32768, 32768, n).astype(np.int16)
        raw[::17] = np.int16(bad) if abs(bad) < 32768 else 123
        a, b = np.empty(n, np.float32), np.empty(n, np.float32)
        oracle.orc_i16_to_interpolation_array(raw.ctypes.data, n, bad, fp(a))
        ref.ref_i16_to_interpolation_array(raw.ctypes.data, n, bad, fp(b))
        assert np.array_equal(a, b, equal_nan=True)
        v = rng.uniform(-40000, 40000, n).astype(np.float32)
        v[:5000] = np.round(v[:5000]) + np.float32(0.5)           # ties: half away from zero
        v[5000:6000] = np.float32(0.49999997) * rng.choice([-1, 1], 1000)
        v[6000:7000] = rng.uniform(-3e9, 3e9, 1000).astype(np.float32)  # beyond short and int: the C casts wrap
        v[7000:8000] = np.nan
        v[8000] = -0.0
        c, d = np.empty(n, np.int16), np.empty(n, np.int16)
        oracle.orc_interpolation_array_to_i16(fp(v), n, bad, c.ctypes.data)
        ref.ref_interpolation_array_to_i16(fp(v), n, bad, d.ctypes.data)
        assert np.array_equal(c, d)
    for fill, scale, offset in ((-32767.0, 0.01, 273.15), (9999.0, 1.0, 0.0), (-32768.0, 1e-3, -5.0)):
        raw = rng.integers(-32768, 32768, n).astype(np.int16)
        raw[::13] = np.int16(fill)
        a, b = np.empty(n, np.float32), np.empty(n, np.float32)
        oracle.orc_i16_scaled_to_float(raw.ctypes.data, n, fill, scale, offset, fp(a))
        ref.ref_i16_scaled_to_float(raw.ctypes.data, n, fill, scale, offset, fp(b))
        assert np.array_equal(a, b, equal_nan=True)
        assert np.isnan(a[::13]).all()


def _fpn(a):
    return None if a is None else fp(a)


def test_vertical_converters_match_reference_functions(oracle, ref):
    """the converter restatement against the reference's own C functions (src/vertical_coordinate_transformations.c, compiled
    into oracle/_ref), composed per column the way the converter classes call them"""
    rng = np.random.default_rng(8)
    nz, nI = 23, 57
    a = np.sort(rng.uniform(0, 300, nz))
    b = np.sort(rng.uniform(0, 1, nz))
    ps = rng.uniform(60000, 104000, (2, nI)).astype(np.float32)
    orog = rng.uniform(0, 2500, (1, nI)).astype(np.float32)
    depth = rng.uniform(5, 4000, (1, nI)).astype(np.float32)
    eta = rng.normal(0, 0.5, (1, nI)).astype(np.float32)
    topo = rng.uniform(0, 2000, nI).astype(np.float32)
    sC = -np.sort(rng.uniform(0, 1, nz))[::-1].copy()
    cases = [
        (ol_k.HYBRID_SIGMA_AP, a, b, 0.0, ps, 0.01, None, lambda s, z: _call(ref.mifi_atmosphere_hybrid_sigma_ap_pressure, nz, s, a, b)),
        (ol_k.HYBRID_SIGMA, a / 1000, b, 1000.0, ps, 0.01, None, lambda s, z: _call(ref.mifi_atmosphere_hybrid_sigma_pressure, nz, 1000.0, s, a / 1000, b)),
        (ol_k.SIGMA, b, None, 10.0, ps, 0.01, None, lambda s, z: _call(ref.mifi_atmosphere_sigma_pressure, nz, 10.0, s, b)),
        (ol_k.LN_PRESSURE, a / 50, None, 1000.0, None, 1.0, None, lambda s, z: _call(ref.mifi_atmosphere_ln_pressure, nz, 1000.0, a / 50)),
        (ol_k.HYBRID_HEIGHT, a * 100, b, 0.0, orog, 1.0, None, lambda s, z: _call(ref.mifi_atmosphere_hybrid_height, nz, s, a * 100, b)),
        (ol_k.OCEAN_S_G1, sC, sC * 0.7, 10.0, depth, 1.0, eta, lambda s, z: -_call(ref.mifi_ocean_s_g1_z, nz, s, 10.0, z, sC, sC * 0.7)),
        (ol_k.OCEAN_S_G2, sC, sC * 0.7, 10.0, depth, 1.0, eta, lambda s, z: -_call(ref.mifi_ocean_s_g2_z, nz, s, 10.0, z, sC, sC * 0.7)),
    ]
    for kind, ka, kb, scalar, surf, scale, e, column in cases:
        for post, factor, tp in ((0, 0.0, None), (1, -1.0, topo), (2, 1.0 / 9.80665, topo)):
            nO = 2 if surf is ps else 1
            out = np.empty((nO, nz, nI), np.float32)
            kbb = ka if kb is None else kb
            rc = oracle.orc_vcoord_levels_f(kind, dp(np.ascontiguousarray(ka)), dp(np.ascontiguousarray(kbb)), scalar, _fpn(surf), nO if surf is not None else 1,
                                            scale, _fpn(e), post, _fpn(tp), factor, nO, nz, nI, fp(out))
            assert rc == 1
            want = np.empty((nO, nz, nI), np.float64)
            for o in range(nO):
                for i in range(nI):
                    s = scale * float(surf[o, i]) if surf is not None else 0.0
                    col = column(s, float(e[0, i]) if e is not None else 0.0)
                    if post == 1:
                        col = _call(ref.mifi_barometric_standard_altitude, nz, col)
                    elif post == 2:
                        col = _call(ref.mifi_barometric_standard_pressure, nz, col)
                    if tp is not None:
                        col = col + factor * float(tp[i])
                    want[o, :, i] = col
            with np.errstate(all="ignore"):
                assert np.array_equal(out, want.astype(np.float32), equal_nan=True), (kind, post)
    # building blocks of the hypsometric integration
    q, T = rng.uniform(0, 0.02, 1000).astype(np.float32), rng.uniform(200, 310, 1000).astype(np.float32)
    pl, ph = rng.uniform(100, 1050, 1000).astype(np.float32), rng.uniform(100, 1050, 1000).astype(np.float32)
    for k in range(1000):
        assert oracle.orc_virtual_temperature(q[k], T[k]) == ref.mifi_virtual_temperature(q[k], T[k])
        assert oracle.orc_barometric_layer_thickness(pl[k], ph[k], T[k]) == ref.mifi_barometric_layer_thickness(pl[k], ph[k], T[k])


def _call(fn, n, *args):
    """call a mifi_* column function: scalars stay, arrays become double pointers, the last argument is the output"""
    out = np.empty(n, np.float64)
    conv = [dp(np.ascontiguousarray(x, dtype=np.float64)) if isinstance(x, np.ndarray) else x for x in args]
    fn(n, *conv, dp(out))
    return out


class ol_k:
    HYBRID_SIGMA_AP, HYBRID_SIGMA, SIGMA, LN_PRESSURE, HYBRID_HEIGHT, OCEAN_S_G1, OCEAN_S_G2 = range(7)


def test_pressure_integration_restatement_against_python_loop(oracle, ref):
    """PressureIntegrationToAltitudeConverter.cc:225-246 replayed in Python over the reference's two float functions"""
    rng = np.random.default_rng(4)
    nz, nI = 17, 29
    for start_high in (1, 0, -1):
        eta = np.linspace(1.0, 0.02, nz) if start_high != 0 else np.linspace(0.02, 1.0, nz)
        sap = rng.uniform(700, 1040, (1, nI)).astype(np.float32)
        p = (eta[:, None] * sap[0][None, :]).astype(np.float32)[None] * np.float32(0.999)
        sgp = rng.uniform(0, 20000, (1, nI)).astype(np.float32)
        T = rng.uniform(210, 300, (1, nz, nI)).astype(np.float32)
        q = rng.uniform(0, 0.02, (1, nz, nI)).astype(np.float32)
        for hum in (q, None):
            out = np.empty((1, nz, nI), np.float32)
            assert oracle.orc_vcoord_pressure_integration_f(fp(p), fp(sap), 1.0, fp(sgp), 1.0, fp(T), _fpn(hum), 1, 1, nz, nI, start_high, None, 0.0,
                                                            fp(out)) == 1
            want = np.empty((nz, nI), np.float64)
            high = bool(p[0, 0, 0] > p[0, -1, 0])
            for i in range(nI):
                acc = float(sgp[0, i]) / 9.80665
                p_low = sap[0, i]
                for n in range(nz):
                    iz = n if high else nz - 1 - n
                    tv = T[0, iz, i] if hum is None else ref.mifi_virtual_temperature(q[0, iz, i], T[0, iz, i])
                    acc += float(ref.mifi_barometric_layer_thickness(p_low, p[0, iz, i], tv))
                    want[iz, i] = acc
                    p_low = p[0, iz, i]
            assert np.array_equal(out[0], want.astype(np.float32))


@pytest.mark.parametrize("dtype", sorted(ol.PACKED_DTYPES))
def test_packed_conversions_all_types_match_reference(oracle, ref, dtype):
    """data2InterpolationArray / getScaledDataSlice / interpolationArray2Data for every CDM integer type, against the reference's
    header templates instantiated for that type"""
    rng = np.random.default_rng(dtype)
    T = ol.PACKED_DTYPES[dtype]
    info = np.iinfo(T)
    n = 100000
    raw = rng.integers(info.min, int(info.max) + 1, n, dtype=np.int64).astype(T)
    for bad in (float(info.min), float(info.max), 17.0, 1.0e30, -1.0e10):
        fillv = np.array([bad]).astype(np.float64)
        if info.min <= bad <= info.max:
            raw[::19] = T(bad)
        a, b = np.empty(n, np.float32), np.empty(n, np.float32)
        assert oracle.orc_unpack_f(dtype, raw.ctypes.data, n, bad, fp(a)) == 1 and ref.ref_unpack_f(dtype, raw.ctypes.data, n, bad, fp(b)) == 1
        assert np.array_equal(a, b, equal_nan=True)
        if info.min <= bad <= info.max:  # casting an out-of-range double to the integer type is undefined in the reference
            assert oracle.orc_unpack_scaled_f(dtype, raw.ctypes.data, n, bad, 0.01, 273.15, fp(a)) == 1
            assert ref.ref_unpack_scaled_f(dtype, raw.ctypes.data, n, bad, 0.01, 273.15, fp(b)) == 1
            assert np.array_equal(a, b, equal_nan=True)
            span = float(info.max) - float(info.min)
            v = rng.uniform(info.min - 0.1 * span, info.max + 0.1 * span, n).astype(np.float32)
            v[:4000] = np.round(v[:4000]) + np.float32(0.5)
            v[4000:5000] = np.nan
            v[5000:5100] = rng.uniform(-2e9, 2e9, 100).astype(np.float32)
            c, d = np.empty(n, T), np.empty(n, T)
            assert oracle.orc_pack_f(dtype, fp(v), n, bad, c.ctypes.data) == 1 and ref.ref_pack_f(dtype, fp(v), n, bad, d.ctypes.data) == 1
            assert np.array_equal(c, d)


def _blob_field(rng, ny, nx, frac=0.15):
    yy, xx = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
    f = (280 + 15 * np.sin(xx / 9.0) * np.cos(yy / 7.0) + rng.normal(0, 1, (ny, nx))).astype(np.float32)
    for _ in range(max(1, int(frac * nx * ny / 60))):
        cy, cx, r = rng.integers(0, ny), rng.integers(0, nx), rng.integers(2, 7)
        f[(yy - cy) ** 2 + (xx - cx) ** 2 <= r * r] = np.nan
    return f


@pytest.mark.parametrize("nx,ny,maxloop", [(61, 47, 100), (33, 80, 7), (20, 20, 3), (90, 12, 31)])
def test_fill2d_restatement_matches_reference(oracle, ref, nx, ny, maxloop):
    rng = np.random.default_rng(nx * ny + maxloop)
    f = _blob_field(rng, ny, nx)
    f[0, :5] = np.nan  # undefined cells on the border and in a corner
    f[-1, -1] = np.nan
    a, b = f.copy(), f.copy()
    na, nb = C.c_size_t(), C.c_size_t()
    assert oracle.orc_fill2d_f(nx, ny, fp(a), 0.01, 1.6, maxloop, C.byref(na)) == 1
    ref.mifi_fill2d_f(nx, ny, fp(b), 0.01, 1.6, maxloop, C.byref(nb))
    assert na.value == nb.value == int(np.isnan(f).sum())
    assert np.array_equal(a, b) and not np.isnan(a).any()


@pytest.mark.parametrize("nx,ny,repeat,weight", [(61, 47, 20, 2), (33, 80, 1, 1), (20, 20, 5, 2), (90, 12, 20, 4), (3, 3, 3, 2)])
def test_creepfill_restatement_matches_reference(oracle, ref, nx, ny, repeat, weight):
    rng = np.random.default_rng(nx * ny + repeat)
    f = _blob_field(rng, ny, nx, frac=0.3)
    f[0, :5] = np.nan
    f[-1, -1] = np.nan
    f[ny // 2, 0] = np.nan
    f[1, 0] = np.nan
    if nx * ny == 9:
        f = rng.normal(280, 5, (3, 3)).astype(np.float32)
        f[1, 1] = f[0, 0] = f[1, 0] = np.nan
    w = bytes([weight])
    for kind in ("mean", "val"):
        a, b = f.copy(), f.copy()
        na, nb = C.c_size_t(), C.c_size_t()
        if kind == "mean":
            assert oracle.orc_creepfill2d_f(nx, ny, fp(a), repeat, w, C.byref(na)) == 1
            ref.mifi_creepfill2d_f(nx, ny, fp(b), repeat, w, C.byref(nb))
        else:
            assert oracle.orc_creepfillval2d_f(nx, ny, fp(a), 271.5, repeat, w, C.byref(na)) == 1
            ref.mifi_creepfillval2d_f(nx, ny, fp(b), 271.5, repeat, w, C.byref(nb))
        assert na.value == nb.value == int(np.isnan(f).sum())
        assert np.array_equal(a, b) and not np.isnan(a).any()
