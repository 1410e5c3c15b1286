"""GPU parity tests of the vertical path (K14/K15), the two-level kernels and the forward aggregation (K11-K13),
through the C ABI.  Checker: oracle/liboracle.so and the committed reference fixtures.
Bar: bit-exact for every method except the log / log_log weights, which go through the device's f64 log():
those values are compared within 1e-5 relative (BASELINE.json), their undefined masks bit-exactly."""
import os

import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import dp, fp

pytestmark = pytest.mark.gpu
NAN = float("nan")
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _eq(a, b):
    a, b = np.asarray(a), np.asarray(b)
    na, nb = np.isnan(a), np.isnan(b)
    return a.shape == b.shape and np.array_equal(na, nb) and np.array_equal(a[~na], b[~nb])


def _close(a, b, rtol=1e-5):
    a, b = np.asarray(a), np.asarray(b)
    na, nb = np.isnan(a), np.isnan(b)
    return a.shape == b.shape and np.array_equal(na, nb) and np.allclose(a[~na], b[~nb], rtol=rtol, atol=0)


@pytest.fixture(scope="module")
def fx():
    import fimex_b200
    assert fimex_b200.lib.fimexcu_device_count() >= 1
    return fimex_b200


def _oracle_vertical(method, fld, vc, levels=None, vout=None, vmin=None, vmax=None, valid_min=NAN, valid_max=NAN):
    """fld, vc: [nO][nzi][nI] -> [nO][nzo][nI] with the oracle (one outer slice at a time)"""
    o = ol.oracle()
    nO, nzi, nI = fld.shape
    nzo = len(levels) if levels is not None else vout.shape[1]
    out = np.empty((nO, nzo, nI), np.float32)
    lv = None if levels is None else np.ascontiguousarray(levels, dtype=np.float64)
    for k in range(nO):
        f, v = np.ascontiguousarray(fld[k]), np.ascontiguousarray(vc[k])
        vo = None if vout is None else np.ascontiguousarray(vout[k])
        mn = None if vmin is None else np.ascontiguousarray(vmin[k])
        mx = None if vmax is None else np.ascontiguousarray(vmax[k])
        res = np.empty((nzo, nI), np.float32)
        rc = o.orc_vertical_interpolate_f(method, fp(f), fp(v), nzi, nI, None if lv is None else dp(lv), None if vo is None else fp(vo), nzo,
                                          None if mn is None else dp(mn), None if mx is None else dp(mx), valid_min, valid_max, fp(res))
        assert rc == 1
        out[k] = res
    return out


def _columns(rng, nO, nzi, nI, kind):
    if kind == "pressure":  # hybrid-like, decreasing upwards, column dependent
        ps = rng.uniform(950, 1040, (nO, 1, nI))
        eta = np.linspace(0.01, 1.0, nzi)[None, :, None]
        vc = (ps * eta + 5 * (1 - eta)).astype(np.float32)
    elif kind == "reversed":
        ps = rng.uniform(950, 1040, (nO, 1, nI))
        eta = np.linspace(1.0, 0.01, nzi)[None, :, None]
        vc = (ps * eta).astype(np.float32)
    else:  # unordered with duplicates and NaN
        vc = rng.choice(np.arange(1, 40, dtype=np.float32) * 25, (nO, nzi, nI)).astype(np.float32)
        vc[rng.random(vc.shape) < 0.02] = np.nan
    fld = rng.normal(280, 20, (nO, nzi, nI)).astype(np.float32)
    fld[rng.random(fld.shape) < 0.03] = np.nan
    return fld, np.ascontiguousarray(vc)


@pytest.mark.parametrize("method", [ol.VINT_LIN, ol.VINT_LOG, ol.VINT_LOGLOG, ol.VINT_NN, ol.VINT_LIN_WEAK, ol.VINT_LIN_NO, ol.VINT_LIN_CONST])
@pytest.mark.parametrize("kind", ["pressure", "reversed", "unordered"])
def test_vertical_fixed_levels_vs_oracle(fx, method, kind):
    from fimex_b200.vertical import verticalInterpolate
    rng = np.random.default_rng(100 * method + len(kind))
    nO, nzi, nI = 3, 65, 517
    fld, vc = _columns(rng, nO, nzi, nI, kind)
    levels = [1000., 925., 850., 700., 500., 400., 300., 250., 200., 150., 100., 50., 1100., 5., float(vc[0, 0, 0]), float(vc[0, -1, 0]),
              float(vc[0, 30, 5]), float(vc[1, 1, 7]), float(vc[2, nzi - 2, 11]), 1e12, -1e12, 3.0e9, 0.0]
    levels = [v for v in levels if not np.isnan(v)]
    want = _oracle_vertical(method, fld, vc, levels=levels)
    got = verticalInterpolate(method, fld, vc, level1=levels)
    if method in (ol.VINT_LOG, ol.VINT_LOGLOG):
        assert _close(got, want)
    else:
        assert _eq(got, want)
    assert np.isnan(want).any() and (~np.isnan(want)).any()


def test_vertical_validity_clamp_bad_fill_and_template_axis(fx, oracle):
    """converter validity range (CDMVerticalInterpolator.cc:505-528), valid_min/valid_max clamp (:551-561), an output vertical
    coordinate instead of fixed levels (template / to-axis) and the fused bad/fill conversions; device-resident."""
    import torch
    from fimex_b200.vertical import verticalInterpolate
    rng = np.random.default_rng(3)
    nO, nzi, nI, nzo = 2, 40, 300, 9
    fld, vc = _columns(rng, nO, nzi, nI, "pressure")
    vout = np.sort(rng.uniform(20, 1050, (nO, nzo, nI)).astype(np.float32), axis=1)
    vmin = rng.uniform(0, 200, (nO, nI))
    vmax = rng.uniform(800, 1100, (nO, nI))
    bad, fill = np.float32(-999.), np.float32(1e30)
    raw = fld.copy()
    raw[np.isnan(raw)] = bad
    for kw in (dict(vmin=vmin, vmax=vmax), dict(vmin=vmin), dict(vmax=vmax), {}):
        want = _oracle_vertical(ol.VINT_LIN, fld, vc, vout=vout, valid_min=250., valid_max=300., **kw)
        oracle.orc_nanf2bad(fp(want), want.size, fill)
        got = verticalInterpolate(ol.VINT_LIN, torch.from_numpy(raw).cuda(), torch.from_numpy(vc).cuda(), verticalOut=torch.from_numpy(vout).cuda(),
                                  validMin=torch.from_numpy(kw["vmin"]).cuda() if "vmin" in kw else None,
                                  validMax=torch.from_numpy(kw["vmax"]).cuda() if "vmax" in kw else None, valid_min=250., valid_max=300.,
                                  badValue=float(bad), fillValue=float(fill))
        torch.cuda.synchronize()
        assert _eq(got.cpu().numpy(), want)


def test_vertical_hybrid_pressure_matches_fimex_output_files(fx):
    """the reference's own expected files (see tests/golden/make_golden.py): pressure formed on the fly from ap, b, ps"""
    from fimex_b200.vertical import verticalHybridPressure
    from test_oracle_golden import cdmcmp_close
    g = np.load(os.path.join(GOLD, "reference_vertical.npz"))
    lev = g["sh_levels"]
    bad = 9.96921e36
    for name in ("air_potential_temperature", "specific_humidity"):
        v, e = g["sh_in_" + name], g["sh_ex_" + name]
        nt, nk, ny, nx = v.shape
        for t in range(nt):
            out = verticalHybridPressure(ol.VINT_LOG, v[t].reshape(1, nk, ny * nx), g["sh_ap_pa"] * 0.01, g["sh_b"], g["sh_ps_raw"][t].reshape(1, -1),
                                         lev, ps_scale=float(g["sh_ps_scale"]), badValue=bad, fillValue=bad)
            assert cdmcmp_close(out[0], e[t].reshape(lev.size, -1)).all(), (name, t)
    lev = g["en_levels"]
    for name in ("specific_humidity_ml", "x_wind_ml", "y_wind_ml", "upward_air_velocity_ml"):
        v, e = g["en_in_" + name], g["en_ex_" + name]
        nt, nk = v.shape[:2]
        ncol = int(np.prod(v.shape[2:]))
        for t in range(nt):
            out = verticalHybridPressure(ol.VINT_LOG, v[t].reshape(1, nk, ncol).astype(np.float32), g["en_ap_pa"] * 0.01, g["en_b"],
                                         g["en_ps_pa"][t].reshape(1, ncol), lev, ps_scale=0.01, badValue=-2147483647.0)[0]
            got = np.where(np.isnan(out), -2147483647, np.sign(out) * np.floor(np.abs(out.astype(np.float64)) + 0.5)).astype(np.int64)
            exp = e[t].reshape(lev.size, ncol).astype(np.int64)
            # the packed integers fimex wrote; a 1-ulp difference of a log weight may move a value sitting on .5 by one count
            assert np.abs(got - exp).max() <= 1 and (got != exp).mean() < 1e-3, (name, t)


def test_vertical_hybrid_pressure_vs_oracle_config5_shape(fx):
    """BASELINE configs[4] layout: (hybrid=65, member, y, x) -> 12 pressure levels, ps broadcast or per member"""
    from fimex_b200.vertical import verticalHybridPressure
    rng = np.random.default_rng(12)
    g = np.load(os.path.join(GOLD, "reference_vertical.npz"))
    ap, b = g["en_ap_pa"] * 0.01, g["en_b"]
    nmem, ny, nx = 4, 23, 31
    nI = nmem * ny * nx
    levels = np.array([1000, 925, 850, 700, 500, 400, 300, 250, 200, 150, 100, 50], dtype=np.float64)
    ps = (101325 - 12 * rng.uniform(0, 2500, nI) + rng.normal(0, 300, nI)).astype(np.float32)
    fld = rng.normal(0, 10, (1, 65, nI)).astype(np.float32)
    p = (ap[:, None] + b[:, None] * (ps.astype(np.float64) * 0.01)[None, :]).astype(np.float32)[None]
    want = _oracle_vertical(ol.VINT_LIN, fld, np.ascontiguousarray(p), levels=levels)
    got = verticalHybridPressure(ol.VINT_LIN, fld, ap, b, ps.reshape(1, nI), levels, ps_scale=0.01)
    assert _eq(got, want)
    wantl = _oracle_vertical(ol.VINT_LOG, fld, np.ascontiguousarray(p), levels=levels)
    gotl = verticalHybridPressure(ol.VINT_LOG, fld, ap, b, ps.reshape(1, nI), levels, ps_scale=0.01)
    assert _close(gotl, wantl)


@pytest.mark.parametrize("name", ["nearest", "linear", "linear_weak_extrapol", "linear_no_extrapol", "linear_const_extrapol", "log", "log_log"])
def test_two_level_kernels(fx, oracle, name):
    import fimex_b200.vertical as V
    rng = np.random.default_rng(7)
    fo, fg = getattr(oracle, f"orc_get_values_{name}_f"), getattr(V, f"mifi_get_values_{name}_f")
    n = 1000
    for _ in range(40):
        A = rng.normal(size=n).astype(np.float32)
        B = rng.normal(size=n).astype(np.float32)
        A[rng.random(n) < 0.1] = np.nan
        xa, xb = rng.uniform(1, 1000, 2)
        x = float(rng.choice([xa, xb, rng.uniform(-100, 1500), (xa + xb) / 2]))
        if rng.random() < 0.1:
            xb = xa
        want = np.full(n, 12345, np.float32)
        rc = fo(fp(A), fp(B), fp(want), n, xa, xb, x)
        got, rcg = fg(A, B, xa, xb, x)
        assert rc == rcg and _eq(got, want)


# ---------------------------------------------------------------------------------- forward aggregation
FWD = [("sum", ol.FWD_SUM), ("mean", ol.FWD_MEAN), ("median", ol.FWD_MEDIAN), ("max", ol.FWD_MAX), ("min", ol.FWD_MIN),
       ("undef_sum", ol.FWD_UNDEF_SUM), ("undef_mean", ol.FWD_UNDEF_MEAN), ("undef_max", ol.FWD_UNDEF_MAX), ("undef_min", ol.FWD_UNDEF_MIN)]


@pytest.mark.parametrize("name,method", FWD)
def test_forward_golden(fx, name, method):
    g = np.load(os.path.join(GOLD, "reference_small.npz"))
    in_x, in_y, out_x, out_y, nz = (int(v) for v in g["f_dims"])
    ci = fx.CachedForwardInterpolation("x", "y", method, g["f_px"], g["f_py"], in_x, in_y, out_x, out_y)
    out, n = ci.interpolateValues(ci.reduceInput(g["f_field"], in_x, in_y))
    assert n == nz * out_x * out_y
    assert _eq(out, g[f"f_{name}_out"])  # the reference's own output (level-0 quirk included)


@pytest.mark.parametrize("method", [m for _, m in FWD] + [ol.FWD_UNDEF_MEDIAN])
@pytest.mark.parametrize("compat", [True, False])
def test_forward_vs_oracle(fx, method, compat):
    rng = np.random.default_rng(method)
    in_x, in_y, out_x, out_y, nz = 203, 157, 41, 33, 4
    pox = rng.uniform(-3, out_x + 2, in_x * in_y)
    poy = rng.uniform(-3, out_y + 2, in_x * in_y)
    pox[:50] = np.round(pox[:50] * 2) / 2
    fld = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    fld[rng.random(fld.shape) < 0.1] = np.nan
    want, info = ol.o_cached_apply(method, pox, poy, in_x, in_y, out_x, out_y, fld, level_bug=1 if compat else 0)
    ci = fx.CachedForwardInterpolation("x", "y", method, pox, poy, in_x, in_y, out_x, out_y, level_compat=compat)
    assert (ci.getInX(), ci.getInY(), ci.reducedDomain() is not None) == (info["inX"], info["inY"], info["reduced"])
    got, _ = ci.interpolateValues(ci.reduceInput(fld, in_x, in_y))
    assert _eq(got, want)
    if not compat:
        assert not _eq(got[0], got[1])


def test_forward_reduced_window_bad_fill_and_swath_shape(fx, oracle):
    """a 2030x1354 swath-sized source onto a small grid: long segments, reduced window, fused bad/fill, device resident"""
    import torch
    rng = np.random.default_rng(5)
    in_x, in_y, out_x, out_y, nz = 1354, 2030, 120, 90, 2
    gx, gy = np.meshgrid(np.arange(in_x), np.arange(in_y))
    pox = (gx - 300) * (out_x / 700.0) + rng.normal(0, 0.2, gx.shape)   # only part of the swath hits the target
    poy = (gy - 500) * (out_y / 900.0) + rng.normal(0, 0.2, gx.shape)
    bad, fill = np.float32(-1e30), np.float32(9.96921e36)
    fld = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    fld[rng.random(fld.shape) < 0.3] = np.nan      # "30 % land-masked"
    for method in (ol.FWD_MEAN, ol.FWD_MAX, ol.FWD_MEDIAN):
        want, info = ol.o_cached_apply(method, pox.ravel(), poy.ravel(), in_x, in_y, out_x, out_y, fld, level_bug=0)
        oracle.orc_nanf2bad(fp(want), want.size, fill)
        ci = fx.CachedForwardInterpolation("x", "y", method, pox.ravel(), poy.ravel(), in_x, in_y, out_x, out_y, level_compat=False)
        assert info["reduced"] and ci.reducedDomain() is not None
        assert (ci.reducedDomain().xMin, ci.reducedDomain().yMin, ci.getInX(), ci.getInY()) == (info["xMin"], info["yMin"], info["inX"], info["inY"])
        raw = fld.copy()
        raw[np.isnan(raw)] = bad
        d_in = torch.from_numpy(ci.reduceInput(raw, in_x, in_y)).cuda()
        got, _ = ci.interpolateValues(d_in, badValue=float(bad), fillValue=float(fill))
        torch.cuda.synchronize()
        assert _eq(got.cpu().numpy(), want)


# ---------------------------------------------------------------------------------- chained horizontal -> vertical (configs[4])
def test_chained_horizontal_then_vertical_on_device(fx):
    """BASELINE configs[4], second half: ensemble hybrid-sigma field (hybrid, member, y, x) regridded bilinearly first, then
    interpolated to pressure levels with the pressure formed from the REGRIDDED surface pressure (the reference's order,
    fimex.cc:862-863 / SURVEY 3.5).  All slabs stay on the device in the padded layout; the oracle runs the same chain."""
    import torch
    from fimex_b200.vertical import verticalHybridPressure
    rng = np.random.default_rng(42)
    g = np.load(os.path.join(GOLD, "reference_vertical.npz"))
    ap, b = g["en_ap_pa"] * 0.01, g["en_b"]
    nk, nmem, in_x, in_y, out_x, out_y = 65, 3, 90, 70, 217, 153
    gx, gy = np.meshgrid(np.arange(out_x), np.arange(out_y))
    px = 2.0 + gx * ((in_x - 5.0) / out_x) + 0.05 * gy * (in_x / out_x)
    py = 1.5 + gy * ((in_y - 12.0) / out_y) + 0.03 * gx * (in_y / out_x)
    fld = rng.normal(0, 10, (nk, nmem, in_y, in_x)).astype(np.float32)
    ps = (101325 - 12 * rng.uniform(0, 2500, (nmem, in_y, in_x)) + rng.normal(0, 300, (nmem, in_y, in_x))).astype(np.float32)
    levels = np.array([1000, 925, 850, 700, 500, 400, 300, 250, 200, 150, 100, 50], dtype=np.float64)
    # oracle chain
    h_f, info = ol.o_cached_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, fld.reshape(-1, in_y, in_x))
    h_ps, _ = ol.o_cached_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, ps)
    ncol = nmem * out_y * out_x
    p = (ap[:, None] + b[:, None] * (h_ps.reshape(1, ncol).astype(np.float64) * 0.01)).astype(np.float32)
    want = _oracle_vertical(ol.VINT_LIN, h_f.reshape(1, nk, ncol), np.ascontiguousarray(p[None]), levels=levels).reshape(12, nmem, out_y, out_x)
    # device chain
    ci = fx.CachedInterpolation("x", "y", ol.BILINEAR, px, py, in_x, in_y, out_x, out_y)
    d_f = ci.toPadded(torch.from_numpy(ci.reduceInput(fld.reshape(-1, in_y, in_x), in_x, in_y)).cuda())
    d_ps = ci.toPadded(torch.from_numpy(ci.reduceInput(ps, in_x, in_y)).cuda())
    r_f = ci.interpolateValuesPadded(d_f)      # [65*3][out_y][pitch]
    r_ps = ci.interpolateValuesPadded(d_ps)    # [3][out_y][pitch]
    pitch = r_f.shape[2]
    n_inner = nmem * out_y * pitch
    out = verticalHybridPressure(ol.VINT_LIN, r_f.reshape(1, nk, n_inner), ap, b, r_ps.reshape(1, n_inner), levels, ps_scale=0.01)
    torch.cuda.synchronize()
    got = out.reshape(12, nmem, out_y, pitch)[..., :out_x].cpu().numpy()
    assert _eq(got, want)
    assert (~np.isnan(want)).mean() > 0.5


# ---------------------------------------------------------------------------------- sweep kernel (sorted fixed levels)
def _sorted_level_list(vc, order):
    lv = [1000., 925., 850., 700., 500., 400., 300., 250., 200., 150., 100., 50., 1100., 5., 1e12, -1e12, 3.0e9, 0.0, 512.0, 512.0]
    picks = [vc[0, 0, 0], vc[0, -1, 0], vc[0, vc.shape[1] // 2, 5 % vc.shape[2]], vc[-1, min(1, vc.shape[1] - 1), 7 % vc.shape[2]],
             vc[-1, max(vc.shape[1] - 2, 0), 11 % vc.shape[2]]]
    lv += [float(v) for v in picks if not np.isnan(v)]
    lv += [float(np.nextafter(np.float32(v), np.float32(np.inf))) for v in picks[:2] if not np.isnan(v)]
    lv = sorted(lv)
    return lv if order == "up" else lv[::-1]


@pytest.mark.parametrize("method", [ol.VINT_LIN, ol.VINT_LOG, ol.VINT_NN, ol.VINT_LIN_WEAK, ol.VINT_LIN_CONST])
@pytest.mark.parametrize("kind", ["pressure", "reversed", "unordered", "mixed", "ties"])
@pytest.mark.parametrize("order", ["up", "down"])
def test_vertical_sweep_sorted_levels_vs_oracle(fx, method, kind, order):
    """sorted target levels take the one-pass sweep kernel: every bracket must be the reference search's, including levels equal to
    a column element, beyond either end, f32 difference ties (-> per-output general path) and non-monotone / NaN columns (-> whole
    column by the general scan)"""
    from fimex_b200.vertical import verticalInterpolate
    rng = np.random.default_rng(1000 * method + 10 * len(kind) + len(order))
    nO, nzi, nI = 2, 65, 389
    if kind == "mixed":
        fld, vc = _columns(rng, nO, nzi, nI, "pressure")
        _, vr = _columns(rng, nO, nzi, nI, "reversed")
        _, vu = _columns(rng, nO, nzi, nI, "unordered")
        sel = rng.integers(0, 4, (nO, 1, nI))
        vc = np.where(sel == 1, vr, np.where(sel == 2, vu, vc))
        eq = sel[:, 0, :] == 3  # strictly monotone except one repeated element
        vc[:, 40, :] = np.where(eq, vc[:, 39, :], vc[:, 40, :])
        vc[0, 64, 3] = np.nan
        vc[1, 0, 4] = np.nan
        vc = np.ascontiguousarray(vc.astype(np.float32))
    elif kind == "ties":  # spacing of a few ulps on a large offset: differences to far levels round to the same float
        fld, _ = _columns(rng, nO, nzi, nI, "pressure")
        step = rng.choice([8.0, 16.0, -8.0], (nO, 1, nI))
        vc = np.ascontiguousarray((1.0e8 + step * np.arange(nzi)[None, :, None]).astype(np.float32))
    else:
        fld, vc = _columns(rng, nO, nzi, nI, kind)
    levels = _sorted_level_list(vc, order)
    if kind == "ties":
        levels = sorted(levels + [1.0e8, 1.0e8 + 16, 1.0e8 + 20, 1.0e8 - 300, 1.0e8 + 8 * 64, 1.0e8 + 8 * 64 + 4, 2.0e8, 1.7e15], reverse=(order == "down"))
    want = _oracle_vertical(method, fld, vc, levels=levels)
    got = verticalInterpolate(method, fld, vc, level1=levels)
    if method == ol.VINT_LOG:
        assert _close(got, want)
    else:
        assert _eq(got, want)


@pytest.mark.parametrize("nzi", [1, 2, 3, 5])
def test_vertical_sweep_short_columns_and_validity(fx, nzi):
    import torch
    from fimex_b200.vertical import verticalInterpolate
    rng = np.random.default_rng(nzi)
    nO, nI = 2, 130
    fld, vc = _columns(rng, nO, nzi, nI, "pressure" if nzi % 2 else "reversed")
    levels = sorted([float(vc[0, 0, 0]), float(vc[1, -1, 3]), 1200., 1., 500., 700., 900.])
    vmin, vmax = rng.uniform(0, 600, (nO, nI)), rng.uniform(650, 1300, (nO, nI))
    for kw in (dict(vmin=vmin, vmax=vmax), dict(vmax=vmax), {}):
        want = _oracle_vertical(ol.VINT_LIN, fld, vc, levels=levels, valid_min=260., valid_max=300., **kw)
        got = verticalInterpolate(ol.VINT_LIN, torch.from_numpy(fld).cuda(), torch.from_numpy(vc).cuda(), level1=levels,
                                  validMin=torch.from_numpy(kw["vmin"]).cuda() if "vmin" in kw else None,
                                  validMax=torch.from_numpy(kw["vmax"]).cuda() if "vmax" in kw else None, valid_min=260., valid_max=300.)
        torch.cuda.synchronize()
        assert _eq(got.cpu().numpy(), want)


@pytest.mark.parametrize("order", ["up", "down"])
def test_vertical_sweep_hybrid_nan_ps_and_level_order(fx, order):
    from fimex_b200.vertical import verticalHybridPressure
    rng = np.random.default_rng(77)
    g = np.load(os.path.join(GOLD, "reference_vertical.npz"))
    ap, b = g["en_ap_pa"] * 0.01, g["en_b"]
    nI = 1500
    levels = np.array([1000, 925, 850, 700, 500, 400, 300, 250, 200, 150, 100, 50, 1, 0.001, 1100], dtype=np.float64)
    levels = np.sort(levels) if order == "up" else np.sort(levels)[::-1].copy()
    ps = (101325 - 12 * rng.uniform(0, 2500, nI) + rng.normal(0, 300, nI)).astype(np.float32)
    ps[::97] = np.nan
    ps[5] = 0.0
    ps[6] = -500.0
    fld = rng.normal(0, 10, (1, 65, nI)).astype(np.float32)
    p = (ap[:, None] + b[:, None] * (ps.astype(np.float64) * 0.01)[None, :]).astype(np.float32)[None]
    for m in (ol.VINT_LIN, ol.VINT_LOG, ol.VINT_LOGLOG):
        want = _oracle_vertical(m, fld, np.ascontiguousarray(p), levels=levels)
        got = verticalHybridPressure(m, fld, ap, b, ps.reshape(1, nI), levels, ps_scale=0.01)
        assert (_eq if m == ol.VINT_LIN else _close)(got, want), m


def test_vertical_hybrid_outlier_surface_pressure(fx):
    """fill-like and garbage surface pressures (1e30, 0, negative, inf, NaN) in a few columns: those columns follow the reference's
    general scan, all others are unaffected"""
    from fimex_b200.vertical import verticalHybridPressure
    rng = np.random.default_rng(9)
    g = np.load(os.path.join(GOLD, "reference_vertical.npz"))
    ap, b = g["en_ap_pa"] * 0.01, g["en_b"]
    nI = 6000
    levels = np.array([1000, 925, 850, 700, 500, 400, 300, 250, 200, 150, 100, 50], dtype=np.float64)
    ps = (101325 - 12 * rng.uniform(0, 2500, nI) + rng.normal(0, 300, nI)).astype(np.float32)
    ps[::50] = np.float32(1e30)
    ps[7::61] = 0.0
    ps[11::67] = -3.0e4
    ps[13::71] = np.inf
    ps[17::73] = np.nan
    ps[19::79] = np.float32(9.96921e36)
    ps[23::83] = np.float32(55000.0)  # an isolated but valid value
    fld = rng.normal(0, 10, (1, 65, nI)).astype(np.float32)
    with np.errstate(all="ignore"):
        p = (ap[:, None] + b[:, None] * (ps.astype(np.float64) * 0.01)[None, :]).astype(np.float32)[None]
    for m in (ol.VINT_LIN, ol.VINT_LOG):
        want = _oracle_vertical(m, fld, np.ascontiguousarray(p), levels=levels)
        got = verticalHybridPressure(m, fld, ap, b, ps.reshape(1, nI), levels, ps_scale=0.01)
        assert (_eq if m == ol.VINT_LIN else _close)(got, want), m


@pytest.mark.parametrize("name,method", [("nearest", ol.VINT_NN), ("linear", ol.VINT_LIN), ("linear_weak_extrapol", ol.VINT_LIN_WEAK),
                                         ("linear_no_extrapol", ol.VINT_LIN_NO), ("linear_const_extrapol", ol.VINT_LIN_CONST), ("log", ol.VINT_LOG),
                                         ("log_log", ol.VINT_LOGLOG)])
def test_two_level_kernels_on_device_slabs(fx, oracle, name, method):
    """CDMTimeInterpolator's use of the two-level kernels (CDMTimeInterpolator.cc:127) on slabs that stay on the device"""
    import torch
    from fimex_b200.vertical import timeInterpolateSlabs
    rng = np.random.default_rng(method)
    n = 100003
    A = rng.normal(280, 10, n).astype(np.float32)
    B = rng.normal(280, 10, n).astype(np.float32)
    A[::101] = np.nan
    B[::103] = np.nan
    fn = getattr(oracle, "orc_get_values_" + name + "_f")
    dA, dB = torch.from_numpy(A).cuda(), torch.from_numpy(B).cuda()
    for a, b, x in ((3.0, 9.0, 4.5), (3.0, 9.0, 3.0), (3.0, 9.0, 9.0), (3.0, 9.0, 20.0), (3.0, 9.0, 0.5), (5.0, 5.0, 5.0), (9.0, 3.0, 13.0)):
        want = np.full(n, -7.0, np.float32)
        rc = fn(fp(A), fp(B), fp(want), n, a, b, x)
        assert rc == 1
        got = timeInterpolateSlabs(dA, dB, a, b, x, method=method)
        torch.cuda.synchronize()
        assert _eq(got.cpu().numpy(), want), (name, a, b, x)
    # in place on the first slab, like a caller that reuses its buffer
    want = np.empty(n, np.float32)
    fn(fp(A), fp(B), fp(want), n, 3.0, 9.0, 4.5)
    from fimex_b200 import _lib
    dC = dA.clone()
    _lib.check(_lib.lib.fimexcu_get_values_two_level_dev_f(method, dC.data_ptr(), dB.data_ptr(), dC.data_ptr(), n, 3.0, 9.0, 4.5, _lib.current_stream_ptr()), "x")
    torch.cuda.synchronize()
    assert _eq(dC.cpu().numpy(), want)
    # slabs that do not start on a 16-byte boundary (views into larger arrays): the element-by-element path
    A1, B1, want1 = A[1:].copy(), B[1:].copy(), np.empty(n - 1, np.float32)
    fn(fp(A1), fp(B1), fp(want1), n - 1, 3.0, 9.0, 4.5)
    got = timeInterpolateSlabs(dA[1:], dB[1:], 3.0, 9.0, 4.5, method=method)
    torch.cuda.synchronize()
    assert _eq(got.cpu().numpy(), want1)


def test_config5_full_size_properties(fx):
    """BASELINE configs[4] at full horizontal size (65 hybrid levels x 4 members x 1069 x 949 -> 12 pressure levels, log): a field
    that is linear in log(p) must come back as the same line evaluated at the target pressures (1e-5 relative, the tolerance of
    BASELINE.json), levels below the surface follow the reference's extrapolation from the two lowest levels, and a random sample
    of columns equals the oracle."""
    import torch
    from fimex_b200.vertical import verticalHybridPressure
    from fimex_b200 import synthetic as S
    g = np.load(os.path.join(GOLD, "reference_vertical.npz"))
    ap, b = g["en_ap_pa"] * 0.01, g["en_b"]
    nmem, ny, nx = 4, S.MEPS_NY, S.MEPS_NX
    nI = nmem * ny * nx
    levels = np.array([1000, 925, 850, 700, 500, 400, 300, 250, 200, 150, 100, 50], dtype=np.float64)
    gen = torch.Generator(device="cuda").manual_seed(11)
    yy = torch.arange(ny, device="cuda", dtype=torch.float32)[:, None]
    xx = torch.arange(nx, device="cuda", dtype=torch.float32)[None, :]
    h = 1250.0 * (1 + torch.sin(xx / 97.0) * torch.cos(yy / 131.0))  # smooth orography 0..2500 m
    ps = (101325.0 - 12.0 * h)[None].repeat(nmem, 1, 1) + 300.0 * torch.randn((nmem, ny, nx), device="cuda", generator=gen)
    ps = ps.reshape(1, nI).contiguous()
    p = (torch.from_numpy(ap).cuda()[:, None] + torch.from_numpy(b).cuda()[:, None] * (ps.double() * 0.01)).float()  # [65][nI], the f32 column
    alpha, beta = 250.0, 7.5
    fld = (alpha + beta * torch.log(p.double())).float().reshape(1, 65, nI).contiguous()
    out = verticalHybridPressure(ol.VINT_LOG, fld, ap, b, ps, levels, ps_scale=0.01)
    torch.cuda.synchronize()
    want = torch.from_numpy(alpha + beta * np.log(levels)).cuda().float()[None, :, None].expand(1, 12, nI)
    # the first hybrid level of this fixture is the constant 10 hPa, so every target level lies inside or below the column
    assert not torch.isnan(out).any()
    rel = ((out - want).abs() / want.abs()).max().item()
    assert rel < 1e-5, rel
    # random sample of columns against the oracle (bit-exact masks, 1e-5 relative values)
    idx = torch.randint(0, nI, (4000,), device="cuda", generator=gen)
    f_s = fld[0][:, idx].cpu().numpy()[None]
    p_s = p[:, idx].cpu().numpy()[None]
    w = _oracle_vertical(ol.VINT_LOG, np.ascontiguousarray(f_s), np.ascontiguousarray(p_s), levels=levels)
    assert _close(out[0][:, idx].cpu().numpy()[None], w)


@pytest.mark.parametrize("env", [{"FIMEXCU_VERTICAL_NO_TABLE": "1"}, {"FIMEXCU_VERTICAL_NO_TABLE": "1", "FIMEXCU_VERTICAL_NO_SWEEP": "1"},
                                 {"FIMEXCU_VERTICAL_NO_SEARCH": "1"}])
def test_vertical_fallback_tiers(env):
    """The hybrid-sigma path has three tiers (launch-wide table -> one-pass sweep -> bisection per level), the explicit coordinate
    two above the general kernel (scan + bisection in shared memory -> sweep); the lower ones serve shapes the upper ones do not
    take.  The switches are read once per process, so the parity tests are re-run in a subprocess with the upper tiers switched off."""
    import subprocess
    import sys
    e = dict(os.environ)
    e.update(env)
    here = os.path.dirname(os.path.abspath(__file__))
    select = "vertical and not fallback_tiers" if "FIMEXCU_VERTICAL_NO_SEARCH" in env else "hybrid or sweep or config5 or chained"
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(here, "test_gpu_vertical_forward.py"), "-m", "gpu", "-x", "-q", "-p", "no:cacheprovider",
                        "-k", select], env=e, capture_output=True, text=True, cwd=os.path.dirname(here))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert " passed" in r.stdout


# ---------------------------------------------------------------------------------- padded rows (slabs straight from the horizontal step)
@pytest.mark.parametrize("kind", ["hybrid", "explicit"])
def test_vertical_padded_rows_skip_the_pad(fx, kind):
    """rows = (width, pitch): the pad elements of every row are neither read (whatever they hold cannot slow or change anything)
    nor written; the columns give what the dense call gives.  Also: columns whose coordinate is NaN at every level."""
    import torch
    from fimex_b200.vertical import verticalHybridPressure, verticalInterpolate
    rng = np.random.default_rng(31)
    g = np.load(os.path.join(GOLD, "reference_vertical.npz"))
    ap, b = g["en_ap_pa"] * 0.01, g["en_b"]
    W, pitch, ny, nO, nzi = 37, 64, 23, 2, ap.size
    nd, npad = ny * W, ny * pitch
    levels = np.array([1000, 925, 850, 700, 500, 400, 300, 250, 200, 150, 100, 50, 1100], dtype=np.float64)
    ps = (101325 - 12 * rng.uniform(0, 2500, (nO, nd)) + rng.normal(0, 300, (nO, nd))).astype(np.float32)
    ps[:, ::29] = np.nan  # outside the source grid of the horizontal step
    ps[0, 5] = np.float32(1e30)
    fld = rng.normal(0, 10, (nO, nzi, nd)).astype(np.float32)
    with np.errstate(all="ignore"):
        p = (ap[None, :, None] + b[None, :, None] * (ps.astype(np.float64) * 0.01)[:, None, :]).astype(np.float32)
    if kind == "explicit":
        p[1, 3:9, 11] = np.nan  # a column that is NaN in part only
    garbage = np.array([0.0, np.nan, 1e30, -5.0, 3.0e4, np.inf, 101325.0], dtype=np.float32)

    def padded(a):  # [..., nd] -> [..., npad] with garbage in the pad
        lead = a.shape[:-1]
        buf = np.resize(garbage, int(np.prod(lead)) * npad).reshape(lead + (ny, pitch)).copy()
        buf[..., :W] = a.reshape(lead + (ny, W))
        return torch.from_numpy(buf.reshape(lead + (npad,))).cuda()

    for m in (ol.VINT_LIN, ol.VINT_LOG):
        want = _oracle_vertical(m, fld, np.ascontiguousarray(p), levels=levels)
        out = torch.full((nO, levels.size, npad), 12345.0, dtype=torch.float32, device="cuda")
        if kind == "hybrid":
            verticalHybridPressure(m, padded(fld), ap, b, padded(ps), levels, ps_scale=0.01, rows=(W, pitch), out=out)
            dense = verticalHybridPressure(m, torch.from_numpy(fld).cuda(), ap, b, torch.from_numpy(ps).cuda(), levels, ps_scale=0.01)
        else:
            verticalInterpolate(m, padded(fld), padded(p), level1=levels, rows=(W, pitch), out=out)
            dense = verticalInterpolate(m, torch.from_numpy(fld).cuda(), torch.from_numpy(p).cuda(), level1=levels)
        torch.cuda.synchronize()
        got = out.cpu().numpy().reshape(nO, levels.size, ny, pitch)
        assert np.all(got[..., W:] == 12345.0), "pad elements were written"
        cols = got[..., :W].reshape(nO, levels.size, nd)
        assert _eq(cols, dense.cpu().numpy()), (kind, m)
        assert (_eq if m == ol.VINT_LIN else _close)(cols, want), (kind, m)
        assert np.isnan(cols[:, :, ::29]).all()


def test_vertical_padded_rows_arguments(fx):
    import torch
    from fimex_b200.vertical import verticalInterpolate
    from fimex_b200._lib import FimexCudaError
    f = torch.zeros((1, 4, 128), dtype=torch.float32, device="cuda")
    c = torch.arange(4, dtype=torch.float32, device="cuda").reshape(1, 4, 1).repeat(1, 1, 128).contiguous()
    lv = np.array([1.5])
    for rows in ((0, 64), (65, 64), (30, 48)):  # no width, width beyond the pitch, pitch that does not divide the inner size
        with pytest.raises(FimexCudaError):
            verticalInterpolate(ol.VINT_LIN, f, c, level1=lv, rows=rows)
    a = verticalInterpolate(ol.VINT_LIN, f, c, level1=lv, rows=(64, 64))  # no pad: the plain form
    b = verticalInterpolate(ol.VINT_LIN, f, c, level1=lv)
    torch.cuda.synchronize()
    assert _eq(a.cpu().numpy(), b.cpu().numpy())
