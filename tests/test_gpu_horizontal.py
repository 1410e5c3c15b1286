"""GPU parity tests of the horizontal path, through the C ABI (ctypes -> libfimexcu.so).

Checker: oracle/liboracle.so (pinned against the reference in test_oracle_*.py) and the committed golden
fixtures produced by the reference's own code (tests/golden/reference_small.npz).
Bar: BIT-EXACT for every value, index and undefined mask (the CUDA kernels reproduce the reference's
rounding order; the 1e-5 relative tolerance of BASELINE.json is not needed and not used)."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import dp, fp

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_small.npz")


def _eq(a, b):
    a, b = np.asarray(a), np.asarray(b)
    na, nb = np.isnan(a), np.isnan(b)
    return a.shape == b.shape and np.array_equal(na, nb) and np.array_equal(a[~na], b[~nb])


@pytest.fixture(scope="module")
def fx():
    import fimex_b200
    assert fimex_b200.lib.fimexcu_device_count() >= 1, "no sm_100 device: the CUDA path must run, there is no fallback"
    return fimex_b200


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLDEN)


def _tables(rng, in_x, in_y, n, margin=3.0):
    px = rng.uniform(-margin, in_x - 1 + margin, n)
    py = rng.uniform(-margin, in_y - 1 + margin, n)
    k = n // 8
    px[:k] = np.round(px[:k] * 2) / 2
    py[:k] = np.round(py[:k] * 2) / 2
    px[k:k + 4] = -999.
    py[k + 2:k + 6] = -1000.
    return px, py


# ---------------------------------------------------------------------------------- points2position
def test_points2position_golden(fx, gold):
    names = sorted({k[:-5] for k in gold.files if k.startswith("p_") and k.endswith("_axis")})
    assert len(names) == 15
    for n in names:
        t = int(n.rsplit("_", 1)[1])
        got = fx.mifi_points2position(gold[n + "_in"], gold[n + "_axis"], t)
        assert _eq(got, gold[n + "_out"]), n


def test_points2position_vs_oracle_large(fx, oracle):
    rng = np.random.default_rng(42)
    ax = np.linspace(-1060084., 1309916., 949)
    pts = rng.uniform(-2e6, 2e6, 300000)
    pts[:1000] = rng.choice(ax, 1000)
    want = pts.copy()
    oracle.orc_points2position(dp(want), want.size, dp(ax), ax.size, ol.PROJ_AXIS)
    assert _eq(fx.mifi_points2position(pts, ax, ol.PROJ_AXIS), want)


# ---------------------------------------------------------------------------------- cached apply
@pytest.mark.parametrize("tag", ["full", "sub"])
@pytest.mark.parametrize("name,method", [("bilinear", ol.BILINEAR), ("bicubic", ol.BICUBIC), ("nearest", ol.NEAREST)])
def test_cached_apply_golden(fx, gold, tag, name, method):
    in_x, in_y, out_x, out_y, nz = (int(v) for v in gold["h_dims"])
    ci = fx.createCachedInterpolation("x", "y", method, gold[f"h_{tag}_px"], gold[f"h_{tag}_py"], in_x, in_y, out_x, out_y)
    info = gold[f"h_{tag}_{name}_info"]
    assert (ci.getInX(), ci.getInY()) == (info[0], info[1])
    assert (ci.reducedDomain() is not None) == bool(info[2])
    if ci.reducedDomain() is not None:
        assert (ci.reducedDomain().xMin, ci.reducedDomain().yMin) == (info[3], info[4])
    sub = ci.reduceInput(gold["h_field"], in_x, in_y)
    out, new_size = ci.interpolateValues(sub)
    assert new_size == nz * out_x * out_y
    assert _eq(out, gold[f"h_{tag}_{name}_out"])


@pytest.mark.parametrize("method", [ol.BILINEAR, ol.BICUBIC, ol.NEAREST, ol.COORD_NN_KD])
@pytest.mark.parametrize("shape", [(61, 47, 53, 39, 5), (300, 200, 517, 333, 7), (5, 4, 1, 1, 1), (2, 2, 40, 30, 2), (64, 64, 1000, 3, 65)])
def test_cached_apply_vs_oracle(fx, method, shape):
    in_x, in_y, out_x, out_y, nz = shape
    rng = np.random.default_rng(sum(shape) + method)
    px, py = _tables(rng, in_x, in_y, out_x * out_y)
    fld = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    fld[rng.random(fld.shape) < 0.03] = np.nan
    want, info = ol.o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, fld)
    ci = fx.createCachedInterpolation("x", "y", method, px, py, in_x, in_y, out_x, out_y)
    assert (ci.getInX(), ci.getInY(), ci.reducedDomain() is not None) == (info["inX"], info["inY"], info["reduced"])
    got, _ = ci.interpolateValues(ci.reduceInput(fld, in_x, in_y))
    assert _eq(got, want)


def test_cached_apply_device_resident_and_fused_bad_fill(fx, oracle):
    """device pointers (torch only allocates) + bad->NaN on load and NaN->fill on store fused into the kernel
    (data2InterpolationArray / interpolationArray2Data of CDMInterpolator.cc:140-151)."""
    import torch
    rng = np.random.default_rng(77)
    in_x, in_y, out_x, out_y, nz = 120, 90, 233, 177, 9
    bad, fill = np.float32(9.96921e36), np.float32(-32767.0)
    for method in (ol.BILINEAR, ol.BICUBIC, ol.NEAREST):
        px, py = _tables(rng, in_x, in_y, out_x * out_y)
        fld = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
        fld[rng.random(fld.shape) < 0.05] = bad
        clean = fld.copy()
        oracle.orc_bad2nanf(fp(clean), clean.size, bad)
        want, _ = ol.o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, clean)
        oracle.orc_nanf2bad(fp(want), want.size, fill)
        ci = fx.createCachedInterpolation("x", "y", method, px, py, in_x, in_y, out_x, out_y)
        d_in = torch.from_numpy(ci.reduceInput(fld, in_x, in_y)).cuda()
        d_out, _ = ci.interpolateValues(d_in, badValue=float(bad), fillValue=float(fill))
        torch.cuda.synchronize()
        assert _eq(d_out.cpu().numpy(), want)
        # host path with the same fused conversions
        h_out, _ = ci.interpolateValues(ci.reduceInput(fld, in_x, in_y), badValue=float(bad), fillValue=float(fill))
        assert _eq(h_out, want)


def test_cache_from_device_tables(fx):
    import torch
    rng = np.random.default_rng(5)
    in_x, in_y, out_x, out_y, nz = 80, 60, 90, 70, 3
    px, py = _tables(rng, in_x, in_y, out_x * out_y)
    fld = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    want, _ = ol.o_cached_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, fld)
    ci = fx.CachedInterpolation("x", "y", ol.BILINEAR, torch.from_numpy(px).cuda(), torch.from_numpy(py).cuda(), in_x, in_y, out_x, out_y)
    got, _ = ci.interpolateValues(ci.reduceInput(fld, in_x, in_y))
    assert _eq(got, want)


def test_unknown_method_is_an_error(fx):
    with pytest.raises(fx.FimexCudaError, match="unknown interpolation function"):
        fx.createCachedInterpolation("x", "y", 99, np.zeros(4), np.zeros(4), 2, 2, 2, 2)
    assert fx.lib.fimexcu_cache_create(ol.FWD_SUM, None, None, 2, 2, 0, 0) is None  # forward goes through fimexcu_forward_create


@pytest.mark.parametrize("method", [ol.NEAREST, ol.BILINEAR, ol.BICUBIC])
def test_single_position_kernels(fx, oracle, method):
    from fimex_b200.interpolation import mifi_get_values
    rng = np.random.default_rng(8)
    ix, iy, iz = 9, 7, 5
    fld = rng.normal(size=(iz, iy, ix)).astype(np.float32)
    fo = {ol.NEAREST: oracle.orc_get_values_f, ol.BILINEAR: oracle.orc_get_values_bilinear_f, ol.BICUBIC: oracle.orc_get_values_bicubic_f}[method]
    want = np.zeros(iz, np.float32)
    for x, y in [(0.3, 0.), (0., 0.3), (4.5, 3.5), (7.99, 5.99), (-0.4, 1), (-0.6, 1), (8.4, 6.4), (2.5, 2.5), (8, 6), (1.25, 1.75)]:
        fo(fp(fld), fp(want), x, y, ix, iy, iz)
        assert _eq(mifi_get_values(method, fld, x, y), want), (x, y)


# ---------------------------------------------------------------------------------- vector rotation
def test_vector_rotation_golden(fx, gold):
    cvr = fx.CachedVectorReprojection(gold["v_matrix"].reshape(19, 23, 3))
    u, v = gold["v_u"].copy(), gold["v_v"].copy()
    cvr.reprojectValues(u, v)
    assert _eq(u, gold["v_u_out"]) and _eq(v, gold["v_v_out"])
    d = gold["v_dir"].copy()
    cvr.reprojectDirectionValues(d)
    assert _eq(d, gold["v_dir_out"])


def test_fused_uv_matches_separate_reference_steps(fx, oracle):
    """one fused pass == interpolate u, interpolate v, rotate (CDMInterpolator.cc:232-259)"""
    from fimex_b200.interpolation import interpolateVectorValues
    rng = np.random.default_rng(21)
    in_x, in_y, out_x, out_y, nz = 70, 50, 131, 97, 6
    ang = rng.uniform(-np.pi, np.pi, out_x * out_y)
    m3 = np.stack([np.cos(ang), np.sin(ang), np.degrees(ang)], axis=1).copy()
    for method in (ol.BILINEAR, ol.BICUBIC, ol.NEAREST):
        px, py = _tables(rng, in_x, in_y, out_x * out_y)
        u = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
        v = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
        u[rng.random(u.shape) < 0.02] = np.nan
        wu, _ = ol.o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, u)
        wv, _ = ol.o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, v)
        oracle.orc_vector_reproject_values_by_matrix_f(dp(m3), out_x * out_y, fp(wu), fp(wv), nz)
        ci = fx.createCachedInterpolation("x", "y", method, px, py, in_x, in_y, out_x, out_y)
        cvr = fx.CachedVectorReprojection(m3, out_x, out_y)
        gu, gv = interpolateVectorValues(ci, cvr, ci.reduceInput(u, in_x, in_y), ci.reduceInput(v, in_x, in_y))
        assert _eq(gu, wu) and _eq(gv, wv)


# ---------------------------------------------------------------------------------- full-size properties (BASELINE configs[1])
@pytest.fixture(scope="module")
def meps_t1(fx):
    from fimex_b200 import synthetic as S
    px, py, ox, oy = S.meps_index_tables("T1")
    ci = fx.CachedInterpolation("x", "y", ol.BILINEAR, px, py, S.MEPS_NX, S.MEPS_NY, ox, oy)
    return S, px, py, ox, oy, ci


def test_meps_t1_full_size_against_oracle_levels(fx, oracle, meps_t1):
    """949x1069 -> 2891x1021 at full horizontal size; 2 of the 65 levels so the scalar oracle finishes in seconds"""
    S, px, py, ox, oy, ci = meps_t1
    fld = S.meps_slab_numpy(t=3, nz=2, nan_fraction=0.001)
    want, info = ol.o_cached_apply(ol.BILINEAR, px, py, S.MEPS_NX, S.MEPS_NY, ox, oy, fld)
    assert (ci.getInX(), ci.getInY()) == (info["inX"], info["inY"])
    got, _ = ci.interpolateValues(ci.reduceInput(fld, S.MEPS_NX, S.MEPS_NY))
    assert _eq(got, want)
    frac_defined = 1.0 - np.isnan(want[0]).mean()
    assert 0.60 < frac_defined < 0.66  # 63.4 % of T1 lies inside the MEPS domain


def test_meps_t1_full_slab_properties(fx, meps_t1):
    """all 65 levels, device resident: size-independent properties instead of the scalar oracle."""
    import torch
    S, px, py, ox, oy, ci = meps_t1
    nz = S.MEPS_NZ
    gx = torch.arange(S.MEPS_NX, device="cuda", dtype=torch.float32)[None, None, :]
    gy = torch.arange(S.MEPS_NY, device="cuda", dtype=torch.float32)[None, :, None]
    k = torch.arange(nz, device="cuda", dtype=torch.float32)[:, None, None]
    # (1) a field that is affine in the grid indices is reproduced by bilinear interpolation: f = 2x + 3y + 10k
    fld = (2 * gx + 3 * gy + 10 * k).contiguous()
    sub = ci.reduceInput(fld, S.MEPS_NX, S.MEPS_NY)
    out, n = ci.interpolateValues(sub)
    torch.cuda.synchronize()
    assert n == nz * ox * oy
    out = out.cpu().numpy()
    inside = (px >= 0) & (px <= S.MEPS_NX - 1) & (py >= 0) & (py <= S.MEPS_NY - 1)
    expect = (2 * px + 3 * py)[None] + 10 * np.arange(nz)[:, None, None]
    err = np.abs(out[:, inside] - expect[:, inside]) / np.abs(expect[:, inside]).clip(1)
    assert err.max() < 1e-5  # BASELINE.json tolerance (float32 rounding of the weights)
    # (2) undefined mask: identical on every level, NaN exactly where the position is >= 0.5 cell outside
    mask0 = np.isnan(out[0])
    assert all(np.array_equal(mask0, np.isnan(out[z])) for z in range(nz))
    far_out = (px < -0.5) | (px > S.MEPS_NX - 0.5) | (py < -0.5) | (py > S.MEPS_NY - 0.5)
    assert mask0[far_out].all() and not mask0[inside].any()
    # (3) a constant field stays constant wherever defined (weights sum to one), and levels do not mix
    const = (k + 1.0).expand(nz, S.MEPS_NY, S.MEPS_NX).contiguous()
    outc, _ = ci.interpolateValues(ci.reduceInput(const, S.MEPS_NX, S.MEPS_NY))
    outc = outc.cpu().numpy()
    for z in (0, 31, 64):
        vals = outc[z][~mask0]
        assert np.abs(vals - (z + 1.0)).max() <= 1e-5 * (z + 1.0)


# ---------------------------------------------------------------------------------- staged (TMA) kernel and strided layouts
# (170 levels: more than one level chunk per tile, the last one shorter -- level_chunk() in fxcu_apply.cu)
@pytest.mark.parametrize("shape", [(300, 200, 1400, 900, 9), (949, 1069, 2891, 1021, 3), (128, 96, 700, 500, 65), (128, 96, 700, 500, 170)])
@pytest.mark.parametrize("special", [False, True])
def test_bilinear_staged_kernel_padded_layout(fx, oracle, shape, special):
    """the headline kernel (TMA-staged, warp-specialised) in the padded device layout, incl. rim tiles, tiles that do
    not fit the box (random tables -> direct fallback inside the kernel) and the fused bad/fill conversions"""
    import torch
    in_x, in_y, out_x, out_y, nz = shape
    rng = np.random.default_rng(sum(shape))
    # a smooth, slightly rotated and magnified map (fits the box) with a strip of scattered positions (does not)
    gx, gy = np.meshgrid(np.arange(out_x), np.arange(out_y))
    px = -3.0 + gx * ((in_x + 5.0) / out_x) + 0.11 * gy * (in_x / out_x)
    py = -2.5 + gy * ((in_y + 4.0) / out_y) - 0.07 * gx * (in_y / out_y)
    k = out_y // 2
    px[k:k + 9] = rng.uniform(-2, in_x + 1, px[k:k + 9].shape)
    py[k:k + 9] = rng.uniform(-2, in_y + 1, py[k:k + 9].shape)
    fld = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    bad, fill = np.float32(-999.0), np.float32(1e20)
    fld[rng.random(fld.shape) < 0.01] = bad if special else np.nan
    clean = fld.copy()
    if special:
        oracle.orc_bad2nanf(fp(clean), clean.size, bad)
    want, info = ol.o_cached_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, clean)
    if special:
        oracle.orc_nanf2bad(fp(want), want.size, fill)
    ci = fx.CachedInterpolation("x", "y", ol.BILINEAR, px, py, in_x, in_y, out_x, out_y)
    sub = torch.from_numpy(ci.reduceInput(fld, in_x, in_y)).cuda()
    kw = dict(badValue=float(bad), fillValue=float(fill)) if special else {}
    out = ci.interpolateValuesPadded(ci.toPadded(sub), **kw)
    torch.cuda.synchronize()
    assert _eq(out[:, :, :out_x].cpu().numpy(), want)
    # the dense entry point must give the same answer (direct or staged kernel, whichever its alignment allows)
    dense, _ = ci.interpolateValues(sub, **kw)
    torch.cuda.synchronize()
    assert _eq(dense.cpu().numpy(), want)


def test_staged_and_direct_kernels_agree_on_meps(fx, meps_t1):
    import torch
    S, px, py, ox, oy, ci = meps_t1
    fld = S.meps_slab_torch(t=5, device="cuda", nz=13)
    pin = ci.toPadded(ci.reduceInput(fld, S.MEPS_NX, S.MEPS_NY))
    a = ci.interpolateValuesPadded(pin)
    torch.cuda.synchronize()
    fx.lib.fimexcu_set_staged_kernels(0)  # direct-gather kernels only
    try:
        b = ci.interpolateValuesPadded(pin)
        torch.cuda.synchronize()
    finally:
        fx.lib.fimexcu_set_staged_kernels(-1)
    assert _eq(a[:, :, :ox].cpu().numpy(), b[:, :, :ox].cpu().numpy())


@pytest.mark.parametrize("method", [ol.BILINEAR, ol.BICUBIC, ol.NEAREST])
def test_output_window_equals_slice_of_full_result(fx, method):
    """getOutputDataSlice (CachedInterpolation.cc:113-130): sliceOutput() of the full result and the windowed apply (which computes
    and copies only the window) give the same cells as the oracle's full result cut the same way -- host and device forms"""
    import torch
    rng = np.random.default_rng(300 + method)
    in_x, in_y, out_x, out_y, nz = 90, 70, 233, 151, 5
    px, py = _tables(rng, in_x, in_y, out_x * out_y)
    fld = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    fld[rng.random(fld.shape) < 0.03] = np.nan
    want, _ = ol.o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, fld)
    ci = fx.createCachedInterpolation("x", "y", method, px, py, in_x, in_y, out_x, out_y)
    sub = ci.reduceInput(fld, in_x, in_y)
    full, _ = ci.interpolateValues(sub)
    for (x0, xs, y0, ys) in [(0, out_x, 0, out_y), (17, 64, 9, 40), (200, 33, 150, 1), (5, 1, 0, out_y), (232, 1, 150, 1)]:
        w = want[:, y0:y0 + ys, x0:x0 + xs]
        assert _eq(ci.sliceOutput(full, x0, xs, y0, ys), w)
        assert _eq(ci.interpolateValuesWindow(sub, x0, xs, y0, ys), w)
        d = ci.interpolateValuesWindow(torch.from_numpy(sub).cuda(), x0, xs, y0, ys, fillValue=-7.0)
        torch.cuda.synchronize()
        wf = w.copy()
        wf[np.isnan(wf)] = -7.0
        assert _eq(d.cpu().numpy(), wf)
    with pytest.raises(fx.FimexCudaError):
        ci.interpolateValuesWindow(sub, 200, 64, 0, 10)   # reaches beyond the grid
    with pytest.raises(fx.FimexCudaError):
        ci.sliceOutput(full, 0, out_x + 1, 0, 1)


@pytest.mark.parametrize("shape", [(300, 200, 1400, 900, 9), (128, 96, 700, 500, 65), (128, 96, 700, 500, 170)])
@pytest.mark.parametrize("special", [False, True])
def test_bicubic_staged_kernel_padded_layout(fx, oracle, shape, special):
    """bicubic on padded device slabs: bit-exact (direct kernel) incl. the fused bad/fill conversions; the opt-in f32 variant (staged
    kernel, 32 x 20 TMA box, tiles that do not fit included): same undefined mask, values within 1e-5 of the largest tap"""
    import torch
    in_x, in_y, out_x, out_y, nz = shape
    rng = np.random.default_rng(sum(shape) + 1)
    gx, gy = np.meshgrid(np.arange(out_x), np.arange(out_y))
    px = -3.0 + gx * ((in_x + 5.0) / out_x) + 0.11 * gy * (in_x / out_x)
    py = -2.5 + gy * ((in_y + 4.0) / out_y) - 0.07 * gx * (in_y / out_y)
    k = out_y // 2
    px[k:k + 9] = rng.uniform(-2, in_x + 1, px[k:k + 9].shape)
    py[k:k + 9] = rng.uniform(-2, in_y + 1, py[k:k + 9].shape)
    fld = rng.normal(280, 10, size=(nz, in_y, in_x)).astype(np.float32)
    bad, fill = np.float32(-999.0), np.float32(1e20)
    fld[rng.random(fld.shape) < 0.01] = bad if special else np.nan
    clean = fld.copy()
    if special:
        oracle.orc_bad2nanf(fp(clean), clean.size, bad)
    want, info = ol.o_cached_apply(ol.BICUBIC, px, py, in_x, in_y, out_x, out_y, clean)
    wf = want.copy()
    if special:
        oracle.orc_nanf2bad(fp(wf), wf.size, fill)
    ci = fx.createCachedInterpolation("x", "y", ol.BICUBIC, px, py, in_x, in_y, out_x, out_y)
    sub = torch.from_numpy(ci.reduceInput(fld, in_x, in_y)).cuda()
    kw = dict(badValue=float(bad), fillValue=float(fill)) if special else {}
    pin = ci.toPadded(sub)
    out = ci.interpolateValuesPadded(pin, **kw)
    torch.cuda.synchronize()
    assert _eq(out[:, :, :out_x].cpu().numpy(), wf)
    assert (~np.isnan(want)).mean() > 0.5
    ci.setFastBicubic(True)
    fast = ci.interpolateValuesPadded(pin, **kw)[:, :, :out_x].cpu().numpy()
    ci.setFastBicubic(False)
    ok = ~np.isnan(want)
    if special:  # undefined where the reference is, and there the fill value
        assert np.array_equal(fast == fill, ~ok) and not np.isnan(fast).any()
    else:
        assert np.array_equal(np.isnan(fast), ~ok)
    assert np.abs(fast[ok] - want[ok]).max() <= 1e-5 * np.nanmax(np.abs(clean))
    assert not np.array_equal(fast[ok], want[ok])  # it really is the other arithmetic


def test_nearest_staged_kernel_padded_layout(fx, oracle):
    """nearest-neighbour tables through the TMA-staged kernel (padded layout) and the direct kernel (dense layout)"""
    import torch
    rng = np.random.default_rng(31)
    in_x, in_y, out_x, out_y, nz = 300, 200, 1400, 900, 7
    gx, gy = np.meshgrid(np.arange(out_x), np.arange(out_y))
    px = -3.0 + gx * ((in_x + 5.0) / out_x) + 0.11 * gy * (in_x / out_x)
    py = -2.5 + gy * ((in_y + 4.0) / out_y) - 0.07 * gx * (in_y / out_y)
    k = out_y // 3
    px[k:k + 9] = rng.uniform(-2, in_x + 1, px[k:k + 9].shape)
    fld = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    fld[rng.random(fld.shape) < 0.02] = np.nan
    want, info = ol.o_cached_apply(ol.NEAREST, px, py, in_x, in_y, out_x, out_y, fld)
    ci = fx.CachedNNInterpolation("x", "y", px, py, in_x, in_y, out_x, out_y)
    sub = torch.from_numpy(ci.reduceInput(fld, in_x, in_y)).cuda()
    out = ci.interpolateValuesPadded(ci.toPadded(sub))
    torch.cuda.synchronize()
    assert _eq(out[:, :, :out_x].cpu().numpy(), want)
    dense, _ = ci.interpolateValues(sub, fillValue=-5.0)
    w2 = want.copy()
    w2[np.isnan(w2)] = -5.0
    assert _eq(dense.cpu().numpy(), w2)


@pytest.mark.parametrize("nz", [5, 90])  # 90: two level chunks per tile
@pytest.mark.parametrize("special", [False, True])
def test_fused_uv_staged_kernel_padded_layout(fx, oracle, special, nz):
    """u,v interpolation + rotation as one TMA-staged kernel (padded layout) == the three separate reference steps"""
    import torch
    from fimex_b200.interpolation import interpolateVectorValuesPadded
    rng = np.random.default_rng(77)
    in_x, in_y, out_x, out_y = 300, 200, 1400, 900
    gx, gy = np.meshgrid(np.arange(out_x), np.arange(out_y))
    px = -3.0 + gx * ((in_x + 5.0) / out_x) + 0.11 * gy * (in_x / out_x)
    py = -2.5 + gy * ((in_y + 4.0) / out_y) - 0.07 * gx * (in_y / out_y)
    k = out_y // 2
    px[k:k + 9] = rng.uniform(-2, in_x + 1, px[k:k + 9].shape)
    ang = rng.uniform(-np.pi, np.pi, out_x * out_y)
    m3 = np.stack([np.cos(ang), np.sin(ang), np.degrees(ang)], axis=1).copy()
    u = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    v = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    bad, fill = np.float32(-999.0), np.float32(1e20)
    u[rng.random(u.shape) < 0.01] = bad if special else np.nan
    cu, cv = u.copy(), v.copy()
    if special:
        oracle.orc_bad2nanf(fp(cu), cu.size, bad)
    wu, _ = ol.o_cached_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, cu)
    wv, _ = ol.o_cached_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, cv)
    oracle.orc_vector_reproject_values_by_matrix_f(dp(m3), out_x * out_y, fp(wu), fp(wv), nz)
    if special:
        oracle.orc_nanf2bad(fp(wu), wu.size, fill)
        oracle.orc_nanf2bad(fp(wv), wv.size, fill)
    ci = fx.CachedInterpolation("x", "y", ol.BILINEAR, px, py, in_x, in_y, out_x, out_y)
    cvr = fx.CachedVectorReprojection(m3, out_x, out_y)
    up = ci.toPadded(torch.from_numpy(ci.reduceInput(u, in_x, in_y)).cuda())
    vp = ci.toPadded(torch.from_numpy(ci.reduceInput(v, in_x, in_y)).cuda())
    kw = dict(badU=float(bad), badV=float(bad), fillU=float(fill), fillV=float(fill)) if special else {}
    gu, gv = interpolateVectorValuesPadded(ci, cvr, up, vp, **kw)
    torch.cuda.synchronize()
    assert _eq(gu[:, :, :out_x].cpu().numpy(), wu) and _eq(gv[:, :, :out_x].cpu().numpy(), wv)
    # without rotation
    gu, gv = interpolateVectorValuesPadded(ci, None, up, vp)
    wu2, _ = ol.o_cached_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, u if not special else u)
    torch.cuda.synchronize()
    assert _eq(gu[:, :, :out_x].cpu().numpy(), wu2)


# ---------------------------------------------------------------------------------- packed 16-bit slabs
@pytest.mark.parametrize("name,method", [("bilinear", ol.BILINEAR), ("bicubic", ol.BICUBIC), ("nearest", ol.NEAREST)])
@pytest.mark.parametrize("window", ["full", "reduced"])
def test_packed_i16_apply_vs_oracle(fx, name, method, window):
    """a CDM_SHORT variable through CDMInterpolator::getDataSlice (CDMInterpolator.cc:140-151): raw values cast to float, fill ->
    NaN, interpolate, NaN -> fill, lround and the C casts back to short.  Bit-exact int16, host and device entry points."""
    import torch
    rng = np.random.default_rng(31 + method)
    in_x, in_y, out_x, out_y, nz = 83, 61, 97, 71, 5
    n = out_x * out_y
    if window == "full":
        px, py = _tables(rng, in_x, in_y, n)
    else:
        px, py = rng.uniform(20.2, 50.7, n), rng.uniform(10.1, 40.9, n)
    bad = -32767.0
    raw = rng.integers(-30000, 30000, (nz, in_y, in_x)).astype(np.int16)
    raw[rng.random(raw.shape) < 0.03] = np.int16(bad)
    raw[0, :, :] = (np.arange(in_x)[None, :] * 3 + np.arange(in_y)[:, None] * 7).astype(np.int16)  # smooth level: many exact .5 results
    want = ol.o_packed_apply(method, px, py, in_x, in_y, out_x, out_y, raw, bad)
    if method == ol.BILINEAR:  # the reference's out-of-bounds corner strip is NaN here: compare where it is defined
        pass
    ci = fx.createCachedInterpolation("x", "y", method, px, py, in_x, in_y, out_x, out_y)
    sub = ci.reduceInput(raw, in_x, in_y)
    got = ci.interpolateValuesPacked(sub, bad)
    assert got.dtype == np.int16 and np.array_equal(got, want)
    assert (want == np.int16(bad)).any() and (want != np.int16(bad)).any()
    got_dev = ci.interpolateValuesPacked(torch.from_numpy(np.ascontiguousarray(sub)).cuda(), bad)
    torch.cuda.synchronize()
    assert np.array_equal(got_dev.cpu().numpy(), want)


def test_packed_i16_rounding_and_wrap(fx):
    """values chosen so that the float result sits on .5 (half away from zero), just below it, and beyond the short range after
    bicubic overshoot (the reference's casts wrap)"""
    rng = np.random.default_rng(77)
    in_x, in_y, out_x, out_y = 40, 30, 64, 48
    gx, gy = np.meshgrid(np.arange(out_x), np.arange(out_y))
    px, py = 0.5 * gx + 2.25, 0.5 * gy + 1.75
    raw = np.empty((3, in_y, in_x), np.int16)
    raw[0] = (np.arange(in_x)[None, :] + 2 * np.arange(in_y)[:, None]).astype(np.int16) - 50
    raw[1] = rng.choice(np.array([-32768, 32767, 32766, -32766], dtype=np.int16), (in_y, in_x))
    raw[2] = rng.integers(-5, 6, (in_y, in_x)).astype(np.int16)
    for method in (ol.BILINEAR, ol.BICUBIC):
        want = ol.o_packed_apply(method, px, py, in_x, in_y, out_x, out_y, raw, 9999.0)
        ci = fx.createCachedInterpolation("x", "y", method, px, py, in_x, in_y, out_x, out_y)
        got = ci.interpolateValuesPacked(ci.reduceInput(raw, in_x, in_y), 9999.0)
        assert np.array_equal(got, want), method


def test_concurrent_host_threads_share_a_cache(fx):
    """The reference calls getDataSlice from OpenMP threads (src/io/netcdf/NetCDF_CDMWriter.cc:599-618): several host threads
    use ONE cache at the same time through the host-pointer entry points (ctypes releases the GIL); every thread must get the
    bit-exact result and its own error state."""
    import threading
    rng = np.random.default_rng(99)
    in_x, in_y, out_x, out_y, nz = 101, 87, 203, 167, 6
    px, py = _tables(rng, in_x, in_y, out_x * out_y)
    ci = fx.createCachedInterpolation("x", "y", ol.BILINEAR, px, py, in_x, in_y, out_x, out_y)
    fields = [rng.normal(size=(nz, in_y, in_x)).astype(np.float32) for _ in range(6)]
    wants = [ol.o_cached_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, f)[0] for f in fields]
    results, errors = [None] * len(fields), []

    def work(k):
        try:
            for _ in range(5):
                out, _ = ci.interpolateValues(ci.reduceInput(fields[k], in_x, in_y))
                packed = ci.interpolateValuesPacked(np.zeros((1, ci.getInY(), ci.getInX()), np.int16), -32767.0)
                assert packed.shape == (1, out_y, out_x)
            results[k] = out
            if k == 0:  # an error raised in one thread must not leak into the others' error state
                try:
                    fx.createCachedInterpolation("x", "y", 99, px, py, in_x, in_y, out_x, out_y)
                except Exception as exc:
                    assert "unknown interpolation" in str(exc)
        except Exception as exc:  # noqa: BLE001
            errors.append((k, repr(exc)))

    threads = [threading.Thread(target=work, args=(k,)) for k in range(len(fields))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for k in range(len(fields)):
        assert _eq(results[k], wants[k]), k


@pytest.mark.parametrize("method", [ol.BICUBIC, ol.NEAREST])
@pytest.mark.parametrize("rotate", [True, False])
def test_fused_uv_bicubic_nearest_bad_fill_padded(fx, oracle, method, rotate):
    """bicubic / nearest wind pairs in one fused pass with bad -> NaN on load and NaN -> fill AFTER the rotation (which must see NaN,
    not the fill value), in the padded device layout and in the dense one"""
    import torch
    from fimex_b200.interpolation import interpolateVectorValues, interpolateVectorValuesPadded
    rng = np.random.default_rng(5 + method)
    in_x, in_y, out_x, out_y, nz = 70, 50, 131, 97, 5
    ang = rng.uniform(-np.pi, np.pi, out_x * out_y)
    m3 = np.stack([np.cos(ang), np.sin(ang), np.degrees(ang)], axis=1).copy()
    px, py = _tables(rng, in_x, in_y, out_x * out_y)
    bad_u, bad_v, fill_u, fill_v = np.float32(-999.0), np.float32(-888.0), np.float32(1e30), np.float32(-1e30)
    u = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    v = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    un, vn = u.copy(), v.copy()
    mu, mv = rng.random(u.shape) < 0.02, rng.random(v.shape) < 0.02
    un[mu], vn[mv] = np.nan, np.nan
    u[mu], v[mv] = bad_u, bad_v
    wu, _ = ol.o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, un)
    wv, _ = ol.o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, vn)
    if rotate:
        oracle.orc_vector_reproject_values_by_matrix_f(dp(m3), out_x * out_y, fp(wu), fp(wv), nz)
    oracle.orc_nanf2bad(fp(wu), wu.size, fill_u)
    oracle.orc_nanf2bad(fp(wv), wv.size, fill_v)
    ci = fx.createCachedInterpolation("x", "y", method, px, py, in_x, in_y, out_x, out_y)
    cvr = fx.CachedVectorReprojection(m3, out_x, out_y) if rotate else None
    ru, rv = ci.reduceInput(u, in_x, in_y), ci.reduceInput(v, in_x, in_y)
    gu, gv = interpolateVectorValues(ci, cvr, ru, rv, float(bad_u), float(bad_v), float(fill_u), float(fill_v))
    assert _eq(gu, wu) and _eq(gv, wv)
    du, dv = ci.toPadded(torch.from_numpy(ru).cuda()), ci.toPadded(torch.from_numpy(rv).cuda())
    pu, pv = interpolateVectorValuesPadded(ci, cvr, du, dv, float(bad_u), float(bad_v), float(fill_u), float(fill_v))
    torch.cuda.synchronize()
    assert _eq(pu[..., :out_x].cpu().numpy(), wu) and _eq(pv[..., :out_x].cpu().numpy(), wv)


# ---------------------------------------------------------------------------------- host results: only the row bands that can hold data
@pytest.mark.parametrize("method", [ol.BILINEAR, ol.BICUBIC, ol.NEAREST])
@pytest.mark.parametrize("fill", [float("nan"), -999.5])
def test_host_result_comes_back_in_row_bands(fx, method, fill):
    """A target much larger than the source, rotated against it: the defined part of every target row is a different span.  The
    host entry point copies back only the bands' rectangles and writes the undefined value elsewhere on the host
    (fimexcu_cache_row_bands); the result must be the reference's, cell for cell, for pageable and for pinned result arrays."""
    rng = np.random.default_rng(400 + method)
    in_x, in_y, out_x, out_y, nz = 120, 90, 1100, 700, 5   # 15.4 MB result: above the 8 MB threshold of the banded copy
    gx, gy = np.meshgrid(np.arange(out_x, dtype=np.float64), np.arange(out_y, dtype=np.float64))
    ang = np.radians(24.0)
    cx, cy = out_x / 2.0, out_y / 2.0
    px = in_x / 2.0 + 0.16 * (np.cos(ang) * (gx - cx) - np.sin(ang) * (gy - cy))
    py = in_y / 2.0 + 0.16 * (np.sin(ang) * (gx - cx) + np.cos(ang) * (gy - cy)) + 2e-5 * (gx - cx) ** 2
    fld = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    fld[rng.random(fld.shape) < 0.02] = np.nan
    ci = fx.createCachedInterpolation("x", "y", method, px, py, in_x, in_y, out_x, out_y)
    bands, frac = ci.rowBands()
    assert 1 <= len(bands) <= 24 and frac < 0.9, (len(bands), frac)
    assert bands[0][0] == 0 and bands[-1][1] == out_y and all(a[1] == b[0] for a, b in zip(bands, bands[1:]))
    want, _ = ol.o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, fld)
    if not np.isnan(fill):
        want = np.where(np.isnan(want), np.float32(fill), want)
    # every cell outside the bands must be undefined in the reference's result too
    mask = np.zeros((out_y, out_x), bool)
    for y0, y1, x0, x1 in bands:
        mask[y0:y1, x0:x1] = x0 < x1
    outside = want[:, ~mask]
    assert np.isnan(outside).all() if np.isnan(fill) else (outside == np.float32(fill)).all()
    sub = ci.reduceInput(fld, in_x, in_y)
    got, _ = ci.interpolateValues(sub, fillValue=fill)          # pageable in, pageable out
    assert _eq(got.reshape(want.shape), want)
    lib = fx.lib
    h_out = lib.fimexcu_host_alloc(want.size * 4)
    assert h_out
    C.memset(h_out, 0x5a, want.size * 4)                         # stale content must not survive anywhere
    new = C.c_size_t()
    fx._lib.check(lib.fimexcu_cache_apply_f(ci._h, sub.ctypes.data, sub.size, h_out, C.byref(new), float("nan"), fill), "apply")
    a = np.frombuffer((C.c_float * want.size).from_address(h_out), dtype=np.float32).reshape(want.shape)
    assert new.value == want.size and _eq(a, want)
    lib.fimexcu_host_free(h_out)


def test_host_packed_result_comes_back_in_row_bands(fx):
    rng = np.random.default_rng(77)
    in_x, in_y, out_x, out_y, nz = 120, 90, 1500, 900, 4       # 10.8 MB of int16
    gx, gy = np.meshgrid(np.arange(out_x, dtype=np.float64), np.arange(out_y, dtype=np.float64))
    px = -40.0 + 0.14 * gx + 0.03 * gy
    py = -30.0 + 0.17 * gy - 0.02 * gx
    bad = -32767.0
    raw = rng.integers(-30000, 30000, (nz, in_y, in_x)).astype(np.int16)
    raw[rng.random(raw.shape) < 0.03] = np.int16(bad)
    ci = fx.createCachedInterpolation("x", "y", ol.BILINEAR, px, py, in_x, in_y, out_x, out_y)
    bands, frac = ci.rowBands()
    assert bands and frac < 0.9
    want = ol.o_packed_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, raw, bad)
    got = ci.interpolateValuesPacked(ci.reduceInput(raw, in_x, in_y), bad)
    assert got.dtype == np.int16 and np.array_equal(got, want)


@pytest.mark.parametrize("rotate", [True, False])
@pytest.mark.parametrize("fills", [(float("nan"), float("nan")), (-999.5, 1e20)])
def test_host_vector_result_comes_back_in_row_bands(fx, oracle, rotate, fills):
    """the vector branch through the host entry point (fimexcu_cache_apply_uv_f): level chunks of both components through the fused
    kernel, row bands on the way back, per-component fill values; several chunks (34 MB of results)"""
    from fimex_b200.interpolation import interpolateVectorValues
    rng = np.random.default_rng(500 + int(rotate))
    in_x, in_y, out_x, out_y, nz = 110, 80, 1000, 600, 7
    gx, gy = np.meshgrid(np.arange(out_x, dtype=np.float64), np.arange(out_y, dtype=np.float64))
    px = -25.0 + 0.16 * gx + 0.02 * gy
    py = -12.0 + 0.17 * gy - 0.015 * gx
    ang = rng.uniform(-np.pi, np.pi, out_x * out_y)
    m3 = np.stack([np.cos(ang), np.sin(ang), np.degrees(ang)], axis=1).copy()
    u = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    v = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    u[rng.random(u.shape) < 0.02] = np.nan
    ci = fx.createCachedInterpolation("x", "y", ol.BILINEAR, px, py, in_x, in_y, out_x, out_y)
    bands, frac = ci.rowBands()
    assert bands and frac < 0.9
    cvr = fx.CachedVectorReprojection(m3, out_x, out_y) if rotate else None
    wu, _ = ol.o_cached_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, u)
    wv, _ = ol.o_cached_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, v)
    if rotate:
        oracle.orc_vector_reproject_values_by_matrix_f(dp(m3), out_x * out_y, fp(wu), fp(wv), nz)
    fu, fv = fills
    if not np.isnan(fu):
        wu = np.where(np.isnan(wu), np.float32(fu), wu)
        wv = np.where(np.isnan(wv), np.float32(fv), wv)
    su, sv = ci.reduceInput(u, in_x, in_y), ci.reduceInput(v, in_x, in_y)
    gu, gv = interpolateVectorValues(ci, cvr, su, sv, fillU=fu, fillV=fv)   # pageable arrays on both sides
    assert _eq(gu, wu) and _eq(gv, wv)
    lib = fx.lib
    n_out = wu.size
    hu, hv = lib.fimexcu_host_alloc(n_out * 4), lib.fimexcu_host_alloc(n_out * 4)
    assert hu and hv
    C.memset(hu, 0x5a, n_out * 4)
    C.memset(hv, 0x5a, n_out * 4)
    new = C.c_size_t()
    fx._lib.check(lib.fimexcu_cache_apply_uv_f(ci._h, cvr._h if cvr else None, su.ctypes.data, sv.ctypes.data, su.size, hu, hv, C.byref(new),
                                               float("nan"), float("nan"), fu, fv), "apply_uv")
    au = np.frombuffer((C.c_float * n_out).from_address(hu), dtype=np.float32).reshape(wu.shape)
    av = np.frombuffer((C.c_float * n_out).from_address(hv), dtype=np.float32).reshape(wv.shape)
    assert new.value == n_out and _eq(au, wu) and _eq(av, wv)
    lib.fimexcu_host_free(hu)
    lib.fimexcu_host_free(hv)


@pytest.mark.parametrize("pitch_extra", [0, 21, 29])  # dense, rows on 16 bytes (float4 path), odd pitch (scalar path)
def test_vector_rotation_on_padded_slabs(fx, oracle, pitch_extra):
    """K7 on slabs with a row pitch: == vector_reproject_values_by_matrix_f on the cells, pad elements untouched"""
    import torch
    rng = np.random.default_rng(5 + pitch_extra)
    ox, oy, nz = 2891 // 7, 131, 11  # 413 columns: not a multiple of four
    pitch = ox + pitch_extra
    if pitch_extra == 21:
        pitch = (ox + 31) // 32 * 32
    ang = rng.uniform(-np.pi, np.pi, ox * oy)
    m3 = np.stack([np.cos(ang), np.sin(ang), np.degrees(ang)], axis=1).copy()
    u = rng.normal(0, 8, (nz, oy, ox)).astype(np.float32)
    v = rng.normal(0, 8, (nz, oy, ox)).astype(np.float32)
    u[rng.random(u.shape) < 0.01] = np.nan
    wu, wv = u.copy(), v.copy()
    oracle.orc_vector_reproject_values_by_matrix_f(dp(m3), ox * oy, fp(wu), fp(wv), nz)
    cvr = fx.CachedVectorReprojection(m3, ox, oy)
    up = torch.full((nz, oy, pitch), 777.0, dtype=torch.float32, device="cuda")
    vp = torch.full((nz, oy, pitch), 888.0, dtype=torch.float32, device="cuda")
    up[:, :, :ox] = torch.from_numpy(u).cuda()
    vp[:, :, :ox] = torch.from_numpy(v).cuda()
    cvr.reprojectValuesPadded(up, vp)
    torch.cuda.synchronize()
    assert _eq(up[:, :, :ox].cpu().numpy(), wu) and _eq(vp[:, :, :ox].cpu().numpy(), wv)
    assert bool((up[:, :, ox:] == 777.0).all()) and bool((vp[:, :, ox:] == 888.0).all())
    # a view into a larger allocation that is NOT on a 16-byte boundary: the scalar path must take it
    big = torch.full((nz * oy * pitch + 3,), 5.0, dtype=torch.float32, device="cuda")
    uq = big[1:1 + nz * oy * pitch].view(nz, oy, pitch)
    uq[:, :, :ox] = torch.from_numpy(u).cuda()
    vq = vp.clone()
    vq[:, :, :ox] = torch.from_numpy(v).cuda()
    fx.lib.fimexcu_vecrot_apply_dev_strided_f(cvr._h, uq.data_ptr(), vq.data_ptr(), pitch, oy * pitch, nz, None)
    torch.cuda.synchronize()
    assert _eq(uq[:, :, :ox].cpu().numpy(), wu) and _eq(vq[:, :, :ox].cpu().numpy(), wv)
    assert float(big[0]) == 5.0 and bool((big[-2:] == 5.0).all())
