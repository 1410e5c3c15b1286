"""GPU tests of the cache build: lattice reprojection (K1), index tables (K2), rotation matrices (K8), through the C ABI.

Checker: oracle/miniproj.c (CPU restatement of the published projection formulas; pinned against stored PROJ output for
ob_tran / lcc / stere on the sphere in tests/test_proj_pins.py, unpinned for the other forms, see oracle/miniproj.h) and analytic identities.  Tolerances: 1e-6 m / 1e-9 deg on coordinates (device vs glibc libm
differ in the last bits), 1e-9 cells on index tables, 1e-9 rad on rotation angles."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import dp

pytestmark = pytest.mark.gpu

PROJS = {
    "meps": "+proj=lcc +lat_0=63.3 +lon_0=15 +lat_1=63.3 +lat_2=63.3 +no_defs +R=6.371e+06",
    "arome_arctic": "+proj=lcc +lat_0=77.5 +lon_0=335 +lat_1=77.5 +lat_2=77.5 +no_defs +R=6.371e+06",
    "lcc_2sp_wgs84": "+proj=lcc +lat_1=60 +lat_2=70 +lat_0=65 +lon_0=10 +x_0=1000 +y_0=-2000 +ellps=WGS84",
    "norkyst": "+proj=stere +ellps=WGS84 +lat_0=90.0 +lat_ts=60.0 +x_0=3192800 +y_0=1784000 +lon_0=70",
    "stere_sphere": "+proj=stere +lat_0=90 +lon_0=-32 +lat_ts=60 +ellps=sphere +a=6371000 +e=0",
    "stere_oblique": "+proj=stere +lat_0=55 +lon_0=12 +R=6371000",
    "stere_oblique_ell": "+proj=stere +lat_0=55 +lon_0=12 +ellps=GRS80 +k=0.9999",
    "hirlam": "+proj=ob_tran +o_proj=longlat +lon_0=-40 +o_lat_p=22 +R=6.371e+06 +no_defs",
    "merc": "+proj=merc +lon_0=5 +lat_ts=60 +ellps=WGS84",
    "merc_sphere": "+proj=merc +R=6371000",
    "utm33": "+proj=utm +zone=33 +datum=WGS84 +units=m +no_defs",
    "tmerc_sphere": "+proj=tmerc +lat_0=60 +lon_0=15 +R=6371000 +k=0.9996",
    "laea_eu": "+proj=laea +lat_0=52 +lon_0=10 +x_0=4321000 +y_0=3210000 +ellps=GRS80 +units=m",
    "laea_polar": "+proj=laea +lat_0=90 +lon_0=0 +R=6371000",
}
LL = "+proj=latlong +R=6.371e+06 +no_defs"
LL_WGS = "+proj=latlong +datum=WGS84 +towgs84=0,0,0"


@pytest.fixture(scope="module")
def rp():
    import fimex_b200
    assert fimex_b200.lib.fimexcu_device_count() >= 1
    from fimex_b200 import reproject
    return reproject


def _oracle_transform(src, dst, x, y):
    xv, yv = np.array(x, dtype=np.float64).ravel().copy(), np.array(y, dtype=np.float64).ravel().copy()
    err = C.create_string_buffer(200)
    rc = ol.oracle().miniproj_transform_str(src.encode(), dst.encode(), xv.size, dp(xv), dp(yv), err, 200)
    assert rc >= 0, err.value
    return xv, yv


@pytest.mark.parametrize("name", sorted(PROJS))
def test_lonlat_to_projection_and_back(rp, name):
    rng = np.random.default_rng(1000 + sorted(PROJS).index(name))  # fixed per projection (hash() varies with PYTHONHASHSEED)
    lon = rng.uniform(-20, 50, 20000)
    lat = rng.uniform(45, 85, 20000)
    lon[:5] = [15, 15, 0, 179.9, -179.9]
    lat[:5] = [63.3, 90, 60, 70, 70]
    x, y = rp.reproject_values(LL, PROJS[name], lon, lat)
    wx, wy = _oracle_transform(LL, PROJS[name], lon, lat)
    fin = np.isfinite(wx)
    assert np.array_equal(fin, np.isfinite(x))
    tol = 1e-9 if name == "hirlam" else 1e-6
    assert np.abs(x[fin] - wx[fin]).max() < tol and np.abs(y[fin] - wy[fin]).max() < tol
    # and back: round trip to 1e-9 degree
    lon2, lat2 = rp.reproject_values(PROJS[name], LL, x[fin], y[fin])
    dlon = (lon2 - lon[fin] + 180) % 360 - 180
    ok = np.abs(lat[fin]) < 89.999
    rt = 2e-8 if name == "laea_eu" else 1e-9  # the authalic-latitude series of the ellipsoidal laea is itself good to ~1e-8 deg
    assert np.abs(dlon[ok]).max() < rt and np.abs(lat2 - lat[fin]).max() < rt


def test_known_coordinates(rp):
    # UTM33 on the central meridian at 60 N: x = 500000, y = 0.9996 * meridian arc (6 654 072.819 m on WGS84)
    x, y = rp.reproject_values(LL_WGS, PROJS["utm33"], [15.0], [60.0])
    assert abs(x[0] - 500000.0) < 1e-6 and abs(y[0] - 0.9996 * 6654072.819) < 2e-3
    # MEPS corners: (0,0) -> 50.32 N 0.28 E, (948,1068) -> 71.58 N 54.24 E
    from fimex_b200 import synthetic as S
    lon, lat = rp.reproject_values(S.MEPS_PROJ, LL, [S.MEPS_X0, S.MEPS_X0 + 948 * 2500], [S.MEPS_Y0, S.MEPS_Y0 + 1068 * 2500])
    assert abs(lat[0] - 50.3196) < 1e-3 and abs(lon[0] - 0.2783) < 1e-3 and abs(lat[1] - 71.5760) < 1e-3 and abs(lon[1] - 54.2413) < 1e-3
    # sphere -> WGS84 lat/lon is a null datum shift
    lon, lat = rp.reproject_values(LL, LL_WGS, [12.5], [61.25])
    assert lon[0] == 12.5 and lat[0] == 61.25
    # unknown projection -> error with the reference's "Proj error" wording
    import fimex_b200
    with pytest.raises(fimex_b200.FimexCudaError, match="Proj error"):
        rp.reproject_values("+proj=fork +lat_1000=123", LL, [0.0], [0.0])
    # parameters whose effect is not implemented (datum shift, grid shift, prime meridian, axis order, unknown keys) are errors,
    # never silently dropped; their identity forms are accepted
    for bad in ("+proj=longlat +ellps=bessel +towgs84=565.4,50.3,465.6", "+proj=longlat +ellps=clrk66 +nadgrids=conus", "+proj=longlat +a=6371000 +pm=paris",
                "+proj=longlat +R=6371000 +axis=neu", "+proj=longlat +datum=potsdam", "+proj=stere +lat_0=90 +R=6371000 +lon_wrap=180"):
        with pytest.raises(fimex_b200.FimexCudaError, match="unsupported"):
            rp.reproject_values(bad, LL, [10.0], [60.0])
    lon, lat = rp.reproject_values("+proj=longlat +R=6371000 +towgs84=0,0,0,0,0,0,0 +nadgrids=@null +pm=greenwich +axis=enu +no_defs +type=crs", LL, [12.5], [61.25])
    assert lon[0] == 12.5 and lat[0] == 61.25


def test_meps_index_tables_built_on_device(rp):
    """fimexcu_reproject_axes + mifi_points2position against the closed-form table generator and the oracle"""
    import fimex_b200
    from fimex_b200 import synthetic as S
    lon, lat = S.target_axes("T1")
    x, y = rp.reproject_axes(S.LATLON_PROJ, S.MEPS_PROJ, lon, lat)
    mx, my = S.meps_axes()
    px = fimex_b200.mifi_points2position(x, mx, 0)
    py = fimex_b200.mifi_points2position(y, my, 0)
    cx, cy, ox, oy = S.meps_index_tables("T1")
    assert px.shape == (oy, ox)
    assert np.abs(px - cx).max() < 1e-9 and np.abs(py - cy).max() < 1e-9
    wx, wy = np.empty(ox * oy), np.empty(ox * oy)
    ol.oracle().orc_reproject_axes(S.LATLON_PROJ.encode(), S.MEPS_PROJ.encode(), dp(lon), dp(lat), lon.size, lat.size, dp(wx), dp(wy))
    assert np.abs(x.ravel() - wx).max() < 1e-6 and np.abs(y.ravel() - wy).max() < 1e-6


def test_cache_from_projection_equals_cache_from_tables(rp):
    """the one-call device build gives the same regridded values as tables -> createCachedInterpolation"""
    import fimex_b200
    from fimex_b200 import synthetic as S
    lon, lat = S.target_axes("T2")
    mx, my = S.meps_axes()
    ci = rp.cachedInterpolationFromProjection(1, S.LATLON_PROJ, S.MEPS_PROJ, lon, lat, mx, my)
    cx, cy, ox, oy = S.meps_index_tables("T2")
    ref = fimex_b200.CachedInterpolation("x", "y", 1, cx, cy, S.MEPS_NX, S.MEPS_NY, ox, oy)
    assert (ci.getInX(), ci.getInY()) == (ref.getInX(), ref.getInY()) == (505, 642)
    assert (ci.reducedDomain().xMin, ci.reducedDomain().yMin) == (172, 205)
    fld = S.meps_slab_numpy(t=1, nz=3)
    a, _ = ci.interpolateValues(ci.reduceInput(fld, S.MEPS_NX, S.MEPS_NY))
    b, _ = ref.interpolateValues(ref.reduceInput(fld, S.MEPS_NX, S.MEPS_NY))
    assert not np.isnan(a).any()
    assert np.allclose(a, b, rtol=1e-5, atol=0)  # tables agree to 1e-12 cells; weights to 1 ulp of f32


@pytest.mark.parametrize("src,dst,kind", [("meps", LL, "deg"), ("meps", "stere_sphere", "m"), ("hirlam", "norkyst", "m"), ("utm33", LL_WGS, "deg"),
                                          ("norkyst", "meps", "m")])
def test_vector_reproject_matrix_vs_oracle(rp, src, dst, kind):
    ps = PROJS[src]
    pd = PROJS.get(dst, dst)
    if kind == "deg":
        ax, ay = np.arange(0., 30., 0.5), np.arange(55., 72., 0.5)
    else:
        cx, cy = _oracle_transform(LL, pd, [15.0], [65.0])
        ax, ay = cx[0] + np.arange(-40, 40) * 10000.0, cy[0] + np.arange(-30, 30) * 10000.0
    m = rp.get_vector_reproject_matrix(ps, pd, ax, ay)
    want = np.empty(ax.size * ay.size * 3)
    assert ol.oracle().orc_get_vector_reproject_matrix(ps.encode(), pd.encode(), dp(ax), dp(ay), ax.size, ay.size, dp(want)) == 0
    want = want.reshape(ay.size, ax.size, 3)
    assert np.abs(m[..., 0] - want[..., 0]).max() < 1e-7 and np.abs(m[..., 1] - want[..., 1]).max() < 1e-7
    d = (m[..., 2] - want[..., 2] + 180) % 360 - 180
    assert np.abs(d).max() < 1e-5
    assert np.abs(m[..., 0] ** 2 + m[..., 1] ** 2 - 1).max() < 1e-12
    # points variant
    gx, gy = np.meshgrid(ax, ay)
    mp = rp.get_vector_reproject_matrix_points(ps, pd, 0 if src == "hirlam" else 1, gx, gy)
    wantp = np.empty(gx.size * 3)
    assert ol.oracle().orc_get_vector_reproject_matrix_points(ps.encode(), pd.encode(), 0 if src == "hirlam" else 1, dp(np.ascontiguousarray(gx)),
                                                              dp(np.ascontiguousarray(gy)), ax.size, ay.size, dp(wantp)) == 0
    dp_ = (mp[..., 2].ravel() - wantp.reshape(-1, 3)[:, 2] + 180) % 360 - 180
    assert np.abs(dp_).max() < 1e-3


def test_rotation_90_degrees_between_polar_stereographic_grids(rp):
    """test/testInterpolation.cc:387-470: stere lon_0=0 -> lon_0=90 turns (u, v) into (v, -u)"""
    import fimex_b200
    a = "+proj=stere +lat_0=90 +lon_0=0 +lat_ts=60 +ellps=sphere +a=6371000 +e=0"
    b = "+proj=stere +lat_0=90 +lon_0=90 +lat_ts=60 +ellps=sphere +a=6371000 +e=0"
    ax = np.arange(-5, 6) * 50000.0
    ay = np.arange(-4, 5) * 50000.0 + 1e6
    m = rp.get_vector_reproject_matrix(a, b, ax, ay)
    cvr = fimex_b200.CachedVectorReprojection(m)
    u = np.full((1, ay.size, ax.size), 3.0, np.float32)
    v = np.full((1, ay.size, ax.size), -2.0, np.float32)
    cvr.reprojectValues(u, v)
    assert np.abs(u - (-2.0)).max() < 1e-4 and np.abs(v - (-3.0)).max() < 1e-4


@pytest.mark.parametrize("tag", ["latlong", "stere"])
def test_config0_hirlam12_end_to_end(rp, tag):
    """BASELINE configs[0]: test/hirlam12.nc -> bilinear, the whole device path (cache built from the projection strings,
    scalars through the cache, x_wind/y_wind through the fused interpolate+rotate kernel) against the values produced by
    the reference's own objects (tests/golden/make_golden.py:hirlam12_fixture)."""
    import os
    import fimex_b200
    from fimex_b200.interpolation import interpolateVectorValues
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_hirlam12.npz"))
    proj = bytes(g[f"{tag}_proj"]).decode()
    src = "+proj=latlong +R=6371000 +no_defs"
    xc, yc, ax, ay = g["xc"], g["yc"], g[f"{tag}_ax"], g[f"{tag}_ay"]
    ci = rp.cachedInterpolationFromProjection(1, proj, src, ax, ay, xc, yc, 1, 2)
    info = g[f"{tag}_info"]
    assert (ci.getInX(), ci.getInY()) == (info[0], info[1])
    m = rp.get_vector_reproject_matrix(src, proj, ax, ay)
    assert np.abs(m.reshape(-1, 3)[:, :2] - g[f"{tag}_matrix"].reshape(-1, 3)[:, :2]).max() < 1e-7
    cvr = fimex_b200.CachedVectorReprojection(m)
    exact = tag == "latlong"  # identical tables -> bit-identical values; stere: tables agree to 1e-12 cells -> 1e-5 relative
    # the reference's out-of-bounds corner strip (interpolation.c:322: x within half a cell outside AND y within half a cell of
    # the last row reads past the level; its fixture values there are whatever memory held) is excluded
    tx, ty = g[f"{tag}_px"] - info[3], g[f"{tag}_py"] - info[4]
    strip = ((ty >= info[1] - 0.5) & (ty < info[1] + 0.5) & ~((np.floor(tx) >= 0) & (np.floor(tx) + 1 < info[0]))).reshape(ay.size, ax.size)
    for name in ("geopotential_height", "air_potential_temperature"):
        fld = g["in_" + name].reshape(-1, yc.size, xc.size)
        got, _ = ci.interpolateValues(ci.reduceInput(fld, xc.size, yc.size))
        want = g[f"{tag}_ex_{name}"]
        assert np.array_equal(np.isnan(got)[:, ~strip], np.isnan(want)[:, ~strip])
        k = ~np.isnan(want) & ~strip[None]
        assert np.array_equal(got[k], want[k]) if exact else np.allclose(got[k], want[k], rtol=1e-5, atol=0)
    u = g["in_x_wind"].reshape(-1, yc.size, xc.size)
    v = g["in_y_wind"].reshape(-1, yc.size, xc.size)
    gu, gv = interpolateVectorValues(ci, cvr, ci.reduceInput(u, xc.size, yc.size), ci.reduceInput(v, xc.size, yc.size))
    for got, want in ((gu, g[f"{tag}_ex_x_wind"]), (gv, g[f"{tag}_ex_y_wind"])):
        assert np.array_equal(np.isnan(got)[:, ~strip], np.isnan(want)[:, ~strip])
        k = ~np.isnan(want) & ~strip[None]
        assert np.allclose(got[k], want[k], rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("method", [ol.BILINEAR, ol.BICUBIC, ol.NEAREST])
def test_point_list_and_cross_section_extraction(rp, method):
    """CDMInterpolator::changeProjection(method, lonVals, latVals) / ...ToLatLonTemplate / ...ToCrossSections
    (src/CDMInterpolator.cc:457-613, 1580-1666): a list of lon/lat points (here: two great-circle-ish sections sampled like
    the reference does, plus isolated stations, some outside the domain) as the target of the cached interpolation, built on
    the device in one call and compared with oracle transform -> points2position -> cached apply."""
    from fimex_b200 import synthetic as S
    rng = np.random.default_rng(method)
    lon = np.concatenate([np.linspace(5.0, 25.0, 301), np.linspace(-20.0, 60.0, 257), rng.uniform(-30, 70, 200)])
    lat = np.concatenate([np.linspace(58.0, 70.5, 301), np.linspace(48.0, 76.0, 257), rng.uniform(45, 80, 200)])
    mx, my = S.meps_axes()
    ci = rp.cachedInterpolationForPoints(method, S.MEPS_PROJ, lon, lat, mx, my)
    assert (ci.getOutX(), ci.getOutY()) == (lon.size, 1)
    x, y = _oracle_transform(rp.LAT_LON_PROJSTR, S.MEPS_PROJ, lon, lat)
    o = ol.oracle()
    o.orc_points2position(dp(x), x.size, dp(mx), mx.size, ol.PROJ_AXIS)
    o.orc_points2position(dp(y), y.size, dp(my), my.size, ol.PROJ_AXIS)
    fld = S.meps_slab_numpy(t=2, nz=4)
    want, info = ol.o_cached_apply(method, x, y, S.MEPS_NX, S.MEPS_NY, lon.size, 1, fld)
    got, _ = ci.interpolateValues(ci.reduceInput(fld, S.MEPS_NX, S.MEPS_NY))
    assert np.array_equal(np.isnan(got), np.isnan(want)) and np.isnan(want).any() and (~np.isnan(want)).any()
    ok = ~np.isnan(want)
    if method == ol.NEAREST:
        assert (got[ok] != want[ok]).mean() < 1e-3  # a cell boundary hit to 1e-12 cells may round the other way
    else:
        assert np.allclose(got[ok], want[ok], rtol=1e-5, atol=1e-4)
    # a 2-D lat/lon template is the same call with a shape
    ci2 = rp.cachedInterpolationForPoints(method, S.MEPS_PROJ, lon[:600], lat[:600], mx, my, out_shape=(20, 30))
    got2, _ = ci2.interpolateValues(ci2.reduceInput(fld, S.MEPS_NX, S.MEPS_NY))
    ok2 = ~np.isnan(got2.reshape(4, -1))
    assert got2.shape == (4, 20, 30) and np.allclose(got2.reshape(4, -1)[ok2], got[:, 0, :600][ok2], rtol=1e-6)


def test_device_resident_reprojection_entry_points(rp):
    """fimexcu_reproject_axes_dev / fimexcu_reproject_values_dev / fimexcu_synchronize: the same numbers as the host-pointer entry
    points, with the results left on the device"""
    import torch
    from fimex_b200 import _lib, synthetic as S
    lib = _lib.lib
    lon, lat = S.target_axes("T2")
    x, y = rp.reproject_axes(S.LATLON_PROJ, S.MEPS_PROJ, lon, lat)
    dx = torch.empty((lat.size, lon.size), dtype=torch.float64, device="cuda")
    dy = torch.empty_like(dx)
    _lib.check(lib.fimexcu_reproject_axes_dev(S.LATLON_PROJ.encode(), S.MEPS_PROJ.encode(), lon.ctypes.data, lat.ctypes.data, lon.size, lat.size,
                                              dx.data_ptr(), dy.data_ptr(), _lib.current_stream_ptr()), "reproject_axes_dev")
    _lib.check(lib.fimexcu_synchronize(), "synchronize")
    assert np.array_equal(dx.cpu().numpy(), x) and np.array_equal(dy.cpu().numpy(), y)
    # points back to lon/lat on the device, in place
    _lib.check(lib.fimexcu_reproject_values_dev(S.MEPS_PROJ.encode(), S.LATLON_PROJ.encode(), dx.data_ptr(), dy.data_ptr(), dx.numel(),
                                                _lib.current_stream_ptr()), "reproject_values_dev")
    _lib.check(lib.fimexcu_synchronize(), "synchronize")
    bx, by = rp.reproject_values(S.MEPS_PROJ, S.LATLON_PROJ, x.ravel().copy(), y.ravel().copy())
    assert np.array_equal(dx.cpu().numpy().ravel(), bx) and np.array_equal(dy.cpu().numpy().ravel(), by)
    lon2, lat2 = np.meshgrid(lon, lat)
    assert np.abs(dx.cpu().numpy() - lon2).max() < 1e-9 and np.abs(dy.cpu().numpy() - lat2).max() < 1e-9
