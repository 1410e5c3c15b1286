"""Coordinate-based nearest neighbour (coord_kdtree K9, coord_nearestneighbor K10).

Bar: the source cell of EVERY target equals the reference's -- np.array_equal, no tolerance, no whitelist -- including what follows
from nanoflann's tree layout (which of several exactly equidistant duplicates wins; which matches a cloud with NaN coordinates
loses) and from the reference's latitude-sorted scan order.  Checkers: the reference's vendored nanoflann driven as
CDMInterpolator.cc:953-1028 drives it (oracle/_ref: ref_kdtree_translate) and its coord_nearestneighbor scan restated on the same
libstdc++ calls (ref_coordnn_translate).
CPU part: the host-side logic of the product (tree build + walk, sorted scan; tests/native harnesses) and the tree-free oracle.
GPU part: libfimexcu through the C ABI."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import dp


def swath(rng, nx, ny, nan_frac=0.005, wrap_column=False):
    """curvilinear lat/lon of a tilted swath, with NaN coordinates at the edges"""
    u = np.linspace(-1, 1, nx)[None, :]
    v = np.linspace(0, 1, ny)[:, None]
    lat = 55 + 20 * v + 2.5 * u + 0.3 * np.sin(6 * v) * u
    lon = 5 + 18 * u / np.cos(np.radians(lat)) * 0.5 + 9 * v
    lat = lat + rng.normal(0, 1e-4, lat.shape)
    lon = lon + rng.normal(0, 1e-4, lon.shape)
    edge = (np.abs(u) > 0.97) & (rng.random(lat.shape) < nan_frac * 30)
    lat[edge] = np.nan
    if wrap_column:
        lon[:, -1], lat[:, -1] = lon[:, 0], lat[:, 0]  # duplicated positions: exact ties
    return np.ascontiguousarray(lon), np.ascontiguousarray(lat)


def targets(rng, n):
    return rng.uniform(-12, 32, n), rng.uniform(52, 78, n)


def test_oracle_kdtree_restatement_matches_nanoflann(oracle, ref):
    """NaN-free cloud: the tree-free restatement and the reference's nanoflann agree on every target.  With NaN coordinates
    in the cloud (which the reference keeps, CDMInterpolator.cc:982-986) nanoflann's splits compare against NaN and the search
    then misses points that do lie inside the radius -- a reference defect that is not reproduced."""
    rng = np.random.default_rng(4)
    lon, lat = swath(rng, 61, 83, nan_frac=0.0)
    qx, qy = targets(rng, 4000)
    for maxdist in (20000.0, 60000.0):
        ax, ay, gap = qx.copy(), qy.copy(), np.empty(qx.size)
        oracle.orc_kdtree_translate(maxdist, dp(ax), dp(ay), ax.size, dp(lon), dp(lat), 61, 83, dp(gap))
        bx, by = qx.copy(), qy.copy()
        assert ref.ref_kdtree_translate(maxdist, dp(bx), dp(by), bx.size, dp(lon), dp(lat), 61, 83) == 0
        assert np.array_equal(ax, bx) and np.array_equal(ay, by)
        assert (ax == -1000).any() and (ax >= 0).any()
    # one NaN coordinate: the reference loses matches, the restatement does not
    lat2 = lat.copy()
    lat2[40, 30] = np.nan
    ax, ay = qx.copy(), qy.copy()
    oracle.orc_kdtree_translate(60000.0, dp(ax), dp(ay), ax.size, dp(lon), dp(lat2), 61, 83, None)
    bx, by = qx.copy(), qy.copy()
    ref.ref_kdtree_translate(60000.0, dp(bx), dp(by), bx.size, dp(lon), dp(lat2), 61, 83)
    assert ((bx == -1000) & (ax >= 0)).sum() > 100 and not ((ax == -1000) & (bx >= 0)).any()


# ---------------------------------------------------------------- CPU: the product's host-side logic against the reference
CLOUDS = ["plain", "edge_nan", "edge_nan_wrap", "scatter_nan", "first_nan", "heavy_nan", "rounded_dups", "one_position"]


def cloud(rng, nx, ny, kind):
    lon, lat = swath(rng, nx, ny, nan_frac=0.005 if kind.startswith("edge_nan") else 0.0, wrap_column=kind.endswith("wrap"))
    if kind == "scatter_nan":
        lat[rng.random(lat.shape) < 0.05] = np.nan
    elif kind == "first_nan":  # the bounding box of the whole tree starts from point 0 (nanoflann.hpp:1020-1039)
        lat[0, 0] = np.nan
    elif kind == "heavy_nan":
        lon[rng.random(lat.shape) < 0.4] = np.nan
        lat[0, 0] = np.nan
    elif kind == "rounded_dups":  # many exactly equal positions: ties everywhere, more than 16 matches per target
        lat[:], lon[:] = np.round(lat), np.round(lon)
    elif kind == "one_position":
        lat[:], lon[:] = 60.0, 10.0
    return lon, lat


def cloud_targets(rng, n, kind):
    qx, qy = targets(rng, n)
    if kind in ("rounded_dups", "one_position"):
        qx, qy = np.round(qx * 2) / 2, np.round(qy * 2) / 2  # exactly between duplicates
    return qx, qy


@pytest.mark.parametrize("kind", CLOUDS)
def test_host_tree_clone_equals_nanoflann(ref, kind):
    """fxcu_kdclone.h (build + walk + std::sort) against the reference's nanoflann, every target, sequential and threaded build"""
    lib = ol.native("kdclone_host")
    rng = np.random.default_rng(CLOUDS.index(kind))
    nx, ny = 61, 83
    lon, lat = cloud(rng, nx, ny, kind)
    qx, qy = cloud_targets(rng, 4000, kind)
    for maxdist, threads in ((60000.0, 1), (400000.0, 1), (60000.0, 8)):
        ax, ay, err = qx.copy(), qy.copy(), C.create_string_buffer(512)
        assert lib.kdclone_translate_host(maxdist, dp(ax), dp(ay), ax.size, dp(lon), dp(lat), nx, ny, threads, err, 512) == 0, err.value
        bx, by = qx.copy(), qy.copy()
        assert ref.ref_kdtree_translate(maxdist, dp(bx), dp(by), bx.size, dp(lon), dp(lat), nx, ny) == 0
        assert np.array_equal(ax, bx) and np.array_equal(ay, by)
        assert (bx >= 0).any()


def test_host_tree_clone_threaded_build_large(ref):
    lib = ol.native("kdclone_host")
    rng = np.random.default_rng(77)
    nx, ny = 403, 335  # > 65536 points: the threaded build really forks
    lon, lat = cloud(rng, nx, ny, "edge_nan_wrap")
    qx, qy = targets(rng, 20000)
    ax, ay, err = qx.copy(), qy.copy(), C.create_string_buffer(512)
    assert lib.kdclone_translate_host(20000.0, dp(ax), dp(ay), ax.size, dp(lon), dp(lat), nx, ny, 8, err, 512) == 0, err.value
    bx, by = qx.copy(), qy.copy()
    ref.ref_kdtree_translate(20000.0, dp(bx), dp(by), bx.size, dp(lon), dp(lat), nx, ny)
    assert np.array_equal(ax, bx) and np.array_equal(ay, by) and (bx >= 0).sum() > 1000


@pytest.mark.parametrize("kind", ["plain", "edge_nan_wrap", "scatter_nan", "rounded_dups"])
def test_host_sorted_scan_equals_reference(ref, kind):
    """fxcu_llscan.h (compact keys, dot-product shortcut) against the scan restated on the reference's own library calls"""
    lib = ol.native("llscan_host")
    rng = np.random.default_rng(20 + CLOUDS.index(kind))
    nx, ny = 97, 71
    lon, lat = cloud(rng, nx, ny, kind)
    if kind == "rounded_dups":
        lon, lat = swath(rng, nx, ny, nan_frac=0.0)
        lat[:], lon[:] = np.round(lat * 2) / 2, np.round(lon * 2) / 2
    qx, qy = cloud_targets(rng, 6000, kind)
    roi = ref.ref_grid_distance(dp(lon), dp(lat), nx, ny)
    ax, ay = qx.copy(), qy.copy()
    lib.llscan_translate_host(dp(ax), dp(ay), ax.size, dp(lon), dp(lat), nx, ny, roi)
    bx, by = qx.copy(), qy.copy()
    ref.ref_coordnn_translate(dp(bx), dp(by), bx.size, dp(lon), dp(lat), nx, ny)
    assert np.array_equal(ax, bx) and np.array_equal(ay, by)
    assert (bx >= 0).any() and (bx == -1).any()


# ---------------------------------------------------------------- GPU: libfimexcu through the C ABI
def _ref_or_skip():
    r = ol.ref(omp=True)
    if r is None:
        pytest.skip("oracle/_ref not built")
    return r


@pytest.mark.gpu
@pytest.mark.parametrize("kind", CLOUDS)
def test_gpu_kdtree_equals_nanoflann(kind):
    import fimex_b200
    from fimex_b200.coordinates import flannTranslatePointsToClosestInputCell as gpu
    assert fimex_b200.lib.fimexcu_device_count() >= 1
    r = _ref_or_skip()
    rng = np.random.default_rng(100 + CLOUDS.index(kind))
    nx, ny = 203, 135
    lon, lat = cloud(rng, nx, ny, kind)
    qx, qy = cloud_targets(rng, 30000, kind)
    ok = ~np.isnan(lat[10, 20:70])
    qx[:50][ok], qy[:50][ok] = lon[10, 20:70][ok], lat[10, 20:70][ok]  # exact hits
    qx[50], qy[51] = np.nan, np.nan                                    # NaN targets find nothing
    for maxdist in (3000.0, 15000.0, 150000.0):                        # the last: far more than 16 matches per target
        gx, gy = gpu(maxdist, qx, qy, lon, lat)
        bx, by = qx.copy(), qy.copy()
        r.ref_kdtree_translate(maxdist, dp(bx), dp(by), bx.size, dp(lon), dp(lat), nx, ny)
        assert np.array_equal(gx, bx) and np.array_equal(gy, by)
        assert gx[50] == -1000 and gx[51] == -1000
    assert (bx >= 0).any()


@pytest.mark.gpu
def test_gpu_kdtree_config4_as_written():
    """2030x1354 swath, 0.5 % NaN coordinates at the edges, one duplicated wrap column (exact ties),
    radius = target spacing of a 0.025 deg grid: every target equals nanoflann's answer"""
    from fimex_b200.coordinates import flannTranslatePointsToClosestInputCell as gpu
    r = _ref_or_skip()
    rng = np.random.default_rng(2)
    nx, ny = 1354, 2030
    lon, lat = swath(rng, nx, ny, nan_frac=0.005, wrap_column=True)
    assert np.isnan(lat).any()
    # the grid covers the swath's western edge, where the NaN coordinates and the duplicated column (0 == nx-1) lie
    tlon, tlat = np.meshgrid(np.arange(-18, 2, 0.025), np.arange(53, 65, 0.025))
    maxdist = 0.025 * np.pi / 180 * 6371000.0
    gx, gy = gpu(maxdist, tlon, tlat, lon, lat)
    bx, by = tlon.ravel().copy(), tlat.ravel().copy()
    r.ref_kdtree_translate(maxdist, dp(bx), dp(by), bx.size, dp(lon), dp(lat), nx, ny)
    assert np.array_equal(gx.ravel(), bx) and np.array_equal(gy.ravel(), by)
    hit = bx >= 0
    assert hit.mean() > 0.2 and (bx[~hit] == -1000).all()
    # the duplicated column is really exercised: targets whose winner sits on column 0 or on its copy
    assert ((bx == 0) | (bx == nx - 1)).sum() > 100


@pytest.mark.gpu
def test_gpu_kdtree_corrected_mode(oracle):
    """nanoflann_compat=0: NaN source points are left out and every target gets its true nearest point (the tree-free oracle),
    which the reference does not deliver on such clouds"""
    from fimex_b200.coordinates import flannTranslatePointsToClosestInputCell as gpu
    rng = np.random.default_rng(9)
    nx, ny = 203, 135
    lon, lat = swath(rng, nx, ny, nan_frac=0.01)
    assert np.isnan(lat).any()
    qx, qy = targets(rng, 30000)
    gx, gy = gpu(15000.0, qx, qy, lon, lat, nanoflann_compat=False)
    wx, wy, gap = qx.copy(), qy.copy(), np.empty(qx.size)
    oracle.orc_kdtree_translate(15000.0, dp(wx), dp(wy), wx.size, dp(lon), dp(lat), nx, ny, dp(gap))
    clear = ~(np.abs(gap) <= 1e-12 * (15000.0 / 6371000.0) ** 2 + 1e-28)  # this mode uses the device's sin/cos; misses have no gap
    assert np.array_equal(gx[clear], wx[clear]) and np.array_equal(gy[clear], wy[clear])
    assert clear.mean() > 0.999 and (wx >= 0).sum() > 1000


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["plain", "edge_nan_wrap", "scatter_nan", "rounded_dups"])
def test_gpu_coord_nearestneighbor_equals_reference(kind):
    from fimex_b200.coordinates import fastTranslatePointsToClosestInputCell as gpu
    r = _ref_or_skip()
    rng = np.random.default_rng(200 + CLOUDS.index(kind))
    nx, ny = 151, 97
    lon, lat = cloud(rng, nx, ny, kind)
    if kind == "rounded_dups":
        lon, lat = swath(rng, nx, ny, nan_frac=0.0)
        lat[:], lon[:] = np.round(lat * 2) / 2, np.round(lon * 2) / 2
    qx, qy = cloud_targets(rng, 20000, kind)
    qx[:40], qy[:40] = lon[5, 10:50], lat[5, 10:50]
    qx[41] = np.nan
    gx, gy = gpu(qx, qy, lon, lat)
    bx, by = qx.copy(), qy.copy()
    r.ref_coordnn_translate(dp(bx), dp(by), bx.size, dp(lon), dp(lat), nx, ny)
    assert np.array_equal(gx, bx) and np.array_equal(gy, by)
    assert (bx == -1).any() and (bx >= 0).any()


@pytest.mark.gpu
def test_gpu_coord_nearestneighbor_config4_as_written():
    from fimex_b200.coordinates import fastTranslatePointsToClosestInputCell as gpu
    r = _ref_or_skip()
    rng = np.random.default_rng(3)
    nx, ny = 1354, 2030
    lon, lat = swath(rng, nx, ny, nan_frac=0.005, wrap_column=True)
    tlon, tlat = np.meshgrid(np.arange(-18, 2, 0.05), np.arange(53, 65, 0.05))
    gx, gy = gpu(tlon, tlat, lon, lat)
    bx, by = tlon.ravel().copy(), tlat.ravel().copy()
    r.ref_coordnn_translate(dp(bx), dp(by), bx.size, dp(lon), dp(lat), nx, ny)
    assert np.array_equal(gx.ravel(), bx) and np.array_equal(gy.ravel(), by)
    assert (bx >= 0).mean() > 0.2 and ((bx == 0) | (bx == nx - 1)).sum() > 100
