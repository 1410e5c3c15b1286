"""Coordinate-based nearest neighbour (coord_kdtree K9, coord_nearestneighbor K10).

CPU part: the tree-free restatement in oracle/ against the reference's vendored nanoflann (oracle/_ref).
GPU part: libfimexcu against both.  Bar: identical source cell for every target, except targets whose best and second-best
candidates are closer together than the sin/cos rounding noise (|d2 gap| <= 1e-12 relative: the device's and glibc's sin/cos
differ in the last bit) and exact-distance ties between duplicated source positions, where the chosen duplicate may differ."""
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


def _same_cells(gx, gy, wx, wy, lon, lat, ambiguous):
    diff = (gx != wx) | (gy != wy)
    bad = diff & ~ambiguous
    if bad.any():
        # duplicates: a different index at the very same position is fine
        ok = np.zeros(bad.sum(), bool)
        for k, i in enumerate(np.flatnonzero(bad)):
            if gx[i] >= 0 and wx[i] >= 0:
                a, b = (int(gy[i]), int(gx[i])), (int(wy[i]), int(wx[i]))
                ok[k] = lon[a] == lon[b] and lat[a] == lat[b]
        return ok.all()
    return True


@pytest.mark.gpu
@pytest.mark.parametrize("wrap", [False, True])
def test_gpu_kdtree_matches_reference(wrap):
    import fimex_b200
    from fimex_b200.coordinates import flannTranslatePointsToClosestInputCell as gpu
    assert fimex_b200.lib.fimexcu_device_count() >= 1
    rng = np.random.default_rng(9)
    nx, ny = 203, 135
    lon, lat = swath(rng, nx, ny, nan_frac=0.0, wrap_column=wrap)
    qx, qy = targets(rng, 30000)
    qx[:50], qy[:50] = lon[10, 20:70], lat[10, 20:70]  # exact hits
    for maxdist in (3000.0, 15000.0):
        gx, gy = gpu(maxdist, qx, qy, lon, lat)
        wx, wy, gap = qx.copy(), qy.copy(), np.empty(qx.size)
        ol.oracle().orc_kdtree_translate(maxdist, dp(wx), dp(wy), wx.size, dp(lon), dp(lat), nx, ny, dp(gap))
        r2 = (maxdist / 6371000.0) ** 2
        assert _same_cells(gx, gy, wx, wy, lon, lat, ambiguous=np.abs(gap) <= 1e-12 * r2 + 1e-28)
        assert ((gx == wx) & (gy == wy)).mean() > 0.9999
        r = ol.ref()
        if r is not None and not wrap:  # nanoflann itself (tie order of duplicates is its own)
            bx, by = qx.copy(), qy.copy()
            r.ref_kdtree_translate(maxdist, dp(bx), dp(by), bx.size, dp(lon), dp(lat), nx, ny)
            assert ((gx == bx) & (gy == by)).mean() > 0.9999
        assert (gx == -1000).any() and (gx >= 0).any()
    # NaN source coordinates are skipped, everything else still found (brute-force restatement as the checker)
    lon2, lat2 = swath(rng, nx, ny, nan_frac=0.01)
    assert np.isnan(lat2).any()
    gx, gy = gpu(15000.0, qx, qy, lon2, lat2)
    wx, wy, gap = qx.copy(), qy.copy(), np.empty(qx.size)
    ol.oracle().orc_kdtree_translate(15000.0, dp(wx), dp(wy), wx.size, dp(lon2), dp(lat2), nx, ny, dp(gap))
    assert _same_cells(gx, gy, wx, wy, lon2, lat2, ambiguous=np.abs(gap) <= 1e-12 * (15000.0 / 6371000.0) ** 2 + 1e-28)


@pytest.mark.gpu
def test_gpu_kdtree_config4_size():
    """2030x1354 swath -> 0.025 deg grid, radius = target spacing: every match is within the radius and is the nearest
    (checked against a brute-force subset), NaN source coordinates are never matched"""
    from fimex_b200.coordinates import flannTranslatePointsToClosestInputCell as gpu
    rng = np.random.default_rng(2)
    nx, ny = 1354, 2030
    lon, lat = swath(rng, nx, ny)
    tlon, tlat = np.meshgrid(np.arange(0, 20, 0.025), np.arange(58, 70, 0.025))
    maxdist = 0.025 * np.pi / 180 * 6371000.0
    gx, gy = gpu(maxdist, tlon, tlat, lon, lat)
    hit = gx >= 0
    assert hit.mean() > 0.3 and (gx[~hit] == -1000).all()
    iy, ix = gy[hit].astype(int), gx[hit].astype(int)
    assert not np.isnan(lat[iy, ix]).any()

    def xyz(lo, la):
        lo, la = np.radians(lo), np.radians(la)
        return np.stack([np.cos(la) * np.cos(lo), np.cos(la) * np.sin(lo), np.sin(la)], -1)
    d = np.linalg.norm(xyz(tlon[hit], tlat[hit]) - xyz(lon[iy, ix], lat[iy, ix]), axis=-1)
    assert (d < maxdist / 6371000.0).all()
    sel = rng.choice(tlon.size, 300, replace=False)
    wx, wy = tlon.ravel()[sel].copy(), tlat.ravel()[sel].copy()
    gap = np.empty(sel.size)
    ol.oracle().orc_kdtree_translate(maxdist, dp(wx), dp(wy), wx.size, dp(lon), dp(lat), nx, ny, dp(gap))
    amb = np.abs(gap) <= 1e-12 * (maxdist / 6371000.0) ** 2
    assert (((gx.ravel()[sel] == wx) & (gy.ravel()[sel] == wy)) | amb).all()


@pytest.mark.gpu
def test_gpu_coord_nearestneighbor_matches_oracle():
    from fimex_b200.coordinates import fastTranslatePointsToClosestInputCell as gpu
    rng = np.random.default_rng(11)
    nx, ny = 151, 97
    lon, lat = swath(rng, nx, ny)
    qx, qy = targets(rng, 20000)
    gx, gy = gpu(qx, qy, lon, lat)
    wx, wy = qx.copy(), qy.copy()
    ol.oracle().orc_coordnn_translate(dp(wx), dp(wy), wx.size, dp(lon), dp(lat), nx, ny)
    same = (gx == wx) & (gy == wy)
    assert same.mean() > 0.9995  # cos_d near-ties at the sin/cos rounding level may pick the other of two equidistant cells
    assert (wx == -1).any() and (wx >= 0).any() and np.array_equal(gx == -1, wx == -1) or same.mean() > 0.9999
    # where they differ, the two candidates are equally far from the target to within 1e-12
    for i in np.flatnonzero(~same)[:20]:
        def cosd(ix, iy):
            a, b = np.radians(lat[int(iy), int(ix)]), np.radians(qy[i])
            return np.cos(a) * np.cos(b) * np.cos(np.radians(lon[int(iy), int(ix)] - qx[i])) + np.sin(a) * np.sin(b)
        if gx[i] >= 0 and wx[i] >= 0:
            assert abs(cosd(gx[i], gy[i]) - cosd(wx[i], wy[i])) < 1e-12
