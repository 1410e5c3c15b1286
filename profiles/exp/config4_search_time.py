"""BASELINE configs[3] (swath 2030x1354 -> grid): one-off cache builds on the GPU against the reference's CPU code (oracle/_ref:
nanoflann radius search as in flannTranslatePointsToClosestInputCell; CachedForwardInterpolation ctor via the oracle tables)."""
import sys, os, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "tests"))
import numpy as np, torch
import oracle_lib as ol
from oracle_lib import dp
from test_coordinate_search import swath
from fimex_b200.coordinates import flannTranslatePointsToClosestInputCell, fastTranslatePointsToClosestInputCell
import fimex_b200
rng = np.random.default_rng(2)
nx, ny = 1354, 2030
lon, lat = swath(rng, nx, ny, nan_frac=0.0)
tlon, tlat = np.meshgrid(np.arange(0, 20, 0.025), np.arange(58, 70, 0.025))
maxdist = 0.025 * np.pi / 180 * 6371000.0
print("source", lon.shape, "targets", tlon.shape, tlon.size)


def t(fn, reps=3):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps


print("GPU coord_kdtree translate (host arrays in/out): %.2f ms" % (1e3 * t(lambda: flannTranslatePointsToClosestInputCell(maxdist, tlon, tlat, lon, lat))))
print("GPU coord_nearestneighbor translate:             %.2f ms" % (1e3 * t(lambda: fastTranslatePointsToClosestInputCell(tlon, tlat, lon, lat))))
# forward cache: source cell positions on the target axes
px = (lon - 0.0) / 0.025
py = (lat - 58.0) / 0.025
print("GPU CachedForwardInterpolation ctor:             %.2f ms" % (1e3 * t(lambda: fimex_b200.CachedForwardInterpolation("x", "y", ol.FWD_MEAN, px, py, nx, ny, tlon.shape[1], tlon.shape[0]).close())))
r = ol.ref()
if r is not None:
    sub = 200000
    sel = rng.choice(tlon.size, sub, replace=False)
    wx, wy = tlon.ravel()[sel].copy(), tlat.ravel()[sel].copy()
    t0 = time.perf_counter()
    r.ref_kdtree_translate(maxdist, dp(wx), dp(wy), wx.size, dp(lon), dp(lat), nx, ny)
    dt = time.perf_counter() - t0
    print("reference nanoflann (tree build + %d of %d targets, 1 core): %.2f s" % (sub, tlon.size, dt))
