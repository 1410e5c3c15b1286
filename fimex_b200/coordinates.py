"""Host-side mirror of the coordinate-based nearest-neighbour searches of CDMInterpolator
(src/CDMInterpolator.cc:953-1028 flannTranslatePointsToClosestInputCell, :1112-1190 fastTranslatePointsToClosestInputCell)."""
import numpy as np

from . import _lib
from ._lib import lib


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def flannTranslatePointsToClosestInputCell(maxDist, pointsOnXAxis, pointsOnYAxis, lonVals, latVals, nanoflann_compat=True):
    """coord_kdtree: target lon/lat (degrees) -> source (ix, iy) as doubles, -1000 where nothing lies within maxDist metres.
    lonVals/latVals: [orgYDimSize][orgXDimSize].  nanoflann_compat=True: the reference's result, tie order and NaN-cloud losses
    included; False: true nearest point, NaN source points left out, lower index on ties."""
    lon, lat = _f64(lonVals), _f64(latVals)
    ny, nx = lon.shape
    px, py = np.array(pointsOnXAxis, dtype=np.float64).ravel().copy(), np.array(pointsOnYAxis, dtype=np.float64).ravel().copy()
    _lib.check(lib.fimexcu_kdtree_translate_mode(float(maxDist), px.ctypes.data, py.ctypes.data, px.size, lon.ctypes.data, lat.ctypes.data, nx, ny,
                                                 1 if nanoflann_compat else 0), "flannTranslatePointsToClosestInputCell")
    return px.reshape(np.shape(pointsOnXAxis)), py.reshape(np.shape(pointsOnYAxis))


def fastTranslatePointsToClosestInputCell(pointsOnXAxis, pointsOnYAxis, lonVals, latVals):
    """coord_nearestneighbor: same with the region of influence estimated from the source grid; misses give -1."""
    lon, lat = _f64(lonVals), _f64(latVals)
    ny, nx = lon.shape
    px, py = np.array(pointsOnXAxis, dtype=np.float64).ravel().copy(), np.array(pointsOnYAxis, dtype=np.float64).ravel().copy()
    _lib.check(lib.fimexcu_coordnn_translate(px.ctypes.data, py.ctypes.data, px.size, lon.ctypes.data, lat.ctypes.data, nx, ny),
               "fastTranslatePointsToClosestInputCell")
    return px.reshape(np.shape(pointsOnXAxis)), py.reshape(np.shape(pointsOnYAxis))
