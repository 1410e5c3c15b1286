"""Host-side mirror of the reference's `reproject` namespace (include/fimex/reproject.h:74-227) on top of the C ABI.
Angles in degrees for geographic CRS (PROJ >= 6 proj_create_crs_to_crs convention), metres otherwise."""
import numpy as np

from . import _lib
from ._lib import lib


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def reproject_axes(proj_input, proj_output, in_x_axis, in_y_axis):
    """reproject::reproject_axes (src/reproject.cc:573-587): returns (out_x, out_y), each [iy][ix]."""
    ax, ay = _f64(in_x_axis).ravel(), _f64(in_y_axis).ravel()
    ox = np.empty((ay.size, ax.size), dtype=np.float64)
    oy = np.empty_like(ox)
    _lib.check(lib.fimexcu_reproject_axes(proj_input.encode(), proj_output.encode(), ax.ctypes.data, ay.ctypes.data, ax.size, ay.size,
                                          ox.ctypes.data, oy.ctypes.data), "reproject_axes")
    return ox, oy


def reproject_values(proj_input, proj_output, x, y):
    """reproject::reproject_values (src/reproject.cc:567-571); returns new arrays (the reference works in place)."""
    xv, yv = np.array(x, dtype=np.float64).ravel().copy(), np.array(y, dtype=np.float64).ravel().copy()
    _lib.check(lib.fimexcu_reproject_values(proj_input.encode(), proj_output.encode(), xv.ctypes.data, yv.ctypes.data, xv.size), "reproject_values")
    return xv.reshape(np.shape(x)), yv.reshape(np.shape(y))


def get_vector_reproject_matrix(proj_input, proj_output, out_x_axis, out_y_axis):
    """reproject::get_vector_reproject_matrix (src/reproject.cc:463-486): [oy][ox][3] = cos, sin, angle in degree."""
    ax, ay = _f64(out_x_axis).ravel(), _f64(out_y_axis).ravel()
    m = np.empty((ay.size, ax.size, 3), dtype=np.float64)
    _lib.check(lib.fimexcu_get_vector_reproject_matrix(proj_input.encode(), proj_output.encode(), ax.ctypes.data, ay.ctypes.data, ax.size, ay.size,
                                                       m.ctypes.data), "get_vector_reproject_matrix")
    return m


def get_vector_reproject_matrix_field(proj_input, proj_output, in_x_field, in_y_field):
    """reproject::get_vector_reproject_matrix_field (src/reproject.cc:449-461); fields are [oy][ox] positions in the input projection."""
    fx, fy = _f64(in_x_field), _f64(in_y_field)
    oy, ox = fx.shape
    m = np.empty((oy, ox, 3), dtype=np.float64)
    _lib.check(lib.fimexcu_get_vector_reproject_matrix_field(proj_input.encode(), proj_output.encode(), fx.ctypes.data, fy.ctypes.data, ox, oy,
                                                             m.ctypes.data), "get_vector_reproject_matrix_field")
    return m


def get_vector_reproject_matrix_points(proj_input, proj_output, inputIsMetric, out_x_points, out_y_points):
    """reproject::get_vector_reproject_matrix_points (src/reproject.cc:432-447); points are [oy][ox] arrays."""
    px, py = _f64(out_x_points), _f64(out_y_points)
    oy, ox = px.shape if px.ndim == 2 else (1, px.size)
    m = np.empty((oy, ox, 3), dtype=np.float64)
    _lib.check(lib.fimexcu_get_vector_reproject_matrix_points(proj_input.encode(), proj_output.encode(), int(bool(inputIsMetric)), px.ctypes.data,
                                                              py.ctypes.data, ox, oy, m.ctypes.data), "get_vector_reproject_matrix_points")
    return m


def cachedInterpolationFromProjection(method, proj_input, org_proj, out_x_axis, out_y_axis, org_x_axis, org_y_axis, org_x_axis_type=0,
                                      org_y_axis_type=0):
    """The numeric part of CDMInterpolator::changeProjectionByProjectionParameters (src/CDMInterpolator.cc:1389-1404) in one call:
    reproject_axes(target -> source) + 2 x mifi_points2position + createCachedInterpolation, index tables never leave the device."""
    from .interpolation import CachedInterpolationInterface
    ox, oy = _f64(out_x_axis).ravel(), _f64(out_y_axis).ravel()
    ix, iy = _f64(org_x_axis).ravel(), _f64(org_y_axis).ravel()
    h = lib.fimexcu_cache_create_from_projection(method, proj_input.encode(), org_proj.encode(), ox.ctypes.data, oy.ctypes.data, ox.size, oy.size,
                                                 ix.ctypes.data, iy.ctypes.data, ix.size, iy.size, org_x_axis_type, org_y_axis_type)
    ci = CachedInterpolationInterface("x", "y")
    ci._h = _lib.check_handle(h, "changeProjectionByProjectionParameters")
    ci._read_info()
    return ci


LAT_LON_PROJSTR = "+proj=latlong +datum=WGS84 +towgs84=0,0,0 +no_defs"  # MIFI_WGS84_LATLON_PROJ4 (include/fimex/mifi_constants.h)


def cachedInterpolationForPoints(method, org_proj, lonVals, latVals, org_x_axis, org_y_axis, org_x_axis_type=0, org_y_axis_type=0, out_shape=None,
                                 proj_points=LAT_LON_PROJSTR):
    """CDMInterpolator::changeProjection(method, lonVals, latVals) / ...ToLatLonTemplate / ...ToCrossSections
    (src/CDMInterpolator.cc:457-490, 1580-1666): the target is a list of points in degrees; out_shape = (outY, outX) for a 2-D
    template, default a single row of len(lonVals) points."""
    from .interpolation import CachedInterpolationInterface
    lon, lat = _f64(lonVals).ravel(), _f64(latVals).ravel()
    if lon.size != lat.size:
        raise _lib.FimexCudaError("number of longitude and latitude values differs")
    oy, ox = (1, lon.size) if out_shape is None else out_shape
    if ox * oy != lon.size:
        raise _lib.FimexCudaError("out_shape does not match the number of points")
    ix, iy = _f64(org_x_axis).ravel(), _f64(org_y_axis).ravel()
    h = lib.fimexcu_cache_create_from_points(method, proj_points.encode(), org_proj.encode(), lon.ctypes.data, lat.ctypes.data, ox, oy, ix.ctypes.data,
                                             iy.ctypes.data, ix.size, iy.size, org_x_axis_type, org_y_axis_type)
    ci = CachedInterpolationInterface("x", "y")
    ci._h = _lib.check_handle(h, "changeProjectionToLatLonTemplate")
    ci._read_info()
    return ci
