"""ctypes access to the CPU checkers in oracle/ (TEST INFRASTRUCTURE).

`oracle()`  -> oracle/liboracle.so, the C restatement (always buildable).
`ref(omp)`  -> oracle/_ref/libfimexref[_omp].so, the reference's own sources compiled by oracle/Makefile
               (present when built in the container; travels to the GPU box as a built file), or None.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
HERE = os.path.dirname(os.path.abspath(__file__))

c_double_p = C.POINTER(C.c_double)
c_float_p = C.POINTER(C.c_float)
c_size_p = C.POINTER(C.c_size_t)

# method enums (mifi_constants.h:52-138, 206-231)
NEAREST, BILINEAR, BICUBIC, COORD_NN, COORD_NN_KD = 0, 1, 2, 3, 4
FWD_SUM, FWD_MEAN, FWD_MEDIAN, FWD_MAX, FWD_MIN = 5, 6, 7, 8, 9
FWD_UNDEF_SUM, FWD_UNDEF_MEAN, FWD_UNDEF_MEDIAN, FWD_UNDEF_MAX, FWD_UNDEF_MIN = 10, 11, 12, 13, 14
VINT_LIN, VINT_LOG, VINT_LOGLOG, VINT_NN, VINT_LIN_WEAK, VINT_LIN_NO, VINT_LIN_CONST = 0, 1, 2, 3, 4, 5, 6
PROJ_AXIS, LONGITUDE, LATITUDE = 0, 1, 2
INVALID_IDX = 0xFFFFFFFF


def build(force=False):
    """make -C oracle (liboracle.so always; _ref only when /root/reference is present)."""
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    if force or not os.path.exists(so):
        subprocess.run(["make", "-C", ORACLE_DIR], check=True, stdout=subprocess.DEVNULL)
    return so


def dp(a):
    return a.ctypes.data_as(c_double_p)


def fp(a):
    return a.ctypes.data_as(c_float_p)


class _CachedInterp(C.Structure):
    _fields_ = [("method", C.c_int), ("inX", C.c_size_t), ("inY", C.c_size_t), ("outX", C.c_size_t), ("outY", C.c_size_t),
                ("reduced", C.c_int), ("xMin", C.c_size_t), ("yMin", C.c_size_t), ("px", c_double_p), ("py", c_double_p)]


class _CachedNN(C.Structure):
    _fields_ = [("inX", C.c_size_t), ("inY", C.c_size_t), ("outX", C.c_size_t), ("outY", C.c_size_t), ("reduced", C.c_int),
                ("xMin", C.c_size_t), ("yMin", C.c_size_t), ("pointsInIn", c_size_p)]


class _CachedFwd(C.Structure):
    _fields_ = [("method", C.c_int), ("undefAggr", C.c_int), ("inX", C.c_size_t), ("inY", C.c_size_t), ("outX", C.c_size_t),
                ("outY", C.c_size_t), ("reduced", C.c_int), ("xMin", C.c_size_t), ("yMin", C.c_size_t), ("maxPointsInIn", C.c_size_t),
                ("segStart", c_size_p), ("segIdx", c_size_p)]


_oracle = None


def oracle():
    global _oracle
    if _oracle is not None:
        return _oracle
    lib = C.CDLL(build())
    two_level = [c_float_p, c_float_p, c_float_p, C.c_size_t, C.c_double, C.c_double, C.c_double]
    point = [c_float_p, c_float_p, C.c_double, C.c_double, C.c_int, C.c_int, C.c_int]
    sig = {
        "orc_string_to_interpolation_method": (C.c_int, [C.c_char_p]),
        "orc_bsearch_double_index": (C.c_int, [C.c_double, c_double_p, C.c_int, C.c_int]),
        "orc_points2position": (C.c_int, [c_double_p, C.c_size_t, c_double_p, C.c_int, C.c_int]),
        "orc_get_values_f": (C.c_int, point),
        "orc_get_values_bilinear_f": (C.c_int, point),
        "orc_get_values_bicubic_f": (C.c_int, point),
        "orc_get_values_nearest_f": (C.c_int, two_level),
        "orc_get_values_linear_f": (C.c_int, two_level),
        "orc_get_values_linear_weak_extrapol_f": (C.c_int, two_level),
        "orc_get_values_linear_no_extrapol_f": (C.c_int, two_level),
        "orc_get_values_linear_const_extrapol_f": (C.c_int, two_level),
        "orc_get_values_log_f": (C.c_int, two_level),
        "orc_get_values_log_log_f": (C.c_int, two_level),
        "orc_bad2nanf": (None, [c_float_p, C.c_size_t, C.c_float]),
        "orc_nanf2bad": (None, [c_float_p, C.c_size_t, C.c_float]),
        "orc_vcoord_levels_f": (C.c_int, [C.c_int, c_double_p, c_double_p, C.c_double, c_float_p, C.c_size_t, C.c_double, c_float_p, C.c_int, c_float_p,
                                         C.c_double, C.c_size_t, C.c_size_t, C.c_size_t, c_float_p]),
        "orc_vcoord_pressure_integration_f": (C.c_int, [c_float_p, c_float_p, C.c_double, c_float_p, C.c_double, c_float_p, c_float_p, C.c_size_t,
                                                       C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, c_float_p, C.c_double, c_float_p]),
        "orc_vcoord_transform_f": (None, [C.c_int, c_float_p, C.c_size_t, c_float_p]),
        "orc_virtual_temperature": (C.c_float, [C.c_float, C.c_float]),
        "orc_barometric_layer_thickness": (C.c_float, [C.c_float, C.c_float, C.c_float]),
        "orc_fill2d_f": (C.c_int, [C.c_size_t, C.c_size_t, c_float_p, C.c_float, C.c_float, C.c_size_t, c_size_p]),
        "orc_creepfill2d_f": (C.c_int, [C.c_size_t, C.c_size_t, c_float_p, C.c_ushort, C.c_char, c_size_p]),
        "orc_creepfillval2d_f": (C.c_int, [C.c_size_t, C.c_size_t, c_float_p, C.c_float, C.c_ushort, C.c_char, c_size_p]),
        "orc_unpack_f": (C.c_int, [C.c_int, C.c_void_p, C.c_size_t, C.c_double, c_float_p]),
        "orc_unpack_scaled_f": (C.c_int, [C.c_int, C.c_void_p, C.c_size_t, C.c_double, C.c_double, C.c_double, c_float_p]),
        "orc_pack_f": (C.c_int, [C.c_int, c_float_p, C.c_size_t, C.c_double, C.c_void_p]),
        "orc_i16_to_interpolation_array": (None, [C.c_void_p, C.c_size_t, C.c_double, c_float_p]),
        "orc_interpolation_array_to_i16": (None, [c_float_p, C.c_size_t, C.c_double, C.c_void_p]),
        "orc_i16_scaled_to_float": (None, [C.c_void_p, C.c_size_t, C.c_double, C.c_double, C.c_double, c_float_p]),
        "orc_round_and_clamp": (C.c_size_t, [C.c_double, C.c_size_t, C.c_size_t, C.c_size_t]),
        "orc_cached_interp_create": (C.POINTER(_CachedInterp), [C.c_int, c_double_p, c_double_p] + [C.c_size_t] * 4),
        "orc_cached_interp_apply": (C.c_int, [C.POINTER(_CachedInterp), c_float_p, C.c_size_t, c_float_p]),
        "orc_cached_interp_destroy": (None, [C.POINTER(_CachedInterp)]),
        "orc_cached_nn_create": (C.POINTER(_CachedNN), [c_double_p, c_double_p] + [C.c_size_t] * 4),
        "orc_cached_nn_apply": (C.c_int, [C.POINTER(_CachedNN), c_float_p, C.c_size_t, c_float_p]),
        "orc_cached_nn_destroy": (None, [C.POINTER(_CachedNN)]),
        "orc_cached_fwd_create": (C.POINTER(_CachedFwd), [C.c_int, c_double_p, c_double_p] + [C.c_size_t] * 4),
        "orc_cached_fwd_apply": (C.c_int, [C.POINTER(_CachedFwd), c_float_p, C.c_size_t, c_float_p, C.c_int]),
        "orc_cached_fwd_destroy": (None, [C.POINTER(_CachedFwd)]),
        "orc_vector_reproject_values_by_matrix_f": (None, [c_double_p, C.c_size_t, c_float_p, c_float_p, C.c_size_t]),
        "orc_vector_reproject_direction_by_matrix_f": (None, [c_double_p, C.c_size_t, c_float_p, C.c_size_t]),
        "orc_find_closest_distinct_elements": (None, [c_float_p, C.c_size_t, C.c_ssize_t, C.c_float, c_size_p, c_size_p]),
        "orc_find_closest_neighbor_distinct_elements": (None, [c_float_p, C.c_size_t, C.c_ssize_t, C.c_float, c_size_p, c_size_p]),
        "orc_atmosphere_hybrid_sigma_ap_pressure": (None, [C.c_size_t, C.c_double, c_double_p, c_double_p, c_double_p]),
        "orc_atmosphere_hybrid_sigma_pressure": (None, [C.c_size_t, C.c_double, C.c_double, c_double_p, c_double_p, c_double_p]),
        "orc_atmosphere_sigma_pressure": (None, [C.c_size_t, C.c_double, C.c_double, c_double_p, c_double_p]),
        "orc_atmosphere_ln_pressure": (None, [C.c_size_t, C.c_double, c_double_p, c_double_p]),
        "orc_atmosphere_hybrid_height": (None, [C.c_size_t, C.c_double, c_double_p, c_double_p, c_double_p]),
        "orc_vertical_interpolate_f": (C.c_int, [C.c_int, c_float_p, c_float_p, C.c_size_t, C.c_size_t, c_double_p, c_float_p, C.c_size_t,
                                                 c_double_p, c_double_p, C.c_double, C.c_double, c_float_p]),
        "orc_kdtree_translate": (C.c_int, [C.c_double, c_double_p, c_double_p, C.c_size_t, c_double_p, c_double_p, C.c_size_t, C.c_size_t, c_double_p]),
        "orc_get_grid_distance": (C.c_double, [c_double_p, c_double_p, C.c_size_t, C.c_size_t]),
        "orc_coordnn_translate": (C.c_int, [c_double_p, c_double_p, C.c_size_t, c_double_p, c_double_p, C.c_size_t, C.c_size_t]),
        "miniproj_transform_str": (C.c_long, [C.c_char_p, C.c_char_p, C.c_size_t, c_double_p, c_double_p, C.c_char_p, C.c_size_t]),
        "orc_reproject_axes": (C.c_int, [C.c_char_p, C.c_char_p, c_double_p, c_double_p, C.c_int, C.c_int, c_double_p, c_double_p]),
        "orc_get_vector_reproject_matrix": (C.c_int, [C.c_char_p, C.c_char_p, c_double_p, c_double_p, C.c_int, C.c_int, c_double_p]),
        "orc_get_vector_reproject_matrix_points": (C.c_int, [C.c_char_p, C.c_char_p, C.c_int, c_double_p, c_double_p, C.c_int, C.c_int, c_double_p]),
    }
    for name, (res, args) in sig.items():
        f = getattr(lib, name)
        f.restype = res
        f.argtypes = args
    _oracle = lib
    return lib


_ref = {}


def ref(omp=False):
    """The reference's own code (oracle/_ref), or None when it has not been built."""
    if omp in _ref:
        return _ref[omp]
    path = os.path.join(ORACLE_DIR, "_ref", "libfimexref_omp.so" if omp else "libfimexref.so")
    if not os.path.exists(path):
        _ref[omp] = None
        return None
    lib = C.CDLL(path)
    two_level = [c_float_p, c_float_p, c_float_p, C.c_size_t, C.c_double, C.c_double, C.c_double]
    point = [c_float_p, c_float_p, C.c_double, C.c_double, C.c_int, C.c_int, C.c_int]
    sig = {
        "ref_num_threads": (C.c_int, []),
        "ref_set_num_threads": (None, [C.c_int]),
        "ref_cache_create": (C.c_void_p, [C.c_int, c_double_p, c_double_p] + [C.c_size_t] * 4),
        "ref_cache_info": (None, [C.c_void_p, c_size_p, c_size_p, c_size_p, c_size_p, C.POINTER(C.c_int), c_size_p, c_size_p]),
        "ref_cache_apply": (C.c_int, [C.c_void_p, c_float_p, C.c_size_t, c_float_p]),
        "ref_cache_apply_seconds": (C.c_double, [C.c_void_p, c_float_p, C.c_size_t, C.c_int]),
        "ref_cache_destroy": (None, [C.c_void_p]),
        "ref_round_and_clamp": (C.c_size_t, [C.c_double, C.c_size_t, C.c_size_t, C.c_size_t]),
        "mifi_fill2d_f": (C.c_int, [C.c_size_t, C.c_size_t, c_float_p, C.c_float, C.c_float, C.c_size_t, c_size_p]),
        "mifi_creepfill2d_f": (C.c_int, [C.c_size_t, C.c_size_t, c_float_p, C.c_ushort, C.c_char, c_size_p]),
        "mifi_creepfillval2d_f": (C.c_int, [C.c_size_t, C.c_size_t, c_float_p, C.c_float, C.c_ushort, C.c_char, c_size_p]),
        "ref_unpack_f": (C.c_int, [C.c_int, C.c_void_p, C.c_size_t, C.c_double, c_float_p]),
        "ref_unpack_scaled_f": (C.c_int, [C.c_int, C.c_void_p, C.c_size_t, C.c_double, C.c_double, C.c_double, c_float_p]),
        "ref_pack_f": (C.c_int, [C.c_int, c_float_p, C.c_size_t, C.c_double, C.c_void_p]),
        "ref_i16_to_interpolation_array": (None, [C.c_void_p, C.c_size_t, C.c_double, c_float_p]),
        "ref_interpolation_array_to_i16": (None, [c_float_p, C.c_size_t, C.c_double, C.c_void_p]),
        "ref_i16_scaled_to_float": (None, [C.c_void_p, C.c_size_t, C.c_double, C.c_double, C.c_double, c_float_p]),
        "ref_find_closest_neighbor_distinct_elements": (None, [c_float_p, C.c_size_t, C.c_long, C.c_double, C.POINTER(C.c_long), C.POINTER(C.c_long)]),
        "ref_find_closest_distinct_elements": (None, [c_float_p, C.c_size_t, C.c_long, C.c_double, C.POINTER(C.c_long), C.POINTER(C.c_long)]),
        "ref_kdtree_translate": (C.c_int, [C.c_double, c_double_p, c_double_p, C.c_size_t, c_double_p, c_double_p, C.c_size_t, C.c_size_t]),
        "ref_coordnn_translate": (C.c_int, [c_double_p, c_double_p, C.c_size_t, c_double_p, c_double_p, C.c_size_t, C.c_size_t]),
        "ref_grid_distance": (C.c_double, [c_double_p, c_double_p, C.c_size_t, C.c_size_t]),
        "ref_vector_reproject_values_by_matrix_f": (None, [c_double_p, C.c_int, C.c_int, c_float_p, c_float_p, C.c_int]),
        "ref_vector_reproject_direction_by_matrix_f": (None, [c_double_p, C.c_int, C.c_int, c_float_p, C.c_int]),
        "ref_reproject_axes": (C.c_int, [C.c_char_p, C.c_char_p, c_double_p, c_double_p, C.c_int, C.c_int, c_double_p, c_double_p]),
        "ref_get_vector_reproject_matrix": (C.c_int, [C.c_char_p, C.c_char_p, c_double_p, c_double_p, C.c_int, C.c_int, c_double_p]),
        "ref_get_vector_reproject_matrix_field": (C.c_int, [C.c_char_p, C.c_char_p, c_double_p, c_double_p, C.c_int, C.c_int, c_double_p]),
        "ref_vector_reproject_values_f": (C.c_int, [C.c_char_p, C.c_char_p, c_float_p, c_float_p, c_double_p, c_double_p, C.c_int, C.c_int, C.c_int,
                                                    C.c_int, C.c_int]),
        "ref_reproject_f": (C.c_int, [C.c_int, C.c_char_p, c_float_p, c_double_p, c_double_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p,
                                      c_float_p, c_double_p, c_double_p, C.c_int, C.c_int, C.c_int, C.c_int]),
        # the reference's own extern "C" kernels (src/interpolation.c), exported as they are
        "mifi_points2position": (C.c_int, [c_double_p, C.c_int, c_double_p, C.c_int, C.c_int]),
        "mifi_get_values_f": (C.c_int, point),
        "mifi_get_values_bilinear_f": (C.c_int, point),
        "mifi_get_values_bicubic_f": (C.c_int, point),
        "mifi_get_values_nearest_f": (C.c_int, two_level),
        "mifi_get_values_linear_f": (C.c_int, two_level),
        "mifi_get_values_linear_weak_extrapol_f": (C.c_int, two_level),
        "mifi_get_values_linear_no_extrapol_f": (C.c_int, two_level),
        "mifi_get_values_linear_const_extrapol_f": (C.c_int, two_level),
        "mifi_get_values_log_f": (C.c_int, two_level),
        "mifi_get_values_log_log_f": (C.c_int, two_level),
        "mifi_string_to_interpolation_method": (C.c_int, [C.c_char_p]),
        "mifi_atmosphere_hybrid_sigma_ap_pressure": (C.c_int, [C.c_size_t, C.c_double, c_double_p, c_double_p, c_double_p]),
        "mifi_atmosphere_hybrid_sigma_pressure": (C.c_int, [C.c_size_t, C.c_double, C.c_double, c_double_p, c_double_p, c_double_p]),
        "mifi_atmosphere_sigma_pressure": (C.c_int, [C.c_size_t, C.c_double, C.c_double, c_double_p, c_double_p]),
        "mifi_atmosphere_ln_pressure": (C.c_int, [C.c_size_t, C.c_double, c_double_p, c_double_p]),
        "mifi_atmosphere_hybrid_height": (C.c_int, [C.c_size_t, C.c_double, c_double_p, c_double_p, c_double_p]),
        "mifi_ocean_s_g1_z": (C.c_int, [C.c_size_t, C.c_double, C.c_double, C.c_double, c_double_p, c_double_p, c_double_p]),
        "mifi_ocean_s_g2_z": (C.c_int, [C.c_size_t, C.c_double, C.c_double, C.c_double, c_double_p, c_double_p, c_double_p]),
        "mifi_barometric_standard_altitude": (C.c_int, [C.c_size_t, c_double_p, c_double_p]),
        "mifi_barometric_standard_pressure": (C.c_int, [C.c_size_t, c_double_p, c_double_p]),
        "mifi_virtual_temperature": (C.c_float, [C.c_float, C.c_float]),
        "mifi_barometric_layer_thickness": (C.c_float, [C.c_float, C.c_float, C.c_float]),
    }
    for name, (res, args) in sig.items():
        f = getattr(lib, name)
        f.restype = res
        f.argtypes = args
    _ref[omp] = lib
    return lib


# ---------------------------------------------------------------- numpy-level helpers over the oracle
def o_points2position(points, axis, axis_type):
    p = np.array(points, dtype=np.float64).ravel().copy()
    ax = np.ascontiguousarray(axis, dtype=np.float64)
    oracle().orc_points2position(dp(p), p.size, dp(ax), ax.size, axis_type)
    return p.reshape(np.shape(points))


def o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, field, level_bug=1):
    """Create the matching oracle cache and apply it.  `field` is the FULL [nz][in_y][in_x] array; the
    reduced window is cut here the way getInputDataSlice() would.  Returns (out[nz][out_y][out_x], info)."""
    lib = oracle()
    px = np.ascontiguousarray(px, dtype=np.float64).ravel()
    py = np.ascontiguousarray(py, dtype=np.float64).ravel()
    field = np.ascontiguousarray(field, dtype=np.float32).reshape(-1, in_y, in_x)
    nz = field.shape[0]
    if method in (BILINEAR, BICUBIC):
        h = lib.orc_cached_interp_create(method, dp(px), dp(py), in_x, in_y, out_x, out_y)
        apply, destroy = lib.orc_cached_interp_apply, lib.orc_cached_interp_destroy
    elif method in (NEAREST, COORD_NN, COORD_NN_KD):
        h = lib.orc_cached_nn_create(dp(px), dp(py), in_x, in_y, out_x, out_y)
        apply, destroy = lib.orc_cached_nn_apply, lib.orc_cached_nn_destroy
    else:
        h = lib.orc_cached_fwd_create(method, dp(px), dp(py), in_x, in_y, out_x, out_y)
        apply = lambda hh, a, b, c: lib.orc_cached_fwd_apply(hh, a, b, c, level_bug)  # noqa: E731
        destroy = lib.orc_cached_fwd_destroy
    assert h, "oracle cache creation failed"
    c = h.contents
    info = dict(inX=c.inX, inY=c.inY, outX=c.outX, outY=c.outY, reduced=bool(c.reduced), xMin=c.xMin, yMin=c.yMin)
    sub = np.ascontiguousarray(field[:, c.yMin:c.yMin + c.inY, c.xMin:c.xMin + c.inX])
    out = np.empty((nz, out_y, out_x), dtype=np.float32)
    apply(h, fp(sub), sub.size, fp(out))
    destroy(h)
    return out, info


def r_cached_apply(method, px, py, in_x, in_y, out_x, out_y, field, omp=False):
    """Same through the reference's own classes (oracle/_ref)."""
    lib = ref(omp)
    px = np.ascontiguousarray(px, dtype=np.float64).ravel()
    py = np.ascontiguousarray(py, dtype=np.float64).ravel()
    field = np.ascontiguousarray(field, dtype=np.float32).reshape(-1, in_y, in_x)
    nz = field.shape[0]
    h = lib.ref_cache_create(method, dp(px), dp(py), in_x, in_y, out_x, out_y)
    assert h, "reference cache creation failed"
    vals = [C.c_size_t() for _ in range(4)]
    red, xm, ym = C.c_int(), C.c_size_t(), C.c_size_t()
    lib.ref_cache_info(h, *[C.byref(v) for v in vals], C.byref(red), C.byref(xm), C.byref(ym))
    info = dict(inX=vals[0].value, inY=vals[1].value, outX=vals[2].value, outY=vals[3].value, reduced=bool(red.value), xMin=xm.value,
                yMin=ym.value)
    sub = np.ascontiguousarray(field[:, info["yMin"]:info["yMin"] + info["inY"], info["xMin"]:info["xMin"] + info["inX"]])
    out = np.empty((nz, out_y, out_x), dtype=np.float32)
    rc = lib.ref_cache_apply(h, fp(sub), sub.size, fp(out))
    lib.ref_cache_destroy(h)
    assert rc == 0
    return out, info


def o_packed_apply(method, px, py, inX, inY, outX, outY, raw, badValue):
    """CDMInterpolator::getDataSlice for a CDM_SHORT variable with the oracle: raw int16 [nz][inY][inX] -> int16 [nz][outY][outX]"""
    o = oracle()
    raw = np.ascontiguousarray(raw, dtype=np.int16)
    f = np.empty(raw.shape, np.float32)
    o.orc_i16_to_interpolation_array(raw.ctypes.data, raw.size, float(badValue), fp(f))
    res, _ = o_cached_apply(method, px, py, inX, inY, outX, outY, f)
    res = np.ascontiguousarray(res, dtype=np.float32)
    out = np.empty(res.shape, np.int16)
    o.orc_interpolation_array_to_i16(fp(res), res.size, float(badValue), out.ctypes.data)
    return out


PACKED_DTYPES = {0: np.int8, 1: np.int16, 2: np.int32, 3: np.uint8, 4: np.uint16, 5: np.uint32}


# ---------------------------------------------------------------- CPU harnesses over the product's host-only headers
_native = {}


def native(name):
    """tests/native/<name>.cc -> lib<name>.so (g++; built on first use).  These drive the HOST-side logic of libfimexcu's
    coordinate searches (fimex_b200/csrc/fxcu_kdclone.h, fxcu_llscan.h) and host entry points (fxcu_hostfill.h, fxcu_hostpar.h) so
    that it can be tested without a GPU."""
    if name in _native:
        return _native[name]
    src = os.path.join(HERE, "native", name + ".cc")
    so = os.path.join(HERE, "native", "lib" + name + ".so")
    deps = [src] + [os.path.join(ROOT, "fimex_b200", "csrc", h) for h in ("fxcu_kdclone.h", "fxcu_llscan.h", "fxcu_hostpar.h", "fxcu_hostfill.h")]
    if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
        subprocess.run(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-pthread", "-o", so, src], check=True)
    lib = C.CDLL(so)
    if name == "kdclone_host":
        lib.kdclone_translate_host.restype = C.c_int
        lib.kdclone_translate_host.argtypes = [C.c_double, c_double_p, c_double_p, C.c_size_t, c_double_p, c_double_p, C.c_size_t, C.c_size_t, C.c_uint,
                                               C.c_char_p, C.c_size_t]
    elif name == "hostfill_host":
        lib.hostfill_fill_f.argtypes = [c_float_p, C.c_size_t, C.c_float]
        lib.hostfill_fill_i16.argtypes = [C.c_void_p, C.c_size_t, C.c_short]
        lib.hostfill_copy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        lib.hostfill_merge.restype = C.c_int
        lib.hostfill_merge.argtypes = [C.POINTER(C.c_int), C.c_size_t, C.c_int, C.c_size_t, C.POINTER(C.c_int), c_double_p]
        lib.hostpar_cover.restype = C.c_size_t
        lib.hostpar_cover.argtypes = [C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_int]
    elif name == "llscan_host":
        lib.llscan_translate_host.restype = C.c_int
        lib.llscan_translate_host.argtypes = [c_double_p, c_double_p, C.c_size_t, c_double_p, c_double_p, C.c_size_t, C.c_size_t, C.c_double]
    _native[name] = lib
    return lib
