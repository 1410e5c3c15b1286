"""ctypes binding of libfimexcu.so (include/fimexcu.h).  Fails loudly when the library is missing."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
lib_path = os.path.join(HERE, "libfimexcu.so")


class FimexCudaError(RuntimeError):
    """Counterpart of MetNoFimex::CDMException for failures inside libfimexcu.so."""


if not os.path.exists(lib_path):
    raise ImportError(
        "fimex_b200/libfimexcu.so is missing: build it with `python -m fimex_b200.build` (needs nvcc). "
        "There is no CPU or PyTorch fallback for the regridding path.")

lib = C.CDLL(lib_path)

c_double_p = C.POINTER(C.c_double)
c_float_p = C.POINTER(C.c_float)
c_size_p = C.POINTER(C.c_size_t)
c_int_p = C.POINTER(C.c_int)
vp = C.c_void_p
sz = C.c_size_t

_two_level = [vp, vp, vp, sz, C.c_double, C.c_double, C.c_double]
_point = [vp, vp, C.c_double, C.c_double, C.c_int, C.c_int, C.c_int]

SIGNATURES = {
    "fimexcu_last_error": (C.c_char_p, []),
    "fimexcu_device_count": (C.c_int, []),
    "fimexcu_set_device": (C.c_int, [C.c_int]),
    "fimexcu_host_alloc": (vp, [sz]),
    "fimexcu_host_free": (None, [vp]),
    "fimexcu_dev_alloc": (vp, [sz]),
    "fimexcu_dev_free": (None, [vp]),
    "fimexcu_memcpy_h2d": (C.c_int, [vp, vp, sz]),
    "fimexcu_memcpy_d2h": (C.c_int, [vp, vp, sz]),
    "fimexcu_memcpy_d2d": (C.c_int, [vp, vp, sz]),
    "fimexcu_memcpy2d_h2d": (C.c_int, [vp, sz, vp, sz, sz, sz, sz, sz]),
    "fimexcu_memcpy2d_d2h": (C.c_int, [vp, sz, vp, sz, sz, sz, sz, sz]),
    "fimexcu_synchronize": (C.c_int, []),
    "fimexcu_cache_create": (vp, [C.c_int, vp, vp, sz, sz, sz, sz]),
    "fimexcu_cache_create_dev": (vp, [C.c_int, vp, vp, sz, sz, sz, sz, vp]),
    "fimexcu_cache_info": (C.c_int, [vp, c_size_p, c_size_p, c_size_p, c_size_p, c_int_p, c_size_p, c_size_p]),
    "fimexcu_cache_apply_f": (C.c_int, [vp, vp, sz, vp, c_size_p, C.c_float, C.c_float]),
    "fimexcu_cache_apply_dev_f": (C.c_int, [vp, vp, sz, vp, C.c_float, C.c_float, vp]),
    "fimexcu_cache_apply_window_f": (C.c_int, [vp, vp, sz, sz, sz, sz, sz, vp, c_size_p, C.c_float, C.c_float]),
    "fimexcu_cache_apply_window_dev_f": (C.c_int, [vp, vp, sz, sz, sz, sz, sz, vp, C.c_float, C.c_float, vp]),
    "fimexcu_cache_replicate": (C.c_int, [vp, C.c_int, c_int_p]),
    "fimexcu_vecrot_replicate": (C.c_int, [vp, C.c_int, c_int_p]),
    "fimexcu_set_staged_kernels": (None, [C.c_int]),
    "fimexcu_cache_set_fast_bicubic": (C.c_int, [vp, C.c_int]),
    "fimexcu_host_pool_trim": (None, []),
    "fimexcu_cache_apply_dev_strided_f": (C.c_int, [vp, vp, sz, sz, sz, vp, sz, sz, C.c_float, C.c_float, vp]),
    "fimexcu_vcoord_levels_dev_f": (C.c_int, [C.c_int, vp, vp, C.c_double, vp, sz, C.c_double, vp, C.c_int, vp, C.c_double, sz, sz, sz, vp, vp]),
    "fimexcu_vcoord_transform_dev_f": (C.c_int, [C.c_int, vp, sz, vp, vp]),
    "fimexcu_get_values_two_level_dev_f": (C.c_int, [C.c_int, vp, vp, vp, sz, C.c_double, C.c_double, C.c_double, vp]),
    "fimexcu_vcoord_pressure_integration_dev_f": (C.c_int, [vp, vp, C.c_double, vp, C.c_double, vp, vp, sz, sz, sz, sz, C.c_int, vp, C.c_double, vp, vp]),
    "fimexcu_fill2d_f": (C.c_int, [sz, sz, vp, C.c_float, C.c_float, sz, c_size_p]),
    "fimexcu_fill2d_dev_f": (C.c_int, [sz, sz, sz, vp, C.c_float, C.c_float, sz, vp, vp]),
    "fimexcu_creepfill2d_f": (C.c_int, [sz, sz, vp, C.c_ushort, C.c_char, c_size_p]),
    "fimexcu_creepfillval2d_f": (C.c_int, [sz, sz, vp, C.c_float, C.c_ushort, C.c_char, c_size_p]),
    "fimexcu_creepfill2d_dev_f": (C.c_int, [sz, sz, sz, vp, C.c_int, C.c_float, C.c_ushort, C.c_char, vp, vp]),
    "fimexcu_unpack_dev_f": (C.c_int, [C.c_int, vp, sz, C.c_double, vp, vp]),
    "fimexcu_unpack_scaled_dev_f": (C.c_int, [C.c_int, vp, sz, C.c_double, C.c_double, C.c_double, vp, vp]),
    "fimexcu_pack_dev_f": (C.c_int, [C.c_int, vp, sz, C.c_double, vp, vp]),
    "fimexcu_cache_apply_i16": (C.c_int, [vp, vp, sz, vp, c_size_p, C.c_double]),
    "fimexcu_cache_apply_dev_i16": (C.c_int, [vp, vp, sz, vp, C.c_double, vp]),
    "fimexcu_cache_padded_layout": (C.c_int, [vp, c_size_p, c_size_p, c_size_p, c_size_p]),
    "fimexcu_cache_row_bands": (C.c_int, [vp, C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_double)]),
    "fimexcu_cache_destroy": (None, [vp]),
    "fimexcu_get_values_f": (C.c_int, _point),
    "fimexcu_get_values_bilinear_f": (C.c_int, _point),
    "fimexcu_get_values_bicubic_f": (C.c_int, _point),
    "fimexcu_vecrot_create": (vp, [vp, sz, sz]),
    "fimexcu_vecrot_create_dev": (vp, [vp, sz, sz, vp]),
    "fimexcu_vecrot_apply_f": (C.c_int, [vp, vp, vp, sz]),
    "fimexcu_vecrot_apply_dev_f": (C.c_int, [vp, vp, vp, sz, vp]),
    "fimexcu_vecrot_apply_dev_strided_f": (C.c_int, [vp, vp, vp, sz, sz, sz, vp]),
    "fimexcu_vecrot_direction_f": (C.c_int, [vp, vp, sz]),
    "fimexcu_vecrot_direction_dev_f": (C.c_int, [vp, vp, sz, vp]),
    "fimexcu_vecrot_destroy": (None, [vp]),
    "fimexcu_cache_apply_uv_f": (C.c_int, [vp, vp, vp, vp, sz, vp, vp, c_size_p] + [C.c_float] * 4),
    "fimexcu_cache_apply_uv_dev_f": (C.c_int, [vp, vp, vp, vp, sz, vp, vp] + [C.c_float] * 4 + [vp]),
    "fimexcu_cache_apply_uv_dev_strided_f": (C.c_int, [vp, vp, vp, vp, sz, sz, sz, vp, vp, sz, sz] + [C.c_float] * 4 + [vp]),
    "fimexcu_points2position": (C.c_int, [vp, sz, vp, C.c_int, C.c_int]),
    "fimexcu_points2position_dev": (C.c_int, [vp, sz, vp, C.c_int, C.c_int, vp]),
    "fimexcu_points2position_dev_noalloc": (C.c_int, [vp, sz, vp, C.c_int, C.c_int, vp]),
    "fimexcu_reproject_axes": (C.c_int, [C.c_char_p, C.c_char_p, vp, vp, C.c_int, C.c_int, vp, vp]),
    "fimexcu_reproject_axes_dev": (C.c_int, [C.c_char_p, C.c_char_p, vp, vp, C.c_int, C.c_int, vp, vp, vp]),
    "fimexcu_reproject_values": (C.c_int, [C.c_char_p, C.c_char_p, vp, vp, C.c_int]),
    "fimexcu_reproject_values_dev": (C.c_int, [C.c_char_p, C.c_char_p, vp, vp, sz, vp]),
    "fimexcu_get_vector_reproject_matrix": (C.c_int, [C.c_char_p, C.c_char_p, vp, vp, C.c_int, C.c_int, vp]),
    "fimexcu_get_vector_reproject_matrix_field": (C.c_int, [C.c_char_p, C.c_char_p, vp, vp, C.c_int, C.c_int, vp]),
    "fimexcu_get_vector_reproject_matrix_points": (C.c_int, [C.c_char_p, C.c_char_p, C.c_int, vp, vp, C.c_int, C.c_int, vp]),
    "fimexcu_cache_create_from_projection": (vp, [C.c_int, C.c_char_p, C.c_char_p, vp, vp, sz, sz, vp, vp, sz, sz, C.c_int, C.c_int]),
    "fimexcu_cache_create_from_points": (vp, [C.c_int, C.c_char_p, C.c_char_p, vp, vp, sz, sz, vp, vp, sz, sz, C.c_int, C.c_int]),
    "fimexcu_kdtree_translate": (C.c_int, [C.c_double, vp, vp, sz, vp, vp, sz, sz]),
    "fimexcu_kdtree_translate_mode": (C.c_int, [C.c_double, vp, vp, sz, vp, vp, sz, sz, C.c_int]),
    "fimexcu_coordnn_translate": (C.c_int, [vp, vp, sz, vp, vp, sz, sz]),
    "fimexcu_forward_create": (vp, [C.c_int, vp, vp, sz, sz, sz, sz]),
    "fimexcu_forward_info": (C.c_int, [vp, c_size_p, c_size_p, c_size_p, c_size_p, c_int_p, c_size_p, c_size_p, c_size_p]),
    "fimexcu_forward_apply_f": (C.c_int, [vp, vp, sz, vp, c_size_p, C.c_float, C.c_float, C.c_int]),
    "fimexcu_forward_apply_dev_f": (C.c_int, [vp, vp, sz, vp, C.c_float, C.c_float, C.c_int, vp]),
    "fimexcu_forward_destroy": (None, [vp]),
    "fimexcu_vertical_interpolate_f": (C.c_int, [C.c_int, vp, vp, sz, sz, sz, vp, vp, sz, vp, vp, C.c_double, C.c_double, C.c_float,
                                                 C.c_float, vp]),
    "fimexcu_vertical_interpolate_dev_f": (C.c_int, [C.c_int, vp, vp, sz, sz, sz, vp, vp, sz, vp, vp, C.c_double, C.c_double, C.c_float,
                                                     C.c_float, vp, vp]),
    "fimexcu_vertical_hybrid_pressure_f": (C.c_int, [C.c_int, vp, sz, sz, sz, vp, vp, vp, sz, C.c_double, vp, sz, C.c_double, C.c_double,
                                                     C.c_float, C.c_float, vp]),
    "fimexcu_vertical_hybrid_pressure_dev_f": (C.c_int, [C.c_int, vp, sz, sz, sz, vp, vp, vp, sz, C.c_double, vp, sz, C.c_double,
                                                         C.c_double, C.c_float, C.c_float, vp, vp]),
    "fimexcu_vertical_interpolate_rows_dev_f": (C.c_int, [C.c_int, vp, vp, sz, sz, sz, vp, vp, sz, vp, vp, C.c_double, C.c_double, C.c_float,
                                                          C.c_float, vp, sz, sz, vp]),
    "fimexcu_vertical_hybrid_pressure_rows_dev_f": (C.c_int, [C.c_int, vp, sz, sz, sz, vp, vp, vp, sz, C.c_double, vp, sz, C.c_double,
                                                              C.c_double, C.c_float, C.c_float, vp, sz, sz, vp]),
    "fimexcu_get_values_nearest_f": (C.c_int, _two_level),
    "fimexcu_get_values_linear_f": (C.c_int, _two_level),
    "fimexcu_get_values_linear_weak_extrapol_f": (C.c_int, _two_level),
    "fimexcu_get_values_linear_no_extrapol_f": (C.c_int, _two_level),
    "fimexcu_get_values_linear_const_extrapol_f": (C.c_int, _two_level),
    "fimexcu_get_values_log_f": (C.c_int, _two_level),
    "fimexcu_get_values_log_log_f": (C.c_int, _two_level),
}

missing_symbols = []
for _name, (_res, _args) in SIGNATURES.items():
    try:
        _f = getattr(lib, _name)
    except AttributeError:
        missing_symbols.append(_name)
        continue
    _f.restype = _res
    _f.argtypes = _args


def last_error():
    msg = lib.fimexcu_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(rc, what):
    """MIFI_OK/MIFI_ERROR -> exception, like the `throw CDMException("error during interpolation")` of the reference."""
    if rc != 1:
        raise FimexCudaError(f"{what}: {last_error() or 'error'}")


def check_handle(h, what):
    if not h:
        raise FimexCudaError(f"{what}: {last_error() or 'error'}")
    return h


def np_ptr(a):
    return a.ctypes.data


def current_stream_ptr():
    """cudaStream_t of torch's current stream (torch is plumbing for device memory and streams only)."""
    import torch
    return torch.cuda.current_stream().cuda_stream
