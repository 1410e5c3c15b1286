"""Host-side mirror of the vertical-interpolation hot loop on top of the C ABI.

  verticalInterpolate        CDMVerticalInterpolator::getLevelDataSlice hot loop (src/CDMVerticalInterpolator.cc:488-561)
  verticalHybridPressure     same with the hybrid-sigma pressure formed on the fly
                             (src/coordSys/verticalTransform/HybridSigmaApToPressureConverter.cc:71-100)
  mifi_get_values_*_f        the two-level kernels of include/fimex/interpolation.h:153-232
Arrays are [nOuter][nz][nInner] (C order): the CDM shape (x, y, [ensemble], k, t) of a time-step slab has nInner = the product
of the dimensions faster than the vertical one and nOuter = the product of the slower ones.
"""
import numpy as np

from . import _lib
from ._lib import lib

_NAN = float("nan")


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _opt(a, dtype):
    return None if a is None else np.ascontiguousarray(a, dtype=dtype)


def _ptr(a):
    return None if a is None else a.ctypes.data


def verticalInterpolate(method, inData, verticalIn, level1=None, verticalOut=None, validMin=None, validMax=None, valid_min=_NAN,
                        valid_max=_NAN, badValue=_NAN, fillValue=_NAN, rows=None, out=None):
    """inData, verticalIn: [nOuter][nzi][nInner] f32; level1: fixed target levels (f64) or verticalOut: [nOuter][nzo][nInner] f32.
    validMin/validMax: per-column f64 [nOuter][nInner] (converter validity range); valid_min/valid_max: the variable's
    valid range (clamp).  Returns [nOuter][nzo][nInner] f32.
    rows = (rowWidth, rowPitch), CUDA tensors only: the inner index runs over padded rows (the result of a device-resident
    horizontal step); the pad elements are neither read nor written.  out: a CUDA tensor to write into (device path)."""
    if _is_torch(inData):
        import torch
        nO, nzi, nI = inData.shape
        lv = _opt(level1, np.float64)
        nzo = lv.size if lv is not None else verticalOut.shape[1]
        if out is None:
            out = torch.empty((nO, nzo, nI), dtype=torch.float32, device=inData.device)
        assert out.is_cuda and out.is_contiguous() and out.dtype == torch.float32 and out.numel() == nO * nzo * nI
        rw, rp = rows if rows is not None else (0, 0)
        _lib.check(lib.fimexcu_vertical_interpolate_rows_dev_f(
            method, inData.data_ptr(), verticalIn.data_ptr(), nO, nzi, nI, _ptr(lv), None if verticalOut is None else verticalOut.data_ptr(), nzo,
            None if validMin is None else validMin.data_ptr(), None if validMax is None else validMax.data_ptr(), valid_min, valid_max,
            badValue, fillValue, out.data_ptr(), rw, rp, _lib.current_stream_ptr()), "verticalInterpolate")
        return out
    if rows is not None:
        raise ValueError("rows: device slabs only")
    a, v = _f32(inData), _f32(verticalIn)
    nO, nzi, nI = a.shape
    assert v.shape == a.shape
    lv, vo = _opt(level1, np.float64), _opt(verticalOut, np.float32)
    nzo = lv.size if lv is not None else vo.shape[1]
    mn, mx = _opt(validMin, np.float64), _opt(validMax, np.float64)
    out = np.empty((nO, nzo, nI), dtype=np.float32)
    _lib.check(lib.fimexcu_vertical_interpolate_f(method, a.ctypes.data, v.ctypes.data, nO, nzi, nI, _ptr(lv), _ptr(vo), nzo, _ptr(mn), _ptr(mx),
                                                  valid_min, valid_max, badValue, fillValue, out.ctypes.data), "verticalInterpolate")
    return out


def verticalHybridPressure(method, inData, ap, b, ps, level1, ps_scale=1.0, valid_min=_NAN, valid_max=_NAN, badValue=_NAN, fillValue=_NAN, rows=None,
                           out=None):
    """inData [nOuter][nzi][nInner] f32; ap, b [nzi] f64 (ap in the unit of level1); ps [nOuter or 1][nInner] f32, multiplied by
    ps_scale (unit conversion); level1 target pressures.  rows = (rowWidth, rowPitch): as verticalInterpolate."""
    apv, bv, lv = _opt(ap, np.float64), _opt(b, np.float64), _opt(level1, np.float64)
    if _is_torch(inData):
        import torch
        nO, nzi, nI = inData.shape
        if out is None:
            out = torch.empty((nO, lv.size, nI), dtype=torch.float32, device=inData.device)
        assert out.is_cuda and out.is_contiguous() and out.dtype == torch.float32 and out.numel() == nO * lv.size * nI
        rw, rp = rows if rows is not None else (0, 0)
        _lib.check(lib.fimexcu_vertical_hybrid_pressure_rows_dev_f(method, inData.data_ptr(), nO, nzi, nI, apv.ctypes.data, bv.ctypes.data, ps.data_ptr(),
                                                                   ps.shape[0], ps_scale, lv.ctypes.data, lv.size, valid_min, valid_max, badValue,
                                                                   fillValue, out.data_ptr(), rw, rp, _lib.current_stream_ptr()),
                   "verticalHybridPressure")
        return out
    if rows is not None:
        raise ValueError("rows: device slabs only")
    a, p = _f32(inData), _f32(ps)
    nO, nzi, nI = a.shape
    p = p.reshape(-1, nI)
    out = np.empty((nO, lv.size, nI), dtype=np.float32)
    _lib.check(lib.fimexcu_vertical_hybrid_pressure_f(method, a.ctypes.data, nO, nzi, nI, apv.ctypes.data, bv.ctypes.data, p.ctypes.data, p.shape[0],
                                                      ps_scale, lv.ctypes.data, lv.size, valid_min, valid_max, badValue, fillValue, out.ctypes.data),
               "verticalHybridPressure")
    return out


def _two_level(fn, A, B, a, b, x):
    A, B = _f32(A).ravel(), _f32(B).ravel()
    out = np.full(A.shape, 12345, dtype=np.float32)
    rc = fn(A.ctypes.data, B.ctypes.data, out.ctypes.data, A.size, float(a), float(b), float(x))
    return out, rc


def mifi_get_values_nearest_f(A, B, a, b, x):
    return _two_level(lib.fimexcu_get_values_nearest_f, A, B, a, b, x)


def mifi_get_values_linear_f(A, B, a, b, x):
    return _two_level(lib.fimexcu_get_values_linear_f, A, B, a, b, x)


def mifi_get_values_linear_weak_extrapol_f(A, B, a, b, x):
    return _two_level(lib.fimexcu_get_values_linear_weak_extrapol_f, A, B, a, b, x)


def mifi_get_values_linear_no_extrapol_f(A, B, a, b, x):
    return _two_level(lib.fimexcu_get_values_linear_no_extrapol_f, A, B, a, b, x)


def mifi_get_values_linear_const_extrapol_f(A, B, a, b, x):
    return _two_level(lib.fimexcu_get_values_linear_const_extrapol_f, A, B, a, b, x)


def mifi_get_values_log_f(A, B, a, b, x):
    return _two_level(lib.fimexcu_get_values_log_f, A, B, a, b, x)


def mifi_get_values_log_log_f(A, B, a, b, x):
    return _two_level(lib.fimexcu_get_values_log_log_f, A, B, a, b, x)


# ---- vertical coordinate converters (device-resident; src/coordSys/verticalTransform/*) ----
VCOORD_HYBRID_SIGMA_AP, VCOORD_HYBRID_SIGMA, VCOORD_SIGMA, VCOORD_LN_PRESSURE, VCOORD_HYBRID_HEIGHT, VCOORD_OCEAN_S_G1, VCOORD_OCEAN_S_G2 = range(7)
VPOST_NONE, VPOST_STANDARD_ALTITUDE, VPOST_STANDARD_PRESSURE = range(3)


def _dev(a):
    """numpy -> CUDA tensor (the converters only exist on the device: there is no CPU path)"""
    import torch
    if a is None:
        return None
    if _is_torch(a):
        if not a.is_cuda or a.dtype != torch.float32 or not a.is_contiguous():
            raise _lib.FimexCudaError("needs a contiguous float32 CUDA tensor")
        return a
    return torch.from_numpy(_f32(a)).cuda()


def _dptr(t):
    return None if t is None else t.data_ptr()


def verticalLevelField(kind, a, b, scalar, surface, nz=None, surface_scale=1.0, eta=None, post=VPOST_NONE, topography=None, topography_factor=0.0,
                       nOuter=None, nInner=None):
    """The converter chain  level formula -> standard-atmosphere step -> +- topography  as one device pass.
    a, b: per-level terms (f64, host); surface: [nOuterS][nInner] f32 (ps / orography / depth) or None for ln_pressure (then give
    nOuter, nInner); returns a CUDA tensor [nOuter][nz][nInner] f32 that can be passed as verticalIn / verticalOut."""
    import torch
    av, bv = _opt(a, np.float64), _opt(b, np.float64)
    nz = av.size if nz is None else nz
    s, e, t = _dev(surface), _dev(eta), _dev(topography)
    if s is not None:
        s = s.reshape(-1, s.shape[-1])
        nOuterS, nI = s.shape
        nO = nOuterS if nOuter is None else nOuter
    else:
        nOuterS, nO, nI = 1, nOuter, nInner
    out = torch.empty((nO, nz, nI), dtype=torch.float32, device="cuda")
    _lib.check(lib.fimexcu_vcoord_levels_dev_f(kind, av.ctypes.data, _ptr(bv), float(scalar), _dptr(s), nOuterS, float(surface_scale), _dptr(e), post,
                                               _dptr(t), float(topography_factor), nO, nz, nI, out.data_ptr(), _lib.current_stream_ptr()),
               "verticalLevelField")
    return out


def verticalStandardAtmosphere(post, field):
    """pressure [hPa] -> standard altitude or altitude [m] -> standard pressure on a device field"""
    import torch
    f = _dev(field)
    out = torch.empty_like(f)
    _lib.check(lib.fimexcu_vcoord_transform_dev_f(post, f.data_ptr(), f.numel(), out.data_ptr(), _lib.current_stream_ptr()), "verticalStandardAtmosphere")
    return out


def pressureIntegrationAltitude(pressure, surfacePressure, surfaceGeopotential, airTemperature, specificHumidity=None, sap_scale=1.0, sgp_scale=1.0,
                                start_high_p=-1, topography=None, topography_factor=0.0):
    """PressureIntegrationToAltitudeConverter::getDataSlice on the device: pressure / airTemperature / specificHumidity
    [nOuter][nz][nInner] f32, surface fields [nOuterS][nInner]; returns the altitude of every level (CUDA tensor)."""
    import torch
    p, T, q = _dev(pressure), _dev(airTemperature), _dev(specificHumidity)
    sap, sgp, t = _dev(surfacePressure), _dev(surfaceGeopotential), _dev(topography)
    nO, nz, nI = p.shape
    sap, sgp = sap.reshape(-1, nI), sgp.reshape(-1, nI)
    out = torch.empty_like(p)
    _lib.check(lib.fimexcu_vcoord_pressure_integration_dev_f(p.data_ptr(), sap.data_ptr(), float(sap_scale), sgp.data_ptr(), float(sgp_scale), T.data_ptr(),
                                                             _dptr(q), nO, sap.shape[0], nz, nI, int(start_high_p), _dptr(t), float(topography_factor),
                                                             out.data_ptr(), _lib.current_stream_ptr()), "pressureIntegrationAltitude")
    return out


def timeInterpolateSlabs(sliceA, sliceB, tA, tB, t, method=4):
    """CDMTimeInterpolator::getDataSlice between two device-resident time slices (src/CDMTimeInterpolator.cc:117-131): the
    two-level kernel selected by `method` (default linear_weak_extrapol like the reference) over whole slabs; CUDA tensors."""
    import torch
    a, b = _dev(sliceA), _dev(sliceB)
    if a.numel() != b.numel():
        raise _lib.FimexCudaError("got slices with different size")
    out = torch.empty_like(a)
    _lib.check(lib.fimexcu_get_values_two_level_dev_f(method, a.data_ptr(), b.data_ptr(), out.data_ptr(), a.numel(), float(tA), float(tB), float(t),
                                                      _lib.current_stream_ptr()), "timeInterpolateSlabs")
    return out
