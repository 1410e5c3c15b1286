"""Host-side mirror of the reference's interpolation-cache interface on top of the C ABI.

Same names, argument meaning and error behaviour as
  include/fimex/CachedInterpolation.h:55-189      CachedInterpolationInterface / CachedInterpolation / CachedNNInterpolation
  src/CachedForwardInterpolation.h:36-50          CachedForwardInterpolation
  include/fimex/CachedVectorReprojection.h:38-63  CachedVectorReprojection
  include/fimex/interpolation.h:76-93             mifi_points2position
so the parity tests read like the reference's own tests.  numpy arrays go through the host entry points
(copies included); CUDA torch tensors go through the *_dev entry points on torch's current stream and stay
resident.  The dimension-name arguments of the reference constructors only serve the CDM slicing helpers
(getInputDataSlice/getOutputDataSlice), which stay with the reference, and are accepted and ignored.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import FimexCudaError, lib
from .constants import (MIFI_INTERPOL_BICUBIC, MIFI_INTERPOL_BILINEAR, MIFI_INTERPOL_COORD_NN, MIFI_INTERPOL_COORD_NN_KD,
                        MIFI_INTERPOL_NEAREST_NEIGHBOR)

_NAN = float("nan")


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class ReducedInterpolationDomain:
    """include/fimex/CachedInterpolation.h:42-51"""

    def __init__(self, xDim, yDim, xMin, yMin):
        self.xDim, self.yDim, self.xMin, self.yMin = xDim, yDim, xMin, yMin


class CachedInterpolationInterface:
    """include/fimex/CachedInterpolation.h:55-112"""

    _destroy = staticmethod(lib.fimexcu_cache_destroy)

    def __init__(self, xDimName="x", yDimName="y"):
        self._xDimName, self._yDimName = xDimName, yDimName
        self._h = None
        self._reduced = None

    def _read_info(self):
        v = [C.c_size_t() for _ in range(4)]
        red, xm, ym = C.c_int(), C.c_size_t(), C.c_size_t()
        _lib.check(lib.fimexcu_cache_info(self._h, *[C.byref(t) for t in v], C.byref(red), C.byref(xm), C.byref(ym)), "cache_info")
        self.inX, self.inY, self.outX, self.outY = (t.value for t in v)
        self._reduced = ReducedInterpolationDomain(self._xDimName, self._yDimName, xm.value, ym.value) if red.value else None

    def getInX(self):
        return self.inX

    def getInY(self):
        return self.inY

    def getOutX(self):
        return self.outX

    def getOutY(self):
        return self.outY

    def reducedDomain(self):
        return self._reduced

    def reduceInput(self, full, fullInX, fullInY):
        """What getInputDataSlice() (src/CachedInterpolation.cc:68-111) asks the upstream reader for: the
        reduced x/y window of a full [.., fullInY, fullInX] slab."""
        a = full.reshape(-1, fullInY, fullInX)
        if self._reduced is None:
            return a
        r = self._reduced
        sub = a[:, r.yMin:r.yMin + self.inY, r.xMin:r.xMin + self.inX]
        return sub.contiguous() if _is_torch(sub) else np.ascontiguousarray(sub)

    def sliceOutput(self, out, xStart=0, xSize=None, yStart=0, ySize=None):
        """What getOutputDataSlice() (src/CachedInterpolation.cc:113-130) hands the caller: the requested x/y sub-window of the
        full interpolated slab [.., outY, outX]; the other dimensions were already cut by getInputDataSlice()."""
        a = out.reshape(-1, self.outY, self.outX)
        xSize = self.outX - xStart if xSize is None else xSize
        ySize = self.outY - yStart if ySize is None else ySize
        if xStart + xSize > self.outX or yStart + ySize > self.outY:
            raise FimexCudaError("requested output window lies outside the target grid")
        sub = a[:, yStart:yStart + ySize, xStart:xStart + xSize]
        return sub.contiguous() if _is_torch(sub) else np.ascontiguousarray(sub)

    def interpolateValues(self, inData, size=None, badValue=_NAN, fillValue=_NAN):
        """interpolateValues(inData, size, newSize) -> (outData, newSize).
        inData: [nz][inY][inX] f32 (numpy: host path; CUDA torch tensor: device path).  badValue/fillValue
        fuse data2InterpolationArray / interpolationArray2Data (CDMInterpolator.cc:140-151); NaN = off."""
        if self._h is None:
            raise FimexCudaError("cache was destroyed")
        if _is_torch(inData):
            import torch
            if not inData.is_cuda or inData.dtype != torch.float32 or not inData.is_contiguous():
                raise FimexCudaError("device path needs a contiguous float32 CUDA tensor")
            size = inData.numel() if size is None else size
            nz = size // (self.inX * self.inY)
            out = torch.empty((nz, self.outY, self.outX), dtype=torch.float32, device=inData.device)
            self._apply_dev(inData.data_ptr(), nz, out.data_ptr(), badValue, fillValue, _lib.current_stream_ptr())
            return out, nz * self.outX * self.outY
        a = np.ascontiguousarray(inData, dtype=np.float32)
        size = a.size if size is None else size
        nz = size // (self.inX * self.inY)
        out = np.empty((nz, self.outY, self.outX), dtype=np.float32)
        newSize = C.c_size_t()
        self._apply_host(a.ctypes.data, size, out.ctypes.data, newSize, badValue, fillValue)
        return out, newSize.value

    def interpolateValuesWindow(self, inData, xStart, xSize, yStart, ySize, badValue=_NAN, fillValue=_NAN):
        """interpolateValues + getOutputDataSlice (src/CachedInterpolation.cc:113-130) for a caller that wants only the cells
        [yStart, yStart+ySize) x [xStart, xStart+xSize) of the target grid: only those are computed and copied back.
        Returns [nz][ySize][xSize] (numpy in -> numpy out, CUDA tensor in -> CUDA tensor out)."""
        if self._h is None:
            raise FimexCudaError("cache was destroyed")
        if _is_torch(inData):
            import torch
            if not inData.is_cuda or inData.dtype != torch.float32 or not inData.is_contiguous():
                raise FimexCudaError("device path needs a contiguous float32 CUDA tensor")
            nz = inData.numel() // (self.inX * self.inY)
            out = torch.empty((nz, ySize, xSize), dtype=torch.float32, device=inData.device)
            _lib.check(lib.fimexcu_cache_apply_window_dev_f(self._h, inData.data_ptr(), nz, xStart, xSize, yStart, ySize, out.data_ptr(), badValue, fillValue,
                                                            _lib.current_stream_ptr()), "interpolateValuesWindow")
            return out
        a = np.ascontiguousarray(inData, dtype=np.float32)
        nz = a.size // (self.inX * self.inY)
        out = np.empty((nz, ySize, xSize), dtype=np.float32)
        newSize = C.c_size_t()
        _lib.check(lib.fimexcu_cache_apply_window_f(self._h, a.ctypes.data, a.size, xStart, xSize, yStart, ySize, out.ctypes.data, C.byref(newSize),
                                                    badValue, fillValue), "interpolateValuesWindow")
        assert newSize.value == out.size
        return out

    def setFastBicubic(self, on=True):
        """opt-in f32 FMA arithmetic of the staged bicubic kernel (1e-5 contract instead of bit-exact); see fimexcu_cache_set_fast_bicubic"""
        _lib.check(lib.fimexcu_cache_set_fast_bicubic(self._h, 1 if on else 0), "set_fast_bicubic")

    def replicate(self, devices):
        """copy the tables to the listed device ordinals: host threads bound to different GPUs then share this one object, as the
        reference's OpenMP writer threads share one CachedInterpolation (src/io/netcdf/NetCDF_CDMWriter.cc:599-618)"""
        ids = (C.c_int * len(devices))(*devices)
        _lib.check(lib.fimexcu_cache_replicate(self._h, len(devices), ids), "replicate")

    def interpolateValuesPacked(self, inData, badValue):
        """CDMInterpolator::getDataSlice for a CDM_SHORT variable around the cached interpolation (CDMInterpolator.cc:140-151):
        inData [nz][inY][inX] int16 RAW packed values, badValue = the variable's _FillValue; returns int16 [nz][outY][outX].
        numpy array: host path (2 bytes per value over PCIe each way); CUDA torch int16 tensor: device path."""
        if self._h is None:
            raise FimexCudaError("cache was destroyed")
        if _is_torch(inData):
            import torch
            if not inData.is_cuda or inData.dtype != torch.int16 or not inData.is_contiguous():
                raise FimexCudaError("device path needs a contiguous int16 CUDA tensor")
            nz = inData.numel() // (self.inX * self.inY)
            out = torch.empty((nz, self.outY, self.outX), dtype=torch.int16, device=inData.device)
            _lib.check(lib.fimexcu_cache_apply_dev_i16(self._h, inData.data_ptr(), nz, out.data_ptr(), float(badValue), _lib.current_stream_ptr()),
                       "cache_apply_dev_i16")
            return out
        a = np.ascontiguousarray(inData, dtype=np.int16)
        nz = a.size // (self.inX * self.inY)
        out = np.empty((nz, self.outY, self.outX), dtype=np.int16)
        newSize = C.c_size_t()
        _lib.check(lib.fimexcu_cache_apply_i16(self._h, a.ctypes.data, a.size, out.ctypes.data, C.byref(newSize), float(badValue)), "cache_apply_i16")
        return out

    def paddedLayout(self):
        """(in_row_pitch, in_level_stride, out_row_pitch, out_level_stride) in elements of the device-resident layout
        that the staged kernels like best: source rows on 32 B, target rows on 128 B (fimexcu_cache_padded_layout)."""
        v = [C.c_size_t() for _ in range(4)]
        _lib.check(lib.fimexcu_cache_padded_layout(self._h, *[C.byref(t) for t in v]), "padded_layout")
        return tuple(t.value for t in v)

    def rowBands(self):
        """([(y0, y1, x0, x1), ...], fraction): the bands of target rows with the column span outside which the result is undefined
        whatever the input; only these rectangles are copied back by the host entry points (fimexcu_cache_row_bands)"""
        buf = (C.c_int * (4 * 64))()
        frac = C.c_double(1.0)
        n = lib.fimexcu_cache_row_bands(self._h, buf, 64, C.byref(frac))
        return [tuple(buf[4 * k:4 * k + 4]) for k in range(n)], frac.value

    def toPadded(self, dense):
        """copy a dense CUDA slab [nz][inY][inX] into a padded one; returns the padded tensor (a view [..., :inX] is the data)"""
        import torch
        ip, il, _, _ = self.paddedLayout()
        nz = dense.numel() // (self.inX * self.inY)
        buf = torch.empty((nz, self.inY, ip), dtype=torch.float32, device=dense.device)
        buf[:, :, :self.inX].copy_(dense.reshape(nz, self.inY, self.inX))
        return buf

    def interpolateValuesPadded(self, paddedIn, out=None, badValue=_NAN, fillValue=_NAN):
        """device-resident apply in the padded layout: paddedIn [nz][inY][in_row_pitch] -> out [nz][outY][out_row_pitch]
        (the defined data is out[..., :outX])."""
        import torch
        ip, il, op, ol = self.paddedLayout()
        nz = paddedIn.shape[0]
        assert paddedIn.is_cuda and paddedIn.is_contiguous() and paddedIn.shape[1:] == (self.inY, ip)
        if out is None:
            out = torch.empty((nz, self.outY, op), dtype=torch.float32, device=paddedIn.device)
        _lib.check(lib.fimexcu_cache_apply_dev_strided_f(self._h, paddedIn.data_ptr(), ip, il, nz, out.data_ptr(), op, ol, badValue, fillValue,
                                                         _lib.current_stream_ptr()), "interpolateValues")
        return out

    def _apply_host(self, pin, size, pout, newSize, bad, fill):
        _lib.check(lib.fimexcu_cache_apply_f(self._h, pin, size, pout, C.byref(newSize), bad, fill), "interpolateValues")

    def _apply_dev(self, pin, nz, pout, bad, fill, stream):
        _lib.check(lib.fimexcu_cache_apply_dev_f(self._h, pin, nz, pout, bad, fill, stream), "interpolateValues")

    def close(self):
        if self._h:
            self._destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class CachedInterpolation(CachedInterpolationInterface):
    """include/fimex/CachedInterpolation.h:124-158 (bilinear, bicubic)"""

    def __init__(self, xDimName, yDimName, funcType, pointsOnXAxis, pointsOnYAxis, inX, inY, outX, outY):
        super().__init__(xDimName, yDimName)
        if funcType not in (MIFI_INTERPOL_BILINEAR, MIFI_INTERPOL_BICUBIC):
            raise FimexCudaError(f"CachedInterpolation supports only bilinear and bicubic, not: {funcType}")
        self._h = _create_cache(funcType, pointsOnXAxis, pointsOnYAxis, inX, inY, outX, outY)
        self._read_info()


class CachedNNInterpolation(CachedInterpolationInterface):
    """include/fimex/CachedInterpolation.h:164-189"""

    def __init__(self, xDimName, yDimName, pointsOnXAxis, pointsOnYAxis, inX, inY, outX, outY):
        super().__init__(xDimName, yDimName)
        self._h = _create_cache(MIFI_INTERPOL_NEAREST_NEIGHBOR, pointsOnXAxis, pointsOnYAxis, inX, inY, outX, outY)
        self._read_info()


def _create_cache(method, px, py, inX, inY, outX, outY):
    if _is_torch(px):
        h = lib.fimexcu_cache_create_dev(method, px.data_ptr(), py.data_ptr(), inX, inY, outX, outY, _lib.current_stream_ptr())
    else:
        px, py = _f64(px).ravel(), _f64(py).ravel()
        if px.size != outX * outY or py.size != outX * outY:
            raise FimexCudaError("index tables must have outX*outY entries")
        h = lib.fimexcu_cache_create(method, px.ctypes.data, py.ctypes.data, inX, inY, outX, outY)
    return _lib.check_handle(h, "createCachedInterpolation")


def createCachedInterpolation(xDimName, yDimName, method, pointsOnXAxis, pointsOnYAxis, inX, inY, outX, outY):
    """src/CachedInterpolation.cc:132-147"""
    if method in (MIFI_INTERPOL_BILINEAR, MIFI_INTERPOL_BICUBIC):
        return CachedInterpolation(xDimName, yDimName, method, pointsOnXAxis, pointsOnYAxis, inX, inY, outX, outY)
    if method in (MIFI_INTERPOL_NEAREST_NEIGHBOR, MIFI_INTERPOL_COORD_NN, MIFI_INTERPOL_COORD_NN_KD):
        return CachedNNInterpolation(xDimName, yDimName, pointsOnXAxis, pointsOnYAxis, inX, inY, outX, outY)
    raise FimexCudaError(f"unknown interpolation function: {method}")


class CachedForwardInterpolation(CachedInterpolationInterface):
    """src/CachedForwardInterpolation.h:36-50.  pointsOnXAxis/pointsOnYAxis hold, for every SOURCE cell, its
    position on the target axes.  level_compat=True reproduces the reference literally (every output level
    aggregates input level 0, src/CachedForwardInterpolation.cc:212-221)."""

    _destroy = staticmethod(getattr(lib, "fimexcu_forward_destroy", None))

    def __init__(self, xDimName, yDimName, funcType, pointsOnXAxis, pointsOnYAxis, inX, inY, outX, outY, level_compat=True):
        super().__init__(xDimName, yDimName)
        px, py = _f64(pointsOnXAxis).ravel(), _f64(pointsOnYAxis).ravel()
        if px.size != inX * inY or py.size != inX * inY:
            raise FimexCudaError("forward tables must have inX*inY entries")
        self.level_compat = 1 if level_compat else 0
        self._h = _lib.check_handle(lib.fimexcu_forward_create(funcType, px.ctypes.data, py.ctypes.data, inX, inY, outX, outY),
                                    "CachedForwardInterpolation")
        self._read_info()

    def _read_info(self):
        v = [C.c_size_t() for _ in range(4)]
        red, xm, ym, mp = C.c_int(), C.c_size_t(), C.c_size_t(), C.c_size_t()
        _lib.check(lib.fimexcu_forward_info(self._h, *[C.byref(t) for t in v], C.byref(red), C.byref(xm), C.byref(ym), C.byref(mp)),
                   "forward_info")
        self.inX, self.inY, self.outX, self.outY = (t.value for t in v)
        self.maxPointsInIn = mp.value
        self._reduced = ReducedInterpolationDomain(self._xDimName, self._yDimName, xm.value, ym.value) if red.value else None

    def interpolateValuesPacked(self, inData, badValue):
        raise FimexCudaError("packed slabs are not supported by the forward aggregation")

    def _apply_host(self, pin, size, pout, newSize, bad, fill):
        _lib.check(lib.fimexcu_forward_apply_f(self._h, pin, size, pout, C.byref(newSize), bad, fill, self.level_compat), "interpolateValues")

    def _apply_dev(self, pin, nz, pout, bad, fill, stream):
        _lib.check(lib.fimexcu_forward_apply_dev_f(self._h, pin, nz, pout, bad, fill, self.level_compat, stream), "interpolateValues")


class CachedVectorReprojection:
    """include/fimex/CachedVectorReprojection.h:38-63; `matrix` is the ox*oy*3 array of reproject::Matrix
    (cos, sin, angle in degree per target cell)."""

    def __init__(self, matrix, ox=None, oy=None):
        if _is_torch(matrix):
            self.ox, self.oy = ox, oy
            h = lib.fimexcu_vecrot_create_dev(matrix.data_ptr(), ox, oy, _lib.current_stream_ptr())
        else:
            m = _f64(matrix)
            if ox is None:
                oy, ox = m.shape[0], m.shape[1]
            self.ox, self.oy = ox, oy
            if m.size != ox * oy * 3:
                raise FimexCudaError("matrix must have ox*oy*3 entries")
            h = lib.fimexcu_vecrot_create(m.ctypes.data, ox, oy)
        self._h = _lib.check_handle(h, "CachedVectorReprojection")

    def getXSize(self):
        return self.ox

    def getYSize(self):
        return self.oy

    def reprojectValues(self, uValues, vValues, size=None):
        """in place, like the reference (src/CachedVectorReprojection.cc:44-48)"""
        if _is_torch(uValues):
            nz = (uValues.numel() if size is None else size) // (self.ox * self.oy)
            _lib.check(lib.fimexcu_vecrot_apply_dev_f(self._h, uValues.data_ptr(), vValues.data_ptr(), nz, _lib.current_stream_ptr()),
                       "reprojectValues")
            return
        assert uValues.dtype == np.float32 and vValues.dtype == np.float32 and uValues.flags.c_contiguous and vValues.flags.c_contiguous
        size = uValues.size if size is None else size
        _lib.check(lib.fimexcu_vecrot_apply_f(self._h, uValues.ctypes.data, vValues.ctypes.data, size), "reprojectValues")

    def reprojectValuesPadded(self, uPadded, vPadded):
        """in place on CUDA slabs [nz][oy][row_pitch] (the padded rows of a device-resident chain); the pad elements are not touched"""
        assert uPadded.is_cuda and uPadded.is_contiguous() and vPadded.is_contiguous() and uPadded.shape == vPadded.shape
        nz, oy, pitch = uPadded.shape
        assert oy == self.oy and pitch >= self.ox
        _lib.check(lib.fimexcu_vecrot_apply_dev_strided_f(self._h, uPadded.data_ptr(), vPadded.data_ptr(), pitch, oy * pitch, nz, _lib.current_stream_ptr()),
                   "reprojectValues")

    def replicate(self, devices):
        """the matrix on every listed device (see CachedInterpolationInterface.replicate)"""
        ids = (C.c_int * len(devices))(*devices)
        _lib.check(lib.fimexcu_vecrot_replicate(self._h, len(devices), ids), "replicate")

    def reprojectDirectionValues(self, angles, size=None):
        """src/CachedVectorReprojection.cc:50-54"""
        if _is_torch(angles):
            nz = (angles.numel() if size is None else size) // (self.ox * self.oy)
            _lib.check(lib.fimexcu_vecrot_direction_dev_f(self._h, angles.data_ptr(), nz, _lib.current_stream_ptr()), "reprojectDirectionValues")
            return
        assert angles.dtype == np.float32 and angles.flags.c_contiguous
        size = angles.size if size is None else size
        _lib.check(lib.fimexcu_vecrot_direction_f(self._h, angles.ctypes.data, size), "reprojectDirectionValues")

    def close(self):
        if self._h:
            lib.fimexcu_vecrot_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def interpolateVectorValues(ci, cvr, uIn, vIn, badU=_NAN, badV=_NAN, fillU=_NAN, fillV=_NAN):
    """The vector branch of CDMInterpolator::getDataSlice (src/CDMInterpolator.cc:232-259) as one fused pass:
    returns (u, v) interpolated with `ci` and rotated with `cvr` (None: no rotation)."""
    rh = cvr._h if cvr is not None else None
    if _is_torch(uIn):
        import torch
        nz = uIn.numel() // (ci.inX * ci.inY)
        uo = torch.empty((nz, ci.outY, ci.outX), dtype=torch.float32, device=uIn.device)
        vo = torch.empty_like(uo)
        _lib.check(lib.fimexcu_cache_apply_uv_dev_f(ci._h, rh, uIn.data_ptr(), vIn.data_ptr(), nz, uo.data_ptr(), vo.data_ptr(), badU, badV,
                                                    fillU, fillV, _lib.current_stream_ptr()), "interpolateVectorValues")
        return uo, vo
    u = np.ascontiguousarray(uIn, dtype=np.float32)
    v = np.ascontiguousarray(vIn, dtype=np.float32)
    nz = u.size // (ci.inX * ci.inY)
    uo = np.empty((nz, ci.outY, ci.outX), dtype=np.float32)
    vo = np.empty_like(uo)
    newSize = C.c_size_t()
    _lib.check(lib.fimexcu_cache_apply_uv_f(ci._h, rh, u.ctypes.data, v.ctypes.data, u.size, uo.ctypes.data, vo.ctypes.data, C.byref(newSize),
                                            badU, badV, fillU, fillV), "interpolateVectorValues")
    return uo, vo


def interpolateVectorValuesPadded(ci, cvr, uPadded, vPadded, badU=_NAN, badV=_NAN, fillU=_NAN, fillV=_NAN):
    """interpolateVectorValues on device slabs in the padded layout of ci.paddedLayout() (the TMA-staged fused kernel);
    returns (u, v) as [nz][outY][out_row_pitch] tensors."""
    import torch
    ip, il, op, ol = ci.paddedLayout()
    nz = uPadded.shape[0]
    uo = torch.empty((nz, ci.outY, op), dtype=torch.float32, device=uPadded.device)
    vo = torch.empty_like(uo)
    rh = cvr._h if cvr is not None else None
    _lib.check(lib.fimexcu_cache_apply_uv_dev_strided_f(ci._h, rh, uPadded.data_ptr(), vPadded.data_ptr(), ip, il, nz, uo.data_ptr(), vo.data_ptr(),
                                                        op, ol, badU, badV, fillU, fillV, _lib.current_stream_ptr()), "interpolateVectorValues")
    return uo, vo


def mifi_points2position(points, axis, axis_type):
    """include/fimex/interpolation.h:76-93: returns the positions (the reference works in place)."""
    ax = _f64(axis).ravel()
    if _is_torch(points):
        _lib.check(lib.fimexcu_points2position_dev(points.data_ptr(), points.numel(), ax.ctypes.data, ax.size, axis_type,
                                                   _lib.current_stream_ptr()), "mifi_points2position")
        return points
    p = np.array(points, dtype=np.float64).ravel().copy()
    _lib.check(lib.fimexcu_points2position(p.ctypes.data, p.size, ax.ctypes.data, ax.size, axis_type), "mifi_points2position")
    return p.reshape(np.shape(points))


def mifi_get_values(method, infield, x, y):
    """mifi_get_values_{f,bilinear_f,bicubic_f} (include/fimex/interpolation.h:95-137): values of all levels at (x, y)."""
    a = np.ascontiguousarray(infield, dtype=np.float32)
    iz, iy, ix = a.shape
    out = np.empty(iz, dtype=np.float32)
    fn = {MIFI_INTERPOL_NEAREST_NEIGHBOR: lib.fimexcu_get_values_f, MIFI_INTERPOL_BILINEAR: lib.fimexcu_get_values_bilinear_f,
          MIFI_INTERPOL_BICUBIC: lib.fimexcu_get_values_bicubic_f}[method]
    _lib.check(fn(a.ctypes.data, out.ctypes.data, float(x), float(y), ix, iy, iz), "mifi_get_values")
    return out


def mifi_fill2d_f(field, relaxCrit=0.01, corrEff=1.6, maxLoop=100):
    """mifi_fill2d_f / InterpolatorFill2d (include/fimex/interpolation.h:253, src/CDMInterpolator.cc:76-80): undefined cells of every
    [ny][nx] level of `field` ([..., ny, nx]) are filled by Laplace relaxation.  numpy array: returns (filled copy, nChanged per level);
    CUDA float32 tensor: filled IN PLACE on the device, returns (field, nChanged per level)."""
    if _is_torch(field):
        import torch
        if not field.is_cuda or field.dtype != torch.float32 or not field.is_contiguous():
            raise FimexCudaError("device path needs a contiguous float32 CUDA tensor")
        ny, nx = field.shape[-2:]
        nz = field.numel() // (nx * ny) if nx * ny else 0
        cnt = (C.c_size_t * max(nz, 1))()
        _lib.check(lib.fimexcu_fill2d_dev_f(nx, ny, nz, field.data_ptr(), relaxCrit, corrEff, maxLoop, cnt, _lib.current_stream_ptr()), "mifi_fill2d_f")
        return field, list(cnt)[:nz]
    a = np.array(field, dtype=np.float32, order="C", copy=True)
    ny, nx = a.shape[-2:]
    lev = a.reshape(-1, ny, nx)
    counts = []
    for k in range(lev.shape[0]):
        n = C.c_size_t()
        _lib.check(lib.fimexcu_fill2d_f(nx, ny, lev[k].ctypes.data, relaxCrit, corrEff, maxLoop, C.byref(n)), "mifi_fill2d_f")
        counts.append(n.value)
    return a, counts


def mifi_creepfill2d_f(field, repeat=20, setWeight=2, defaultValue=None):
    """mifi_creepfill2d_f (defaultValue None: first guess = the level's mean) / mifi_creepfillval2d_f
    (include/fimex/interpolation.h:275, 298; InterpolatorCreepFill2d / InterpolatorCreepFillVal2d, src/CDMInterpolator.cc:82-92) on every
    [ny][nx] level of `field`.  numpy array: returns (filled copy, nChanged per level); CUDA float32 tensor: filled in place."""
    use_mean = 1 if defaultValue is None else 0
    dv = 0.0 if defaultValue is None else float(defaultValue)
    w = bytes([setWeight & 0xff])
    if _is_torch(field):
        import torch
        if not field.is_cuda or field.dtype != torch.float32 or not field.is_contiguous():
            raise FimexCudaError("device path needs a contiguous float32 CUDA tensor")
        ny, nx = field.shape[-2:]
        nz = field.numel() // (nx * ny) if nx * ny else 0
        cnt = (C.c_size_t * max(nz, 1))()
        _lib.check(lib.fimexcu_creepfill2d_dev_f(nx, ny, nz, field.data_ptr(), use_mean, dv, repeat, w, cnt, _lib.current_stream_ptr()),
                   "mifi_creepfill2d_f")
        return field, list(cnt)[:nz]
    a = np.array(field, dtype=np.float32, order="C", copy=True)
    ny, nx = a.shape[-2:]
    lev = a.reshape(-1, ny, nx)
    counts = []
    for k in range(lev.shape[0]):
        n = C.c_size_t()
        if use_mean:
            _lib.check(lib.fimexcu_creepfill2d_f(nx, ny, lev[k].ctypes.data, repeat, w, C.byref(n)), "mifi_creepfill2d_f")
        else:
            _lib.check(lib.fimexcu_creepfillval2d_f(nx, ny, lev[k].ctypes.data, dv, repeat, w, C.byref(n)), "mifi_creepfillval2d_f")
        counts.append(n.value)
    return a, counts
