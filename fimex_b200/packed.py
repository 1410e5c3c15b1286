"""Slab feed / sink for packed integer variables: the conversions CDMInterpolator / CDMVerticalInterpolator wrap
around their kernels (src/CDMInterpolator.cc:140-151, src/CDMReader.cc:177-192) as device passes over libfimexcu.so.
Packed slabs cross PCIe in their file type; every operator of the path then runs on f32 device tensors."""
import numpy as np

from . import _lib
from ._lib import lib

DTYPE_I8, DTYPE_I16, DTYPE_I32, DTYPE_U8, DTYPE_U16, DTYPE_U32 = range(6)


def _code(t):
    import torch
    table = {torch.int8: DTYPE_I8, torch.int16: DTYPE_I16, torch.int32: DTYPE_I32, torch.uint8: DTYPE_U8}
    for name, code in (("uint16", DTYPE_U16), ("uint32", DTYPE_U32)):
        if hasattr(torch, name):
            table[getattr(torch, name)] = code
    if t.dtype not in table:
        raise _lib.FimexCudaError(f"unsupported packed type {t.dtype}")
    return table[t.dtype]


def _cuda(a):
    import torch
    t = a if type(a).__module__.startswith("torch") else torch.from_numpy(np.ascontiguousarray(a))
    return t.cuda().contiguous()


def unpack(raw, badValue):
    """data2InterpolationArray: raw packed values as f32, the fill value as NaN (CUDA tensor)"""
    import torch
    t = _cuda(raw)
    out = torch.empty(t.shape, dtype=torch.float32, device=t.device)
    _lib.check(lib.fimexcu_unpack_dev_f(_code(t), t.data_ptr(), t.numel(), float(badValue), out.data_ptr(), _lib.current_stream_ptr()), "unpack")
    return out


def unpackScaled(raw, fillValue, scale_factor=1.0, add_offset=0.0):
    """getScaledDataSlice(...)->asFloat(): fill -> NaN, scale_factor*v + add_offset in f64, narrowed to f32 (CUDA tensor)"""
    import torch
    t = _cuda(raw)
    out = torch.empty(t.shape, dtype=torch.float32, device=t.device)
    _lib.check(lib.fimexcu_unpack_scaled_dev_f(_code(t), t.data_ptr(), t.numel(), float(fillValue), float(scale_factor), float(add_offset),
                                               out.data_ptr(), _lib.current_stream_ptr()), "unpackScaled")
    return out


def pack(values, dtype, badValue):
    """interpolationArray2Data: NaN -> fill, lround and the reference's casts; `dtype` a torch integer dtype (CUDA tensor out)"""
    import torch
    v = _cuda(values)
    if v.dtype != torch.float32:
        raise _lib.FimexCudaError("pack needs float32 values")
    out = torch.empty(v.shape, dtype=dtype, device=v.device)
    _lib.check(lib.fimexcu_pack_dev_f(_code(out), v.data_ptr(), v.numel(), float(badValue), out.data_ptr(), _lib.current_stream_ptr()), "pack")
    return out
