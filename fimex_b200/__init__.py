"""fimex_b200 -- B200-native regridding hot path of metno/fimex.

The product is `libfimexcu.so` (hand-written CUDA for sm_100a behind the C ABI of include/fimexcu.h).
This package is the thin Python host side used by the tests and bench.py: it mirrors the reference's
operator interface (CachedInterpolation, CachedNNInterpolation, CachedForwardInterpolation,
CachedVectorReprojection, the mifi_* functions) and forwards every call to the C ABI.
There is no CPU or PyTorch fallback: importing works without a GPU (so the library's symbols can be
checked), any compute call without a usable sm_100 device raises FimexCudaError.
"""
from ._lib import FimexCudaError, lib, lib_path  # noqa: F401
from .interpolation import (  # noqa: F401
    CachedForwardInterpolation,
    CachedInterpolation,
    CachedNNInterpolation,
    CachedVectorReprojection,
    createCachedInterpolation,
    mifi_points2position,
)
from . import constants  # noqa: F401

__all__ = [
    "FimexCudaError", "lib", "lib_path", "constants",
    "CachedInterpolation", "CachedNNInterpolation", "CachedForwardInterpolation", "CachedVectorReprojection",
    "createCachedInterpolation", "mifi_points2position",
]
