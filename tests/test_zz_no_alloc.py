"""GPU test of the allocation-free axis look-up, fimexcu_points2position_dev_noalloc.

Some embedders of the regridding library cannot let it allocate device memory while it runs: the surrounding
application owns the whole GPU through its own pool, or the calling thread must not hit the device-wide stall an
allocation can cause.  For them the library keeps the device copy of the axis in a fixed 128 KiB array inside its own
image instead of a freshly allocated buffer.  This test drives that entry point through the C ABI, twice in a row with
different axes (the second one shorter, so left-overs of the first call would show), and requires results identical to
the allocating entry point fimexcu_points2position_dev (bit-exact, the bar of the other parity tests).  It lives in
its own file, collected last, and passes both in a fresh process and after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_AXIS = 16384  # FIMEXCU_NOALLOC_MAX_AXIS in include/fimexcu.h


def _same(a, b):
    na, nb = np.isnan(a), np.isnan(b)
    return np.array_equal(na, nb) and np.array_equal(a[~na], b[~nb])


def _run(lib, fn, pts, axis, axis_type):
    """points -> device buffer of the library's own allocator -> entry point under test -> host"""
    from fimex_b200 import _lib
    n = pts.size
    d = lib.fimexcu_dev_alloc(n * 8)
    assert d
    try:
        _lib.check(lib.fimexcu_memcpy_h2d(d, pts.ctypes.data, n * 8), "h2d")
        rc = fn(d, n, axis.ctypes.data, axis.size, axis_type, None)
        out = np.empty_like(pts)
        _lib.check(lib.fimexcu_memcpy_d2h(out.ctypes.data, d, n * 8), "d2h")
    finally:
        lib.fimexcu_dev_free(d)
    return rc, out


def test_noalloc_matches_allocating_entry_point_twice_in_a_row():
    import fimex_b200
    lib = fimex_b200.lib
    assert lib.fimexcu_device_count() >= 1
    rng = np.random.default_rng(2026)
    cases = [
        (np.arange(0., 360., 0.05), 1),                   # 7200 values, circular longitude axis
        (np.linspace(75.25, 49.75, 1021), 2),             # shorter, descending latitude axis
    ]
    for axis, axis_type in cases:
        axis = np.ascontiguousarray(axis, dtype=np.float64)
        lo, hi = axis.min(), axis.max()
        pts = rng.uniform(lo - 0.2 * (hi - lo), hi + 0.2 * (hi - lo), 200000)
        pts[:500] = rng.choice(axis, 500)
        pts[500:504] = [np.nan, np.inf, -np.inf, 0.0]
        rc_ref, want = _run(lib, lib.fimexcu_points2position_dev, pts, axis, axis_type)
        rc_new, got = _run(lib, lib.fimexcu_points2position_dev_noalloc, pts, axis, axis_type)
        assert rc_ref == 1 and rc_new == 1, fimex_b200._lib.last_error()
        assert _same(got, want)


def test_noalloc_rejects_axes_beyond_the_fixed_scratch():
    import fimex_b200
    lib = fimex_b200.lib
    axis = np.arange(MAX_AXIS + 1, dtype=np.float64)
    pts = np.array([1.5, 2.5])
    rc, _ = _run(lib, lib.fimexcu_points2position_dev_noalloc, pts, axis, 0)
    assert rc == -1
    assert "at most" in fimex_b200._lib.last_error()
    # the limit itself is accepted
    axis = np.arange(MAX_AXIS, dtype=np.float64)
    rc, out = _run(lib, lib.fimexcu_points2position_dev_noalloc, pts, axis, 0)
    assert rc == 1 and np.array_equal(out, pts)
