"""Host-side pieces of the host entry points, without a GPU (tests/native/hostfill_host.cc over fimex_b200/csrc/fxcu_hostfill.h and
fxcu_hostpar.h): the streaming fill / copy that write the caller's result array around the row bands, the merging of per-row spans
into bands, and the persistent worker pool."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol


@pytest.fixture(scope="module")
def hf():
    return ol.native("hostfill_host")


def test_fill_lines_every_alignment_and_length(hf):
    """ordinary stores up to the first 64-byte boundary, whole lines streamed, ordinary stores for the rest: exactly n values
    written whatever the start alignment, nothing outside"""
    buf = np.zeros(4096 + 64, np.float32)
    for start in range(0, 40):
        for n in (0, 1, 3, 15, 16, 17, 31, 64, 100, 1000, 2891):
            buf[:] = 7.0
            view = buf[start:start + n]
            hf.hostfill_fill_f(view.ctypes.data_as(ol.c_float_p), n, float("nan"))
            assert np.isnan(buf[start:start + n]).all()
            assert (buf[:start] == 7.0).all() and (buf[start + n:] == 7.0).all(), (start, n)
    sbuf = np.zeros(5000, np.int16)
    for start in range(0, 70, 3):
        for n in (0, 1, 31, 32, 33, 65, 1000, 2891):
            sbuf[:] = 5
            hf.hostfill_fill_i16(sbuf[start:].ctypes.data, n, -32767)
            assert (sbuf[start:start + n] == -32767).all() and (sbuf[:start] == 5).all() and (sbuf[start + n:] == 5).all(), (start, n)


def test_copy_lines_every_alignment_and_length(hf):
    rng = np.random.default_rng(1)
    src = rng.integers(0, 255, 20000, dtype=np.uint8)
    dst = np.zeros(20000 + 256, np.uint8)
    for doff in range(0, 70, 7):
        for soff in (0, 1, 4, 13):
            for n in (0, 1, 63, 64, 65, 127, 4 * 1900, 7601):
                dst[:] = 0xAA
                hf.hostfill_copy(dst[doff:].ctypes.data, src[soff:].ctypes.data, n)
                assert np.array_equal(dst[doff:doff + n], src[soff:soff + n])
                assert (dst[:doff] == 0xAA).all() and (dst[doff + n:] == 0xAA).all(), (doff, soff, n)


def _merge(hf, lo, hi, out_x, max_bands):
    spans = np.ascontiguousarray(np.stack([lo, hi], axis=1), dtype=np.int32)
    out = (C.c_int * (4 * 600))()
    frac = C.c_double()
    n = hf.hostfill_merge(spans.ctypes.data_as(C.POINTER(C.c_int)), len(lo), out_x, max_bands, out, C.byref(frac))
    return [tuple(out[4 * k:4 * k + 4]) for k in range(n)], frac.value


def test_row_bands_cover_every_defined_cell(hf):
    """the bands tile [0, outY) in order, every row's own span lies inside its band's span, at most max_bands of them, and the
    fraction is the area of the rectangles"""
    rng = np.random.default_rng(2)
    for out_y, out_x in ((1021, 2891), (7, 50), (1, 1), (5000, 300)):
        centre = out_x / 2 + 0.3 * out_x * np.sin(np.linspace(0, 3, out_y))
        half = 0.25 * out_x * (1 - np.linspace(-1, 1, out_y) ** 2) + 1
        lo = np.clip((centre - half).astype(int), 0, out_x)
        hi = np.clip((centre + half).astype(int), 0, out_x)
        empty = rng.random(out_y) < 0.05
        lo[empty], hi[empty] = out_x, 0
        for max_bands in (1, 3, 24):
            bands, frac = _merge(hf, lo, hi, out_x, max_bands)
            assert 1 <= len(bands) <= max(max_bands, 1)
            assert bands[0][0] == 0 and bands[-1][1] == out_y and all(a[1] == b[0] for a, b in zip(bands, bands[1:]))
            area = 0
            for y0, y1, x0, x1 in bands:
                rows = np.arange(y0, y1)
                live = rows[lo[rows] < hi[rows]]
                if live.size:
                    assert x0 <= lo[live].min() and x1 >= hi[live].max()
                    assert x0 == lo[live].min() and x1 == hi[live].max()          # and no wider than needed
                else:
                    assert x0 >= x1
                area += (y1 - y0) * max(0, x1 - x0)
            assert abs(frac - area / (out_x * out_y)) < 1e-12
        # more bands never cost more area
        fr = [_merge(hf, lo, hi, out_x, m)[1] for m in (1, 2, 4, 8, 24, 100)]
        assert all(a >= b - 1e-12 for a, b in zip(fr, fr[1:]))
    assert _merge(hf, np.zeros(0, int), np.zeros(0, int), 10, 24) == ([], 1.0)


def test_worker_pool_covers_each_index_once(hf):
    """every index of a range is handed to exactly one task, over many rounds (the workers persist), for grains and part limits
    of every kind, and from inside a task (a task that fans out again runs its share alone: workers never wait for workers)"""
    for n, grain, parts in ((0, 1, 4), (1, 1, 4), (1000, 1, 4), (1000, 7, 64), (1 << 16, 1 << 10, 8), (12345, 100, 1), (999, 1000, 16)):
        assert hf.hostpar_cover(n, grain, parts, 20, 0) == 0, (n, grain, parts)
    assert hf.hostpar_cover(5000, 10, 8, 10, 1) == 0
