"""The C++ reference-side glue (fimex_b200/host/CudaCachedInterpolation.h: CachedInterpolationInterface subclasses and
CachedVectorReprojection over the C ABI) driven the way CDMInterpolator::getDataSlice drives the reference classes."""
import os
import subprocess

import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import dp, fp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "fimex_b200", "host", "host_selftest")


def test_glue_compiles_against_the_reference_headers():
    inc = "/root/reference/include"
    if not os.path.isdir(inc):
        pytest.skip("reference headers not present")
    r = subprocess.run(["g++", "-std=c++20", "-fsyntax-only", "-DFIMEXCU_WITH_FIMEX_HEADERS", "-I" + inc,
                        os.path.join(ROOT, "fimex_b200", "host", "CudaCachedInterpolation.h")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("method", [ol.BILINEAR, ol.BICUBIC, ol.NEAREST])
def test_cpp_glue_matches_oracle(tmp_path, method):
    assert os.path.exists(EXE), "fimex_b200/host/host_selftest not built (python -m fimex_b200.build)"
    rng = np.random.default_rng(method)
    in_x, in_y, out_x, out_y, nz = 57, 43, 83, 61, 4
    # tables that reach all four borders, so the reduced window is the whole grid (offset 0)
    px = rng.uniform(-2, in_x + 1, out_x * out_y)
    py = rng.uniform(-2, in_y + 1, out_x * out_y)
    px[:4], py[:4] = [0.2, in_x - 1.2, 0.2, in_x - 1.2], [0.2, 0.2, in_y - 1.2, in_y - 1.2]
    strip = (py >= in_y - 0.5) & (py < in_y + 0.5) & ~((np.floor(px) >= 0) & (np.floor(px) + 1 < in_x))
    py[strip] += 1.0
    u = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    v = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    u[rng.random(u.shape) < 0.03] = np.nan
    ang = rng.uniform(-3, 3, out_x * out_y)
    m3 = np.stack([np.cos(ang), np.sin(ang), np.degrees(ang)], axis=1).copy()
    case, res = tmp_path / "case.bin", tmp_path / "res.bin"
    with open(case, "wb") as f:
        f.write(np.array([method, in_x, in_y, out_x, out_y, nz], dtype=np.int32).tobytes())
        for a in (px, py, u, v, m3):
            f.write(np.ascontiguousarray(a).tobytes())
    r = subprocess.run([EXE, str(case), str(res)], capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    got = np.fromfile(res, dtype=np.float32).reshape(2, nz, out_y, out_x)
    wu, info = ol.o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, u)
    wv, _ = ol.o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, v)
    assert (info["inX"], info["inY"]) == (in_x, in_y)
    ol.oracle().orc_vector_reproject_values_by_matrix_f(dp(m3), out_x * out_y, fp(wu), fp(wv), nz)
    for g, w in ((got[0], wu), (got[1], wv)):
        assert np.array_equal(np.isnan(g), np.isnan(w)) and np.array_equal(g[~np.isnan(g)], w[~np.isnan(w)])


@pytest.mark.gpu
@pytest.mark.parametrize("method", [ol.BILINEAR, ol.NEAREST])
def test_cpp_glue_one_cache_many_writer_threads(tmp_path, method):
    """The reference's process model (src/io/netcdf/NetCDF_CDMWriter.cc:599-618): ONE cache object, several host threads
    calling interpolateValues concurrently with different (pageable) slabs.  Each thread is bound to a GPU round robin over all
    devices of the box (tables replicated by fimexcu_cache_replicate); with a single GPU the threads share it.  Every slab must
    equal the oracle, whichever device regridded it."""
    import fimex_b200
    rng = np.random.default_rng(40 + method)
    in_x, in_y, out_x, out_y, nz, nslabs, nthreads = 120, 90, 331, 207, 6, 11, 5
    px = rng.uniform(-2, in_x + 1, out_x * out_y)
    py = rng.uniform(-2, in_y + 1, out_x * out_y)
    px[:4], py[:4] = [0.2, in_x - 1.2, 0.2, in_x - 1.2], [0.2, 0.2, in_y - 1.2, in_y - 1.2]
    strip = (py >= in_y - 0.5) & (py < in_y + 0.5) & ~((np.floor(px) >= 0) & (np.floor(px) + 1 < in_x))
    py[strip] += 1.0
    slabs = rng.normal(size=(nslabs, nz, in_y, in_x)).astype(np.float32)
    slabs[rng.random(slabs.shape) < 0.02] = np.nan
    case, res = tmp_path / "case.bin", tmp_path / "res.bin"
    with open(case, "wb") as f:
        f.write(np.array([method, in_x, in_y, out_x, out_y, nz, nslabs], dtype=np.int32).tobytes())
        for a in (px, py, slabs):
            f.write(np.ascontiguousarray(a).tobytes())
    r = subprocess.run([EXE, "threads", str(nthreads), str(case), str(res)], capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    raw = np.fromfile(res, dtype=np.float32)
    n = nslabs * nz * out_y * out_x
    got = raw[:n].reshape(nslabs, nz, out_y, out_x)
    dev = raw[n:].view(np.int32)
    for k in range(nslabs):
        w, _ = ol.o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, slabs[k])
        assert np.array_equal(np.isnan(got[k]), np.isnan(w)) and np.array_equal(got[k][~np.isnan(w)], w[~np.isnan(w)]), k
    ndev = fimex_b200.lib.fimexcu_device_count()
    assert len(set(dev.tolist())) == min(ndev, nthreads)  # every GPU of the box (up to the thread count) took part


@pytest.mark.gpu
@pytest.mark.parametrize("vmethod", [ol.VINT_LIN, ol.VINT_LOG])
def test_cpp_device_slab_chain(tmp_path, vmethod):
    """fimex_b200/host/CudaSlab.h: BASELINE configs[4] as a device-resident reader chain -- packed short field up, unpack,
    bilinear S->T (staged kernel, padded layout), ps the same way, hybrid -> pressure levels with the pressure from the REGRIDDED
    ps, fused wind rotation, results down -- against the oracle chain, every value."""
    import os as _os
    rng = np.random.default_rng(50 + vmethod)
    g = np.load(_os.path.join(ROOT, "tests", "golden", "reference_vertical.npz"))
    ap, b = g["en_ap_pa"] * 0.01, g["en_b"]
    nlev, nmem, in_x, in_y, out_x, out_y = 65, 2, 300, 200, 1400, 900   # 1232 target tiles: the staged kernels run
    gx, gy = np.meshgrid(np.arange(out_x), np.arange(out_y))
    px = -3.0 + gx * ((in_x + 5.0) / out_x) + 0.11 * gy * (in_x / out_x)
    py = -2.5 + gy * ((in_y + 4.0) / out_y) - 0.07 * gx * (in_y / out_y)
    fill = -32767.0
    raw = rng.integers(-3000, 3000, (nlev * nmem, in_y, in_x)).astype(np.int16)
    raw[rng.random(raw.shape) < 0.01] = int(fill)
    ps = (101325 - 12 * rng.uniform(0, 2500, (nmem, in_y, in_x)) + rng.normal(0, 300, (nmem, in_y, in_x))).astype(np.float32)
    levels = np.array([1000, 925, 850, 700, 500, 400, 300, 250, 200, 150, 100, 50], dtype=np.float64)
    ang = rng.uniform(-3, 3, out_x * out_y)
    m3 = np.stack([np.cos(ang), np.sin(ang), np.degrees(ang)], axis=1).copy()
    case, res = tmp_path / "case.bin", tmp_path / "res.bin"
    with open(case, "wb") as f:
        f.write(np.array([vmethod, in_x, in_y, out_x, out_y, nlev, nmem, levels.size], dtype=np.int32).tobytes())
        f.write(px.tobytes()); f.write(py.tobytes()); f.write(raw.tobytes()); f.write(np.array([fill]).tobytes())
        f.write(ps.tobytes()); f.write(ap.astype(np.float64).tobytes()); f.write(b.astype(np.float64).tobytes()); f.write(levels.tobytes())
        f.write(m3.tobytes())
    r = subprocess.run([EXE, "chain", str(case), str(res)], capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    out = np.fromfile(res, dtype=np.float32)
    ncol = nmem * out_y * out_x
    got_v = out[:levels.size * ncol].reshape(levels.size, ncol)
    got_u = out[levels.size * ncol:levels.size * ncol + ncol].reshape(nmem, out_y, out_x)
    got_w = out[levels.size * ncol + ncol:].reshape(nmem, out_y, out_x)
    # oracle chain: unpack (asFloat + bad2nan), bilinear, pressure from the regridded ps, vertical
    o = ol.oracle()
    fld = np.empty(raw.shape, np.float32)
    o.orc_i16_to_interpolation_array(raw.ctypes.data, raw.size, fill, fp(fld))
    h_f, _ = ol.o_cached_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, fld)
    h_ps, _ = ol.o_cached_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, ps)
    p = np.ascontiguousarray((ap[:, None] + b[:, None] * (h_ps.reshape(1, ncol).astype(np.float64) * 0.01)).astype(np.float32))
    want = np.empty((levels.size, ncol), np.float32)
    hf2 = np.ascontiguousarray(h_f.reshape(nlev, ncol))
    assert o.orc_vertical_interpolate_f(vmethod, fp(hf2), fp(p), nlev, ncol, dp(levels), None, levels.size, None, None, float("nan"), float("nan"),
                                        fp(want)) == 1
    assert np.array_equal(np.isnan(got_v), np.isnan(want))
    k = ~np.isnan(want)
    if vmethod == ol.VINT_LIN:
        assert np.array_equal(got_v[k], want[k])
    else:
        assert (np.abs(got_v[k] - want[k]) <= 1e-5 * np.abs(want[k])).all() and (got_v[k] == want[k]).mean() > 0.999
    wu, wv = h_ps.copy(), np.ascontiguousarray(h_f[:nmem])
    o.orc_vector_reproject_values_by_matrix_f(dp(m3), out_x * out_y, fp(wu), fp(wv), nmem)
    for gch, w in ((got_u, wu), (got_w, wv)):
        assert np.array_equal(np.isnan(gch), np.isnan(w)) and np.array_equal(gch[~np.isnan(w)], w[~np.isnan(w)])


def test_slab_glue_compiles_against_the_reference_headers():
    inc = "/root/reference/include"
    if not os.path.isdir(inc):
        pytest.skip("reference headers not present")
    r = subprocess.run(["g++", "-std=c++20", "-fsyntax-only", "-DFIMEXCU_WITH_FIMEX_HEADERS", "-I" + inc, "-x", "c++",
                        os.path.join(ROOT, "fimex_b200", "host", "CudaSlab.h")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_process_glue_compiles_against_the_reference_headers():
    inc = "/root/reference/include"
    if not os.path.isdir(inc):
        pytest.skip("reference headers not present")
    r = subprocess.run(["g++", "-std=c++20", "-fsyntax-only", "-DFIMEXCU_WITH_FIMEX_HEADERS", "-I" + inc, "-I" + os.path.join(ROOT, "include"),
                        "-x", "c++", os.path.join(ROOT, "fimex_b200", "host", "CudaInterpolatorProcess2d.h")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("kind,params", [(0, (0.01, 1.6, 100)), (1, (20, 2, 0)), (2, (5, 3, 271.5))])
def test_cpp_fill_functors_match_oracle(tmp_path, kind, params):
    """InterpolatorFill2d / InterpolatorCreepFill2d / InterpolatorCreepFillVal2d over the C ABI, level by level (operator()) and
    as one slab (processArray_), against the oracle restatements"""
    import ctypes as C
    assert os.path.exists(EXE)
    rng = np.random.default_rng(kind)
    nz, ny, nx = 3, 41, 67
    yy, xx = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
    fld = (280 + 10 * np.sin(xx / 9.0) * np.cos(yy / 7.0))[None].repeat(nz, 0).astype(np.float32) + rng.normal(0, 1, (nz, ny, nx)).astype(np.float32)
    for k in range(nz):
        for _ in range(6):
            cy, cx, r = rng.integers(0, ny), rng.integers(0, nx), rng.integers(2, 7)
            fld[k][(yy - cy) ** 2 + (xx - cx) ** 2 <= r * r] = np.nan
    want = fld.copy()
    o = ol.oracle()
    for k in range(nz):
        n = C.c_size_t()
        if kind == 0:
            o.orc_fill2d_f(nx, ny, fp(want[k]), params[0], params[1], int(params[2]), C.byref(n))
        elif kind == 1:
            o.orc_creepfill2d_f(nx, ny, fp(want[k]), int(params[0]), bytes([int(params[1])]), C.byref(n))
        else:
            o.orc_creepfillval2d_f(nx, ny, fp(want[k]), params[2], int(params[0]), bytes([int(params[1])]), C.byref(n))
    case, res = tmp_path / "fill.bin", tmp_path / "fill_out.bin"
    with open(case, "wb") as f:
        np.array([kind, nx, ny, nz], dtype=np.int32).tofile(f)
        np.array(params, dtype=np.float32).tofile(f)
        fld.tofile(f)
    r = subprocess.run([EXE, "fill", str(case), str(res)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    out = np.fromfile(res, dtype=np.float32).reshape(2, nz, ny, nx)
    assert np.array_equal(out[0], want) and np.array_equal(out[1], want)
