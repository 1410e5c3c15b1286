"""One cache handle, several GPUs, driven from host threads of one process -- the reference's own process model
(OpenMP writer threads sharing one CachedInterpolation, src/io/netcdf/NetCDF_CDMWriter.cc:599-618).
fimexcu_cache_replicate copies the tables built on device 0 to the other devices; every apply entry point then uses the copy on
the calling thread's current device.  Needs >= 2 GPUs (`gpurun --gpus 2`); the single-GPU parts still run on one."""
import ctypes as C
import threading

import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu


def _eq(a, b):
    na, nb = np.isnan(a), np.isnan(b)
    return a.shape == b.shape and np.array_equal(na, nb) and np.array_equal(a[~na], b[~nb])


@pytest.fixture(scope="module")
def fx():
    import fimex_b200
    assert fimex_b200.lib.fimexcu_device_count() >= 1
    return fimex_b200


def _case(rng, method):
    in_x, in_y, out_x, out_y, nz = 300, 200, 1400, 900, 6  # enough tiles for the staged kernel
    gx, gy = np.meshgrid(np.arange(out_x), np.arange(out_y))
    px = -3.0 + gx * ((in_x + 5.0) / out_x) + 0.11 * gy * (in_x / out_x)
    py = -2.5 + gy * ((in_y + 4.0) / out_y) - 0.07 * gx * (in_y / out_y)
    return in_x, in_y, out_x, out_y, nz, px, py


def test_wrong_device_without_replica_is_a_clean_error(fx):
    """a handle used while another device is current must fail with a message, not launch on foreign tables"""
    import torch
    if fx.lib.fimexcu_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    rng = np.random.default_rng(1)
    in_x, in_y, out_x, out_y, nz, px, py = _case(rng, ol.BILINEAR)
    fx._lib.check(fx.lib.fimexcu_set_device(0), "set_device")
    ci = fx.CachedInterpolation("x", "y", ol.BILINEAR, px, py, in_x, in_y, out_x, out_y)
    fld = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    sub = ci.reduceInput(fld, in_x, in_y)
    try:
        fx._lib.check(fx.lib.fimexcu_set_device(1), "set_device")
        with pytest.raises(fx.FimexCudaError, match="replicate"):
            ci.interpolateValues(sub)
        ci.replicate([1])
        got, _ = ci.interpolateValues(sub)   # now served by the copy on device 1
    finally:
        fx._lib.check(fx.lib.fimexcu_set_device(0), "set_device")
        torch.cuda.set_device(0)
    want, _ = ol.o_cached_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, fld)
    assert _eq(got.reshape(want.shape), want)


@pytest.mark.parametrize("method", [ol.BILINEAR, ol.NEAREST, ol.BICUBIC])
def test_one_handle_all_devices_from_host_threads(fx, method):
    import torch
    ndev = fx.lib.fimexcu_device_count()
    if ndev < 2:
        pytest.skip("needs 2 GPUs")
    rng = np.random.default_rng(10 + method)
    in_x, in_y, out_x, out_y, nz, px, py = _case(rng, method)
    fx._lib.check(fx.lib.fimexcu_set_device(0), "set_device")
    ci = fx.createCachedInterpolation("x", "y", method, px, py, in_x, in_y, out_x, out_y)
    ci.replicate(list(range(ndev)))
    ang = rng.uniform(-3, 3, out_x * out_y)
    m3 = np.stack([np.cos(ang), np.sin(ang), np.degrees(ang)], axis=1).copy()
    cvr = fx.CachedVectorReprojection(m3, out_x, out_y)
    cvr.replicate(list(range(ndev)))
    nslabs = 3 * ndev
    slabs = rng.normal(size=(nslabs, nz, in_y, in_x)).astype(np.float32)
    slabs[rng.random(slabs.shape) < 0.02] = np.nan
    host_out, dev_out, uv_out, errors = [None] * nslabs, [None] * nslabs, [None] * nslabs, []
    ip, il, op, olv = ci.paddedLayout()
    nan = float("nan")

    def worker(d):
        try:
            fx._lib.check(fx.lib.fimexcu_set_device(d), "set_device")  # binds THIS thread
            with torch.cuda.device(d):
                for k in range(d, nslabs, ndev):                        # slabs dealt t % N, as the reference's MPI writer does
                    sub = ci.reduceInput(slabs[k], in_x, in_y)
                    host_out[k], _ = ci.interpolateValues(sub)          # host-pointer ABI (pageable numpy memory)
                    pin = torch.zeros((nz, ci.getInY(), ip), dtype=torch.float32, device=f"cuda:{d}")
                    pin[:, :, :ci.getInX()] = torch.from_numpy(sub).to(f"cuda:{d}")
                    out = torch.empty((nz, out_y, op), dtype=torch.float32, device=f"cuda:{d}")
                    fx._lib.check(fx.lib.fimexcu_cache_apply_dev_strided_f(ci._h, pin.data_ptr(), ip, il, nz, out.data_ptr(), op, olv, nan, nan, None),
                                  "apply")                              # staged kernel on device d's copy of the tables
                    u2 = torch.empty_like(out)
                    v2 = torch.empty_like(out)
                    fx._lib.check(fx.lib.fimexcu_cache_apply_uv_dev_strided_f(ci._h, cvr._h, pin.data_ptr(), pin.data_ptr(), ip, il, nz, u2.data_ptr(),
                                                                              v2.data_ptr(), op, olv, nan, nan, nan, nan, None), "uv")
                    torch.cuda.synchronize(d)
                    dev_out[k] = out[:, :, :out_x].cpu().numpy()
                    uv_out[k] = (u2[:, :, :out_x].cpu().numpy(), v2[:, :, :out_x].cpu().numpy())
        except Exception as exc:  # noqa: BLE001
            errors.append((d, repr(exc)))

    threads = [threading.Thread(target=worker, args=(d,)) for d in range(ndev)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    fx._lib.check(fx.lib.fimexcu_set_device(0), "set_device")
    assert not errors, errors
    o = ol.oracle()
    for k in range(nslabs):
        want, _ = ol.o_cached_apply(method, px, py, in_x, in_y, out_x, out_y, slabs[k])
        assert _eq(host_out[k].reshape(want.shape), want), k
        assert _eq(dev_out[k], want), k
        wu, wv = want.copy(), want.copy()
        o.orc_vector_reproject_values_by_matrix_f(ol.dp(m3), out_x * out_y, ol.fp(wu), ol.fp(wv), nz)
        assert _eq(uv_out[k][0], wu) and _eq(uv_out[k][1], wv), k


def test_pinned_pool_reuses_blocks(fx):
    """fimexcu_host_alloc / fimexcu_host_free: a freed block of the same size class comes back without a new pinning"""
    lib = fx.lib
    a = lib.fimexcu_host_alloc(50 << 20)
    assert a
    lib.fimexcu_host_free(a)
    b = lib.fimexcu_host_alloc((50 << 20) - 12345)  # same 2 MiB class
    assert b == a
    c = lib.fimexcu_host_alloc(50 << 20)            # a second live block of the class is a different one
    assert c and c != b
    lib.fimexcu_host_free(b)
    lib.fimexcu_host_free(c)
    lib.fimexcu_host_free(c)                        # double free of a parked block is ignored
    lib.fimexcu_host_pool_trim()
    d = lib.fimexcu_host_alloc(1000)
    buf = (C.c_char * 1000).from_address(d)
    buf[0] = b"x"
    lib.fimexcu_host_free(d)


def test_pageable_and_pinned_host_slabs_give_the_same_result(fx):
    """the host entry point with pageable memory on either side (bounce buffers) and with pooled pinned memory"""
    rng = np.random.default_rng(5)
    in_x, in_y, out_x, out_y, nz, px, py = _case(rng, ol.BILINEAR)
    nz = 40  # several level chunks (traffic > 16 MB per chunk boundary)
    ci = fx.CachedInterpolation("x", "y", ol.BILINEAR, px, py, in_x, in_y, out_x, out_y)
    fld = rng.normal(size=(nz, in_y, in_x)).astype(np.float32)
    fld[rng.random(fld.shape) < 0.02] = np.nan
    sub = ci.reduceInput(fld, in_x, in_y)
    want, _ = ol.o_cached_apply(ol.BILINEAR, px, py, in_x, in_y, out_x, out_y, fld)
    got, _ = ci.interpolateValues(sub)  # numpy arrays: pageable in, pageable out
    assert _eq(got.reshape(want.shape), want)
    lib = fx.lib
    n_in, n_out = sub.size, want.size
    h_in, h_out = lib.fimexcu_host_alloc(n_in * 4), lib.fimexcu_host_alloc(n_out * 4)
    assert h_in and h_out
    C.memmove(h_in, sub.ctypes.data, n_in * 4)
    new = C.c_size_t()
    nan = float("nan")
    # pinned in, pinned out
    fx._lib.check(lib.fimexcu_cache_apply_f(ci._h, h_in, n_in, h_out, C.byref(new), nan, nan), "apply")
    a = np.frombuffer((C.c_float * n_out).from_address(h_out), dtype=np.float32).reshape(want.shape)
    assert new.value == n_out and _eq(a, want)
    # pageable in, pinned out (what the C++ glue does)
    a[:] = 0
    fx._lib.check(lib.fimexcu_cache_apply_f(ci._h, sub.ctypes.data, n_in, h_out, C.byref(new), nan, nan), "apply")
    assert _eq(a, want)
    # pinned in, pageable out
    out = np.zeros(want.shape, np.float32)
    fx._lib.check(lib.fimexcu_cache_apply_f(ci._h, h_in, n_in, out.ctypes.data, C.byref(new), nan, nan), "apply")
    assert _eq(out, want)
    lib.fimexcu_host_free(h_in)
    lib.fimexcu_host_free(h_out)
