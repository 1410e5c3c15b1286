"""GPU parity of mifi_fill2d_f through the C ABI against the oracle restatement (pinned bit for bit against
the reference's own mifi_fill2d_f in test_oracle_vs_ref.py).  Bar: bit-exact, including the number of sweeps taken."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import fp

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fx():
    import fimex_b200
    assert fimex_b200.lib.fimexcu_device_count() >= 1
    from fimex_b200 import interpolation
    return interpolation


def _blob_field(rng, ny, nx, frac=0.15):
    yy, xx = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
    f = (280 + 15 * np.sin(xx / 9.0) * np.cos(yy / 7.0) + rng.normal(0, 1, (ny, nx))).astype(np.float32)
    for _ in range(max(1, int(frac * nx * ny / 60))):
        cy, cx, r = rng.integers(0, ny), rng.integers(0, nx), rng.integers(2, 7)
        f[(yy - cy) ** 2 + (xx - cx) ** 2 <= r * r] = np.nan
    return f


def _oracle_fill(f, relax, corr, maxloop):
    a = f.copy()
    n = C.c_size_t()
    assert ol.oracle().orc_fill2d_f(f.shape[1], f.shape[0], fp(a), relax, corr, maxloop, C.byref(n)) == 1
    return a, n.value


@pytest.mark.parametrize("nx,ny,maxloop,relax", [(61, 47, 100, 0.01), (33, 80, 7, 0.01), (20, 20, 3, 0.01), (90, 12, 31, 0.01), (3, 3, 100, 0.01),
                                                  (128, 97, 100, 4.0), (75, 64, 26, 1e-6), (40, 40, 1, 0.01), (1100, 37, 100, 0.05)])
def test_fill2d_vs_oracle(fx, nx, ny, maxloop, relax):
    rng = np.random.default_rng(nx * ny + maxloop)
    f = _blob_field(rng, ny, nx)
    f[0, :2] = np.nan
    f[-1, -1] = np.nan
    f[ny // 2, 0] = np.nan
    if nx * ny == 9:
        f = rng.normal(280, 5, (3, 3)).astype(np.float32)
        f[1, 1] = f[0, 2] = np.nan
    want, nwant = _oracle_fill(f, relax, 1.6, maxloop)
    got, counts = fx.mifi_fill2d_f(f, relax, 1.6, maxloop)
    assert counts == [nwant] and np.array_equal(got, want, equal_nan=True) and not np.isnan(want).any()


def test_fill2d_many_levels_on_device_and_trivial_levels(fx):
    """processArray_ runs the process on every level; levels without undefined cells or without defined cells stay as they are"""
    import torch
    rng = np.random.default_rng(4)
    nz, ny, nx = 7, 59, 83
    fld = np.stack([_blob_field(rng, ny, nx, frac=0.05 + 0.05 * k) for k in range(nz)])
    fld[2] = np.nan
    fld[3] = np.float32(5.0)
    fld[4, 10:50, 20:70] = np.nan  # a large hole: many sweeps
    want = np.empty_like(fld)
    nwant = []
    for k in range(nz):
        want[k], n = _oracle_fill(fld[k], 0.01, 1.6, 100)
        nwant.append(n)
    d = torch.from_numpy(fld).cuda()
    out, counts = fx.mifi_fill2d_f(d, 0.01, 1.6, 100)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    assert counts == nwant
    assert np.array_equal(got, want, equal_nan=True)
    assert np.isnan(got[2]).all() and not np.isnan(got[[0, 1, 3, 4, 5, 6]]).any()


def test_fill2d_rejects_degenerate_grids(fx):
    with pytest.raises(Exception):
        fx.mifi_fill2d_f(np.full((2, 40), np.nan, np.float32))


def _oracle_creep(f, repeat, weight, default=None):
    a = f.copy()
    n = C.c_size_t()
    w = bytes([weight])
    if default is None:
        assert ol.oracle().orc_creepfill2d_f(f.shape[1], f.shape[0], fp(a), repeat, w, C.byref(n)) == 1
    else:
        assert ol.oracle().orc_creepfillval2d_f(f.shape[1], f.shape[0], fp(a), default, repeat, w, C.byref(n)) == 1
    return a, n.value


@pytest.mark.parametrize("nx,ny,repeat,weight", [(61, 47, 20, 2), (33, 80, 1, 1), (20, 20, 5, 2), (90, 12, 20, 4), (3, 3, 3, 2), (128, 97, 20, 2),
                                                  (75, 64, 0, 2), (1100, 37, 7, 3)])
@pytest.mark.parametrize("default", [None, 271.5])
def test_creepfill_vs_oracle(fx, nx, ny, repeat, weight, default):
    rng = np.random.default_rng(nx * ny + repeat)
    f = _blob_field(rng, ny, nx, frac=0.3)
    f[0, :5] = np.nan
    f[-1, -1] = np.nan
    f[ny // 2, 0] = np.nan
    f[1, 0] = np.nan
    f[-2, -1] = np.nan
    if nx * ny == 9:
        f = rng.normal(280, 5, (3, 3)).astype(np.float32)
        f[1, 1] = f[0, 0] = f[1, 0] = np.nan
    want, nwant = _oracle_creep(f, repeat, weight, default)
    got, counts = fx.mifi_creepfill2d_f(f, repeat, weight, default)
    assert counts == [nwant] and np.array_equal(got, want, equal_nan=True)


def test_creepfill_many_levels_on_device(fx):
    import torch
    rng = np.random.default_rng(14)
    nz, ny, nx = 6, 59, 83
    fld = np.stack([_blob_field(rng, ny, nx, frac=0.1 + 0.08 * k) for k in range(nz)])
    fld[2] = np.nan
    fld[3] = np.float32(5.0)
    fld[4, 10:50, 20:70] = np.nan
    fld[5, :, :30] = np.nan  # everything west of a line: the fill creeps a long way, border cells included
    want = np.empty_like(fld)
    nwant = []
    for k in range(nz):
        want[k], n = _oracle_creep(fld[k], 20, 2)
        nwant.append(n)
    out, counts = fx.mifi_creepfill2d_f(torch.from_numpy(fld).cuda(), 20, 2)
    torch.cuda.synchronize()
    assert counts == nwant and np.array_equal(out.cpu().numpy(), want, equal_nan=True)
