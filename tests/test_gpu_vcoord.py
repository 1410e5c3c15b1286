"""GPU parity of the vertical coordinate converters through the C ABI, against the oracle restatement
(pinned against the reference's own C functions in test_oracle_vs_ref.py).
Bar: bit-exact for the arithmetic-only formulas; 2 f32 ulps where the device's exp / log replace glibc's."""
import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import dp, fp

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def V():
    import fimex_b200
    assert fimex_b200.lib.fimexcu_device_count() >= 1
    from fimex_b200 import vertical
    return vertical


def _fpn(a):
    return None if a is None else fp(a)


def _ulps(a, b):
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    nan = np.isnan(a) | np.isnan(b)
    assert np.array_equal(np.isnan(a), np.isnan(b))
    ia, ib = a.view(np.int32).astype(np.int64), b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7fffffff), ia)
    ib = np.where(ib < 0, -(ib & 0x7fffffff), ib)
    return int(np.abs(np.where(nan, 0, ia - ib)).max())


def _oracle_levels(kind, a, b, scalar, surf, scale, eta, post, topo, factor, nO, nz, nI):
    out = np.empty((nO, nz, nI), np.float32)
    bb = a if b is None else b
    rc = ol.oracle().orc_vcoord_levels_f(kind, dp(np.ascontiguousarray(a)), dp(np.ascontiguousarray(bb)), scalar, _fpn(surf),
                                         1 if surf is None else surf.shape[0], scale, _fpn(eta), post, _fpn(topo), factor, nO, nz, nI, fp(out))
    assert rc == 1
    return out


@pytest.mark.parametrize("post", [0, 1, 2])
@pytest.mark.parametrize("kind", range(7))
def test_level_fields_vs_oracle(V, kind, post):
    rng = np.random.default_rng(100 * kind + post)
    nz, nI, nO = 65, 1234, 2
    g = np.load(__import__("os").path.join(__import__("os").path.dirname(__file__), "golden", "reference_vertical.npz"))
    topo = rng.uniform(0, 2500, nI).astype(np.float32)
    eta = None
    if kind in (V.VCOORD_HYBRID_SIGMA_AP, V.VCOORD_HYBRID_SIGMA):
        a, b = g["en_ap_pa"] * 0.01, g["en_b"]
        scalar = 0.0
        if kind == V.VCOORD_HYBRID_SIGMA:
            a, scalar = a / 1000.0, 1000.0
        surf, scale = (101325 - 12 * rng.uniform(0, 2500, (nO, nI))).astype(np.float32), 0.01
    elif kind == V.VCOORD_SIGMA:
        a, b, scalar = np.linspace(0.01, 1, nz), None, 10.0
        surf, scale = rng.uniform(600, 1040, (1, nI)).astype(np.float32), 1.0
    elif kind == V.VCOORD_LN_PRESSURE:
        a, b, scalar, surf, scale = np.linspace(0.0, 7.0, nz), None, 1000.0, None, 1.0
    elif kind == V.VCOORD_HYBRID_HEIGHT:
        a, b, scalar = np.linspace(10, 40000, nz), np.linspace(1, 0, nz) ** 2, 0.0
        surf, scale = rng.uniform(0, 3000, (1, nI)).astype(np.float32), 1.0
    else:
        a = -np.linspace(1, 0.01, nz)
        b, scalar = -(np.linspace(1, 0.01, nz) ** 2), 10.0
        surf, scale = rng.uniform(5, 5000, (nO, nI)).astype(np.float32), 1.0
        eta = rng.normal(0, 0.4, (nO, nI)).astype(np.float32)
    nOs = nO if (surf is not None and surf.shape[0] == nO) else 1
    nOut = nO
    for tp, factor in ((None, 0.0), (topo, -1.0 / 9.80665)):
        want = _oracle_levels(kind, a, b, scalar, surf, scale, eta, post, tp, factor, nOut, nz, nI)
        got = V.verticalLevelField(kind, a, b, scalar, surf, surface_scale=scale, eta=eta, post=post, topography=tp, topography_factor=factor,
                                   nOuter=nOut, nInner=nI).cpu().numpy()
        with np.errstate(all="ignore"):
            if post == 0 and kind != V.VCOORD_LN_PRESSURE:
                assert np.array_equal(got, want, equal_nan=True), (kind, post)
            else:
                assert _ulps(got, want) <= 2, (kind, post)


@pytest.mark.parametrize("start_high", [1, 0, -1])
@pytest.mark.parametrize("humidity", [True, False])
def test_pressure_integration_vs_oracle(V, start_high, humidity):
    rng = np.random.default_rng(7 + start_high)
    nO, nz, nI = 2, 65, 777
    eta = np.linspace(0.998, 0.01, nz) if start_high != 0 else np.linspace(0.01, 0.998, nz)
    sap = rng.uniform(65000, 104000, (nO, nI)).astype(np.float32)
    p = (eta[None, :, None] * (sap[:, None, :] * np.float32(0.01))).astype(np.float32)
    if start_high == -1:
        p[0, 0, :5] = -1.0  # the reference skips columns without positive ends when it looks for the direction
    sgp = rng.uniform(0, 25000, (1, nI)).astype(np.float32)
    T = rng.uniform(205, 305, (nO, nz, nI)).astype(np.float32)
    q = rng.uniform(0, 0.02, (nO, nz, nI)).astype(np.float32) if humidity else None
    topo = (sgp[0] / np.float32(9.80665)).astype(np.float32)
    for tp, factor in ((None, 0.0), (topo, -1.0)):
        want = np.empty((nO, nz, nI), np.float32)
        # surface geopotential is broadcast over the outer dimension, surface pressure is per outer slab: the oracle takes ONE
        # nOuterS, so both are expanded for it
        sgp2 = np.ascontiguousarray(np.broadcast_to(sgp, (nO, nI)))
        with np.errstate(all="ignore"):
            assert ol.oracle().orc_vcoord_pressure_integration_f(fp(p), fp(sap), 0.01, fp(sgp2), 1.0, fp(T), _fpn(q), nO, nO, nz, nI, start_high, _fpn(tp),
                                                                 factor, fp(want)) == 1
        got = V.pressureIntegrationAltitude(p, sap, sgp2, T, q, sap_scale=0.01, start_high_p=start_high, topography=tp, topography_factor=factor).cpu().numpy()
        ok = ~(np.isnan(want) | np.isnan(got))
        assert np.array_equal(np.isnan(want), np.isnan(got))
        # the device's log differs from glibc's in the last bit now and then; the f32 layer thickness then moves by one ulp and the
        # sum by that much: 1e-6 relative
        assert np.allclose(got[ok], want[ok], rtol=1e-6, atol=1e-3)
        assert (np.abs(got[ok] - want[ok]) > 0).mean() < 0.05


def test_pressure_integration_without_direction_fails_loudly(V):
    import fimex_b200
    p = -np.ones((1, 5, 40), np.float32)
    z = np.ones((1, 40), np.float32)
    with pytest.raises(Exception):
        V.pressureIntegrationAltitude(p, z, z, np.ones_like(p), start_high_p=-1)


def test_height_levels_chain_on_device(V):
    """verticalInterpolate.type=height: hybrid pressure -> standard altitude -> minus topography as the input coordinate of K15,
    all device-resident; compared with the oracle chain"""
    import os
    import torch
    import test_gpu_vertical_forward as T
    rng = np.random.default_rng(3)
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_vertical.npz"))
    ap, b = g["en_ap_pa"] * 0.01, g["en_b"]
    nz, nI = 65, 2000
    ps = (101325 - 12 * rng.uniform(0, 2500, (1, nI))).astype(np.float32)
    topo = rng.uniform(0, 1500, nI).astype(np.float32)
    fld = rng.normal(280, 10, (1, nz, nI)).astype(np.float32)
    levels = np.array([0., 10., 50., 100., 250., 500., 1000., 2000., 5000., 10000.])
    want_h = _oracle_levels(V.VCOORD_HYBRID_SIGMA_AP, ap, b, 0.0, ps, 0.01, None, 1, topo, -1.0, 1, nz, nI)
    h = V.verticalLevelField(V.VCOORD_HYBRID_SIGMA_AP, ap, b, 0.0, ps, surface_scale=0.01, post=V.VPOST_STANDARD_ALTITUDE, topography=topo,
                             topography_factor=-1.0)
    assert _ulps(h.cpu().numpy(), want_h) <= 2
    got = V.verticalInterpolate(ol.VINT_LIN, torch.from_numpy(fld).cuda(), h, level1=levels)
    torch.cuda.synchronize()
    want = T._oracle_vertical(ol.VINT_LIN, fld, h.cpu().numpy(), levels=levels)  # same coordinate: the interpolation itself is bit-exact
    assert T._eq(got.cpu().numpy(), want)
