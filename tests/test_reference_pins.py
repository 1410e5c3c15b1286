"""The reference's own value-level pins that touch projection arithmetic, replayed AS WRITTEN.

fimex keeps no projection code of its own: src/reproject.cc hands proj4 strings to PROJ (un-vendored, absent here).  What the
reference's tests pin at that boundary is replayed here twice -- CPU: the reference's reproject.cc compiled over this repo's
mini-PROJ (oracle/_ref); GPU: the product's device projections through the C ABI:
  test/testInterpolation.cc:387-428   mifi_vector_reproject_values_rotate_90    (u,v) -> (v,-u) within 1e-4
  test/testInterpolation.cc:430-471   ..._rotate_180                            (u,v) -> (-u,-v) within 1e-4
  test/testInterpolation.cc:473-540   mifi_vector_reproject_keep_size           |(u,v)| kept within 1e-3
  test/testInterpolation.cc:542-605   mifi_vector_reproject_directions          315/270/225, 0, 180, 45/90/135 within 1 degree
  test/testProjections.cc:219-239     test_common_projections                   the production proj4 strings must be usable; the
        reference only checks that they parse -- here each gets an explicit forward pin for (10 E, 60 N) and the inverse back, the
        expected numbers coming from the published closed forms (Snyder 1987: 15-1..15-4 lcc, 21-33..21-34 + 15-9 polar stere;
        the pole rotation of ob_tran) evaluated independently in numpy below -- NOT from PROJ, which no test here can run.
PROJ's own numbers for ob_tran / lcc / stere on the sphere are pinned separately, from the lon/lat fields stored in the reference's
fixtures (tests/test_proj_pins.py); see DESIGN.md section 2 for what remains unpinned.
"""
import ctypes as C
import math

import numpy as np
import pytest

import oracle_lib as ol
from oracle_lib import dp, fp

EMEP0 = "+proj=stere +ellps=WGS84 +lat_0=90 +lon_0=0 +lat_ts=60"
LATLONG = "+proj=latlong +ellps=WGS84"  # latlongProj of test/testInterpolation.cc


# ------------------------------------------------------------------ the two implementations behind one small interface
class RefImpl:
    """reproject.cc of the reference over mini-PROJ (CPU)"""
    def __init__(self):
        self.r = ol.ref()
        if self.r is None:
            pytest.skip("oracle/_ref not built")

    def regrid_nn(self, pin, fld, ax, ay, tin, pout, ox_axis, oy_axis, tout):
        out = np.full((1, oy_axis.size, ox_axis.size), np.nan, np.float32)
        assert self.r.ref_reproject_f(ol.NEAREST, pin.encode(), fp(fld), dp(ax), dp(ay), tin[0], tin[1], ax.size, ay.size, 1, pout.encode(), fp(out),
                                      dp(ox_axis), dp(oy_axis), tout[0], tout[1], ox_axis.size, oy_axis.size) == 0
        return out

    def rotate(self, pin, pout, u, v, ox_axis, oy_axis, tout):
        assert self.r.ref_vector_reproject_values_f(pin.encode(), pout.encode(), fp(u), fp(v), dp(ox_axis), dp(oy_axis), tout[0], tout[1], ox_axis.size,
                                                    oy_axis.size, 1) == 0

    def directions(self, pin, pout, fx, fy, ang):
        m = np.empty(fx.size * 3)
        assert self.r.ref_get_vector_reproject_matrix_field(pin.encode(), pout.encode(), dp(fx), dp(fy), fx.shape[1], fx.shape[0], dp(m)) == 0
        self.r.ref_vector_reproject_direction_by_matrix_f(dp(m), fx.shape[1], fx.shape[0], fp(ang), 1)

    def transform(self, pin, pout, x, y):
        xv, yv = np.array(x, dtype=np.float64), np.array(y, dtype=np.float64)
        err = C.create_string_buffer(200)
        assert ol.oracle().miniproj_transform_str(pin.encode(), pout.encode(), xv.size, dp(xv), dp(yv), err, 200) >= 0, err.value
        return xv, yv


class GpuImpl:
    """libfimexcu through the C ABI"""
    def __init__(self):
        import fimex_b200
        from fimex_b200 import reproject
        assert fimex_b200.lib.fimexcu_device_count() >= 1
        self.fx, self.rp = fimex_b200, reproject

    def regrid_nn(self, pin, fld, ax, ay, tin, pout, ox_axis, oy_axis, tout):
        ci = self.rp.cachedInterpolationFromProjection(ol.NEAREST, pout, pin, ox_axis, oy_axis, ax, ay, tin[0], tin[1])
        out, _ = ci.interpolateValues(ci.reduceInput(fld, ax.size, ay.size))
        return np.asarray(out).reshape(1, oy_axis.size, ox_axis.size)

    def rotate(self, pin, pout, u, v, ox_axis, oy_axis, tout):
        m = self.rp.get_vector_reproject_matrix(pin, pout, ox_axis, oy_axis)
        self.fx.CachedVectorReprojection(m).reprojectValues(u, v)

    def directions(self, pin, pout, fx, fy, ang):
        m = self.rp.get_vector_reproject_matrix_field(pin, pout, fx, fy)
        self.fx.CachedVectorReprojection(m).reprojectDirectionValues(ang)

    def transform(self, pin, pout, x, y):
        return self.rp.reproject_values(pin, pout, np.array(x, dtype=np.float64), np.array(y, dtype=np.float64))


@pytest.fixture(params=["reference_over_miniproj", pytest.param("gpu", marks=pytest.mark.gpu)])
def impl(request):
    return RefImpl() if request.param == "reference_over_miniproj" else GpuImpl()


# ------------------------------------------------------------------ test/testInterpolation.cc:387-471
@pytest.mark.parametrize("lon0", [90, 180])
def test_vector_reproject_values_rotate(impl, lon0):
    emep2 = f"+proj=stere +ellps=WGS84 +lat_0=90 +lon_0={lon0} +lat_ts=60"
    ax = np.arange(5) - 2.0  # "initialize axes around northpole": metres
    u = np.arange(25, dtype=np.float32).reshape(1, 5, 5)
    v = (25 - np.arange(25)).astype(np.float32).reshape(1, 5, 5)
    uo = impl.regrid_nn(EMEP0, u, ax, ax, (0, 0), emep2, ax, ax, (0, 0))
    vo = impl.regrid_nn(EMEP0, v, ax, ax, (0, 0), emep2, ax, ax, (0, 0))
    ur, vr = uo.copy(), vo.copy()
    impl.rotate(EMEP0, emep2, uo, vo, ax, ax, (0, 0))
    assert not np.isnan(uo).any()
    if lon0 == 90:  # "rotation of 90deg -> u->-v, v->u"
        assert np.abs(vr - uo).max() < 1e-4 and np.abs(ur + vo).max() < 1e-4
    else:           # "rotation of 180deg"
        assert np.abs(vr + vo).max() < 1e-4 and np.abs(ur + uo).max() < 1e-4


# ------------------------------------------------------------------ test/testInterpolation.cc:473-540
def test_vector_reproject_keep_size(impl):
    emep = "+proj=stere +ellps=WGS84 +lat_0=90 +lon_0=-32 +lat_ts=60 +x_0=7 +y_0=109"
    ia, ja = np.arange(4) + 6.0, np.arange(4) + 108.0
    lon, lat = np.arange(4) * 60.0, np.arange(4) / 2.0 + 88.5
    u = np.arange(16, dtype=np.float32).reshape(1, 4, 4)
    v = (-16 + np.arange(16)).astype(np.float32).reshape(1, 4, 4)
    uo = impl.regrid_nn(emep, u, ia, ja, (0, 0), LATLONG, lon, lat, (1, 2))
    vo = impl.regrid_nn(emep, v, ia, ja, (0, 0), LATLONG, lon, lat, (1, 2))
    ur, vr = uo.copy(), vo.copy()
    impl.rotate(emep, LATLONG, uo, vo, lon, lat, (1, 2))
    diff2 = uo.astype(np.float64) ** 2 + vo.astype(np.float64) ** 2 - ur.astype(np.float64) ** 2 - vr.astype(np.float64) ** 2
    ok = ~np.isnan(diff2)
    assert (np.abs(diff2[ok]) < 1e-3).all()


# ------------------------------------------------------------------ test/testInterpolation.cc:542-605
def test_vector_reproject_directions(impl):
    ax = (np.arange(5) - 2) * 1000.0
    fx, fy = np.meshgrid(ax, ax)
    fx, fy = np.ascontiguousarray(fx), np.ascontiguousarray(fy)
    ang = np.zeros((1, 5, 5), np.float32)  # "all point to y-axis"
    impl.directions(EMEP0, LATLONG, fx, fy, ang)
    a = ang[0]  # a[j][i] = angles[oy*j + i]

    def close(want, got):
        return abs(got - want) <= 1.0 * max(abs(want), abs(got)) / 100.0 or abs(got - want) < 1.0  # TEST4FIMEX_CHECK_CLOSE(.., .., 1) is 1 %

    assert close(315, a[0][0]) and close(270, a[2][0]) and close(225, a[4][0])
    for j in (0, 1):
        out = a[j][2]
        if out > 300:
            out -= 360
        assert close(10, 10 + out)
    assert close(180, a[3][2]) and close(180, a[4][2])
    assert close(45, a[0][4]) and close(90, a[2][4]) and close(135, a[4][4])


# ------------------------------------------------------------------ test/testProjections.cc:219-239 with explicit pins
R = 6.371e6
WGS84_A, WGS84_F = 6378137.0, 1 / 298.257223563


def _lcc_sphere(lon, lat, lat0, lon0, lat1):
    """Snyder (15-1)..(15-4), one standard parallel, sphere"""
    phi, phi0, phi1 = math.radians(lat), math.radians(lat0), math.radians(lat1)
    n = math.sin(phi1)
    F = math.cos(phi1) * math.tan(math.pi / 4 + phi1 / 2) ** n / n
    rho = R * F / math.tan(math.pi / 4 + phi / 2) ** n
    rho0 = R * F / math.tan(math.pi / 4 + phi0 / 2) ** n
    th = n * math.radians(((lon - lon0 + 180) % 360) - 180)
    return rho * math.sin(th), rho0 - rho * math.cos(th)


def _polar_stere(lon, lat, lon0, lat_ts, a, e, x0=0.0, y0=0.0):
    """Snyder (21-33), (21-34) with t from (15-9): north polar aspect, true scale at lat_ts (e = 0: sphere)"""
    def t_of(phi):
        return math.tan(math.pi / 4 - phi / 2) / ((1 - e * math.sin(phi)) / (1 + e * math.sin(phi))) ** (e / 2)
    phi, phic = math.radians(lat), math.radians(lat_ts)
    mc = math.cos(phic) / math.sqrt(1 - e * e * math.sin(phic) ** 2)
    rho = a * mc * t_of(phi) / t_of(phic)
    lam = math.radians(lon - lon0)
    return x0 + rho * math.sin(lam), y0 - rho * math.cos(lam)


def _rotated_pole(lon, lat, lon0, o_lat_p):
    """general oblique transformation with o_proj=longlat, o_lon_p = 0: coordinates relative to a pole at latitude o_lat_p"""
    lam, phi, phip = math.radians(lon - lon0), math.radians(lat), math.radians(o_lat_p)
    rl = math.atan2(math.cos(phi) * math.sin(lam), math.sin(phip) * math.cos(phi) * math.cos(lam) + math.cos(phip) * math.sin(phi))
    rp = math.asin(math.sin(phip) * math.sin(phi) - math.cos(phip) * math.cos(phi) * math.cos(lam))
    return math.degrees(rl), math.degrees(rp)


E_WGS84 = math.sqrt(WGS84_F * (2 - WGS84_F))
COMMON = [  # (proj4 string of test_common_projections, expected image of (10 E, 60 N), tolerance)
    ("+proj=lcc +lat_0=77.5 +lon_0=335 +lat_1=77.5 +lat_2=77.5 +no_defs +R=6.371e+06", _lcc_sphere(10, 60, 77.5, 335, 77.5), 1e-5),
    ("+proj=lcc +lat_0=77.5 +lon_0=-25 +lat_1=77.5 +lat_2=77.5 +no_defs +R=6.371e+06", _lcc_sphere(10, 60, 77.5, -25, 77.5), 1e-5),
    ("+proj=lcc +lat_0=63.3 +lon_0=15 +lat_1=63.3 +lat_2=63.3 +no_defs +R=6.371e+06", _lcc_sphere(10, 60, 63.3, 15, 63.3), 1e-5),
    ("+proj=lcc +lon_0=15 +lat_0=63 +lat_1=63 +lat_2=63 +R=6.371e+06 +units=m +no_defs", _lcc_sphere(10, 60, 63, 15, 63), 1e-5),
    ("+proj=stere +ellps=WGS84 +lat_0=90.0 +lat_ts=60.0 +x_0=3192800 +y_0=1784000 +lon_0=70",
     _polar_stere(10, 60, 70, 60, WGS84_A, E_WGS84, 3192800, 1784000), 1e-5),
    ("+proj=stere +lat_0=90 +lon_0=70 +lat_ts=60 +units=m +a=6.371e+06 +e=0 +no_defs", _polar_stere(10, 60, 70, 60, R, 0.0), 1e-5),
    ("+proj=ob_tran +o_proj=longlat +lon_0=-40 +o_lat_p=22 +R=6.371e+06 +no_defs", _rotated_pole(10, 60, -40, 22), 1e-10),
    ("+proj=longlat +a=6367470 +e=0 +no_defs", (10.0, 60.0), 1e-12),
    ("+proj=lcc +lat_0=77.5 +lon_0=-25 +lat_1=77.5 +lat_2=77.5 +no_defs +R=6.371e+06", _lcc_sphere(10, 60, 77.5, -25, 77.5), 1e-5),
]


@pytest.mark.parametrize("k", range(len(COMMON)))
def test_common_projections_forward_and_inverse(impl, k):
    p4, (wx, wy), tol = COMMON[k]
    # Projection::convertFromLonLat (src/coordSys/Projection.cc): lon/lat on the projection's own sphere/ellipsoid -> projection
    geo = "+proj=longlat " + " ".join(t for t in p4.split() if t.startswith(("+R=", "+a=", "+e=", "+ellps=")))
    x, y = impl.transform(geo, p4, [10.0], [60.0])
    assert abs(x[0] - wx) <= tol and abs(y[0] - wy) <= tol, (p4, x[0], wx, y[0], wy)
    lon, lat = impl.transform(p4, geo, [wx], [wy])
    assert abs(lon[0] - 10.0) < 1e-9 and abs(lat[0] - 60.0) < 1e-9
