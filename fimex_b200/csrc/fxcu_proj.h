// fxcu_proj.h -- cartographic projections evaluated on the device (K1), f64 throughout.
//
// Replaces the part of the third-party PROJ library that fimex reaches through src/reproject.cc:56-116
// (proj_create_crs_to_crs / proj_trans_generic).  PROJ is not part of the reference tree; the formulas are the published
// ones PROJ documents for each operation (Snyder, USGS PP 1395; Poder/Engsager for the transverse Mercator series):
// longlat, lcc (1SP/2SP, sphere + ellipsoid), stere (polar / oblique / equatorial, sphere + ellipsoid), ob_tran over longlat
// (rotated lat/lon), merc, tmerc / utm (extended series on the ellipsoid, closed form on the sphere), laea; fxcu_proj3.h adds
// sinu, cea, aea, aeqd, ortho, nsper, geos and omerc.
// One `Crs` is a plain struct passed to kernels by value; every function is __host__ __device__ so that the few scalar
// evaluations the host needs use the same code.
#pragma once

#include <cmath>

#ifdef __CUDACC__
#define FXP_HD __host__ __device__ __forceinline__
#else
#define FXP_HD inline
#endif

namespace fxproj {

enum Kind { LONGLAT = 0, LCC, STERE, OB_TRAN, MERC, TMERC, LAEA, SINU, CEA, AEA, AEQD, ORTHO, NSPER, GEOS, OMERC };
enum Mode { S_POLE = 0, N_POLE = 1, OBLIQ = 2, EQUIT = 3 };

struct Crs {
    int kind;
    int geographic; // coordinates are degrees (longlat and rotated longlat)
    int ellips;     // es != 0
    int mode;
    int over;
    double a, ra, es, e, one_es;
    double lam0, phi0, k0, x0, y0, to_meter, fr_meter;
    double q[32]; // per-projection constants, filled by the host-side set-up in fxcu_proj.cu
};

constexpr double PI = 3.14159265358979323846;
constexpr double HALFPI = 1.57079632679489661923;
constexpr double FORTPI = 0.78539816339744830962;
constexpr double TWOPI = 6.28318530717958647693;
constexpr double D2R = 0.017453292519943295769;
constexpr double R2D = 57.295779513082320877;
constexpr double EPS10 = 1e-10;

FXP_HD double adjlon(double lon)
{
    if (fabs(lon) < PI + 1e-12)
        return lon;
    lon += PI;
    lon -= TWOPI * floor(lon / TWOPI);
    lon -= PI;
    return lon;
}
FXP_HD double msfn(double sinphi, double cosphi, double es) { return cosphi / sqrt(1. - es * sinphi * sinphi); }
FXP_HD double tsfn(double phi, double sinphi, double e)
{
    sinphi *= e;
    return tan(.5 * (HALFPI - phi)) / pow((1. - sinphi) / (1. + sinphi), .5 * e);
}
FXP_HD double phi2(double ts, double e, bool& ok)
{
    const double eccnth = .5 * e;
    double phi = HALFPI - 2. * atan(ts);
    ok = true;
    for (int i = 15; i; --i) {
        const double con = e * sin(phi);
        const double dphi = HALFPI - 2. * atan(ts * pow((1. - con) / (1. + con), eccnth)) - phi;
        phi += dphi;
        if (fabs(dphi) <= 1e-10)
            return phi;
    }
    ok = false;
    return phi;
}
FXP_HD double aasin(double v) { return fabs(v) >= 1. ? (v < 0. ? -HALFPI : HALFPI) : asin(v); }
FXP_HD double aatan2(double n, double d) { return (fabs(n) < 1e-50 && fabs(d) < 1e-50) ? 0. : atan2(n, d); }
FXP_HD double ssfn(double phit, double sinphi, double eccen)
{
    sinphi *= eccen;
    return tan(.5 * (HALFPI + phit)) * pow((1. - sinphi) / (1. + sinphi), .5 * eccen);
}
FXP_HD double qsfn(double sinphi, double e, double one_es)
{
    if (e >= 1e-7) {
        const double con = e * sinphi;
        return one_es * (sinphi / (1.0 - con * con) - (.5 / e) * log((1. - con) / (1.0 + con)));
    }
    return sinphi + sinphi;
}

// ------------------------------------------------------------------ lcc: q0 = n, q1 = c, q2 = rho0
FXP_HD bool lcc_fwd(const Crs& c, double lam, double phi, double& x, double& y)
{
    const double n = c.q[0], cc = c.q[1], rho0 = c.q[2];
    double rho;
    if (fabs(fabs(phi) - HALFPI) < EPS10) {
        if (phi * n <= 0.)
            return false;
        rho = 0.;
    } else {
        rho = cc * (c.ellips ? pow(tsfn(phi, sin(phi), c.e), n) : pow(tan(FORTPI + .5 * phi), -n));
    }
    lam *= n;
    x = c.k0 * (rho * sin(lam));
    y = c.k0 * (rho0 - rho * cos(lam));
    return true;
}
FXP_HD bool lcc_inv(const Crs& c, double x, double y, double& lam, double& phi)
{
    const double n = c.q[0], cc = c.q[1], rho0 = c.q[2];
    x /= c.k0;
    y /= c.k0;
    y = rho0 - y;
    double rho = hypot(x, y);
    if (rho != 0.) {
        if (n < 0.) {
            rho = -rho;
            x = -x;
            y = -y;
        }
        if (c.ellips) {
            bool ok;
            phi = phi2(pow(rho / cc, 1. / n), c.e, ok);
            if (!ok)
                return false;
        } else {
            phi = 2. * atan(pow(cc / rho, 1. / n)) - HALFPI;
        }
        lam = atan2(x, y) / n;
    } else {
        lam = 0.;
        phi = n > 0. ? HALFPI : -HALFPI;
    }
    return true;
}

// ------------------------------------------------------------------ stere: q0 = akm1, q1 = sinX1, q2 = cosX1
FXP_HD bool stere_fwd(const Crs& c, double lam, double phi, double& ox, double& oy)
{
    const double akm1 = c.q[0], sinX1 = c.q[1], cosX1 = c.q[2];
    double coslam = cos(lam), sinlam = sin(lam), sinphi = sin(phi);
    double x = 0, y = 0;
    if (c.ellips) {
        double sinX = 0, cosX = 0, A;
        if (c.mode == OBLIQ || c.mode == EQUIT) {
            const double X = 2. * atan(ssfn(phi, sinphi, c.e)) - HALFPI;
            sinX = sin(X);
            cosX = cos(X);
        }
        if (c.mode == OBLIQ) {
            A = akm1 / (cosX1 * (1. + sinX1 * sinX + cosX1 * cosX * coslam));
            y = A * (cosX1 * sinX - sinX1 * cosX * coslam);
            x = A * cosX;
        } else if (c.mode == EQUIT) {
            A = 1. + cosX * coslam;
            if (A == 0.)
                return false;
            A = akm1 / A;
            y = A * sinX;
            x = A * cosX;
        } else {
            if (c.mode == S_POLE) {
                phi = -phi;
                coslam = -coslam;
                sinphi = -sinphi;
            }
            x = akm1 * tsfn(phi, sinphi, c.e);
            y = -x * coslam;
        }
        x = x * sinlam;
    } else {
        const double cosphi = cos(phi);
        if (c.mode == EQUIT || c.mode == OBLIQ) {
            y = (c.mode == EQUIT) ? 1. + cosphi * coslam : 1. + sinX1 * sinphi + cosX1 * cosphi * coslam;
            if (y <= EPS10)
                return false;
            y = akm1 / y;
            x = y * cosphi * sinlam;
            y *= (c.mode == EQUIT) ? sinphi : cosX1 * sinphi - sinX1 * cosphi * coslam;
        } else {
            if (c.mode == N_POLE) {
                coslam = -coslam;
                phi = -phi;
            }
            if (fabs(phi - HALFPI) < 1e-8)
                return false;
            y = akm1 * tan(FORTPI + .5 * phi);
            x = sinlam * y;
            y *= coslam;
        }
    }
    ox = x;
    oy = y;
    return true;
}
FXP_HD bool stere_inv(const Crs& c, double x, double y, double& lam, double& phi)
{
    const double akm1 = c.q[0], sinX1 = c.q[1], cosX1 = c.q[2];
    if (c.ellips) {
        double tp, phi_l, halfe, halfpi;
        const double rho = hypot(x, y);
        if (c.mode == OBLIQ || c.mode == EQUIT) {
            tp = 2. * atan2(rho * cosX1, akm1);
            const double cosphi = cos(tp), sinphi = sin(tp);
            phi_l = (rho == 0.0) ? asin(cosphi * sinX1) : asin(cosphi * sinX1 + (y * sinphi * cosX1 / rho));
            tp = tan(.5 * (HALFPI + phi_l));
            x *= sinphi;
            y = rho * cosX1 * cosphi - y * sinX1 * sinphi;
            halfpi = HALFPI;
            halfe = .5 * c.e;
        } else {
            if (c.mode == N_POLE)
                y = -y;
            tp = -rho / akm1;
            phi_l = HALFPI - 2. * atan(tp);
            halfpi = -HALFPI;
            halfe = -.5 * c.e;
        }
        for (int i = 8; i--;) {
            const double sinphi = c.e * sin(phi_l);
            double p = 2. * atan(tp * pow((1. + sinphi) / (1. - sinphi), halfe)) - halfpi;
            if (fabs(phi_l - p) < 1e-10) {
                phi = (c.mode == S_POLE) ? -p : p;
                lam = (x == 0. && y == 0.) ? 0. : atan2(x, y);
                return true;
            }
            phi_l = p;
        }
        return false;
    }
    const double rh = hypot(x, y);
    double cang = 2. * atan(rh / akm1);
    const double sinc = sin(cang), cosc = cos(cang);
    lam = 0.;
    if (c.mode == EQUIT) {
        phi = (fabs(rh) <= EPS10) ? 0. : asin(y * sinc / rh);
        if (cosc != 0. || x != 0.)
            lam = atan2(x * sinc, cosc * rh);
    } else if (c.mode == OBLIQ) {
        phi = (fabs(rh) <= EPS10) ? c.phi0 : asin(cosc * sinX1 + y * sinc * cosX1 / rh);
        cang = cosc - sinX1 * sin(phi);
        if (cang != 0. || x != 0.)
            lam = atan2(x * sinc * cosX1, cang * rh);
    } else {
        if (c.mode == N_POLE)
            y = -y;
        phi = (fabs(rh) <= EPS10) ? c.phi0 : asin(c.mode == S_POLE ? -cosc : cosc);
        lam = (x == 0. && y == 0.) ? 0. : atan2(x, y);
    }
    return true;
}

// (continued in fxcu_proj2.h: ob_tran, merc, tmerc, laea and the crs->crs dispatch)
} // namespace fxproj

#include "fxcu_proj2.h"
