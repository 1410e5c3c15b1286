// fxcu_proj3.h -- the remaining projections fimex's Projection classes can emit (src/coordSys/*Projection.cc): sinusoidal,
// Lambert cylindrical equal area, Albers conical equal area, azimuthal equidistant, orthographic, geostationary, oblique
// Mercator (Hotine, centre-line form) and vertical perspective.  Included by fxcu_proj2.h before the dispatch.
// Formulas: Snyder, "Map Projections - A Working Manual" (USGS PP 1395) ch. 14, 10, 25, 20, 9, 23, 30; the geostationary view as
// in the CGMS LRIT/HRIT global specification 4.4.3.2.  PARITY UNPINNED against PROJ (no output of PROJ for these exists in the
// reference tree); pinned by closed forms, limiting cases and round trips (tests/test_more_projections.py).
// Not covered (an error, never a silent approximation): oblique azimuthal equidistant on an ellipsoid (PROJ solves the geodesic
// problem there), orthographic on an ellipsoid (PROJ < 7.2 ignores the flattening, >= 7.2 does not), the two-point form of omerc.
#pragma once

namespace fxproj {

// ------------------------------------------------------------------ meridional distance (series in e^2, Snyder 3-21) and its inverse
FXP_HD void enfn(double es, double* en)
{
    const double C00 = 1., C02 = .25, C04 = .046875, C06 = .01953125, C08 = .01068115234375, C22 = .75, C44 = .46875, C46 = .01302083333333333333,
                 C48 = .00712076822916666666, C66 = .36458333333333333333, C68 = .00569661458333333333, C88 = .3076171875;
    double t;
    en[0] = C00 - es * (C02 + es * (C04 + es * (C06 + es * C08)));
    en[1] = es * (C22 - es * (C04 + es * (C06 + es * C08)));
    en[2] = (t = es * es) * (C44 - es * (C46 + es * C48));
    en[3] = (t *= es) * (C66 - es * C68);
    en[4] = t * es * C88;
}
FXP_HD double mlfn(double phi, double sphi, double cphi, const double* en)
{
    cphi *= sphi;
    sphi *= sphi;
    return en[0] * phi - cphi * (en[1] + sphi * (en[2] + sphi * (en[3] + sphi * en[4])));
}
FXP_HD double inv_mlfn(double arg, double es, const double* en, bool& ok)
{
    const double k = 1. / (1. - es);
    double phi = arg;
    ok = true;
    for (int i = 10; i; --i) { // Newton, rarely more than two steps
        const double s = sin(phi);
        double t = 1. - es * s * s;
        t = (mlfn(phi, s, cos(phi), en) - arg) * (t * sqrt(t)) * k;
        phi -= t;
        if (fabs(t) < 1e-11)
            return phi;
    }
    ok = false;
    return phi;
}
// authalic latitude -> geodetic latitude (series, Snyder 3-18); apa as set up for laea
FXP_HD void authset(double es, double* apa)
{
    double t = es * es;
    apa[0] = es * (1. / 3.) + t * (31. / 180.);
    apa[1] = t * (23. / 360.);
    t *= es;
    apa[0] += t * (517. / 5040.);
    apa[1] += t * (251. / 3780.);
    apa[2] = t * (761. / 45360.);
}
FXP_HD double authlat(double beta, const double* apa)
{
    const double t = beta + beta;
    return beta + apa[0] * sin(t) + apa[1] * sin(t + t) + apa[2] * sin(t + t + t);
}
FXP_HD int azimuthal_mode(double phi0)
{
    if (fabs(fabs(phi0) - HALFPI) < EPS10)
        return phi0 < 0. ? S_POLE : N_POLE;
    return fabs(phi0) < EPS10 ? EQUIT : OBLIQ;
}

// ------------------------------------------------------------------ sinu: q0..4 en
FXP_HD bool sinu_fwd(const Crs& c, double lam, double phi, double& x, double& y)
{
    if (c.ellips) {
        const double s = sin(phi), co = cos(phi);
        y = mlfn(phi, s, co, c.q);
        x = lam * co / sqrt(1. - c.es * s * s);
    } else {
        x = lam * cos(phi);
        y = phi;
    }
    return true;
}
FXP_HD bool sinu_inv(const Crs& c, double x, double y, double& lam, double& phi)
{
    if (c.ellips) {
        bool ok;
        phi = inv_mlfn(y, c.es, c.q, ok);
        if (!ok)
            return false;
        double s = fabs(phi);
        if (s < HALFPI) {
            s = sin(phi);
            lam = x * sqrt(1. - c.es * s * s) / cos(phi);
        } else if (s - EPS10 < HALFPI) {
            lam = 0.;
        } else {
            return false;
        }
    } else {
        if (fabs(y) - EPS10 > HALFPI)
            return false;
        phi = y;
        const double co = cos(y);
        lam = co != 0. ? x / co : 0.;
    }
    return true;
}

// ------------------------------------------------------------------ cea: k0 = cos(lat_ts) [/ sqrt(1 - es sin^2)], q0 qp, q1..3 apa
FXP_HD bool cea_fwd(const Crs& c, double lam, double phi, double& x, double& y)
{
    x = c.k0 * lam;
    y = c.ellips ? .5 * qsfn(sin(phi), c.e, c.one_es) / c.k0 : sin(phi) / c.k0;
    return true;
}
FXP_HD bool cea_inv(const Crs& c, double x, double y, double& lam, double& phi)
{
    if (c.ellips) {
        phi = authlat(aasin(2. * y * c.k0 / c.q[0]), c.q + 1);
    } else {
        y *= c.k0;
        const double t = fabs(y);
        if (t - EPS10 > 1.)
            return false;
        phi = t >= 1. ? (y < 0. ? -HALFPI : HALFPI) : asin(y);
    }
    lam = x / c.k0;
    return true;
}

// ------------------------------------------------------------------ aea: q0 n, q1 c, q2 dd = 1/n, q3 rho0, q4 n2 (sphere), q5 ec
FXP_HD double aea_phi1(double qs, double e, double one_es, bool& ok)
{
    double phi = asin(.5 * qs);
    ok = true;
    if (e < 1e-7)
        return phi;
    for (int i = 15; i; --i) {
        const double sinpi = sin(phi), cospi = cos(phi), con = e * sinpi, com = 1. - con * con;
        const double dphi = .5 * com * com / cospi * (qs / one_es - sinpi / com + .5 / e * log((1. - con) / (1. + con)));
        phi += dphi;
        if (fabs(dphi) <= 1e-10)
            return phi;
    }
    ok = false;
    return phi;
}
FXP_HD bool aea_fwd(const Crs& c, double lam, double phi, double& x, double& y)
{
    double rho = c.q[1] - (c.ellips ? c.q[0] * qsfn(sin(phi), c.e, c.one_es) : c.q[4] * sin(phi));
    if (rho < 0.)
        return false;
    rho = c.q[2] * sqrt(rho);
    lam *= c.q[0];
    x = rho * sin(lam);
    y = c.q[3] - rho * cos(lam);
    return true;
}
FXP_HD bool aea_inv(const Crs& c, double x, double y, double& lam, double& phi)
{
    const double n = c.q[0];
    y = c.q[3] - y;
    double rho = hypot(x, y);
    if (rho != 0.) {
        if (n < 0.) {
            rho = -rho;
            x = -x;
            y = -y;
        }
        phi = rho / c.q[2];
        if (c.ellips) {
            phi = (c.q[1] - phi * phi) / n;
            if (fabs(c.q[5] - fabs(phi)) > 1e-7) {
                bool ok;
                phi = aea_phi1(phi, c.e, c.one_es, ok);
                if (!ok)
                    return false;
            } else {
                phi = phi < 0. ? -HALFPI : HALFPI;
            }
        } else {
            phi = (c.q[1] - phi * phi) / c.q[4];
            phi = fabs(phi) <= 1. ? asin(phi) : (phi < 0. ? -HALFPI : HALFPI);
        }
        lam = atan2(x, y) / n;
    } else {
        lam = 0.;
        phi = n > 0. ? HALFPI : -HALFPI;
    }
    return true;
}

// ------------------------------------------------------------------ aeqd: q0 sin(phi0), q1 cos(phi0), q2..6 en, q7 Mp (polar, ellipsoid)
FXP_HD bool aeqd_fwd(const Crs& c, double lam, double phi, double& x, double& y)
{
    const double sinphi = sin(phi), cosphi = cos(phi);
    double coslam = cos(lam);
    if (c.ellips) { // polar aspects only (checked at set-up)
        if (c.mode == N_POLE)
            coslam = -coslam;
        const double rho = fabs(c.q[7] - mlfn(phi, sinphi, cosphi, c.q + 2));
        x = rho * sin(lam);
        y = rho * coslam;
        return true;
    }
    if (c.mode == EQUIT || c.mode == OBLIQ) {
        y = c.mode == EQUIT ? cosphi * coslam : c.q[0] * sinphi + c.q[1] * cosphi * coslam;
        if (fabs(fabs(y) - 1.) < 1e-14) {
            if (y < 0.)
                return false; // the antipode is a circle
            x = y = 0.;
            return true;
        }
        y = acos(y);
        y /= sin(y);
        x = y * cosphi * sin(lam);
        y *= c.mode == EQUIT ? sinphi : c.q[1] * sinphi - c.q[0] * cosphi * coslam;
        return true;
    }
    if (c.mode == N_POLE) {
        phi = -phi;
        coslam = -coslam;
    }
    if (fabs(phi - HALFPI) < EPS10)
        return false;
    y = HALFPI + phi;
    x = y * sin(lam);
    y *= coslam;
    return true;
}
FXP_HD bool aeqd_inv(const Crs& c, double x, double y, double& lam, double& phi)
{
    double c_rh = hypot(x, y);
    if (c.ellips) {
        if (c_rh < EPS10) {
            phi = c.phi0;
            lam = 0.;
            return true;
        }
        bool ok;
        phi = inv_mlfn(c.mode == N_POLE ? c.q[7] - c_rh : c.q[7] + c_rh, c.es, c.q + 2, ok);
        lam = atan2(x, c.mode == N_POLE ? -y : y);
        return ok;
    }
    if (c_rh > PI) {
        if (c_rh - EPS10 > PI)
            return false;
        c_rh = PI;
    } else if (c_rh < EPS10) {
        phi = c.phi0;
        lam = 0.;
        return true;
    }
    if (c.mode == OBLIQ || c.mode == EQUIT) {
        const double sinc = sin(c_rh), cosc = cos(c_rh);
        if (c.mode == EQUIT) {
            phi = aasin(y * sinc / c_rh);
            x *= sinc;
            y = cosc * c_rh;
        } else {
            phi = aasin(cosc * c.q[0] + y * sinc * c.q[1] / c_rh);
            y = (cosc - c.q[0] * sin(phi)) * c_rh;
            x *= sinc * c.q[1];
        }
        lam = y == 0. ? 0. : atan2(x, y);
    } else if (c.mode == N_POLE) {
        phi = HALFPI - c_rh;
        lam = atan2(x, -y);
    } else {
        phi = c_rh - HALFPI;
        lam = atan2(x, y);
    }
    return true;
}

// ------------------------------------------------------------------ ortho (sphere): q0 sin(phi0), q1 cos(phi0)
FXP_HD bool ortho_fwd(const Crs& c, double lam, double phi, double& x, double& y)
{
    const double cosphi = cos(phi), sinphi = sin(phi);
    double coslam = cos(lam);
    switch (c.mode) {
    case EQUIT:
        if (cosphi * coslam < -EPS10)
            return false;
        y = sinphi;
        break;
    case OBLIQ:
        if (c.q[0] * sinphi + c.q[1] * cosphi * coslam < -EPS10)
            return false;
        y = c.q[1] * sinphi - c.q[0] * cosphi * coslam;
        break;
    default:
        if (c.mode == N_POLE)
            coslam = -coslam;
        if (fabs(phi - c.phi0) - EPS10 > HALFPI)
            return false;
        y = cosphi * coslam;
        break;
    }
    x = cosphi * sin(lam);
    return true;
}
FXP_HD bool ortho_inv(const Crs& c, double x, double y, double& lam, double& phi)
{
    const double rh = hypot(x, y);
    double sinc = rh;
    if (sinc > 1.) {
        if (sinc - 1. > EPS10)
            return false;
        sinc = 1.;
    }
    const double cosc = sqrt(1. - sinc * sinc);
    if (fabs(rh) <= EPS10) {
        phi = c.phi0;
        lam = 0.;
        return true;
    }
    switch (c.mode) {
    case N_POLE:
        y = -y;
        phi = acos(sinc);
        break;
    case S_POLE: phi = -acos(sinc); break;
    case EQUIT:
        phi = y * sinc / rh;
        x *= sinc;
        y = cosc * rh;
        phi = fabs(phi) >= 1. ? (phi < 0. ? -HALFPI : HALFPI) : asin(phi);
        break;
    default:
        phi = cosc * c.q[0] + y * sinc * c.q[1] / rh;
        y = (cosc - c.q[0] * phi) * rh;
        x *= sinc * c.q[1];
        phi = fabs(phi) >= 1. ? (phi < 0. ? -HALFPI : HALFPI) : asin(phi);
        break;
    }
    if (y == 0. && (c.mode == OBLIQ || c.mode == EQUIT))
        lam = x == 0. ? 0. : (x < 0. ? -HALFPI : HALFPI);
    else
        lam = atan2(x, y);
    return true;
}

// ------------------------------------------------------------------ nsper (sphere): q0 sin(phi0), q1 cos(phi0), q2 p, q3 rp, q4 pn1, q5 pfact
FXP_HD bool nsper_fwd(const Crs& c, double lam, double phi, double& x, double& y)
{
    const double sinphi = sin(phi), cosphi = cos(phi);
    double coslam = cos(lam);
    switch (c.mode) {
    case OBLIQ: y = c.q[0] * sinphi + c.q[1] * cosphi * coslam; break;
    case EQUIT: y = cosphi * coslam; break;
    case S_POLE: y = -sinphi; break;
    default: y = sinphi; break;
    }
    if (y < c.q[3])
        return false; // beyond the horizon
    y = c.q[4] / (c.q[2] - y);
    x = y * cosphi * sin(lam);
    switch (c.mode) {
    case OBLIQ: y *= c.q[1] * sinphi - c.q[0] * cosphi * coslam; break;
    case EQUIT: y *= sinphi; break;
    case N_POLE: y *= -(cosphi * coslam); break;
    default: y *= cosphi * coslam; break;
    }
    return true;
}
FXP_HD bool nsper_inv(const Crs& c, double x, double y, double& lam, double& phi)
{
    const double rh = hypot(x, y);
    double sinz = 1. - rh * rh * c.q[5];
    if (sinz < 0.)
        return false;
    if (fabs(rh) <= EPS10) {
        lam = 0.;
        phi = c.phi0;
        return true;
    }
    sinz = (c.q[2] - sqrt(sinz)) / (c.q[4] / rh + rh / c.q[4]);
    const double cosz = sqrt(1. - sinz * sinz);
    switch (c.mode) {
    case OBLIQ:
        phi = asin(cosz * c.q[0] + y * sinz * c.q[1] / rh);
        y = (cosz - c.q[0] * sin(phi)) * rh;
        x *= sinz * c.q[1];
        break;
    case EQUIT:
        phi = asin(y * sinz / rh);
        y = cosz * rh;
        x *= sinz;
        break;
    case N_POLE:
        phi = asin(cosz);
        y = -y;
        break;
    default: phi = -asin(cosz); break;
    }
    lam = atan2(x, y);
    return true;
}

// ------------------------------------------------------------------ geos: q0 radius_p, q1 radius_p^2, q2 1/radius_p^2, q3 radius_g, q4 radius_g - 1, q5 C; mode 1: sweep x
FXP_HD bool geos_fwd(const Crs& c, double lam, double phi, double& x, double& y)
{
    double Vx, Vy, Vz;
    if (c.ellips) {
        phi = atan(c.q[1] * tan(phi)); // geocentric latitude
        const double r = c.q[0] / hypot(c.q[0] * cos(phi), sin(phi));
        Vx = r * cos(lam) * cos(phi);
        Vy = r * sin(lam) * cos(phi);
        Vz = r * sin(phi);
        if ((c.q[3] - Vx) * Vx - Vy * Vy - Vz * Vz * c.q[2] < 0.)
            return false; // not visible from the satellite
    } else {
        const double co = cos(phi);
        Vx = cos(lam) * co;
        Vy = sin(lam) * co;
        Vz = sin(phi);
        if ((c.q[3] - Vx) * Vx - Vy * Vy - Vz * Vz < 0.)
            return false;
    }
    const double tmp = c.q[3] - Vx;
    if (c.mode) {
        x = c.q[4] * atan(Vy / hypot(Vz, tmp));
        y = c.q[4] * atan(Vz / tmp);
    } else {
        x = c.q[4] * atan(Vy / tmp);
        y = c.q[4] * atan(Vz / hypot(Vy, tmp));
    }
    return true;
}
FXP_HD bool geos_inv(const Crs& c, double x, double y, double& lam, double& phi)
{
    double Vx = -1., Vy, Vz;
    if (c.mode) {
        Vz = tan(y / c.q[4]);
        Vy = tan(x / c.q[4]) * hypot(1., Vz);
    } else {
        Vy = tan(x / c.q[4]);
        Vz = tan(y / c.q[4]) * hypot(1., Vy);
    }
    double a = c.ellips ? Vz / c.q[0] : Vz;
    a = Vy * Vy + a * a + Vx * Vx;
    const double b = 2. * c.q[3] * Vx;
    const double det = b * b - 4. * a * c.q[5];
    if (det < 0.)
        return false; // the line of sight misses the earth
    const double k = (-b - sqrt(det)) / (2. * a);
    Vx = c.q[3] + k * Vx;
    Vy *= k;
    Vz *= k;
    lam = atan2(Vy, Vx);
    phi = atan(Vz * cos(lam) / Vx);
    if (c.ellips)
        phi = atan(c.q[2] * tan(phi));
    return true;
}

// ------------------------------------------------------------------ omerc (Hotine, Snyder 9-11 ff.): q0 A, q1 B, q2 E, q3 ArB, q4 BrA, q5 rB, q6 singam,
// q7 cosgam, q8 sinrot, q9 cosrot, q10 v_pole_n, q11 v_pole_s, q12 u_0; mode 1: +no_rot.  lam is relative to the lam0 the set-up derives from +lonc.
FXP_HD bool omerc_fwd(const Crs& c, double lam, double phi, double& x, double& y)
{
    const double A = c.q[0], B = c.q[1], E = c.q[2], ArB = c.q[3];
    double u, v;
    if (fabs(fabs(phi) - HALFPI) > EPS10) {
        const double W = E / pow(tsfn(phi, sin(phi), c.e), B);
        const double S = .5 * (W - 1. / W), T = .5 * (W + 1. / W), V = sin(B * lam);
        const double U = (S * c.q[6] - V * c.q[7]) / T;
        if (fabs(fabs(U) - 1.) < EPS10)
            return false;
        v = .5 * ArB * log((1. - U) / (1. + U));
        const double t = cos(B * lam);
        u = fabs(t) < 1e-7 ? A * lam : ArB * atan2(S * c.q[7] + V * c.q[6], t);
    } else {
        v = phi > 0. ? c.q[10] : c.q[11];
        u = ArB * phi;
    }
    if (c.mode) {
        x = u;
        y = v;
    } else {
        u -= c.q[12];
        x = v * c.q[9] + u * c.q[8];
        y = u * c.q[9] - v * c.q[8];
    }
    return true;
}
FXP_HD bool omerc_inv(const Crs& c, double x, double y, double& lam, double& phi)
{
    double u, v;
    if (c.mode) {
        v = y;
        u = x;
    } else {
        v = x * c.q[9] - y * c.q[8];
        u = y * c.q[9] + x * c.q[8] + c.q[12];
    }
    const double Qp = exp(-c.q[4] * v);
    const double Sp = .5 * (Qp - 1. / Qp), Tp = .5 * (Qp + 1. / Qp), Vp = sin(c.q[4] * u);
    const double Up = (Vp * c.q[7] + Sp * c.q[6]) / Tp;
    if (fabs(fabs(Up) - 1.) < EPS10) {
        lam = 0.;
        phi = Up < 0. ? -HALFPI : HALFPI;
        return true;
    }
    bool ok;
    phi = phi2(pow(c.q[2] / sqrt((1. + Up) / (1. - Up)), 1. / c.q[1]), c.e, ok);
    lam = -c.q[5] * atan2(Sp * c.q[7] - Vp * c.q[6], cos(c.q[4] * u));
    return ok;
}

} // namespace fxproj
