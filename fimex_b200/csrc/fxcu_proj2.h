// fxcu_proj2.h -- second half of fxcu_proj.h (included from there): ob_tran, merc, tmerc/utm, laea, dispatch.
#pragma once

namespace fxproj {

// ------------------------------------------------------------------ ob_tran over longlat: q0 = lamp, q1 = cos(phip), q2 = sin(phip); mode 0 oblique, 1 transverse
FXP_HD bool obtran_fwd(const Crs& c, double lam, double phi, double& x, double& y)
{
    const double lamp = c.q[0], cphip = c.q[1], sphip = c.q[2];
    const double coslam = cos(lam), sinphi = sin(phi), cosphi = cos(phi);
    if (c.mode == 0) {
        x = adjlon(aatan2(cosphi * sin(lam), sphip * cosphi * coslam + cphip * sinphi) + lamp);
        y = aasin(sphip * sinphi - cphip * cosphi * coslam);
    } else {
        x = adjlon(aatan2(cosphi * sin(lam), sinphi) + lamp);
        y = aasin(-cosphi * coslam);
    }
    return true;
}
FXP_HD bool obtran_inv(const Crs& c, double x, double y, double& lam, double& phi)
{
    const double lamp = c.q[0], cphip = c.q[1], sphip = c.q[2];
    const double l = x - lamp;
    const double coslam = cos(l), sinphi = sin(y), cosphi = cos(y);
    if (c.mode == 0) {
        phi = aasin(sphip * sinphi + cphip * cosphi * coslam);
        lam = aatan2(cosphi * sin(l), sphip * cosphi * coslam - cphip * sinphi);
    } else {
        lam = aatan2(cosphi * sin(l), -sinphi);
        phi = aasin(cosphi * coslam);
    }
    return true;
}

// ------------------------------------------------------------------ merc
FXP_HD bool merc_fwd(const Crs& c, double lam, double phi, double& x, double& y)
{
    if (fabs(fabs(phi) - HALFPI) <= EPS10)
        return false;
    x = c.k0 * lam;
    y = c.ellips ? -c.k0 * log(tsfn(phi, sin(phi), c.e)) : c.k0 * log(tan(FORTPI + .5 * phi));
    return true;
}
FXP_HD bool merc_inv(const Crs& c, double x, double y, double& lam, double& phi)
{
    if (c.ellips) {
        bool ok;
        phi = phi2(exp(-y / c.k0), c.e, ok);
        if (!ok)
            return false;
    } else {
        phi = HALFPI - 2. * atan(exp(-y / c.k0));
    }
    lam = x / c.k0;
    return true;
}

// ------------------------------------------------------------------ tmerc / utm
// ellipsoid: q0 = Qn, q1 = Zb, q2..7 = cgb, q8..13 = cbg, q14..19 = utg, q20..25 = gtu; sphere: q2 = k0, q3 = k0/2
FXP_HD double gatg(const double* p1, int len, double B, double cos_2B, double sin_2B)
{
    double h = 0, h1, h2 = 0;
    const double two_cos_2B = 2 * cos_2B;
    int i = len - 1;
    h1 = p1[i];
    while (i > 0) {
        --i;
        h = -h2 + two_cos_2B * h1 + p1[i];
        h2 = h1;
        h1 = h;
    }
    return B + h * sin_2B;
}
FXP_HD void clenS(const double* a, int size, double sin_arg_r, double cos_arg_r, double sinh_arg_i, double cosh_arg_i, double& R, double& I)
{
    double r = 2 * cos_arg_r * cosh_arg_i, i = -2 * sin_arg_r * sinh_arg_i;
    double hr, hr1 = 0, hr2, hi = 0, hi1 = 0, hi2;
    int k = size - 1;
    hr = a[k];
    while (k > 0) {
        --k;
        hr2 = hr1;
        hi2 = hi1;
        hr1 = hr;
        hi1 = hi;
        hr = -hr2 + r * hr1 - i * hi1 + a[k];
        hi = -hi2 + i * hr1 + r * hi1;
    }
    r = sin_arg_r * cosh_arg_i;
    i = cos_arg_r * sinh_arg_i;
    R = r * hr - i * hi;
    I = r * hi + i * hr;
}
FXP_HD double clens(const double* a, int size, double arg_r)
{
    const double r = 2 * cos(arg_r);
    double hr, hr1 = 0, hr2;
    int k = size - 1;
    hr = a[k];
    while (k > 0) {
        --k;
        hr2 = hr1;
        hr1 = hr;
        hr = -hr2 + r * hr1 + a[k];
    }
    return sin(arg_r) * hr;
}
FXP_HD bool tmerc_fwd(const Crs& c, double lam, double phi, double& x, double& y)
{
    if (!c.ellips) {
        const double cosphi = cos(phi);
        double b = cosphi * sin(lam);
        if (fabs(fabs(b) - 1.) <= EPS10)
            return false;
        x = c.q[3] * log((1. + b) / (1. - b));
        double yy = cosphi * cos(lam) / sqrt(1. - b * b);
        b = fabs(yy);
        if (b >= 1.) {
            if ((b - 1.) > EPS10)
                return false;
            yy = 0.;
        } else {
            yy = acos(yy);
        }
        if (phi < 0.)
            yy = -yy;
        y = c.q[2] * (yy - c.phi0);
        return true;
    }
    const double* cbg = c.q + 8;
    const double* gtu = c.q + 20;
    double Cn = gatg(cbg, 6, phi, cos(2 * phi), sin(2 * phi));
    const double sin_Cn = sin(Cn), cos_Cn = cos(Cn), sin_Ce = sin(lam), cos_Ce = cos(lam);
    const double cos_Cn_cos_Ce = cos_Cn * cos_Ce;
    Cn = atan2(sin_Cn, cos_Cn_cos_Ce);
    const double inv_denom = 1. / hypot(sin_Cn, cos_Cn_cos_Ce);
    const double tan_Ce = sin_Ce * cos_Cn * inv_denom;
    double Ce = asinh(tan_Ce);
    const double two_inv = 2 * inv_denom;
    const double two_inv_sq = two_inv * inv_denom;
    const double tmp_r = cos_Cn_cos_Ce * two_inv_sq;
    double dCn, dCe;
    clenS(gtu, 6, sin_Cn * tmp_r, cos_Cn_cos_Ce * tmp_r - 1, tan_Ce * two_inv, two_inv_sq - 1, dCn, dCe);
    Cn += dCn;
    Ce += dCe;
    if (!(fabs(Ce) <= 2.623395162778))
        return false;
    y = c.q[0] * Cn + c.q[1];
    x = c.q[0] * Ce;
    return true;
}
FXP_HD bool tmerc_inv(const Crs& c, double x, double y, double& lam, double& phi)
{
    if (!c.ellips) {
        double h = exp(x / c.q[2]);
        const double g = .5 * (h - 1. / h);
        h = cos(c.phi0 + y / c.q[2]);
        phi = asin(sqrt((1. - h * h) / (1. + g * g)));
        if (y < 0. && -phi + c.phi0 < 0.0)
            phi = -phi;
        lam = (g != 0.0 || h != 0.0) ? atan2(g, h) : 0.;
        return true;
    }
    const double* cgb = c.q + 2;
    const double* utg = c.q + 14;
    double Cn = (y - c.q[1]) / c.q[0];
    double Ce = x / c.q[0];
    if (!(fabs(Ce) <= 2.623395162778))
        return false;
    const double exp_2_Ce = exp(2 * Ce);
    const double half_inv = 0.5 / exp_2_Ce;
    double dCn, dCe;
    clenS(utg, 6, sin(2 * Cn), cos(2 * Cn), 0.5 * exp_2_Ce - half_inv, 0.5 * exp_2_Ce + half_inv, dCn, dCe);
    Cn += dCn;
    Ce += dCe;
    const double sin_Cn = sin(Cn), cos_Cn = cos(Cn), sinhCe = sinh(Ce);
    Ce = atan2(sinhCe, cos_Cn);
    const double modulus = hypot(sinhCe, cos_Cn);
    Cn = atan2(sin_Cn, modulus);
    const double tmp = 2 * modulus / (sinhCe * sinhCe + 1);
    phi = gatg(cgb, 6, Cn, tmp * modulus - 1., sin_Cn * tmp);
    lam = Ce;
    return true;
}

// ------------------------------------------------------------------ laea: q0 sinb1, q1 cosb1, q2 xmf, q3 ymf, q5 qp, q6 dd, q7 rq, q8..10 apa
FXP_HD bool laea_fwd(const Crs& c, double lam, double phi, double& ox, double& oy)
{
    double x = 0, y = 0;
    const double coslam = cos(lam), sinlam = sin(lam);
    if (c.ellips) {
        const double sinb1 = c.q[0], cosb1 = c.q[1], xmf = c.q[2], ymf = c.q[3], qp = c.q[5];
        double qq = qsfn(sin(phi), c.e, c.one_es);
        double sinb = 0, cosb = 0, b;
        if (c.mode == OBLIQ || c.mode == EQUIT) {
            sinb = qq / qp;
            const double cosb2 = 1. - sinb * sinb;
            cosb = cosb2 > 0 ? sqrt(cosb2) : 0;
        }
        if (c.mode == OBLIQ)
            b = 1. + sinb1 * sinb + cosb1 * cosb * coslam;
        else if (c.mode == EQUIT)
            b = 1. + cosb * coslam;
        else if (c.mode == N_POLE) {
            b = HALFPI + phi;
            qq = qp - qq;
        } else {
            b = phi - HALFPI;
            qq = qp + qq;
        }
        if (fabs(b) < EPS10)
            return false;
        if (c.mode == OBLIQ) {
            b = sqrt(2. / b);
            y = ymf * b * (cosb1 * sinb - sinb1 * cosb * coslam);
            x = xmf * b * cosb * sinlam;
        } else if (c.mode == EQUIT) {
            b = sqrt(2. / (1. + cosb * coslam));
            y = b * sinb * ymf;
            x = xmf * b * cosb * sinlam;
        } else if (qq >= 1e-15) {
            b = sqrt(qq);
            x = b * sinlam;
            y = coslam * (c.mode == S_POLE ? b : -b);
        }
    } else {
        const double sinphi = sin(phi), cosphi = cos(phi);
        double cl = coslam;
        if (c.mode == EQUIT || c.mode == OBLIQ) {
            y = (c.mode == EQUIT) ? 1. + cosphi * cl : 1. + c.q[0] * sinphi + c.q[1] * cosphi * cl;
            if (y <= EPS10)
                return false;
            y = sqrt(2. / y);
            x = y * cosphi * sinlam;
            y *= (c.mode == EQUIT) ? sinphi : c.q[1] * sinphi - c.q[0] * cosphi * cl;
        } else {
            if (c.mode == N_POLE)
                cl = -cl;
            if (fabs(phi + c.phi0) < EPS10)
                return false;
            y = FORTPI - phi * .5;
            y = 2. * (c.mode == S_POLE ? cos(y) : sin(y));
            x = y * sinlam;
            y *= cl;
        }
    }
    ox = x;
    oy = y;
    return true;
}
FXP_HD bool laea_inv(const Crs& c, double x, double y, double& lam, double& phi)
{
    if (c.ellips) {
        const double sinb1 = c.q[0], cosb1 = c.q[1], qp = c.q[5], dd = c.q[6], rq = c.q[7];
        double ab;
        if (c.mode == OBLIQ || c.mode == EQUIT) {
            x /= dd;
            y *= dd;
            const double rho = hypot(x, y);
            if (rho < EPS10) {
                lam = 0.;
                phi = c.phi0;
                return true;
            }
            const double sCe = 2. * asin(.5 * rho / rq);
            const double cCe = cos(sCe), ssCe = sin(sCe);
            x *= ssCe;
            if (c.mode == OBLIQ) {
                ab = cCe * sinb1 + y * ssCe * cosb1 / rho;
                y = rho * cosb1 * cCe - y * sinb1 * ssCe;
            } else {
                ab = y * ssCe / rho;
                y = rho * cCe;
            }
        } else {
            if (c.mode == N_POLE)
                y = -y;
            const double qq = x * x + y * y;
            if (qq == 0.0) {
                lam = 0.;
                phi = c.phi0;
                return true;
            }
            ab = 1. - qq / qp;
            if (c.mode == S_POLE)
                ab = -ab;
        }
        lam = atan2(x, y);
        const double beta = asin(ab), t = beta + beta;
        phi = beta + c.q[8] * sin(t) + c.q[9] * sin(t + t) + c.q[10] * sin(t + t + t);
        return true;
    }
    const double rh = hypot(x, y);
    double p = rh * .5;
    if (p > 1.)
        return false;
    p = 2. * asin(p);
    double sinz = 0, cosz = 0;
    if (c.mode == OBLIQ || c.mode == EQUIT) {
        sinz = sin(p);
        cosz = cos(p);
    }
    if (c.mode == EQUIT) {
        p = fabs(rh) <= EPS10 ? 0. : asin(y * sinz / rh);
        x *= sinz;
        y = cosz * rh;
    } else if (c.mode == OBLIQ) {
        p = fabs(rh) <= EPS10 ? c.phi0 : asin(cosz * c.q[0] + y * sinz * c.q[1] / rh);
        x *= sinz * c.q[1];
        y = (cosz - sin(p) * c.q[0]) * rh;
    } else if (c.mode == N_POLE) {
        y = -y;
        p = HALFPI - p;
    } else {
        p -= HALFPI;
    }
    phi = p;
    lam = (y == 0. && (c.mode == EQUIT || c.mode == OBLIQ)) ? 0. : atan2(x, y);
    return true;
}

} // namespace fxproj
#include "fxcu_proj3.h"
namespace fxproj {

// ------------------------------------------------------------------ dispatch
// geographic radians -> CRS coordinates (metres, or degrees for geographic CRS)
FXP_HD bool crs_fwd(const Crs& c, double lam, double phi, double& x, double& y)
{
    if (!isfinite(lam) || !isfinite(phi))
        return false;
    if (c.kind == LONGLAT) {
        x = lam * R2D;
        y = phi * R2D;
        return true;
    }
    if (fabs(phi) - HALFPI > 1e-12)
        return false;
    if (fabs(phi) > HALFPI)
        phi = phi < 0 ? -HALFPI : HALFPI;
    lam -= c.lam0;
    if (!c.over)
        lam = adjlon(lam);
    double px = 0, py = 0;
    bool ok;
    switch (c.kind) {
    case LCC: ok = lcc_fwd(c, lam, phi, px, py); break;
    case STERE: ok = stere_fwd(c, lam, phi, px, py); break;
    case OB_TRAN: ok = obtran_fwd(c, lam, phi, px, py); break;
    case MERC: ok = merc_fwd(c, lam, phi, px, py); break;
    case TMERC: ok = tmerc_fwd(c, lam, phi, px, py); break;
    case LAEA: ok = laea_fwd(c, lam, phi, px, py); break;
    case SINU: ok = sinu_fwd(c, lam, phi, px, py); break;
    case CEA: ok = cea_fwd(c, lam, phi, px, py); break;
    case AEA: ok = aea_fwd(c, lam, phi, px, py); break;
    case AEQD: ok = aeqd_fwd(c, lam, phi, px, py); break;
    case ORTHO: ok = ortho_fwd(c, lam, phi, px, py); break;
    case NSPER: ok = nsper_fwd(c, lam, phi, px, py); break;
    case GEOS: ok = geos_fwd(c, lam, phi, px, py); break;
    case OMERC: ok = omerc_fwd(c, lam, phi, px, py); break;
    default: ok = false; break;
    }
    if (!ok || !isfinite(px) || !isfinite(py))
        return false;
    if (c.kind == OB_TRAN) { // rotated lon/lat stays angular
        x = px * R2D;
        y = py * R2D;
    } else {
        x = c.fr_meter * (c.a * px + c.x0);
        y = c.fr_meter * (c.a * py + c.y0);
    }
    return true;
}
// CRS coordinates -> geographic radians
FXP_HD bool crs_inv(const Crs& c, double x, double y, double& lam, double& phi)
{
    if (!isfinite(x) || !isfinite(y))
        return false;
    if (c.kind == LONGLAT) {
        lam = x * D2R;
        phi = y * D2R;
        return true;
    }
    double l = 0, p = 0;
    bool ok;
    if (c.kind == OB_TRAN) {
        ok = obtran_inv(c, x * D2R, y * D2R, l, p);
    } else {
        const double px = (x * c.to_meter - c.x0) * c.ra, py = (y * c.to_meter - c.y0) * c.ra;
        switch (c.kind) {
        case LCC: ok = lcc_inv(c, px, py, l, p); break;
        case STERE: ok = stere_inv(c, px, py, l, p); break;
        case MERC: ok = merc_inv(c, px, py, l, p); break;
        case TMERC: ok = tmerc_inv(c, px, py, l, p); break;
        case LAEA: ok = laea_inv(c, px, py, l, p); break;
        case SINU: ok = sinu_inv(c, px, py, l, p); break;
        case CEA: ok = cea_inv(c, px, py, l, p); break;
        case AEA: ok = aea_inv(c, px, py, l, p); break;
        case AEQD: ok = aeqd_inv(c, px, py, l, p); break;
        case ORTHO: ok = ortho_inv(c, px, py, l, p); break;
        case NSPER: ok = nsper_inv(c, px, py, l, p); break;
        case GEOS: ok = geos_inv(c, px, py, l, p); break;
        case OMERC: ok = omerc_inv(c, px, py, l, p); break;
        default: ok = false; break;
        }
    }
    if (!ok || !isfinite(l) || !isfinite(p))
        return false;
    l += c.lam0;
    if (!c.over)
        l = adjlon(l);
    lam = l;
    phi = p;
    return true;
}
// crs -> crs: source inverse, no datum shift, destination forward; failures give +inf like PROJ's HUGE_VAL
FXP_HD void crs_transform(const Crs& src, const Crs& dst, double& x, double& y)
{
    if (src.kind == LONGLAT && dst.kind == LONGLAT) // geographic -> geographic without datum shift: degrees pass through untouched
        return;
    double lam, phi, ox, oy;
    if (crs_inv(src, x, y, lam, phi) && crs_fwd(dst, lam, phi, ox, oy)) {
        x = ox;
        y = oy;
    } else {
        x = INFINITY;
        y = INFINITY;
    }
}

} // namespace fxproj
