/*
 * miniproj.c -- see miniproj.h.  TEST INFRASTRUCTURE ONLY.
 * Build: gcc -std=gnu99 -O2 -ffp-contract=off.
 */
#include "miniproj.h"

#include <ctype.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define MP_PI 3.14159265358979323846
#define MP_HALFPI 1.57079632679489661923
#define MP_FORTPI 0.78539816339744830962
#define MP_TWOPI 6.28318530717958647693
#define MP_D2R 0.017453292519943295769
#define MP_R2D 57.295779513082320877
#define EPS10 1e-10

/* ------------------------------------------------------------------ string parameters */
typedef struct {
    char key[24];
    char val[64];
} mp_kv;

static int mp_tokenize(const char* def, mp_kv* kv, int maxkv)
{
    int n = 0;
    const char* p = def;
    while (*p && n < maxkv) {
        while (*p && (isspace((unsigned char)*p) || *p == '+'))
            ++p;
        if (!*p)
            break;
        size_t k = 0, v = 0;
        while (*p && !isspace((unsigned char)*p) && *p != '=') {
            if (k + 1 < sizeof(kv[n].key))
                kv[n].key[k++] = *p;
            ++p;
        }
        kv[n].key[k] = 0;
        if (*p == '=') {
            ++p;
            while (*p && !isspace((unsigned char)*p)) {
                if (v + 1 < sizeof(kv[n].val))
                    kv[n].val[v++] = *p;
                ++p;
            }
        }
        kv[n].val[v] = 0;
        ++n;
    }
    return n;
}

static const char* mp_get(const mp_kv* kv, int n, const char* key)
{
    for (int i = 0; i < n; ++i)
        if (strcmp(kv[i].key, key) == 0)
            return kv[i].val;
    return NULL;
}

static int mp_has(const mp_kv* kv, int n, const char* key) { return mp_get(kv, n, key) != NULL; }

static double mp_num(const mp_kv* kv, int n, const char* key, double dflt)
{
    const char* s = mp_get(kv, n, key);
    return (s && *s) ? strtod(s, NULL) : dflt;
}

/* angle in degrees with optional hemisphere suffix -> radians */
static double mp_ang(const mp_kv* kv, int n, const char* key, double dflt_rad)
{
    const char* s = mp_get(kv, n, key);
    if (!s || !*s)
        return dflt_rad;
    char* end = NULL;
    double d = strtod(s, &end);
    if (end && (*end == 'S' || *end == 's' || *end == 'W' || *end == 'w'))
        d = -d;
    return d * MP_D2R;
}

/* ------------------------------------------------------------------ shared helper functions */
static double mp_adjlon(double lon)
{
    if (fabs(lon) < MP_PI + 1e-12)
        return lon;
    lon += MP_PI;
    lon -= MP_TWOPI * floor(lon / MP_TWOPI);
    lon -= MP_PI;
    return lon;
}

static double mp_msfn(double sinphi, double cosphi, double es) { return cosphi / sqrt(1. - es * sinphi * sinphi); }

static double mp_tsfn(double phi, double sinphi, double e)
{
    sinphi *= e;
    return tan(.5 * (MP_HALFPI - phi)) / pow((1. - sinphi) / (1. + sinphi), .5 * e);
}

static double mp_phi2(double ts, double e)
{
    const double eccnth = .5 * e;
    double phi = MP_HALFPI - 2. * atan(ts);
    for (int i = 15; i; --i) {
        const double con = e * sin(phi);
        const double dphi = MP_HALFPI - 2. * atan(ts * pow((1. - con) / (1. + con), eccnth)) - phi;
        phi += dphi;
        if (fabs(dphi) <= 1e-10)
            return phi;
    }
    return HUGE_VAL;
}

static double mp_aasin(double v)
{
    if (fabs(v) >= 1.)
        return v < 0. ? -MP_HALFPI : MP_HALFPI;
    return asin(v);
}

static double mp_aatan2(double n, double d) { return (fabs(n) < 1e-50 && fabs(d) < 1e-50) ? 0. : atan2(n, d); }

/* ------------------------------------------------------------------ ellipsoid */
static int mp_ellipsoid(const mp_kv* kv, int n, miniproj_crs* c, char* err, size_t errlen)
{
    static const struct { const char* id; double a, rf, b; } ell[] = {
        {"WGS84", 6378137.0, 298.257223563, 0},   {"GRS80", 6378137.0, 298.257222101, 0}, {"intl", 6378388.0, 297., 0},
        {"bessel", 6377397.155, 299.1528128, 0},  {"clrk66", 6378206.4, 0, 6356583.8},     {"WGS72", 6378135.0, 298.26, 0},
        {"sphere", 6370997.0, 0, 6370997.0},      {"airy", 6377563.396, 299.3249646, 0},   {"krass", 6378245.0, 298.3, 0},
    };
    double a = 0, es = 0;
    int have = 0;
    if (mp_has(kv, n, "R")) {
        a = mp_num(kv, n, "R", 0);
        es = 0;
        have = 1;
    } else {
        const char* name = mp_get(kv, n, "ellps");
        const char* datum = mp_get(kv, n, "datum");
        if (!name && datum && strcmp(datum, "WGS84") == 0)
            name = "WGS84";
        if (!name && datum && (strcmp(datum, "NAD83") == 0))
            name = "GRS80";
        if (name) {
            for (size_t i = 0; i < sizeof(ell) / sizeof(ell[0]); ++i)
                if (strcmp(ell[i].id, name) == 0) {
                    a = ell[i].a;
                    if (ell[i].rf != 0) {
                        const double f = 1. / ell[i].rf;
                        es = f * (2. - f);
                    } else {
                        const double b = ell[i].b;
                        es = 1. - (b * b) / (a * a);
                    }
                    have = 1;
                }
            if (!have) {
                snprintf(err, errlen, "unknown ellipsoid '%s'", name);
                return -1;
            }
        }
        if (mp_has(kv, n, "a")) {
            a = mp_num(kv, n, "a", 0);
            if (!have)
                es = 0;
            have = 1;
        }
        if (have) {
            if (mp_has(kv, n, "es"))
                es = mp_num(kv, n, "es", 0);
            else if (mp_has(kv, n, "e")) {
                const double e = mp_num(kv, n, "e", 0);
                es = e * e;
            } else if (mp_has(kv, n, "rf")) {
                const double rf = mp_num(kv, n, "rf", 0);
                const double f = (rf != 0) ? 1. / rf : 0;
                es = f * (2. - f);
            } else if (mp_has(kv, n, "f")) {
                const double f = mp_num(kv, n, "f", 0);
                es = f * (2. - f);
            } else if (mp_has(kv, n, "b")) {
                const double b = mp_num(kv, n, "b", 0);
                es = 1. - (b * b) / (a * a);
            }
        }
    }
    if (!have) { /* PROJ >= 6 default */
        a = 6378137.0;
        const double f = 1. / 298.257222101;
        es = f * (2. - f);
    }
    if (!(a > 0) || es < 0 || es >= 1) {
        snprintf(err, errlen, "invalid ellipsoid (a=%g es=%g)", a, es);
        return -1;
    }
    if (es < 1e-18)
        es = 0;
    c->a = a;
    c->ra = 1. / a;
    c->es = es;
    c->e = sqrt(es);
    c->one_es = 1. - es;
    c->rone_es = 1. / c->one_es;
    c->ellips = es != 0;
    return 0;
}

/* ------------------------------------------------------------------ lcc (q: 0 n, 1 c, 2 rho0) */
static int mp_lcc_setup(const mp_kv* kv, int n, miniproj_crs* c, char* err, size_t errlen)
{
    double phi1 = mp_ang(kv, n, "lat_1", 0.);
    double phi2;
    if (mp_has(kv, n, "lat_2")) {
        phi2 = mp_ang(kv, n, "lat_2", 0.);
    } else {
        phi2 = phi1;
        if (!mp_has(kv, n, "lat_0"))
            c->phi0 = phi1;
    }
    if (fabs(phi1 + phi2) < EPS10) {
        snprintf(err, errlen, "lcc: lat_1 = -lat_2");
        return -1;
    }
    double sinphi = sin(phi1), nn = sinphi;
    const double cosphi = cos(phi1);
    const int secant = fabs(phi1 - phi2) >= EPS10;
    double cc, rho0;
    if (c->ellips) {
        const double m1 = mp_msfn(sinphi, cosphi, c->es);
        const double ml1 = mp_tsfn(phi1, sinphi, c->e);
        if (secant) {
            sinphi = sin(phi2);
            nn = log(m1 / mp_msfn(sinphi, cos(phi2), c->es));
            const double ml2 = mp_tsfn(phi2, sinphi, c->e);
            nn /= log(ml1 / ml2);
        }
        cc = rho0 = m1 * pow(ml1, -nn) / nn;
        rho0 *= (fabs(fabs(c->phi0) - MP_HALFPI) < EPS10) ? 0. : pow(mp_tsfn(c->phi0, sin(c->phi0), c->e), nn);
    } else {
        if (secant)
            nn = log(cosphi / cos(phi2)) / log(tan(MP_FORTPI + .5 * phi2) / tan(MP_FORTPI + .5 * phi1));
        cc = cosphi * pow(tan(MP_FORTPI + .5 * phi1), nn) / nn;
        rho0 = (fabs(fabs(c->phi0) - MP_HALFPI) < EPS10) ? 0. : cc * pow(tan(MP_FORTPI + .5 * c->phi0), -nn);
    }
    c->q[0] = nn;
    c->q[1] = cc;
    c->q[2] = rho0;
    return 0;
}

static int mp_lcc_fwd(const miniproj_crs* c, double lam, double phi, double* x, double* y)
{
    const double nn = c->q[0], cc = c->q[1], rho0 = c->q[2];
    double rho;
    if (fabs(fabs(phi) - MP_HALFPI) < EPS10) {
        if (phi * nn <= 0.)
            return -1;
        rho = 0.;
    } else {
        rho = cc * (c->ellips ? pow(mp_tsfn(phi, sin(phi), c->e), nn) : pow(tan(MP_FORTPI + .5 * phi), -nn));
    }
    lam *= nn;
    *x = c->k0 * (rho * sin(lam));
    *y = c->k0 * (rho0 - rho * cos(lam));
    return 0;
}

static int mp_lcc_inv(const miniproj_crs* c, double x, double y, double* lam, double* phi)
{
    const double nn = c->q[0], cc = c->q[1], rho0 = c->q[2];
    x /= c->k0;
    y /= c->k0;
    y = rho0 - y;
    double rho = hypot(x, y);
    if (rho != 0.) {
        if (nn < 0.) {
            rho = -rho;
            x = -x;
            y = -y;
        }
        if (c->ellips) {
            *phi = mp_phi2(pow(rho / cc, 1. / nn), c->e);
            if (*phi == HUGE_VAL)
                return -1;
        } else {
            *phi = 2. * atan(pow(cc / rho, 1. / nn)) - MP_HALFPI;
        }
        *lam = atan2(x, y) / nn;
    } else {
        *lam = 0.;
        *phi = nn > 0. ? MP_HALFPI : -MP_HALFPI;
    }
    return 0;
}

/* ------------------------------------------------------------------ stere (q: 0 akm1, 1 sinX1, 2 cosX1, 3 phits) */
enum { ST_S_POLE = 0, ST_N_POLE = 1, ST_OBLIQ = 2, ST_EQUIT = 3 };

static double mp_ssfn(double phit, double sinphi, double eccen)
{
    sinphi *= eccen;
    return tan(.5 * (MP_HALFPI + phit)) * pow((1. - sinphi) / (1. + sinphi), .5 * eccen);
}

static int mp_stere_setup(const mp_kv* kv, int n, miniproj_crs* c)
{
    double phits = mp_has(kv, n, "lat_ts") ? mp_ang(kv, n, "lat_ts", 0.) : MP_HALFPI;
    double t = fabs(c->phi0);
    if (fabs(t - MP_HALFPI) < EPS10)
        c->mode = c->phi0 < 0. ? ST_S_POLE : ST_N_POLE;
    else
        c->mode = t > EPS10 ? ST_OBLIQ : ST_EQUIT;
    phits = fabs(phits);
    double akm1 = 0, sinX1 = 0, cosX1 = 0;
    if (c->ellips) {
        if (c->mode == ST_N_POLE || c->mode == ST_S_POLE) {
            if (fabs(phits - MP_HALFPI) < EPS10) {
                akm1 = 2. * c->k0 / sqrt(pow(1 + c->e, 1 + c->e) * pow(1 - c->e, 1 - c->e));
            } else {
                t = sin(phits);
                akm1 = cos(phits) / mp_tsfn(phits, t, c->e);
                t *= c->e;
                akm1 /= sqrt(1. - t * t);
            }
        } else {
            t = sin(c->phi0);
            const double X = 2. * atan(mp_ssfn(c->phi0, t, c->e)) - MP_HALFPI;
            t *= c->e;
            akm1 = 2. * c->k0 * cos(c->phi0) / sqrt(1. - t * t);
            sinX1 = sin(X);
            cosX1 = cos(X);
        }
    } else {
        if (c->mode == ST_OBLIQ) {
            sinX1 = sin(c->phi0);
            cosX1 = cos(c->phi0);
            akm1 = 2. * c->k0;
        } else if (c->mode == ST_EQUIT) {
            akm1 = 2. * c->k0;
        } else {
            akm1 = fabs(phits - MP_HALFPI) >= EPS10 ? cos(phits) / tan(MP_FORTPI - .5 * phits) : 2. * c->k0;
        }
    }
    c->q[0] = akm1;
    c->q[1] = sinX1;
    c->q[2] = cosX1;
    c->q[3] = phits;
    return 0;
}

static int mp_stere_fwd(const miniproj_crs* c, double lam, double phi, double* px, double* py)
{
    const double akm1 = c->q[0], sinX1 = c->q[1], cosX1 = c->q[2];
    double coslam = cos(lam), sinlam = sin(lam), sinphi = sin(phi);
    double x = 0, y = 0;
    if (c->ellips) {
        double sinX = 0, cosX = 0, A;
        if (c->mode == ST_OBLIQ || c->mode == ST_EQUIT) {
            const double X = 2. * atan(mp_ssfn(phi, sinphi, c->e)) - MP_HALFPI;
            sinX = sin(X);
            cosX = cos(X);
        }
        switch (c->mode) {
        case ST_OBLIQ:
            A = akm1 / (cosX1 * (1. + sinX1 * sinX + cosX1 * cosX * coslam));
            y = A * (cosX1 * sinX - sinX1 * cosX * coslam);
            x = A * cosX;
            break;
        case ST_EQUIT:
            A = 1. + cosX * coslam;
            if (A == 0.) {
                y = HUGE_VAL;
            } else {
                A = akm1 / A;
                y = A * sinX;
            }
            x = A * cosX;
            break;
        case ST_S_POLE:
            phi = -phi;
            coslam = -coslam;
            sinphi = -sinphi;
            /* fall through */
        default:
            x = akm1 * mp_tsfn(phi, sinphi, c->e);
            y = -x * coslam;
            break;
        }
        x = x * sinlam;
    } else {
        const double cosphi = cos(phi);
        switch (c->mode) {
        case ST_EQUIT:
        case ST_OBLIQ:
            y = (c->mode == ST_EQUIT) ? 1. + cosphi * coslam : 1. + sinX1 * sinphi + cosX1 * cosphi * coslam;
            if (y <= EPS10)
                return -1;
            y = akm1 / y;
            x = y * cosphi * sinlam;
            y *= (c->mode == ST_EQUIT) ? sinphi : cosX1 * sinphi - sinX1 * cosphi * coslam;
            break;
        case ST_N_POLE:
            coslam = -coslam;
            phi = -phi;
            /* fall through */
        default:
            if (fabs(phi - MP_HALFPI) < 1e-8)
                return -1;
            y = akm1 * tan(MP_FORTPI + .5 * phi);
            x = sinlam * y;
            y *= coslam;
            break;
        }
    }
    *px = x;
    *py = y;
    return 0;
}

static int mp_stere_inv(const miniproj_crs* c, double x, double y, double* lam, double* phi)
{
    const double akm1 = c->q[0], sinX1 = c->q[1], cosX1 = c->q[2];
    if (c->ellips) {
        double tp = 0, phi_l = 0, halfe = 0, halfpi = 0;
        const double rho = hypot(x, y);
        if (c->mode == ST_OBLIQ || c->mode == ST_EQUIT) {
            tp = 2. * atan2(rho * cosX1, akm1);
            const double cosphi = cos(tp), sinphi = sin(tp);
            if (rho == 0.0)
                phi_l = asin(cosphi * sinX1);
            else
                phi_l = asin(cosphi * sinX1 + (y * sinphi * cosX1 / rho));
            tp = tan(.5 * (MP_HALFPI + phi_l));
            x *= sinphi;
            y = rho * cosX1 * cosphi - y * sinX1 * sinphi;
            halfpi = MP_HALFPI;
            halfe = .5 * c->e;
        } else {
            if (c->mode == ST_N_POLE)
                y = -y;
            tp = -rho / akm1;
            phi_l = MP_HALFPI - 2. * atan(tp);
            halfpi = -MP_HALFPI;
            halfe = -.5 * c->e;
        }
        for (int i = 8; i--;) {
            const double sinphi = c->e * sin(phi_l);
            double p = 2. * atan(tp * pow((1. + sinphi) / (1. - sinphi), halfe)) - halfpi;
            if (fabs(phi_l - p) < 1e-10) {
                if (c->mode == ST_S_POLE)
                    p = -p;
                *phi = p;
                *lam = (x == 0. && y == 0.) ? 0. : atan2(x, y);
                return 0;
            }
            phi_l = p;
        }
        return -1;
    }
    const double rh = hypot(x, y);
    double cang = 2. * atan(rh / akm1);
    const double sinc = sin(cang), cosc = cos(cang);
    *lam = 0.;
    switch (c->mode) {
    case ST_EQUIT:
        *phi = (fabs(rh) <= EPS10) ? 0. : asin(y * sinc / rh);
        if (cosc != 0. || x != 0.)
            *lam = atan2(x * sinc, cosc * rh);
        break;
    case ST_OBLIQ:
        *phi = (fabs(rh) <= EPS10) ? c->phi0 : asin(cosc * sinX1 + y * sinc * cosX1 / rh);
        cang = cosc - sinX1 * sin(*phi);
        if (cang != 0. || x != 0.)
            *lam = atan2(x * sinc * cosX1, cang * rh);
        break;
    case ST_N_POLE:
        y = -y;
        /* fall through */
    default:
        *phi = (fabs(rh) <= EPS10) ? c->phi0 : asin(c->mode == ST_S_POLE ? -cosc : cosc);
        *lam = (x == 0. && y == 0.) ? 0. : atan2(x, y);
        break;
    }
    return 0;
}

/* ------------------------------------------------------------------ ob_tran with o_proj=longlat
 * q: 0 lamp, 1 cphip, 2 sphip; mode 0 oblique, 1 transverse */
static int mp_obtran_setup(const mp_kv* kv, int n, miniproj_crs* c, char* err, size_t errlen)
{
    const char* o = mp_get(kv, n, "o_proj");
    if (!o || !(strcmp(o, "longlat") == 0 || strcmp(o, "latlong") == 0 || strcmp(o, "latlon") == 0 || strcmp(o, "lonlat") == 0)) {
        snprintf(err, errlen, "ob_tran: only o_proj=longlat is supported");
        return -1;
    }
    double lamp, phip;
    if (mp_has(kv, n, "o_alpha")) {
        const double lamc = mp_ang(kv, n, "o_lon_c", 0.), phic = mp_ang(kv, n, "o_lat_c", 0.), alpha = mp_ang(kv, n, "o_alpha", 0.);
        if (fabs(fabs(phic) - MP_HALFPI) <= EPS10) {
            snprintf(err, errlen, "ob_tran: invalid o_lat_c");
            return -1;
        }
        lamp = lamc + mp_aatan2(-cos(alpha), -sin(alpha) * sin(phic));
        phip = mp_aasin(cos(phic) * sin(alpha));
    } else if (mp_has(kv, n, "o_lat_p")) {
        lamp = mp_ang(kv, n, "o_lon_p", 0.);
        phip = mp_ang(kv, n, "o_lat_p", 0.);
    } else {
        snprintf(err, errlen, "ob_tran: need o_lat_p or o_alpha");
        return -1;
    }
    c->q[0] = lamp;
    if (fabs(phip) > EPS10) {
        c->q[1] = cos(phip);
        c->q[2] = sin(phip);
        c->mode = 0;
    } else {
        c->mode = 1;
    }
    c->es = 0; /* the rotation is done on a sphere */
    c->e = 0;
    c->one_es = c->rone_es = 1;
    c->ellips = 0;
    c->geographic = 1;
    return 0;
}

static int mp_obtran_fwd(const miniproj_crs* c, double lam, double phi, double* x, double* y)
{
    const double lamp = c->q[0], cphip = c->q[1], sphip = c->q[2];
    const double coslam = cos(lam), sinphi = sin(phi), cosphi = cos(phi);
    if (c->mode == 0) {
        *x = mp_adjlon(mp_aatan2(cosphi * sin(lam), sphip * cosphi * coslam + cphip * sinphi) + lamp);
        *y = mp_aasin(sphip * sinphi - cphip * cosphi * coslam);
    } else {
        *x = mp_adjlon(mp_aatan2(cosphi * sin(lam), sinphi) + lamp);
        *y = mp_aasin(-cosphi * coslam);
    }
    return 0;
}

static int mp_obtran_inv(const miniproj_crs* c, double x, double y, double* lam, double* phi)
{
    const double lamp = c->q[0], cphip = c->q[1], sphip = c->q[2];
    const double l = x - lamp;
    const double coslam = cos(l), sinphi = sin(y), cosphi = cos(y);
    if (c->mode == 0) {
        *phi = mp_aasin(sphip * sinphi + cphip * cosphi * coslam);
        *lam = mp_aatan2(cosphi * sin(l), sphip * cosphi * coslam - cphip * sinphi);
    } else {
        const double t = x - lamp;
        *lam = mp_aatan2(cosphi * sin(t), -sin(y));
        *phi = mp_aasin(cosphi * cos(t));
    }
    return 0;
}

/* ------------------------------------------------------------------ merc */
static int mp_merc_setup(const mp_kv* kv, int n, miniproj_crs* c, char* err, size_t errlen)
{
    if (mp_has(kv, n, "lat_ts")) {
        const double phits = fabs(mp_ang(kv, n, "lat_ts", 0.));
        if (phits >= MP_HALFPI) {
            snprintf(err, errlen, "merc: lat_ts >= 90");
            return -1;
        }
        c->k0 = c->ellips ? mp_msfn(sin(phits), cos(phits), c->es) : cos(phits);
    }
    return 0;
}

static int mp_merc_fwd(const miniproj_crs* c, double lam, double phi, double* x, double* y)
{
    if (fabs(fabs(phi) - MP_HALFPI) <= EPS10)
        return -1;
    *x = c->k0 * lam;
    *y = c->ellips ? -c->k0 * log(mp_tsfn(phi, sin(phi), c->e)) : c->k0 * log(tan(MP_FORTPI + .5 * phi));
    return 0;
}

static int mp_merc_inv(const miniproj_crs* c, double x, double y, double* lam, double* phi)
{
    if (c->ellips) {
        *phi = mp_phi2(exp(-y / c->k0), c->e);
        if (*phi == HUGE_VAL)
            return -1;
    } else {
        *phi = MP_HALFPI - 2. * atan(exp(-y / c->k0));
    }
    *lam = x / c->k0;
    return 0;
}

/* ------------------------------------------------------------------ tmerc / utm
 * ellipsoid: Poder/Engsager extended series (6th order in the third flattening n), the algorithm PROJ
 * documents for tmerc/utm; q: 0 Qn, 1 Zb, 2..7 cgb, 8..13 cbg, 14..19 utg, 20..25 gtu.
 * sphere: Snyder (8-1)..(8-3) closed forms; q: 0 esp(unused) 1 ml0(=phi0) 2 aks0 3 aks5 */
static double mp_gatg(const double* p1, int len, double B, double cos_2B, double sin_2B)
{
    double h = 0, h1, h2 = 0;
    const double two_cos_2B = 2 * cos_2B;
    const double* p = p1 + len;
    h1 = *--p;
    while (p - p1) {
        h = -h2 + two_cos_2B * h1 + *--p;
        h2 = h1;
        h1 = h;
    }
    return B + h * sin_2B;
}

static double mp_clenS(const double* a, int size, double sin_arg_r, double cos_arg_r, double sinh_arg_i, double cosh_arg_i, double* R, double* I)
{
    double r, i, hr, hr1, hr2, hi, hi1, hi2;
    const double* p = a + size;
    r = 2 * cos_arg_r * cosh_arg_i;
    i = -2 * sin_arg_r * sinh_arg_i;
    hi1 = hr1 = hi = 0;
    hr = *--p;
    for (; a - p;) {
        hr2 = hr1;
        hi2 = hi1;
        hr1 = hr;
        hi1 = hi;
        hr = -hr2 + r * hr1 - i * hi1 + *--p;
        hi = -hi2 + i * hr1 + r * hi1;
    }
    r = sin_arg_r * cosh_arg_i;
    i = cos_arg_r * sinh_arg_i;
    *R = r * hr - i * hi;
    *I = r * hi + i * hr;
    return *R;
}

static double mp_clens(const double* a, int size, double arg_r)
{
    double r, hr, hr1, hr2, cos_arg_r;
    const double* p = a + size;
    cos_arg_r = cos(arg_r);
    r = 2 * cos_arg_r;
    hr1 = 0;
    hr = *--p;
    for (; a - p;) {
        hr2 = hr1;
        hr1 = hr;
        hr = -hr2 + r * hr1 + *--p;
    }
    return sin(arg_r) * hr;
}

static int mp_tmerc_setup(miniproj_crs* c)
{
    if (!c->ellips) {
        c->q[2] = c->k0;      /* aks0 */
        c->q[3] = .5 * c->k0; /* aks5 */
        return 0;
    }
    const double f = 1. - sqrt(c->one_es); /* es = f(2-f) */
    const double n = f / (2. - f);
    double* cgb = c->q + 2;
    double* cbg = c->q + 8;
    double* utg = c->q + 14;
    double* gtu = c->q + 20;
    double np = n;
    cgb[0] = n * (2 + n * (-2 / 3.0 + n * (-2 + n * (116 / 45.0 + n * (26 / 45.0 + n * (-2854 / 675.0))))));
    cbg[0] = n * (-2 + n * (2 / 3.0 + n * (4 / 3.0 + n * (-82 / 45.0 + n * (32 / 45.0 + n * (4642 / 4725.0))))));
    np *= n;
    cgb[1] = np * (7 / 3.0 + n * (-8 / 5.0 + n * (-227 / 45.0 + n * (2704 / 315.0 + n * (2323 / 945.0)))));
    cbg[1] = np * (5 / 3.0 + n * (-16 / 15.0 + n * (-13 / 9.0 + n * (904 / 315.0 + n * (-1522 / 945.0)))));
    np *= n;
    cgb[2] = np * (56 / 15.0 + n * (-136 / 35.0 + n * (-1262 / 105.0 + n * (73814 / 2835.0))));
    cbg[2] = np * (-26 / 15.0 + n * (34 / 21.0 + n * (8 / 5.0 + n * (-12686 / 2835.0))));
    np *= n;
    cgb[3] = np * (4279 / 630.0 + n * (-332 / 35.0 + n * (-399572 / 14175.0)));
    cbg[3] = np * (1237 / 630.0 + n * (-12 / 5.0 + n * (-24832 / 14175.0)));
    np *= n;
    cgb[4] = np * (4174 / 315.0 + n * (-144838 / 6237.0));
    cbg[4] = np * (-734 / 315.0 + n * (109598 / 31185.0));
    np *= n;
    cgb[5] = np * (601676 / 22275.0);
    cbg[5] = np * (444337 / 155925.0);

    np = n * n;
    const double Qn = c->k0 / (1 + n) * (1 + np * (1 / 4.0 + np * (1 / 64.0 + np / 256.0)));
    utg[0] = n * (-0.5 + n * (2 / 3.0 + n * (-37 / 96.0 + n * (1 / 360.0 + n * (81 / 512.0 + n * (-96199 / 604800.0))))));
    gtu[0] = n * (0.5 + n * (-2 / 3.0 + n * (5 / 16.0 + n * (41 / 180.0 + n * (-127 / 288.0 + n * (7891 / 37800.0))))));
    utg[1] = np * (-1 / 48.0 + n * (-1 / 15.0 + n * (437 / 1440.0 + n * (-46 / 105.0 + n * (1118711 / 3870720.0)))));
    gtu[1] = np * (13 / 48.0 + n * (-3 / 5.0 + n * (557 / 1440.0 + n * (281 / 630.0 + n * (-1983433 / 1935360.0)))));
    np *= n;
    utg[2] = np * (-17 / 480.0 + n * (37 / 840.0 + n * (209 / 4480.0 + n * (-5569 / 90720.0))));
    gtu[2] = np * (61 / 240.0 + n * (-103 / 140.0 + n * (15061 / 26880.0 + n * (167603 / 181440.0))));
    np *= n;
    utg[3] = np * (-4397 / 161280.0 + n * (11 / 504.0 + n * (830251 / 7257600.0)));
    gtu[3] = np * (49561 / 161280.0 + n * (-179 / 168.0 + n * (6601661 / 7257600.0)));
    np *= n;
    utg[4] = np * (-4583 / 161280.0 + n * (108847 / 3991680.0));
    gtu[4] = np * (34729 / 80640.0 + n * (-3418889 / 1995840.0));
    np *= n;
    utg[5] = np * (-20648693 / 638668800.0);
    gtu[5] = np * (212378941 / 319334400.0);

    const double Z = mp_gatg(cbg, 6, c->phi0, cos(2 * c->phi0), sin(2 * c->phi0));
    c->q[0] = Qn;
    c->q[1] = -Qn * (Z + mp_clens(gtu, 6, 2 * Z));
    return 0;
}

static int mp_tmerc_fwd(const miniproj_crs* c, double lam, double phi, double* x, double* y)
{
    if (!c->ellips) {
        const double cosphi = cos(phi);
        double b = cosphi * sin(lam);
        if (fabs(fabs(b) - 1.) <= EPS10)
            return -1;
        *x = c->q[3] * log((1. + b) / (1. - b));
        double yy = cosphi * cos(lam) / sqrt(1. - b * b);
        b = fabs(yy);
        if (b >= 1.) {
            if ((b - 1.) > EPS10)
                return -1;
            yy = 0.;
        } else {
            yy = acos(yy);
        }
        if (phi < 0.)
            yy = -yy;
        *y = c->q[2] * (yy - c->phi0);
        return 0;
    }
    const double* cbg = c->q + 8;
    const double* gtu = c->q + 20;
    double Cn = mp_gatg(cbg, 6, phi, cos(2 * phi), sin(2 * phi));
    const double sin_Cn = sin(Cn), cos_Cn = cos(Cn), sin_Ce = sin(lam), cos_Ce = cos(lam);
    const double cos_Cn_cos_Ce = cos_Cn * cos_Ce;
    Cn = atan2(sin_Cn, cos_Cn_cos_Ce);
    const double inv_denom = 1. / hypot(sin_Cn, cos_Cn_cos_Ce);
    const double tan_Ce = sin_Ce * cos_Cn * inv_denom;
    double Ce = asinh(tan_Ce);
    const double two_inv = 2 * inv_denom;
    const double two_inv_sq = two_inv * inv_denom;
    const double tmp_r = cos_Cn_cos_Ce * two_inv_sq;
    const double sin_arg_r = sin_Cn * tmp_r;
    const double cos_arg_r = cos_Cn_cos_Ce * tmp_r - 1;
    const double sinh_arg_i = tan_Ce * two_inv;
    const double cosh_arg_i = two_inv_sq - 1;
    double dCn, dCe;
    Cn += mp_clenS(gtu, 6, sin_arg_r, cos_arg_r, sinh_arg_i, cosh_arg_i, &dCn, &dCe);
    Ce += dCe;
    if (fabs(Ce) <= 2.623395162778)
        ;
    else
        return -1;
    *y = c->q[0] * Cn + c->q[1];
    *x = c->q[0] * Ce;
    return 0;
}

static int mp_tmerc_inv(const miniproj_crs* c, double x, double y, double* lam, double* phi)
{
    if (!c->ellips) {
        double h = exp(x / c->q[2]);
        const double g = .5 * (h - 1. / h);
        h = cos(c->phi0 + y / c->q[2]);
        *phi = asin(sqrt((1. - h * h) / (1. + g * g)));
        if (y < 0. && -(*phi) + c->phi0 < 0.0)
            *phi = -(*phi);
        *lam = (g != 0.0 || h != 0.0) ? atan2(g, h) : 0.;
        return 0;
    }
    const double* cgb = c->q + 2;
    const double* utg = c->q + 14;
    double Cn = (y - c->q[1]) / c->q[0];
    double Ce = x / c->q[0];
    if (fabs(Ce) > 2.623395162778)
        return -1;
    const double sin_arg_r = sin(2 * Cn), cos_arg_r = cos(2 * Cn);
    const double exp_2_Ce = exp(2 * Ce);
    const double half_inv = 0.5 / exp_2_Ce;
    const double sinh_arg_i = 0.5 * exp_2_Ce - half_inv;
    const double cosh_arg_i = 0.5 * exp_2_Ce + half_inv;
    double dCn, dCe;
    Cn += mp_clenS(utg, 6, sin_arg_r, cos_arg_r, sinh_arg_i, cosh_arg_i, &dCn, &dCe);
    Ce += dCe;
    const double sin_Cn = sin(Cn), cos_Cn = cos(Cn);
    const double sinhCe = sinh(Ce);
    Ce = atan2(sinhCe, cos_Cn);
    const double modulus = hypot(sinhCe, cos_Cn);
    Cn = atan2(sin_Cn, modulus);
    const double tmp = 2 * modulus / (sinhCe * sinhCe + 1);
    const double sin_2Cn = sin_Cn * tmp;
    const double cos_2Cn = tmp * modulus - 1.;
    *phi = mp_gatg(cgb, 6, Cn, cos_2Cn, sin_2Cn);
    *lam = Ce;
    return 0;
}

/* ------------------------------------------------------------------ laea (sphere + ellipsoid)
 * q: 0 sinb1 1 cosb1 2 xmf 3 ymf 4 mmf(unused) 5 qp 6 dd 7 rq 8..10 apa */
static double mp_qsfn(double sinphi, double e, double one_es)
{
    if (e >= 1e-7) {
        const double con = e * sinphi;
        const double div1 = 1.0 - con * con, div2 = 1.0 + con;
        return one_es * (sinphi / div1 - (.5 / e) * log((1. - con) / div2));
    }
    return sinphi + sinphi;
}

static int mp_laea_setup(miniproj_crs* c)
{
    const double t = fabs(c->phi0);
    if (fabs(t - MP_HALFPI) < EPS10)
        c->mode = c->phi0 < 0. ? ST_S_POLE : ST_N_POLE;
    else if (fabs(t) < EPS10)
        c->mode = ST_EQUIT;
    else
        c->mode = ST_OBLIQ;
    if (c->ellips) {
        c->q[5] = mp_qsfn(1., c->e, c->one_es);
        /* authalic latitude series (Snyder 3-18) */
        const double es = c->es;
        double tt = es * es;
        c->q[8] = es * (1. / 3.) + tt * (31. / 180.);
        c->q[9] = tt * (23. / 360.);
        tt *= es;
        c->q[8] += tt * (517. / 5040.);
        c->q[9] += tt * (251. / 3780.);
        c->q[10] = tt * (761. / 45360.);
        double sinphi;
        switch (c->mode) {
        case ST_N_POLE:
        case ST_S_POLE:
            c->q[6] = 1.;
            break;
        case ST_EQUIT:
            c->q[7] = sqrt(.5 * c->q[5]);
            c->q[6] = 1. / c->q[7];
            c->q[2] = 1.;
            c->q[3] = .5 * c->q[5];
            break;
        default:
            c->q[7] = sqrt(.5 * c->q[5]);
            sinphi = sin(c->phi0);
            c->q[0] = mp_qsfn(sinphi, c->e, c->one_es) / c->q[5];
            c->q[1] = sqrt(1. - c->q[0] * c->q[0]);
            c->q[6] = cos(c->phi0) / (sqrt(1. - c->es * sinphi * sinphi) * c->q[7] * c->q[1]);
            c->q[3] = (c->q[2] = c->q[7]) / c->q[6];
            c->q[2] *= c->q[6];
            break;
        }
    } else if (c->mode == ST_OBLIQ) {
        c->q[0] = sin(c->phi0);
        c->q[1] = cos(c->phi0);
    }
    return 0;
}

static int mp_laea_fwd(const miniproj_crs* c, double lam, double phi, double* px, double* py)
{
    double x = 0, y = 0;
    const double coslam = cos(lam), sinlam = sin(lam);
    if (c->ellips) {
        const double sinb1 = c->q[0], cosb1 = c->q[1], xmf = c->q[2], ymf = c->q[3], qp = c->q[5];
        const double sinphi = sin(phi);
        double qq = mp_qsfn(sinphi, c->e, c->one_es);
        double sinb = 0, cosb = 0, b = 0;
        if (c->mode == ST_OBLIQ || c->mode == ST_EQUIT) {
            sinb = qq / qp;
            const double cosb2 = 1. - sinb * sinb;
            cosb = cosb2 > 0 ? sqrt(cosb2) : 0;
        }
        switch (c->mode) {
        case ST_OBLIQ: b = 1. + sinb1 * sinb + cosb1 * cosb * coslam; break;
        case ST_EQUIT: b = 1. + cosb * coslam; break;
        case ST_N_POLE: b = MP_HALFPI + phi; qq = qp - qq; break;
        default: b = phi - MP_HALFPI; qq = qp + qq; break;
        }
        if (fabs(b) < EPS10)
            return -1;
        switch (c->mode) {
        case ST_OBLIQ:
            b = sqrt(2. / b);
            y = ymf * b * (cosb1 * sinb - sinb1 * cosb * coslam);
            x = xmf * b * cosb * sinlam;
            break;
        case ST_EQUIT:
            b = sqrt(2. / (1. + cosb * coslam));
            y = b * sinb * ymf;
            x = xmf * b * cosb * sinlam;
            break;
        default:
            if (qq >= 1e-15) {
                b = sqrt(qq);
                x = b * sinlam;
                y = coslam * (c->mode == ST_S_POLE ? b : -b);
            } else {
                x = y = 0.;
            }
            break;
        }
    } else {
        const double sinphi = sin(phi), cosphi = cos(phi);
        double cl = coslam;
        switch (c->mode) {
        case ST_EQUIT:
        case ST_OBLIQ:
            y = (c->mode == ST_EQUIT) ? 1. + cosphi * cl : 1. + c->q[0] * sinphi + c->q[1] * cosphi * cl;
            if (y <= EPS10)
                return -1;
            y = sqrt(2. / y);
            x = y * cosphi * sinlam;
            y *= (c->mode == ST_EQUIT) ? sinphi : c->q[1] * sinphi - c->q[0] * cosphi * cl;
            break;
        case ST_N_POLE:
            cl = -cl;
            /* fall through */
        default:
            if (fabs(phi + c->phi0) < EPS10)
                return -1;
            y = MP_FORTPI - phi * .5;
            y = 2. * (c->mode == ST_S_POLE ? cos(y) : sin(y));
            x = y * sinlam;
            y *= cl;
            break;
        }
    }
    *px = x;
    *py = y;
    return 0;
}

static double mp_authlat(double beta, const double* apa)
{
    const double t = beta + beta;
    return beta + apa[0] * sin(t) + apa[1] * sin(t + t) + apa[2] * sin(t + t + t);
}

static int mp_laea_inv(const miniproj_crs* c, double x, double y, double* lam, double* phi)
{
    if (c->ellips) {
        const double sinb1 = c->q[0], cosb1 = c->q[1], qp = c->q[5], dd = c->q[6], rq = c->q[7];
        double ab = 0;
        if (c->mode == ST_OBLIQ || c->mode == ST_EQUIT) {
            x /= dd;
            y *= dd;
            const double rho = hypot(x, y);
            if (rho < EPS10) {
                *lam = 0.;
                *phi = c->phi0;
                return 0;
            }
            const double sCe = 2. * asin(.5 * rho / rq);
            const double cCe = cos(sCe), ssCe = sin(sCe);
            x *= ssCe;
            if (c->mode == ST_OBLIQ) {
                ab = cCe * sinb1 + y * ssCe * cosb1 / rho;
                y = rho * cosb1 * cCe - y * sinb1 * ssCe;
            } else {
                ab = y * ssCe / rho;
                y = rho * cCe;
            }
        } else {
            if (c->mode == ST_N_POLE)
                y = -y;
            const double qq = x * x + y * y;
            if (qq == 0.0) {
                *lam = 0.;
                *phi = c->phi0;
                return 0;
            }
            ab = 1. - qq / qp;
            if (c->mode == ST_S_POLE)
                ab = -ab;
        }
        *lam = atan2(x, y);
        *phi = mp_authlat(asin(ab), c->q + 8);
        return 0;
    }
    const double rh = hypot(x, y);
    double p = rh * .5;
    if (p > 1.)
        return -1;
    p = 2. * asin(p);
    double sinz = 0, cosz = 0;
    if (c->mode == ST_OBLIQ || c->mode == ST_EQUIT) {
        sinz = sin(p);
        cosz = cos(p);
    }
    switch (c->mode) {
    case ST_EQUIT:
        p = fabs(rh) <= EPS10 ? 0. : asin(y * sinz / rh);
        x *= sinz;
        y = cosz * rh;
        break;
    case ST_OBLIQ:
        p = fabs(rh) <= EPS10 ? c->phi0 : asin(cosz * c->q[0] + y * sinz * c->q[1] / rh);
        x *= sinz * c->q[1];
        y = (cosz - sin(p) * c->q[0]) * rh;
        break;
    case ST_N_POLE:
        y = -y;
        p = MP_HALFPI - p;
        break;
    default:
        p -= MP_HALFPI;
        break;
    }
    *phi = p;
    *lam = (y == 0. && (c->mode == ST_EQUIT || c->mode == ST_OBLIQ)) ? 0. : atan2(x, y);
    return 0;
}

/* ------------------------------------------------------------------ parse */
#include "miniproj_more.inc"

int miniproj_parse(const char* def, miniproj_crs* c, char* err, size_t errlen)
{
    mp_kv kv[48];
    char dummy[8];
    if (!err) {
        err = dummy;
        errlen = sizeof(dummy);
    }
    err[0] = 0;
    memset(c, 0, sizeof(*c));
    if (strncmp(def, "EPSG:4326", 9) == 0)
        def = "+proj=longlat +datum=WGS84";
    const int n = mp_tokenize(def, kv, 48);
    const char* proj = mp_get(kv, n, "proj");
    if (!proj) {
        snprintf(err, errlen, "no +proj in '%s'", def);
        return -1;
    }
    if (mp_ellipsoid(kv, n, c, err, errlen))
        return -1;
    c->lam0 = mp_ang(kv, n, "lon_0", 0.);
    c->phi0 = mp_ang(kv, n, "lat_0", 0.);
    c->x0 = mp_num(kv, n, "x_0", 0.);
    c->y0 = mp_num(kv, n, "y_0", 0.);
    c->k0 = mp_has(kv, n, "k_0") ? mp_num(kv, n, "k_0", 1.) : mp_num(kv, n, "k", 1.);
    c->over = mp_has(kv, n, "over");
    c->to_meter = 1.;
    if (mp_has(kv, n, "to_meter"))
        c->to_meter = mp_num(kv, n, "to_meter", 1.);
    else if (mp_has(kv, n, "units")) {
        const char* u = mp_get(kv, n, "units");
        if (strcmp(u, "km") == 0)
            c->to_meter = 1000.;
        else if (strcmp(u, "m") == 0)
            c->to_meter = 1.;
        else if (strcmp(u, "dm") == 0)
            c->to_meter = 0.1;
        else if (strcmp(u, "cm") == 0)
            c->to_meter = 0.01;
        else if (strcmp(u, "ft") == 0)
            c->to_meter = 0.3048;
        else {
            snprintf(err, errlen, "unsupported +units=%s", u);
            return -1;
        }
    }
    c->fr_meter = 1. / c->to_meter;

    if (strcmp(proj, "longlat") == 0 || strcmp(proj, "latlong") == 0 || strcmp(proj, "lonlat") == 0 || strcmp(proj, "latlon") == 0) {
        c->kind = MP_LONGLAT;
        c->geographic = 1;
        return 0;
    }
    if (strcmp(proj, "lcc") == 0) {
        c->kind = MP_LCC;
        return mp_lcc_setup(kv, n, c, err, errlen);
    }
    if (strcmp(proj, "stere") == 0 || strcmp(proj, "ups") == 0) {
        c->kind = MP_STERE;
        if (strcmp(proj, "ups") == 0) {
            c->phi0 = mp_has(kv, n, "south") ? -MP_HALFPI : MP_HALFPI;
            c->k0 = .994;
            c->x0 = c->y0 = 2000000.;
            c->lam0 = 0;
        }
        return mp_stere_setup(kv, n, c);
    }
    if (strcmp(proj, "ob_tran") == 0) {
        c->kind = MP_OB_TRAN;
        return mp_obtran_setup(kv, n, c, err, errlen);
    }
    if (strcmp(proj, "merc") == 0) {
        c->kind = MP_MERC;
        return mp_merc_setup(kv, n, c, err, errlen);
    }
    if (strcmp(proj, "tmerc") == 0 || strcmp(proj, "etmerc") == 0) {
        c->kind = MP_TMERC;
        return mp_tmerc_setup(c);
    }
    if (strcmp(proj, "utm") == 0) {
        c->kind = MP_TMERC;
        if (!c->ellips) {
            snprintf(err, errlen, "utm needs an ellipsoid");
            return -1;
        }
        long zone;
        if (mp_has(kv, n, "zone")) {
            zone = (long)mp_num(kv, n, "zone", 0);
            if (zone < 1 || zone > 60) {
                snprintf(err, errlen, "utm: invalid zone");
                return -1;
            }
            zone -= 1;
        } else {
            zone = lround(floor((mp_adjlon(c->lam0) + MP_PI) * 30. / MP_PI));
            if (zone < 0)
                zone = 0;
            else if (zone >= 60)
                zone = 59;
        }
        c->lam0 = (zone + .5) * MP_PI / 30. - MP_PI;
        c->k0 = 0.9996;
        c->phi0 = 0.;
        c->x0 = 500000.;
        c->y0 = mp_has(kv, n, "south") ? 10000000. : 0.;
        return mp_tmerc_setup(c);
    }
    if (strcmp(proj, "laea") == 0) {
        c->kind = MP_LAEA;
        return mp_laea_setup(c);
    }
    if (strcmp(proj, "sinu") == 0) {
        c->kind = MP_SINU;
        return 0;
    }
    if (strcmp(proj, "cea") == 0) {
        c->kind = MP_CEA;
        return mq_cea_setup(kv, n, c, err, errlen);
    }
    if (strcmp(proj, "aea") == 0) {
        c->kind = MP_AEA;
        return mq_aea_setup(kv, n, c, err, errlen);
    }
    if (strcmp(proj, "aeqd") == 0 || strcmp(proj, "ortho") == 0 || strcmp(proj, "nsper") == 0) {
        c->kind = proj[0] == 'o' ? MP_ORTHO : (proj[0] == 'n' ? MP_NSPER : MP_AEQD);
        c->mode = mq_aspect(c->phi0);
        if (c->kind == MP_NSPER) {
            const double h = mp_num(kv, n, "h", 0.);
            if (!(h > 0)) {
                snprintf(err, errlen, "nsper: +h must be positive");
                return -1;
            }
            c->q[0] = 1. + h * c->ra;
            c->es = c->e = 0.; /* spherical by definition: PROJ drops the flattening here */
            c->one_es = c->rone_es = 1.;
            c->ellips = 0;
        } else if (c->ellips && (c->kind == MP_ORTHO || c->mode == ST_OBLIQ || c->mode == ST_EQUIT)) {
            snprintf(err, errlen, "%s: this aspect is not implemented on an ellipsoid", proj);
            return -1;
        }
        return 0;
    }
    if (strcmp(proj, "geos") == 0) {
        c->kind = MP_GEOS;
        const double h = mp_num(kv, n, "h", 0.);
        const char* sweep = mp_get(kv, n, "sweep");
        if (!(h > 0) || c->phi0 != 0. || (sweep && strcmp(sweep, "x") != 0 && strcmp(sweep, "y") != 0)) {
            snprintf(err, errlen, "geos: needs +h > 0, lat_0 = 0, +sweep=x|y");
            return -1;
        }
        c->mode = (sweep && sweep[0] == 'x') ? 1 : 0;
        c->q[0] = 1. + h * c->ra;
        c->q[1] = c->one_es;
        return 0;
    }
    if (strcmp(proj, "omerc") == 0) {
        c->kind = MP_OMERC;
        return mq_omerc_setup(kv, n, c, err, errlen);
    }
    snprintf(err, errlen, "unsupported projection '%s'", proj);
    return -1;
}

/* ------------------------------------------------------------------ fwd / inv / transform */
int miniproj_fwd(const miniproj_crs* c, double lam, double phi, double* x, double* y)
{
    if (!isfinite(lam) || !isfinite(phi))
        return -1;
    if (c->kind == MP_LONGLAT) {
        *x = lam * MP_R2D;
        *y = phi * MP_R2D;
        return 0;
    }
    if (fabs(phi) - MP_HALFPI > 1e-12)
        return -1;
    if (fabs(phi) > MP_HALFPI)
        phi = phi < 0 ? -MP_HALFPI : MP_HALFPI;
    lam -= c->lam0;
    if (!c->over)
        lam = mp_adjlon(lam);
    double px = 0, py = 0;
    int rc;
    switch (c->kind) {
    case MP_LCC: rc = mp_lcc_fwd(c, lam, phi, &px, &py); break;
    case MP_STERE: rc = mp_stere_fwd(c, lam, phi, &px, &py); break;
    case MP_OB_TRAN: rc = mp_obtran_fwd(c, lam, phi, &px, &py); break;
    case MP_MERC: rc = mp_merc_fwd(c, lam, phi, &px, &py); break;
    case MP_TMERC: rc = mp_tmerc_fwd(c, lam, phi, &px, &py); break;
    case MP_LAEA: rc = mp_laea_fwd(c, lam, phi, &px, &py); break;
    case MP_SINU: rc = mq_sinu_fwd(c, lam, phi, &px, &py); break;
    case MP_CEA: rc = mq_cea_fwd(c, lam, phi, &px, &py); break;
    case MP_AEA: rc = mq_aea_fwd(c, lam, phi, &px, &py); break;
    case MP_AEQD: rc = mq_aeqd_fwd(c, lam, phi, &px, &py); break;
    case MP_ORTHO: rc = mq_ortho_fwd(c, lam, phi, &px, &py); break;
    case MP_NSPER: rc = mq_nsper_fwd(c, lam, phi, &px, &py); break;
    case MP_GEOS: rc = mq_geos_fwd(c, lam, phi, &px, &py); break;
    case MP_OMERC: rc = mq_omerc_fwd(c, lam, phi, &px, &py); break;
    default: rc = -1; break;
    }
    if (rc || !isfinite(px) || !isfinite(py))
        return -1;
    if (c->kind == MP_OB_TRAN) { /* rotated lon/lat: angular output, no scaling by a */
        *x = px * MP_R2D;
        *y = py * MP_R2D;
    } else {
        *x = c->fr_meter * (c->a * px + c->x0);
        *y = c->fr_meter * (c->a * py + c->y0);
    }
    return 0;
}

int miniproj_inv(const miniproj_crs* c, double x, double y, double* lam, double* phi)
{
    if (!isfinite(x) || !isfinite(y))
        return -1;
    if (c->kind == MP_LONGLAT) {
        *lam = x * MP_D2R;
        *phi = y * MP_D2R;
        return 0;
    }
    double l = 0, p = 0;
    int rc;
    if (c->kind == MP_OB_TRAN) {
        rc = mp_obtran_inv(c, x * MP_D2R, y * MP_D2R, &l, &p);
    } else {
        const double px = (x * c->to_meter - c->x0) * c->ra;
        const double py = (y * c->to_meter - c->y0) * c->ra;
        switch (c->kind) {
        case MP_LCC: rc = mp_lcc_inv(c, px, py, &l, &p); break;
        case MP_STERE: rc = mp_stere_inv(c, px, py, &l, &p); break;
        case MP_MERC: rc = mp_merc_inv(c, px, py, &l, &p); break;
        case MP_TMERC: rc = mp_tmerc_inv(c, px, py, &l, &p); break;
        case MP_LAEA: rc = mp_laea_inv(c, px, py, &l, &p); break;
        case MP_SINU: rc = mq_sinu_inv(c, px, py, &l, &p); break;
        case MP_CEA: rc = mq_cea_inv(c, px, py, &l, &p); break;
        case MP_AEA: rc = mq_aea_inv(c, px, py, &l, &p); break;
        case MP_AEQD: rc = mq_aeqd_inv(c, px, py, &l, &p); break;
        case MP_ORTHO: rc = mq_ortho_inv(c, px, py, &l, &p); break;
        case MP_NSPER: rc = mq_nsper_inv(c, px, py, &l, &p); break;
        case MP_GEOS: rc = mq_geos_inv(c, px, py, &l, &p); break;
        case MP_OMERC: rc = mq_omerc_inv(c, px, py, &l, &p); break;
        default: rc = -1; break;
        }
    }
    if (rc || !isfinite(l) || !isfinite(p))
        return -1;
    l += c->lam0;
    if (!c->over)
        l = mp_adjlon(l);
    *lam = l;
    *phi = p;
    return 0;
}

size_t miniproj_transform(const miniproj_crs* src, const miniproj_crs* dst, size_t n, double* x, double* y)
{
    size_t failed = 0;
    if (src->kind == MP_LONGLAT && dst->kind == MP_LONGLAT)
        return 0; /* geographic -> geographic, no datum shift: degrees pass through untouched */
    for (size_t i = 0; i < n; ++i) {
        double lam, phi, ox, oy;
        if (miniproj_inv(src, x[i], y[i], &lam, &phi) || miniproj_fwd(dst, lam, phi, &ox, &oy)) {
            x[i] = HUGE_VAL;
            y[i] = HUGE_VAL;
            ++failed;
        } else {
            x[i] = ox;
            y[i] = oy;
        }
    }
    return failed;
}

long miniproj_transform_str(const char* src, const char* dst, size_t n, double* x, double* y, char* err, size_t errlen)
{
    miniproj_crs s, d;
    if (miniproj_parse(src, &s, err, errlen) || miniproj_parse(dst, &d, err, errlen))
        return -1;
    return (long)miniproj_transform(&s, &d, n, x, y);
}

/* ------------------------------------------------------------------ src/reproject.cc restated */
int orc_reproject_axes(const char* proj_in, const char* proj_out, const double* in_x, const double* in_y, int ix, int iy, double* out_x, double* out_y)
{
    miniproj_crs s, d;
    if (miniproj_parse(proj_in, &s, NULL, 0) || miniproj_parse(proj_out, &d, NULL, 0))
        return -1;
    for (int y = 0; y < iy; ++y)
        for (int x = 0; x < ix; ++x) {
            out_x[(size_t)y * ix + x] = in_x[x];
            out_y[(size_t)y * ix + x] = in_y[y];
        }
    miniproj_transform(&s, &d, (size_t)ix * iy, out_x, out_y);
    return 0;
}

static double orc_bearing_deg(double lat0, double lon0, double lat1, double lon1)
{
    /* reproject.cc:241-264; DEG_TO_RAD there is (M_PI / 180) */
    const double d2r = MP_PI / 180;
    lat0 *= d2r; lon0 *= d2r; lat1 *= d2r; lon1 *= d2r;
    const double dlon = lon0 - lon1;
    return atan2(sin(dlon) * cos(lat1), cos(lat0) * sin(lat1) - sin(lat0) * cos(lat1) * cos(dlon));
}

static void orc_store_rotation(double* m, double phi)
{
    m[0] = cos(phi);
    m[1] = sin(phi);
    m[2] = (180 / MP_PI) * phi;
}

/* reproject.cc:310-357.  c2c.fwd == src -> dst */
static int orc_matrix_points_delta(const miniproj_crs* src, const miniproj_crs* dst, const double* in_x, const double* in_y, const double* out_x,
                                   const double* out_y, double dX, double dY, size_t count, double* mtx)
{
    double* tx = (double*)malloc((count ? count : 1) * sizeof(double));
    double* ty = (double*)malloc((count ? count : 1) * sizeof(double));
    const int dst_latlon = dst->geographic;
    if (!dst_latlon) {
        for (size_t i = 0; i < count; ++i) {
            tx[i] = in_x[i] + dX;
            ty[i] = in_y[i];
        }
        miniproj_transform(src, dst, count, tx, ty);
        const double add_x = dX > 0 ? 0 : MP_PI;
        for (size_t i = 0; i < count; ++i)
            mtx[3 * i] = atan2(ty[i] - out_y[i], tx[i] - out_x[i]) + add_x;
    }
    for (size_t i = 0; i < count; ++i) {
        tx[i] = in_x[i];
        ty[i] = in_y[i] + dY;
    }
    miniproj_transform(src, dst, count, tx, ty);
    const double add_y = dY > 0 ? 0 : MP_PI;
    for (size_t i = 0; i < count; ++i) {
        double phi0;
        if (dst_latlon) {
            phi0 = orc_bearing_deg(out_y[i], out_x[i], ty[i], tx[i]) + add_y;
        } else {
            const double phi_x = mtx[3 * i];
            const double phi_y = -atan2(tx[i] - out_x[i], ty[i] - out_y[i]) + add_y;
            phi0 = (phi_y + phi_x) / 2;
        }
        orc_store_rotation(mtx + 3 * i, phi0);
    }
    free(tx);
    free(ty);
    return 0;
}

int orc_get_vector_reproject_matrix(const char* proj_in, const char* proj_out, const double* out_x_axis, const double* out_y_axis, int ox, int oy,
                                    double* matrix3)
{
    miniproj_crs s, d;
    if (miniproj_parse(proj_in, &s, NULL, 0) || miniproj_parse(proj_out, &d, NULL, 0))
        return -1;
    const size_t n = (size_t)ox * oy;
    double* in_x = (double*)malloc((n ? n : 1) * sizeof(double));
    double* in_y = (double*)malloc((n ? n : 1) * sizeof(double));
    double* o_x = (double*)malloc((n ? n : 1) * sizeof(double));
    double* o_y = (double*)malloc((n ? n : 1) * sizeof(double));
    for (int y = 0; y < oy; ++y)
        for (int x = 0; x < ox; ++x) {
            const size_t i = (size_t)y * ox + x;
            in_x[i] = o_x[i] = out_x_axis[x];
            in_y[i] = o_y[i] = out_y_axis[y];
        }
    miniproj_transform(&d, &s, n, in_x, in_y); /* c2c.inv */
    /* reproject.cc:363-397 */
    const double dflt = 1e-3;
    double delta;
    if (ox > 1) {
        if (oy > 1) {
            delta = dflt * (in_x[(size_t)1 * ox + 1] - in_x[0]);
            const size_t ox2 = ox / 2, oy2 = oy / 2;
            const double d2 = dflt * (in_x[(oy2 + 1) * ox + (ox2 + 1)] - in_x[oy2 * ox + ox2]);
            delta += d2;
            delta /= 2;
        } else {
            delta = dflt * (in_x[1] - in_x[0]);
        }
    } else {
        if (oy > 1)
            delta = dflt * (in_x[(size_t)1 * ox] - in_x[0]);
        else
            delta = (in_x[0] > 1) ? (in_x[0] * dflt) : dflt;
    }
    if (fabs(delta) < 1e-9)
        delta = dflt;
    orc_matrix_points_delta(&s, &d, in_x, in_y, o_x, o_y, delta, delta, n, matrix3);
    free(in_x);
    free(in_y);
    free(o_x);
    free(o_y);
    return 0;
}

int orc_get_vector_reproject_matrix_points(const char* proj_in, const char* proj_out, int inputIsMetric, const double* out_x_points,
                                           const double* out_y_points, int ox, int oy, double* matrix3)
{
    miniproj_crs s, d;
    if (miniproj_parse(proj_in, &s, NULL, 0) || miniproj_parse(proj_out, &d, NULL, 0))
        return -1;
    const size_t n = (size_t)ox * oy;
    double* in_x = (double*)malloc((n ? n : 1) * sizeof(double));
    double* in_y = (double*)malloc((n ? n : 1) * sizeof(double));
    memcpy(in_x, out_x_points, n * sizeof(double));
    memcpy(in_y, out_y_points, n * sizeof(double));
    miniproj_transform(&d, &s, n, in_x, in_y);
    const double delta = inputIsMetric ? 100 : 0.00001;
    orc_matrix_points_delta(&s, &d, in_x, in_y, out_x_points, out_y_points, delta, delta, n, matrix3);
    free(in_x);
    free(in_y);
    return 0;
}
