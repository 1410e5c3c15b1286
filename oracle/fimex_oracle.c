/*
 * fimex_oracle.c -- CPU restatement of the fimex regridding hot path (TEST INFRASTRUCTURE ONLY,
 * see fimex_oracle.h).  Build: gcc -std=gnu99 -O2 -ffp-contract=off (oracle/Makefile).
 *
 * Citations are fimex 2.4.1 paths relative to /root/reference.
 */
#include "fimex_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

static float orc_nanf(void) { return nanf(""); }

/* ------------------------------------------------------------------------------------------
 * src/interpolation.c:72-107 -- method names; note "forward_undef_min" maps to FORWARD_MIN
 * (reference quirk, :103-104).
 * ------------------------------------------------------------------------------------------ */
int orc_string_to_interpolation_method(const char* s)
{
    static const struct { const char* name; int method; } table[] = {
        {"bilinear", ORC_INTERPOL_BILINEAR},
        {"nearestneighbor", ORC_INTERPOL_NEAREST_NEIGHBOR},
        {"bicubic", ORC_INTERPOL_BICUBIC},
        {"coord_nearestneighbor", ORC_INTERPOL_COORD_NN},
        {"coord_kdtree", ORC_INTERPOL_COORD_NN_KD},
        {"forward_sum", ORC_INTERPOL_FORWARD_SUM},
        {"forward_mean", ORC_INTERPOL_FORWARD_MEAN},
        {"forward_median", ORC_INTERPOL_FORWARD_MEDIAN},
        {"forward_max", ORC_INTERPOL_FORWARD_MAX},
        {"forward_min", ORC_INTERPOL_FORWARD_MIN},
        {"forward_undef_sum", ORC_INTERPOL_FORWARD_UNDEF_SUM},
        {"forward_undef_mean", ORC_INTERPOL_FORWARD_UNDEF_MEAN},
        {"forward_undef_median", ORC_INTERPOL_FORWARD_UNDEF_MEDIAN},
        {"forward_undef_max", ORC_INTERPOL_FORWARD_UNDEF_MAX},
        {"forward_undef_min", ORC_INTERPOL_FORWARD_MIN}, /* sic */
    };
    for (size_t i = 0; i < sizeof(table) / sizeof(table[0]); ++i)
        if (strcmp(table[i].name, s) == 0)
            return table[i].method;
    return -1;
}

/* ------------------------------------------------------------------------------------------
 * src/interpolation.c:129-152 -- binary search returning the index, or -(insertion point)-1.
 * ------------------------------------------------------------------------------------------ */
int orc_bsearch_double_index(double key, const double* base, int num, int ascending)
{
    int lo = 0, hi = num - 1, mid = 0, cmp = 0;
    while (lo <= hi) {
        mid = (lo + hi) / 2;
        const double b = base[mid];
        cmp = (key == b) ? 0 : ((key > b) ? 1 : -1);
        if (!ascending)
            cmp = -cmp;
        if (cmp == 0)
            return mid;
        if (cmp > 0)
            lo = mid + 1;
        else
            hi = mid - 1;
    }
    return (cmp > 0) ? -(mid + 1) - 1 : -mid - 1;
}

/* ------------------------------------------------------------------------------------------
 * src/interpolation.c:154-227 -- coordinate -> fractional index on an axis.
 * ------------------------------------------------------------------------------------------ */
int orc_points2position(double* pts, size_t n, const double* axis, int num, int axis_type)
{
    const int ascending = axis[0] < axis[num - 1];
    int circular = 0;
    if (axis_type == ORC_LONGITUDE) {
        if (axis[0] < 0 || axis[num - 1] < 0) { /* axis uses -180..180 */
            for (size_t i = 0; i < n; ++i)
                if (pts[i] > 180.)
                    pts[i] -= 360.;
        } else { /* axis uses 0..360 */
            for (size_t i = 0; i < n; ++i)
                if (pts[i] < 0)
                    pts[i] += 360.;
        }
        double next = axis[num - 1] + (axis[1] - axis[0]) * 1.01;
        if (ascending) {
            next -= 360.;
            if (next >= axis[0])
                circular = 1;
        } else {
            next += 360.;
            if (next <= axis[0])
                circular = 1;
        }
    }
    for (size_t i = 0; i < n; ++i) {
        if (!isfinite(pts[i])) {
            pts[i] = -999.;
            continue;
        }
        const int pos = orc_bsearch_double_index(pts[i], axis, num, ascending);
        if (pos >= 0) {
            pts[i] = (double)pos;
            continue;
        }
        int k = -(pos + 1); /* first element beyond the key */
        if (k == num)
            k--; /* extrapolate with the last interval */
        else if (k == 0)
            k++; /* extrapolate with the first interval */
        const double slope = axis[k] - axis[k - 1];
        const double offset = axis[k] - (slope * k);
        double p = (pts[i] - offset) / slope;
        if (circular && p <= -0.5)
            p += num;
        if (circular && p > (num - 0.5))
            p -= num;
        pts[i] = p;
    }
    return ORC_OK;
}

/* Index tables are finite by construction (points2position maps non-finite to -999, the kd-tree to
 * -1000).  The reference's casts of NaN / out-of-int-range doubles are undefined behaviour; this
 * restatement (and the CUDA path) define such entries as "outside". */
static int orc_table_entry_usable(double x, double y)
{
    return isfinite(x) && isfinite(y) && fabs(x) < 2147483000. && fabs(y) < 2147483000.;
}

static void orc_fill_nan(float* out, int iz)
{
    for (int z = 0; z < iz; ++z)
        out[z] = orc_nanf();
}

/* src/interpolation.c:251-265 */
int orc_get_values_f(const float* in, float* out, double x, double y, int ix, int iy, int iz)
{
    if (!orc_table_entry_usable(x, y)) {
        orc_fill_nan(out, iz);
        return ORC_OK;
    }
    const int rx = (int)lround(x), ry = (int)lround(y);
    if (rx >= 0 && rx < ix && ry >= 0 && ry < iy) {
        for (int z = 0; z < iz; ++z)
            out[z] = in[((size_t)z * iy + ry) * ix + rx];
    } else {
        orc_fill_nan(out, iz);
    }
    return ORC_OK;
}

/* ------------------------------------------------------------------------------------------
 * src/interpolation.c:267-343 -- bilinear with the reference's edge rules:
 *   x,y interior           : 4-tap formula, all float
 *   x interior, y outside  : row lround(y) if inside, linear in x          (:289-305)
 *   x outside              : column lround(x) if inside; y interior -> linear in y (:311-319),
 *                            else row lround(y), accepted when 0 <= y0 <= iy (sic, :322).  y0 == iy
 *                            happens for iy-0.5 <= y < iy+0.5 with x within half a cell outside:
 *                            the reference then reads row 0 of the NEXT level (past the end of the
 *                            array on the last level: undefined behaviour).  Defined as NaN here
 *                            and in the CUDA path; parity tests exclude that corner strip.
 * ------------------------------------------------------------------------------------------ */
int orc_get_values_bilinear_f(const float* in, float* out, double x, double y, int ix, int iy, int iz)
{
    if (!orc_table_entry_usable(x, y)) {
        orc_fill_nan(out, iz);
        return ORC_OK;
    }
    int x0 = (int)floor(x);
    const int x1 = x0 + 1;
    const float xfrac = (float)(x - x0);
    int y0 = (int)floor(y);
    const int y1 = y0 + 1;
    const float yfrac = (float)(y - y0);
    const size_t layer = (size_t)ix * iy;
    if (0 <= x0 && x1 < ix) {
        if (0 <= y0 && y1 < iy) {
            size_t pos = (size_t)y0 * ix + x0;
            for (int z = 0; z < iz; ++z, pos += layer) {
                const float s00 = in[pos], s01 = in[pos + 1], s10 = in[pos + ix], s11 = in[pos + ix + 1];
                out[z] = (1.f - yfrac) * ((1.f - xfrac) * s00 + xfrac * s01) + yfrac * ((1.f - xfrac) * s10 + xfrac * s11);
            }
        } else {
            y0 = (int)lround(y);
            if (0 <= y0 && y0 < iy) {
                size_t pos = (size_t)y0 * ix + x0;
                for (int z = 0; z < iz; ++z, pos += layer) {
                    const float s00 = in[pos], s01 = in[pos + 1];
                    out[z] = (1.f - xfrac) * s00 + xfrac * s01;
                }
            } else {
                orc_fill_nan(out, iz);
            }
        }
    } else {
        x0 = (int)lround(x);
        if (0 <= x0 && x0 < ix) {
            if (0 <= y0 && y1 < iy) {
                size_t pos = (size_t)y0 * ix + x0;
                for (int z = 0; z < iz; ++z, pos += layer) {
                    const float s00 = in[pos], s10 = in[pos + ix];
                    out[z] = (1 - yfrac) * s00 + (yfrac * s10);
                }
            } else {
                y0 = (int)lround(y);
                if (0 <= y0 && y0 < iy) { /* reference tests y0 <= iy; y0 == iy is its out-of-bounds read */
                    size_t pos = (size_t)y0 * ix + x0;
                    for (int z = 0; z < iz; ++z, pos += layer)
                        out[z] = in[pos];
                } else {
                    orc_fill_nan(out, iz);
                }
            }
        } else {
            orc_fill_nan(out, iz);
        }
    }
    return ORC_OK;
}

/* ------------------------------------------------------------------------------------------
 * src/interpolation.c:345-414 -- bicubic convolution (a = -0.5): weights in double, the four
 * column sums are added INTO the float output one at a time (narrowing after every add).
 * ------------------------------------------------------------------------------------------ */
static void orc_cubic_weights(double frac, double w[4])
{
    /* rows of the a=-0.5 convolution matrix, already halved (:348-354) */
    static const double M[4][4] = {{0.0, 1.0, 0.0, 0.0}, {-0.5, 0.0, 0.5, 0.0}, {1.0, -2.5, 2.0, -0.5}, {-0.5, 1.5, -1.5, 0.5}};
    double p[4];
    p[0] = 1;
    p[1] = frac;
    p[2] = frac * frac;
    p[3] = p[2] * frac;
    for (int i = 0; i < 4; ++i) {
        double acc = 0;
        for (int j = 0; j < 4; ++j)
            acc += p[j] * M[j][i];
        w[i] = acc;
    }
}

int orc_get_values_bicubic_f(const float* in, float* out, double x, double y, int ix, int iy, int iz)
{
    if (!orc_table_entry_usable(x, y)) {
        orc_fill_nan(out, iz);
        return ORC_OK;
    }
    const int x0 = (int)floor(x), y0 = (int)floor(y);
    const double xfrac = x - x0, yfrac = y - y0;
    if (!(1 <= x0 && x0 + 2 < ix && 1 <= y0 && y0 + 2 < iy)) {
        orc_fill_nan(out, iz);
        return ORC_OK;
    }
    double wx[4], wy[4];
    orc_cubic_weights(xfrac, wx);
    orc_cubic_weights(yfrac, wy);
    const size_t layer = (size_t)ix * iy;
    for (int z = 0; z < iz; ++z) {
        const float* p = in + (size_t)z * layer;
        float acc = 0;
        for (int j = 0; j < 4; ++j) { /* j: y-tap */
            double col = 0;
            for (int i = 0; i < 4; ++i) /* i: x-tap */
                col += wx[i] * (double)p[(size_t)(y0 + j - 1) * ix + (x0 + i - 1)];
            acc = (float)(acc + col * wy[j]);
        }
        out[z] = acc;
    }
    return ORC_OK;
}

/* ------------------------------------------------------------------------------------------
 * src/interpolation.c:416-546 -- two-level 1-D kernels (vertical / time).
 * ------------------------------------------------------------------------------------------ */
int orc_get_values_nearest_f(const float* a, const float* b, float* out, size_t n, double xa, double xb, double x)
{
    (void)b; (void)xa; (void)xb; (void)x;
    memcpy(out, a, n * sizeof(float));
    return ORC_OK;
}

static float orc_lin_factor(double xa, double xb, double x)
{
    return (xa == xb) ? 0.f : (float)((x - xa) / (xb - xa));
}

static void orc_lin_simple(const float* a, const float* b, float* out, size_t n, float f)
{
    for (size_t i = 0; i < n; ++i)
        out[i] = a[i] + f * (b[i] - a[i]);
}

int orc_get_values_linear_f(const float* a, const float* b, float* out, size_t n, double xa, double xb, double x)
{
    const float f = orc_lin_factor(xa, xb, x);
    if (f == 0)
        memcpy(out, a, n * sizeof(float)); /* exact copies avoid 0*NaN (:442-447) */
    else if (f == 1)
        memcpy(out, b, n * sizeof(float));
    else
        orc_lin_simple(a, b, out, n, f);
    return ORC_OK;
}

static int orc_lin_limited(float lo, float hi, const float* a, const float* b, float* out, size_t n, double xa, double xb, double x)
{
    const float f = orc_lin_factor(xa, xb, x);
    if (f == 0)
        memcpy(out, a, n * sizeof(float));
    else if (f == 1)
        memcpy(out, b, n * sizeof(float));
    else if (f >= lo && f <= hi)
        orc_lin_simple(a, b, out, n, f);
    else
        for (size_t i = 0; i < n; ++i)
            out[i] = orc_nanf();
    return ORC_OK;
}

int orc_get_values_linear_weak_extrapol_f(const float* a, const float* b, float* out, size_t n, double xa, double xb, double x)
{
    return orc_lin_limited(-1.f, 2.f, a, b, out, n, xa, xb, x);
}

int orc_get_values_linear_no_extrapol_f(const float* a, const float* b, float* out, size_t n, double xa, double xb, double x)
{
    return orc_lin_limited(0.f, 1.f, a, b, out, n, xa, xb, x);
}

int orc_get_values_linear_const_extrapol_f(const float* a, const float* b, float* out, size_t n, double xa, double xb, double x)
{
    const float f = orc_lin_factor(xa, xb, x);
    if (f >= 1)
        memcpy(out, b, n * sizeof(float));
    else if (f <= 0)
        memcpy(out, a, n * sizeof(float));
    else
        orc_lin_simple(a, b, out, n, f);
    return ORC_OK;
}

int orc_get_values_log_f(const float* a, const float* b, float* out, size_t n, double xa, double xb, double x)
{
    if (xa <= 0 || xb <= 0 || x <= 0)
        return ORC_ERROR; /* output untouched, as in the reference (:524-526) */
    return orc_get_values_linear_f(a, b, out, n, log(xa), log(xb), log(x));
}

int orc_get_values_log_log_f(const float* a, const float* b, float* out, size_t n, double xa, double xb, double x)
{
    if (xa <= 0 || xb <= 0 || x <= 0)
        return ORC_ERROR;
    const double e = 2.7182818284590452354; /* M_E */
    return orc_get_values_log_f(a, b, out, n, log(xa + e), log(xb + e), log(x + e));
}

void orc_bad2nanf(float* v, size_t n, float bad)
{
    if (isnan(bad))
        return;
    for (size_t i = 0; i < n; ++i)
        if (v[i] == bad)
            v[i] = orc_nanf();
}

void orc_nanf2bad(float* v, size_t n, float bad)
{
    if (isnan(bad))
        return;
    for (size_t i = 0; i < n; ++i)
        if (isnan(v[i]))
            v[i] = bad;
}

/* src/CDMInterpolator.cc:140-145 */
void orc_i16_to_interpolation_array(const short* in, size_t n, double badValue, float* out)
{
    size_t i;
    for (i = 0; i < n; ++i)
        out[i] = (float)in[i]; /* data_caster<float,short>: static_cast, no rounding involved */
    orc_bad2nanf(out, n, (float)badValue); /* the double fill value narrows at the call (interpolation.h:310) */
}

/* src/CDMInterpolator.cc:147-151 */
void orc_interpolation_array_to_i16(const float* in, size_t n, double badValue, short* out)
{
    /* ScaleValue<float,short>(oldFill = NaN, 1, 0, newFill = badValue, 1, 0): oldScale/newScale = 1, (oldOffset-newOffset)/newScale = 0 */
    const short newFill = (short)badValue;
    size_t i;
    for (i = 0; i < n; ++i) {
        const float v = in[i];
        if (v != v) { /* in == NaN is never true; mifi_isnan(in) is */
            out[i] = newFill;
        } else {
            const double d = 1.0 * v + 0.0;
            const int r = (int)lround(d); /* MathUtils.h:40-43: int round(double) { return ::lround(num); } */
            out[i] = (short)r;            /* data_rounder: static_cast<OUT>(round(in)) */
        }
    }
}

/* src/CDMReader.cc:177-192 + Data::asFloat() */
void orc_i16_scaled_to_float(const short* in, size_t n, double fill, double scale, double offset, float* out)
{
    const short oldFill = (short)fill;
    const double a = scale / 1.0, b = (offset - 0.0) / 1.0;
    size_t i;
    for (i = 0; i < n; ++i) {
        const double d = (in[i] == oldFill) ? NAN : a * in[i] + b;
        out[i] = (float)d;
    }
}

/* ------------------------------------------------------------------------------------------
 * src/MathUtils.cc:30-46 -- round half away from zero, then range test.  The reference converts
 * the rounded double to size_t first (negative -> huge -> invalid, -0.0 -> 0 -> valid, NaN -> invalid
 * on x86-64); restated as a comparison on the rounded double.
 * ------------------------------------------------------------------------------------------ */
size_t orc_round_and_clamp(double d, size_t mini, size_t maxi, size_t invalid)
{
    const double r = round(d);
    if (!(r >= 0.0) || r > 9.0e18)
        return invalid;
    const size_t u = (size_t)r;
    return (u >= mini && u <= maxi) ? u : invalid;
}

/* ------------------------------------------------------------------------------------------
 * src/CachedInterpolation.cc:149-171, 222-253 -- bilinear/bicubic cache with reduced domain.
 * ------------------------------------------------------------------------------------------ */
static long long orc_clamp_ex(long long low, double v, long long extend, long long high)
{
    long long value = (long long)(extend > 0 ? ceil(v) : floor(v)) + extend;
    if (value < low)
        value = low;
    if (value > high)
        value = high;
    return value;
}

orc_cached_interp* orc_cached_interp_create(int method, const double* px, const double* py, size_t inX, size_t inY, size_t outX, size_t outY)
{
    if (method != ORC_INTERPOL_BILINEAR && method != ORC_INTERPOL_BICUBIC)
        return NULL;
    orc_cached_interp* c = (orc_cached_interp*)calloc(1, sizeof(*c));
    const size_t n = outX * outY;
    c->method = method;
    c->inX = inX; c->inY = inY; c->outX = outX; c->outY = outY;
    c->px = (double*)malloc(n * sizeof(double));
    c->py = (double*)malloc(n * sizeof(double));
    memcpy(c->px, px, n * sizeof(double));
    memcpy(c->py, py, n * sizeof(double));
    if (n == 0)
        return c;
    double xlo = px[0], xhi = px[0], ylo = py[0], yhi = py[0];
    for (size_t i = 1; i < n; ++i) { /* std::min_element / max_element semantics (strict <) */
        if (px[i] < xlo) xlo = px[i];
        if (xhi < px[i]) xhi = px[i];
        if (py[i] < ylo) ylo = py[i];
        if (yhi < py[i]) yhi = py[i];
    }
    const long long EXT = 2;
    const long long minX = orc_clamp_ex(0, xlo, -EXT, (long long)inX - 1);
    const long long minY = orc_clamp_ex(0, ylo, -EXT, (long long)inY - 1);
    const long long maxX = orc_clamp_ex(0, xhi, +EXT, (long long)inX - 1);
    const long long maxY = orc_clamp_ex(0, yhi, +EXT, (long long)inY - 1);
    if ((maxX - minX) < 1 || (maxY - minY) < 1)
        return c;
    for (size_t i = 0; i < n; ++i) {
        c->px[i] -= minX;
        c->py[i] -= minY;
    }
    c->reduced = 1;
    c->xMin = (size_t)minX;
    c->yMin = (size_t)minY;
    c->inX = (size_t)(maxX - minX + 1);
    c->inY = (size_t)(maxY - minY + 1);
    return c;
}

int orc_cached_interp_apply(const orc_cached_interp* c, const float* in, size_t size, float* out)
{
    const size_t layer = c->outX * c->outY;
    const size_t nz = size / (c->inX * c->inY);
    float* zv = (float*)malloc((nz ? nz : 1) * sizeof(float));
    for (size_t xy = 0; xy < layer; ++xy) {
        if (c->method == ORC_INTERPOL_BILINEAR)
            orc_get_values_bilinear_f(in, zv, c->px[xy], c->py[xy], (int)c->inX, (int)c->inY, (int)nz);
        else
            orc_get_values_bicubic_f(in, zv, c->px[xy], c->py[xy], (int)c->inX, (int)c->inY, (int)nz);
        for (size_t z = 0; z < nz; ++z)
            out[z * layer + xy] = zv[z];
    }
    free(zv);
    return ORC_OK;
}

void orc_cached_interp_destroy(orc_cached_interp* c)
{
    if (!c)
        return;
    free(c->px);
    free(c->py);
    free(c);
}

/* ------------------------------------------------------------------------------------------
 * src/CachedInterpolation.cc:262-337 -- nearest-neighbour cache.
 * ------------------------------------------------------------------------------------------ */
orc_cached_nn* orc_cached_nn_create(const double* px, const double* py, size_t inX, size_t inY, size_t outX, size_t outY)
{
    orc_cached_nn* c = (orc_cached_nn*)calloc(1, sizeof(*c));
    const size_t n = outX * outY;
    c->inX = inX; c->inY = inY; c->outX = outX; c->outY = outY;
    c->pointsInIn = (size_t*)malloc((n ? n : 1) * sizeof(size_t));
    size_t minInX = inX, maxInX = 0, minInY = inY, maxInY = 0;
    for (size_t o = 0; o < n; ++o) {
        const size_t ix = orc_round_and_clamp(px[o], 0, inX - 1, ORC_INVALID_IDX);
        const size_t iy = orc_round_and_clamp(py[o], 0, inY - 1, ORC_INVALID_IDX);
        if (ix != ORC_INVALID_IDX && iy != ORC_INVALID_IDX) {
            c->pointsInIn[o] = iy * inX + ix;
            if (iy > maxInY) maxInY = iy;
            if (iy < minInY) minInY = iy;
            if (ix > maxInX) maxInX = ix;
            if (ix < minInX) minInX = ix;
        } else {
            c->pointsInIn[o] = ORC_INVALID_IDX;
        }
    }
    const size_t EXT = 2;
    minInX = (minInX > EXT) ? minInX - EXT : 0;
    minInY = (minInY > EXT) ? minInY - EXT : 0;
    maxInX = (inX - 1 < maxInX + EXT) ? inX - 1 : maxInX + EXT;
    maxInY = (inY - 1 < maxInY + EXT) ? inY - 1 : maxInY + EXT;
    if ((minInX > 0 || minInY > 0 || maxInX < inX - 1 || maxInY < inY - 1) && (minInX + 2 * EXT <= maxInX) && (minInY + 2 * EXT <= maxInY)) {
        const size_t redX = maxInX - minInX + 1, redY = maxInY - minInY + 1;
        for (size_t o = 0; o < n; ++o) {
            size_t i = c->pointsInIn[o];
            if (i != ORC_INVALID_IDX) {
                const size_t iy = i / inX - minInY, ix = i % inX - minInX;
                c->pointsInIn[o] = iy * redX + ix;
            }
        }
        c->reduced = 1;
        c->xMin = minInX;
        c->yMin = minInY;
        c->inX = redX;
        c->inY = redY;
    }
    return c;
}

int orc_cached_nn_apply(const orc_cached_nn* c, const float* in, size_t size, float* out)
{
    const size_t olayer = c->outX * c->outY, ilayer = c->inX * c->inY;
    const size_t nz = size / ilayer;
    for (size_t z = 0; z < nz; ++z) {
        const float* iz = in + z * ilayer;
        float* oz = out + z * olayer;
        for (size_t o = 0; o < olayer; ++o) {
            const size_t i = c->pointsInIn[o];
            oz[o] = (i != ORC_INVALID_IDX) ? iz[i] : orc_nanf();
        }
    }
    return ORC_OK;
}

void orc_cached_nn_destroy(orc_cached_nn* c)
{
    if (!c)
        return;
    free(c->pointsInIn);
    free(c);
}

/* ------------------------------------------------------------------------------------------
 * src/CachedForwardInterpolation.cc:133-224 -- forward (scatter/aggregate) cache.
 * The reference keeps a vector<size_t> per target cell, filled while walking the source row-major,
 * so every list is ascending in source index; a CSR built by counting sort preserves that order.
 * ------------------------------------------------------------------------------------------ */
orc_cached_fwd* orc_cached_fwd_create(int method, const double* pOnX, const double* pOnY, size_t inX, size_t inY, size_t outX, size_t outY)
{
    if (method < ORC_INTERPOL_FORWARD_SUM || method > ORC_INTERPOL_FORWARD_UNDEF_MIN)
        return NULL;
    orc_cached_fwd* c = (orc_cached_fwd*)calloc(1, sizeof(*c));
    const size_t nout = outX * outY, nin = inX * inY;
    c->method = method;
    c->undefAggr = method >= ORC_INTERPOL_FORWARD_UNDEF_SUM;
    c->inX = inX; c->inY = inY; c->outX = outX; c->outY = outY;
    size_t* target = (size_t*)malloc((nin ? nin : 1) * sizeof(size_t));
    c->segStart = (size_t*)calloc(nout + 1, sizeof(size_t));
    size_t minInX = 0, maxInX = 0, minInY = 0, maxInY = 0, nvalid = 0;
    for (size_t iy = 0; iy < inY; ++iy) {
        for (size_t ix = 0; ix < inX; ++ix) {
            const size_t i = iy * inX + ix;
            const size_t px = orc_round_and_clamp(pOnX[i], 0, outX - 1, ORC_INVALID_IDX);
            const size_t py = orc_round_and_clamp(pOnY[i], 0, outY - 1, ORC_INVALID_IDX);
            target[i] = ORC_INVALID_IDX;
            if (px != ORC_INVALID_IDX && py != ORC_INVALID_IDX) {
                if (nvalid == 0) { /* reference keys on maxPointsInIn == 0, equivalent: first valid point */
                    minInX = maxInX = ix;
                    minInY = maxInY = iy;
                } else {
                    if (iy < minInY) minInY = iy; else if (maxInY < iy) maxInY = iy;
                    if (ix < minInX) minInX = ix; else if (maxInX < ix) maxInX = ix;
                }
                target[i] = py * outX + px;
                c->segStart[target[i] + 1]++;
                nvalid++;
            }
        }
    }
    for (size_t o = 0; o < nout; ++o) {
        if (c->segStart[o + 1] > c->maxPointsInIn)
            c->maxPointsInIn = c->segStart[o + 1];
        c->segStart[o + 1] += c->segStart[o];
    }
    c->segIdx = (size_t*)malloc((nvalid ? nvalid : 1) * sizeof(size_t));
    size_t* cursor = (size_t*)malloc((nout ? nout : 1) * sizeof(size_t));
    memcpy(cursor, c->segStart, nout * sizeof(size_t));
    for (size_t i = 0; i < nin; ++i)
        if (target[i] != ORC_INVALID_IDX)
            c->segIdx[cursor[target[i]]++] = i;
    free(cursor);
    free(target);

    const size_t EXT = 2;
    minInX = (minInX > EXT) ? minInX - EXT : 0;
    minInY = (minInY > EXT) ? minInY - EXT : 0;
    maxInX = (inX - 1 < maxInX + EXT) ? inX - 1 : maxInX + EXT;
    maxInY = (inY - 1 < maxInY + EXT) ? inY - 1 : maxInY + EXT;
    if ((minInX > 0 || minInY > 0 || maxInX < inX - 1 || maxInY < inY - 1) && (minInX + 2 * EXT <= maxInX) && (minInY + 2 * EXT <= maxInY)) {
        const size_t redX = maxInX - minInX + 1, redY = maxInY - minInY + 1;
        for (size_t k = 0; k < nvalid; ++k) {
            const size_t i = c->segIdx[k];
            const size_t iy = i / inX - minInY, ix = i % inX - minInX;
            c->segIdx[k] = iy * redX + ix;
        }
        c->reduced = 1;
        c->xMin = minInX;
        c->yMin = minInY;
        c->inX = redX;
        c->inY = redY;
    }
    return c;
}

static int orc_cmp_float(const void* a, const void* b)
{
    const float fa = *(const float*)a, fb = *(const float*)b;
    return (fa > fb) - (fa < fb);
}

int orc_cached_fwd_apply(const orc_cached_fwd* c, const float* in, size_t size, float* out, int level_bug)
{
    const size_t olayer = c->outX * c->outY, ilayer = c->inX * c->inY;
    const size_t nz = size / ilayer;
    int kind; /* 0 sum, 1 mean, 2 median, 3 max, 4 min */
    switch (c->method) {
    case ORC_INTERPOL_FORWARD_SUM: case ORC_INTERPOL_FORWARD_UNDEF_SUM: kind = 0; break;
    case ORC_INTERPOL_FORWARD_MEAN: case ORC_INTERPOL_FORWARD_UNDEF_MEAN: kind = 1; break;
    case ORC_INTERPOL_FORWARD_MEDIAN: case ORC_INTERPOL_FORWARD_UNDEF_MEDIAN: kind = 2; break;
    case ORC_INTERPOL_FORWARD_MAX: case ORC_INTERPOL_FORWARD_UNDEF_MAX: kind = 3; break;
    default: kind = 4; break;
    }
    float* buf = (float*)malloc((c->maxPointsInIn ? c->maxPointsInIn : 1) * sizeof(float));
    for (size_t z = 0; z < nz; ++z) {
        const float* src = level_bug ? in : in + z * ilayer;
        for (size_t o = 0; o < olayer; ++o) {
            size_t count = 0;
            float val = 0;
            int has_nan = 0;
            for (size_t k = c->segStart[o]; k < c->segStart[o + 1]; ++k) {
                const float f = src[c->segIdx[k]];
                if (!c->undefAggr && isnan(f))
                    continue;
                switch (kind) {
                case 0: case 1: val += f; break;                               /* AggSum::push :79-86 */
                case 2: buf[count] = f; has_nan |= isnan(f); break;            /* AggMedian::push */
                case 3: if (count == 0 || f > val) val = f; break;             /* AggMax::push :99-106 */
                default: if (count == 0 || f < val) val = f; break;            /* AggMin::push :89-96 */
                }
                count++;
            }
            float r;
            if (count == 0) {
                r = orc_nanf();
            } else if (kind == 1) {
                r = val / (float)count; /* float / size_t -> float division (:90) */
            } else if (kind == 2) {
                /* nth_element at n/2 == element n/2 of the sorted list (upper median). With NaNs among
                 * the values (undef variant) the reference's comparison order is unspecified: defined as NaN. */
                if (has_nan) {
                    r = orc_nanf();
                } else {
                    qsort(buf, count, sizeof(float), orc_cmp_float);
                    r = buf[count / 2];
                }
            } else {
                r = val;
            }
            out[z * olayer + o] = r;
        }
    }
    free(buf);
    return ORC_OK;
}

void orc_cached_fwd_destroy(orc_cached_fwd* c)
{
    if (!c)
        return;
    free(c->segStart);
    free(c->segIdx);
    free(c);
}

/* ------------------------------------------------------------------------------------------
 * src/reproject.cc:488-530 -- rotation by the cached matrix (stride 3: cos, sin, angle in degree).
 * ------------------------------------------------------------------------------------------ */
void orc_vector_reproject_values_by_matrix_f(const double* m, size_t layer, float* u, float* v, size_t nz)
{
    for (size_t z = 0; z < nz; ++z) {
        float* uz = u + z * layer;
        float* vz = v + z * layer;
        for (size_t i = 0; i < layer; ++i) {
            const double c = m[3 * i], s = m[3 * i + 1];
            const double un = uz[i] * c - vz[i] * s;
            const double vn = uz[i] * s + vz[i] * c;
            uz[i] = (float)un;
            vz[i] = (float)vn;
        }
    }
}

void orc_vector_reproject_direction_by_matrix_f(const double* m, size_t layer, float* angle, size_t nz)
{
    for (size_t z = 0; z < nz; ++z) {
        float* az = angle + z * layer;
        for (size_t i = 0; i < layer; ++i) {
            double a = az[i] - m[3 * i + 2];
            if (a < 0)
                a += 360;
            if (a > 360)
                a -= 360;
            az[i] = (float)a;
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * include/fimex/FindNeighborElements.h -- value_type is float (leap_iterator<const float*>).
 * ------------------------------------------------------------------------------------------ */
void orc_find_closest_distinct_elements(const float* col, size_t n, ptrdiff_t stride, float x, size_t* first, size_t* second)
{
    size_t r1 = 0, r2 = 0;
    if (n == 0) {
        *first = *second = 0;
        return;
    }
    float v1 = col[0];
    float d1 = fabsf(x - col[0]);
    float d2 = d1;
    for (size_t k = 0; k < n; ++k) {
        const float cur = col[(ptrdiff_t)k * stride];
        const float d = fabsf(x - cur);
        if (d <= d2) {
            if (d < d1) {
                r2 = r1;
                d2 = d1;
                v1 = cur;
                r1 = k;
                d1 = d;
            } else if (cur != v1) {
                r2 = k;
                d2 = d;
            }
        }
    }
    *first = r1;
    *second = r2;
}

void orc_find_closest_neighbor_distinct_elements(const float* col, size_t n, ptrdiff_t stride, float x, size_t* first, size_t* second)
{
    if (n == 0) {
        *first = *second = 0;
        return;
    }
    size_t lowest = 0, highest = 0;
    float lowDiff = x - col[0];
    float highDiff = -lowDiff;
    if (lowDiff < 0)
        lowDiff = FLT_MAX;
    if (highDiff < 0)
        highDiff = FLT_MAX;
    for (size_t k = 1; k < n; ++k) {
        const float diff = x - col[(ptrdiff_t)k * stride];
        if (diff >= 0) {
            if (diff < lowDiff) {
                lowDiff = diff;
                lowest = k;
            }
        } else {
            if (-diff < highDiff) {
                highDiff = -diff;
                highest = k;
            }
        }
    }
    if (lowDiff == FLT_MAX || highDiff == FLT_MAX) {
        orc_find_closest_distinct_elements(col, n, stride, x, first, second);
        return;
    }
    *first = lowest;
    *second = highest;
}

/* ------------------------------------------------------------------------------------------
 * src/vertical_coordinate_transformations.c:29-77
 * ------------------------------------------------------------------------------------------ */
void orc_atmosphere_ln_pressure(size_t n, double p0, const double* lev, double* p)
{
    for (size_t i = 0; i < n; ++i)
        p[i] = p0 * exp(-lev[i]);
}

void orc_atmosphere_sigma_pressure(size_t n, double ptop, double ps, const double* sigma, double* p)
{
    const double d = ps - ptop;
    for (size_t i = 0; i < n; ++i)
        p[i] = ptop + sigma[i] * d;
}

void orc_atmosphere_hybrid_sigma_pressure(size_t n, double p0, double ps, const double* a, const double* b, double* p)
{
    for (size_t i = 0; i < n; ++i)
        p[i] = (a[i] * p0) + (b[i] * ps);
}

void orc_atmosphere_hybrid_sigma_ap_pressure(size_t n, double ps, const double* ap, const double* b, double* p)
{
    for (size_t i = 0; i < n; ++i)
        p[i] = ap[i] + (b[i] * ps);
}

void orc_atmosphere_hybrid_height(size_t n, double orog, const double* a, const double* b, double* h)
{
    for (size_t i = 0; i < n; ++i)
        h[i] = a[i] + (b[i] * orog);
}

/* ------------------------------------------------------------------------------------------
 * src/CDMVerticalInterpolator.cc:488-561
 * ------------------------------------------------------------------------------------------ */
int orc_vertical_interpolate_f(int method, const float* in, const float* vcoord_in, size_t nzi, size_t ncol, const double* levels,
                               const float* vcoord_out, size_t nzo, const double* vmin, const double* vmax, double valid_min,
                               double valid_max, float* out)
{
    int (*fn)(const float*, const float*, float*, size_t, double, double, double);
    switch (method) {
    case ORC_VINT_METHOD_LIN: fn = orc_get_values_linear_f; break;
    case ORC_VINT_METHOD_LIN_WEAK_EXTRA: fn = orc_get_values_linear_weak_extrapol_f; break;
    case ORC_VINT_METHOD_LIN_NO_EXTRA: fn = orc_get_values_linear_no_extrapol_f; break;
    case ORC_VINT_METHOD_LIN_CONST_EXTRA: fn = orc_get_values_linear_const_extrapol_f; break;
    case ORC_VINT_METHOD_LOG: fn = orc_get_values_log_f; break;
    case ORC_VINT_METHOD_LOGLOG: fn = orc_get_values_log_log_f; break;
    case ORC_VINT_METHOD_NN: fn = orc_get_values_nearest_f; break;
    default: return ORC_ERROR;
    }
    for (size_t k = 0; k < nzo; ++k) {
        for (size_t c = 0; c < ncol; ++c) {
            const double vout = vcoord_out ? (double)vcoord_out[k * ncol + c] : levels[k];
            float* dst = out + k * ncol + c;
            int range = 1;
            if (vmin && vmax)
                range = (vout >= vmin[c]) && (vout <= vmax[c]);
            else if (vmin)
                range = (vout >= vmin[c]);
            else if (vmax)
                range = (vout <= vmax[c]);
            *dst = orc_nanf(); /* reference leaves this unwritten when the log kernels refuse (see header) */
            if (!range)
                continue;
            size_t p0, p1;
            orc_find_closest_neighbor_distinct_elements(vcoord_in + c, nzi, (ptrdiff_t)ncol, (float)vout, &p0, &p1);
            if (p0 == p1)
                continue;
            const float a = in[p0 * ncol + c], b = in[p1 * ncol + c];
            const float va = vcoord_in[p0 * ncol + c], vb = vcoord_in[p1 * ncol + c];
            fn(&a, &b, dst, 1, va, vb, vout);
        }
    }
    const size_t total = nzo * ncol;
    if (!isnan(valid_min)) {
        const float lo = (float)valid_min;
        for (size_t i = 0; i < total; ++i)
            if (out[i] < lo)
                out[i] = lo;
    }
    if (!isnan(valid_max)) {
        const float hi = (float)valid_max;
        for (size_t i = 0; i < total; ++i)
            if (out[i] > hi)
                out[i] = hi;
    }
    return ORC_OK;
}

/* ------------------------------------------------------------------------------------------
 * src/CDMInterpolator.cc:953-1028 -- coord_kdtree without the tree (see header).
 * ------------------------------------------------------------------------------------------ */
#define ORC_D2R 0.017453292519943296 /* deg_to_rad<double>, MathUtils.h:158-162 */

int orc_kdtree_translate(double maxDist_m, double* px, double* py, size_t nPoints, const double* lon, const double* lat, size_t nx, size_t ny,
                         double* gap)
{
    const size_t n = nx * ny;
    double* c = (double*)malloc((n ? n : 1) * 3 * sizeof(double));
    for (size_t i = 0; i < n; ++i) {
        if (!(isnan(lat[i]) || isnan(lon[i]))) {
            const double la = lat[i] * ORC_D2R, lo = lon[i] * ORC_D2R;
            c[3 * i] = cos(la) * cos(lo);
            c[3 * i + 1] = cos(la) * sin(lo);
            c[3 * i + 2] = sin(la);
        } else {
            c[3 * i] = c[3 * i + 1] = c[3 * i + 2] = nan("");
        }
    }
    const double r = maxDist_m / 6371000.;
    const double r2 = r * r;
    for (size_t q = 0; q < nPoints; ++q) {
        const double la = py[q] * ORC_D2R, lo = px[q] * ORC_D2R;
        const double q0 = cos(la) * cos(lo), q1 = cos(la) * sin(lo), q2 = sin(la);
        double best = HUGE_VAL, second = HUGE_VAL;
        size_t bi = (size_t)-1;
        for (size_t i = 0; i < n; ++i) {
            const double d0 = q0 - c[3 * i], d1 = q1 - c[3 * i + 1], d2 = q2 - c[3 * i + 2];
            double d = 0;
            d += d0 * d0;
            d += d1 * d1;
            d += d2 * d2;
            if (!(d < r2))
                continue;
            if (d < best) {
                second = best;
                best = d;
                bi = i;
            } else if (d < second) {
                second = d;
            }
        }
        if (gap)
            gap[q] = second - best;
        if (bi != (size_t)-1) {
            px[q] = (double)(bi % nx);
            py[q] = (double)(bi / nx);
        } else {
            px[q] = -1000;
            py[q] = -1000;
        }
    }
    free(c);
    return ORC_OK;
}

/* src/CDMInterpolator.cc:1030-1090 */
double orc_get_grid_distance(const double* lon, const double* lat, size_t nx, size_t ny)
{
    const size_t n = nx * ny;
    int steps;
    size_t stepSize;
    if (n > 1000) {
        steps = 53;
        stepSize = n / steps;
    } else {
        stepSize = 1;
        steps = (int)n;
    }
    double min_sample = 2;
    for (int ik = 0; ik < steps; ++ik) {
        const size_t sp = ik * stepSize;
        const double lon0 = lon[sp] * ORC_D2R, lat0 = lat[sp] * ORC_D2R;
        if (isnan(lon0) || isnan(lat0))
            continue;
        double best = -2;
        for (size_t pos = 0; pos < n; ++pos) {
            if (pos == sp)
                continue;
            const double lon1 = lon[pos] * ORC_D2R, lat1 = lat[pos] * ORC_D2R;
            if (isnan(lon1) || isnan(lat1))
                continue;
            const double dlon = lon0 - lon1;
            const double cd = cos(lat0) * cos(lat1) * cos(dlon) + sin(lat0) * sin(lat1);
            if (cd > best)
                best = cd;
        }
        if (best < min_sample)
            min_sample = best;
    }
    const double d = acos(min_sample) * 1.414;
    return d < 3.1415926535897932384626433832795 ? d : 3.1415926535897932384626433832795;
}

typedef struct { double lat, lon, x, y; } orc_llpoint;

static int orc_cmp_ll(const void* a, const void* b)
{
    const double la = ((const orc_llpoint*)a)->lat, lb = ((const orc_llpoint*)b)->lat;
    return (la > lb) - (la < lb);
}

/* src/CDMInterpolator.cc:1112-1190 */
int orc_coordnn_translate(double* px, double* py, size_t nPoints, const double* lon, const double* lat, size_t nx, size_t ny)
{
    const double max_grid_d = orc_get_grid_distance(lon, lat, nx, ny);
    const double min_grid_cos_d = cos(max_grid_d);
    orc_llpoint* ll = (orc_llpoint*)malloc((nx * ny != 0 ? nx * ny : 1) * sizeof(orc_llpoint));
    size_t m = 0;
    for (size_t ix = 0; ix < nx; ++ix)
        for (size_t iy = 0; iy < ny; ++iy) {
            const size_t pos = ix + iy * nx;
            if (!(isnan(lon[pos]) || isnan(lat[pos]))) {
                ll[m].lat = lat[pos] * ORC_D2R;
                ll[m].lon = lon[pos] * ORC_D2R;
                ll[m].x = (double)ix;
                ll[m].y = (double)iy;
                ++m;
            }
        }
    qsort(ll, m, sizeof(orc_llpoint), orc_cmp_ll);
    for (size_t q = 0; q < nPoints; ++q) {
        const double plat = py[q] * ORC_D2R, plon = px[q] * ORC_D2R;
        double rx = -1., ry = -1.;
        double min_cos_d = min_grid_cos_d, min_d = acos(min_cos_d);
        size_t lo = 0, hi = m; /* lower_bound on latitude */
        while (lo < hi) {
            const size_t mid = (lo + hi) / 2;
            if (ll[mid].lat < plat)
                lo = mid + 1;
            else
                hi = mid;
        }
        for (size_t it = lo; it < m; ++it) {
            if (fabs(ll[it].lat - plat) > min_d)
                break;
            const double dlon = ll[it].lon - plon;
            const double cd = cos(ll[it].lat) * cos(plat) * cos(dlon) + sin(ll[it].lat) * sin(plat);
            if (cd > min_cos_d) {
                min_cos_d = cd;
                min_d = acos(min_cos_d);
                rx = ll[it].x;
                ry = ll[it].y;
            }
        }
        for (size_t it = lo; it-- > 0;) {
            if (fabs(ll[it].lat - plat) > min_d)
                break;
            const double dlon = ll[it].lon - plon;
            const double cd = cos(ll[it].lat) * cos(plat) * cos(dlon) + sin(ll[it].lat) * sin(plat);
            if (cd > min_cos_d) {
                min_cos_d = cd;
                min_d = acos(min_cos_d);
                rx = ll[it].x;
                ry = ll[it].y;
            }
        }
        px[q] = rx;
        py[q] = ry;
    }
    free(ll);
    return ORC_OK;
}

/* ---- vertical coordinate converters ---------------------------------------------------------------------------------------- */
#define ORC_GAS_CONSTANT 8.31432        /* mifi_constants.h:241 */
#define ORC_EARTH_GRAVITY 9.80665       /* :235 */
#define ORC_MOLAR_MASS_DRY_AIR 0.0289644 /* :238 */
static const double ORC_BAROMETRIC_FACTOR = ORC_GAS_CONSTANT / (ORC_EARTH_GRAVITY * ORC_MOLAR_MASS_DRY_AIR); /* v_c_t.c:80-81 */
static const double ORC_Z_MOL_WEIGHT_RATIO = .60771704180064308681;                                          /* :84 */

float orc_virtual_temperature(float spec_humidity, float T) { return (1 + ORC_Z_MOL_WEIGHT_RATIO * spec_humidity) * T; }

float orc_barometric_layer_thickness(float p_low_alti, float p_high_alti, float T)
{
    return log(p_low_alti / p_high_alti) * T * ORC_BAROMETRIC_FACTOR;
}

static double orc_vpost(int post, double v)
{
    if (post == ORC_VPOST_STANDARD_ALTITUDE) { /* mifi_barometric_height(n, 1013.25, p, 288.15, .), :101-113 */
        const double K = -ORC_BAROMETRIC_FACTOR * 288.15;
        return K * log(v / 1013.25);
    }
    if (post == ORC_VPOST_STANDARD_PRESSURE) { /* mifi_barometric_pressure(n, 1013.25, h, 288.15, .), :86-99 */
        const double Cc = -1 / (ORC_BAROMETRIC_FACTOR * 288.15);
        return 1013.25 * exp(Cc * v);
    }
    return v;
}

int orc_vcoord_levels_f(int kind, const double* a, const double* b, double scalar, const float* surf, size_t nOuterS, double surf_scale,
                        const float* eta, int post, const float* topo, double topo_factor, size_t nOuter, size_t nz, size_t nInner, float* out)
{
    size_t o, k, i;
    if (kind < 0 || kind > ORC_VCOORD_OCEAN_S_G2 || !(nOuterS == 1 || nOuterS == nOuter))
        return 0;
    for (o = 0; o < nOuter; ++o)
        for (k = 0; k < nz; ++k)
            for (i = 0; i < nInner; ++i) {
                const size_t si = (nOuterS == 1 ? 0 : o) * nInner + i;
                const double s = surf ? surf_scale * surf[si] : 0.; /* ScaleValue: oldScale * in (+ 0), DataUtils.h:57 */
                double v;
                switch (kind) {
                case ORC_VCOORD_HYBRID_SIGMA_AP: v = a[k] + (b[k] * s); break;            /* v_c_t.c:64-70 */
                case ORC_VCOORD_HYBRID_SIGMA: v = (a[k] * scalar) + (b[k] * s); break;   /* :56-62 */
                case ORC_VCOORD_SIGMA: { const double pDiff = (s - scalar); v = scalar + a[k] * pDiff; break; } /* :37-44 */
                case ORC_VCOORD_LN_PRESSURE: v = scalar * exp(-a[k]); break;              /* :29-35 */
                case ORC_VCOORD_HYBRID_HEIGHT: v = a[k] + (b[k] * s); break;              /* :72-78 */
                case ORC_VCOORD_OCEAN_S_G1: {                                            /* :166-174 */
                    const double zeta = eta ? (double)eta[si] : 0., h_inv = 1 / s;
                    const double S = scalar * a[k] + (s - scalar) * b[k];
                    v = -(S + zeta * (1 + S * h_inv)); /* positive down, OceanSCoordinateGToDepthConverter.cc:100-101 */
                    break;
                }
                default: {                                                               /* :175-183 */
                    const double zeta = eta ? (double)eta[si] : 0., hph_c_inv = 1 / (s + scalar);
                    const double S = hph_c_inv * (scalar * a[k] + s * b[k]);
                    v = -(zeta + (zeta + s) * S);
                    break;
                }
                }
                v = orc_vpost(post, v);
                if (topo)
                    v = v + topo_factor * topo[i]; /* AltitudeHeightConverter.cc:125-127 */
                out[(o * nz + k) * nInner + i] = (float)v;
            }
    return 1;
}

void orc_vcoord_transform_f(int post, const float* in, size_t n, float* out)
{
    size_t i;
    for (i = 0; i < n; ++i)
        out[i] = (float)orc_vpost(post, (double)in[i]);
}

int orc_vcoord_pressure_integration_f(const float* p, const float* sap, double sap_scale, const float* sgp, double sgp_scale, const float* airt,
                                      const float* shum, size_t nOuter, size_t nOuterS, size_t nz, size_t nInner, int start_high_p,
                                      const float* topo, double topo_factor, float* out)
{
    size_t o, i, n;
    if (!(nOuterS == 1 || nOuterS == nOuter) || nz == 0)
        return 0;
    if (start_high_p < 0) { /* PressureIntegrationToAltitudeConverter.cc:141-164: first column with positive ends decides */
        int have = 0;
        for (o = 0; o < nOuter && !have; ++o)
            for (i = 0; i < nInner && !have; ++i) {
                const float p_first = p[(o * nz) * nInner + i], p_last = p[(o * nz + nz - 1) * nInner + i];
                if (p_first > 0 && p_last > 0) {
                    start_high_p = (p_first > p_last);
                    have = 1;
                }
            }
        if (!have)
            return 0;
    }
    for (o = 0; o < nOuter; ++o)
        for (i = 0; i < nInner; ++i) { /* :225-246 */
            const size_t si = (nOuterS == 1 ? 0 : o) * nInner + i;
            const float sgp_v = (float)(sgp_scale * sgp[si]), sap_v = (float)(sap_scale * sap[si]); /* VarFloat: asFloat of the scaled double */
            double a = sgp_v / ORC_EARTH_GRAVITY;
            float p_low_alti = sap_v;
            for (n = 0; n < nz; ++n) {
                const size_t iz = start_high_p ? n : nz - 1 - n;
                const size_t idx = (o * nz + iz) * nInner + i;
                const float p_high_alti = p[idx];
                float Tv = airt[idx];
                double lt, v;
                if (shum)
                    Tv = orc_virtual_temperature(shum[idx], Tv);
                lt = orc_barometric_layer_thickness(p_low_alti, p_high_alti, Tv);
                a += lt;
                v = a;
                if (topo)
                    v = v + topo_factor * topo[i];
                out[idx] = (float)v;
                p_low_alti = p_high_alti;
            }
        }
    return 1;
}

/* ---- packed integer slabs, generic over the CDM integer types ---- */
#define ORC_PACK_FUNCS(NAME, T)                                                                                          \
    static void orc_unpack_##NAME(const T* in, size_t n, double badValue, float* out)                                    \
    {                                                                                                                    \
        size_t i;                                                                                                        \
        for (i = 0; i < n; ++i)                                                                                          \
            out[i] = (float)in[i];                                                                                       \
        orc_bad2nanf(out, n, (float)badValue);                                                                           \
    }                                                                                                                    \
    static void orc_unpack_scaled_##NAME(const T* in, size_t n, double fill, double scale, double offset, float* out)    \
    {                                                                                                                    \
        const T oldFill = (T)fill;                                                                                       \
        const double a = scale / 1.0, b = (offset - 0.0) / 1.0;                                                          \
        size_t i;                                                                                                        \
        for (i = 0; i < n; ++i) {                                                                                        \
            const double d = (in[i] == oldFill) ? NAN : a * in[i] + b;                                                   \
            out[i] = (float)d;                                                                                           \
        }                                                                                                                \
    }                                                                                                                    \
    static void orc_pack_##NAME(const float* in, size_t n, double badValue, T* out)                                      \
    {                                                                                                                    \
        const T newFill = (T)badValue;                                                                                   \
        size_t i;                                                                                                        \
        for (i = 0; i < n; ++i) {                                                                                        \
            const float v = in[i];                                                                                       \
            if (v != v) {                                                                                                \
                out[i] = newFill;                                                                                        \
            } else {                                                                                                     \
                const int r = (int)lround(1.0 * v + 0.0);                                                                \
                out[i] = (T)r;                                                                                           \
            }                                                                                                            \
        }                                                                                                                \
    }
ORC_PACK_FUNCS(i8, signed char)
ORC_PACK_FUNCS(i16, short)
ORC_PACK_FUNCS(i32, int)
ORC_PACK_FUNCS(u8, unsigned char)
ORC_PACK_FUNCS(u16, unsigned short)
ORC_PACK_FUNCS(u32, unsigned int)

int orc_unpack_f(int dtype, const void* in, size_t n, double badValue, float* out)
{
    switch (dtype) {
    case 0: orc_unpack_i8((const signed char*)in, n, badValue, out); return 1;
    case 1: orc_unpack_i16((const short*)in, n, badValue, out); return 1;
    case 2: orc_unpack_i32((const int*)in, n, badValue, out); return 1;
    case 3: orc_unpack_u8((const unsigned char*)in, n, badValue, out); return 1;
    case 4: orc_unpack_u16((const unsigned short*)in, n, badValue, out); return 1;
    case 5: orc_unpack_u32((const unsigned int*)in, n, badValue, out); return 1;
    default: return 0;
    }
}

int orc_unpack_scaled_f(int dtype, const void* in, size_t n, double fill, double scale, double offset, float* out)
{
    switch (dtype) {
    case 0: orc_unpack_scaled_i8((const signed char*)in, n, fill, scale, offset, out); return 1;
    case 1: orc_unpack_scaled_i16((const short*)in, n, fill, scale, offset, out); return 1;
    case 2: orc_unpack_scaled_i32((const int*)in, n, fill, scale, offset, out); return 1;
    case 3: orc_unpack_scaled_u8((const unsigned char*)in, n, fill, scale, offset, out); return 1;
    case 4: orc_unpack_scaled_u16((const unsigned short*)in, n, fill, scale, offset, out); return 1;
    case 5: orc_unpack_scaled_u32((const unsigned int*)in, n, fill, scale, offset, out); return 1;
    default: return 0;
    }
}

int orc_pack_f(int dtype, const float* in, size_t n, double badValue, void* out)
{
    switch (dtype) {
    case 0: orc_pack_i8(in, n, badValue, (signed char*)out); return 1;
    case 1: orc_pack_i16(in, n, badValue, (short*)out); return 1;
    case 2: orc_pack_i32(in, n, badValue, (int*)out); return 1;
    case 3: orc_pack_u8(in, n, badValue, (unsigned char*)out); return 1;
    case 4: orc_pack_u16(in, n, badValue, (unsigned short*)out); return 1;
    case 5: orc_pack_u32(in, n, badValue, (unsigned int*)out); return 1;
    default: return 0;
    }
}

/* src/interpolation.c:548-678 */
int orc_fill2d_f(size_t nx, size_t ny, float* field, float relaxCrit, float corrEff, size_t maxLoop, size_t* nChanged)
{
    const size_t total = nx * ny;
    size_t i, x, y, n, nUnchanged;
    double sum = 0, average, stddev = 0, crit;
    float *w, *e;
    if (total == 0)
        return 1;
    *nChanged = 0;
    for (i = 0; i < total; ++i) { /* :556-563 */
        if (isnan(field[i]))
            (*nChanged)++;
        else
            sum += field[i];
    }
    nUnchanged = total - *nChanged;
    if (nUnchanged == 0 || *nChanged == 0)
        return 1;
    w = (float*)malloc(total * sizeof(float));
    e = (float*)malloc(total * sizeof(float));
    average = sum / nUnchanged; /* :580 */
    for (i = 0; i < total; ++i) { /* :586-596 */
        if (isnan(field[i])) {
            w[i] = 1.;
            field[i] = average;
        } else {
            stddev += fabs(field[i] - average);
            w[i] = 0.;
        }
    }
    stddev /= nUnchanged;
    crit = relaxCrit * stddev;
    for (y = 1; y + 1 < ny; y++) /* :608-612 */
        for (x = 1; x + 1 < nx; x++)
            w[y * nx + x] *= corrEff;
    for (n = 0; n < maxLoop; n++) { /* :621-669 */
        for (y = 1; y + 1 < ny; y++)
            for (x = 1; x + 1 < nx; x++) {
                float* f = &field[y * nx + x];
                e[y * nx + x] = (*(f + 1) + *(f - 1) + *(f + nx) + *(f - nx)) * 0.25 - *f;
                (*f) += e[y * nx + x] * w[y * nx + x];
            }
        if ((n < (maxLoop - 5)) && (n % 10 == 0)) { /* size_t arithmetic: maxLoop < 5 wraps, as in the reference */
            const float crtest = crit * corrEff;
            size_t nbad = 0;
            for (y = 1; y + 1 < ny && !nbad; y++)
                for (x = 1; x + 1 < nx; x++)
                    if (fabs(e[y * nx + x] * w[y * nx + x]) > crtest)
                        nbad = 1;
            if (nbad == 0) {
                free(e);
                free(w);
                return 1;
            }
        }
        for (y = 1; y + 1 < ny; y++) { /* :660-663 */
            field[y * nx + 0] += (field[y * nx + 1] - field[y * nx + 0]) * w[y * nx + 0];
            field[y * nx + (nx - 1)] += (field[y * nx + (nx - 2)] - field[y * nx + (nx - 1)]) * w[y * nx + (nx - 1)];
        }
        for (x = 0; x < nx; x++) { /* :664-667 */
            field[0 * nx + x] += (field[1 * nx + x] - field[0 * nx + x]) * w[0 * nx + x];
            field[(ny - 1) * nx + x] += (field[(ny - 2) * nx + x] - field[(ny - 1) * nx + x]) * w[(ny - 1) * nx + x];
        }
    }
    free(e);
    free(w);
    return 1;
}

/* src/interpolation.c:680-775 */
static int orc_creepfill_impl(size_t nx, size_t ny, float* field, float defaultVal, unsigned short repeat, char setWeight, size_t nChanged)
{
    const size_t total = nx * ny;
    size_t i, x, y, l, nUnchanged, changedInLoop;
    char* w;
    unsigned short* r;
    if (total == 0)
        return 1;
    nUnchanged = total - nChanged;
    if (nUnchanged == 0 || nChanged == 0)
        return 1;
    w = (char*)malloc(total);
    r = (unsigned short*)malloc(total * sizeof(unsigned short));
    for (i = 0; i < total; ++i) { /* :703-715 */
        if (isnan(field[i])) {
            w[i] = 0;
            r[i] = 0;
            field[i] = defaultVal;
        } else {
            w[i] = setWeight;
            r[i] = repeat;
        }
    }
    l = 0;
    changedInLoop = 1;
    while ((changedInLoop > 0) && (l < nUnchanged)) { /* :724-754 */
        changedInLoop = 0;
        l++;
        for (y = 1; y + 1 < ny; y++)
            for (x = 1; x + 1 < nx; x++) {
                float* f = &field[y * nx + x];
                unsigned short* rr = &r[y * nx + x];
                char* ww = &w[y * nx + x];
                if (*rr < repeat) {
                    size_t wFieldSum = *(ww + 1) + *(ww - 1) + *(ww + nx) + *(ww - nx);
                    if (wFieldSum != 0) {
                        (*f) += *(ww + 1) * *(f + 1) + *(ww - 1) * *(f - 1) + *(ww + nx) * *(f + nx) + *(ww - nx) * *(f - nx);
                        (*f) /= (1 + wFieldSum);
                        (*ww) = 1;
                        (*rr)++;
                        changedInLoop++;
                    }
                }
            }
    }
    for (l = 0; l < repeat; l++) { /* :756-781 */
        for (y = 1; y + 1 < ny; y++) {
            if (r[y * nx + 0] < repeat) {
                field[y * nx + 0] += field[y * nx + 1] * w[y * nx + 1];
                field[y * nx + 0] /= (1 + w[y * nx + 1]);
                w[y * nx + 0] = 1;
            }
            if (r[y * nx + (nx - 1)] < repeat) {
                field[y * nx + (nx - 1)] += field[y * nx + (nx - 2)] * w[y * nx + (nx - 2)];
                field[y * nx + (nx - 1)] /= (1 + w[y * nx + (nx - 2)]);
                w[y * nx + (nx - 1)] = 1;
            }
        }
        for (x = 0; x < nx; x++) {
            if (r[0 * nx + x] < repeat) {
                field[0 * nx + x] += field[1 * nx + x] * w[1 * nx + x];
                field[0 * nx + x] /= (1 + w[1 * nx + x]);
                w[0 * nx + x] = 1;
            }
            if (r[(ny - 1) * nx + x] < repeat) {
                field[(ny - 1) * nx + x] += field[(ny - 2) * nx + x] * w[(ny - 2) * nx + x];
                field[(ny - 1) * nx + x] /= (1 + w[(ny - 2) * nx + x]);
                w[(ny - 1) * nx + x] = 1;
            }
        }
    }
    free(r);
    free(w);
    return 1;
}

/* :777-797 */
int orc_creepfill2d_f(size_t nx, size_t ny, float* field, unsigned short repeat, char setWeight, size_t* nChanged)
{
    const size_t total = nx * ny;
    size_t i, nUnchanged;
    double sum = 0;
    float average;
    if (total == 0)
        return 1;
    *nChanged = 0;
    for (i = 0; i < total; ++i) {
        if (isnan(field[i]))
            (*nChanged)++;
        else
            sum += field[i];
    }
    nUnchanged = total - *nChanged;
    if (nUnchanged == 0)
        return 1;
    average = sum / nUnchanged;
    return orc_creepfill_impl(nx, ny, field, average, repeat, setWeight, *nChanged);
}

/* :799-810 */
int orc_creepfillval2d_f(size_t nx, size_t ny, float* field, float defaultVal, unsigned short repeat, char setWeight, size_t* nChanged)
{
    const size_t total = nx * ny;
    size_t i;
    if (total == 0)
        return 1;
    *nChanged = 0;
    for (i = 0; i < total; ++i)
        if (isnan(field[i]))
            (*nChanged)++;
    return orc_creepfill_impl(nx, ny, field, defaultVal, repeat, setWeight, *nChanged);
}
