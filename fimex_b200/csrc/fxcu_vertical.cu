// fxcu_vertical.cu -- vertical interpolation (K14 + K15) and the two-level 1-D kernels.
//
// Reference behaviour restated here (fimex 2.4.1):
//   CDMVerticalInterpolator::getLevelDataSlice, hot loop        src/CDMVerticalInterpolator.cc:488-561
//   find_closest_neighbor_distinct_elements / _distinct_elements include/fimex/FindNeighborElements.h:46-125 (value_type float)
//   mifi_get_values_{nearest,linear*,log,log_log}_f             src/interpolation.c:416-546
//   mifi_atmosphere_hybrid_sigma_ap_pressure                    src/vertical_coordinate_transformations.c:62-68
//   HybridSigmaApToPressureConverter::getDataSlice              src/coordSys/verticalTransform/HybridSigmaApToPressureConverter.cc:71-100
//
// HBM layout: in / vcoord_in [nOuter][nzi][nInner] f32, out [nOuter][nzo][nInner] f32; a column is (outer, inner), consecutive
// threads own consecutive inner indices, so every level access of a warp is one coalesced 128 B line.  A block keeps the
// vertical coordinate of its columns in shared memory ([nzi][threads], conflict free) and scans it once per output level, as
// the reference does; for hybrid-sigma sources the coordinate is formed on the fly (p = ap + b*ps) and never exists in HBM.
// Algorithmic bytes per column: 4*nzi (values) + 4*nzi (coordinate) or 4 (ps) + 4*nzo.
//
// Rounding: every f32/f64 operation is a separately rounded _rn intrinsic like the reference's x86-64 code.  The log / log_log
// methods call the device's log(), which may differ from glibc's in the last bit of the f64 result; after the narrowing to the
// f32 factor this shows, rarely, as a 1-ulp difference of the weight (parity within 1e-5 relative; all other methods bit-exact).
#include "fxcu_common.h"
#include "fxcu_logtable.h"

#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

namespace fxcu {

constexpr int VT = 128; // threads (columns) per block

struct VertParams {
    int method;
    const float* in;
    const float* vcoord_in;   // nullptr -> hybrid pressure from ap, b, ps
    const double* ap;         // [nzi] device
    const double* b;          // [nzi] device
    const float* ps;          // [nOuterPs][nInner]
    size_t nOuterPs;
    double ps_scale;
    size_t nOuter, nzi, nInner, nzo;
    const double* levels;     // [nzo] device, or nullptr
    int levels_dir;           // fixed levels: +1 non-decreasing, -1 non-increasing (both for one level), 0 unordered or NaN
    int dir_hint;             // hybrid mode: +1 / -1 when every column of the launch is provably strictly monotone, 0 = check per column
    const double* xlog;       // [nzo] device: host-computed log / log-log of the fixed levels (LOG / LOGLOG), or nullptr
    const float* vcoord_out;  // [nOuter][nzo][nInner] or nullptr
    const double* vmin;       // per column or nullptr
    const double* vmax;
    float valid_min, valid_max; // NaN = none (already narrowed like CDMVerticalInterpolator.cc:553,558)
    float bad, fill;
    float* out;
    const double2* logtab; // 128-entry logarithm table in device memory (fxcu_logtable.h), made once per device
    // rows of a padded slab (fimexcu_cache_padded_layout): of every rowPitch consecutive inner indices only the first rowWidth are
    // columns; the pad is neither read nor written.  0 / 0: every inner index is a column.  (nInner < 2^32 when set.)
    unsigned rowWidth, rowPitch;
};

__device__ __forceinline__ bool pad_column(const VertParams& P, size_t inner)
{
    return P.rowPitch != 0 && static_cast<unsigned>(inner) % P.rowPitch >= P.rowWidth;
}

// 1/d to ~1e-12 (one Newton step on the hardware's approximation) or to full f64 accuracy (two steps); d normal and positive
__device__ __forceinline__ double rcp_newton(double d, bool full)
{
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
    r = __fma_rn(r, __fma_rn(-d, r, 1.0), r);
    if (full)
        r = __fma_rn(r, __fma_rn(-d, r, 1.0), r);
    return r;
}

__constant__ double c_inv_odd[12] = {1., 1. / 3, 1. / 5, 1. / 7, 1. / 9, 1. / 11, 1. / 13, 1. / 15, 1. / 17, 1. / 19, 1. / 21, 1. / 23};

// The weight of the LOG method, f = (log x - log a) / (log b - log a) narrowed to f32 (interpolation.c:521-535), without a
// logarithm: log(x/a) = 2 atanh(s), s = (x - a)/(x + a), as a series that is exact to < 2e-17 relative for |s| < 0.2.  The
// quotient of the two series is accurate to ~1e-15, an order better than the reference's own difference of rounded logarithms,
// whose error bound (2^-52 * max|log| * (1 + |f|) / |log b - log a|) is the tolerance below: the fast value is returned only if
// every real number within that tolerance narrows to the SAME f32 -- then the reference's f is that float.  Otherwise (about
// once in 10^4 values, or brackets wider than a factor 1.5) the caller evaluates the logarithms.  a, b, x positive, a != b.
__device__ __forceinline__ bool log_weight_fast(double a, double b, double x, float& f)
{
    const double u1 = __dsub_rn(x, a), v1 = __dadd_rn(x, a), u2 = __dsub_rn(b, a), v2 = __dadd_rn(b, a);
    // a and b were born as positive finite f32 (callers test a, b, x > 0), so only the level x can be out of the range in which the
    // products below stay clear of f64 overflow and underflow: v1 < 1e100 bounds x; an infinite a or b fails the ratio tests
    if (!(fabs(u1) < 0.2 * v1 && fabs(u2) < 0.2 * v2 && v1 < 1e100))
        return false;
    const double s1 = u1 * rcp_newton(v1, false), s2 = u2 * rcp_newton(v2, false);
    const double z1 = s1 * s1, z2 = s2 * s2;
    double p1, p2; // operands straight from the constant bank: no immediates to materialise
    if (z1 < 0.004 && z2 < 0.004) {
        // brackets within a factor 1.13 (the bulk of a 65-level hybrid column): z^7 / 15 < 1.1e-18, seven terms are as exact as twelve
        p1 = p2 = c_inv_odd[6];
#pragma unroll
        for (int i = 5; i >= 0; --i) {
            p1 = __fma_rn(p1, z1, c_inv_odd[i]);
            p2 = __fma_rn(p2, z2, c_inv_odd[i]);
        }
    } else {
        p1 = p2 = c_inv_odd[11];
#pragma unroll
        for (int i = 10; i >= 0; --i) {
            p1 = __fma_rn(p1, z1, c_inv_odd[i]);
            p2 = __fma_rn(p2, z2, c_inv_odd[i]);
        }
    }
    const double num = (u1 * v2) * p1, den = (v1 * u2) * p2; // f = num / den; |den| > 0 because a != b
    // magnitudes: |u|, v within [1e-45, 1e100] * f32 range, p within [1, 1.02]: |num|, |den| lie in [1e-90, 1e140], no range test needed
    const double ad = fabs(den);
    const double rd = rcp_newton(ad, true);
    double q = fabs(num) * rd;
    q = __fma_rn(__fma_rn(-ad, q, fabs(num)), rd, q); // one correction step with the residual
    // |log b - log a| = 2 |s2| p2 >= 2 |s2|; |log| < 105 for f32-born a, b and for a level within a factor 1.5 of them
    // the bound only has to be an upper one: the division by |s2| is a multiplication by a reciprocal rounded up by 1e-6
    double rs2;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(rs2) : "d"(fabs(s2)));
    const double tol = (1. + q) * (2.0001e-14 * rs2) + 1e-15 * q;
    const float lo = __double2float_rn(q - tol), hi = __double2float_rn(q + tol);
    if (lo != hi)
        return false;
    f = ((num < 0) != (den < 0)) ? -lo : lo;
    return true;
}

// The same weight from table logarithms (fxcu_logtable.h) when log(x) is at hand (fixed target levels: computed by the host's
// libm, i.e. the reference's own number): f = (log x - log a) / (log b - log a) with two 10-operation logarithms and one
// reciprocal.  Each table logarithm is within 2e-15 of the true one, the reference's rounded logarithms within 2^-53 |log| (< 1.2e-14
// for any f32-born argument), so the quotient is within (1 + |f|) * 4e-14 / |log b - log a| of the reference's f64 quotient; it is
// returned only if every real number in that interval narrows to the SAME f32.  a, b positive f32-born, a != b.
// NORMAL: a and b are known to be positive normal numbers (their logarithms then differ by less than 178 in magnitude).
template <bool NORMAL = false> __device__ __forceinline__ bool log_weight_table(const double2* tab, float a, float b, double xlog, float& f)
{
    const double la = log_table_f32<NORMAL>(tab, a), lb = log_table_f32<NORMAL>(tab, b);
    const double num = xlog - la, den = lb - la;
    const double ad = fabs(den);
    if (!(ad > 1e-12 && (NORMAL || ad < 1e3) && fabs(num) < 1e6)) // brackets closer than the table's error, or junk: let the caller decide
        return false;
    double rd; // the hardware's reciprocal (relative error < 2^-23); the quotient gets a correction step, which squares that error
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(rd) : "d"(ad));
    double q = fabs(num) * rd;
    q = __fma_rn(__fma_rn(-ad, q, fabs(num)), rd, q);          // relative error < 2^-46 + 2^-52
    const double k = __fma_rn(rd, 4.0002e-14, 3e-14);          // (1 + q) k >= (1 + q) 4.0001e-14 / |den| + (1e-15 + 2^-45) q
    const double tol = __fma_rn(q, k, k);
    const float lo = __double2float_rn(q - tol), hi = __double2float_rn(q + tol);
    if (lo != hi)
        return false;
    f = ((num < 0) != (den < 0)) ? -lo : lo;
    return true;
}

// mifi_get_values_linear_f & friends for n == 1; returns false where the reference leaves the output untouched
// xlog: log(x) (LOG) or log(log(x + e)) (LOGLOG) computed on the host for fixed target levels, NaN = compute here
__device__ __forceinline__ float two_level(int method, float A, float B, double a, double b, double x, double xlog)
{
    if (method == FIMEXCU_VINT_METHOD_NN)
        return A; // interpolation.c:416-425
    bool have_f = false;
    float f = 0.f;
    if (method == FIMEXCU_VINT_METHOD_LOG || method == FIMEXCU_VINT_METHOD_LOGLOG) {
        if (a <= 0 || b <= 0 || x <= 0)
            return nan_f(); // reference: MIFI_ERROR, output unwritten
        // the same function of the same argument: exact weights whatever the logarithm's last bit (f == 0 / 1 copy A / B even
        // next to an undefined value)
        if (a == b || x == a) {
            have_f = true;
            f = 0.f;
        } else if (x == b) {
            have_f = true;
            f = 1.f;
        } else if (method == FIMEXCU_VINT_METHOD_LOG) {
            have_f = log_weight_fast(a, b, x, f);
        }
        if (!have_f && method == FIMEXCU_VINT_METHOD_LOGLOG) { // log(v + e) > 1 for v > 0: the inner checks cannot fail otherwise
            const double e = 2.7182818284590452354;
            a = log(__dadd_rn(a, e));
            b = log(__dadd_rn(b, e));
            if (isnan(xlog))
                x = log(__dadd_rn(x, e));
            if (a <= 0 || b <= 0 || (isnan(xlog) && x <= 0))
                return nan_f();
        }
        if (!have_f) {
            a = log(a);
            b = log(b);
            x = isnan(xlog) ? log(x) : xlog;
        }
    }
    if (!have_f)
        f = (a == b) ? 0.f : __double2float_rn(__ddiv_rn(__dsub_rn(x, a), __dsub_rn(b, a)));
    switch (method) {
    case FIMEXCU_VINT_METHOD_LIN_CONST_EXTRA:
        if (f >= 1.f)
            return B;
        if (f <= 0.f)
            return A;
        break;
    case FIMEXCU_VINT_METHOD_LIN_WEAK_EXTRA:
    case FIMEXCU_VINT_METHOD_LIN_NO_EXTRA: {
        if (f == 0.f)
            return A;
        if (f == 1.f)
            return B;
        const float lo = method == FIMEXCU_VINT_METHOD_LIN_WEAK_EXTRA ? -1.f : 0.f;
        const float hi = method == FIMEXCU_VINT_METHOD_LIN_WEAK_EXTRA ? 2.f : 1.f;
        if (!(f >= lo && f <= hi))
            return nan_f();
        break;
    }
    default:
        if (f == 0.f)
            return A; // exact copies avoid 0*NaN (interpolation.c:442-447)
        if (f == 1.f)
            return B;
        break;
    }
    return __fadd_rn(A, __fmul_rn(f, __fsub_rn(B, A)));
}

// FindNeighborElements.h:91-125 with the fallback :46-77, on a column stored with stride `st` floats
template <class Col>
__device__ __forceinline__ void find_neighbors(const Col& c, int n, float x, int& first, int& second)
{
    int lowest = 0, highest = 0;
    float lowDiff = __fsub_rn(x, c(0));
    float highDiff = -lowDiff;
    if (lowDiff < 0)
        lowDiff = FLT_MAX;
    if (highDiff < 0)
        highDiff = FLT_MAX;
    for (int k = 1; k < n; ++k) {
        const float diff = __fsub_rn(x, c(k));
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
        // extrapolating: the two closest distinct values (closer first)
        int r1 = 0, r2 = 0;
        float v1 = c(0);
        float d1 = fabsf(__fsub_rn(x, c(0)));
        float d2 = d1;
        for (int k = 0; k < n; ++k) {
            const float cur = c(k);
            const float d = fabsf(__fsub_rn(x, cur));
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
        first = r1;
        second = r2;
        return;
    }
    first = lowest;
    second = highest;
}

// The same answer as find_neighbors() for a column that is STRICTLY monotone and NaN-free (the usual pressure / height
// column), in O(log n).  Derived from the reference algorithm, including its first-element quirk:
//   x == c(0)                           -> (0,0)      (highDiff starts as -0.0 and is never improved)
//   x beyond the end that holds c(0)    -> (0,0)      (the fallback cannot find a second element)
//   x beyond the other end              -> (n-1,n-2)  (fallback: closest, then second closest)
//   otherwise                           -> (largest element <= x, smallest element > x)
// dir = +1 increasing, -1 decreasing.  Returns false when x is NaN (caller uses the general scan).
template <class Col>
__device__ __forceinline__ bool find_neighbors_monotone(const Col& c, int n, int dir, float x, int& first, int& second)
{
    if (isnan(x))
        return false;
    const float c0 = c(0), cl = c(n - 1);
    if (x == c0) {
        first = second = 0;
        return true;
    }
    if (dir > 0 ? x < c0 : x > c0) { // beyond the first element: (0,0) as long as the second element is measurably farther
        if (n >= 2 && !(fabsf(__fsub_rn(x, c0)) < fabsf(__fsub_rn(x, c(1)))))
            return false;
        first = second = 0;
        return true;
    }
    if (n == 1) { // single level, x on the far side: fallback finds nothing else
        first = second = 0;
        return true;
    }
    if (dir > 0 ? x >= cl : x <= cl) {
        // far end: (n-1, n-2), provided the f32 distances of the last two elements are distinguishable (they are unless |x|
        // exceeds 2^24 level spacings; then the reference's tie handling decides and the general scan is used)
        const float dl = fabsf(__fsub_rn(x, cl)), d2 = fabsf(__fsub_rn(x, c(n - 2)));
        if (!(dl < d2))
            return false;
        // the fallback walks upwards and keeps the FIRST element of a run of equal distances as runner-up: n-2 only if n-3 is
        // measurably farther
        if (n >= 3 && !(d2 < fabsf(__fsub_rn(x, c(n - 3)))))
            return false;
        first = n - 1;
        second = n - 2;
        return true;
    }
    // c0 < x < cl (increasing) or c0 > x > cl (decreasing): bracket by bisection; invariant: x lies between c[lo] and c[hi]
    int lo = 0, hi = n - 1;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        const float v = c(mid);
        if (dir > 0 ? v <= x : v >= x)
            lo = mid;
        else
            hi = mid;
    }
    // the reference minimises the ROUNDED differences x - c[k]; they order like the exact ones unless two of them round to the
    // same float -- then leave the decision to the general scan
    if (lo > 0 && __fsub_rn(x, c(lo)) == __fsub_rn(x, c(lo - 1)))
        return false;
    // a difference that overflows to FLT_MAX or beyond looks like "no neighbour on this side" to the reference
    if (!(fabsf(__fsub_rn(x, c(lo))) < FLT_MAX && fabsf(__fsub_rn(x, c(hi))) < FLT_MAX))
        return false;
    if (hi < n - 1 && __fsub_rn(x, c(hi)) == __fsub_rn(x, c(hi + 1)))
        return false;
    // increasing: c[lo] <= x < c[hi] -> lowest = lo, highest = hi.  decreasing: c[lo] >= x > c[hi]:
    //   largest element <= x is c[hi] unless c[lo] == x; smallest element > x is c[lo] unless c[lo] == x, then c[lo-1]
    if (dir > 0) {
        first = lo;
        second = hi;
    } else if (c(lo) == x) {
        first = lo;
        second = lo - 1; // lo >= 1 here because x != c(0)
    } else {
        first = hi;
        second = lo;
    }
    return true;
}

// column accessors: the vertical coordinate of one column as c(k)
struct StridedCol { // shared memory ([nzi][VT]) or global memory ([nzi][nInner])
    const float* p;
    int st;
    __device__ __forceinline__ float operator()(int k) const { return p[k * st]; }
};
struct HybridCol { // p = ap + b*ps formed on the fly: vertical_coordinate_transformations.c:62-68, then asFloat() (CDMVerticalInterpolator.cc:483)
    const double* ap;
    const double* b;
    double ps;
    __device__ __forceinline__ float operator()(int k) const { return __double2float_rn(__dadd_rn(ap[k], __dmul_rn(b[k], ps))); }
};

// CDMVerticalInterpolator.cc:505-528: is the target level inside the column's [vmin, vmax]?
__device__ __forceinline__ bool level_in_range(const VertParams& P, double vout, double lo, double hi)
{
    if (P.vmin && P.vmax)
        return (vout >= lo) && (vout <= hi);
    if (P.vmin)
        return vout >= lo;
    if (P.vmax)
        return vout <= hi;
    return true;
}

// what one output value needs besides its column: passed by value to the out-of-line helpers of the sweep kernel
struct OutCfg {
    int method;
    float bad, fill, valid_min, valid_max;
    size_t nInner;
    const double2* logtab; // shared-memory logarithm table of the kernel (fxcu_logtable.h), or nullptr
};
__device__ __forceinline__ OutCfg out_cfg(const VertParams& P) { return OutCfg{P.method, P.bad, P.fill, P.valid_min, P.valid_max, P.nInner, nullptr}; }

// valid_min / valid_max clamp (CDMVerticalInterpolator.cc:551-561; NaN compares false and stays), fill value, streaming store
__device__ __forceinline__ void store_result(const OutCfg& C, float* dst, float r)
{
    if (r < C.valid_min)
        r = C.valid_min;
    if (r > C.valid_max)
        r = C.valid_max;
    if (!isnan(C.fill) && isnan(r))
        r = C.fill;
    __stcs(dst, r);
}

// A column whose coordinate is NaN at every level (a hybrid column over an undefined surface pressure, an explicit column outside
// the source grid of the horizontal step before): every difference the reference's scan forms is NaN, no comparison holds, it
// ends at (0,0) for every target level (FindNeighborElements.h:91-125) and the result is undefined whatever the values are.
__device__ __forceinline__ void column_undefined(const OutCfg& C, int nzo, size_t nInner, float* dout)
{
    for (int jj = 0; jj < nzo; ++jj)
        store_result(C, dout + static_cast<size_t>(jj) * nInner, nan_f());
}

// out-of-line copy for the rare cases the table kernel's LOG path hands back (keeps its unrolled column loop small)
__device__ __noinline__ float two_level_out_of_line(int method, float A, float B, double a, double b, double x, double xlog)
{
    return two_level(method, A, B, a, b, x, xlog);
}

// the two bracketing levels' values through the two-level kernel
// TRUSTED (table kernel): the caller has established, for the LOG method, that the level is positive with a host-computed logarithm,
// that a and b are positive normal numbers and that the level's f32 value differs from both (so the level itself does too)
template <int METHOD = -1, bool TRUSTED = false> // METHOD >= 0: compile-time constant (the other methods' arithmetic drops out of the caller)
__device__ __forceinline__ float blend_levels(const OutCfg& C, float A, float B, float a, float b, double vout, double xlog)
{
    const int method = METHOD >= 0 ? METHOD : C.method;
    if (!isnan(C.bad)) {
        if (A == C.bad)
            A = nan_f();
        if (B == C.bad)
            B = nan_f();
    }
    if (C.logtab && method == FIMEXCU_VINT_METHOD_LOG && (TRUSTED || (!isnan(xlog) && a > 0.f && b > 0.f && a != b))) {
        // fixed target level with its host-computed logarithm, f32 bracket: table logarithms (the exact-weight cases x == a / x == b
        // are the quotient's own 0 and 1 only approximately, so they stay with two_level)
        float f;
        const double ad = static_cast<double>(a), bd = static_cast<double>(b);
        if ((TRUSTED || (vout > 0 && vout != ad && vout != bd)) && log_weight_table<TRUSTED>(C.logtab, a, b, xlog, f)) {
            if (f == 0.f)
                return A; // interpolation.c:442-447
            if (f == 1.f)
                return B;
            return __fadd_rn(A, __fmul_rn(f, __fsub_rn(B, A)));
        }
    }
    if (C.logtab)
        return two_level_out_of_line(method, A, B, static_cast<double>(a), static_cast<double>(b), vout, xlog);
    return two_level(method, A, B, static_cast<double>(a), static_cast<double>(b), vout, xlog);
}

// one output value by the reference's neighbour search (dir: +1 / -1 column known strictly monotone, 0 general scan)
template <class Col>
__device__ __forceinline__ float general_output(const OutCfg& C, const Col& c, int nzi, int dir, const float* din, double vout, double xlog)
{
    int p0, p1;
    const float xf = __double2float_rn(vout);
    if (!(dir != 0 && find_neighbors_monotone(c, nzi, dir, xf, p0, p1)))
        find_neighbors(c, nzi, xf, p0, p1);
    if (p0 == p1)
        return nan_f();
    return blend_levels(C, __ldg(din + static_cast<size_t>(p0) * C.nInner), __ldg(din + static_cast<size_t>(p1) * C.nInner), c(p0), c(p1), vout, xlog);
}

// +1 strictly increasing, -1 strictly decreasing (both NaN-free), 0 otherwise
template <class Col>
__device__ __forceinline__ int column_direction(const Col& c, int nzi)
{
    if (nzi < 2)
        return 0;
    bool inc = true, dec = true;
    float prev = c(0);
    for (int k = 1; k < nzi; ++k) {
        const float cur = c(k);
        inc = inc && (cur > prev);
        dec = dec && (cur < prev);
        prev = cur;
    }
    return inc ? 1 : (dec ? -1 : 0);
}

template <class Col>
__device__ __forceinline__ void vertical_column(const VertParams& P, const Col& c, size_t cidx, size_t outer, size_t inner, int dir_known)
{
    const int nzi = static_cast<int>(P.nzi);
    const int dir = dir_known != 0 ? dir_known : column_direction(c, nzi);
    const float* din = P.in + outer * P.nzi * P.nInner + inner;
    float* dout = P.out + outer * P.nzo * P.nInner + inner;
    const double lo = P.vmin ? P.vmin[cidx] : 0., hi = P.vmax ? P.vmax[cidx] : 0.;
    const OutCfg C = out_cfg(P);
    for (size_t k = 0; k < P.nzo; ++k) {
        const double vout = P.vcoord_out ? static_cast<double>(__ldg(P.vcoord_out + (outer * P.nzo + k) * P.nInner + inner)) : __ldg(P.levels + k);
        float r = nan_f();
        if (level_in_range(P, vout, lo, hi))
            r = general_output(C, c, nzi, dir, din, vout, P.xlog ? __ldg(P.xlog + k) : nan(""));
        store_result(C, dout + k * P.nInner, r);
    }
}

// MODE 0: coordinate column staged in shared memory; 1: hybrid pressure on the fly (ap, b in shared memory, no column storage,
// full occupancy); 2: coordinate scanned in global memory (very long columns)
template <int MODE>
__global__ void __launch_bounds__(VT) vertical_kernel(const VertParams P)
{
    extern __shared__ double smem_d[];
    const int nzi = static_cast<int>(P.nzi);
    if (MODE == 1) { // whole block first: coefficients to shared memory
        for (int k = threadIdx.x; k < nzi; k += VT) {
            smem_d[k] = P.ap[k];
            smem_d[nzi + k] = P.b[k];
        }
        __syncthreads();
    }
    const size_t ncol = P.nOuter * P.nInner;
    const size_t cidx = blockIdx.x * static_cast<size_t>(VT) + threadIdx.x;
    if (cidx >= ncol)
        return;
    const size_t outer = cidx / P.nInner, inner = cidx - outer * P.nInner;
    if (pad_column(P, inner))
        return;
    if (MODE == 0) {
        float* mine = reinterpret_cast<float*>(smem_d) + threadIdx.x;
        const float* g = P.vcoord_in + outer * P.nzi * P.nInner + inner;
        for (int k = 0; k < nzi; ++k)
            mine[k * VT] = __ldg(g + static_cast<size_t>(k) * P.nInner);
        vertical_column(P, StridedCol{mine, VT}, cidx, outer, inner, 0);
    } else if (MODE == 1) {
        const size_t po = (P.nOuterPs == 1 ? 0 : outer) * P.nInner + inner;
        const double ps = __dmul_rn(static_cast<double>(__ldg(P.ps + po)), P.ps_scale);
        if (isnan(ps)) // ap + b * NaN is NaN at every level
            column_undefined(out_cfg(P), static_cast<int>(P.nzo), P.nInner, P.out + outer * P.nzo * P.nInner + inner);
        else
            vertical_column(P, HybridCol{smem_d, smem_d + nzi, ps}, cidx, outer, inner, P.dir_hint);
    } else {
        vertical_column(P, StridedCol{P.vcoord_in + outer * P.nzi * P.nInner + inner, static_cast<int>(P.nInner)}, cidx, outer, inner, 0);
    }
}

// ---- fast paths for strictly monotone columns and FIXED target levels ------------------------------------------------------------
// For such a column the reference's search returns (see find_neighbors_monotone), with n(x) = number of elements on the near
// side of x in column order (c(k) <= x for an increasing, c(k) >= x for a decreasing column) and step = (x == c(0)) ? 0 : n(x):
//   step == 0      (0,0): nothing to interpolate between;  step == nzi   (nzi-1, nzi-2);  else lo = step-1, hi = step.
// value_from_step() turns a step into the output value; f32 difference ties and the other corner cases in which the reference's
// scan is not fully determined by the bracket go to the out-of-line general search (rare).
template <class Col>
__device__ __noinline__ float sweep_slow(OutCfg C, Col c, int nzi, int dir, const float* din, double vout, double xlog)
{
    return general_output(C, c, nzi, dir, din, vout, xlog);
}

struct PairCol { // hybrid pressure from interleaved (ap, b) pairs
    const double2* apb;
    double ps;
    __device__ __forceinline__ float operator()(int k) const
    {
        const double2 q = apb[k];
        return __double2float_rn(__dadd_rn(q.x, __dmul_rn(q.y, ps)));
    }
};

struct ColumnEnds { // c(0), c(1), c(nzi-1), c(nzi-2), c(nzi-3)
    float c0, c1, cl, cl2, cl3;
};

// plo = c(step-1), phi = c(step) for an interior step (ignored otherwise)
struct GlobalVal { // the column's values where they lie
    const float* din;
    size_t nInner;
    __device__ __forceinline__ float operator()(int k) const { return __ldg(din + static_cast<size_t>(k) * nInner); }
};
template <class Col, class Val>
__device__ __forceinline__ float value_from_step(const OutCfg& C, const Col& c, const Val& val, int nzi, bool inc, const ColumnEnds& E, int step, float x,
                                                 float plo, float phi, const float* din, double vout, double xlog)
{
    bool fast, distinct = true;
    int first = 0, second = 0;
    float a = 0.f, b = 0.f;
    if (step == 0) {
        distinct = false;
        fast = (x == E.c0) || fabsf(__fsub_rn(x, E.c0)) < fabsf(__fsub_rn(x, E.c1));
    } else if (step == nzi) {
        first = nzi - 1;
        second = nzi - 2;
        a = E.cl;
        b = E.cl2;
        const float d2 = fabsf(__fsub_rn(x, E.cl2));
        fast = fabsf(__fsub_rn(x, E.cl)) < d2 && (nzi < 3 || d2 < fabsf(__fsub_rn(x, E.cl3)));
    } else {
        const float dlo = __fsub_rn(x, plo), dhi = __fsub_rn(x, phi);
        fast = fabsf(dlo) < FLT_MAX && fabsf(dhi) < FLT_MAX && (inc || x != plo);
        if (fast && step > 1)
            fast = dlo != __fsub_rn(x, c(step - 2));
        first = inc ? step - 1 : step;
        second = inc ? step : step - 1;
        a = inc ? plo : phi;
        b = inc ? phi : plo;
    }
    if (!fast)
        return sweep_slow(C, c, nzi, inc ? 1 : -1, din, vout, xlog);
    if (!distinct)
        return nan_f();
    return blend_levels(C, val(first), val(second), a, b, vout, xlog);
}

// every output of a column by the reference's scan (NaN / unordered / single-level columns)
template <class Col>
__device__ __forceinline__ void column_by_general_scan(const VertParams& P, const OutCfg& C, const Col& c, int nzi, int nzo, const double* s_xd,
                                                       const double* s_xl, double rlo, double rhi, const float* din, float* dout)
{
    // direction (as column_direction) and "nothing but NaN" in one pass
    bool inc = nzi >= 2, dec = nzi >= 2;
    float prev = c(0);
    bool allnan = isnan(prev);
    for (int k = 1; k < nzi; ++k) {
        const float cur = c(k);
        inc = inc && (cur > prev);
        dec = dec && (cur < prev);
        allnan = allnan && isnan(cur);
        prev = cur;
    }
    if (allnan) {
        column_undefined(C, nzo, P.nInner, dout);
        return;
    }
    const int dir = inc ? 1 : (dec ? -1 : 0);
    for (int jj = 0; jj < nzo; ++jj) {
        const double vout = s_xd[jj];
        float r = nan_f();
        if (level_in_range(P, vout, rlo, rhi))
            r = sweep_slow(C, c, nzi, dir, din, vout, s_xl[jj]);
        store_result(C, dout + static_cast<size_t>(jj) * P.nInner, r);
    }
}

// Sweep kernel, for target levels that are SORTED (either way).  Phase 1 walks the column once, level by level (uniform,
// broadcast accesses; an explicit coordinate streams through registers CH levels ahead), and notes for every target level the
// step at which the walk passes it; monotonicity is verified on the way.  Phase 2 runs over the target levels with all lanes in
// step.  Shared memory: hybrid (ap, b) pairs, the level list (f64, log, f32 with a NaN sentinel either side), notes [nzo][VT] u16.
constexpr int LG = 4; // target levels handled together in straight-line code (their 2 * LG gathers are in flight at once)
constexpr int CH = 4;

template <bool HYBRID, int METHOD = -1> // METHOD >= 0 (explicit coordinate only): the interpolation method as a compile-time constant
__global__ void __launch_bounds__(VT, 8) vertical_sweep_kernel(const VertParams P)
{
    extern __shared__ double smem_d[];
    const int nzi = static_cast<int>(P.nzi), nzo = static_cast<int>(P.nzo);
    double2* s_apb = reinterpret_cast<double2*>(smem_d);
    double* s_xd = smem_d + (HYBRID ? 2 * nzi : 0);
    double* s_xl = s_xd + nzo;
    float* s_xf = reinterpret_cast<float*>(s_xl + nzo) + 1; // [-1] and [nzo] are sentinels
    unsigned short* s_k = reinterpret_cast<unsigned short*>(s_xf + nzo + 1 + (nzo & 1)) + threadIdx.x; // [nzo][VT], this thread's column
    // explicit coordinate: the walk also leaves c(step-2), c(step-1), c(step) of every target level here, [3][nzo][VT]
    float* s_c = reinterpret_cast<float*>(s_k - threadIdx.x + static_cast<size_t>(nzo + (nzo & 1)) * VT) + threadIdx.x;
    if (HYBRID) {
        for (int k = threadIdx.x; k < nzi; k += VT)
            s_apb[k] = make_double2(P.ap[k], P.b[k]);
    }
    for (int k = threadIdx.x; k < nzo; k += VT) {
        const double x = P.levels[k];
        s_xd[k] = x;
        s_xl[k] = P.xlog ? P.xlog[k] : nan("");
        s_xf[k] = __double2float_rn(x);
    }
    if (threadIdx.x == 0) {
        s_xf[-1] = nan_f();
        s_xf[nzo] = nan_f();
    }
    __shared__ double2 s_log[(!HYBRID && METHOD == FIMEXCU_VINT_METHOD_LOG) ? 128 : 1];
    if (!HYBRID && METHOD == FIMEXCU_VINT_METHOD_LOG && P.logtab)
        for (int i = threadIdx.x; i < 128; i += VT)
            s_log[i] = P.logtab[i]; // made once per device (log_table_device)
    __syncthreads();
    const size_t ncol = P.nOuter * P.nInner;
    const size_t cidx = blockIdx.x * static_cast<size_t>(VT) + threadIdx.x;
    if (cidx >= ncol)
        return;
    const size_t outer = cidx / P.nInner, inner = cidx - outer * P.nInner;
    if (pad_column(P, inner))
        return;
    const float* din = P.in + outer * P.nzi * P.nInner + inner;
    float* dout = P.out + outer * P.nzo * P.nInner + inner;
    const float* gc = HYBRID ? nullptr : P.vcoord_in + outer * P.nzi * P.nInner + inner;
    double ps = 0.;
    if (HYBRID)
        ps = __dmul_rn(static_cast<double>(__ldg(P.ps + (P.nOuterPs == 1 ? 0 : outer) * P.nInner + inner)), P.ps_scale);
    const PairCol hc{s_apb, ps};
    const StridedCol sc{gc, static_cast<int>(P.nInner)};
    auto coord = [&](int k) -> float { return HYBRID ? hc(k) : sc(k); };
    OutCfg C = out_cfg(P);
    if (!HYBRID && METHOD == FIMEXCU_VINT_METHOD_LOG && P.logtab)
        C.logtab = s_log;
    const double rlo = P.vmin ? P.vmin[cidx] : 0., rhi = P.vmax ? P.vmax[cidx] : 0.;

    // ---- phase 1: one walk along the column, in the order in which it increases after multiplication by sg ----
    bool ok = nzi >= 2;
    ColumnEnds E{};
    bool inc = false;
    if (ok) {
        E.c0 = coord(0);
        E.c1 = coord(1);
        inc = E.c1 > E.c0;
        ok = inc || (E.c1 < E.c0);
    }
    if (ok) {
        const float sg = inc ? 1.f : -1.f; // exact
        const bool forward = (inc ? 1 : -1) == P.levels_dir || nzo == 1;
        const int xstep = forward ? 1 : -1, kstep = xstep * VT;
        const float* xp = s_xf + (forward ? 0 : nzo - 1);
        unsigned short* kp = s_k + (forward ? 0 : nzo - 1) * VT;
        float x = __fmul_rn(sg, *xp); // NaN once the list is exhausted
        float pm1 = __fmul_rn(sg, E.c0), pm2 = pm1, pcur = pm1; // walk-order values of c(k-1), c(k-2), c(k)
        float* cp = s_c + (forward ? 0 : nzo - 1) * VT;
        const int cplane = nzo * VT;
        auto note = [&](int step) {
            *kp = static_cast<unsigned short>(step);
            if (!HYBRID) { // sg is +-1: multiplying again gives the stored values back exactly
                cp[0] = __fmul_rn(sg, pm2);
                cp[cplane] = __fmul_rn(sg, pm1);
                cp[2 * cplane] = __fmul_rn(sg, pcur);
                cp += kstep;
            }
            xp += xstep;
            kp += kstep;
            x = __fmul_rn(sg, *xp);
        };
        while (x <= pm1) // at or beyond the end that holds element 0
            note(0);
        bool mono = true;
        if (HYBRID) {
            for (int k = 1; k < nzi; ++k) {
                const float pk = __fmul_rn(sg, hc(k));
                mono = mono && (pk > pm1);
                while (x < pk) // c(k-1) <= x < c(k) in walk order
                    note(k);
                pm1 = pk;
            }
        } else {
            // whole chunks of CH levels, the next chunk's loads issued before this one is used; level offsets are 32-bit
            // (sweep_applies) and the read pointer is bumped, so a level costs a load, a compare and a branch
            const unsigned lstride = static_cast<unsigned>(P.nInner);
            const float* gp = gc + lstride; // level 1
            const int nchunks = (nzi - 1) / CH;
            float ccur[CH];
#pragma unroll
            for (int i = 0; i < CH; ++i)
                ccur[i] = nchunks > 0 ? __ldg(gp + static_cast<unsigned>(i) * lstride) : 0.f;
            int k = 1;
            for (int c = 0; c < nchunks; ++c, k += CH) {
                gp += CH * lstride;
                float cnx[CH];
                const bool more = c + 1 < nchunks;
#pragma unroll
                for (int i = 0; i < CH; ++i)
                    cnx[i] = more ? __ldg(gp + static_cast<unsigned>(i) * lstride) : 0.f;
#pragma unroll
                for (int i = 0; i < CH; ++i) {
                    const float pk = __fmul_rn(sg, ccur[i]);
                    mono = mono && (pk > pm1);
                    pcur = pk;
                    if (__any_sync(__activemask(), x < pk)) // most levels pass no target level in any column of the warp
                        while (x < pk)
                            note(k + i);
                    pm2 = pm1;
                    pm1 = pk;
                }
#pragma unroll
                for (int i = 0; i < CH; ++i)
                    ccur[i] = cnx[i];
            }
            for (; k < nzi; ++k, gp += lstride) { // fewer than CH levels left (gp stands at level k)
                const float pk = __fmul_rn(sg, __ldg(gp));
                mono = mono && (pk > pm1);
                pcur = pk;
                while (x < pk)
                    note(k);
                pm2 = pm1;
                pm1 = pk;
            }
        }
        ok = mono;
        if (ok)
            while (!isnan(x)) // at or beyond the far end
                note(nzi);
    }
    if (!ok) {
        if (HYBRID && isnan(ps)) // NaN surface pressure: ap + b * NaN is NaN at every level
            column_undefined(C, nzo, P.nInner, dout);
        else if (HYBRID)
            column_by_general_scan(P, C, hc, nzi, nzo, s_xd, s_xl, rlo, rhi, din, dout);
        else
            column_by_general_scan(P, C, sc, nzi, nzo, s_xd, s_xl, rlo, rhi, din, dout);
        return;
    }

    // ---- phase 2: the target levels, all lanes in step ----
    E.cl = coord(nzi - 1);
    E.cl2 = coord(nzi - 2);
    E.cl3 = nzi >= 3 ? coord(nzi - 3) : 0.f;
    if (HYBRID) {
        for (int jj = 0; jj < nzo; ++jj) {
            const double vout = s_xd[jj];
            float r = nan_f();
            if (level_in_range(P, vout, rlo, rhi)) {
                const int step = s_k[jj * VT];
                float plo = 0.f, phi = 0.f;
                if (step > 0 && step < nzi) {
                    plo = hc(step - 1);
                    phi = hc(step);
                }
                r = value_from_step(C, hc, GlobalVal{din, P.nInner}, nzi, inc, E, step, s_xf[jj], plo, phi, din, vout, s_xl[jj]);
            }
            store_result(C, dout + static_cast<size_t>(jj) * P.nInner, r);
        }
        return;
    }
    // explicit coordinate: the walk left the bracket of every level in shared memory, so a level costs two gathers (the values);
    // the decision is predicate logic and selects as in the table kernel, LG levels in straight-line code
    constexpr bool LOGM = METHOD == FIMEXCU_VINT_METHOD_LOG;
    const unsigned stride = static_cast<unsigned>(P.nInner); // sweep_applies(): nzi * nInner and nzo * nInner < 2^31
    const bool ranged = P.vmin || P.vmax;
    const bool plain = isnan(C.valid_min) && isnan(C.valid_max) && isnan(C.fill);
    // LOG weights from the table logarithms need positive normal brackets: the column is monotone, so its ends decide
    bool trusted = LOGM && C.logtab != nullptr;
    if (LOGM) {
        const unsigned b0 = __float_as_uint(E.c0), b1 = __float_as_uint(E.cl);
        trusted = trusted && !(b0 - 0x00800000u >= 0x7f000000u || b1 - 0x00800000u >= 0x7f000000u);
    }
    const int cplane = nzo * VT;
    for (int j0 = 0; j0 < nzo; j0 += LG) {
        float A[LG], B[LG], ca[LG], cb[LG]; // ca: NaN = general search, +inf = the result is NaN (undefined level)
#pragma unroll
        for (int u = 0; u < LG; ++u) {
            const int jj = min(j0 + u, nzo - 1);
            const int step = s_k[jj * VT];
            const float x = s_xf[jj];
            const float g = s_c[jj * VT], plo = s_c[cplane + jj * VT], phi = s_c[2 * cplane + jj * VT];
            const float dlo = __fsub_rn(x, plo), dhi = __fsub_rn(x, phi), dg = __fsub_rn(x, g);
            // interior step (see value_from_step): finite differences, x != c(step-1) when walking down, no f32 tie with c(step-2)
            bool ok_in = fabsf(dlo) < FLT_MAX && fabsf(dhi) < FLT_MAX && (inc || x != plo) && (step <= 1 || dlo != dg);
            const bool far = step >= nzi, zero = step == 0;
            const float d2 = fabsf(__fsub_rn(x, E.cl2));
            const bool ok_far = fabsf(__fsub_rn(x, E.cl)) < d2 && (nzi < 3 || d2 < fabsf(__fsub_rn(x, E.cl3)));
            const bool zero_nan = (x == E.c0) || fabsf(__fsub_rn(x, E.c0)) < fabsf(__fsub_rn(x, E.c1));
            bool ok = !zero && (far ? ok_far : ok_in);
            const float a = far ? E.cl : (inc ? plo : phi);
            const float b = far ? E.cl2 : (inc ? phi : plo);
            if (LOGM) // exact hits and levels without a usable host logarithm are the general search's business
                ok = ok && (!trusted || (x != a && x != b && s_xd[jj] > 0. && !isnan(s_xl[jj])));
            bool undefined = zero && zero_nan;
            if (ranged)
                undefined = undefined || !level_in_range(P, s_xd[jj], rlo, rhi);
            int first = far ? nzi - 1 : (inc ? step - 1 : step);
            int second = far ? nzi - 2 : (inc ? step : step - 1);
            first = (ok && !undefined) ? first : 0;
            second = (ok && !undefined) ? second : 0;
            A[u] = __ldg(din + static_cast<unsigned>(first) * stride);
            B[u] = __ldg(din + static_cast<unsigned>(second) * stride);
            ca[u] = undefined ? INFINITY : (ok ? a : nan_f());
            cb[u] = b;
        }
#pragma unroll
        for (int u = 0; u < LG; ++u) {
            const int jj = j0 + u;
            if (jj < nzo) {
                float r;
                if (isnan(ca[u]))
                    r = sweep_slow(C, sc, nzi, inc ? 1 : -1, din, s_xd[jj], s_xl[jj]);
                else if (ca[u] == INFINITY)
                    r = nan_f();
                else if (LOGM && trusted)
                    r = blend_levels<METHOD, true>(C, A[u], B[u], ca[u], cb[u], s_xd[jj], s_xl[jj]);
                else
                    r = blend_levels<METHOD, false>(C, A[u], B[u], ca[u], cb[u], s_xd[jj], s_xl[jj]);
                float* dst = dout + static_cast<unsigned>(jj) * stride;
                if (plain)
                    __stcs(dst, r);
                else
                    store_result(C, dst, r);
            }
        }
    }
}

// ---- explicit coordinate, fixed target levels: scan + bisection -----------------------------------------------------------------
// The walk of the sweep kernel notes, for every lane, the input level at which it passes each target level -- and a warp pays for
// every distinct such level among its 32 columns (285 of the sweep's 455 instructions per output value, ncu).  Here the column is
// only SCANNED once (a load and two compares per level: is it strictly monotone, NaN-free?), and every target level then finds its
// step by bisection (a fixed number of rounds; the column is hot in L1/L2 after the scan, the probes of a group of LG levels are
// independent of each other).  The step is the same number the walk notes -- the count of elements at or before the level in
// walk order, 0 at or before the first -- so everything downstream is the sweep kernel's second phase.  Target levels need not
// be sorted.  Shared memory: the level list only (f64, log, f32).
constexpr int SLG = 4; // target levels per group of the search kernel (their probes and 2 * SLG gathers are independent)
// The block's coordinate columns are staged in shared memory ([nzi][VT], one bank per column: conflict-free whatever level a lane
// asks for) while they are scanned, with sixteen loads in flight per thread; the probes of the bisection and the bracket values
// then come from there instead of paying a trip to L1 / L2 per round (a first form that probed global memory was latency-bound:
// 41 % of the stalls on the scoreboard, slower than the sweep).
template <int METHOD> // METHOD >= 0: compile-time interpolation method; -1: the method of the launch
__global__ void __launch_bounds__(VT, 6) vertical_search_kernel(const VertParams P, int rounds)
{
    extern __shared__ double smem_d[];
    const int nzi = static_cast<int>(P.nzi), nzo = static_cast<int>(P.nzo);
    double* s_xd = smem_d;
    double* s_xl = s_xd + nzo;
    float* s_xf = reinterpret_cast<float*>(s_xl + nzo);
    float* s_col = s_xf + nzo + (nzo & 1) + threadIdx.x; // [nzi][VT], this thread's column
    constexpr bool LOGM = METHOD == FIMEXCU_VINT_METHOD_LOG;
    __shared__ double2 s_log[LOGM ? 128 : 1];
    for (int k = threadIdx.x; k < nzo; k += VT) {
        const double x = P.levels[k];
        s_xd[k] = x;
        s_xl[k] = P.xlog ? P.xlog[k] : nan("");
        s_xf[k] = __double2float_rn(x);
    }
    if (LOGM && P.logtab)
        for (int i = threadIdx.x; i < 128; i += VT)
            s_log[i] = P.logtab[i];
    __syncthreads();
    const size_t ncol = P.nOuter * P.nInner;
    const size_t cidx = blockIdx.x * static_cast<size_t>(VT) + threadIdx.x;
    if (cidx >= ncol)
        return;
    const size_t outer = cidx / P.nInner, inner = cidx - outer * P.nInner;
    if (pad_column(P, inner))
        return;
    const float* din = P.in + outer * P.nzi * P.nInner + inner;
    float* dout = P.out + outer * P.nzo * P.nInner + inner;
    const float* gc = P.vcoord_in + outer * P.nzi * P.nInner + inner;
    const unsigned stride = static_cast<unsigned>(P.nInner); // the launch guarantees nzi * nInner and nzo * nInner < 2^31
    const StridedCol sc{gc, static_cast<int>(P.nInner)};
    OutCfg C = out_cfg(P);
    if (LOGM && P.logtab)
        C.logtab = s_log;
    const double rlo = P.vmin ? P.vmin[cidx] : 0., rhi = P.vmax ? P.vmax[cidx] : 0.;

    // ---- the scan: direction from the first two elements, then strictly monotone all the way (NaN fails every compare) ----
    ColumnEnds E{};
    E.c0 = __ldg(gc);
    E.c1 = __ldg(gc + stride);
    s_col[0] = E.c0;
    s_col[VT] = E.c1;
    const bool inc = E.c1 > E.c0;
    bool ok = inc || (E.c1 < E.c0);
    const float sg = inc ? 1.f : -1.f; // exact
    {
        float prev = __fmul_rn(sg, E.c1);
        const float* gp = gc + 2 * stride;
        int k = 2;
        constexpr int SC = 16; // loads in flight per thread: the scan is all memory latency
        for (; k + SC <= nzi; k += SC, gp += SC * stride) {
            float v[SC];
#pragma unroll
            for (int i = 0; i < SC; ++i)
                v[i] = __ldg(gp + static_cast<unsigned>(i) * stride);
#pragma unroll
            for (int i = 0; i < SC; ++i) {
                s_col[(k + i) * VT] = v[i];
                const float w = __fmul_rn(sg, v[i]);
                ok = ok && (w > prev);
                prev = w;
            }
        }
        for (; k < nzi; ++k, gp += stride) {
            const float c = __ldg(gp);
            s_col[k * VT] = c;
            const float v = __fmul_rn(sg, c);
            ok = ok && (v > prev);
            prev = v;
        }
    }
    if (!ok) {
        column_by_general_scan(P, C, sc, nzi, nzo, s_xd, s_xl, rlo, rhi, din, dout);
        return;
    }
    E.cl = s_col[(nzi - 1) * VT];
    E.cl2 = s_col[(nzi - 2) * VT];
    E.cl3 = nzi >= 3 ? s_col[(nzi - 3) * VT] : 0.f;
    const float w0 = __fmul_rn(sg, E.c0);
    const bool ranged = P.vmin || P.vmax;
    const bool plain = isnan(C.valid_min) && isnan(C.valid_max) && isnan(C.fill);
    // LOG weights from the table logarithms need positive normal brackets: the column is monotone, so its ends decide
    bool trusted = LOGM && C.logtab != nullptr;
    if (LOGM) {
        const unsigned b0 = __float_as_uint(E.c0), b1 = __float_as_uint(E.cl);
        trusted = trusted && !(b0 - 0x00800000u >= 0x7f000000u || b1 - 0x00800000u >= 0x7f000000u);
    }

    // ---- the target levels, SLG at a time ----
    for (int j0 = 0; j0 < nzo; j0 += SLG) {
        int lo[SLG], hi[SLG];
        float xs[SLG];
#pragma unroll
        for (int u = 0; u < SLG; ++u) {
            lo[u] = 0;
            hi[u] = nzi;
            xs[u] = __fmul_rn(sg, s_xf[min(j0 + u, nzo - 1)]);
        }
        for (int r = 0; r < rounds; ++r) { // count of elements <= the level in walk order (a NaN level counts none)
#pragma unroll
            for (int u = 0; u < SLG; ++u) {
                const int mid = min((lo[u] + hi[u]) >> 1, nzi - 1);
                const float v = __fmul_rn(sg, s_col[mid * VT]);
                const bool le = v <= xs[u];
                lo[u] = le ? mid + 1 : lo[u];
                hi[u] = le ? hi[u] : min(hi[u], mid);
            }
        }
        float A[SLG], B[SLG], ca[SLG], cb[SLG]; // ca: NaN = general search, +inf = the result is NaN (undefined level)
#pragma unroll
        for (int u = 0; u < SLG; ++u) {
            const int jj = min(j0 + u, nzo - 1);
            const float x = s_xf[jj];
            const int step = (xs[u] <= w0) ? 0 : lo[u]; // at or before the first element: step 0, as the walk has it
            const float g = s_col[max(step - 2, 0) * VT], plo = s_col[max(step - 1, 0) * VT], phi = s_col[min(step, nzi - 1) * VT];
            const float dlo = __fsub_rn(x, plo), dhi = __fsub_rn(x, phi), dg = __fsub_rn(x, g);
            // interior step (see value_from_step): finite differences, x != c(step-1) when walking down, no f32 tie with c(step-2)
            const bool ok_in = fabsf(dlo) < FLT_MAX && fabsf(dhi) < FLT_MAX && (inc || x != plo) && (step <= 1 || dlo != dg);
            const bool far = step >= nzi, zero = step == 0;
            const float d2 = fabsf(__fsub_rn(x, E.cl2));
            const bool ok_far = fabsf(__fsub_rn(x, E.cl)) < d2 && (nzi < 3 || d2 < fabsf(__fsub_rn(x, E.cl3)));
            const bool zero_nan = (x == E.c0) || fabsf(__fsub_rn(x, E.c0)) < fabsf(__fsub_rn(x, E.c1));
            bool good = !zero && (far ? ok_far : ok_in);
            const float a = far ? E.cl : (inc ? plo : phi);
            const float b = far ? E.cl2 : (inc ? phi : plo);
            if (LOGM) // exact hits and levels without a usable host logarithm are the general search's business
                good = good && (!trusted || (x != a && x != b && s_xd[jj] > 0. && !isnan(s_xl[jj])));
            bool undefined = zero && zero_nan;
            if (ranged)
                undefined = undefined || !level_in_range(P, s_xd[jj], rlo, rhi);
            int first = far ? nzi - 1 : (inc ? step - 1 : step);
            int second = far ? nzi - 2 : (inc ? step : step - 1);
            first = (good && !undefined) ? first : 0;
            second = (good && !undefined) ? second : 0;
            A[u] = __ldg(din + static_cast<unsigned>(first) * stride);
            B[u] = __ldg(din + static_cast<unsigned>(second) * stride);
            ca[u] = undefined ? INFINITY : (good ? a : nan_f());
            cb[u] = b;
        }
#pragma unroll
        for (int u = 0; u < SLG; ++u) {
            const int jj = j0 + u;
            if (jj < nzo) {
                float r;
                if (isnan(ca[u]))
                    r = sweep_slow(C, sc, nzi, inc ? 1 : -1, din, s_xd[jj], s_xl[jj]);
                else if (ca[u] == INFINITY)
                    r = nan_f();
                else if (LOGM && trusted)
                    r = blend_levels<METHOD, true>(C, A[u], B[u], ca[u], cb[u], s_xd[jj], s_xl[jj]);
                else
                    r = blend_levels<METHOD, false>(C, A[u], B[u], ca[u], cb[u], s_xd[jj], s_xl[jj]);
                float* dst = dout + static_cast<unsigned>(jj) * stride;
                if (plain)
                    __stcs(dst, r);
                else
                    store_result(C, dst, r);
            }
        }
    }
}

// ---- hybrid-sigma source, launch-wide table -----------------------------------------------------------------------------------
// p_k(ps) = ap_k + b_k*ps depends on the column only through ps, so n(x_j) is a step function of ps.  One small kernel proves
// from the launch's ps range that every column is strictly monotone (as hybrid_direction() does on the host) and tabulates
// n(x_j) on a ps grid; the main kernel looks its column's bin up, corrects the count by +-1 steps against the exactly evaluated
// f32 column and emits.  No per-column walk at all: ~100 instructions per output value.  Target levels need not be sorted.
struct HybridTableHeader {
    double ps_lo, ps_hi, inv_width; // scaled surface pressure range the table (and the proof) covers
    int dir, nbins;
};

__host__ __device__ inline int hybrid_direction(const double* ap, const double* b, size_t nzi, double ps_lo, double ps_hi)
{
    // p_{k+1} - p_k is linear in ps: it keeps its sign on [ps_lo, ps_hi] if it has it at both ends; a margin of a few f32 ulps
    // makes the statement hold for the ROUNDED f32 column the kernels work on
    if (!(ps_lo <= ps_hi) || !(fabs(ps_lo) <= DBL_MAX) || !(fabs(ps_hi) <= DBL_MAX) || nzi < 2)
        return 0;
    bool inc = true, dec = true;
    for (size_t k = 0; k + 1 < nzi; ++k) {
        for (int e = 0; e < 2; ++e) {
            const double ps = e ? ps_hi : ps_lo;
            const double p0 = ap[k] + b[k] * ps, p1 = ap[k + 1] + b[k + 1] * ps;
            const double margin = 8.0 * 1.1920929e-7 * fmax(fabs(p0), fabs(p1)) + 1e-30;
            inc = inc && (p1 - p0 > margin);
            dec = dec && (p0 - p1 > margin);
        }
    }
    return inc ? 1 : (dec ? -1 : 0);
}

__device__ __forceinline__ float decode_ordered(int e) { return e >= 0 ? __int_as_float(e) : -__int_as_float(-e); }

// Histogram of the launch's surface pressure by the top bits of its ordered-int encoding (sign, exponent, 2 mantissa bits:
// quarter-octave bins).  NaN is skipped.  The table covers the populated region around the bulk of the data, so that a few
// outliers (fill values, uninitialised pad columns) cannot stretch it; columns outside go through the general search.
constexpr int PS_HIST_BINS = 2048, PS_HIST_SHIFT = 21;
__device__ __forceinline__ int ps_hist_key(float v)
{
    const int bits = __float_as_int(v);
    const int enc = bits >= 0 ? bits : -(bits & 0x7fffffff);
    return (enc >> PS_HIST_SHIFT) + PS_HIST_BINS / 2;
}

// `every`: one block of 256 values in `every` is looked at.  The histogram only has to find where the bulk of the data lies; a column
// whose ps falls outside the range found takes the general search (slower, never wrong), so a sample serves large launches.
// rowWidth / rowPitch: the pad columns of padded rows are skipped (rowPitch divides the inner size, so it divides the index too)
__global__ void __launch_bounds__(256) ps_hist_kernel(const float* __restrict__ ps, size_t n, unsigned* __restrict__ hist, unsigned every, unsigned rowWidth,
                                                      unsigned rowPitch)
{
    __shared__ unsigned sh[PS_HIST_BINS];
    for (int i = threadIdx.x; i < PS_HIST_BINS; i += 256)
        sh[i] = 0;
    __syncthreads();
    for (size_t base = blockIdx.x * 256ull * every; base < n; base += 256ull * every * gridDim.x) { // whole warps stay in step
        const size_t i = base + threadIdx.x;
        const float v = i < n ? __ldg(ps + i) : nan_f();
        const bool valid = !isnan(v) && !(rowPitch != 0 && static_cast<unsigned>(i % rowPitch) >= rowWidth);
        const int key = valid ? ps_hist_key(v) : -1;
        const unsigned peers = __match_any_sync(0xffffffffu, key); // the data sit in a handful of bins: one atomic per bin and warp
        if (valid && static_cast<int>(threadIdx.x & 31) == __ffs(peers) - 1)
            atomicAdd(&sh[key], static_cast<unsigned>(__popc(peers)));
    }
    __syncthreads();
    for (int i = threadIdx.x; i < PS_HIST_BINS; i += 256)
        if (sh[i])
            atomicAdd(&hist[i], sh[i]);
}

__global__ void __launch_bounds__(256) hybrid_table_kernel(const double* __restrict__ ap, const double* __restrict__ b, int nzi,
                                                           const double* __restrict__ levels, int nzo, const unsigned* __restrict__ hist,
                                                           double ps_scale, int nbins, HybridTableHeader* __restrict__ hdr,
                                                           unsigned short* __restrict__ tab)
{
    __shared__ unsigned sh[PS_HIST_BINS];
    __shared__ int s_dir;
    __shared__ double s_lo, s_w;
    __shared__ int s_first, s_last;
    __shared__ unsigned long long s_total;
    // the coefficients are read many times by one thread (the monotonicity proof) and by all (the searches): staged once, so that
    // the serial part does not pay a trip to the L2 per level (it made this kernel 24 us long: longer than a member's main kernel)
    extern __shared__ double s_coef[]; // ap[nzi], b[nzi]
    double* s_ap = s_coef;
    double* s_b = s_coef + nzi;
    for (int k = threadIdx.x; k < nzi; k += 256) {
        s_ap[k] = ap[k];
        s_b[k] = b[k];
    }
    if (threadIdx.x == 0) {
        s_first = PS_HIST_BINS;
        s_last = -1;
        s_total = 0;
    }
    __syncthreads();
    {
        unsigned long long part = 0;
        int first = PS_HIST_BINS, last = -1;
        for (int i = threadIdx.x; i < PS_HIST_BINS; i += 256) {
            const unsigned c = hist[i];
            sh[i] = c;
            if (c) {
                part += c;
                first = min(first, i);
                last = max(last, i);
            }
        }
        if (part) {
            atomicAdd(&s_total, part);
            atomicMin(&s_first, first);
            atomicMax(&s_last, last);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned long long total = s_total;
        int dir = 0;
        double lo = 0., hi = 0.;
        if (total) {
            // smallest run of bins holding 90 % of the values, then outwards over every neighbouring populated bin
            const unsigned long long need = total - total / 10;
            int best_l = s_first, best_r = s_last, l = s_first;
            unsigned long long acc = 0;
            for (int r = s_first; r <= s_last; ++r) {
                acc += sh[r];
                while (acc - sh[l] >= need) {
                    acc -= sh[l];
                    ++l;
                }
                if (acc >= need && r - l < best_r - best_l) {
                    best_l = l;
                    best_r = r;
                }
            }
            while (best_l > 0 && sh[best_l - 1])
                --best_l;
            while (best_r < PS_HIST_BINS - 1 && sh[best_r + 1])
                ++best_r;
            // bin edges back to floats: the smallest and the largest encoding of the run
            const long long e_lo = (static_cast<long long>(best_l) - PS_HIST_BINS / 2) << PS_HIST_SHIFT;
            const long long e_hi = ((static_cast<long long>(best_r) - PS_HIST_BINS / 2 + 1) << PS_HIST_SHIFT) - 1;
            const long long cap = 0x7f800000ll;
            const int i_lo = static_cast<int>(e_lo < -cap ? -cap : e_lo), i_hi = static_cast<int>(e_hi > cap ? cap : e_hi);
            const double a0 = static_cast<double>(decode_ordered(i_lo)) * ps_scale, a1 = static_cast<double>(decode_ordered(i_hi)) * ps_scale;
            lo = fmin(a0, a1);
            hi = fmax(a0, a1);
            dir = hybrid_direction(s_ap, s_b, nzi, lo, hi);
        }
        const double w = dir ? (hi - lo) / nbins : 0.;
        s_dir = dir;
        s_lo = lo;
        s_w = w;
        hdr->ps_lo = lo;
        hdr->ps_hi = hi;
        hdr->inv_width = (w > 0 && 1. / w <= DBL_MAX) ? 1. / w : 0.;
        hdr->dir = dir;
        hdr->nbins = nbins;
    }
    __syncthreads();
    const int dir = s_dir;
    for (int idx = threadIdx.x; idx < nzo * nbins; idx += 256) {
        const int j = idx / nbins, i = idx - j * nbins;
        int n = 0;
        if (dir) {
            const double ps = s_lo + (i + 0.5) * s_w;
            const float sg = dir > 0 ? 1.f : -1.f;
            const float x = sg * __double2float_rn(levels[j]);
            int lo = 0, hi = nzi; // count of k with sg*c(k) <= x: c is increasing in walk order
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                const float v = sg * __double2float_rn(__dadd_rn(s_ap[mid], __dmul_rn(s_b[mid], ps)));
                if (v <= x)
                    lo = mid + 1;
                else
                    hi = mid;
            }
            n = lo;
        }
        tab[idx] = static_cast<unsigned short>(n);
    }
}

constexpr int HT = 128;       // threads (columns) per block of the table kernel's consumer
constexpr int HT_CHUNKS = 4;  // runs of HT columns one block works through with the same staged tables
constexpr unsigned CODE_NAN = 0u, CODE_BLEND = 1u << 14, CODE_SLOW = 2u << 14, CODE_UP = 1u << 13; // | first (13 bits)

// what value_from_step() decides, without fetching anything: NaN / general search / blend(first, first +- 1)
__device__ __forceinline__ unsigned classify_step(const PairCol& c, int nzi, bool inc, const ColumnEnds& E, int step, float x, float plo, float phi)
{
    if (step == 0)
        return ((x == E.c0) || fabsf(__fsub_rn(x, E.c0)) < fabsf(__fsub_rn(x, E.c1))) ? CODE_NAN : CODE_SLOW;
    if (step == nzi) {
        const float d2 = fabsf(__fsub_rn(x, E.cl2));
        const bool fast = fabsf(__fsub_rn(x, E.cl)) < d2 && (nzi < 3 || d2 < fabsf(__fsub_rn(x, E.cl3)));
        return fast ? (CODE_BLEND | static_cast<unsigned>(nzi - 1)) : CODE_SLOW; // second = first - 1
    }
    const float dlo = __fsub_rn(x, plo), dhi = __fsub_rn(x, phi);
    bool fast = fabsf(dlo) < FLT_MAX && fabsf(dhi) < FLT_MAX && (inc || x != plo);
    if (fast && step > 1)
        fast = dlo != __fsub_rn(x, c(step - 2));
    if (!fast)
        return CODE_SLOW;
    return inc ? (CODE_BLEND | CODE_UP | static_cast<unsigned>(step - 1)) : (CODE_BLEND | static_cast<unsigned>(step));
}

// The table kernel's column loop.  Per target level: the table's count m for the column's ps bin, the two f32 column values around
// it, ONE correction step against them (the table is off by a step at most next to a bin's own transitions; the value fetched for
// the correction doubles as the tie guard's c(step-2)), the decision of classify_step() inlined for the two cases that make up the
// bulk -- an interior bracket and the extrapolation beyond the column's last level -- and the two bracket values fetched with plain
// loads.  LG levels are handled together in straight-line code so that their 2 * LG gathers are in flight at once.  Everything
// else (level before or equal to the first element, f32 difference ties, a second correction step, NaN levels) goes to the
// out-of-line general search, which decides by the reference's own scan.

// per-column constants of table_column()
struct TableColumn {
    float c0, cl3;      // first element, third from the end (0 for nzi < 3)
    unsigned stride;    // nInner (the launch guarantees nzi * nInner and nzo * nInner < 2^31: 32-bit offsets)
    bool plain;         // no valid range, no fill value: results are stored as they are
};

// LG target levels starting at j0 (FULL: all of them exist; otherwise the group is cut at nzo).  Straight-line code: the bracket
// decision is all predicate logic and selects, so that the 2 * LG gathers are in flight together.
template <int METHOD, bool INC, bool FULL>
__device__ __forceinline__ float* table_group(const OutCfg& C, const TableColumn& T, const PairCol& hc, int nzi, int nzo, int nbins, int dir, int j0,
                                              const unsigned short* mytab, const float* s_xf, const double* s_xd, const double* s_xl,
                                              const float* __restrict__ din, float* __restrict__ dst)
{
    constexpr bool LOGM = METHOD == FIMEXCU_VINT_METHOD_LOG;
    float A[LG], B[LG], ca[LG], cb[LG]; // ca = NaN marks a level for the general search (a bracket value is never NaN)
#pragma unroll
    for (int u = 0; u < LG; ++u) {
        const int jj = FULL ? j0 + u : min(j0 + u, nzo - 1); // a short last group repeats its last level (stores are skipped below)
        const float x = s_xf[jj];
        const int m = mytab[jj * nbins]; // clamped to [1, nzi - 1] when the table was staged
        const float pm1 = hc(m - 1), pm = hc(m);
        const bool up = INC ? !(x < pm) : !(x > pm);             // x lies at or after c(m) (also true for NaN: sent to the general search)
        const bool down = !up && (INC ? x < pm1 : x > pm1);      // x lies before c(m-1) in walk order
        const float ce = hc(min(max(up ? m + 1 : m - 2, 0), nzi - 1));
        // one correction step: the bracket c(step-1), c(step) and, unless the step went down, c(step-2)
        const int step = m + (up ? 1 : 0) - (down ? 1 : 0);
        const float plo = up ? pm : (down ? ce : pm1);
        const float phi = up ? ce : (down ? pm1 : pm);
        const float guard = up ? pm1 : ce;
        const float dlo = __fsub_rn(x, plo), dhi = __fsub_rn(x, phi), dg = __fsub_rn(x, guard);
        // interior: the bracket must hold x (c(step-1) <= x < c(step) in walk order), the differences must be finite, and the rounded
        // differences to c(step-1) and c(step-2) must differ -- without c(step-2) at hand they do whenever |x - c(step-1)| <
        // |c(step-1)| / 8, because c(step-2) is at least one f32 spacing of c(step-1) away
        bool ok_in = INC ? (plo <= x && x < phi) : (plo > x && x > phi);
        ok_in = ok_in && fabsf(dlo) < FLT_MAX && fabsf(dhi) < FLT_MAX;
        const bool tie_free = down ? fabsf(dlo) < 0.125f * fabsf(plo) : dlo != dg;
        ok_in = ok_in && (step <= 1 || tie_free);
        // beyond the last element (step == nzi, reached only by a step up from m == nzi - 1: plo = c(nzi-1), guard = c(nzi-2))
        const bool far = step >= nzi;
        const bool ok_far = fabsf(dlo) < fabsf(dg) && (nzi < 3 || fabsf(dg) < fabsf(__fsub_rn(x, T.cl3)));
        bool ok = !isnan(x) && x != T.c0 && step >= 1 && (far ? ok_far : ok_in);
        const float a = (INC || far) ? plo : phi;
        const float b = far ? guard : (INC ? phi : plo);
        if (LOGM)
            ok = ok && x != a && x != b; // an exact hit has the weight 0 or 1 exactly: not the table quotient's business
        int first = (INC || far) ? step - 1 : step;
        int second = far ? step - 2 : (INC ? step : step - 1);
        first = ok ? first : 0;
        second = ok ? second : 0;
        A[u] = __ldg(din + static_cast<unsigned>(first) * T.stride);
        B[u] = __ldg(din + static_cast<unsigned>(second) * T.stride);
        ca[u] = ok ? a : nan_f();
        cb[u] = b;
    }
#pragma unroll
    for (int u = 0; u < LG; ++u) {
        const int jj = j0 + u;
        if (FULL || jj < nzo) {
            float r;
            if (isnan(ca[u]))
                r = sweep_slow(C, hc, nzi, dir, din, s_xd[jj], s_xl[jj]);
            else
                r = blend_levels<METHOD, LOGM>(C, A[u], B[u], ca[u], cb[u], s_xd[jj], s_xl[jj]);
            if (T.plain)
                __stcs(dst, r);
            else
                store_result(C, dst, r);
            dst += T.stride;
        }
    }
    return dst;
}

template <int METHOD, bool INC>
__device__ __forceinline__ void table_column(const VertParams& P, const OutCfg& C, const PairCol& hc, int nzi, int nzo, int nbins, int dir,
                                             const unsigned short* mytab, const float* s_xf, const double* s_xd, const double* s_xl,
                                             const float* __restrict__ din, float* __restrict__ dout)
{
    TableColumn T;
    T.c0 = hc(0);
    T.cl3 = nzi >= 3 ? hc(nzi - 3) : 0.f;
    T.stride = static_cast<unsigned>(P.nInner);
    T.plain = isnan(C.valid_min) && isnan(C.valid_max) && isnan(C.fill);
    if (METHOD == FIMEXCU_VINT_METHOD_LOG) {
        // the column is strictly monotone (the launch's proof), so every element lies between its ends: if both are positive normal
        // numbers, all are, and the per-level tests of the logarithm drop out (blend_levels<.., TRUSTED>)
        const unsigned b0 = __float_as_uint(T.c0), b1 = __float_as_uint(hc(nzi - 1));
        if (b0 - 0x00800000u >= 0x7f000000u || b1 - 0x00800000u >= 0x7f000000u) {
            column_by_general_scan(P, C, hc, nzi, nzo, s_xd, s_xl, 0., 0., din, dout);
            return;
        }
    }
    int j0 = 0;
    for (; j0 + LG <= nzo; j0 += LG)
        dout = table_group<METHOD, INC, true>(C, T, hc, nzi, nzo, nbins, dir, j0, mytab, s_xf, s_xd, s_xl, din, dout);
    if (j0 < nzo)
        table_group<METHOD, INC, false>(C, T, hc, nzi, nzo, nbins, dir, j0, mytab, s_xf, s_xd, s_xl, din, dout);
}

// Per column: bin of the table, then table_column() over the target levels.
template <int METHOD> // compile-time interpolation method: the other methods' arithmetic is not in the kernel at all
__global__ void __launch_bounds__(HT, 8) vertical_hybrid_table_kernel(const VertParams P, const HybridTableHeader* __restrict__ hdr,
                                                                     const unsigned short* __restrict__ tab)
{
    extern __shared__ double smem_d[];
    const int nzi = static_cast<int>(P.nzi), nzo = static_cast<int>(P.nzo);
    const int nbins = hdr->nbins, dir = hdr->dir;
    double2* s_apb = reinterpret_cast<double2*>(smem_d);
    double* s_xd = smem_d + 2 * nzi;
    double* s_xl = s_xd + nzo;
    float* s_xf = reinterpret_cast<float*>(s_xl + nzo);
    unsigned short* s_tab = reinterpret_cast<unsigned short*>(s_xf + nzo + (nzo & 1));
    __shared__ double2 s_log[METHOD == FIMEXCU_VINT_METHOD_LOG ? 128 : 1];
    if (METHOD == FIMEXCU_VINT_METHOD_LOG)
        for (int i = threadIdx.x; i < 128; i += HT)
            s_log[i] = P.logtab[i]; // made once per device (log_table_device)
    for (int k = threadIdx.x; k < nzi; k += HT)
        s_apb[k] = make_double2(P.ap[k], P.b[k]);
    for (int k = threadIdx.x; k < nzo; k += HT) {
        const double x = P.levels[k];
        const double xl = P.xlog ? P.xlog[k] : nan("");
        s_xd[k] = x;
        s_xl[k] = xl;
        // LOG: a level without a usable host logarithm (non-positive, NaN) is marked NaN here, which sends it to the general search
        s_xf[k] = (METHOD == FIMEXCU_VINT_METHOD_LOG && !(x > 0. && !isnan(xl))) ? nan_f() : __double2float_rn(x);
    }
    if (dir) {
        const int words = (nzo * nbins + 1) / 2; // the table buffer is padded to a whole number of 32-bit words
        const unsigned lo2 = 0x00010001u, hi2 = static_cast<unsigned>(nzi - 1) * 0x00010001u; // counts clamped to [1, nzi - 1] once, here
        for (int i = threadIdx.x; i < words; i += HT)
            reinterpret_cast<unsigned*>(s_tab)[i] = __vminu2(__vmaxu2(__ldg(reinterpret_cast<const unsigned*>(tab) + i), lo2), hi2);
    }
    __syncthreads();
    OutCfg C = out_cfg(P);
    C.method = METHOD;
    C.logtab = METHOD == FIMEXCU_VINT_METHOD_LOG ? s_log : nullptr;
    const double ps_lo = hdr->ps_lo, ps_hi = hdr->ps_hi, inv_width = hdr->inv_width;
    const size_t outer = blockIdx.y; // grid: (blocks of HT_CHUNKS * HT columns, outer)
    // the staged tables serve HT_CHUNKS runs of HT neighbouring columns (staging them is a sixth of a single run's instructions)
#pragma unroll 1
    for (int chunk = 0; chunk < HT_CHUNKS; ++chunk) {
        const size_t inner = (blockIdx.x * static_cast<size_t>(HT_CHUNKS) + chunk) * HT + threadIdx.x;
        if (inner >= P.nInner)
            return;
        if (pad_column(P, inner))
            continue;
        const float* din = P.in + outer * P.nzi * P.nInner + inner;
        float* dout = P.out + outer * P.nzo * P.nInner + inner;
        const double ps = __dmul_rn(static_cast<double>(__ldg(P.ps + (P.nOuterPs == 1 ? 0 : outer) * P.nInner + inner)), P.ps_scale);
        const PairCol hc{s_apb, ps};
        if (dir == 0 || !(ps >= ps_lo && ps <= ps_hi)) { // no proof for this launch, or a column outside its range (NaN, outlier)
            if (isnan(ps)) // ap + b * NaN is NaN at every level (the targets of a regridded slab that lie outside its source grid)
                column_undefined(C, nzo, P.nInner, dout);
            else
                column_by_general_scan(P, C, hc, nzi, nzo, s_xd, s_xl, 0., 0., din, dout);
            continue;
        }
        const double t = (ps - ps_lo) * inv_width;
        const int bin = min(max(static_cast<int>(t), 0), nbins - 1);
        if (dir > 0)
            table_column<METHOD, true>(P, C, hc, nzi, nzo, nbins, dir, s_tab + bin, s_xf, s_xd, s_xl, din, dout);
        else
            table_column<METHOD, false>(P, C, hc, nzi, nzo, nbins, dir, s_tab + bin, s_xf, s_xd, s_xl, din, dout);
    }
}

// min / max of the surface pressure of a launch (NaN ignored): lets the host prove monotonicity for all columns at once
__global__ void __launch_bounds__(256) ps_minmax_kernel(const float* __restrict__ ps, size_t n, float* __restrict__ mm, unsigned rowWidth, unsigned rowPitch)
{
    float lo = INFINITY, hi = -INFINITY;
    for (size_t i = blockIdx.x * 256ull + threadIdx.x; i < n; i += 256ull * gridDim.x) {
        if (rowPitch != 0 && static_cast<unsigned>(i % rowPitch) >= rowWidth)
            continue; // pad column of a padded row
        const float v = ps[i];
        lo = fminf(lo, v);
        hi = fmaxf(hi, v);
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, s));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, s));
    }
    if ((threadIdx.x & 31) == 0) { // floats order like their sign-magnitude integers when handled per sign: use CAS-free int trick on non-negative / negative
        atomicMin(reinterpret_cast<int*>(mm), lo >= 0 ? __float_as_int(lo) : -__float_as_int(-lo));
        atomicMax(reinterpret_cast<int*>(mm) + 1, hi >= 0 ? __float_as_int(hi) : -__float_as_int(-hi));
    }
}

// sorted fixed levels -> sweep kernel (FIMEXCU_VERTICAL_NO_SWEEP=1 keeps the per-level search kernels, for A/B measurements)
static size_t sweep_smem(const VertParams& P)
{
    return ((P.vcoord_in ? 0 : 2 * P.nzi) + 2 * P.nzo) * sizeof(double) + (P.nzo + 2 + (P.nzo & 1)) * sizeof(float) +
           (P.nzo + (P.nzo & 1)) * VT * sizeof(unsigned short) + (P.vcoord_in ? 3 * P.nzo * VT * sizeof(float) : 0);
}
static bool sweep_applies(const VertParams& P)
{
    static const bool disabled = std::getenv("FIMEXCU_VERTICAL_NO_SWEEP") != nullptr;
    return !disabled && P.levels && !P.vcoord_out && P.levels_dir != 0 && P.nzi < 65535 && P.nzo < 65535 && sweep_smem(P) <= 40 * 1024 &&
           P.nzi * P.nInner < (size_t(1) << 31) && P.nzo * P.nInner < (size_t(1) << 31); // 32-bit level offsets in the kernel
}

static size_t search_smem(const VertParams& P) // level list (f64, log, f32) + the block's coordinate columns [nzi][VT]
{
    return 2 * P.nzo * sizeof(double) + (P.nzo + (P.nzo & 1)) * sizeof(float) + P.nzi * VT * sizeof(float);
}
static bool search_applies(const VertParams& P)
{
    static const bool disabled = std::getenv("FIMEXCU_VERTICAL_NO_SEARCH") != nullptr;
    return !disabled && P.vcoord_in && P.levels && !P.vcoord_out && P.nzi >= 2 && P.nzo >= 1 && search_smem(P) <= 96 * 1024 &&
           P.nzi * P.nInner < (size_t(1) << 31) && P.nzo * P.nInner < (size_t(1) << 31); // 32-bit level offsets in the kernel
}

// direction of a fixed level list: +1 non-decreasing, -1 non-increasing, 0 unordered / NaN
static int levels_direction(const double* lv, size_t n)
{
    bool up = true, down = true;
    for (size_t k = 0; k < n; ++k) {
        if (std::isnan(lv[k]))
            return 0;
        if (k) {
            up = up && lv[k] >= lv[k - 1];
            down = down && lv[k] <= lv[k - 1];
        }
    }
    return up ? 1 : (down ? -1 : 0);
}

// the 128-entry logarithm table of fxcu_logtable.h in device memory: made once per device from the host's libm, never freed
static const double2* log_table_device()
{
    static std::mutex mu;
    static double2* tabs[64] = {};
    int dev = -1;
    if (!cuda_ok(cudaGetDevice(&dev), "cudaGetDevice", __FILE__, __LINE__) || dev < 0 || dev >= 64)
        return nullptr;
    std::lock_guard<std::mutex> g(mu);
    if (!tabs[dev]) {
        double2 h[128];
        for (int i = 0; i < 128; ++i) {
            const double c = 1.0 + (i + 0.5) / 128.0;
            h[i] = make_double2(1.0 / c, std::log(c));
        }
        double2* d = nullptr;
        if (!cuda_ok(cudaMalloc(&d, sizeof(h)), "cudaMalloc", __FILE__, __LINE__))
            return nullptr;
        if (!cuda_ok(cudaMemcpy(d, h, sizeof(h), cudaMemcpyHostToDevice), "cudaMemcpy", __FILE__, __LINE__)) {
            cudaFree(d);
            return nullptr;
        }
        tabs[dev] = d;
    }
    return tabs[dev];
}

// device copy of a coefficient set (hybrid ap, b, target levels and their logarithms), per host thread and device; nullptr when the
// set cannot be kept (first sight of it: it is uploaded here, synchronously, and kept; more than 8 sets alive: not cached)
static const double* cached_coefficients(const std::vector<double>& h)
{
    struct Entry {
        int device;
        std::vector<double> host;
        double* dev;
    };
    thread_local std::vector<Entry> cache;
    int dev = -1;
    if (h.empty() || cudaGetDevice(&dev) != cudaSuccess)
        return nullptr;
    for (const Entry& e : cache)
        if (e.device == dev && e.host.size() == h.size() && std::memcmp(e.host.data(), h.data(), h.size() * sizeof(double)) == 0)
            return e.dev;
    if (cache.size() >= 8)
        return nullptr;
    double* d = nullptr;
    if (cudaMalloc(&d, h.size() * sizeof(double)) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    if (cudaMemcpy(d, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess) {
        cudaGetLastError();
        cudaFree(d);
        return nullptr;
    }
    cache.push_back(Entry{dev, h, d});
    return d;
}

static int launch_vertical(const VertParams& P, cudaStream_t s)
{
    const size_t ncol = P.nOuter * P.nInner;
    if (ncol == 0 || P.nzo == 0)
        return FIMEXCU_OK;
    if (P.nzi == 0 || P.nzi > 0x3fffffffull / VT)
        FXCU_FAIL("fimexcu_vertical: invalid number of input levels");
    if (P.method < 0 || P.method > FIMEXCU_VINT_METHOD_LIN_CONST_EXTRA)
        FXCU_FAIL("fimexcu_vertical: unknown vertical interpolation method");
    const unsigned blocks = static_cast<unsigned>((ncol + VT - 1) / VT);
    if (search_applies(P)) { // explicit coordinate, fixed levels (sorted or not): scan + bisection
        int rounds = 1;
        while ((size_t(1) << rounds) <= P.nzi)
            ++rounds;
        const size_t smem = search_smem(P);
        if (smem > 40 * 1024) { // (static + dynamic beyond 48 KB needs the opt-in; the limit is a property of the function ON A DEVICE)
            static std::atomic<unsigned long long> raised{0};
            int dev = 0;
            FXCU_CUDA(cudaGetDevice(&dev));
            const unsigned long long bit = 1ull << (dev & 63);
            if (!(raised.fetch_or(bit) & bit)) {
                cudaFuncSetAttribute(vertical_search_kernel<FIMEXCU_VINT_METHOD_LOG>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
                cudaFuncSetAttribute(vertical_search_kernel<FIMEXCU_VINT_METHOD_LIN>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
                cudaFuncSetAttribute(vertical_search_kernel<-1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
            }
        }
        if (P.method == FIMEXCU_VINT_METHOD_LOG)
            vertical_search_kernel<FIMEXCU_VINT_METHOD_LOG><<<blocks, VT, smem, s>>>(P, rounds);
        else if (P.method == FIMEXCU_VINT_METHOD_LIN)
            vertical_search_kernel<FIMEXCU_VINT_METHOD_LIN><<<blocks, VT, smem, s>>>(P, rounds);
        else
            vertical_search_kernel<-1><<<blocks, VT, smem, s>>>(P, rounds);
        FXCU_CUDA(cudaGetLastError());
        return FIMEXCU_OK;
    }
    if (sweep_applies(P)) {
        const size_t smem = sweep_smem(P);
        if (!P.vcoord_in)
            vertical_sweep_kernel<true><<<blocks, VT, smem, s>>>(P);
        else if (P.method == FIMEXCU_VINT_METHOD_LOG)
            vertical_sweep_kernel<false, FIMEXCU_VINT_METHOD_LOG><<<blocks, VT, smem, s>>>(P);
        else if (P.method == FIMEXCU_VINT_METHOD_LIN)
            vertical_sweep_kernel<false, FIMEXCU_VINT_METHOD_LIN><<<blocks, VT, smem, s>>>(P);
        else
            vertical_sweep_kernel<false><<<blocks, VT, smem, s>>>(P);
        FXCU_CUDA(cudaGetLastError());
        return FIMEXCU_OK;
    }
    if (!P.vcoord_in) {
        const size_t smem = 2 * P.nzi * sizeof(double);
        if (smem > 48 * 1024)
            FXCU_FAIL("fimexcu_vertical_hybrid_pressure: more than 3000 input levels are not supported");
        vertical_kernel<1><<<blocks, VT, smem, s>>>(P);
    } else {
        const size_t smem = P.nzi * VT * sizeof(float);
        if (smem <= 200 * 1024) {
            if (smem > 48 * 1024)
                FXCU_CUDA(cudaFuncSetAttribute(vertical_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
            vertical_kernel<0><<<blocks, VT, smem, s>>>(P);
        } else {
            vertical_kernel<2><<<blocks, VT, 0, s>>>(P);
        }
    }
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}

// log / log-log of a fixed target level with the host's libm, exactly the scalars the reference computes
// (interpolation.c:521-546); NaN where the kernel has to decide itself (non-positive level -> undefined result)
static double host_level_log(int method, double x)
{
    if (!(x > 0))
        return std::numeric_limits<double>::quiet_NaN();
    if (method == FIMEXCU_VINT_METHOD_LOG)
        return std::log(x);
    if (method == FIMEXCU_VINT_METHOD_LOGLOG)
        return std::log(std::log(x + 2.7182818284590452354));
    return std::numeric_limits<double>::quiet_NaN();
}

static float narrow_or_nan(double v) { return std::isnan(v) ? std::numeric_limits<float>::quiet_NaN() : static_cast<float>(v); }

// ---- two-level kernels on whole fields (interpolation.c:416-546) ----
// VEC (all three arrays on 16 bytes): four values per thread and access; `out` may be A (in place: a thread reads its values
// before it writes them, and nobody else touches them)
template <bool VEC>
__global__ void __launch_bounds__(256) lin_kernel(const float* A, const float* B, float* out, size_t n, float f)
{
    const size_t nv = VEC ? n / 4 : 0;
    if constexpr (VEC)
        for (size_t k = blockIdx.x * 256ull + threadIdx.x; k < nv; k += 256ull * gridDim.x) {
            const float4 a = reinterpret_cast<const float4*>(A)[k], b = reinterpret_cast<const float4*>(B)[k];
            reinterpret_cast<float4*>(out)[k] = make_float4(__fadd_rn(a.x, __fmul_rn(f, __fsub_rn(b.x, a.x))), __fadd_rn(a.y, __fmul_rn(f, __fsub_rn(b.y, a.y))),
                                                            __fadd_rn(a.z, __fmul_rn(f, __fsub_rn(b.z, a.z))), __fadd_rn(a.w, __fmul_rn(f, __fsub_rn(b.w, a.w))));
        }
    for (size_t i = nv * 4 + blockIdx.x * 256ull + threadIdx.x; i < n; i += 256ull * gridDim.x)
        out[i] = __fadd_rn(A[i], __fmul_rn(f, __fsub_rn(B[i], A[i])));
}

// kind: 0 linear, 1 weak, 2 no, 3 const.  Decides on the host exactly like the reference which of copy-A / copy-B / blend / NaN
// applies; device pointers, enqueued on `s`
static int two_level_dev(int kind, const float* dA, const float* dB, float* dO, size_t n, double a, double b, double x, cudaStream_t s)
{
    const float f = (a == b) ? 0.f : static_cast<float>((x - a) / (b - a));
    int action; // 0 copy A, 1 copy B, 2 blend, 3 NaN
    if (kind == 3)
        action = (f >= 1) ? 1 : ((f <= 0) ? 0 : 2);
    else if (f == 0)
        action = 0;
    else if (f == 1)
        action = 1;
    else if (kind == 1)
        action = (f >= -1.f && f <= 2.f) ? 2 : 3;
    else if (kind == 2)
        action = (f >= 0.f && f <= 1.f) ? 2 : 3;
    else
        action = 2;
    if (n == 0)
        return FIMEXCU_OK;
    const size_t bytes = n * sizeof(float);
    const unsigned blocks = static_cast<unsigned>(std::min<size_t>((n + 255) / 256, static_cast<size_t>(sm_count()) * 16));
    switch (action) {
    case 0:
        if (dO != dA)
            FXCU_CUDA(cudaMemcpyAsync(dO, dA, bytes, cudaMemcpyDeviceToDevice, s));
        break;
    case 1:
        if (dO != dB)
            FXCU_CUDA(cudaMemcpyAsync(dO, dB, bytes, cudaMemcpyDeviceToDevice, s));
        break;
    case 2:
        if (((reinterpret_cast<uintptr_t>(dA) | reinterpret_cast<uintptr_t>(dB) | reinterpret_cast<uintptr_t>(dO)) & 15u) == 0)
            lin_kernel<true><<<std::max(1u, (blocks + 3) / 4), 256, 0, s>>>(dA, dB, dO, n, f);
        else
            lin_kernel<false><<<blocks, 256, 0, s>>>(dA, dB, dO, n, f);
        break;
    default: FXCU_CUDA(cudaMemsetAsync(dO, 0xff, bytes, s)); break; // all-ones is a quiet NaN
    }
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}

// vertical-interpolation method -> (kind, transformed coordinates) of the reference's seven two-level functions
// (interpolation.c:416-546); returns false (error set) where the reference returns MIFI_ERROR and leaves the output untouched
static bool two_level_plan(int method, double& a, double& b, double& x, int& kind)
{
    switch (method) {
    case FIMEXCU_VINT_METHOD_NN: kind = 3; a = 0.; b = 1.; x = -1.; return true; // f < 0 with the "const" rule: copy A
    case FIMEXCU_VINT_METHOD_LIN: kind = 0; return true;
    case FIMEXCU_VINT_METHOD_LIN_WEAK_EXTRA: kind = 1; return true;
    case FIMEXCU_VINT_METHOD_LIN_NO_EXTRA: kind = 2; return true;
    case FIMEXCU_VINT_METHOD_LIN_CONST_EXTRA: kind = 3; return true;
    case FIMEXCU_VINT_METHOD_LOGLOG:
        if (a <= 0 || b <= 0 || x <= 0) {
            set_error("mifi_get_values_log_log_f: non-positive coordinate");
            return false;
        }
        a = std::log(a + 2.7182818284590452354);
        b = std::log(b + 2.7182818284590452354);
        x = std::log(x + 2.7182818284590452354);
        [[fallthrough]]; // mifi_get_values_log_log_f calls mifi_get_values_log_f (interpolation.c:537-546)
    case FIMEXCU_VINT_METHOD_LOG:
        if (a <= 0 || b <= 0 || x <= 0) {
            set_error("mifi_get_values_log_f: non-positive coordinate");
            return false; // output untouched, like interpolation.c:524-526
        }
        // the three logarithms are host-side scalars (as in the reference); the field arithmetic runs on the device
        a = std::log(a);
        b = std::log(b);
        x = std::log(x);
        kind = 0;
        return true;
    default: set_error("fimexcu_get_values_two_level: unknown method"); return false;
    }
}

static int two_level_host(int method, const float* A, const float* B, float* out, size_t n, double a, double b, double x)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    int kind = 0;
    if (!two_level_plan(method, a, b, x, kind))
        return FIMEXCU_ERROR;
    if (n == 0)
        return FIMEXCU_OK;
    const size_t bytes = n * sizeof(float);
    DevBuf dA, dB, dO;
    if (!dA.alloc(bytes, nullptr) || !dB.alloc(bytes, nullptr) || !dO.alloc(bytes, nullptr))
        return FIMEXCU_ERROR;
    // the product has no CPU path: even the plain copies travel through the device
    FXCU_CUDA(cudaMemcpyAsync(dA.p, A, bytes, cudaMemcpyHostToDevice, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(dB.p, B, bytes, cudaMemcpyHostToDevice, nullptr));
    if (two_level_dev(kind, dA.as<float>(), dB.as<float>(), dO.as<float>(), n, a, b, x, nullptr) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(out, dO.p, bytes, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

} // namespace fxcu

using namespace fxcu;

// rows of a padded slab: validates and stores the pair (0 / 0 = every inner index is a column)
static int set_rows(VertParams& P, size_t nInner, size_t rowWidth, size_t rowPitch)
{
    if (rowWidth == 0 && rowPitch == 0)
        return FIMEXCU_OK;
    if (rowWidth == 0 || rowWidth > rowPitch || nInner % rowPitch != 0 || nInner >= (size_t(1) << 32))
        FXCU_FAIL("vertical interpolation: rowWidth / rowPitch must satisfy 0 < rowWidth <= rowPitch, rowPitch dividing nInner (< 2^32)");
    P.rowWidth = static_cast<unsigned>(rowWidth);
    P.rowPitch = rowWidth == rowPitch ? 0u : static_cast<unsigned>(rowPitch); // no pad: nothing to test per column
    if (P.rowPitch == 0)
        P.rowWidth = 0;
    return FIMEXCU_OK;
}

extern "C" {

int fimexcu_vertical_interpolate_rows_dev_f(int method, const float* d_in, const float* d_vcoord_in, size_t nOuter, size_t nzi, size_t nInner,
                                       const double* levels_host, const float* d_vcoord_out, size_t nzo, const double* d_vmin,
                                       const double* d_vmax, double valid_min, double valid_max, float badIn, float fillOut, float* d_out,
                                       size_t rowWidth, size_t rowPitch, void* stream)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    {
        VertParams chk{};
        if (set_rows(chk, nInner, rowWidth, rowPitch) != FIMEXCU_OK)
            return FIMEXCU_ERROR;
    }
    if (!levels_host && !d_vcoord_out)
        FXCU_FAIL("fimexcu_vertical_interpolate: need either fixed levels or an output vertical coordinate");
    cudaStream_t s = as_stream(stream);
    DevBuf lv;
    VertParams P{};
    if (levels_host) {
        if (!lv.alloc(std::max<size_t>(2 * nzo, 1) * sizeof(double), s))
            return FIMEXCU_ERROR;
        std::vector<double> hl(2 * nzo);
        for (size_t k = 0; k < nzo; ++k) {
            hl[k] = levels_host[k];
            hl[nzo + k] = host_level_log(method, levels_host[k]);
        }
        FXCU_CUDA(cudaMemcpyAsync(lv.p, hl.data(), 2 * nzo * sizeof(double), cudaMemcpyHostToDevice, s));
        FXCU_CUDA(cudaStreamSynchronize(s)); // `hl` is a temporary
        P.levels = lv.as<double>();
        P.levels_dir = levels_direction(levels_host, nzo);
        if (method == FIMEXCU_VINT_METHOD_LOG || method == FIMEXCU_VINT_METHOD_LOGLOG)
            P.xlog = lv.as<double>() + nzo;
        P.vcoord_out = nullptr;
    } else {
        P.vcoord_out = d_vcoord_out;
    }
    set_rows(P, nInner, rowWidth, rowPitch);
    P.method = method;
    P.in = d_in;
    P.vcoord_in = d_vcoord_in;
    P.nOuter = nOuter;
    P.nzi = nzi;
    P.nInner = nInner;
    P.nzo = nzo;
    P.vmin = d_vmin;
    P.vmax = d_vmax;
    P.valid_min = narrow_or_nan(valid_min);
    P.valid_max = narrow_or_nan(valid_max);
    P.bad = badIn;
    P.fill = fillOut;
    P.out = d_out;
    P.logtab = (method == FIMEXCU_VINT_METHOD_LOG && levels_host) ? log_table_device() : nullptr; // nullptr: the kernels use the series / log() forms
    if (!d_vcoord_in)
        FXCU_FAIL("fimexcu_vertical_interpolate: null input vertical coordinate");
    const int rc = launch_vertical(P, s);
    if (levels_host)
        FXCU_CUDA(cudaStreamSynchronize(s)); // levels_host may be a temporary of the caller
    return rc;
}

int fimexcu_vertical_interpolate_dev_f(int method, const float* d_in, const float* d_vcoord_in, size_t nOuter, size_t nzi, size_t nInner,
                                       const double* levels_host, const float* d_vcoord_out, size_t nzo, const double* d_vmin,
                                       const double* d_vmax, double valid_min, double valid_max, float badIn, float fillOut, float* d_out,
                                       void* stream)
{
    return fimexcu_vertical_interpolate_rows_dev_f(method, d_in, d_vcoord_in, nOuter, nzi, nInner, levels_host, d_vcoord_out, nzo, d_vmin, d_vmax, valid_min,
                                                   valid_max, badIn, fillOut, d_out, 0, 0, stream);
}

int fimexcu_vertical_interpolate_f(int method, const float* in, const float* vcoord_in, size_t nOuter, size_t nzi, size_t nInner,
                                   const double* levels, const float* vcoord_out, size_t nzo, const double* vmin, const double* vmax,
                                   double valid_min, double valid_max, float badIn, float fillOut, float* out)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    const size_t ncol = nOuter * nInner;
    const size_t nin = ncol * nzi, nout = ncol * nzo;
    if (nout == 0)
        return FIMEXCU_OK;
    DevBuf din, dvc, dvo, dmin, dmax, dout;
    if (!din.alloc(nin * 4, nullptr) || !dvc.alloc(nin * 4, nullptr) || !dout.alloc(nout * 4, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(din.p, in, nin * 4, cudaMemcpyHostToDevice, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(dvc.p, vcoord_in, nin * 4, cudaMemcpyHostToDevice, nullptr));
    if (!levels && vcoord_out) {
        if (!dvo.alloc(nout * 4, nullptr))
            return FIMEXCU_ERROR;
        FXCU_CUDA(cudaMemcpyAsync(dvo.p, vcoord_out, nout * 4, cudaMemcpyHostToDevice, nullptr));
    }
    if (vmin) {
        if (!dmin.alloc(ncol * 8, nullptr))
            return FIMEXCU_ERROR;
        FXCU_CUDA(cudaMemcpyAsync(dmin.p, vmin, ncol * 8, cudaMemcpyHostToDevice, nullptr));
    }
    if (vmax) {
        if (!dmax.alloc(ncol * 8, nullptr))
            return FIMEXCU_ERROR;
        FXCU_CUDA(cudaMemcpyAsync(dmax.p, vmax, ncol * 8, cudaMemcpyHostToDevice, nullptr));
    }
    if (fimexcu_vertical_interpolate_dev_f(method, din.as<float>(), dvc.as<float>(), nOuter, nzi, nInner, levels, levels ? nullptr : dvo.as<float>(), nzo,
                                           vmin ? dmin.as<double>() : nullptr, vmax ? dmax.as<double>() : nullptr, valid_min, valid_max, badIn, fillOut,
                                           dout.as<float>(), nullptr) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(out, dout.p, nout * 4, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

int fimexcu_vertical_hybrid_pressure_rows_dev_f(int method, const float* d_in, size_t nOuter, size_t nzi, size_t nInner, const double* ap_host,
                                                const double* b_host, const float* d_ps, size_t nOuterPs, double ps_scale, const double* levels_host,
                                                size_t nzo, double valid_min, double valid_max, float badIn, float fillOut, float* d_out,
                                                size_t rowWidth, size_t rowPitch, void* stream)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    {
        VertParams chk{};
        if (set_rows(chk, nInner, rowWidth, rowPitch) != FIMEXCU_OK)
            return FIMEXCU_ERROR;
    }
    if (!(nOuterPs == 1 || nOuterPs == nOuter))
        FXCU_FAIL("fimexcu_vertical_hybrid_pressure: ps must have 1 or nOuter outer entries");
    if (!levels_host || !ap_host || !b_host || !d_ps)
        FXCU_FAIL("fimexcu_vertical_hybrid_pressure: null argument");
    cudaStream_t s = as_stream(stream);
    const size_t nc = 2 * nzi + 2 * nzo;
    std::vector<double> h(nc);
    std::copy(ap_host, ap_host + nzi, h.begin());
    std::copy(b_host, b_host + nzi, h.begin() + nzi);
    std::copy(levels_host, levels_host + nzo, h.begin() + 2 * nzi);
    for (size_t k = 0; k < nzo; ++k)
        h[2 * nzi + nzo + k] = host_level_log(method, levels_host[k]);
    // A reader chain calls this once per variable and time step with the same coefficients and levels: the device copy of the
    // last few coefficient sets is kept per host thread and device (a few KB each, never freed), so that a call neither uploads
    // from pageable memory nor has to wait for its own completion before `h` goes away.
    const double* d_coef = cached_coefficients(h);
    DevBuf coef;
    const bool cached = d_coef != nullptr;
    if (!cached) {
        if (!coef.alloc(std::max<size_t>(nc, 1) * sizeof(double), s))
            return FIMEXCU_ERROR;
        FXCU_CUDA(cudaMemcpyAsync(coef.p, h.data(), nc * sizeof(double), cudaMemcpyHostToDevice, s));
        d_coef = coef.as<double>();
    }
    VertParams P{};
    set_rows(P, nInner, rowWidth, rowPitch);
    P.method = method;
    P.in = d_in;
    P.vcoord_in = nullptr;
    P.ap = d_coef;
    P.b = d_coef + nzi;
    P.levels = d_coef + 2 * nzi;
    P.levels_dir = levels_direction(levels_host, nzo);
    if (method == FIMEXCU_VINT_METHOD_LOG || method == FIMEXCU_VINT_METHOD_LOGLOG)
        P.xlog = d_coef + 2 * nzi + nzo;
    P.ps = d_ps;
    P.nOuterPs = nOuterPs;
    P.ps_scale = ps_scale;
    P.nOuter = nOuter;
    P.nzi = nzi;
    P.nInner = nInner;
    P.nzo = nzo;
    P.valid_min = narrow_or_nan(valid_min);
    P.valid_max = narrow_or_nan(valid_max);
    P.bad = badIn;
    P.fill = fillOut;
    P.out = d_out;
    P.logtab = method == FIMEXCU_VINT_METHOD_LOG ? log_table_device() : nullptr;
    if (method == FIMEXCU_VINT_METHOD_LOG && !P.logtab)
        return FIMEXCU_ERROR;
    const size_t ncol = nOuter * nInner, nps = nOuterPs * nInner;
    static const bool no_table = std::getenv("FIMEXCU_VERTICAL_NO_TABLE") != nullptr;
    const int nbins = static_cast<int>(std::min<size_t>(512, std::max<size_t>(8, 2048 / std::max<size_t>(nzo, 1))));
    const size_t tab_words = (nzo * nbins + 1) / 2;
    const size_t tab_smem = (2 * nzi + 2 * nzo) * sizeof(double) + (nzo + (nzo & 1)) * sizeof(float) + tab_words * 4;
    const int init[2] = {0x7f800000, static_cast<int>(0x80800000u)}; // +inf / "-inf" in the ordered-int encoding
    // (the consumer kernel addresses a column's levels with 32-bit offsets and takes the outer index from blockIdx.y)
    if (!no_table && ncol && nzo && nzi >= 2 && nzi < 8192 && nzo < 65535 && tab_smem <= 28 * 1024 && nOuter <= 65535 &&
        nzi * nInner < (size_t(1) << 31) && nzo * nInner < (size_t(1) << 31)) {
        // launch-wide table path: ps range -> monotonicity proof + table (device side, no host round trip) -> consumer kernel
        DevBuf scratch;
        const size_t hdr_off = PS_HIST_BINS * sizeof(unsigned), tab_off = hdr_off + sizeof(HybridTableHeader);
        if (!scratch.alloc(tab_off + tab_words * 4, s))
            return FIMEXCU_ERROR;
        char* base = static_cast<char*>(scratch.p);
        FXCU_CUDA(cudaMemsetAsync(base, 0, hdr_off, s));
        if (nps) {
            const unsigned every = nps >= (size_t(1) << 22) ? 4u : 1u; // >= 4 M values: every fourth block of 256
            const size_t hblocks = (nps + 256ull * every - 1) / (256ull * every);
            ps_hist_kernel<<<static_cast<unsigned>(std::min<size_t>(hblocks, static_cast<size_t>(sm_count()) * 8)), 256, 0, s>>>(
                d_ps, nps, reinterpret_cast<unsigned*>(base), every, P.rowWidth, P.rowPitch);
        }
        hybrid_table_kernel<<<1, 256, 2 * nzi * sizeof(double), s>>>(P.ap, P.b, static_cast<int>(nzi), P.levels, static_cast<int>(nzo), reinterpret_cast<const unsigned*>(base),
                                               ps_scale, nbins, reinterpret_cast<HybridTableHeader*>(base + hdr_off),
                                               reinterpret_cast<unsigned short*>(base + tab_off));
        const dim3 blocks(static_cast<unsigned>((nInner + HT * HT_CHUNKS - 1) / (HT * HT_CHUNKS)), static_cast<unsigned>(nOuter));
        const HybridTableHeader* hdr = reinterpret_cast<const HybridTableHeader*>(base + hdr_off);
        const unsigned short* tab = reinterpret_cast<const unsigned short*>(base + tab_off);
        switch (method) {
#define FXCU_TABLE_CASE(M) \
    case M: vertical_hybrid_table_kernel<M><<<blocks, HT, tab_smem, s>>>(P, hdr, tab); break;
            FXCU_TABLE_CASE(FIMEXCU_VINT_METHOD_LIN)
            FXCU_TABLE_CASE(FIMEXCU_VINT_METHOD_LOG)
            FXCU_TABLE_CASE(FIMEXCU_VINT_METHOD_LOGLOG)
            FXCU_TABLE_CASE(FIMEXCU_VINT_METHOD_NN)
            FXCU_TABLE_CASE(FIMEXCU_VINT_METHOD_LIN_WEAK_EXTRA)
            FXCU_TABLE_CASE(FIMEXCU_VINT_METHOD_LIN_NO_EXTRA)
            FXCU_TABLE_CASE(FIMEXCU_VINT_METHOD_LIN_CONST_EXTRA)
#undef FXCU_TABLE_CASE
        default: FXCU_FAIL("fimexcu_vertical_hybrid_pressure: unknown vertical interpolation method");
        }
        FXCU_CUDA(cudaGetLastError());
        if (!cached)
            FXCU_CUDA(cudaStreamSynchronize(s)); // `h` is a temporary the upload may still be reading
        return FIMEXCU_OK;
    }
    if (!sweep_applies(P)) { // the per-level search kernel profits from a launch-wide monotonicity proof
        DevBuf mm;
        if (!mm.alloc(2 * sizeof(float), s))
            return FIMEXCU_ERROR;
        FXCU_CUDA(cudaMemcpyAsync(mm.p, init, sizeof(init), cudaMemcpyHostToDevice, s));
        if (nps)
            ps_minmax_kernel<<<static_cast<unsigned>(std::min<size_t>((nps + 255) / 256, 2048)), 256, 0, s>>>(d_ps, nps, mm.as<float>(), P.rowWidth, P.rowPitch);
        FXCU_CUDA(cudaGetLastError());
        int enc[2];
        FXCU_CUDA(cudaMemcpyAsync(enc, mm.p, sizeof(enc), cudaMemcpyDeviceToHost, s));
        FXCU_CUDA(cudaStreamSynchronize(s));
        auto dec = [](int e) {
            float f;
            const int bits = e >= 0 ? e : -e;
            std::memcpy(&f, &bits, sizeof(f));
            return static_cast<double>(e >= 0 ? f : -f);
        };
        const double a0 = dec(enc[0]) * ps_scale, a1 = dec(enc[1]) * ps_scale;
        P.dir_hint = hybrid_direction(ap_host, b_host, nzi, std::min(a0, a1), std::max(a0, a1));
    }
    const int rc = launch_vertical(P, s);
    FXCU_CUDA(cudaStreamSynchronize(s)); // `h` is a temporary
    return rc;
}

int fimexcu_vertical_hybrid_pressure_dev_f(int method, const float* d_in, size_t nOuter, size_t nzi, size_t nInner, const double* ap_host,
                                           const double* b_host, const float* d_ps, size_t nOuterPs, double ps_scale, const double* levels_host,
                                           size_t nzo, double valid_min, double valid_max, float badIn, float fillOut, float* d_out,
                                           void* stream)
{
    return fimexcu_vertical_hybrid_pressure_rows_dev_f(method, d_in, nOuter, nzi, nInner, ap_host, b_host, d_ps, nOuterPs, ps_scale, levels_host, nzo, valid_min,
                                                       valid_max, badIn, fillOut, d_out, 0, 0, stream);
}

int fimexcu_vertical_hybrid_pressure_f(int method, const float* in, size_t nOuter, size_t nzi, size_t nInner, const double* ap, const double* b,
                                       const float* ps, size_t nOuterPs, double ps_scale, const double* levels, size_t nzo, double valid_min,
                                       double valid_max, float badIn, float fillOut, float* out)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    const size_t ncol = nOuter * nInner, nin = ncol * nzi, nout = ncol * nzo, nps = nOuterPs * nInner;
    if (nout == 0)
        return FIMEXCU_OK;
    DevBuf din, dps, dout;
    if (!din.alloc(nin * 4, nullptr) || !dps.alloc(nps * 4, nullptr) || !dout.alloc(nout * 4, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(din.p, in, nin * 4, cudaMemcpyHostToDevice, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(dps.p, ps, nps * 4, cudaMemcpyHostToDevice, nullptr));
    if (fimexcu_vertical_hybrid_pressure_dev_f(method, din.as<float>(), nOuter, nzi, nInner, ap, b, dps.as<float>(), nOuterPs, ps_scale, levels, nzo,
                                               valid_min, valid_max, badIn, fillOut, dout.as<float>(), nullptr) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(out, dout.p, nout * 4, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

int fimexcu_get_values_nearest_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x)
{
    return two_level_host(FIMEXCU_VINT_METHOD_NN, infieldA, infieldB, outfield, n, a, b, x);
}

int fimexcu_get_values_linear_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x)
{
    return two_level_host(FIMEXCU_VINT_METHOD_LIN, infieldA, infieldB, outfield, n, a, b, x);
}

int fimexcu_get_values_linear_weak_extrapol_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x)
{
    return two_level_host(FIMEXCU_VINT_METHOD_LIN_WEAK_EXTRA, infieldA, infieldB, outfield, n, a, b, x);
}

int fimexcu_get_values_linear_no_extrapol_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x)
{
    return two_level_host(FIMEXCU_VINT_METHOD_LIN_NO_EXTRA, infieldA, infieldB, outfield, n, a, b, x);
}

int fimexcu_get_values_linear_const_extrapol_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x)
{
    return two_level_host(FIMEXCU_VINT_METHOD_LIN_CONST_EXTRA, infieldA, infieldB, outfield, n, a, b, x);
}

int fimexcu_get_values_log_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x)
{
    return two_level_host(FIMEXCU_VINT_METHOD_LOG, infieldA, infieldB, outfield, n, a, b, x);
}

int fimexcu_get_values_log_log_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x)
{
    return two_level_host(FIMEXCU_VINT_METHOD_LOGLOG, infieldA, infieldB, outfield, n, a, b, x);
}

int fimexcu_get_values_two_level_dev_f(int method, const float* d_A, const float* d_B, float* d_out, size_t n, double a, double b, double x,
                                       void* stream)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    int kind = 0;
    if (!two_level_plan(method, a, b, x, kind))
        return FIMEXCU_ERROR;
    return two_level_dev(kind, d_A, d_B, d_out, n, a, b, x, as_stream(stream));
}

} // extern "C"
