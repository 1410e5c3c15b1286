// fxcu_logtable.h -- a table-driven natural logarithm for device code whose result feeds an f32: x = 2^e * m, m in [1, 2);
// c = the centre of m's 1/128 interval, r = m/c - 1 (|r| < 2^-8), log x = e ln2 + log c + (r - r^2/2 + r^3/3 - r^4/4 + r^5/5).
// Absolute error < 2e-15 (truncation 6e-16, table and final roundings < 1e-15): a 16-byte table look-up and 10 FP64 operations
// instead of the ~45 of log().  tab[i] = (1/c_i, log c_i), 128 entries, meant for shared memory.
// Zero, negative, subnormal, infinite and NaN arguments go to log().
#pragma once

#include <cuda_runtime.h>

namespace fxcu {

constexpr double LOG_TABLE_ABS_ERR = 2e-15;

__device__ __forceinline__ void log_table_fill(double2* tab, int tid, int nthreads)
{
    for (int i = tid; i < 128; i += nthreads) {
        const double c = 1.0 + (i + 0.5) / 128.0;
        tab[i] = make_double2(1.0 / c, log(c));
    }
}

__device__ __forceinline__ double log_table(const double2* tab, double x)
{
    const long long bits = __double_as_longlong(x);
    const int ex = static_cast<int>(bits >> 52); // sign and biased exponent
    if (ex <= 0 || ex >= 0x7ff)
        return log(x);
    const double m = __longlong_as_double((bits & 0x000fffffffffffffll) | 0x3ff0000000000000ll);
    const double2 t = tab[static_cast<int>(bits >> 45) & 127];
    const double r = __fma_rn(m, t.x, -1.0);
    double p = __fma_rn(r, 0.2, -0.25);
    p = __fma_rn(p, r, 1.0 / 3);
    p = __fma_rn(p, r, -0.5);
    p = __fma_rn(p * r, r, r);
    return __fma_rn(static_cast<double>(ex - 1023), 0.693147180559945309417232, t.y + p);
}

// the same for an argument that was born as a positive, normal f32 (a coordinate column narrowed by Data::asFloat()): everything
// comes from the 32-bit pattern -- no 64-bit shifts; the caller guarantees 0 < a < inf, subnormals go to log()
// NORMAL: the caller has already established that a is a positive normal number
template <bool NORMAL = false> __device__ __forceinline__ double log_table_f32(const double2* tab, float a)
{
    const unsigned bits = __float_as_uint(a);
    if (!NORMAL && bits - 0x00800000u >= 0x7f000000u) // zero, subnormal, negative, inf, NaN
        return log(static_cast<double>(a));
    const unsigned mant = bits & 0x007fffffu;
    const double m = __hiloint2double(static_cast<int>(0x3ff00000u | (mant >> 3)), static_cast<int>(mant << 29)); // 1.mant, exact
    const double2 t = tab[mant >> 16];
    const double r = __fma_rn(m, t.x, -1.0);
    double p = __fma_rn(r, 0.2, -0.25);
    p = __fma_rn(p, r, 1.0 / 3);
    p = __fma_rn(p, r, -0.5);
    p = __fma_rn(p * r, r, r);
    return __fma_rn(__int2double_rn(static_cast<int>(bits >> 23) - 127), 0.693147180559945309417232, t.y + p);
}

} // namespace fxcu
