// fxcu_apply.cu -- horizontal interpolation cache: table build (K2/K3) and the per-slab apply kernels
// (K4 bilinear, K5 bicubic, K6 nearest) with bad->NaN fused on load and NaN->fill fused on store.
//
// Reference behaviour restated here (fimex 2.4.1):
//   CachedInterpolation ctor + createReducedDomain   src/CachedInterpolation.cc:149-171, 222-253
//   CachedInterpolation::interpolateValues           src/CachedInterpolation.cc:173-209
//   CachedNNInterpolation ctor / interpolateValues   src/CachedInterpolation.cc:262-337
//   mifi_get_values_{f,bilinear_f,bicubic_f}         src/interpolation.c:251-414
//   RoundAndClamp                                    src/MathUtils.cc:30-46
//
// Data layout in HBM
//   source slab  [nz][inY][inX] f32, dense (the reduced window the cache reports)
//   target slab  [nz][outY][outX] f32, dense
//   bilinear table, SoA over the outX*outY targets: off (int32: linear offset of the first tap inside a
//     level, -1 = undefined), wx, wy (f32 fractions; a negative value switches that direction off, which
//     encodes the reference's edge rules: x-only, y-only and nearest fallbacks) -> 12 B per target,
//     read once per slab and reused for every level.
//   bicubic table: off (int32, tap (x0-1, y0-1)), fx, fy (f64 fractions) -> 20 B per target.
//   nearest table: idx (int32, -1 = undefined) -> 4 B per target.
//
// Arithmetic: the reference is compiled for x86-64 without FMA contraction, so every product and sum is
// rounded separately; the kernels use the _rn intrinsics (never contracted by nvcc) and reproduce the
// reference's results bit for bit.
#include "fxcu_handles.h"

#include <algorithm>
#include <memory>
#include <vector>

namespace fxcu {

// ------------------------------------------------------------------------------------------------
// table entries
// ------------------------------------------------------------------------------------------------
// Index tables hold finite values by construction (points2position writes -999 for non-finite input, the
// kd-tree search -1000).  Casting NaN or values beyond the int range is undefined behaviour in the
// reference; such entries are treated as "outside".
__device__ __forceinline__ bool entry_usable(double x, double y)
{
    return isfinite(x) && isfinite(y) && fabs(x) < 2147483000. && fabs(y) < 2147483000.;
}

// lround(): half away from zero
__device__ __forceinline__ long long lround_dev(double v) { return llround(v); }

// interpolation.c:267-343 -> (off, wx, wy)
__device__ __forceinline__ void bilinear_entry(double x, double y, int ix, int iy, int& off, float& wx, float& wy)
{
    off = -1;
    wx = wy = -1.f;
    if (!entry_usable(x, y))
        return;
    const double fx = floor(x), fy = floor(y);
    int x0 = static_cast<int>(fx), y0 = static_cast<int>(fy);
    const float xfrac = static_cast<float>(x - fx);
    const float yfrac = static_cast<float>(y - fy);
    const bool x_in = (0 <= x0) && (x0 + 1 < ix);
    const bool y_in = (0 <= y0) && (y0 + 1 < iy);
    if (x_in) {
        wx = xfrac;
        if (y_in) {
            wy = yfrac;
        } else {
            const long long yr = lround_dev(y);
            if (!(0 <= yr && yr < iy))
                return;
            y0 = static_cast<int>(yr);
        }
    } else {
        const long long xr = lround_dev(x);
        if (!(0 <= xr && xr < ix))
            return;
        x0 = static_cast<int>(xr);
        if (y_in) {
            wy = yfrac;
        } else {
            const long long yr = lround_dev(y);
            // the reference accepts yr == iy here (interpolation.c:322) and then reads outside the level;
            // that corner strip is defined as undefined (NaN) instead
            if (!(0 <= yr && yr < iy))
                return;
            y0 = static_cast<int>(yr);
        }
    }
    off = y0 * ix + x0;
}

// interpolation.c:345-414 -> (off of tap (x0-1,y0-1), fractions)
__device__ __forceinline__ void bicubic_entry(double x, double y, int ix, int iy, int& off, double& fx, double& fy)
{
    off = -1;
    fx = fy = 0.;
    if (!entry_usable(x, y))
        return;
    const double flx = floor(x), fly = floor(y);
    const int x0 = static_cast<int>(flx), y0 = static_cast<int>(fly);
    if (!((1 <= x0) && (x0 + 2 < ix) && (1 <= y0) && (y0 + 2 < iy)))
        return;
    fx = x - flx;
    fy = y - fly;
    off = (y0 - 1) * ix + (x0 - 1);
}

// MathUtils.cc:30-46 with mini = 0: round half away from zero, valid iff 0 <= r <= maxi (-0.0 counts as 0)
__device__ __forceinline__ long long round_and_clamp(double d, long long maxi)
{
    const double r = round(d);
    if (!(r >= 0.0) || r > static_cast<double>(maxi))
        return -1;
    return static_cast<long long>(r);
}

// cubic convolution weights, a = -0.5 (interpolation.c:348-386), products and sums rounded one by one
__device__ __forceinline__ void cubic_weights(double f, double w[4])
{
    const double M[4][4] = {{0.0, 1.0, 0.0, 0.0}, {-0.5, 0.0, 0.5, 0.0}, {1.0, -2.5, 2.0, -0.5}, {-0.5, 1.5, -1.5, 0.5}};
    double p[4];
    p[0] = 1.;
    p[1] = f;
    p[2] = __dmul_rn(f, f);
    p[3] = __dmul_rn(p[2], f);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        double acc = 0.;
#pragma unroll
        for (int j = 0; j < 4; ++j)
            acc = __dadd_rn(acc, __dmul_rn(p[j], M[j][i]));
        w[i] = acc;
    }
}

// ------------------------------------------------------------------------------------------------
// build kernels
// ------------------------------------------------------------------------------------------------
// K3: min/max of both tables (std::min_element / max_element at CachedInterpolation.cc:229-230).
// partial[4*b + {0,1,2,3}] = xmin, xmax, ymin, ymax of block b
__global__ void __launch_bounds__(256) minmax_kernel(const double* __restrict__ px, const double* __restrict__ py, size_t n, double* __restrict__ partial)
{
    double xlo = INFINITY, xhi = -INFINITY, ylo = INFINITY, yhi = -INFINITY;
    for (size_t i = blockIdx.x * 256ull + threadIdx.x; i < n; i += 256ull * gridDim.x) {
        const double x = px[i], y = py[i];
        xlo = fmin(xlo, x);
        xhi = fmax(xhi, x);
        ylo = fmin(ylo, y);
        yhi = fmax(yhi, y);
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        xlo = fmin(xlo, __shfl_xor_sync(0xffffffffu, xlo, s));
        xhi = fmax(xhi, __shfl_xor_sync(0xffffffffu, xhi, s));
        ylo = fmin(ylo, __shfl_xor_sync(0xffffffffu, ylo, s));
        yhi = fmax(yhi, __shfl_xor_sync(0xffffffffu, yhi, s));
    }
    __shared__ double sm[8][4];
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) {
        sm[w][0] = xlo; sm[w][1] = xhi; sm[w][2] = ylo; sm[w][3] = yhi;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; ++k) {
            xlo = fmin(xlo, sm[k][0]); xhi = fmax(xhi, sm[k][1]);
            ylo = fmin(ylo, sm[k][2]); yhi = fmax(yhi, sm[k][3]);
        }
        partial[4 * blockIdx.x + 0] = xlo;
        partial[4 * blockIdx.x + 1] = xhi;
        partial[4 * blockIdx.x + 2] = ylo;
        partial[4 * blockIdx.x + 3] = yhi;
    }
}

__global__ void __launch_bounds__(256) build_bilinear_kernel(const double* __restrict__ px, const double* __restrict__ py, size_t n, int ix, int iy,
                                                             double minX, double minY, int* __restrict__ off, float* __restrict__ wx,
                                                             float* __restrict__ wy)
{
    const size_t i = blockIdx.x * 256ull + threadIdx.x;
    if (i >= n)
        return;
    // rebasing exactly as CachedInterpolation.cc:243-246 (pointsOnXAxis[xy] -= minX)
    const double x = px[i] - minX, y = py[i] - minY;
    int o;
    float fx, fy;
    bilinear_entry(x, y, ix, iy, o, fx, fy);
    off[i] = o;
    wx[i] = fx;
    wy[i] = fy;
}

__global__ void __launch_bounds__(256) build_bicubic_kernel(const double* __restrict__ px, const double* __restrict__ py, size_t n, int ix, int iy,
                                                            double minX, double minY, int* __restrict__ off, double* __restrict__ fxo,
                                                            double* __restrict__ fyo)
{
    const size_t i = blockIdx.x * 256ull + threadIdx.x;
    if (i >= n)
        return;
    const double x = px[i] - minX, y = py[i] - minY;
    int o;
    double fx, fy;
    bicubic_entry(x, y, ix, iy, o, fx, fy);
    off[i] = o;
    fxo[i] = fx;
    fyo[i] = fy;
}

// CachedNNInterpolation ctor, first loop (CachedInterpolation.cc:270-286): idx + used range
// range[0..3] = minInX, maxInX, minInY, maxInY (initialised by the host to inX, 0, inY, 0)
__global__ void __launch_bounds__(256) build_nn_kernel(const double* __restrict__ px, const double* __restrict__ py, size_t n, int ix, int iy,
                                                       int* __restrict__ idx, int* __restrict__ range)
{
    const size_t i = blockIdx.x * 256ull + threadIdx.x;
    int lox = ix, hix = 0, loy = iy, hiy = 0;
    if (i < n) {
        const long long rx = round_and_clamp(px[i], ix - 1), ry = round_and_clamp(py[i], iy - 1);
        int v = -1;
        if (rx >= 0 && ry >= 0) {
            v = static_cast<int>(ry) * ix + static_cast<int>(rx);
            lox = hix = static_cast<int>(rx);
            loy = hiy = static_cast<int>(ry);
        }
        idx[i] = v;
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        lox = min(lox, __shfl_xor_sync(0xffffffffu, lox, s));
        hix = max(hix, __shfl_xor_sync(0xffffffffu, hix, s));
        loy = min(loy, __shfl_xor_sync(0xffffffffu, loy, s));
        hiy = max(hiy, __shfl_xor_sync(0xffffffffu, hiy, s));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(&range[0], lox);
        atomicMax(&range[1], hix);
        atomicMin(&range[2], loy);
        atomicMax(&range[3], hiy);
    }
}

// CachedInterpolation.cc:299-306: re-express the indices in the reduced window
__global__ void __launch_bounds__(256) rebase_nn_kernel(int* __restrict__ idx, size_t n, int ix, int minX, int minY, int redX)
{
    const size_t i = blockIdx.x * 256ull + threadIdx.x;
    if (i >= n)
        return;
    const int v = idx[i];
    if (v >= 0) {
        const int y = v / ix - minY, x = v % ix - minX;
        idx[i] = y * redX + x;
    }
}

// ------------------------------------------------------------------------------------------------
// apply kernels
// ------------------------------------------------------------------------------------------------
template <bool BAD>
__device__ __forceinline__ float load_in(const float* __restrict__ p, float bad)
{
    float v = __ldg(p);
    if (BAD && v == bad) // mifi_bad2nanf, interpolation.c:1064-1072
        v = nan_f();
    return v;
}

template <bool FILL>
__device__ __forceinline__ void store_out(float* __restrict__ p, float v, float fill)
{
    if (FILL && isnan(v)) // mifi_nanf2bad, interpolation.c:1074-1082
        v = fill;
    __stcs(p, v); // written once, never re-read by this kernel: keep L2 for the source planes
}

// K4.  One thread per target xy, blockIdx.y selects a chunk of levels; table entry in registers for all
// its levels.  Consecutive threads -> consecutive targets -> coalesced 128 B stores per warp and level.
template <bool BAD, bool FILL>
__global__ void __launch_bounds__(256) bilinear_apply_kernel(const int* __restrict__ off, const float* __restrict__ wx, const float* __restrict__ wy,
                                                             const float* __restrict__ in, float* __restrict__ out, size_t nxy, int nz, int zchunk,
                                                             size_t inLayer, int inX, float bad, float fill)
{
    const size_t o = blockIdx.x * 256ull + threadIdx.x;
    if (o >= nxy)
        return;
    const int z0 = blockIdx.y * zchunk;
    const int z1 = min(nz, z0 + zchunk);
    const int f = off[o];
    const float fx = wx[o], fy = wy[o];
    float* po = out + static_cast<size_t>(z0) * nxy + o;
    if (f < 0) {
        for (int z = z0; z < z1; ++z, po += nxy)
            store_out<FILL>(po, nan_f(), fill);
        return;
    }
    const float* p = in + static_cast<size_t>(z0) * inLayer + f;
    if (fx >= 0.f && fy >= 0.f) {
        const float omx = __fsub_rn(1.f, fx), omy = __fsub_rn(1.f, fy);
#pragma unroll 4
        for (int z = z0; z < z1; ++z, p += inLayer, po += nxy) {
            const float s00 = load_in<BAD>(p, bad), s01 = load_in<BAD>(p + 1, bad);
            const float s10 = load_in<BAD>(p + inX, bad), s11 = load_in<BAD>(p + inX + 1, bad);
            const float top = __fadd_rn(__fmul_rn(omx, s00), __fmul_rn(fx, s01));
            const float bot = __fadd_rn(__fmul_rn(omx, s10), __fmul_rn(fx, s11));
            store_out<FILL>(po, __fadd_rn(__fmul_rn(omy, top), __fmul_rn(fy, bot)), fill);
        }
    } else if (fx >= 0.f) { // linear in x on the nearest row (interpolation.c:289-299)
        const float omx = __fsub_rn(1.f, fx);
#pragma unroll 4
        for (int z = z0; z < z1; ++z, p += inLayer, po += nxy) {
            const float s00 = load_in<BAD>(p, bad), s01 = load_in<BAD>(p + 1, bad);
            store_out<FILL>(po, __fadd_rn(__fmul_rn(omx, s00), __fmul_rn(fx, s01)), fill);
        }
    } else if (fy >= 0.f) { // linear in y on the nearest column (:311-319)
        const float omy = __fsub_rn(1.f, fy);
#pragma unroll 4
        for (int z = z0; z < z1; ++z, p += inLayer, po += nxy) {
            const float s00 = load_in<BAD>(p, bad), s10 = load_in<BAD>(p + inX, bad);
            store_out<FILL>(po, __fadd_rn(__fmul_rn(omy, s00), __fmul_rn(fy, s10)), fill);
        }
    } else { // nearest cell (:321-328)
#pragma unroll 4
        for (int z = z0; z < z1; ++z, p += inLayer, po += nxy)
            store_out<FILL>(po, load_in<BAD>(p, bad), fill);
    }
}

// K5.  Same mapping; 16 taps, weights in f64, the four column sums added into the f32 result one at a
// time (interpolation.c:388-406).
template <bool BAD, bool FILL>
__global__ void __launch_bounds__(256) bicubic_apply_kernel(const int* __restrict__ off, const double* __restrict__ fxs, const double* __restrict__ fys,
                                                            const float* __restrict__ in, float* __restrict__ out, size_t nxy, int nz, int zchunk,
                                                            size_t inLayer, int inX, float bad, float fill)
{
    const size_t o = blockIdx.x * 256ull + threadIdx.x;
    if (o >= nxy)
        return;
    const int z0 = blockIdx.y * zchunk;
    const int z1 = min(nz, z0 + zchunk);
    const int f = off[o];
    float* po = out + static_cast<size_t>(z0) * nxy + o;
    if (f < 0) {
        for (int z = z0; z < z1; ++z, po += nxy)
            store_out<FILL>(po, nan_f(), fill);
        return;
    }
    double wxv[4], wyv[4];
    cubic_weights(fxs[o], wxv);
    cubic_weights(fys[o], wyv);
    const float* p = in + static_cast<size_t>(z0) * inLayer + f;
    for (int z = z0; z < z1; ++z, p += inLayer, po += nxy) {
        float v[4][4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int i = 0; i < 4; ++i)
                v[j][i] = load_in<BAD>(p + j * inX + i, bad);
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            double col = 0.;
#pragma unroll
            for (int i = 0; i < 4; ++i)
                col = __dadd_rn(col, __dmul_rn(wxv[i], static_cast<double>(v[j][i])));
            acc = __double2float_rn(__dadd_rn(static_cast<double>(acc), __dmul_rn(col, wyv[j])));
        }
        store_out<FILL>(po, acc, fill);
    }
}

// K6.  Pure gather.
template <bool BAD, bool FILL>
__global__ void __launch_bounds__(256) nearest_apply_kernel(const int* __restrict__ idx, const float* __restrict__ in, float* __restrict__ out,
                                                            size_t nxy, int nz, int zchunk, size_t inLayer, float bad, float fill)
{
    const size_t o = blockIdx.x * 256ull + threadIdx.x;
    if (o >= nxy)
        return;
    const int z0 = blockIdx.y * zchunk;
    const int z1 = min(nz, z0 + zchunk);
    const int f = idx[o];
    float* po = out + static_cast<size_t>(z0) * nxy + o;
    if (f < 0) {
        for (int z = z0; z < z1; ++z, po += nxy)
            store_out<FILL>(po, nan_f(), fill);
        return;
    }
    const float* p = in + static_cast<size_t>(z0) * inLayer + f;
#pragma unroll 4
    for (int z = z0; z < z1; ++z, p += inLayer, po += nxy)
        store_out<FILL>(po, load_in<BAD>(p, bad), fill);
}

// Fused vector pass: bilinear u and v with one table read, then the rotation of reproject.cc:488-507, then the
// stores.  Replaces four interpolateValues() calls and two reprojectValues() calls of the reference
// (CDMInterpolator.cc:232-259 runs once for x_wind and once for y_wind).
__device__ __forceinline__ float bilinear_value(const float s00, const float s01, const float s10, const float s11, float fx, float fy)
{
    if (fx >= 0.f && fy >= 0.f) {
        const float omx = __fsub_rn(1.f, fx), omy = __fsub_rn(1.f, fy);
        const float top = __fadd_rn(__fmul_rn(omx, s00), __fmul_rn(fx, s01));
        const float bot = __fadd_rn(__fmul_rn(omx, s10), __fmul_rn(fx, s11));
        return __fadd_rn(__fmul_rn(omy, top), __fmul_rn(fy, bot));
    }
    if (fx >= 0.f)
        return __fadd_rn(__fmul_rn(__fsub_rn(1.f, fx), s00), __fmul_rn(fx, s01));
    if (fy >= 0.f)
        return __fadd_rn(__fmul_rn(__fsub_rn(1.f, fy), s00), __fmul_rn(fy, s10));
    return s00;
}

template <bool BAD, bool FILL, bool ROT>
__global__ void __launch_bounds__(256) bilinear_uv_kernel(const int* __restrict__ off, const float* __restrict__ wx, const float* __restrict__ wy,
                                                          const double* __restrict__ cs, const double* __restrict__ sn, const float* __restrict__ uin,
                                                          const float* __restrict__ vin, float* __restrict__ uout, float* __restrict__ vout, size_t nxy,
                                                          int nz, int zchunk, size_t inLayer, int inX, float badU, float badV, float fillU, float fillV)
{
    const size_t o = blockIdx.x * 256ull + threadIdx.x;
    if (o >= nxy)
        return;
    const int z0 = blockIdx.y * zchunk;
    const int z1 = min(nz, z0 + zchunk);
    const int f = off[o];
    const float fx = wx[o], fy = wy[o];
    float* pu = uout + static_cast<size_t>(z0) * nxy + o;
    float* pv = vout + static_cast<size_t>(z0) * nxy + o;
    if (f < 0) {
        for (int z = z0; z < z1; ++z, pu += nxy, pv += nxy) {
            store_out<FILL>(pu, nan_f(), fillU);
            store_out<FILL>(pv, nan_f(), fillV);
        }
        return;
    }
    double c = 1., s = 0.;
    if (ROT) {
        c = cs[o];
        s = sn[o];
    }
    // taps that a switched-off direction does not use alias the first tap (never read out of range)
    const int dx = fx >= 0.f ? 1 : 0;
    const int dy = fy >= 0.f ? inX : 0;
    const float* qu = uin + static_cast<size_t>(z0) * inLayer + f;
    const float* qv = vin + static_cast<size_t>(z0) * inLayer + f;
#pragma unroll 2
    for (int z = z0; z < z1; ++z, qu += inLayer, qv += inLayer, pu += nxy, pv += nxy) {
        const float u00 = load_in<BAD>(qu, badU), u01 = load_in<BAD>(qu + dx, badU);
        const float u10 = load_in<BAD>(qu + dy, badU), u11 = load_in<BAD>(qu + dy + dx, badU);
        const float v00 = load_in<BAD>(qv, badV), v01 = load_in<BAD>(qv + dx, badV);
        const float v10 = load_in<BAD>(qv + dy, badV), v11 = load_in<BAD>(qv + dy + dx, badV);
        float u = bilinear_value(u00, u01, u10, u11, fx, fy);
        float v = bilinear_value(v00, v01, v10, v11, fx, fy);
        if (ROT) {
            const double du = static_cast<double>(u), dv = static_cast<double>(v);
            u = __double2float_rn(__dsub_rn(__dmul_rn(du, c), __dmul_rn(dv, s)));
            v = __double2float_rn(__dadd_rn(__dmul_rn(du, s), __dmul_rn(dv, c)));
        }
        store_out<FILL>(pu, u, fillU);
        store_out<FILL>(pv, v, fillV);
    }
}

// single-position variants with the reference signatures (one launch; levels across threads)
__global__ void point_kernel(int method, const float* __restrict__ in, float* __restrict__ out, double x, double y, int ix, int iy, int iz)
{
    const size_t layer = static_cast<size_t>(ix) * iy;
    if (method == FIMEXCU_INTERPOL_BILINEAR) {
        int off;
        float fx, fy;
        bilinear_entry(x, y, ix, iy, off, fx, fy);
        for (int z = threadIdx.x; z < iz; z += blockDim.x) {
            float r = nan_f();
            if (off >= 0) {
                const float* p = in + z * layer + off;
                if (fx >= 0.f && fy >= 0.f) {
                    const float omx = __fsub_rn(1.f, fx), omy = __fsub_rn(1.f, fy);
                    const float top = __fadd_rn(__fmul_rn(omx, p[0]), __fmul_rn(fx, p[1]));
                    const float bot = __fadd_rn(__fmul_rn(omx, p[ix]), __fmul_rn(fx, p[ix + 1]));
                    r = __fadd_rn(__fmul_rn(omy, top), __fmul_rn(fy, bot));
                } else if (fx >= 0.f) {
                    r = __fadd_rn(__fmul_rn(__fsub_rn(1.f, fx), p[0]), __fmul_rn(fx, p[1]));
                } else if (fy >= 0.f) {
                    r = __fadd_rn(__fmul_rn(__fsub_rn(1.f, fy), p[0]), __fmul_rn(fy, p[ix]));
                } else {
                    r = p[0];
                }
            }
            out[z] = r;
        }
    } else if (method == FIMEXCU_INTERPOL_BICUBIC) {
        int off;
        double fx, fy;
        bicubic_entry(x, y, ix, iy, off, fx, fy);
        double wxv[4], wyv[4];
        cubic_weights(fx, wxv);
        cubic_weights(fy, wyv);
        for (int z = threadIdx.x; z < iz; z += blockDim.x) {
            float acc = nan_f();
            if (off >= 0) {
                const float* p = in + z * layer + off;
                acc = 0.f;
                for (int j = 0; j < 4; ++j) {
                    double col = 0.;
                    for (int i = 0; i < 4; ++i)
                        col = __dadd_rn(col, __dmul_rn(wxv[i], static_cast<double>(p[j * ix + i])));
                    acc = __double2float_rn(__dadd_rn(static_cast<double>(acc), __dmul_rn(col, wyv[j])));
                }
            }
            out[z] = acc;
        }
    } else { // mifi_get_values_f: lround, in range -> copy (interpolation.c:251-265)
        long long rx = -1, ry = -1;
        if (entry_usable(x, y)) {
            rx = lround_dev(x);
            ry = lround_dev(y);
        }
        const bool ok = rx >= 0 && rx < ix && ry >= 0 && ry < iy;
        for (int z = threadIdx.x; z < iz; z += blockDim.x)
            out[z] = ok ? in[z * layer + ry * ix + rx] : nan_f();
    }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static long long clamp_ex(long long low, double v, long long extend, long long high)
{
    // CachedInterpolation.cc:215-219
    long long value = static_cast<long long>(extend > 0 ? std::ceil(v) : std::floor(v)) + extend;
    return std::max(low, std::min(value, high));
}

// picks the number of level chunks (gridDim.y) so that small targets still fill the machine
static int pick_zchunk(size_t nxy, int nz)
{
    const size_t blocks_xy = (nxy + 255) / 256;
    const size_t want = static_cast<size_t>(sm_count()) * 8 * 2; // two waves of 8 resident CTAs per SM
    if (blocks_xy >= want || nz <= 1)
        return nz;
    size_t chunks = (want + blocks_xy - 1) / blocks_xy;
    chunks = std::min<size_t>(chunks, static_cast<size_t>(nz));
    return static_cast<int>((nz + chunks - 1) / chunks);
}

static fimexcu_cache* cache_build(int method, const double* d_px, const double* d_py, size_t inX, size_t inY, size_t outX, size_t outY, cudaStream_t s)
{
    if (!require_device())
        return nullptr;
    const size_t n = outX * outY;
    if (inX == 0 || inY == 0 || inX * inY >= 2147483647ull || n >= (1ull << 40)) {
        set_error("fimexcu_cache_create: grid sizes out of range (need 0 < inX*inY < 2^31)");
        return nullptr;
    }
    std::unique_ptr<fimexcu_cache> c(new fimexcu_cache());
    c->method = method;
    c->inX = inX;
    c->inY = inY;
    c->outX = outX;
    c->outY = outY;
    FXCU_CUDA_NULL(cudaGetDevice(&c->device));
    switch (method) {
    case FIMEXCU_INTERPOL_BILINEAR: c->kind = KIND_BILINEAR; break;
    case FIMEXCU_INTERPOL_BICUBIC: c->kind = KIND_BICUBIC; break;
    case FIMEXCU_INTERPOL_NEAREST_NEIGHBOR:
    case FIMEXCU_INTERPOL_COORD_NN:
    case FIMEXCU_INTERPOL_COORD_NN_KD: c->kind = KIND_NN; break;
    default:
        // createCachedInterpolation: "unknown interpolation function" (CachedInterpolation.cc:145)
        set_error("unknown interpolation function: " + std::to_string(method));
        return nullptr;
    }
    if (!c->off.alloc_sync(std::max<size_t>(n, 1) * sizeof(int)))
        return nullptr;
    if (n == 0)
        return c.release();
    const unsigned nb = blocks_for(n, 256);

    if (c->kind == KIND_NN) {
        DevBuf range;
        if (!range.alloc(4 * sizeof(int), s))
            return nullptr;
        const int init[4] = {static_cast<int>(inX), 0, static_cast<int>(inY), 0};
        FXCU_CUDA_NULL(cudaMemcpyAsync(range.p, init, sizeof(init), cudaMemcpyHostToDevice, s));
        build_nn_kernel<<<nb, 256, 0, s>>>(d_px, d_py, n, static_cast<int>(inX), static_cast<int>(inY), c->off.as<int>(), range.as<int>());
        FXCU_CUDA_NULL(cudaGetLastError());
        int r[4];
        FXCU_CUDA_NULL(cudaMemcpyAsync(r, range.p, sizeof(r), cudaMemcpyDeviceToHost, s));
        FXCU_CUDA_NULL(cudaStreamSynchronize(s));
        // CachedInterpolation.cc:288-314
        size_t minInX = static_cast<size_t>(r[0]), maxInX = static_cast<size_t>(r[1]);
        size_t minInY = static_cast<size_t>(r[2]), maxInY = static_cast<size_t>(r[3]);
        const size_t EXT = 2;
        minInX = (minInX > EXT) ? minInX - EXT : 0;
        minInY = (minInY > EXT) ? minInY - EXT : 0;
        maxInX = std::min(inX - 1, maxInX + EXT);
        maxInY = std::min(inY - 1, maxInY + EXT);
        if ((minInX > 0 || minInY > 0 || maxInX < inX - 1 || maxInY < inY - 1) && (minInX + 2 * EXT <= maxInX) && (minInY + 2 * EXT <= maxInY)) {
            const size_t redX = maxInX - minInX + 1, redY = maxInY - minInY + 1;
            rebase_nn_kernel<<<nb, 256, 0, s>>>(c->off.as<int>(), n, static_cast<int>(inX), static_cast<int>(minInX), static_cast<int>(minInY),
                                                 static_cast<int>(redX));
            FXCU_CUDA_NULL(cudaGetLastError());
            c->reduced = true;
            c->xMin = minInX;
            c->yMin = minInY;
            c->inX = redX;
            c->inY = redY;
        }
        FXCU_CUDA_NULL(cudaStreamSynchronize(s));
        return c.release();
    }

    // bilinear / bicubic: reduced domain from the table extrema (CachedInterpolation.cc:222-253)
    const unsigned rb = std::min<unsigned>(nb, 1024u);
    DevBuf partial;
    if (!partial.alloc(static_cast<size_t>(rb) * 4 * sizeof(double), s))
        return nullptr;
    minmax_kernel<<<rb, 256, 0, s>>>(d_px, d_py, n, partial.as<double>());
    FXCU_CUDA_NULL(cudaGetLastError());
    std::vector<double> hp(static_cast<size_t>(rb) * 4);
    FXCU_CUDA_NULL(cudaMemcpyAsync(hp.data(), partial.p, hp.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
    FXCU_CUDA_NULL(cudaStreamSynchronize(s));
    double xlo = INFINITY, xhi = -INFINITY, ylo = INFINITY, yhi = -INFINITY;
    for (unsigned b = 0; b < rb; ++b) {
        xlo = std::fmin(xlo, hp[4 * b]);
        xhi = std::fmax(xhi, hp[4 * b + 1]);
        ylo = std::fmin(ylo, hp[4 * b + 2]);
        yhi = std::fmax(yhi, hp[4 * b + 3]);
    }
    long long minX = 0, minY = 0;
    if (std::isfinite(xlo) && std::isfinite(xhi) && std::isfinite(ylo) && std::isfinite(yhi)) {
        const long long EXT = 2;
        const long long loX = clamp_ex(0, xlo, -EXT, static_cast<long long>(inX) - 1);
        const long long loY = clamp_ex(0, ylo, -EXT, static_cast<long long>(inY) - 1);
        const long long hiX = clamp_ex(0, xhi, +EXT, static_cast<long long>(inX) - 1);
        const long long hiY = clamp_ex(0, yhi, +EXT, static_cast<long long>(inY) - 1);
        if (!((hiX - loX) < 1 || (hiY - loY) < 1)) {
            minX = loX;
            minY = loY;
            c->reduced = true;
            c->xMin = static_cast<size_t>(loX);
            c->yMin = static_cast<size_t>(loY);
            c->inX = static_cast<size_t>(hiX - loX + 1);
            c->inY = static_cast<size_t>(hiY - loY + 1);
        }
    }
    if (c->kind == KIND_BILINEAR) {
        if (!c->wx.alloc_sync(n * sizeof(float)) || !c->wy.alloc_sync(n * sizeof(float)))
            return nullptr;
        build_bilinear_kernel<<<nb, 256, 0, s>>>(d_px, d_py, n, static_cast<int>(c->inX), static_cast<int>(c->inY), static_cast<double>(minX),
                                                  static_cast<double>(minY), c->off.as<int>(), c->wx.as<float>(), c->wy.as<float>());
    } else {
        if (!c->wx.alloc_sync(n * sizeof(double)) || !c->wy.alloc_sync(n * sizeof(double)))
            return nullptr;
        build_bicubic_kernel<<<nb, 256, 0, s>>>(d_px, d_py, n, static_cast<int>(c->inX), static_cast<int>(c->inY), static_cast<double>(minX),
                                                 static_cast<double>(minY), c->off.as<int>(), c->wx.as<double>(), c->wy.as<double>());
    }
    FXCU_CUDA_NULL(cudaGetLastError());
    FXCU_CUDA_NULL(cudaStreamSynchronize(s));
    return c.release();
}

template <bool BAD, bool FILL>
static void launch_cache_apply(const fimexcu_cache* c, const float* d_in, int nz, float* d_out, float bad, float fill, cudaStream_t s)
{
    const size_t nxy = c->outX * c->outY;
    const size_t inLayer = c->inX * c->inY;
    const int zchunk = pick_zchunk(nxy, nz);
    const dim3 grid(blocks_for(nxy, 256), static_cast<unsigned>((nz + zchunk - 1) / zchunk));
    switch (c->kind) {
    case KIND_BILINEAR:
        bilinear_apply_kernel<BAD, FILL><<<grid, 256, 0, s>>>(c->off.as<int>(), c->wx.as<float>(), c->wy.as<float>(), d_in, d_out, nxy, nz, zchunk,
                                                               inLayer, static_cast<int>(c->inX), bad, fill);
        break;
    case KIND_BICUBIC:
        bicubic_apply_kernel<BAD, FILL><<<grid, 256, 0, s>>>(c->off.as<int>(), c->wx.as<double>(), c->wy.as<double>(), d_in, d_out, nxy, nz, zchunk,
                                                              inLayer, static_cast<int>(c->inX), bad, fill);
        break;
    default:
        nearest_apply_kernel<BAD, FILL><<<grid, 256, 0, s>>>(c->off.as<int>(), d_in, d_out, nxy, nz, zchunk, inLayer, bad, fill);
        break;
    }
}

int cache_apply_dev(const fimexcu_cache* c, const float* d_in, size_t nz, float* d_out, float bad, float fill, cudaStream_t s)
{
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply: null cache");
    if (nz == 0 || c->outX * c->outY == 0)
        return FIMEXCU_OK;
    if (nz > 0x7fffffffull / 2)
        FXCU_FAIL("fimexcu_cache_apply: too many levels");
    const bool has_bad = !std::isnan(bad), has_fill = !std::isnan(fill);
    const int inz = static_cast<int>(nz);
    if (has_bad && has_fill)
        launch_cache_apply<true, true>(c, d_in, inz, d_out, bad, fill, s);
    else if (has_bad)
        launch_cache_apply<true, false>(c, d_in, inz, d_out, bad, fill, s);
    else if (has_fill)
        launch_cache_apply<false, true>(c, d_in, inz, d_out, bad, fill, s);
    else
        launch_cache_apply<false, false>(c, d_in, inz, d_out, bad, fill, s);
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}

} // namespace fxcu

using namespace fxcu;

extern "C" {

fimexcu_cache* fimexcu_cache_create_dev(int method, const double* d_px, const double* d_py, size_t inX, size_t inY, size_t outX, size_t outY,
                                        void* stream)
{
    clear_error();
    return cache_build(method, d_px, d_py, inX, inY, outX, outY, as_stream(stream));
}

fimexcu_cache* fimexcu_cache_create(int method, const double* px, const double* py, size_t inX, size_t inY, size_t outX, size_t outY)
{
    clear_error();
    if (!require_device())
        return nullptr;
    const size_t n = outX * outY;
    DevBuf dx, dy;
    if (!dx.alloc_sync(n * sizeof(double)) || !dy.alloc_sync(n * sizeof(double)))
        return nullptr;
    FXCU_CUDA_NULL(cudaMemcpy(dx.p, px, n * sizeof(double), cudaMemcpyHostToDevice));
    FXCU_CUDA_NULL(cudaMemcpy(dy.p, py, n * sizeof(double), cudaMemcpyHostToDevice));
    return cache_build(method, dx.as<double>(), dy.as<double>(), inX, inY, outX, outY, nullptr);
}

int fimexcu_cache_info(const fimexcu_cache* c, size_t* inX, size_t* inY, size_t* outX, size_t* outY, int* reduced, size_t* xMin, size_t* yMin)
{
    if (!c)
        FXCU_FAIL("fimexcu_cache_info: null cache");
    if (inX) *inX = c->inX;
    if (inY) *inY = c->inY;
    if (outX) *outX = c->outX;
    if (outY) *outY = c->outY;
    if (reduced) *reduced = c->reduced ? 1 : 0;
    if (xMin) *xMin = c->xMin;
    if (yMin) *yMin = c->yMin;
    return FIMEXCU_OK;
}

int fimexcu_cache_apply_dev_f(const fimexcu_cache* c, const float* d_in, size_t nz, float* d_out, float badIn, float fillOut, void* stream)
{
    clear_error();
    return cache_apply_dev(c, d_in, nz, d_out, badIn, fillOut, as_stream(stream));
}

int fimexcu_cache_apply_f(const fimexcu_cache* c, const float* in, size_t size, float* out, size_t* newSize, float badIn, float fillOut)
{
    clear_error();
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply_f: null cache");
    if (!require_device())
        return FIMEXCU_ERROR;
    const size_t inLayer = c->inX * c->inY, outLayer = c->outX * c->outY;
    const size_t nz = size / inLayer; // CachedInterpolation.cc:176
    if (newSize)
        *newSize = nz * outLayer;
    if (nz == 0 || outLayer == 0)
        return FIMEXCU_OK;
    // Levels are independent: split the slab into chunks and run copy-in, kernel and copy-out of
    // neighbouring chunks on different streams so that PCIe traffic in both directions overlaps.
    const int NS = 3;
    const size_t chunk = std::max<size_t>(1, std::min<size_t>(nz, (nz + 7) / 8));
    cudaStream_t st[NS];
    for (int i = 0; i < NS; ++i)
        FXCU_CUDA(cudaStreamCreateWithFlags(&st[i], cudaStreamNonBlocking));
    int rc = FIMEXCU_OK;
    {
        DevBuf din[NS], dout[NS];
        for (int i = 0; i < NS && rc == FIMEXCU_OK; ++i)
            if (!din[i].alloc(chunk * inLayer * sizeof(float), st[i]) || !dout[i].alloc(chunk * outLayer * sizeof(float), st[i]))
                rc = FIMEXCU_ERROR;
        size_t k = 0;
        for (size_t z0 = 0; z0 < nz && rc == FIMEXCU_OK; z0 += chunk, ++k) {
            const size_t zn = std::min(chunk, nz - z0);
            const int i = static_cast<int>(k % NS);
            if (!cuda_ok(cudaMemcpyAsync(din[i].p, in + z0 * inLayer, zn * inLayer * sizeof(float), cudaMemcpyHostToDevice, st[i]), "H2D", __FILE__,
                         __LINE__)) {
                rc = FIMEXCU_ERROR;
                break;
            }
            rc = cache_apply_dev(c, din[i].as<float>(), zn, dout[i].as<float>(), badIn, fillOut, st[i]);
            if (rc != FIMEXCU_OK)
                break;
            if (!cuda_ok(cudaMemcpyAsync(out + z0 * outLayer, dout[i].p, zn * outLayer * sizeof(float), cudaMemcpyDeviceToHost, st[i]), "D2H",
                         __FILE__, __LINE__))
                rc = FIMEXCU_ERROR;
        }
        for (int i = 0; i < NS; ++i)
            if (!cuda_ok(cudaStreamSynchronize(st[i]), "cudaStreamSynchronize", __FILE__, __LINE__))
                rc = FIMEXCU_ERROR;
    }
    for (int i = 0; i < NS; ++i)
        cudaStreamDestroy(st[i]);
    return rc;
}

int fimexcu_cache_apply_uv_dev_f(const fimexcu_cache* c, const fimexcu_vecrot* r, const float* d_u_in, const float* d_v_in, size_t nz,
                                 float* d_u_out, float* d_v_out, float badU, float badV, float fillU, float fillV, void* stream)
{
    clear_error();
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply_uv: null cache");
    const size_t nxy = c->outX * c->outY;
    if (r && r->ox * r->oy != nxy)
        FXCU_FAIL("fimexcu_cache_apply_uv: rotation matrix and cache have different target grids");
    if (nz == 0 || nxy == 0)
        return FIMEXCU_OK;
    cudaStream_t s = as_stream(stream);
    if (c->kind != KIND_BILINEAR) {
        // bicubic / nearest: the general kernels per component, then the stand-alone rotation; the
        // rotation sees NaN (not fill), so the fill is applied by the rotation-free path only
        if (r && (!std::isnan(fillU) || !std::isnan(fillV)))
            FXCU_FAIL("fimexcu_cache_apply_uv: fill values with rotation are only supported for bilinear caches");
        if (cache_apply_dev(c, d_u_in, nz, d_u_out, badU, fillU, s) != FIMEXCU_OK || cache_apply_dev(c, d_v_in, nz, d_v_out, badV, fillV, s) != FIMEXCU_OK)
            return FIMEXCU_ERROR;
        return r ? fimexcu_vecrot_apply_dev_f(r, d_u_out, d_v_out, nz, stream) : FIMEXCU_OK;
    }
    const int inz = static_cast<int>(nz);
    const size_t inLayer = c->inX * c->inY;
    const int zchunk = pick_zchunk(nxy, inz);
    const dim3 grid(blocks_for(nxy, 256), static_cast<unsigned>((inz + zchunk - 1) / zchunk));
    const bool bad = !std::isnan(badU) || !std::isnan(badV), fill = !std::isnan(fillU) || !std::isnan(fillV);
    // a NaN bad value never compares equal, a NaN fill value re-writes NaN: the <true,true> variant is always correct
    const double* cs = r ? r->c.as<double>() : nullptr;
    const double* sn = r ? r->s.as<double>() : nullptr;
#define FXCU_UV(B, F, R)                                                                                                                      \
    bilinear_uv_kernel<B, F, R><<<grid, 256, 0, s>>>(c->off.as<int>(), c->wx.as<float>(), c->wy.as<float>(), cs, sn, d_u_in, d_v_in, d_u_out, \
                                                     d_v_out, nxy, inz, zchunk, inLayer, static_cast<int>(c->inX), badU, badV, fillU, fillV)
    if (r) {
        if (bad || fill)
            FXCU_UV(true, true, true);
        else
            FXCU_UV(false, false, true);
    } else {
        if (bad || fill)
            FXCU_UV(true, true, false);
        else
            FXCU_UV(false, false, false);
    }
#undef FXCU_UV
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}

int fimexcu_cache_apply_uv_f(const fimexcu_cache* c, const fimexcu_vecrot* r, const float* u_in, const float* v_in, size_t size, float* u_out,
                             float* v_out, size_t* newSize, float badU, float badV, float fillU, float fillV)
{
    clear_error();
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply_uv_f: null cache");
    if (!require_device())
        return FIMEXCU_ERROR;
    const size_t inLayer = c->inX * c->inY, outLayer = c->outX * c->outY;
    const size_t nz = size / inLayer;
    if (newSize)
        *newSize = nz * outLayer;
    if (nz == 0 || outLayer == 0)
        return FIMEXCU_OK;
    DevBuf du, dv, ou, ov;
    const size_t ib = nz * inLayer * sizeof(float), ob = nz * outLayer * sizeof(float);
    if (!du.alloc(ib, nullptr) || !dv.alloc(ib, nullptr) || !ou.alloc(ob, nullptr) || !ov.alloc(ob, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(du.p, u_in, ib, cudaMemcpyHostToDevice, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(dv.p, v_in, ib, cudaMemcpyHostToDevice, nullptr));
    if (fimexcu_cache_apply_uv_dev_f(c, r, du.as<float>(), dv.as<float>(), nz, ou.as<float>(), ov.as<float>(), badU, badV, fillU, fillV, nullptr) !=
        FIMEXCU_OK)
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(u_out, ou.p, ob, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(v_out, ov.p, ob, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

void fimexcu_cache_destroy(fimexcu_cache* c) { delete c; }

static int point_values(int method, const float* infield, float* outvalues, double x, double y, int ix, int iy, int iz)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    if (iz <= 0)
        return FIMEXCU_OK;
    const size_t nin = static_cast<size_t>(ix) * iy * iz;
    DevBuf din, dout;
    if (!din.alloc(nin * sizeof(float), nullptr) || !dout.alloc(iz * sizeof(float), nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(din.p, infield, nin * sizeof(float), cudaMemcpyHostToDevice, nullptr));
    point_kernel<<<1, 128>>>(method, din.as<float>(), dout.as<float>(), x, y, ix, iy, iz);
    FXCU_CUDA(cudaGetLastError());
    FXCU_CUDA(cudaMemcpyAsync(outvalues, dout.p, iz * sizeof(float), cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

int fimexcu_get_values_f(const float* infield, float* outvalues, double x, double y, int ix, int iy, int iz)
{
    return point_values(FIMEXCU_INTERPOL_NEAREST_NEIGHBOR, infield, outvalues, x, y, ix, iy, iz);
}

int fimexcu_get_values_bilinear_f(const float* infield, float* outvalues, double x, double y, int ix, int iy, int iz)
{
    return point_values(FIMEXCU_INTERPOL_BILINEAR, infield, outvalues, x, y, ix, iy, iz);
}

int fimexcu_get_values_bicubic_f(const float* infield, float* outvalues, double x, double y, int ix, int iy, int iz)
{
    return point_values(FIMEXCU_INTERPOL_BICUBIC, infield, outvalues, x, y, ix, iy, iz);
}

} // extern "C"
