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
#include "fxcu_hostfill.h"
#include "fxcu_hostpar.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <limits>
#include <emmintrin.h>
#include <functional>
#include <memory>
#include <thread>
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
// All apply kernels share one parameter block.  Layouts are given by element strides so that the same kernels
// serve the dense layout of the reference ([z][y][x], pitch == width) and the padded device-resident layout
// (rows on 128 B boundaries) that lets every warp store be one aligned line and lets TMA describe the source.
struct ApplyParams {
    const int* off;
    const void* wx;
    const void* wy;
    const void* in; // f32, or i16 for the packed kernels
    void* out;
    int outX, outY, nz, zchunk;
    int l2hint; // staged kernels: L2 eviction priority of the source boxes (tma_l2_policy)
    int inX;       // row length the table offsets refer to (the cache's reduced window)
    int inPitch;   // elements between source rows
    size_t inLevel;  // elements between source levels
    int outPitch;  // elements between target rows
    size_t outLevel; // elements between target levels
    float bad, fill;
    const int4* tile_meta; // per 32 x 8 target tile, see fimexcu_cache (staged kernels only)
    // direct-gather kernels only: outX x outY may be a window of the target grid (getOutputDataSlice, CachedInterpolation.cc:113-130);
    // tabX = 0 means the whole grid, else the tables' row length with the window starting at (winX0, winY0)
    int tabX = 0, winX0 = 0, winY0 = 0;
};

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

// Packed 16-bit slabs, the way CDMInterpolator::getDataSlice treats a `short` variable (CDMInterpolator.cc:140-151): the RAW
// values are cast to float (Data::asFloat), the variable's fill value becomes NaN (mifi_bad2nanf), and the result goes back
// through ScaleValue<float,short>(NaN, 1, 0, fill, 1, 0) (DataUtils.h:38-60): NaN -> (short)fill, else
// static_cast<short>(int(lround(double(v)))) (MathUtils.h:40-96) -- half away from zero, wrapping like the C casts.
// `bad` carries float(fill) on the way in and the bit pattern of short(fill) on the way out.
template <bool BAD>
__device__ __forceinline__ float load_in(const short* __restrict__ p, float bad)
{
    float v = static_cast<float>(__ldg(p));
    if (v == bad)
        v = nan_f();
    return v;
}

__device__ __forceinline__ short pack_i16(float v, float fill_bits)
{
    if (isnan(v))
        return static_cast<short>(__float_as_int(fill_bits));
    float t = truncf(v);
    if (fabsf(__fsub_rn(v, t)) >= 0.5f) // v - trunc(v) is exact
        t = __fadd_rn(t, copysignf(1.f, v));
    return static_cast<short>(static_cast<int>(__float2ll_rz(t)));
}

template <bool FILL>
__device__ __forceinline__ void store_out(short* __restrict__ p, float v, float fill_bits)
{
    *p = pack_i16(v, fill_bits);
}

// thread -> target cell of a 32 x 8 tile; returns false outside the grid (or the requested window of it)
__device__ __forceinline__ bool tile_cell(const ApplyParams& P, int& ox, int& oy)
{
    ox = blockIdx.x * 32 + (threadIdx.x & 31);
    oy = blockIdx.y * 8 + (threadIdx.x >> 5);
    return ox < P.outX && oy < P.outY;
}
// index of that cell in the cache's tables
__device__ __forceinline__ size_t table_index(const ApplyParams& P, int ox, int oy)
{
    if (P.tabX == 0)
        return static_cast<size_t>(oy) * P.outX + ox;
    return static_cast<size_t>(oy + P.winY0) * P.tabX + (ox + P.winX0);
}

// table offset (relative to a dense inX-wide level) -> offset in the pitched source level
__device__ __forceinline__ int source_offset(const ApplyParams& P, int f)
{
    if (P.inPitch == P.inX)
        return f;
    const int y = f / P.inX;
    return y * P.inPitch + (f - y * P.inX);
}

__device__ __forceinline__ float blend4(float s00, float s01, float s10, float s11, float fx, float fy, float omx, float omy)
{
    const float top = __fadd_rn(__fmul_rn(omx, s00), __fmul_rn(fx, s01));
    const float bot = __fadd_rn(__fmul_rn(omx, s10), __fmul_rn(fx, s11));
    return __fadd_rn(__fmul_rn(omy, top), __fmul_rn(fy, bot));
}

// value from 4 taps where a switched-off direction aliases the first tap (edge rules of interpolation.c:289-328)
__device__ __forceinline__ float bilinear_value(float s00, float s01, float s10, float s11, float fx, float fy)
{
    if (fx >= 0.f && fy >= 0.f)
        return blend4(s00, s01, s10, s11, fx, fy, __fsub_rn(1.f, fx), __fsub_rn(1.f, fy));
    if (fx >= 0.f)
        return __fadd_rn(__fmul_rn(__fsub_rn(1.f, fx), s00), __fmul_rn(fx, s01));
    if (fy >= 0.f)
        return __fadd_rn(__fmul_rn(__fsub_rn(1.f, fy), s00), __fmul_rn(fy, s10));
    return s00;
}

// K4, direct-gather form.  One thread per target cell of a 32 x 8 tile, blockIdx.z selects a chunk of levels; the
// table entry stays in registers for all its levels.  Used for small grids, for sources TMA cannot describe
// (unaligned base or pitch) and for tiles whose source footprint does not fit the staged kernel's box.
template <bool BAD, bool FILL, class T = float>
__global__ void __launch_bounds__(256) bilinear_apply_kernel(const ApplyParams P)
{
    int ox, oy;
    if (!tile_cell(P, ox, oy))
        return;
    const size_t o = table_index(P, ox, oy);
    const int z0 = blockIdx.z * P.zchunk;
    const int z1 = min(P.nz, z0 + P.zchunk);
    const int f = P.off[o];
    const float fx = static_cast<const float*>(P.wx)[o], fy = static_cast<const float*>(P.wy)[o];
    T* po = static_cast<T*>(P.out) + static_cast<size_t>(z0) * P.outLevel + static_cast<size_t>(oy) * P.outPitch + ox;
    if (f < 0) {
        for (int z = z0; z < z1; ++z, po += P.outLevel)
            store_out<FILL>(po, nan_f(), P.fill);
        return;
    }
    const T* p = static_cast<const T*>(P.in) + static_cast<size_t>(z0) * P.inLevel + source_offset(P, f);
    const int inX = P.inPitch;
    if (fx >= 0.f && fy >= 0.f) {
        const float omx = __fsub_rn(1.f, fx), omy = __fsub_rn(1.f, fy);
#pragma unroll 8
        for (int z = z0; z < z1; ++z, p += P.inLevel, po += P.outLevel) {
            const float s00 = load_in<BAD>(p, P.bad), s01 = load_in<BAD>(p + 1, P.bad);
            const float s10 = load_in<BAD>(p + inX, P.bad), s11 = load_in<BAD>(p + inX + 1, P.bad);
            store_out<FILL>(po, blend4(s00, s01, s10, s11, fx, fy, omx, omy), P.fill);
        }
    } else if (fx >= 0.f) { // linear in x on the nearest row (interpolation.c:289-299)
        const float omx = __fsub_rn(1.f, fx);
#pragma unroll 8
        for (int z = z0; z < z1; ++z, p += P.inLevel, po += P.outLevel) {
            const float s00 = load_in<BAD>(p, P.bad), s01 = load_in<BAD>(p + 1, P.bad);
            store_out<FILL>(po, __fadd_rn(__fmul_rn(omx, s00), __fmul_rn(fx, s01)), P.fill);
        }
    } else if (fy >= 0.f) { // linear in y on the nearest column (:311-319)
        const float omy = __fsub_rn(1.f, fy);
#pragma unroll 8
        for (int z = z0; z < z1; ++z, p += P.inLevel, po += P.outLevel) {
            const float s00 = load_in<BAD>(p, P.bad), s10 = load_in<BAD>(p + inX, P.bad);
            store_out<FILL>(po, __fadd_rn(__fmul_rn(omy, s00), __fmul_rn(fy, s10)), P.fill);
        }
    } else { // nearest cell (:321-328)
#pragma unroll 8
        for (int z = z0; z < z1; ++z, p += P.inLevel, po += P.outLevel)
            store_out<FILL>(po, load_in<BAD>(p, P.bad), P.fill);
    }
}

// 16 taps around p (tap (x0-1, y0-1)), weights in f64, the four column sums added into the f32 result one at a time
// (interpolation.c:388-406)
template <bool BAD, class T>
__device__ __forceinline__ float bicubic_sample(const T* __restrict__ p, int pitch, const double wxv[4], const double wyv[4], float bad)
{
    float v[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i)
            v[j][i] = load_in<BAD>(p + j * pitch + i, bad);
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        double col = 0.;
#pragma unroll
        for (int i = 0; i < 4; ++i)
            col = __dadd_rn(col, __dmul_rn(wxv[i], static_cast<double>(v[j][i])));
        acc = __double2float_rn(__dadd_rn(static_cast<double>(acc), __dmul_rn(col, wyv[j])));
    }
    return acc;
}

// K5.  Same mapping; 16 taps, weights in f64, the four column sums added into the f32 result one at a
// time (interpolation.c:388-406).
template <bool BAD, bool FILL, class T = float>
__global__ void __launch_bounds__(256) bicubic_apply_kernel(const ApplyParams P)
{
    int ox, oy;
    if (!tile_cell(P, ox, oy))
        return;
    const size_t o = table_index(P, ox, oy);
    const int z0 = blockIdx.z * P.zchunk;
    const int z1 = min(P.nz, z0 + P.zchunk);
    const int f = P.off[o];
    T* po = static_cast<T*>(P.out) + static_cast<size_t>(z0) * P.outLevel + static_cast<size_t>(oy) * P.outPitch + ox;
    if (f < 0) {
        for (int z = z0; z < z1; ++z, po += P.outLevel)
            store_out<FILL>(po, nan_f(), P.fill);
        return;
    }
    double wxv[4], wyv[4];
    cubic_weights(static_cast<const double*>(P.wx)[o], wxv);
    cubic_weights(static_cast<const double*>(P.wy)[o], wyv);
    const T* p = static_cast<const T*>(P.in) + static_cast<size_t>(z0) * P.inLevel + source_offset(P, f);
    const int inX = P.inPitch;
    for (int z = z0; z < z1; ++z, p += P.inLevel, po += P.outLevel)
        store_out<FILL>(po, bicubic_sample<BAD>(p, inX, wxv, wyv, P.bad), P.fill);
}

// K6.  Pure gather.
template <bool BAD, bool FILL, class T = float>
__global__ void __launch_bounds__(256) nearest_apply_kernel(const ApplyParams P)
{
    int ox, oy;
    if (!tile_cell(P, ox, oy))
        return;
    const size_t o = table_index(P, ox, oy);
    const int z0 = blockIdx.z * P.zchunk;
    const int z1 = min(P.nz, z0 + P.zchunk);
    const int f = P.off[o];
    T* po = static_cast<T*>(P.out) + static_cast<size_t>(z0) * P.outLevel + static_cast<size_t>(oy) * P.outPitch + ox;
    if (f < 0) {
        for (int z = z0; z < z1; ++z, po += P.outLevel)
            store_out<FILL>(po, nan_f(), P.fill);
        return;
    }
    const T* p = static_cast<const T*>(P.in) + static_cast<size_t>(z0) * P.inLevel + source_offset(P, f);
#pragma unroll 4
    for (int z = z0; z < z1; ++z, p += P.inLevel, po += P.outLevel)
        store_out<FILL>(po, load_in<BAD>(p, P.bad), P.fill);
}

// ------------------------------------------------------------------------------------------------
// K4, staged form (the headline kernel): warp-specialised TMA pipeline.
//   CTA = 8 consumer warps + 1 producer warp, target tile 32 x 8 (one row segment per consumer warp, so a warp
//   store is one 128 B line when the target rows sit on 128 B boundaries).  The CTA finds the bounding box of
//   the source cells its tile needs; the producer warp streams that box, TMA_ZB levels per stage, into a ring of
//   shared-memory stages with cp.async.bulk.tensor (3-D tensor map over [level][row][col]) and full/empty
//   mbarriers; consumers gather their four taps with LDS and never touch the global-load path.
//   Tiles whose footprint exceeds the box (strong shear / coarsening) use direct gathers instead.
// ------------------------------------------------------------------------------------------------
constexpr int TMA_BW = 32;    // box width  (source columns): one shared-memory row == 32 banks
constexpr int TMA_BH = 16;    // box height (source rows)
constexpr int TMA_ZB = 4;     // levels per stage
constexpr int TMA_STAGES = 4; // ring depth (32 KB per CTA, 6 CTAs per SM)
constexpr int TMA_TH = 8;     // consumer warps == tile rows
constexpr int TMA_THREADS = 32 * TMA_TH + 32;
// staged bicubic: 4 x 4 footprints need 3 more rows and columns than the bilinear box offers
constexpr int CUB_BH = 20;    // box height
constexpr int CUB_ZB = 4;     // levels per stage (10 KB)
constexpr int CUB_STAGES = 3;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity)
{
    uint32_t ok = 0;
    const uint32_t addr = smem_u32(bar);
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(ok)
            : "r"(addr), "r"(parity)
            : "memory");
    } while (!ok);
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int x, int y, int z)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(smem_u32(dst)),
                 "l"(map), "r"(smem_u32(bar)), "r"(x), "r"(y), "r"(z)
                 : "memory");
}

// L2 eviction priority of the staged source boxes: neighbouring tiles re-read each other's cells (the boxes over-fetch 2-4x), the
// result is written once with streaming stores -- boxes marked evict_last stay in L2 until their neighbours have had them:
// S->T1 0.2093 -> 0.2059 ms (80.0 %); evict_first, for comparison, 0.267 ms (profiles/exp/l2hint_ab.sh).  hint: 0 none, 1 evict_last,
// 2 evict_first (FIMEXCU_TMA_L2HINT, read once; default 1).
__device__ __forceinline__ uint64_t tma_l2_policy(int hint)
{
    uint64_t policy = 0;
    if (hint == 1)
        asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(policy));
    else if (hint == 2)
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
    return policy;
}
__device__ __forceinline__ void tma_load_3d_hint(void* dst, const CUtensorMap* map, uint64_t* bar, int x, int y, int z, uint64_t policy)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(smem_u32(dst)),
                 "l"(map), "r"(smem_u32(bar)), "r"(x), "r"(y), "r"(z), "l"(policy)
                 : "memory");
}

template <bool BAD>
__device__ __forceinline__ float tap(const float* t, int i, float bad)
{
    float v = t[i];
    if (BAD && v == bad)
        v = nan_f();
    return v;
}

// NN = true: nearest-neighbour tables (one tap, no weights) through the same pipeline
template <bool BAD, bool FILL, bool NN>
__global__ void __launch_bounds__(TMA_THREADS, 6) staged_apply_tma_kernel(const __grid_constant__ CUtensorMap tmap, const ApplyParams P)
{
    constexpr int STAGE_FLOATS = TMA_BW * TMA_BH * TMA_ZB;
    constexpr int LVL = TMA_BW * TMA_BH;
    __shared__ __align__(128) float tile[TMA_STAGES][STAGE_FLOATS];
    __shared__ __align__(8) uint64_t full[TMA_STAGES];
    __shared__ __align__(8) uint64_t empty[TMA_STAGES];
    static_assert(TMA_TH == 8, "tile_fit_kernel describes 32 x 8 tiles");
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const bool consumer = warp < TMA_TH;
    const int ox = blockIdx.x * 32 + lane, oy = blockIdx.y * TMA_TH + warp;
    const bool in_grid = consumer && ox < P.outX && oy < P.outY;
    // the tile's box and path come from the cache (computed once at build time): the producer can start streaming before a
    // single table entry of this tile has arrived, and no reduction over the tile is needed here
    const int4 meta = __ldg(P.tile_meta + blockIdx.y * gridDim.x + blockIdx.x);
    int f = -1;
    float fx = -1.f, fy = -1.f;
    if (in_grid) {
        const size_t o = static_cast<size_t>(oy) * P.outX + ox;
        f = P.off[o];
        if (!NN) {
            fx = static_cast<const float*>(P.wx)[o];
            fy = static_cast<const float*>(P.wy)[o];
        }
    }
    if (tid == 0) {
        for (int s = 0; s < TMA_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], TMA_TH);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // Programmatic dependent launch: the next launch on the stream may place its CTAs as soon as every CTA of this one has started,
    // so that its table loads and barrier set-up run while this launch drains; it touches no slab before the launch before it is
    // complete (the wait below).  The tables read above belong to the cache and are final when a cache handle exists.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    __syncthreads();
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int bx0 = meta.x, by0 = meta.y;
    const bool all_interior = (meta.z & TILE_ALL_INTERIOR) != 0;
    // blockIdx.z selects a chunk of levels (a multiple of TMA_ZB): few tiles x many levels would otherwise leave the last wave
    // of CTAs mostly empty
    const int zb = blockIdx.z * P.zchunk, nzc = min(P.nz, zb + P.zchunk) - zb;
    float* po = static_cast<float*>(P.out) + static_cast<size_t>(zb) * P.outLevel + static_cast<size_t>(oy) * P.outPitch + ox;
    if (!(meta.z & TILE_HAS_DATA)) { // nothing defined in this tile
        if (in_grid)
            for (int z = 0; z < nzc; ++z, po += P.outLevel)
                store_out<FILL>(po, nan_f(), P.fill);
        return;
    }
    const int y0 = (f >= 0) ? f / P.inX : 0, x0 = (f >= 0) ? f - y0 * P.inX : 0;
    if (!(meta.z & TILE_FITS)) { // footprint too large: direct gathers
        if (!in_grid)
            return;
        if (f < 0) {
            for (int z = 0; z < nzc; ++z, po += P.outLevel)
                store_out<FILL>(po, nan_f(), P.fill);
            return;
        }
        const int dxg = fx >= 0.f ? 1 : 0, dyg = fy >= 0.f ? P.inPitch : 0;
        const float* p = static_cast<const float*>(P.in) + static_cast<size_t>(zb) * P.inLevel + static_cast<size_t>(y0) * P.inPitch + x0;
#pragma unroll 4
        for (int z = 0; z < nzc; ++z, p += P.inLevel, po += P.outLevel)
            store_out<FILL>(po, bilinear_value(load_in<BAD>(p, P.bad), load_in<BAD>(p + dxg, P.bad), load_in<BAD>(p + dyg, P.bad),
                                               load_in<BAD>(p + dyg + dxg, P.bad), fx, fy),
                            P.fill);
        return;
    }
    constexpr unsigned BYTES = STAGE_FLOATS * sizeof(float);
    const int niter = (nzc + TMA_ZB - 1) / TMA_ZB;
    if (!consumer) { // producer warp: one lane keeps the ring full
        if (lane == 0) {
            const uint64_t policy = tma_l2_policy(P.l2hint);
            for (int it = 0; it < niter; ++it) {
                const int s = it % TMA_STAGES;
                mbar_wait(&empty[s], ((it / TMA_STAGES) & 1) ^ 1); // passes at once the first time round
                mbar_expect_tx(&full[s], BYTES);
                if (P.l2hint)
                    tma_load_3d_hint(&tile[s][0], &tmap, &full[s], bx0, by0, zb + it * TMA_ZB, policy);
                else
                    tma_load_3d(&tile[s][0], &tmap, &full[s], bx0, by0, zb + it * TMA_ZB);
            }
        }
        return;
    }
    const int r = (f >= 0) ? (y0 - by0) * TMA_BW + (x0 - bx0) : 0;
    if (NN) { // one tap per target cell
        for (int it = 0; it < niter; ++it) {
            const int s = it % TMA_STAGES;
            mbar_wait(&full[s], (it / TMA_STAGES) & 1);
            const float* t = &tile[s][r];
#pragma unroll
            for (int k = 0; k < TMA_ZB; ++k) {
                const float v = (f >= 0) ? tap<BAD>(t, k * LVL, P.bad) : nan_f();
                if (in_grid && it * TMA_ZB + k < nzc)
                    store_out<FILL>(po + k * P.outLevel, v, P.fill);
            }
            po += TMA_ZB * P.outLevel;
            __syncwarp();
            if (lane == 0)
                mbar_arrive(&empty[s]);
        }
    } else if (all_interior) {
        // undefined targets carry NaN weights: every product is NaN, no select in the loop
        const float wfx = (f >= 0) ? fx : nan_f(), wfy = (f >= 0) ? fy : nan_f();
        const float omx = __fsub_rn(1.f, wfx), omy = __fsub_rn(1.f, wfy);
        // tiles that lie wholly inside the target grid store whole stages without a test per level
        const bool whole_tile = blockIdx.x * 32 + 32 <= P.outX && blockIdx.y * TMA_TH + TMA_TH <= P.outY;
        for (int it = 0; it < niter; ++it) {
            const int s = it % TMA_STAGES;
            mbar_wait(&full[s], (it / TMA_STAGES) & 1);
            const float* t = &tile[s][r];
            if (whole_tile && (it + 1) * TMA_ZB <= nzc) {
#pragma unroll
                for (int k = 0; k < TMA_ZB; ++k)
                    store_out<FILL>(po + k * P.outLevel,
                                    blend4(tap<BAD>(t, k * LVL, P.bad), tap<BAD>(t, k * LVL + 1, P.bad), tap<BAD>(t, k * LVL + TMA_BW, P.bad),
                                           tap<BAD>(t, k * LVL + TMA_BW + 1, P.bad), wfx, wfy, omx, omy),
                                    P.fill);
            } else
#pragma unroll
            for (int k = 0; k < TMA_ZB; ++k) {
                const float v = blend4(tap<BAD>(t, k * LVL, P.bad), tap<BAD>(t, k * LVL + 1, P.bad), tap<BAD>(t, k * LVL + TMA_BW, P.bad),
                                       tap<BAD>(t, k * LVL + TMA_BW + 1, P.bad), wfx, wfy, omx, omy);
                if (in_grid && it * TMA_ZB + k < nzc)
                    store_out<FILL>(po + k * P.outLevel, v, P.fill);
            }
            po += TMA_ZB * P.outLevel;
            __syncwarp();
            if (lane == 0)
                mbar_arrive(&empty[s]);
        }
    } else { // tiles on the rim of the source grid: per-thread edge rules
        const int dxs = fx >= 0.f ? 1 : 0, dys = fy >= 0.f ? TMA_BW : 0;
        for (int it = 0; it < niter; ++it) {
            const int s = it % TMA_STAGES;
            mbar_wait(&full[s], (it / TMA_STAGES) & 1);
            const float* t = &tile[s][r];
#pragma unroll
            for (int k = 0; k < TMA_ZB; ++k) {
                float v = nan_f();
                if (f >= 0)
                    v = bilinear_value(tap<BAD>(t, k * LVL, P.bad), tap<BAD>(t, k * LVL + dxs, P.bad), tap<BAD>(t, k * LVL + dys, P.bad),
                                       tap<BAD>(t, k * LVL + dys + dxs, P.bad), fx, fy);
                if (in_grid && it * TMA_ZB + k < nzc)
                    store_out<FILL>(po + k * P.outLevel, v, P.fill);
            }
            po += TMA_ZB * P.outLevel;
            __syncwarp();
            if (lane == 0)
                mbar_arrive(&empty[s]);
        }
    }
}

// K5, staged form, opt-in arithmetic (fimexcu_cache_set_fast_bicubic).  Same ring as K4 with a 32 x 20 box (4 x 4 footprints).
// What bounds the reference's arithmetic on this machine is not memory: per value it is 16 f64 products + 16 f64 sums (the column
// sums XMF) and 4 + 4 more for the row combination, each rounded on its own (interpolation.c:395-406) = 40 FP64 operations --
// 0.41 ms per S->T1 slab at 64 FP64 lanes per SM and clock, so 40 % of the HBM roofline is the ceiling of ANY bit-exact form -- plus
// 16 f32->f64 widenings of the taps: on the conversion (XU) pipe in the direct kernel (87 % busy, profiles/r1_bicubic_ncu_raw.txt:
// 0.73 ms), or, tried in round 2, once per staged value into an f64 copy of the box in shared memory, where the 8-byte taps then
// cost two shared-memory wavefronts each (0.89 ms: slower, not kept; profiles/r2_kernel_variants.md).  The bit-exact bicubic
// therefore stays with the direct kernel.  This kernel is the variant inside BASELINE.json's 1e-5: weights narrowed to f32, 20
// fused multiply-adds per value straight from the f32 box; 16 four-byte taps per value make it shared-memory-bound (one wavefront
// per tap and warp: 0.33 ms per slab).  Undefined masks are identical to the reference's (NaN and Inf taps poison both forms alike).
template <bool BAD, bool FILL>
__global__ void __launch_bounds__(TMA_THREADS, 5) staged_bicubic_tma_kernel(const __grid_constant__ CUtensorMap tmap, const ApplyParams P)
{
    constexpr int LVL = TMA_BW * CUB_BH;
    constexpr int STAGE_FLOATS = LVL * CUB_ZB;
    __shared__ __align__(128) float raw[CUB_STAGES][STAGE_FLOATS];
    __shared__ __align__(8) uint64_t full[CUB_STAGES];
    __shared__ __align__(8) uint64_t empty[CUB_STAGES];
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const bool consumer = warp < TMA_TH;
    const int ox = blockIdx.x * 32 + lane, oy = blockIdx.y * TMA_TH + warp;
    const bool in_grid = consumer && ox < P.outX && oy < P.outY;
    const int4 meta = __ldg(P.tile_meta + blockIdx.y * gridDim.x + blockIdx.x);
    int f = -1;
    double fxd = 0., fyd = 0.;
    if (in_grid) {
        const size_t o = static_cast<size_t>(oy) * P.outX + ox;
        f = P.off[o];
        if (f >= 0) {
            fxd = static_cast<const double*>(P.wx)[o];
            fyd = static_cast<const double*>(P.wy)[o];
        }
    }
    if (tid == 0) {
        for (int s = 0; s < CUB_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], TMA_TH);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int bx0 = meta.x, by0 = meta.y;
    const int zb = blockIdx.z * P.zchunk, nzc = min(P.nz, zb + P.zchunk) - zb;
    float* po = static_cast<float*>(P.out) + static_cast<size_t>(zb) * P.outLevel + static_cast<size_t>(oy) * P.outPitch + ox;
    float wxf[4], wyf[4];
    {
        double wxv[4], wyv[4];
        cubic_weights(fxd, wxv);
        cubic_weights(fyd, wyv);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            wxf[i] = static_cast<float>(wxv[i]);
            wyf[i] = static_cast<float>(wyv[i]);
        }
    }
    const int y0 = (f >= 0) ? f / P.inX : by0, x0 = (f >= 0) ? f - y0 * P.inX : bx0;
    if (!(meta.z & TILE_HAS_DATA) || !(meta.z & TILE_FITS)) { // nothing defined, or a footprint beyond the box: no staging
        if (!in_grid)
            return;
        if (f < 0) {
            for (int z = 0; z < nzc; ++z, po += P.outLevel)
                store_out<FILL>(po, nan_f(), P.fill);
            return;
        }
        const float* p = static_cast<const float*>(P.in) + static_cast<size_t>(zb) * P.inLevel + static_cast<size_t>(y0) * P.inPitch + x0;
        for (int z = 0; z < nzc; ++z, p += P.inLevel, po += P.outLevel) {
            float acc = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float col = __fmul_rn(wxf[0], load_in<BAD>(p + j * P.inPitch, P.bad));
#pragma unroll
                for (int i = 1; i < 4; ++i)
                    col = __fmaf_rn(wxf[i], load_in<BAD>(p + j * P.inPitch + i, P.bad), col);
                acc = __fmaf_rn(col, wyf[j], acc);
            }
            store_out<FILL>(po, acc, P.fill);
        }
        return;
    }
    constexpr unsigned BYTES = STAGE_FLOATS * sizeof(float);
    const int niter = (nzc + CUB_ZB - 1) / CUB_ZB;
    if (!consumer) {
        if (lane == 0) {
            for (int it = 0; it < niter; ++it) {
                const int s = it % CUB_STAGES;
                mbar_wait(&empty[s], ((it / CUB_STAGES) & 1) ^ 1);
                mbar_expect_tx(&full[s], BYTES);
                if (P.l2hint)
                    tma_load_3d_hint(&raw[s][0], &tmap, &full[s], bx0, by0, zb + it * CUB_ZB, tma_l2_policy(P.l2hint));
                else
                    tma_load_3d(&raw[s][0], &tmap, &full[s], bx0, by0, zb + it * CUB_ZB);
            }
        }
        return;
    }
    const int r = (y0 - by0) * TMA_BW + (x0 - bx0);
    for (int it = 0; it < niter; ++it) {
        const int s = it % CUB_STAGES;
        mbar_wait(&full[s], (it / CUB_STAGES) & 1);
#pragma unroll
        for (int k = 0; k < CUB_ZB; ++k) {
            const float* t = &raw[s][k * LVL + r];
            float acc = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float col = __fmul_rn(wxf[0], tap<BAD>(t, j * TMA_BW, P.bad));
#pragma unroll
                for (int i = 1; i < 4; ++i)
                    col = __fmaf_rn(wxf[i], tap<BAD>(t, j * TMA_BW + i, P.bad), col);
                acc = __fmaf_rn(col, wyf[j], acc);
            }
            if (in_grid && it * CUB_ZB + k < nzc)
                store_out<FILL>(po + k * P.outLevel, f >= 0 ? acc : nan_f(), P.fill);
        }
        po += CUB_ZB * P.outLevel;
        __syncwarp();
        if (lane == 0)
            mbar_arrive(&empty[s]);
    }
}

// Staged form of the fused vector pass: u and v boxes of the same footprint travel through the same ring (two TMA loads per
// stage on one mbarrier), consumers interpolate both components, rotate (reproject.cc:488-507) and store.
constexpr int UV_ZB = 2;     // levels per stage (two fields per level)
constexpr int UV_STAGES = 3; // ring depth
struct UvStagedParams {
    const double* cs;
    const double* sn;
    const float* vin;
    float* vout;
    float badV, fillV;
};

template <bool BAD, bool FILL, bool ROT>
__global__ void __launch_bounds__(TMA_THREADS, 5) staged_uv_tma_kernel(const __grid_constant__ CUtensorMap umap, const __grid_constant__ CUtensorMap vmap,
                                                                       const ApplyParams P, const UvStagedParams Q)
{
    constexpr int LVL = TMA_BW * TMA_BH;
    constexpr int FIELD = LVL * UV_ZB;
    __shared__ __align__(128) float tile[UV_STAGES][2 * FIELD];
    __shared__ __align__(8) uint64_t full[UV_STAGES];
    __shared__ __align__(8) uint64_t empty[UV_STAGES];
    __shared__ int red[4][TMA_TH];
    __shared__ int bbox[4];

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const bool consumer = warp < TMA_TH;
    const int ox = blockIdx.x * 32 + lane, oy = blockIdx.y * TMA_TH + warp;
    const bool in_grid = consumer && ox < P.outX && oy < P.outY;
    int f = -1;
    float fx = -1.f, fy = -1.f;
    size_t o = 0;
    if (in_grid) {
        o = static_cast<size_t>(oy) * P.outX + ox;
        f = P.off[o];
        fx = static_cast<const float*>(P.wx)[o];
        fy = static_cast<const float*>(P.wy)[o];
    }
    int x0 = 0x3fffffff, y0 = 0x3fffffff, x1 = -1, y1 = -1;
    if (f >= 0) {
        y0 = f / P.inX;
        x0 = f - y0 * P.inX;
        x1 = x0 + (fx >= 0.f ? 1 : 0);
        y1 = y0 + (fy >= 0.f ? 1 : 0);
    }
    int lox = x0, loy = y0, hix = x1, hiy = y1;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        lox = min(lox, __shfl_xor_sync(0xffffffffu, lox, s));
        loy = min(loy, __shfl_xor_sync(0xffffffffu, loy, s));
        hix = max(hix, __shfl_xor_sync(0xffffffffu, hix, s));
        hiy = max(hiy, __shfl_xor_sync(0xffffffffu, hiy, s));
    }
    if (consumer && lane == 0) {
        red[0][warp] = lox;
        red[1][warp] = loy;
        red[2][warp] = hix;
        red[3][warp] = hiy;
    }
    if (tid == 0) {
        for (int s = 0; s < UV_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], TMA_TH);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0) {
        int a = red[0][0], b = red[1][0], c = red[2][0], d = red[3][0];
        for (int w = 1; w < TMA_TH; ++w) {
            a = min(a, red[0][w]);
            b = min(b, red[1][w]);
            c = max(c, red[2][w]);
            d = max(d, red[3][w]);
        }
        bbox[0] = a & ~7;
        bbox[1] = b;
        bbox[2] = c;
        bbox[3] = d;
    }
    __syncthreads();
    const int bx0 = bbox[0], by0 = bbox[1], bx1 = bbox[2], by1 = bbox[3];
    const int zb = blockIdx.z * P.zchunk, nzc = min(P.nz, zb + P.zchunk) - zb; // this CTA's chunk of levels (a multiple of UV_ZB)
    const size_t obase = static_cast<size_t>(zb) * P.outLevel + static_cast<size_t>(oy) * P.outPitch + ox;
    float* pu = static_cast<float*>(P.out) + obase;
    float* pv = Q.vout + obase;
    if (bx1 < 0) {
        if (in_grid)
            for (int z = 0; z < nzc; ++z, pu += P.outLevel, pv += P.outLevel) {
                store_out<FILL>(pu, nan_f(), P.fill);
                store_out<FILL>(pv, nan_f(), Q.fillV);
            }
        return;
    }
    double c = 1., s = 0.;
    if (ROT && in_grid) {
        c = Q.cs[o];
        s = Q.sn[o];
    }
    const bool staged = (bx1 - bx0 + 1 <= TMA_BW) && (by1 - by0 + 1 <= TMA_BH);
    const int niter = staged ? (nzc + UV_ZB - 1) / UV_ZB : 0;
    if (!consumer) {
        if (staged && lane == 0) {
            for (int it = 0; it < niter; ++it) {
                const int st = it % UV_STAGES;
                mbar_wait(&empty[st], ((it / UV_STAGES) & 1) ^ 1);
                mbar_expect_tx(&full[st], 2 * FIELD * sizeof(float));
                if (P.l2hint) {
                    const uint64_t policy = tma_l2_policy(P.l2hint);
                    tma_load_3d_hint(&tile[st][0], &umap, &full[st], bx0, by0, zb + it * UV_ZB, policy);
                    tma_load_3d_hint(&tile[st][FIELD], &vmap, &full[st], bx0, by0, zb + it * UV_ZB, policy);
                } else {
                    tma_load_3d(&tile[st][0], &umap, &full[st], bx0, by0, zb + it * UV_ZB);
                    tma_load_3d(&tile[st][FIELD], &vmap, &full[st], bx0, by0, zb + it * UV_ZB);
                }
            }
        }
        return;
    }
    // tap offsets: a switched-off direction aliases the first tap
    const int dxs = fx >= 0.f ? 1 : 0;
    if (!staged) { // direct gathers for this tile
        if (!in_grid)
            return;
        if (f < 0) {
            for (int z = 0; z < nzc; ++z, pu += P.outLevel, pv += P.outLevel) {
                store_out<FILL>(pu, nan_f(), P.fill);
                store_out<FILL>(pv, nan_f(), Q.fillV);
            }
            return;
        }
        const int dyg = fy >= 0.f ? P.inPitch : 0;
        const float* qu = static_cast<const float*>(P.in) + static_cast<size_t>(zb) * P.inLevel + static_cast<size_t>(y0) * P.inPitch + x0;
        const float* qv2 = Q.vin + static_cast<size_t>(zb) * P.inLevel + static_cast<size_t>(y0) * P.inPitch + x0;
        for (int z = 0; z < nzc; ++z, qu += P.inLevel, qv2 += P.inLevel, pu += P.outLevel, pv += P.outLevel) {
            float u = bilinear_value(load_in<BAD>(qu, P.bad), load_in<BAD>(qu + dxs, P.bad), load_in<BAD>(qu + dyg, P.bad),
                                     load_in<BAD>(qu + dyg + dxs, P.bad), fx, fy);
            float v = bilinear_value(load_in<BAD>(qv2, Q.badV), load_in<BAD>(qv2 + dxs, Q.badV), load_in<BAD>(qv2 + dyg, Q.badV),
                                     load_in<BAD>(qv2 + dyg + dxs, Q.badV), fx, fy);
            if (ROT) {
                const double du = static_cast<double>(u), dv = static_cast<double>(v);
                u = __double2float_rn(__dsub_rn(__dmul_rn(du, c), __dmul_rn(dv, s)));
                v = __double2float_rn(__dadd_rn(__dmul_rn(du, s), __dmul_rn(dv, c)));
            }
            store_out<FILL>(pu, u, P.fill);
            store_out<FILL>(pv, v, Q.fillV);
        }
        return;
    }
    const int dys = fy >= 0.f ? TMA_BW : 0;
    const int r = (f >= 0) ? (y0 - by0) * TMA_BW + (x0 - bx0) : 0;
    for (int it = 0; it < niter; ++it) {
        const int st = it % UV_STAGES;
        mbar_wait(&full[st], (it / UV_STAGES) & 1);
        const float* tu = &tile[st][r];
        const float* tv = tu + FIELD;
#pragma unroll
        for (int k = 0; k < UV_ZB; ++k) {
            float u = nan_f(), v = nan_f();
            if (f >= 0) {
                u = bilinear_value(tap<BAD>(tu, k * LVL, P.bad), tap<BAD>(tu, k * LVL + dxs, P.bad), tap<BAD>(tu, k * LVL + dys, P.bad),
                                   tap<BAD>(tu, k * LVL + dys + dxs, P.bad), fx, fy);
                v = bilinear_value(tap<BAD>(tv, k * LVL, Q.badV), tap<BAD>(tv, k * LVL + dxs, Q.badV), tap<BAD>(tv, k * LVL + dys, Q.badV),
                                   tap<BAD>(tv, k * LVL + dys + dxs, Q.badV), fx, fy);
                if (ROT) {
                    const double du = static_cast<double>(u), dv = static_cast<double>(v);
                    u = __double2float_rn(__dsub_rn(__dmul_rn(du, c), __dmul_rn(dv, s)));
                    v = __double2float_rn(__dadd_rn(__dmul_rn(du, s), __dmul_rn(dv, c)));
                }
            }
            if (in_grid && it * UV_ZB + k < nzc) {
                store_out<FILL>(pu + k * P.outLevel, u, P.fill);
                store_out<FILL>(pv + k * P.outLevel, v, Q.fillV);
            }
        }
        pu += UV_ZB * P.outLevel;
        pv += UV_ZB * P.outLevel;
        __syncwarp();
        if (lane == 0)
            mbar_arrive(&empty[st]);
    }
}

// fraction of 32 x 8 target tiles whose source footprint fits the staged kernel's box (decides, once per cache,
// whether the staged kernel is worth launching).  counts[0] = tiles with data, counts[1] = of those, tiles that fit
// cubic = true: `off` is the tap (x0-1, y0-1) of a 4 x 4 footprint (bicubic tables), box height CUB_BH
__global__ void __launch_bounds__(256) tile_fit_kernel(const int* __restrict__ off, const float* __restrict__ wx, const float* __restrict__ wy, int outX,
                                                       int outY, int inX, unsigned* __restrict__ counts, int4* __restrict__ meta, bool cubic = false)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ox = blockIdx.x * 32 + lane, oy = blockIdx.y * 8 + warp;
    int x0 = 0x3fffffff, y0 = 0x3fffffff, x1 = -1, y1 = -1;
    bool interior = wx != nullptr; // undefined cells and cells outside the grid do not spoil a tile's fast path
    if (ox < outX && oy < outY) {
        const size_t o = static_cast<size_t>(oy) * outX + ox;
        const int f = off[o];
        if (f >= 0) {
            y0 = f / inX;
            x0 = f - y0 * inX;
            const bool ex = cubic || (wx && wx[o] >= 0.f), ey = cubic || (wy && wy[o] >= 0.f);
            x1 = x0 + (cubic ? 3 : (ex ? 1 : 0));
            y1 = y0 + (cubic ? 3 : (ey ? 1 : 0));
            interior = ex && ey;
        }
    }
    const int all_interior = __syncthreads_and(interior ? 1 : 0);
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        x0 = min(x0, __shfl_xor_sync(0xffffffffu, x0, s));
        y0 = min(y0, __shfl_xor_sync(0xffffffffu, y0, s));
        x1 = max(x1, __shfl_xor_sync(0xffffffffu, x1, s));
        y1 = max(y1, __shfl_xor_sync(0xffffffffu, y1, s));
    }
    __shared__ int red[4][8];
    if (lane == 0) {
        red[0][warp] = x0;
        red[1][warp] = y0;
        red[2][warp] = x1;
        red[3][warp] = y1;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) {
            x0 = min(x0, red[0][w]);
            y0 = min(y0, red[1][w]);
            x1 = max(x1, red[2][w]);
            y1 = max(y1, red[3][w]);
        }
        int flags = all_interior ? TILE_ALL_INTERIOR : 0;
        if (x1 >= 0) {
            flags |= TILE_HAS_DATA;
            atomicAdd(&counts[0], 1u);
            if ((x1 - (x0 & ~7) + 1 <= TMA_BW) && (y1 - y0 + 1 <= (cubic ? CUB_BH : TMA_BH))) {
                flags |= TILE_FITS;
                atomicAdd(&counts[1], 1u);
            }
        }
        meta[blockIdx.y * gridDim.x + blockIdx.x] = make_int4(x0 & ~7, y0, flags, 0); // box starts on a 32 B sector
    }
}

// ------------------------------------------------------------------------------------------------
// fused vector pass (direct-gather form): bilinear u and v with one table read, the rotation of
// reproject.cc:488-507, then the stores.  Replaces four interpolateValues() calls and two reprojectValues()
// calls of the reference (CDMInterpolator.cc:232-259 runs once for x_wind and once for y_wind).
// ------------------------------------------------------------------------------------------------
struct UvParams {
    const double* cs;
    const double* sn;
    const float* vin;
    float* vout;
    float badV, fillV;
};

template <bool BAD, bool FILL, bool ROT>
__global__ void __launch_bounds__(256) bilinear_uv_kernel(const ApplyParams P, const UvParams Q)
{
    int ox, oy;
    if (!tile_cell(P, ox, oy))
        return;
    const size_t o = table_index(P, ox, oy);
    const int z0 = blockIdx.z * P.zchunk;
    const int z1 = min(P.nz, z0 + P.zchunk);
    const int f = P.off[o];
    const float fx = static_cast<const float*>(P.wx)[o], fy = static_cast<const float*>(P.wy)[o];
    const size_t obase = static_cast<size_t>(z0) * P.outLevel + static_cast<size_t>(oy) * P.outPitch + ox;
    float* pu = static_cast<float*>(P.out) + obase;
    float* pv = Q.vout + obase;
    if (f < 0) {
        for (int z = z0; z < z1; ++z, pu += P.outLevel, pv += P.outLevel) {
            store_out<FILL>(pu, nan_f(), P.fill);
            store_out<FILL>(pv, nan_f(), Q.fillV);
        }
        return;
    }
    double c = 1., s = 0.;
    if (ROT) {
        c = Q.cs[o];
        s = Q.sn[o];
    }
    // taps that a switched-off direction does not use alias the first tap (never read out of range)
    const int dx = fx >= 0.f ? 1 : 0;
    const int dy = fy >= 0.f ? P.inPitch : 0;
    const size_t ibase = static_cast<size_t>(z0) * P.inLevel + source_offset(P, f);
    const float* qu = static_cast<const float*>(P.in) + ibase;
    const float* qv = Q.vin + ibase;
#pragma unroll 2
    for (int z = z0; z < z1; ++z, qu += P.inLevel, qv += P.inLevel, pu += P.outLevel, pv += P.outLevel) {
        const float u00 = load_in<BAD>(qu, P.bad), u01 = load_in<BAD>(qu + dx, P.bad);
        const float u10 = load_in<BAD>(qu + dy, P.bad), u11 = load_in<BAD>(qu + dy + dx, P.bad);
        const float v00 = load_in<BAD>(qv, Q.badV), v01 = load_in<BAD>(qv + dx, Q.badV);
        const float v10 = load_in<BAD>(qv + dy, Q.badV), v11 = load_in<BAD>(qv + dy + dx, Q.badV);
        float u = bilinear_value(u00, u01, u10, u11, fx, fy);
        float v = bilinear_value(v00, v01, v10, v11, fx, fy);
        if (ROT) {
            const double du = static_cast<double>(u), dv = static_cast<double>(v);
            u = __double2float_rn(__dsub_rn(__dmul_rn(du, c), __dmul_rn(dv, s)));
            v = __double2float_rn(__dadd_rn(__dmul_rn(du, s), __dmul_rn(dv, c)));
        }
        store_out<FILL>(pu, u, P.fill);
        store_out<FILL>(pv, v, Q.fillV);
    }
}

// K7 for bicubic and nearest-neighbour tables: both components sampled and rotated in one pass (the reference runs
// interpolateValues twice and then reprojectValues over both results, src/CDMInterpolator.cc:232-259)
template <bool CUBIC, bool ROT>
__global__ void __launch_bounds__(256) sampled_uv_kernel(const ApplyParams P, const UvParams Q)
{
    int ox, oy;
    if (!tile_cell(P, ox, oy))
        return;
    const size_t o = table_index(P, ox, oy);
    const int z0 = blockIdx.z * P.zchunk;
    const int z1 = min(P.nz, z0 + P.zchunk);
    const int f = P.off[o];
    const size_t obase = static_cast<size_t>(z0) * P.outLevel + static_cast<size_t>(oy) * P.outPitch + ox;
    float* pu = static_cast<float*>(P.out) + obase;
    float* pv = Q.vout + obase;
    if (f < 0) {
        for (int z = z0; z < z1; ++z, pu += P.outLevel, pv += P.outLevel) {
            store_out<true>(pu, nan_f(), P.fill);
            store_out<true>(pv, nan_f(), Q.fillV);
        }
        return;
    }
    double c = 1., s = 0.;
    if (ROT) {
        c = Q.cs[o];
        s = Q.sn[o];
    }
    double wxv[4], wyv[4];
    if (CUBIC) {
        cubic_weights(static_cast<const double*>(P.wx)[o], wxv);
        cubic_weights(static_cast<const double*>(P.wy)[o], wyv);
    }
    const size_t ibase = static_cast<size_t>(z0) * P.inLevel + source_offset(P, f);
    const float* qu = static_cast<const float*>(P.in) + ibase;
    const float* qv = Q.vin + ibase;
    for (int z = z0; z < z1; ++z, qu += P.inLevel, qv += P.inLevel, pu += P.outLevel, pv += P.outLevel) {
        float u, v;
        if (CUBIC) {
            u = bicubic_sample<true>(qu, P.inPitch, wxv, wyv, P.bad);
            v = bicubic_sample<true>(qv, P.inPitch, wxv, wyv, Q.badV);
        } else {
            u = load_in<true>(qu, P.bad);
            v = load_in<true>(qv, Q.badV);
        }
        if (ROT) {
            const double du = static_cast<double>(u), dv = static_cast<double>(v);
            u = __double2float_rn(__dsub_rn(__dmul_rn(du, c), __dmul_rn(dv, s)));
            v = __double2float_rn(__dadd_rn(__dmul_rn(du, s), __dmul_rn(dv, c)));
        }
        store_out<true>(pu, u, P.fill);
        store_out<true>(pv, v, Q.fillV);
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
        const int dx = fx >= 0.f ? 1 : 0, dy = fy >= 0.f ? ix : 0;
        for (int z = threadIdx.x; z < iz; z += blockDim.x) {
            float r = nan_f();
            if (off >= 0) {
                const float* p = in + z * layer + off;
                r = bilinear_value(p[0], p[dx], p[dy], p[dy + dx], fx, fy);
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

// Levels per blockIdx.z chunk: at least `min_chunks` chunks, none longer than ~80 levels, each a multiple of `zb` levels.
// The cap: CTAs that walk hundreds of levels drift apart, and the source cells that neighbouring tiles share (the staged kernels'
// boxes over-fetch 2-4x) stop meeting in L2 -- S->T1 with 650 levels in one launch ran at 5.26 us per level with one chunk per
// tile against 3.09 us with chunks of 64 (T2, 1950 levels: 0.81 against 0.53; profiles/exp/zchunk_probe.py).  FIMEXCU_ZCHUNK_MAX
// overrides the cap (read once).
static int level_chunk(int nz, size_t min_chunks, int zb)
{
    static const int cap = []() {
        const char* e = getenv("FIMEXCU_ZCHUNK_MAX");
        const int v = e ? atoi(e) : 0;
        return v > 0 ? v : 80;
    }();
    if (nz <= 1)
        return std::max(nz, 1);
    size_t chunks = std::max<size_t>(std::max<size_t>(min_chunks, 1), (static_cast<size_t>(nz) + cap - 1) / cap);
    chunks = std::min<size_t>(chunks, static_cast<size_t>(nz));
    const size_t per = (static_cast<size_t>(nz) + chunks - 1) / chunks;
    return static_cast<int>((per + zb - 1) / zb * zb);
}

// direct kernels: enough chunks that small targets still fill the machine
static int pick_zchunk(size_t tiles, int nz)
{
    const size_t want = static_cast<size_t>(sm_count()) * 8 * 2; // two waves of 8 resident CTAs per SM
    return level_chunk(nz, tiles >= want ? 1 : (want + tiles - 1) / tiles, 1);
}

// per target row: first and one-past-last column whose table entry is defined (lo >= hi: none)
__global__ void __launch_bounds__(128) row_span_kernel(const int* __restrict__ off, int outX, int2* __restrict__ span)
{
    const int* row = off + static_cast<size_t>(blockIdx.x) * outX;
    int lo = outX, hi = 0;
    for (int x = threadIdx.x; x < outX; x += 128)
        if (__ldg(row + x) >= 0) {
            lo = min(lo, x);
            hi = max(hi, x + 1);
        }
    __shared__ int s_lo, s_hi;
    if (threadIdx.x == 0) {
        s_lo = outX;
        s_hi = 0;
    }
    __syncthreads();
    if (lo < hi) {
        atomicMin(&s_lo, lo);
        atomicMax(&s_hi, hi);
    }
    __syncthreads();
    if (threadIdx.x == 0)
        span[blockIdx.x] = make_int2(s_lo, s_hi);
}

// Row bands of a finished table (fimexcu_cache::bands): the per-row spans from the device, merged on the host into at most
// MAX_BANDS bands, always joining the two neighbours whose common rectangle adds the least area.
constexpr size_t MAX_BANDS = 24;
static bool build_row_bands(fimexcu_cache* c, cudaStream_t s)
{
    const size_t outX = c->outX, outY = c->outY;
    c->bands.clear();
    c->band_fraction = 1.0;
    if (outX == 0 || outY == 0 || outY > (1u << 20))
        return true;
    DevBuf dspan;
    if (!dspan.alloc(outY * sizeof(int2), s))
        return false;
    row_span_kernel<<<static_cast<unsigned>(outY), 128, 0, s>>>(c->off.as<int>(), static_cast<int>(outX), dspan.as<int2>());
    std::vector<int2> span(outY);
    if (!cuda_ok(cudaGetLastError(), "row_span_kernel", __FILE__, __LINE__) ||
        !cuda_ok(cudaMemcpyAsync(span.data(), dspan.p, outY * sizeof(int2), cudaMemcpyDeviceToHost, s), "cudaMemcpyAsync", __FILE__, __LINE__) ||
        !cuda_ok(cudaStreamSynchronize(s), "cudaStreamSynchronize", __FILE__, __LINE__))
        return false;
    std::vector<hostfill::Span> rows(outY);
    for (size_t y = 0; y < outY; ++y)
        rows[y] = hostfill::Span{span[y].x, span[y].y};
    double fraction = 1.0;
    const std::vector<hostfill::Band> merged = hostfill::merge_row_spans(rows, static_cast<int>(outX), MAX_BANDS, fraction);
    c->bands.clear();
    for (const hostfill::Band& b : merged)
        c->bands.push_back(fimexcu_cache::RowBand{b.y0, b.y1, b.x0, b.x1});
    c->band_fraction = fraction;
    return true;
}

fimexcu_cache* cache_build(int method, const double* d_px, const double* d_py, size_t inX, size_t inY, size_t outX, size_t outY, cudaStream_t s)
{
    if (!require_device())
        return nullptr;
    const size_t n = outX * outY;
    if (inX == 0 || inY == 0 || inX * inY >= 2147483647ull || outX >= (1ull << 30) || outY >= (1ull << 30) || n >= (1ull << 40)) {
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
        {
            DevBuf counts;
            if (!counts.alloc(2 * sizeof(unsigned), s))
                return nullptr;
            FXCU_CUDA_NULL(cudaMemsetAsync(counts.p, 0, 2 * sizeof(unsigned), s));
            const dim3 tg(static_cast<unsigned>((outX + 31) / 32), static_cast<unsigned>((outY + 7) / 8));
            if (!c->tile_meta.alloc_sync(static_cast<size_t>(tg.x) * tg.y * sizeof(int4)))
                return nullptr;
            tile_fit_kernel<<<tg, 256, 0, s>>>(c->off.as<int>(), nullptr, nullptr, static_cast<int>(outX), static_cast<int>(outY),
                                               static_cast<int>(c->inX), counts.as<unsigned>(), c->tile_meta.as<int4>());
            FXCU_CUDA_NULL(cudaGetLastError());
            unsigned hc[2] = {0, 0};
            FXCU_CUDA_NULL(cudaMemcpyAsync(hc, counts.p, sizeof(hc), cudaMemcpyDeviceToHost, s));
            FXCU_CUDA_NULL(cudaStreamSynchronize(s));
            c->tiles_with_data = hc[0];
            c->tiles_fitting = hc[1];
        }
        FXCU_CUDA_NULL(cudaStreamSynchronize(s));
        if (!build_row_bands(c.get(), s))
            return nullptr;
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
        FXCU_CUDA_NULL(cudaGetLastError());
        // how many 32 x 8 target tiles have a source footprint the staged (TMA) kernel can hold
        DevBuf counts;
        if (!counts.alloc(2 * sizeof(unsigned), s))
            return nullptr;
        FXCU_CUDA_NULL(cudaMemsetAsync(counts.p, 0, 2 * sizeof(unsigned), s));
        const dim3 tg(static_cast<unsigned>((outX + 31) / 32), static_cast<unsigned>((outY + 7) / 8));
        if (!c->tile_meta.alloc_sync(static_cast<size_t>(tg.x) * tg.y * sizeof(int4)))
            return nullptr;
        tile_fit_kernel<<<tg, 256, 0, s>>>(c->off.as<int>(), c->wx.as<float>(), c->wy.as<float>(), static_cast<int>(outX), static_cast<int>(outY),
                                           static_cast<int>(c->inX), counts.as<unsigned>(), c->tile_meta.as<int4>());
        FXCU_CUDA_NULL(cudaGetLastError());
        unsigned hc[2] = {0, 0};
        FXCU_CUDA_NULL(cudaMemcpyAsync(hc, counts.p, sizeof(hc), cudaMemcpyDeviceToHost, s));
        FXCU_CUDA_NULL(cudaStreamSynchronize(s));
        c->tiles_with_data = hc[0];
        c->tiles_fitting = hc[1];
    } else {
        if (!c->wx.alloc_sync(n * sizeof(double)) || !c->wy.alloc_sync(n * sizeof(double)))
            return nullptr;
        build_bicubic_kernel<<<nb, 256, 0, s>>>(d_px, d_py, n, static_cast<int>(c->inX), static_cast<int>(c->inY), static_cast<double>(minX),
                                                 static_cast<double>(minY), c->off.as<int>(), c->wx.as<double>(), c->wy.as<double>());
        FXCU_CUDA_NULL(cudaGetLastError());
        // 32 x 8 target tiles whose 4 x 4 footprints fit the staged bicubic kernel's 32 x 20 box
        DevBuf counts;
        if (!counts.alloc(2 * sizeof(unsigned), s))
            return nullptr;
        FXCU_CUDA_NULL(cudaMemsetAsync(counts.p, 0, 2 * sizeof(unsigned), s));
        const dim3 tg(static_cast<unsigned>((outX + 31) / 32), static_cast<unsigned>((outY + 7) / 8));
        if (!c->tile_meta.alloc_sync(static_cast<size_t>(tg.x) * tg.y * sizeof(int4)))
            return nullptr;
        tile_fit_kernel<<<tg, 256, 0, s>>>(c->off.as<int>(), nullptr, nullptr, static_cast<int>(outX), static_cast<int>(outY), static_cast<int>(c->inX),
                                           counts.as<unsigned>(), c->tile_meta.as<int4>(), true);
        FXCU_CUDA_NULL(cudaGetLastError());
        unsigned hc[2] = {0, 0};
        FXCU_CUDA_NULL(cudaMemcpyAsync(hc, counts.p, sizeof(hc), cudaMemcpyDeviceToHost, s));
        FXCU_CUDA_NULL(cudaStreamSynchronize(s));
        c->tiles_with_data = hc[0];
        c->tiles_fitting = hc[1];
    }
    FXCU_CUDA_NULL(cudaStreamSynchronize(s));
    if (!build_row_bands(c.get(), s))
        return nullptr;
    return c.release();
}

// ---- TMA tensor map over the source slab [nz][inY][inX] with row pitch / level stride (elements) ----
typedef CUresult (*tensor_map_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                         const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static tensor_map_encode_fn tensor_map_encoder()
{
    static tensor_map_encode_fn fn = []() -> tensor_map_encode_fn {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            return nullptr;
        }
        return reinterpret_cast<tensor_map_encode_fn>(p);
    }();
    return fn;
}

static bool make_source_map(CUtensorMap* m, const float* d_in, size_t inX, size_t inY, size_t nz, int inPitch, size_t inLevel, int depth = TMA_ZB,
                            int height = TMA_BH)
{
    tensor_map_encode_fn enc = tensor_map_encoder();
    if (!enc)
        return false;
    // TMA needs a 16 B aligned base and strides that are multiples of 16 B
    if ((reinterpret_cast<uintptr_t>(d_in) & 15u) != 0 || (inPitch & 3) != 0 || (inLevel & 3) != 0)
        return false;
    cuuint64_t dims[3] = {inX, inY, nz};
    cuuint64_t strides[2] = {static_cast<cuuint64_t>(inPitch) * 4, static_cast<cuuint64_t>(inLevel) * 4};
    cuuint32_t box[3] = {TMA_BW, static_cast<cuuint32_t>(height), static_cast<cuuint32_t>(depth)};
    cuuint32_t es[3] = {1, 1, 1};
    // L2 promotion of the box rows: 256 B measured 0.7 % faster than 128 B or none on S->T1 (0.2097 against 0.2112 ms), 64 B 3 % slower
    // (profiles/exp/l2promo_ab.sh); FIMEXCU_TMA_L2PROMO = 0 / 1 / 2 / 3 selects none / 64 / 128 / 256 B (read once)
    static const CUtensorMapL2promotion promo = []() {
        const char* e = getenv("FIMEXCU_TMA_L2PROMO");
        const int v = e ? atoi(e) : 3;
        return v == 0 ? CU_TENSOR_MAP_L2_PROMOTION_NONE : v == 1 ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B : v == 3 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B;
    }();
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(d_in), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static int tma_l2hint()
{
    static const int v = []() { const char* e = getenv("FIMEXCU_TMA_L2HINT"); return e ? atoi(e) : 1; }();
    return v;
}

// A/B switch between the staged (TMA) and the direct-gather kernels: the environment variable
// FIMEXCU_NO_TMA is read once per process; fimexcu_set_staged_kernels() overrides it at run time
static std::atomic<int> g_tma_override{-1};
static bool tma_allowed()
{
    static const bool env = getenv("FIMEXCU_NO_TMA") == nullptr;
    const int o = g_tma_override.load(std::memory_order_relaxed);
    return o < 0 ? env : o != 0;
}

struct Layout {
    int inPitch;
    size_t inLevel;
    int outPitch;
    size_t outLevel;
};

template <bool BAD, bool FILL>
static void launch_cache_apply(const fimexcu_cache* c, const ApplyParams& P0, cudaStream_t s, bool allow_tma)
{
    ApplyParams P = P0;
    const size_t tiles = static_cast<size_t>((P.outX + 31) / 32) * ((P.outY + 7) / 8);
    if ((c->kind == KIND_BILINEAR || c->kind == KIND_NN) && allow_tma && P.tabX == 0) {
        // staged kernel: worth it when most tiles fit its box and there are enough tiles to fill the GPU without
        // splitting the level loop
        CUtensorMap map;
        const bool enough = tiles >= static_cast<size_t>(sm_count()) * 6;
        const bool fits = c->tiles_with_data > 0 && 10ull * c->tiles_fitting >= 8ull * c->tiles_with_data;
        if (enough && fits && make_source_map(&map, static_cast<const float*>(P.in), c->inX, c->inY, static_cast<size_t>(P.nz), P.inPitch, P.inLevel)) {
            // at least ~8 waves of CTAs (6 per SM): split the levels when the tiles alone are too few
            const size_t slots = static_cast<size_t>(sm_count()) * 6;
            size_t zsplit = std::max<size_t>(1, (8 * slots + tiles - 1) / tiles); // 16 / 24 waves measured slower on T1
            zsplit = std::min<size_t>(zsplit, std::max<size_t>(1, static_cast<size_t>(P.nz) / (4 * TMA_ZB)));
            P.zchunk = level_chunk(P.nz, zsplit, TMA_ZB);
            P.l2hint = tma_l2hint();
            const dim3 grid(static_cast<unsigned>((c->outX + 31) / 32), static_cast<unsigned>((c->outY + TMA_TH - 1) / TMA_TH),
                            static_cast<unsigned>((P.nz + P.zchunk - 1) / P.zchunk));
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = grid;
            cfg.blockDim = dim3(TMA_THREADS);
            cfg.stream = s;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            static const bool pdl = getenv("FIMEXCU_NO_PDL") == nullptr;
            cfg.attrs = attr;
            cfg.numAttrs = pdl ? 1 : 0;
            if (c->kind == KIND_NN)
                cudaLaunchKernelEx(&cfg, staged_apply_tma_kernel<BAD, FILL, true>, map, P);
            else
                cudaLaunchKernelEx(&cfg, staged_apply_tma_kernel<BAD, FILL, false>, map, P);
            return;
        }
    }
    if (c->kind == KIND_BICUBIC && c->fast_bicubic && allow_tma && P.tabX == 0 && c->tile_meta.p) {
        CUtensorMap map;
        const bool enough = tiles >= static_cast<size_t>(sm_count()) * 3;
        const bool fits = c->tiles_with_data > 0 && 10ull * c->tiles_fitting >= 8ull * c->tiles_with_data;
        if (enough && fits &&
            make_source_map(&map, static_cast<const float*>(P.in), c->inX, c->inY, static_cast<size_t>(P.nz), P.inPitch, P.inLevel, CUB_ZB, CUB_BH)) {
            const size_t slots = static_cast<size_t>(sm_count()) * 5;
            size_t zsplit = std::max<size_t>(1, (8 * slots + tiles - 1) / tiles);
            zsplit = std::min<size_t>(zsplit, std::max<size_t>(1, static_cast<size_t>(P.nz) / (4 * CUB_ZB)));
            P.zchunk = level_chunk(P.nz, zsplit, CUB_ZB);
            P.l2hint = tma_l2hint();
            const dim3 grid(static_cast<unsigned>((c->outX + 31) / 32), static_cast<unsigned>((c->outY + TMA_TH - 1) / TMA_TH),
                            static_cast<unsigned>((P.nz + P.zchunk - 1) / P.zchunk));
            staged_bicubic_tma_kernel<BAD, FILL><<<grid, TMA_THREADS, 0, s>>>(map, P);
            return;
        }
    }
    P.zchunk = pick_zchunk(tiles, P.nz);
    const dim3 grid(static_cast<unsigned>((P.outX + 31) / 32), static_cast<unsigned>((P.outY + 7) / 8),
                    static_cast<unsigned>((P.nz + P.zchunk - 1) / P.zchunk));
    switch (c->kind) {
    case KIND_BILINEAR: bilinear_apply_kernel<BAD, FILL><<<grid, 256, 0, s>>>(P); break;
    case KIND_BICUBIC: bicubic_apply_kernel<BAD, FILL><<<grid, 256, 0, s>>>(P); break;
    default: nearest_apply_kernel<BAD, FILL><<<grid, 256, 0, s>>>(P); break;
    }
}

static int check_layout(const fimexcu_cache* c, const Layout& L, size_t nz)
{
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev != c->device) // the tables live in that device's memory
        FXCU_FAIL("fimexcu_cache_apply: the cache was built on device " + std::to_string(c->device) + ", the current device is " +
                  std::to_string(dev) + " (call fimexcu_set_device first)");
    if (L.inPitch < static_cast<long long>(c->inX) || L.inLevel < static_cast<size_t>(L.inPitch) * (c->inY - 1) + c->inX ||
        L.outPitch < static_cast<long long>(c->outX) || L.outLevel < static_cast<size_t>(L.outPitch) * (c->outY - 1) + c->outX)
        FXCU_FAIL("fimexcu_cache_apply: row pitch / level stride smaller than the grid");
    if (nz > 0x3fffffffull)
        FXCU_FAIL("fimexcu_cache_apply: too many levels");
    return FIMEXCU_OK;
}

// a window of the target grid (getOutputDataSlice, CachedInterpolation.cc:113-130); xSize == 0: the whole grid
struct Window {
    size_t xStart = 0, xSize = 0, yStart = 0, ySize = 0;
};

int cache_apply_strided(const fimexcu_cache* c, const float* d_in, const Layout& L, size_t nz, float* d_out, float bad, float fill, cudaStream_t s,
                        const Window& w = Window())
{
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply: null cache");
    if (nz == 0 || c->outX * c->outY == 0)
        return FIMEXCU_OK;
    if (w.xSize == 0 && check_layout(c, L, nz) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    ApplyParams P;
    P.off = c->off.as<int>();
    P.wx = c->wx.p;
    P.wy = c->wy.p;
    P.tile_meta = c->tile_meta.as<int4>();
    P.in = d_in;
    P.out = d_out;
    P.outX = static_cast<int>(c->outX);
    P.outY = static_cast<int>(c->outY);
    P.nz = static_cast<int>(nz);
    P.zchunk = P.nz;
    P.inX = static_cast<int>(c->inX);
    P.inPitch = L.inPitch;
    P.inLevel = L.inLevel;
    P.outPitch = L.outPitch;
    P.outLevel = L.outLevel;
    P.bad = bad;
    P.fill = fill;
    if (w.xSize != 0) {
        P.tabX = static_cast<int>(c->outX);
        P.winX0 = static_cast<int>(w.xStart);
        P.winY0 = static_cast<int>(w.yStart);
        P.outX = static_cast<int>(w.xSize);
        P.outY = static_cast<int>(w.ySize);
    }
    const bool has_bad = !std::isnan(bad), has_fill = !std::isnan(fill);
    const bool allow_tma = tma_allowed();
    if (has_bad && has_fill)
        launch_cache_apply<true, true>(c, P, s, allow_tma);
    else if (has_bad)
        launch_cache_apply<true, false>(c, P, s, allow_tma);
    else if (has_fill)
        launch_cache_apply<false, true>(c, P, s, allow_tma);
    else
        launch_cache_apply<false, false>(c, P, s, allow_tma);
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}

// packed 16-bit slabs (see load_in / pack_i16): direct-gather kernels on the raw values, fill <-> NaN fused on both sides
int cache_apply_strided_i16(const fimexcu_cache* c, const short* d_in, const Layout& L, size_t nz, short* d_out, double badValue, cudaStream_t s)
{
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply_i16: null cache");
    if (nz == 0 || c->outX * c->outY == 0)
        return FIMEXCU_OK;
    if (check_layout(c, L, nz) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    ApplyParams P;
    P.off = c->off.as<int>();
    P.wx = c->wx.p;
    P.wy = c->wy.p;
    P.tile_meta = c->tile_meta.as<int4>();
    P.in = d_in;
    P.out = d_out;
    P.outX = static_cast<int>(c->outX);
    P.outY = static_cast<int>(c->outY);
    P.nz = static_cast<int>(nz);
    P.inX = static_cast<int>(c->inX);
    P.inPitch = L.inPitch;
    P.inLevel = L.inLevel;
    P.outPitch = L.outPitch;
    P.outLevel = L.outLevel;
    P.bad = static_cast<float>(badValue); // mifi_bad2nanf's float parameter (CDMInterpolator.cc:143)
    const int fill_bits = static_cast<short>(badValue); // ScaleValue's newFill_ = static_cast<short>(newFill) (DataUtils.h:49)
    std::memcpy(&P.fill, &fill_bits, sizeof(float));
    const size_t tiles = static_cast<size_t>((c->outX + 31) / 32) * ((c->outY + 7) / 8);
    P.zchunk = pick_zchunk(tiles, P.nz);
    const dim3 grid(static_cast<unsigned>((c->outX + 31) / 32), static_cast<unsigned>((c->outY + 7) / 8),
                    static_cast<unsigned>((P.nz + P.zchunk - 1) / P.zchunk));
    switch (c->kind) {
    case KIND_BILINEAR: bilinear_apply_kernel<true, true, short><<<grid, 256, 0, s>>>(P); break;
    case KIND_BICUBIC: bicubic_apply_kernel<true, true, short><<<grid, 256, 0, s>>>(P); break;
    default: nearest_apply_kernel<true, true, short><<<grid, 256, 0, s>>>(P); break;
    }
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}

static Layout dense_layout(const fimexcu_cache* c)
{
    return Layout{static_cast<int>(c->inX), c->inX * c->inY, static_cast<int>(c->outX), c->outX * c->outY};
}

// rows of the device-resident layout start on 32 B (source: TMA + sector aligned boxes) and 128 B (target: one line per warp store)
static Layout padded_layout(const fimexcu_cache* c)
{
    const int ip = static_cast<int>((c->inX + 7) & ~size_t(7));
    const int op = static_cast<int>((c->outX + 31) & ~size_t(31));
    return Layout{ip, static_cast<size_t>(ip) * c->inY, op, static_cast<size_t>(op) * c->outY};
}

// streams of the host-pointer entry points: kept per host thread AND per device (a thread of the reference's OpenMP writer
// loop may serve different devices over its life), created on first use, never destroyed (the context may be gone at thread exit)
constexpr int NS_MAX = 8;
struct ThreadStreams {
    bool ready = false;
    cudaStream_t st[NS_MAX] = {};
    cudaEvent_t h2d[NS_MAX] = {}; // "the host-to-device copy of this stream's current chunk is done"
};
static ThreadStreams* thread_streams(int device)
{
    thread_local ThreadStreams pools[64];
    if (device < 0 || device >= 64) {
        set_error("device ordinal out of range");
        return nullptr;
    }
    ThreadStreams& p = pools[device];
    if (!p.ready) {
        cudaStream_t made[NS_MAX] = {};
        for (int i = 0; i < NS_MAX; ++i)
            if (!cuda_ok(cudaStreamCreateWithFlags(&made[i], cudaStreamNonBlocking), "cudaStreamCreateWithFlags", __FILE__, __LINE__)) {
                for (int k = 0; k < i; ++k)
                    cudaStreamDestroy(made[k]);
                return nullptr;
            }
        for (int i = 0; i < NS_MAX; ++i) {
            p.st[i] = made[i];
            if (!cuda_ok(cudaEventCreateWithFlags(&p.h2d[i], cudaEventDisableTiming), "cudaEventCreateWithFlags", __FILE__, __LINE__))
                return nullptr;
        }
        p.ready = true;
    }
    return &p;
}

static int e2e_streams()
{
    static const int v = []() { const char* e = getenv("FIMEXCU_E2E_STREAMS"); const int n = e ? atoi(e) : 3; return std::max(1, std::min(NS_MAX, n)); }();
    return v;
}
static int e2e_chunks()
{
    static const int v = []() { const char* e = getenv("FIMEXCU_E2E_CHUNKS"); const int n = e ? atoi(e) : 8; return std::max(1, n); }();
    return v;
}

static size_t e2e_copy_threads()
{
    static const size_t v = []() { const char* e = getenv("FIMEXCU_E2E_COPY_THREADS"); const int n = e ? atoi(e) : 4; return static_cast<size_t>(std::max(1, std::min(64, n))); }();
    return v;
}
static size_t e2e_fill_threads()
{
    static const size_t v = []() { const char* e = getenv("FIMEXCU_E2E_FILL_THREADS"); const int n = e ? atoi(e) : 4; return static_cast<size_t>(std::max(1, std::min(64, n))); }();
    return v;
}
// FIMEXCU_E2E_TRACE=1: one line per host-pointer call on stderr with where its wall time went (for integrators: which of the
// host copy, the PCIe queue and the margin fill bounds a slab on their machine)
static bool e2e_trace()
{
    static const bool v = []() { const char* e = getenv("FIMEXCU_E2E_TRACE"); return e && atoi(e) != 0; }();
    return v;
}
struct E2ETrace {
    typedef std::chrono::steady_clock clock;
    bool on = e2e_trace();
    clock::time_point t0 = clock::now();
    double copy_in = 0, wait = 0, enqueue = 0, drain = 0, fill = 0;
    static double ms(clock::time_point a, clock::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); }
    template <class F> void timed(double& acc, F&& f)
    {
        if (!on) {
            f();
            return;
        }
        const clock::time_point a = clock::now();
        f();
        acc += ms(a, clock::now());
    }
    void report(size_t nz, bool banded, bool bounce_in, bool bounce_out) const
    {
        if (on)
            fprintf(stderr, "fimexcu e2e: %zu levels%s%s%s: total %.2f ms = host copy-in %.2f + waits for buffers %.2f + enqueue %.2f + drain %.2f (margin fill, concurrent: %.2f)\n",
                    nz, banded ? ", row bands" : "", bounce_in ? ", pageable in" : "", bounce_out ? ", pageable out" : "", ms(t0, clock::now()), copy_in, wait, enqueue, drain,
                    fill);
    }
};

static bool e2e_bands() // FIMEXCU_E2E_BANDS=0: copy whole result slabs back (for measurements)
{
    static const bool v = []() { const char* e = getenv("FIMEXCU_E2E_BANDS"); return !(e && atoi(e) == 0); }();
    return v;
}

// pageable <-> pinned copies of one level chunk on the host cores (a single memcpy thread runs at a fraction of PCIe speed)
static void host_copy(void* dst, const void* src, size_t bytes)
{
    const size_t threads = e2e_copy_threads();
    const size_t grain = std::max<size_t>(size_t(1) << 20, (bytes + threads - 1) / threads);
    parallel_ranges(bytes, grain, [&](size_t a, size_t b) { std::memcpy(static_cast<char*>(dst) + a, static_cast<const char*>(src) + a, b - a); });
}

// host slab -> host slab through the device; T = float or short, `apply` = the strided device apply for that element type,
// `undef` = what that apply stores where the table has no source cell.  NF fields share the call (1, or 2 for the components of a
// vector: each chunk carries both through one fused kernel).  c is the table set on the CURRENT device.
template <class T, int NF, class ApplyFn>
static int cache_apply_host(const fimexcu_cache* c, const T* const (&in)[NF], size_t size, T* const (&out)[NF], size_t* newSize, const T (&undef)[NF],
                            ApplyFn apply)
{
    const size_t inLayer = c->inX * c->inY, outLayer = c->outX * c->outY;
    const size_t nz = size / inLayer; // CachedInterpolation.cc:176
    if (newSize)
        *newSize = nz * outLayer;
    if (nz == 0 || outLayer == 0)
        return FIMEXCU_OK;
    // Levels are independent: the slab is cut into level chunks; copy-in, kernel and copy-out of neighbouring chunks
    // run on different streams so that PCIe traffic in both directions overlaps the kernels.  This path is bound by the
    // PCIe copies (the 767 MB result alone takes ~14 ms, the kernels 0.4 ms), so the chunks keep the dense host layout on
    // the device: plain 1-D copies run at full DMA speed, strided 2-D copies into the padded layout measured 1.5x slower.
    const Layout L = dense_layout(c);
    const int NS = e2e_streams();
    // small slabs are latency-bound: one chunk per 16 MB of traffic (at most FIMEXCU_E2E_CHUNKS)
    const size_t traffic = NF * nz * (inLayer + outLayer) * sizeof(T);
    const size_t nchunk = std::max<size_t>(1, std::min<size_t>(static_cast<size_t>(e2e_chunks()), traffic / (16u << 20)));
    const size_t chunk = std::max<size_t>(1, std::min<size_t>(nz, (nz + nchunk - 1) / nchunk));
    ThreadStreams* ts = thread_streams(c->device);
    if (!ts)
        return FIMEXCU_ERROR;
    cudaStream_t* st = ts->st;
    // The reference hands ordinary `new float[]` memory (include/fimex/SharedArray.h:42-48).  The DMA engines cannot read or
    // write pageable memory: such a slab goes through pinned bounce buffers from the library's pool, one per stream, filled /
    // drained by the host cores while the neighbouring chunk is on the bus.
    bool bounce_in = false, bounce_out = false;
    for (int f = 0; f < NF; ++f) {
        bounce_in = bounce_in || is_pageable_host(in[f]);
        bounce_out = bounce_out || is_pageable_host(out[f]);
    }
    // Only the row bands of the table that can hold data cross the bus (pitched device->host copies run at the speed of plain
    // ones: profiles/exp/d2h_probe.cu); the rest of the result is the undefined value, written by host cores meanwhile.
    const bool banded = e2e_bands() && !c->bands.empty() && c->band_fraction < 0.9 && NF * nz * outLayer * sizeof(T) >= (size_t(8) << 20);
    const size_t outX = c->outX, outY = c->outY;
    E2ETrace trace;
    std::thread filler;
    std::function<void()> start_filler;
    std::vector<int> row_x0, row_x1; // per target row: the span that is copied back (empty: x0 == x1 == 0)
    if (banded) {
        row_x0.assign(outY, 0);
        row_x1.assign(outY, 0);
        for (const fimexcu_cache::RowBand& B : c->bands)
            if (B.x0 < B.x1)
                for (int y = B.y0; y < B.y1; ++y) {
                    row_x0[static_cast<size_t>(y)] = B.x0;
                    row_x1[static_cast<size_t>(y)] = B.x1;
                }
        // everything between the spans, as runs that continue across row ends (the right margin of a row and the left margin of
        // the next one are one piece of memory); each thread takes a range of (field, level, row) triples
        auto fill_margins = [&]() {
            const E2ETrace::clock::time_point f0 = E2ETrace::clock::now();
            const size_t rows = nz * outY;
            parallel_ranges(NF * rows, 256, [&](size_t a, size_t b) {
                while (a < b) {
                    const size_t f = a / rows, e = std::min(b, (f + 1) * rows); // the part of [a, b) inside field f
                    T* const base = out[f];
                    T* run = base + (a - f * rows) * outX; // start of the current undefined run
                    for (size_t k = a - f * rows; k < e - f * rows; ++k) {
                        const size_t y = k % outY;
                        T* row = base + k * outX;
                        if (row_x0[y] < row_x1[y]) {
                            hostfill::fill_lines(run, static_cast<size_t>(row + row_x0[y] - run), undef[f]);
                            run = row + row_x1[y];
                        }
                    }
                    hostfill::fill_lines(run, static_cast<size_t>(base + (e - f * rows) * outX - run), undef[f]);
                    a = e;
                }
                _mm_sfence();
            }, e2e_fill_threads());
            trace.fill = E2ETrace::ms(f0, E2ETrace::clock::now());
        };
        start_filler = [&, fill_margins]() {
            try {
                filler = std::thread(fill_margins);
            } catch (...) { // no thread to be had: write the margins here
                fill_margins();
            }
        };
        // a pageable input keeps the host cores busy with its bounce copies until the last chunk is queued, and those copies are
        // what the device waits for: the margins are then written afterwards, while the queue drains
        if (!bounce_in)
            start_filler();
    }
    struct JoinOnExit {
        std::thread& t;
        ~JoinOnExit()
        {
            if (t.joinable())
                t.join();
        }
    } join_filler{filler};
    struct Pinned {
        void* p = nullptr;
        ~Pinned() { host_pool_free(p); }
    };
    Pinned hin[NS_MAX], hout[NS_MAX]; // per stream: the NF fields of a chunk one after the other
    struct Pending {                  // a chunk whose result still sits in hout[i]
        bool live = false;
        size_t z0 = 0, zn = 0;
    } pend[NS_MAX];
    const size_t inChunk = chunk * L.inLevel, outChunk = chunk * L.outLevel; // elements of one field of a chunk (dense: = chunk * layer)
    int rc = FIMEXCU_OK;
    {
        DevBuf din[NS_MAX], dout[NS_MAX];
        for (int i = 0; i < NS && rc == FIMEXCU_OK; ++i) {
            if (!din[i].alloc(NF * inChunk * sizeof(T), st[i]) || !dout[i].alloc(NF * outChunk * sizeof(T), st[i]))
                rc = FIMEXCU_ERROR;
            if (rc == FIMEXCU_OK && bounce_in && !(hin[i].p = host_pool_alloc(NF * inChunk * sizeof(T))))
                rc = FIMEXCU_ERROR;
            if (rc == FIMEXCU_OK && bounce_out && !(hout[i].p = host_pool_alloc(NF * outChunk * sizeof(T))))
                rc = FIMEXCU_ERROR;
        }
        // device result of a chunk of one field -> host (the caller's array or this stream's bounce buffer, same dense layout)
        auto copy_back = [&](int i, const T* src, T* dst, size_t zn) {
            if (!banded)
                return cuda_ok(cudaMemcpyAsync(dst, src, zn * outLayer * sizeof(T), cudaMemcpyDeviceToHost, st[i]), "D2H", __FILE__, __LINE__);
            for (const fimexcu_cache::RowBand& B : c->bands) {
                if (B.x0 >= B.x1)
                    continue;
                cudaMemcpy3DParms p = {};
                p.srcPtr = make_cudaPitchedPtr(const_cast<T*>(src), outX * sizeof(T), outX * sizeof(T), outY);
                p.dstPtr = make_cudaPitchedPtr(dst, outX * sizeof(T), outX * sizeof(T), outY);
                p.srcPos = make_cudaPos(static_cast<size_t>(B.x0) * sizeof(T), static_cast<size_t>(B.y0), 0);
                p.dstPos = p.srcPos;
                p.extent = make_cudaExtent(static_cast<size_t>(B.x1 - B.x0) * sizeof(T), static_cast<size_t>(B.y1 - B.y0), zn);
                p.kind = cudaMemcpyDeviceToHost;
                if (!cuda_ok(cudaMemcpy3DAsync(&p, st[i]), "D2H (row band)", __FILE__, __LINE__))
                    return false;
            }
            return true;
        };
        auto drain = [&](int i) { // wait for stream i; move a finished bounced result to the caller's arrays
            if (!cuda_ok(cudaStreamSynchronize(st[i]), "cudaStreamSynchronize", __FILE__, __LINE__))
                rc = FIMEXCU_ERROR;
            for (int f = 0; f < NF && pend[i].live && rc == FIMEXCU_OK; ++f) {
                T* dst = out[f] + pend[i].z0 * outLayer;
                const T* src = static_cast<const T*>(hout[i].p) + f * outChunk;
                if (!banded)
                    host_copy(dst, src, pend[i].zn * outLayer * sizeof(T));
                else
                    parallel_ranges(pend[i].zn * outY, 64, [&](size_t a, size_t b) {
                        for (size_t k = a; k < b; ++k) {
                            const size_t y = k % outY;
                            if (row_x0[y] < row_x1[y]) {
                                const size_t o = k * outX + static_cast<size_t>(row_x0[y]);
                                hostfill::copy_lines(dst + o, src + o, static_cast<size_t>(row_x1[y] - row_x0[y]) * sizeof(T));
                            }
                        }
                        _mm_sfence();
                    }, e2e_copy_threads());
            }
            pend[i].live = false;
        };
        size_t k = 0;
        // (shorter first and last chunks -- to start and to end the bus traffic sooner -- measured no different: 15.2 / 15.0 / 15.1 ms)
        for (size_t z0 = 0, zn = 0; z0 < nz && rc == FIMEXCU_OK; z0 += zn, ++k) {
            zn = std::min(chunk, nz - z0);
            const int i = static_cast<int>(k % NS);
            if (k >= static_cast<size_t>(NS)) { // the bounce buffers of this stream must be free again
                trace.timed(trace.wait, [&]() {
                    if (bounce_out)
                        drain(i);
                    else if (bounce_in && !cuda_ok(cudaEventSynchronize(ts->h2d[i]), "cudaEventSynchronize", __FILE__, __LINE__))
                        rc = FIMEXCU_ERROR;
                });
                if (rc != FIMEXCU_OK)
                    break;
            }
            const T* din_f[NF];
            T* dout_f[NF];
            for (int f = 0; f < NF && rc == FIMEXCU_OK; ++f) {
                const T* src = in[f] + z0 * inLayer;
                if (bounce_in) {
                    T* hb = static_cast<T*>(hin[i].p) + f * inChunk;
                    trace.timed(trace.copy_in, [&]() { host_copy(hb, src, zn * inLayer * sizeof(T)); });
                    src = hb;
                }
                const E2ETrace::clock::time_point e0 = trace.on ? E2ETrace::clock::now() : E2ETrace::clock::time_point();
                T* d = din[i].template as<T>() + f * inChunk;
                if (!cuda_ok(cudaMemcpyAsync(d, src, zn * inLayer * sizeof(T), cudaMemcpyHostToDevice, st[i]), "H2D", __FILE__, __LINE__))
                    rc = FIMEXCU_ERROR;
                din_f[f] = d;
                dout_f[f] = dout[i].template as<T>() + f * outChunk;
                if (trace.on)
                    trace.enqueue += E2ETrace::ms(e0, E2ETrace::clock::now());
            }
            if (rc != FIMEXCU_OK)
                break;
            const E2ETrace::clock::time_point e0 = trace.on ? E2ETrace::clock::now() : E2ETrace::clock::time_point();
            if (bounce_in && !cuda_ok(cudaEventRecord(ts->h2d[i], st[i]), "cudaEventRecord", __FILE__, __LINE__)) {
                rc = FIMEXCU_ERROR;
                break;
            }
            rc = apply(din_f, L, zn, dout_f, st[i]);
            if (rc != FIMEXCU_OK)
                break;
            for (int f = 0; f < NF && rc == FIMEXCU_OK; ++f)
                if (!copy_back(i, dout_f[f], bounce_out ? static_cast<T*>(hout[i].p) + f * outChunk : out[f] + z0 * outLayer, zn))
                    rc = FIMEXCU_ERROR;
            if (bounce_out) {
                pend[i].live = true;
                pend[i].z0 = z0;
                pend[i].zn = zn;
            }
            if (trace.on)
                trace.enqueue += E2ETrace::ms(e0, E2ETrace::clock::now());
        }
        if (banded && bounce_in && rc == FIMEXCU_OK)
            start_filler();
        trace.timed(trace.drain, [&]() {
            for (int i = 0; i < NS; ++i)
                drain(i);
        });
    }
    if (filler.joinable())
        filler.join();
    trace.report(nz, banded, bounce_in, bounce_out);
    return rc;
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

int fimexcu_cache_padded_layout(const fimexcu_cache* c, size_t* in_row_pitch, size_t* in_level_stride, size_t* out_row_pitch, size_t* out_level_stride)
{
    if (!c)
        FXCU_FAIL("fimexcu_cache_padded_layout: null cache");
    const Layout L = padded_layout(c);
    if (in_row_pitch) *in_row_pitch = static_cast<size_t>(L.inPitch);
    if (in_level_stride) *in_level_stride = L.inLevel;
    if (out_row_pitch) *out_row_pitch = static_cast<size_t>(L.outPitch);
    if (out_level_stride) *out_level_stride = L.outLevel;
    return FIMEXCU_OK;
}

int fimexcu_cache_row_bands(const fimexcu_cache* c, int* bands4, int max_bands, double* fraction)
{
    clear_error();
    if (!c)
        return 0;
    if (fraction)
        *fraction = c->band_fraction;
    const int n = static_cast<int>(c->bands.size());
    for (int k = 0; bands4 && k < n && k < max_bands; ++k) {
        bands4[4 * k] = c->bands[k].y0;
        bands4[4 * k + 1] = c->bands[k].y1;
        bands4[4 * k + 2] = c->bands[k].x0;
        bands4[4 * k + 3] = c->bands[k].x1;
    }
    return n;
}

int fimexcu_cache_apply_dev_f(const fimexcu_cache* c, const float* d_in, size_t nz, float* d_out, float badIn, float fillOut, void* stream)
{
    clear_error();
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply: null cache");
    if (!(c = on_current_device(c, "fimexcu_cache_apply")))
        return FIMEXCU_ERROR;
    return cache_apply_strided(c, d_in, dense_layout(c), nz, d_out, badIn, fillOut, as_stream(stream));
}

int fimexcu_cache_apply_dev_strided_f(const fimexcu_cache* c, const float* d_in, size_t in_row_pitch, size_t in_level_stride, size_t nz, float* d_out,
                                      size_t out_row_pitch, size_t out_level_stride, float badIn, float fillOut, void* stream)
{
    clear_error();
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply: null cache");
    if (in_row_pitch > 0x7fffffffull || out_row_pitch > 0x7fffffffull)
        FXCU_FAIL("fimexcu_cache_apply: row pitch out of range");
    if (!(c = on_current_device(c, "fimexcu_cache_apply")))
        return FIMEXCU_ERROR;
    const Layout L{static_cast<int>(in_row_pitch), in_level_stride, static_cast<int>(out_row_pitch), out_level_stride};
    return cache_apply_strided(c, d_in, L, nz, d_out, badIn, fillOut, as_stream(stream));
}

static int check_window(const fimexcu_cache* c, const Layout& L, size_t nz, size_t xStart, size_t xSize, size_t yStart, size_t ySize)
{
    if (xSize == 0 || ySize == 0 || xStart + xSize > c->outX || yStart + ySize > c->outY)
        FXCU_FAIL("fimexcu_cache_apply_window: the window must be a non-empty part of the target grid");
    const Layout full{L.inPitch, L.inLevel, static_cast<int>(c->outX), c->outX * c->outY}; // the source side is checked as usual
    return check_layout(c, full, nz);
}

int fimexcu_cache_apply_window_dev_f(const fimexcu_cache* c, const float* d_in, size_t nz, size_t xStart, size_t xSize, size_t yStart, size_t ySize,
                                     float* d_out, float badIn, float fillOut, void* stream)
{
    clear_error();
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply_window: null cache");
    if (!(c = on_current_device(c, "fimexcu_cache_apply_window")))
        return FIMEXCU_ERROR;
    Layout L = dense_layout(c);
    if (check_window(c, L, nz, xStart, xSize, yStart, ySize) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    L.outPitch = static_cast<int>(xSize);
    L.outLevel = xSize * ySize;
    return cache_apply_strided(c, d_in, L, nz, d_out, badIn, fillOut, as_stream(stream), Window{xStart, xSize, yStart, ySize});
}

int fimexcu_cache_apply_window_f(const fimexcu_cache* c, const float* in, size_t size, size_t xStart, size_t xSize, size_t yStart, size_t ySize, float* out,
                                 size_t* newSize, float badIn, float fillOut)
{
    clear_error();
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply_window: null cache");
    if (!require_device() || !(c = on_current_device(c, "fimexcu_cache_apply_window")))
        return FIMEXCU_ERROR;
    const size_t inLayer = c->inX * c->inY;
    const size_t nz = size / inLayer;
    if (newSize)
        *newSize = nz * xSize * ySize;
    if (check_window(c, dense_layout(c), nz, xStart, xSize, yStart, ySize) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    if (nz == 0)
        return FIMEXCU_OK;
    DevBuf din, dout;
    const size_t ib = nz * inLayer * sizeof(float), ob = nz * xSize * ySize * sizeof(float);
    if (!din.alloc(ib, nullptr) || !dout.alloc(ob, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(din.p, in, ib, cudaMemcpyHostToDevice, nullptr));
    if (fimexcu_cache_apply_window_dev_f(c, din.as<float>(), nz, xStart, xSize, yStart, ySize, dout.as<float>(), badIn, fillOut, nullptr) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(out, dout.p, ob, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

int fimexcu_cache_apply_f(const fimexcu_cache* c, const float* in, size_t size, float* out, size_t* newSize, float badIn, float fillOut)
{
    clear_error();
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply_f: null cache");
    if (!require_device() || !(c = on_current_device(c, "fimexcu_cache_apply_f")))
        return FIMEXCU_ERROR;
    const float undef = std::isnan(fillOut) ? std::numeric_limits<float>::quiet_NaN() : fillOut; // store_out<FILL>(nan_f(), fill)
    const float* const ins[1] = {in};
    float* const outs[1] = {out};
    const float undefs[1] = {undef};
    return cache_apply_host<float, 1>(c, ins, size, outs, newSize, undefs, [&](const float* const* di, const Layout& L, size_t zn, float* const* dout, cudaStream_t st) {
        return cache_apply_strided(c, di[0], L, zn, dout[0], badIn, fillOut, st);
    });
}

int fimexcu_cache_apply_i16(const fimexcu_cache* c, const int16_t* in, size_t size, int16_t* out, size_t* newSize, double badValue)
{
    clear_error();
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply_i16: null cache");
    if (!require_device() || !(c = on_current_device(c, "fimexcu_cache_apply_i16")))
        return FIMEXCU_ERROR;
    const short undef = static_cast<short>(badValue); // pack_i16(NaN): ScaleValue's newFill_ (DataUtils.h:49)
    const short* const ins[1] = {in};
    short* const outs[1] = {out};
    const short undefs[1] = {undef};
    return cache_apply_host<short, 1>(c, ins, size, outs, newSize, undefs, [&](const short* const* di, const Layout& L, size_t zn, short* const* dout, cudaStream_t st) {
        return cache_apply_strided_i16(c, di[0], L, zn, dout[0], badValue, st);
    });
}

int fimexcu_cache_apply_dev_i16(const fimexcu_cache* c, const int16_t* d_in, size_t nz, int16_t* d_out, double badValue, void* stream)
{
    clear_error();
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply_dev_i16: null cache");
    if (!(c = on_current_device(c, "fimexcu_cache_apply_dev_i16")))
        return FIMEXCU_ERROR;
    return cache_apply_strided_i16(c, d_in, dense_layout(c), nz, d_out, badValue, as_stream(stream));
}

static int cache_apply_uv_strided(const fimexcu_cache* c, const fimexcu_vecrot* r, const float* d_u_in, const float* d_v_in, const Layout& L, size_t nz,
                                  float* d_u_out, float* d_v_out, float badU, float badV, float fillU, float fillV, cudaStream_t s)
{
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply_uv: null cache");
    if (!(c = on_current_device(c, "fimexcu_cache_apply_uv")) || (r && !(r = on_current_device(r, "fimexcu_cache_apply_uv (rotation matrix)"))))
        return FIMEXCU_ERROR;
    const size_t nxy = c->outX * c->outY;
    if (r && r->ox * r->oy != nxy)
        FXCU_FAIL("fimexcu_cache_apply_uv: rotation matrix and cache have different target grids");
    if (nz == 0 || nxy == 0)
        return FIMEXCU_OK;
    if (check_layout(c, L, nz) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    if (c->kind != KIND_BILINEAR) { // bicubic / nearest: one fused pass (a NaN bad / fill value switches the conversion off by itself)
        ApplyParams P;
        P.off = c->off.as<int>();
        P.wx = c->wx.p;
        P.wy = c->wy.p;
        P.tile_meta = nullptr;
        P.in = d_u_in;
        P.out = d_u_out;
        P.outX = static_cast<int>(c->outX);
        P.outY = static_cast<int>(c->outY);
        P.nz = static_cast<int>(nz);
        P.inX = static_cast<int>(c->inX);
        P.inPitch = L.inPitch;
        P.inLevel = L.inLevel;
        P.outPitch = L.outPitch;
        P.outLevel = L.outLevel;
        P.bad = badU;
        P.fill = fillU;
        UvParams Q;
        Q.cs = r ? r->c.as<double>() : nullptr;
        Q.sn = r ? r->s.as<double>() : nullptr;
        Q.vin = d_v_in;
        Q.vout = d_v_out;
        Q.badV = badV;
        Q.fillV = fillV;
        const size_t tiles = static_cast<size_t>((c->outX + 31) / 32) * ((c->outY + 7) / 8);
        P.zchunk = pick_zchunk(tiles, P.nz);
        const dim3 grid(static_cast<unsigned>((c->outX + 31) / 32), static_cast<unsigned>((c->outY + 7) / 8),
                        static_cast<unsigned>((P.nz + P.zchunk - 1) / P.zchunk));
        if (c->kind == KIND_BICUBIC) {
            if (r)
                sampled_uv_kernel<true, true><<<grid, 256, 0, s>>>(P, Q);
            else
                sampled_uv_kernel<true, false><<<grid, 256, 0, s>>>(P, Q);
        } else {
            if (r)
                sampled_uv_kernel<false, true><<<grid, 256, 0, s>>>(P, Q);
            else
                sampled_uv_kernel<false, false><<<grid, 256, 0, s>>>(P, Q);
        }
        FXCU_CUDA(cudaGetLastError());
        return FIMEXCU_OK;
    }
    ApplyParams P;
    P.off = c->off.as<int>();
    P.wx = c->wx.p;
    P.wy = c->wy.p;
    P.tile_meta = c->tile_meta.as<int4>();
    P.in = d_u_in;
    P.out = d_u_out;
    P.outX = static_cast<int>(c->outX);
    P.outY = static_cast<int>(c->outY);
    P.nz = static_cast<int>(nz);
    P.zchunk = P.nz;
    P.inX = static_cast<int>(c->inX);
    P.inPitch = L.inPitch;
    P.inLevel = L.inLevel;
    P.outPitch = L.outPitch;
    P.outLevel = L.outLevel;
    P.bad = badU;
    P.fill = fillU;
    const bool special = !std::isnan(badU) || !std::isnan(badV) || !std::isnan(fillU) || !std::isnan(fillV);
    const size_t tiles = static_cast<size_t>((c->outX + 31) / 32) * ((c->outY + 7) / 8);
    // staged (TMA) form when both sources can be described and the tiles fit
    CUtensorMap umap, vmap;
    const bool enough = tiles >= static_cast<size_t>(sm_count()) * 5;
    const bool fits = c->tiles_with_data > 0 && 10ull * c->tiles_fitting >= 8ull * c->tiles_with_data;
    if (tma_allowed() && enough && fits &&
        make_source_map(&umap, d_u_in, c->inX, c->inY, nz, L.inPitch, L.inLevel, UV_ZB) &&
        make_source_map(&vmap, d_v_in, c->inX, c->inY, nz, L.inPitch, L.inLevel, UV_ZB)) {
        UvStagedParams Q;
        Q.cs = r ? r->c.as<double>() : nullptr;
        Q.sn = r ? r->s.as<double>() : nullptr;
        Q.vin = d_v_in;
        Q.vout = d_v_out;
        Q.badV = badV;
        Q.fillV = fillV;
        const size_t slots = static_cast<size_t>(sm_count()) * 5; // see launch_cache_apply
        size_t zsplit = std::max<size_t>(1, (8 * slots + tiles - 1) / tiles);
        zsplit = std::min<size_t>(zsplit, std::max<size_t>(1, nz / (4 * UV_ZB)));
        P.zchunk = level_chunk(P.nz, zsplit, UV_ZB);
        P.l2hint = tma_l2hint();
        const dim3 grid(static_cast<unsigned>((c->outX + 31) / 32), static_cast<unsigned>((c->outY + TMA_TH - 1) / TMA_TH),
                        static_cast<unsigned>((P.nz + P.zchunk - 1) / P.zchunk));
        // a NaN bad value never compares equal and a NaN fill value re-writes NaN: the <true,true> variant is always correct
        if (r) {
            if (special)
                staged_uv_tma_kernel<true, true, true><<<grid, TMA_THREADS, 0, s>>>(umap, vmap, P, Q);
            else
                staged_uv_tma_kernel<false, false, true><<<grid, TMA_THREADS, 0, s>>>(umap, vmap, P, Q);
        } else {
            if (special)
                staged_uv_tma_kernel<true, true, false><<<grid, TMA_THREADS, 0, s>>>(umap, vmap, P, Q);
            else
                staged_uv_tma_kernel<false, false, false><<<grid, TMA_THREADS, 0, s>>>(umap, vmap, P, Q);
        }
        FXCU_CUDA(cudaGetLastError());
        return FIMEXCU_OK;
    }
    UvParams Q;
    Q.cs = r ? r->c.as<double>() : nullptr;
    Q.sn = r ? r->s.as<double>() : nullptr;
    Q.vin = d_v_in;
    Q.vout = d_v_out;
    Q.badV = badV;
    Q.fillV = fillV;
    P.zchunk = pick_zchunk(tiles, P.nz);
    const dim3 grid(static_cast<unsigned>((c->outX + 31) / 32), static_cast<unsigned>((c->outY + 7) / 8),
                    static_cast<unsigned>((P.nz + P.zchunk - 1) / P.zchunk));
    if (r) {
        if (special)
            bilinear_uv_kernel<true, true, true><<<grid, 256, 0, s>>>(P, Q);
        else
            bilinear_uv_kernel<false, false, true><<<grid, 256, 0, s>>>(P, Q);
    } else {
        if (special)
            bilinear_uv_kernel<true, true, false><<<grid, 256, 0, s>>>(P, Q);
        else
            bilinear_uv_kernel<false, false, false><<<grid, 256, 0, s>>>(P, Q);
    }
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}

int fimexcu_cache_apply_uv_dev_f(const fimexcu_cache* c, const fimexcu_vecrot* r, const float* d_u_in, const float* d_v_in, size_t nz,
                                 float* d_u_out, float* d_v_out, float badU, float badV, float fillU, float fillV, void* stream)
{
    clear_error();
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply_uv: null cache");
    return cache_apply_uv_strided(c, r, d_u_in, d_v_in, dense_layout(c), nz, d_u_out, d_v_out, badU, badV, fillU, fillV, as_stream(stream));
}

int fimexcu_cache_apply_uv_dev_strided_f(const fimexcu_cache* c, const fimexcu_vecrot* r, const float* d_u_in, const float* d_v_in,
                                         size_t in_row_pitch, size_t in_level_stride, size_t nz, float* d_u_out, float* d_v_out,
                                         size_t out_row_pitch, size_t out_level_stride, float badU, float badV, float fillU, float fillV,
                                         void* stream)
{
    clear_error();
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply_uv: null cache");
    if (in_row_pitch > 0x7fffffffull || out_row_pitch > 0x7fffffffull)
        FXCU_FAIL("fimexcu_cache_apply_uv: row pitch out of range");
    const Layout L{static_cast<int>(in_row_pitch), in_level_stride, static_cast<int>(out_row_pitch), out_level_stride};
    return cache_apply_uv_strided(c, r, d_u_in, d_v_in, L, nz, d_u_out, d_v_out, badU, badV, fillU, fillV, as_stream(stream));
}

int fimexcu_cache_apply_uv_f(const fimexcu_cache* c, const fimexcu_vecrot* r, const float* u_in, const float* v_in, size_t size, float* u_out,
                             float* v_out, size_t* newSize, float badU, float badV, float fillU, float fillV)
{
    clear_error();
    if (!c)
        FXCU_FAIL("fimexcu_cache_apply_uv_f: null cache");
    if (!require_device() || !(c = on_current_device(c, "fimexcu_cache_apply_uv_f")))
        return FIMEXCU_ERROR;
    if (r && !(r = on_current_device(r, "fimexcu_cache_apply_uv_f (rotation matrix)")))
        return FIMEXCU_ERROR;
    // the level chunks of both components go through the fused kernel together (same pipelining, bounce buffers and row bands as
    // the scalar entry point)
    const float* const ins[2] = {u_in, v_in};
    float* const outs[2] = {u_out, v_out};
    const float nanv = std::numeric_limits<float>::quiet_NaN();
    const float undefs[2] = {std::isnan(fillU) ? nanv : fillU, std::isnan(fillV) ? nanv : fillV};
    return cache_apply_host<float, 2>(c, ins, size, outs, newSize, undefs,
                                      [&](const float* const* di, const Layout& L, size_t zn, float* const* dout, cudaStream_t st) {
                                          return cache_apply_uv_strided(c, r, di[0], di[1], L, zn, dout[0], dout[1], badU, badV, fillU, fillV, st);
                                      });
}

void fimexcu_cache_destroy(fimexcu_cache* c) { delete c; }

int fimexcu_cache_set_fast_bicubic(fimexcu_cache* c, int on)
{
    clear_error();
    if (!c)
        FXCU_FAIL("fimexcu_cache_set_fast_bicubic: null cache");
    c->fast_bicubic = on != 0;
    for (fimexcu_cache* r : c->replicas)
        if (r)
            r->fast_bicubic = on != 0;
    return FIMEXCU_OK;
}

void fimexcu_set_staged_kernels(int mode) { g_tma_override.store(mode < 0 ? -1 : (mode ? 1 : 0), std::memory_order_relaxed); }

int fimexcu_cache_replicate(fimexcu_cache* c, int n_devices, const int* device_ids)
{
    clear_error();
    if (!c)
        FXCU_FAIL("fimexcu_cache_replicate: null cache");
    if (n_devices < 0 || (n_devices > 0 && !device_ids))
        FXCU_FAIL("fimexcu_cache_replicate: bad device list");
    int before = -1;
    FXCU_CUDA(cudaGetDevice(&before));
    int rc = FIMEXCU_OK;
    for (int k = 0; k < n_devices && rc == FIMEXCU_OK; ++k) {
        const int d = device_ids[k];
        if (d < 0 || d >= 64) {
            set_error("fimexcu_cache_replicate: device ordinal out of range");
            rc = FIMEXCU_ERROR;
            break;
        }
        if (d == c->device || (static_cast<size_t>(d) < c->replicas.size() && c->replicas[d]))
            continue;
        if (!cuda_ok(cudaSetDevice(d), "cudaSetDevice", __FILE__, __LINE__) || !require_device()) {
            rc = FIMEXCU_ERROR;
            break;
        }
        std::unique_ptr<fimexcu_cache> r(new fimexcu_cache());
        r->method = c->method;
        r->kind = c->kind;
        r->device = d;
        r->inX = c->inX;
        r->inY = c->inY;
        r->outX = c->outX;
        r->outY = c->outY;
        r->reduced = c->reduced;
        r->xMin = c->xMin;
        r->yMin = c->yMin;
        r->fast_bicubic = c->fast_bicubic;
        r->tiles_with_data = c->tiles_with_data;
        r->tiles_fitting = c->tiles_fitting;
        r->bands = c->bands;
        r->band_fraction = c->band_fraction;
        if (!replicate_buf(r->off, d, c->off, c->device, nullptr) || !replicate_buf(r->wx, d, c->wx, c->device, nullptr) ||
            !replicate_buf(r->wy, d, c->wy, c->device, nullptr) || !replicate_buf(r->tile_meta, d, c->tile_meta, c->device, nullptr) ||
            !cuda_ok(cudaStreamSynchronize(nullptr), "cudaStreamSynchronize", __FILE__, __LINE__)) {
            rc = FIMEXCU_ERROR;
            break;
        }
        if (c->replicas.size() <= static_cast<size_t>(d))
            c->replicas.resize(d + 1, nullptr);
        c->replicas[d] = r.release();
    }
    cudaSetDevice(before);
    return rc;
}

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
