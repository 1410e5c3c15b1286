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

#include <cfloat>
#include <cstring>
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
    int dir_hint;             // hybrid mode: +1 / -1 when every column of the launch is provably strictly monotone, 0 = check per column
    const double* xlog;       // [nzo] device: host-computed log / log-log of the fixed levels (LOG / LOGLOG), or nullptr
    const float* vcoord_out;  // [nOuter][nzo][nInner] or nullptr
    const double* vmin;       // per column or nullptr
    const double* vmax;
    float valid_min, valid_max; // NaN = none (already narrowed like CDMVerticalInterpolator.cc:553,558)
    float bad, fill;
    float* out;
};

// mifi_get_values_linear_f & friends for n == 1; returns false where the reference leaves the output untouched
// xlog: log(x) (LOG) or log(log(x + e)) (LOGLOG) computed on the host for fixed target levels, NaN = compute here
__device__ __forceinline__ float two_level(int method, float A, float B, double a, double b, double x, double xlog)
{
    if (method == FIMEXCU_VINT_METHOD_NN)
        return A; // interpolation.c:416-425
    if (method == FIMEXCU_VINT_METHOD_LOG || method == FIMEXCU_VINT_METHOD_LOGLOG) {
        if (a <= 0 || b <= 0 || x <= 0)
            return nan_f(); // reference: MIFI_ERROR, output unwritten
        if (method == FIMEXCU_VINT_METHOD_LOGLOG) {
            const double e = 2.7182818284590452354;
            a = log(__dadd_rn(a, e));
            b = log(__dadd_rn(b, e));
            if (isnan(xlog))
                x = log(__dadd_rn(x, e));
            if (a <= 0 || b <= 0 || (isnan(xlog) && x <= 0))
                return nan_f();
        }
        a = log(a);
        b = log(b);
        x = isnan(xlog) ? log(x) : xlog;
    }
    const float f = (a == b) ? 0.f : __double2float_rn(__ddiv_rn(__dsub_rn(x, a), __dsub_rn(b, a)));
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
        if (!(fabsf(__fsub_rn(x, cl)) < fabsf(__fsub_rn(x, c(n - 2)))))
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

template <class Col>
__device__ __forceinline__ void vertical_column(const VertParams& P, const Col& c, size_t cidx, size_t outer, size_t inner, int dir_known)
{
    const int nzi = static_cast<int>(P.nzi);
    // strictly monotone and NaN-free?  (+1 increasing, -1 decreasing, 0 -> general scan)
    int dir = dir_known;
    if (dir == 0 && nzi >= 2) {
        bool inc = true, dec = true;
        float prev = c(0);
        for (int k = 1; k < nzi; ++k) {
            const float cur = c(k);
            inc = inc && (cur > prev);
            dec = dec && (cur < prev);
            prev = cur;
        }
        dir = inc ? 1 : (dec ? -1 : 0);
    }
    const float* din = P.in + outer * P.nzi * P.nInner + inner;
    float* dout = P.out + outer * P.nzo * P.nInner + inner;
    const bool has_bad = !isnan(P.bad), has_fill = !isnan(P.fill);
    const double lo = P.vmin ? P.vmin[cidx] : 0., hi = P.vmax ? P.vmax[cidx] : 0.;
    for (size_t k = 0; k < P.nzo; ++k) {
        const double vout = P.vcoord_out ? static_cast<double>(__ldg(P.vcoord_out + (outer * P.nzo + k) * P.nInner + inner)) : __ldg(P.levels + k);
        bool range = true; // CDMVerticalInterpolator.cc:505-528
        if (P.vmin && P.vmax)
            range = (vout >= lo) && (vout <= hi);
        else if (P.vmin)
            range = vout >= lo;
        else if (P.vmax)
            range = vout <= hi;
        float r = nan_f();
        if (range) {
            int p0, p1;
            const float xf = __double2float_rn(vout);
            if (!(dir != 0 && find_neighbors_monotone(c, nzi, dir, xf, p0, p1)))
                find_neighbors(c, nzi, xf, p0, p1);
            if (p0 != p1) {
                float A = __ldg(din + static_cast<size_t>(p0) * P.nInner), B = __ldg(din + static_cast<size_t>(p1) * P.nInner);
                if (has_bad) {
                    if (A == P.bad)
                        A = nan_f();
                    if (B == P.bad)
                        B = nan_f();
                }
                r = two_level(P.method, A, B, static_cast<double>(c(p0)), static_cast<double>(c(p1)), vout, P.xlog ? __ldg(P.xlog + k) : nan(""));
            }
        }
        // valid_min / valid_max clamp (CDMVerticalInterpolator.cc:551-561); NaN compares false and stays
        if (r < P.valid_min)
            r = P.valid_min;
        if (r > P.valid_max)
            r = P.valid_max;
        if (has_fill && isnan(r))
            r = P.fill;
        __stcs(dout + k * P.nInner, r);
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
    if (MODE == 0) {
        float* mine = reinterpret_cast<float*>(smem_d) + threadIdx.x;
        const float* g = P.vcoord_in + outer * P.nzi * P.nInner + inner;
        for (int k = 0; k < nzi; ++k)
            mine[k * VT] = __ldg(g + static_cast<size_t>(k) * P.nInner);
        vertical_column(P, StridedCol{mine, VT}, cidx, outer, inner, 0);
    } else if (MODE == 1) {
        const size_t po = (P.nOuterPs == 1 ? 0 : outer) * P.nInner + inner;
        const double ps = __dmul_rn(static_cast<double>(__ldg(P.ps + po)), P.ps_scale);
        vertical_column(P, HybridCol{smem_d, smem_d + nzi, ps}, cidx, outer, inner, isnan(ps) ? 0 : P.dir_hint);
    } else {
        vertical_column(P, StridedCol{P.vcoord_in + outer * P.nzi * P.nInner + inner, static_cast<int>(P.nInner)}, cidx, outer, inner, 0);
    }
}

// min / max of the surface pressure of a launch (NaN ignored): lets the host prove monotonicity for all columns at once
__global__ void __launch_bounds__(256) ps_minmax_kernel(const float* __restrict__ ps, size_t n, float* __restrict__ mm)
{
    float lo = INFINITY, hi = -INFINITY;
    for (size_t i = blockIdx.x * 256ull + threadIdx.x; i < n; i += 256ull * gridDim.x) {
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

// p_k(ps) = ap_k + b_k*ps is linear in ps, so p_{k+1} - p_k keeps its sign on [ps_lo, ps_hi] if it has it at both ends; a margin
// of a few f32 ulps makes the statement hold for the ROUNDED f32 column the kernel works on.
static int hybrid_direction(const double* ap, const double* b, size_t nzi, double ps_lo, double ps_hi)
{
    if (!(ps_lo <= ps_hi) || !std::isfinite(ps_lo) || !std::isfinite(ps_hi) || nzi < 2)
        return 0;
    bool inc = true, dec = true;
    for (size_t k = 0; k + 1 < nzi; ++k) {
        for (double ps : {ps_lo, ps_hi}) {
            const double p0 = ap[k] + b[k] * ps, p1 = ap[k + 1] + b[k + 1] * ps;
            const double margin = 8.0 * 1.1920929e-7 * std::max(std::fabs(p0), std::fabs(p1)) + 1e-30;
            inc = inc && (p1 - p0 > margin);
            dec = dec && (p0 - p1 > margin);
        }
    }
    return inc ? 1 : (dec ? -1 : 0);
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
__global__ void __launch_bounds__(256) lin_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ out, size_t n, float f)
{
    for (size_t i = blockIdx.x * 256ull + threadIdx.x; i < n; i += 256ull * gridDim.x)
        out[i] = __fadd_rn(A[i], __fmul_rn(f, __fsub_rn(B[i], A[i])));
}

// kind: 0 linear, 1 weak, 2 no, 3 const.  Decides on the host exactly like the reference which of copy-A / copy-B / blend / NaN applies.
static int two_level_host(int kind, const float* A, const float* B, float* out, size_t n, double a, double b, double x)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
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
    DevBuf dA, dB, dO;
    if (!dA.alloc(bytes, nullptr) || !dB.alloc(bytes, nullptr) || !dO.alloc(bytes, nullptr))
        return FIMEXCU_ERROR;
    // the product has no CPU path: even the plain copies travel through the device
    FXCU_CUDA(cudaMemcpyAsync(dA.p, A, bytes, cudaMemcpyHostToDevice, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(dB.p, B, bytes, cudaMemcpyHostToDevice, nullptr));
    const unsigned blocks = static_cast<unsigned>(std::min<size_t>((n + 255) / 256, static_cast<size_t>(sm_count()) * 16));
    switch (action) {
    case 0: FXCU_CUDA(cudaMemcpyAsync(dO.p, dA.p, bytes, cudaMemcpyDeviceToDevice, nullptr)); break;
    case 1: FXCU_CUDA(cudaMemcpyAsync(dO.p, dB.p, bytes, cudaMemcpyDeviceToDevice, nullptr)); break;
    case 2: lin_kernel<<<blocks, 256>>>(dA.as<float>(), dB.as<float>(), dO.as<float>(), n, f); break;
    default: FXCU_CUDA(cudaMemsetAsync(dO.p, 0xff, bytes, nullptr)); break; // all-ones is a quiet NaN
    }
    FXCU_CUDA(cudaGetLastError());
    FXCU_CUDA(cudaMemcpyAsync(out, dO.p, bytes, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

} // namespace fxcu

using namespace fxcu;

extern "C" {

int fimexcu_vertical_interpolate_dev_f(int method, const float* d_in, const float* d_vcoord_in, size_t nOuter, size_t nzi, size_t nInner,
                                       const double* levels_host, const float* d_vcoord_out, size_t nzo, const double* d_vmin,
                                       const double* d_vmax, double valid_min, double valid_max, float badIn, float fillOut, float* d_out,
                                       void* stream)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
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
        if (method == FIMEXCU_VINT_METHOD_LOG || method == FIMEXCU_VINT_METHOD_LOGLOG)
            P.xlog = lv.as<double>() + nzo;
        P.vcoord_out = nullptr;
    } else {
        P.vcoord_out = d_vcoord_out;
    }
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
    if (!d_vcoord_in)
        FXCU_FAIL("fimexcu_vertical_interpolate: null input vertical coordinate");
    const int rc = launch_vertical(P, s);
    if (levels_host)
        FXCU_CUDA(cudaStreamSynchronize(s)); // levels_host may be a temporary of the caller
    return rc;
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

int fimexcu_vertical_hybrid_pressure_dev_f(int method, const float* d_in, size_t nOuter, size_t nzi, size_t nInner, const double* ap_host,
                                           const double* b_host, const float* d_ps, size_t nOuterPs, double ps_scale, const double* levels_host,
                                           size_t nzo, double valid_min, double valid_max, float badIn, float fillOut, float* d_out,
                                           void* stream)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    if (!(nOuterPs == 1 || nOuterPs == nOuter))
        FXCU_FAIL("fimexcu_vertical_hybrid_pressure: ps must have 1 or nOuter outer entries");
    if (!levels_host || !ap_host || !b_host || !d_ps)
        FXCU_FAIL("fimexcu_vertical_hybrid_pressure: null argument");
    cudaStream_t s = as_stream(stream);
    DevBuf coef;
    const size_t nc = 2 * nzi + 2 * nzo;
    if (!coef.alloc(std::max<size_t>(nc, 1) * sizeof(double), s))
        return FIMEXCU_ERROR;
    std::vector<double> h(nc);
    std::copy(ap_host, ap_host + nzi, h.begin());
    std::copy(b_host, b_host + nzi, h.begin() + nzi);
    std::copy(levels_host, levels_host + nzo, h.begin() + 2 * nzi);
    for (size_t k = 0; k < nzo; ++k)
        h[2 * nzi + nzo + k] = host_level_log(method, levels_host[k]);
    FXCU_CUDA(cudaMemcpyAsync(coef.p, h.data(), nc * sizeof(double), cudaMemcpyHostToDevice, s));
    VertParams P{};
    P.method = method;
    P.in = d_in;
    P.vcoord_in = nullptr;
    P.ap = coef.as<double>();
    P.b = coef.as<double>() + nzi;
    P.levels = coef.as<double>() + 2 * nzi;
    if (method == FIMEXCU_VINT_METHOD_LOG || method == FIMEXCU_VINT_METHOD_LOGLOG)
        P.xlog = coef.as<double>() + 2 * nzi + nzo;
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
    {
        DevBuf mm;
        if (!mm.alloc(2 * sizeof(float), s))
            return FIMEXCU_ERROR;
        const int init[2] = {0x7f800000, static_cast<int>(0x80800000u)}; // +inf / "-inf" in the ordered-int encoding
        FXCU_CUDA(cudaMemcpyAsync(mm.p, init, sizeof(init), cudaMemcpyHostToDevice, s));
        const size_t nps = nOuterPs * nInner;
        if (nps)
            ps_minmax_kernel<<<static_cast<unsigned>(std::min<size_t>((nps + 255) / 256, 2048)), 256, 0, s>>>(d_ps, nps, mm.as<float>());
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
    (void)a; (void)b; (void)x;
    return two_level_host(3, infieldA, infieldB, outfield, n, 0., 1., -1.); // f < 0 with the "const" rule: copy A
}

int fimexcu_get_values_linear_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x)
{
    return two_level_host(0, infieldA, infieldB, outfield, n, a, b, x);
}

int fimexcu_get_values_linear_weak_extrapol_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x)
{
    return two_level_host(1, infieldA, infieldB, outfield, n, a, b, x);
}

int fimexcu_get_values_linear_no_extrapol_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x)
{
    return two_level_host(2, infieldA, infieldB, outfield, n, a, b, x);
}

int fimexcu_get_values_linear_const_extrapol_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x)
{
    return two_level_host(3, infieldA, infieldB, outfield, n, a, b, x);
}

int fimexcu_get_values_log_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x)
{
    if (a <= 0 || b <= 0 || x <= 0) {
        set_error("mifi_get_values_log_f: non-positive coordinate");
        return FIMEXCU_ERROR; // output untouched, like interpolation.c:524-526
    }
    // the three logarithms are host-side scalars (as in the reference); the field arithmetic runs on the device
    return two_level_host(0, infieldA, infieldB, outfield, n, std::log(a), std::log(b), std::log(x));
}

int fimexcu_get_values_log_log_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x)
{
    if (a <= 0 || b <= 0 || x <= 0) {
        set_error("mifi_get_values_log_log_f: non-positive coordinate");
        return FIMEXCU_ERROR;
    }
    const double e = 2.7182818284590452354;
    return fimexcu_get_values_log_f(infieldA, infieldB, outfield, n, std::log(a + e), std::log(b + e), std::log(x + e));
}

} // extern "C"
