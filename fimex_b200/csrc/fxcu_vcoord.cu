// fxcu_vcoord.cu -- vertical coordinate converters on the device: the fields that feed K15 as the
// input / output vertical coordinate when the target is pressure, height, altitude or depth.
//
// Reference behaviour restated here (fimex 2.4.1):
//   mifi_atmosphere_{ln_pressure,sigma_pressure,hybrid_sigma_pressure,hybrid_sigma_ap_pressure,hybrid_height}
//                                                        src/vertical_coordinate_transformations.c:29-78
//   mifi_barometric_{pressure,height,standard_*}          :86-113
//   mifi_virtual_temperature, mifi_barometric_layer_thickness   :115-118, 161-164
//   mifi_ocean_s_g1_z, mifi_ocean_s_g2_z                   :166-183
//   converter classes that call them                      src/coordSys/verticalTransform/{HybridSigmaApToPressure,HybridSigmaToPressure,
//       SigmaToPressure,LnPressureToPressure,OceanSCoordinateGToDepth,PressureToStandardAltitude,AltitudeStandardToPressure,
//       AltitudeHeight,PressureIntegrationToAltitude}Converter.cc
//
// The reference's converters hand DOUBLE arrays to one another and the interpolator narrows the last one with Data::asFloat()
// (CDMVerticalInterpolator.cc:483); here a whole chain (level formula -> standard-atmosphere step -> +- topography) is one
// kernel in f64 registers with one f32 store.  Every f64 operation is a separately rounded _rn intrinsic like the reference's
// x86-64 code; exp / log are the device's (last-bit differences to glibc show, rarely, as 1 ulp of the f32 result).
// HBM layout: fields [nOuter][nz][nInner] f32, surface fields [nOuterS][nInner] (nOuterS = 1 or nOuter), consecutive threads
// own consecutive inner indices.  Algorithmic bytes per column: level formulas 4 (surface) + 4*nz (out); hypsometric
// integration 8 (sap, sgp) + 4*nz*(p, T[, q]) + 4*nz (out).
#include "fxcu_common.h"
#include "fxcu_logtable.h"

#include <cmath>
#include <vector>

namespace fxcu {

constexpr double GAS_CONSTANT = 8.31432, EARTH_GRAVITY = 9.80665, MOLAR_MASS_DRY_AIR = 0.0289644; // mifi_constants.h:235-241
constexpr double BAROMETRIC_FACTOR = GAS_CONSTANT / (EARTH_GRAVITY * MOLAR_MASS_DRY_AIR);           // v_c_t.c:80-81
constexpr double Z_MOL_WEIGHT_RATIO = .60771704180064308681;                                        // :84

struct VcoordParams {
    int kind, post;
    const double* a; // [nz] device
    const double* b; // [nz] device
    double scalar;
    const float* surf;
    size_t nOuterS;
    double surf_scale;
    const float* eta;
    const float* topo;
    double topo_factor;
    size_t nOuter, nz, nInner;
    float* out;
};

// the standard-atmosphere step narrows its result to f32 (contract: 2 f32 ulps; the device's own log already differs from glibc's in
// the last f64 bit), so its logarithm is the table form of fxcu_logtable.h
__device__ __forceinline__ double standard_step(int post, double v, const double2* logtab = nullptr)
{
    if (post == FIMEXCU_VPOST_STANDARD_ALTITUDE) { // mifi_barometric_height(n, 1013.25, p, 288.15, .)
        const double K = -BAROMETRIC_FACTOR * 288.15;
        if (logtab) // log(v / 1013.25) = log v - log 1013.25
            return __dmul_rn(K, log_table(logtab, v) - 6.920918265508418);
        return __dmul_rn(K, log(__ddiv_rn(v, 1013.25)));
    }
    if (post == FIMEXCU_VPOST_STANDARD_PRESSURE) { // mifi_barometric_pressure(n, 1013.25, h, 288.15, .)
        const double Cc = -1 / (BAROMETRIC_FACTOR * 288.15);
        return __dmul_rn(1013.25, exp(__dmul_rn(Cc, v)));
    }
    return v;
}

// one thread per column: the surface terms are read once, the level loop streams the output.  The converter and the
// standard-atmosphere step are compile-time constants: with both decided per level at run time the loop was issue-bound (91 % of
// the issue slots, ncu) at a quarter of the memory roofline.
template <int KIND, int POST>
__global__ void __launch_bounds__(256) vcoord_levels_kernel(const VcoordParams P)
{
    extern __shared__ double s_ab[]; // a[nz], b[nz]
    __shared__ double2 s_log[128];
    const int nz = static_cast<int>(P.nz);
    for (int k = threadIdx.x; k < nz; k += 256) {
        s_ab[k] = P.a[k];
        s_ab[nz + k] = P.b ? P.b[k] : 0.;
    }
    if (POST == FIMEXCU_VPOST_STANDARD_ALTITUDE)
        log_table_fill(s_log, threadIdx.x, 256);
    __syncthreads();
    const size_t ncol = P.nOuter * P.nInner;
    const size_t cidx = blockIdx.x * 256ull + threadIdx.x;
    if (cidx >= ncol)
        return;
    const size_t outer = cidx / P.nInner, inner = cidx - outer * P.nInner;
    const size_t si = (P.nOuterS == 1 ? 0 : outer) * P.nInner + inner;
    const double s = P.surf ? __dmul_rn(P.surf_scale, static_cast<double>(__ldg(P.surf + si))) : 0.;
    const double zeta = P.eta ? static_cast<double>(__ldg(P.eta + si)) : 0.;
    const double tp = P.topo ? __dmul_rn(P.topo_factor, static_cast<double>(__ldg(P.topo + inner))) : 0.;
    const double sc = P.scalar;
    // per-column invariants of the formulas
    const double pDiff = __dsub_rn(s, sc);                    // sigma
    const double h_inv = __ddiv_rn(1., s);                    // ocean g1
    const double hph_c_inv = __ddiv_rn(1., __dadd_rn(s, sc)); // ocean g2
    float* po = P.out + outer * P.nz * P.nInner + inner;
    for (int k = 0; k < nz; ++k, po += P.nInner) {
        const double ak = s_ab[k], bk = s_ab[nz + k];
        double v;
        if (KIND == FIMEXCU_VCOORD_HYBRID_SIGMA_AP || KIND == FIMEXCU_VCOORD_HYBRID_HEIGHT) {
            v = __dadd_rn(ak, __dmul_rn(bk, s));
        } else if (KIND == FIMEXCU_VCOORD_HYBRID_SIGMA) {
            v = __dadd_rn(__dmul_rn(ak, sc), __dmul_rn(bk, s));
        } else if (KIND == FIMEXCU_VCOORD_SIGMA) {
            v = __dadd_rn(sc, __dmul_rn(ak, pDiff));
        } else if (KIND == FIMEXCU_VCOORD_LN_PRESSURE) {
            v = __dmul_rn(sc, exp(-ak));
        } else if (KIND == FIMEXCU_VCOORD_OCEAN_S_G1) {
            const double S = __dadd_rn(__dmul_rn(sc, ak), __dmul_rn(__dsub_rn(s, sc), bk));
            v = -__dadd_rn(S, __dmul_rn(zeta, __dadd_rn(1., __dmul_rn(S, h_inv))));
        } else {
            const double S = __dmul_rn(hph_c_inv, __dadd_rn(__dmul_rn(sc, ak), __dmul_rn(s, bk)));
            v = -__dadd_rn(zeta, __dmul_rn(__dadd_rn(zeta, s), S));
        }
        v = standard_step(POST, v, s_log);
        if (P.topo)
            v = __dadd_rn(v, tp);
        __stcs(po, __double2float_rn(v));
    }
}

__global__ void __launch_bounds__(256) vcoord_transform_kernel(int post, const float* __restrict__ in, size_t n, float* __restrict__ out)
{
    for (size_t i = blockIdx.x * 256ull + threadIdx.x; i < n; i += 256ull * gridDim.x)
        out[i] = __double2float_rn(standard_step(post, static_cast<double>(in[i])));
}

__constant__ double c_atanh_odd[12] = {1., 1. / 3, 1. / 5, 1. / 7, 1. / 9, 1. / 11, 1. / 13, 1. / 15, 1. / 17, 1. / 19, 1. / 21, 1. / 23};

// mifi_barometric_layer_thickness (v_c_t.c:161-164): (float)(log((double)(p_low / p_high)) * T * BAROMETRIC_FACTOR).  For layer
// ratios within a factor 1.5 the logarithm is evaluated as 2 atanh((r-1)/(r+1)) (series exact to < 2e-17 relative for |s| < 0.2,
// a third of the instructions of log()); the result is taken only if every real number within 1e-15 relative -- more than the
// error of either evaluation -- narrows to the same f32, otherwise (and for thicker layers) log() is called.
__device__ __forceinline__ float layer_thickness(float p_low, float p_high, float Tv)
{
    const double r = static_cast<double>(__fdiv_rn(p_low, p_high));
    const double T = static_cast<double>(Tv);
    if (r > 0.667 && r < 1.5) {
        const double den = __dadd_rn(r, 1.);
        double q;
        asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(q) : "d"(den));
        q = __fma_rn(q, __fma_rn(-den, q, 1.0), q); // 2^-23 -> 2^-46; the quotient below gets its own correction step
        const double num = __dsub_rn(r, 1.); // exact (Sterbenz)
        double sv = num * q;
        sv = __fma_rn(__fma_rn(-den, sv, num), q, sv); // one correction: s to ~1 ulp
        const double z = sv * sv;
        // terms z^n / (2n + 1): eight of them leave < 1e-19 relative for |s| < 0.07 (layer ratios within 0.87 .. 1.15, the usual
        // spacing of model levels), twelve < 2e-17 for |s| < 0.2
        double pl;
        if (z < 0.0049) {
            pl = c_atanh_odd[7];
#pragma unroll
            for (int i = 6; i >= 0; --i)
                pl = __fma_rn(pl, z, c_atanh_odd[i]);
        } else {
            pl = c_atanh_odd[11];
#pragma unroll
            for (int i = 10; i >= 0; --i)
                pl = __fma_rn(pl, z, c_atanh_odd[i]);
        }
        const double lg = 2. * sv * pl;
        const double m = __dmul_rn(__dmul_rn(lg, T), BAROMETRIC_FACTOR);
        const float lo = __double2float_rn(m * (1. - 1e-15)), hi = __double2float_rn(m * (1. + 1e-15));
        if (lo == hi)
            return lo;
    }
    return __double2float_rn(__dmul_rn(__dmul_rn(log(r), T), BAROMETRIC_FACTOR));
}

// first column (in the reference's loop order) whose first and last pressure are positive decides the direction of ALL
// columns (PressureIntegrationToAltitudeConverter.cc:141-164): dirflag[0] = its index (atomicMin), resolved by the consumer
__global__ void __launch_bounds__(256) vcoord_direction_kernel(const float* __restrict__ p, size_t nOuter, size_t nz, size_t nInner,
                                                               unsigned long long* __restrict__ first_col)
{
    const size_t ncol = nOuter * nInner;
    unsigned long long best = ~0ull;
    for (size_t c = blockIdx.x * 256ull + threadIdx.x; c < ncol; c += 256ull * gridDim.x) {
        const size_t outer = c / nInner, inner = c - outer * nInner;
        const float pf = p[(outer * nz) * nInner + inner], pl = p[(outer * nz + nz - 1) * nInner + inner];
        if (pf > 0 && pl > 0) {
            best = c;
            break; // later columns of this thread have larger indices
        }
    }
    if (best != ~0ull)
        atomicMin(first_col, best);
}

struct IntegrationParams {
    const float* p;
    const float* sap;
    double sap_scale;
    const float* sgp;
    double sgp_scale;
    const float* airt;
    const float* shum;
    size_t nOuter, nOuterS, nz, nInner;
    int start_high_p; // 1 / 0, or -1: from first_col
    const unsigned long long* first_col;
    const float* topo;
    double topo_factor;
    float* out;
    int* status; // set to 1 when the direction cannot be determined
};

// One thread per column walks from the surface upwards (a serial prefix sum in f64, as the reference's loop); loads of a
// level are coalesced across the warp's columns and issued PF levels ahead of their use.
constexpr int PF = 4;
__global__ void __launch_bounds__(128) vcoord_integration_kernel(const IntegrationParams P)
{
    const size_t ncol = P.nOuter * P.nInner;
    const size_t cidx = blockIdx.x * 128ull + threadIdx.x;
    if (cidx >= ncol)
        return;
    int high = P.start_high_p;
    if (high < 0) {
        const unsigned long long c = *P.first_col;
        if (c == ~0ull) { // "unable to determine pressure axis direction ... giving up"
            if (cidx == 0)
                *P.status = 1;
            return;
        }
        const size_t o = c / P.nInner, i = c - o * P.nInner;
        high = P.p[(o * P.nz) * P.nInner + i] > P.p[(o * P.nz + P.nz - 1) * P.nInner + i];
    }
    const size_t outer = cidx / P.nInner, inner = cidx - outer * P.nInner;
    const size_t si = (P.nOuterS == 1 ? 0 : outer) * P.nInner + inner;
    const float sgp = __double2float_rn(__dmul_rn(P.sgp_scale, static_cast<double>(__ldg(P.sgp + si))));
    float p_low = __double2float_rn(__dmul_rn(P.sap_scale, static_cast<double>(__ldg(P.sap + si))));
    double a = __ddiv_rn(static_cast<double>(sgp), EARTH_GRAVITY);
    const double tp = P.topo ? __dmul_rn(P.topo_factor, static_cast<double>(__ldg(P.topo + inner))) : 0.;
    const int nz = static_cast<int>(P.nz);
    const long long step = high ? static_cast<long long>(P.nInner) : -static_cast<long long>(P.nInner);
    const size_t base = (outer * P.nz + (high ? 0 : P.nz - 1)) * P.nInner + inner;
    const float* pp = P.p + base;
    const float* pt = P.airt + base;
    const float* pq = P.shum ? P.shum + base : nullptr;
    float* po = P.out + base;
    float bp[PF], bt[PF], bq[PF];
#pragma unroll
    for (int i = 0; i < PF; ++i) {
        const bool in = i < nz;
        bp[i] = in ? __ldg(pp + i * step) : 1.f;
        bt[i] = in ? __ldg(pt + i * step) : 0.f;
        bq[i] = (in && pq) ? __ldg(pq + i * step) : 0.f;
    }
    // the read pointers run PF levels ahead of the level being integrated; all are bumped by one block of PF levels per round
    const long long block_step = PF * step;
    pp += block_step;
    pt += block_step;
    if (pq)
        pq += block_step;
    for (int n0 = 0; n0 < nz; n0 += PF, pp += block_step, pt += block_step, po += block_step) {
        float np[PF], nt[PF], nq[PF];
#pragma unroll
        for (int i = 0; i < PF; ++i) {
            const bool in = n0 + PF + i < nz;
            np[i] = in ? __ldg(pp + i * step) : 1.f;
            nt[i] = in ? __ldg(pt + i * step) : 0.f;
            nq[i] = (in && pq) ? __ldg(pq + i * step) : 0.f;
        }
        if (pq)
            pq += block_step;
#pragma unroll
        for (int i = 0; i < PF; ++i) {
            const int n = n0 + i;
            if (n < nz) {
                const float p_high = bp[i];
                float Tv = bt[i];
                if (pq) // mifi_virtual_temperature: (1 + Z*q) * T in double, returned as float
                    Tv = __double2float_rn(__dmul_rn(__dadd_rn(1., __dmul_rn(Z_MOL_WEIGHT_RATIO, static_cast<double>(bq[i]))), static_cast<double>(Tv)));
                const float lt = layer_thickness(p_low, p_high, Tv);
                a = __dadd_rn(a, static_cast<double>(lt));
                __stcs(po + i * step, __double2float_rn(P.topo ? __dadd_rn(a, tp) : a));
                p_low = p_high;
            }
        }
#pragma unroll
        for (int i = 0; i < PF; ++i) {
            bp[i] = np[i];
            bt[i] = nt[i];
            bq[i] = nq[i];
        }
    }
}

} // namespace fxcu

using namespace fxcu;

extern "C" {

int fimexcu_vcoord_levels_dev_f(int kind, const double* a_host, const double* b_host, double scalar, const float* d_surf, size_t nOuterS,
                                double surf_scale, const float* d_eta, int post, const float* d_topo, double topo_factor, size_t nOuter, size_t nz,
                                size_t nInner, float* d_out, void* stream)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    if (kind < 0 || kind > FIMEXCU_VCOORD_OCEAN_S_G2)
        FXCU_FAIL("fimexcu_vcoord_levels: unknown converter kind");
    if (post < 0 || post > FIMEXCU_VPOST_STANDARD_PRESSURE)
        FXCU_FAIL("fimexcu_vcoord_levels: unknown standard-atmosphere step");
    if (!a_host || !d_out)
        FXCU_FAIL("fimexcu_vcoord_levels: null argument");
    const bool needs_b = kind != FIMEXCU_VCOORD_SIGMA && kind != FIMEXCU_VCOORD_LN_PRESSURE;
    if (needs_b && !b_host)
        FXCU_FAIL("fimexcu_vcoord_levels: this converter needs both per-level terms");
    if (kind != FIMEXCU_VCOORD_LN_PRESSURE && !d_surf)
        FXCU_FAIL("fimexcu_vcoord_levels: this converter needs a surface field");
    if (!(nOuterS == 1 || nOuterS == nOuter))
        FXCU_FAIL("fimexcu_vcoord_levels: surface fields must have 1 or nOuter outer entries");
    const size_t ncol = nOuter * nInner;
    if (ncol == 0 || nz == 0)
        return FIMEXCU_OK;
    if (nz > 3000)
        FXCU_FAIL("fimexcu_vcoord_levels: more than 3000 levels are not supported");
    cudaStream_t s = as_stream(stream);
    DevBuf coef;
    if (!coef.alloc(2 * nz * sizeof(double), s))
        return FIMEXCU_ERROR;
    std::vector<double> h(2 * nz, 0.);
    std::copy(a_host, a_host + nz, h.begin());
    if (b_host)
        std::copy(b_host, b_host + nz, h.begin() + nz);
    FXCU_CUDA(cudaMemcpyAsync(coef.p, h.data(), 2 * nz * sizeof(double), cudaMemcpyHostToDevice, s));
    VcoordParams P{};
    P.kind = kind;
    P.post = post;
    P.a = coef.as<double>();
    P.b = coef.as<double>() + nz;
    P.scalar = scalar;
    P.surf = kind == FIMEXCU_VCOORD_LN_PRESSURE ? nullptr : d_surf;
    P.nOuterS = nOuterS;
    P.surf_scale = surf_scale;
    P.eta = (kind == FIMEXCU_VCOORD_OCEAN_S_G1 || kind == FIMEXCU_VCOORD_OCEAN_S_G2) ? d_eta : nullptr;
    P.topo = d_topo;
    P.topo_factor = topo_factor;
    P.nOuter = nOuter;
    P.nz = nz;
    P.nInner = nInner;
    P.out = d_out;
    const unsigned nblocks = blocks_for(ncol, 256);
    const size_t smem = 2 * nz * sizeof(double);
#define FXCU_VC_POST(K) \
    switch (post) { \
    case FIMEXCU_VPOST_NONE: vcoord_levels_kernel<K, FIMEXCU_VPOST_NONE><<<nblocks, 256, smem, s>>>(P); break; \
    case FIMEXCU_VPOST_STANDARD_ALTITUDE: vcoord_levels_kernel<K, FIMEXCU_VPOST_STANDARD_ALTITUDE><<<nblocks, 256, smem, s>>>(P); break; \
    default: vcoord_levels_kernel<K, FIMEXCU_VPOST_STANDARD_PRESSURE><<<nblocks, 256, smem, s>>>(P); break; \
    }
    switch (kind) {
    case FIMEXCU_VCOORD_HYBRID_SIGMA_AP: FXCU_VC_POST(FIMEXCU_VCOORD_HYBRID_SIGMA_AP) break;
    case FIMEXCU_VCOORD_HYBRID_SIGMA: FXCU_VC_POST(FIMEXCU_VCOORD_HYBRID_SIGMA) break;
    case FIMEXCU_VCOORD_SIGMA: FXCU_VC_POST(FIMEXCU_VCOORD_SIGMA) break;
    case FIMEXCU_VCOORD_LN_PRESSURE: FXCU_VC_POST(FIMEXCU_VCOORD_LN_PRESSURE) break;
    case FIMEXCU_VCOORD_HYBRID_HEIGHT: FXCU_VC_POST(FIMEXCU_VCOORD_HYBRID_HEIGHT) break;
    case FIMEXCU_VCOORD_OCEAN_S_G1: FXCU_VC_POST(FIMEXCU_VCOORD_OCEAN_S_G1) break;
    default: FXCU_VC_POST(FIMEXCU_VCOORD_OCEAN_S_G2) break;
    }
#undef FXCU_VC_POST
    FXCU_CUDA(cudaGetLastError());
    FXCU_CUDA(cudaStreamSynchronize(s)); // `h` is a temporary
    return FIMEXCU_OK;
}

int fimexcu_vcoord_transform_dev_f(int post, const float* d_in, size_t n, float* d_out, void* stream)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    if (post < 0 || post > FIMEXCU_VPOST_STANDARD_PRESSURE)
        FXCU_FAIL("fimexcu_vcoord_transform: unknown standard-atmosphere step");
    if (n == 0)
        return FIMEXCU_OK;
    const unsigned blocks = static_cast<unsigned>(std::min<size_t>((n + 255) / 256, static_cast<size_t>(sm_count()) * 16));
    vcoord_transform_kernel<<<blocks, 256, 0, as_stream(stream)>>>(post, d_in, n, d_out);
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}

int fimexcu_vcoord_pressure_integration_dev_f(const float* d_p, const float* d_sap, double sap_scale, const float* d_sgp, double sgp_scale,
                                              const float* d_airt, const float* d_shum, size_t nOuter, size_t nOuterS, size_t nz, size_t nInner,
                                              int start_high_p, const float* d_topo, double topo_factor, float* d_out, void* stream)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    if (!d_p || !d_sap || !d_sgp || !d_airt || !d_out)
        FXCU_FAIL("fimexcu_vcoord_pressure_integration: null argument");
    if (!(nOuterS == 1 || nOuterS == nOuter))
        FXCU_FAIL("fimexcu_vcoord_pressure_integration: surface fields must have 1 or nOuter outer entries");
    const size_t ncol = nOuter * nInner;
    if (ncol == 0 || nz == 0)
        return FIMEXCU_OK;
    if (nz > 0x3fffffffull)
        FXCU_FAIL("fimexcu_vcoord_pressure_integration: too many levels");
    cudaStream_t s = as_stream(stream);
    DevBuf scratch; // [0] first column with positive ends, [1] status
    if (!scratch.alloc(16, s))
        return FIMEXCU_ERROR;
    const unsigned long long init[2] = {~0ull, 0ull};
    FXCU_CUDA(cudaMemcpyAsync(scratch.p, init, sizeof(init), cudaMemcpyHostToDevice, s));
    if (start_high_p < 0)
        vcoord_direction_kernel<<<static_cast<unsigned>(std::min<size_t>((ncol + 255) / 256, 1024)), 256, 0, s>>>(d_p, nOuter, nz, nInner,
                                                                                                                  scratch.as<unsigned long long>());
    IntegrationParams P{};
    P.p = d_p;
    P.sap = d_sap;
    P.sap_scale = sap_scale;
    P.sgp = d_sgp;
    P.sgp_scale = sgp_scale;
    P.airt = d_airt;
    P.shum = d_shum;
    P.nOuter = nOuter;
    P.nOuterS = nOuterS;
    P.nz = nz;
    P.nInner = nInner;
    P.start_high_p = start_high_p < 0 ? -1 : (start_high_p ? 1 : 0);
    P.first_col = scratch.as<unsigned long long>();
    P.topo = d_topo;
    P.topo_factor = topo_factor;
    P.out = d_out;
    P.status = reinterpret_cast<int*>(scratch.as<unsigned long long>() + 1);
    vcoord_integration_kernel<<<blocks_for(ncol, 128), 128, 0, s>>>(P);
    FXCU_CUDA(cudaGetLastError());
    int status = 0;
    FXCU_CUDA(cudaMemcpyAsync(&status, P.status, sizeof(int), cudaMemcpyDeviceToHost, s));
    FXCU_CUDA(cudaStreamSynchronize(s)); // `init` is a temporary; the status decides the return value
    if (status)
        FXCU_FAIL("fimexcu_vcoord_pressure_integration: unable to determine the pressure axis direction (no column with positive ends)");
    return FIMEXCU_OK;
}

} // extern "C"
