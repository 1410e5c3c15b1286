// fxcu_fill.cu -- mifi_fill2d_f on the device: the `fill2d` pre/post-process of
// CDMInterpolator::getDataSlice (src/CDMInterpolator.cc:76-80, 163-186), bit-exact.
//
// Reference behaviour restated here (fimex 2.4.1, src/interpolation.c:548-678): undefined cells are set to the field mean and
// relaxed towards the solution of Laplace's equation by SEQUENTIAL Gauss-Seidel sweeps in row-major order (every cell sees the
// already updated west and north neighbours of the same sweep), over-relaxed by corrEff; after every sweep the four borders are
// extrapolated from their inner neighbours; after sweeps 0, 10, 20, ... (below maxLoop-5) the largest correction is compared
// with relaxCrit * mean absolute deviation and the routine returns when all are below.
//
// Parallel form with the SAME dependency graph.  Cell (x,y) of sweep s needs (x-1,y), (x,y-1) of sweep s and (x+1,y), (x,y+1)
// of sweep s-1: with the wavefront index t = x + y + 2s every dependency has a smaller t, so all cells of equal t -- one
// anti-diagonal of each sweep in flight -- are independent (and, by parity, no cell written at step t is read at step t).  The
// border cells fit the same schedule one step behind their inner neighbour.  A group of up to ten sweeps (up to the next
// convergence test) is therefore ONE pass of nx+ny+20 steps instead of ten passes of nx+ny steps, each step a barrier of the
// CTA that owns the level.  One CTA per level (levels are independent, processArray_ runs them in an OpenMP loop).  The sweeps
// work on a diagonal-major copy of the level (cell (x,y) stored at [x+y][x]): an anti-diagonal and its four neighbour
// diagonals are then contiguous rows, every warp access one line (in the row-major layout every lane hits its own sector and
// the step is bound by the L1 tag stage: 13 us per step against 5.7 us); the ~150 KB a step touches stay in L1 / L2.
// The two double-precision sums of the set-up (mean, mean absolute deviation) are added in the reference's sequential order by
// one thread from shared-memory stages filled by the others: their roundings decide the result, so the order cannot be changed.
#include "fxcu_common.h"

#include <vector>

namespace fxcu {

constexpr int FT = 512;      // threads per CTA (one CTA per level)
constexpr int GMAX = 10;     // sweeps in flight: the convergence test comes every 10 sweeps
constexpr int FCHUNK = 2048; // doubles per stage of the sequential sums

struct Fill2dParams {
    float* field;  // [nz][ny][nx], in place
    float* sf;     // [nz][nx+ny-1][nx] scratch: the level in DIAGONAL-MAJOR order, cell (x,y) at [(x+y)][x]
    float* sw;     // same layout: the variational weights
    int nx, ny;
    size_t level;  // elements per level
    size_t slevel; // elements per skewed level
    float relaxCrit, corrEff;
    unsigned long long maxLoop;
    unsigned long long* nChanged; // [nz]
};

// sum of term(i) over the level in index order, every addition rounded like the reference's loop
template <class Term>
__device__ __forceinline__ double sequential_sum(size_t total, double* stage, Term term)
{
    double sum = 0.;
    for (size_t base = 0; base < total; base += FCHUNK) {
        const int m = static_cast<int>(min(static_cast<size_t>(FCHUNK), total - base));
        for (int i = threadIdx.x; i < m; i += FT)
            stage[i] = term(base + i); // +0.0 where the reference adds nothing: s + 0.0 == s
        __syncthreads();
        if (threadIdx.x == 0) {
#pragma unroll 8
            for (int i = 0; i < m; ++i)
                sum = __dadd_rn(sum, stage[i]);
        }
        __syncthreads();
    }
    return sum; // valid in thread 0
}

// one border cell: field[i] += (field[j] - field[i]) * w[i]  (:660-667)
__device__ __forceinline__ void border_cell(float* f, const float* w, size_t i, size_t j)
{
    f[i] = __fadd_rn(f[i], __fmul_rn(__fsub_rn(f[j], f[i]), w[i]));
}

__global__ void __launch_bounds__(FT) fill2d_kernel(const Fill2dParams P)
{
    __shared__ double stage[FCHUNK];
    __shared__ double s_avg, s_crit;
    const int nx = P.nx, ny = P.ny, tid = threadIdx.x;
    const size_t total = static_cast<size_t>(nx) * ny;
    float* f = P.field + blockIdx.x * P.level;
    float* sf = P.sf + blockIdx.x * P.slevel;
    float* sw = P.sw + blockIdx.x * P.slevel;
    auto sk = [nx](int x, int y) { return static_cast<size_t>(x + y) * nx + x; }; // diagonal-major index

    // ---- set-up (:556-612) ----
    unsigned long long mine = 0;
    for (size_t i = tid; i < total; i += FT)
        mine += isnan(f[i]) ? 1u : 0u;
    __shared__ unsigned long long s_nchg;
    if (tid == 0)
        s_nchg = 0;
    __syncthreads();
    if (mine)
        atomicAdd(&s_nchg, mine);
    __syncthreads();
    const unsigned long long nChanged = s_nchg, nUnchanged = total - nChanged;
    if (tid == 0)
        P.nChanged[blockIdx.x] = nChanged;
    if (nUnchanged == 0 || nChanged == 0)
        return; // nothing to do (:565-567)
    const double sum = sequential_sum(total, stage, [&](size_t i) {
        const float v = f[i];
        return isnan(v) ? 0. : static_cast<double>(v);
    });
    if (tid == 0)
        s_avg = __ddiv_rn(sum, static_cast<double>(nUnchanged));
    __syncthreads();
    const double average = s_avg;
    const double dev = sequential_sum(total, stage, [&](size_t i) {
        const float v = f[i];
        return isnan(v) ? 0. : fabs(__dsub_rn(static_cast<double>(v), average));
    });
    if (tid == 0)
        s_crit = __dmul_rn(static_cast<double>(P.relaxCrit), __ddiv_rn(dev, static_cast<double>(nUnchanged)));
    const float avgf = __double2float_rn(average);
    for (size_t i = tid; i < total; i += FT) {
        const int y = static_cast<int>(i / nx), x = static_cast<int>(i - static_cast<size_t>(y) * nx);
        const bool interior = x >= 1 && x < nx - 1 && y >= 1 && y < ny - 1;
        float wv = 0.f, fv = f[i];
        if (isnan(fv)) {
            wv = 1.f;
            fv = avgf;
        }
        sf[sk(x, y)] = fv;
        sw[sk(x, y)] = interior ? __fmul_rn(wv, P.corrEff) : wv;
    }
    __syncthreads();
    const float crtest = __double2float_rn(__dmul_rn(s_crit, static_cast<double>(P.corrEff)));

    // ---- sweeps, in groups that end at the next convergence test (:621-669) ----
    const unsigned long long maxLoop = P.maxLoop;
    const int last = ny - 1;
    unsigned long long n = 0;
    bool converged = false;
    while (n < maxLoop && !converged) {
        unsigned long long n1 = (n % 10 == 0) ? n : n + (10 - n % 10);
        if (n1 > maxLoop - 1)
            n1 = maxLoop - 1;
        const int G = static_cast<int>(n1 - n + 1);
        const bool tested = (n1 < (maxLoop - 5)) && (n1 % 10 == 0); // unsigned wrap for maxLoop < 5, as in the reference
        int bad = 0;
        const int tmax = (nx + ny - 2) + 2 * (G - 1);
        for (int t = 2; t <= tmax; ++t) {
            // interior cells: one anti-diagonal per sweep in flight.  All cells of a step are independent, so a thread first
            // sends for the operands of its cell on every diagonal (up to GMAX x 6 loads in flight), then computes and stores.
            for (int x0 = tid; x0 < nx; x0 += FT) {
                float c0[GMAX], cE[GMAX], cW[GMAX], cS[GMAX], cN[GMAX], cw[GMAX];
                size_t idx[GMAX];
#pragma unroll
                for (int g = 0; g < GMAX; ++g) {
                    const int d = t - 2 * g; // this sweep's anti-diagonal x + y == d
                    const int x = max(1, d - (ny - 2)) + x0;
                    const bool on = g < G && d >= 2 && x <= min(nx - 2, d - 1);
                    idx[g] = on ? static_cast<size_t>(d) * nx + x : 0;
                    if (on) {
                        const float* q = sf + idx[g];
                        c0[g] = q[0];
                        cE[g] = q[nx + 1];  // (x+1, y): next diagonal
                        cW[g] = q[-nx - 1]; // (x-1, y): previous diagonal
                        cS[g] = q[nx];      // (x, y+1)
                        cN[g] = q[-nx];     // (x, y-1)
                        cw[g] = sw[idx[g]];
                    }
                }
#pragma unroll
                for (int g = 0; g < GMAX; ++g) {
                    if (idx[g]) { // interpolation.c:629-631
                        const float sum4 = __fadd_rn(__fadd_rn(__fadd_rn(cE[g], cW[g]), cS[g]), cN[g]);
                        const float e = __double2float_rn(__dsub_rn(__dmul_rn(static_cast<double>(sum4), 0.25), static_cast<double>(c0[g])));
                        const float c = __fmul_rn(e, cw[g]);
                        sf[idx[g]] = __fadd_rn(c0[g], c);
                        if (tested && g == G - 1 && fabsf(c) > crtest)
                            bad = 1;
                    }
                }
            }
            if (tid < 8) {
                for (int g = 0; g < G - 1; ++g) { // borders of a sweep that is followed by another one of this group, one step behind
                    const int d = t - 2 * g;
                    if (d < 2)
                        break;
                    if (tid == 0 && d - 2 >= 1 && d - 2 <= ny - 2) // west border (0, y)
                        border_cell(sf, sw, sk(0, d - 2), sk(1, d - 2));
                    if (tid == 1 && d - (nx - 1) >= 1 && d - (nx - 1) <= ny - 2) // east border (nx-1, y)
                        border_cell(sf, sw, sk(nx - 1, d - (nx - 1)), sk(nx - 2, d - (nx - 1)));
                    if (tid == 2 && d - 2 >= 1 && d - 2 <= nx - 2) // north border (x, 0)
                        border_cell(sf, sw, sk(d - 2, 0), sk(d - 2, 1));
                    if (tid == 3 && d - last >= 1 && d - last <= nx - 2) // south border (x, ny-1)
                        border_cell(sf, sw, sk(d - last, last), sk(d - last, last - 1));
                    if (tid == 4 && d == 4) // corners follow the border cell they read (the row loop runs after the column loop)
                        border_cell(sf, sw, sk(0, 0), sk(0, 1));
                    if (tid == 5 && d == nx + 1)
                        border_cell(sf, sw, sk(nx - 1, 0), sk(nx - 1, 1));
                    if (tid == 6 && d == ny + 1)
                        border_cell(sf, sw, sk(0, last), sk(0, last - 1));
                    if (tid == 7 && d == nx + ny - 2)
                        border_cell(sf, sw, sk(nx - 1, last), sk(nx - 1, last - 1));
                }
            }
            __syncthreads();
        }
        if (tested && !__syncthreads_or(bad)) {
            converged = true; // the borders of this sweep stay as they are (:652-656)
        } else {
            // borders of the group's last sweep (:660-667): columns, then rows
            for (int y = 1 + tid; y < ny - 1; y += FT) {
                border_cell(sf, sw, sk(0, y), sk(1, y));
                border_cell(sf, sw, sk(nx - 1, y), sk(nx - 2, y));
            }
            __syncthreads();
            for (int x = tid; x < nx; x += FT) {
                border_cell(sf, sw, sk(x, 0), sk(x, 1));
                border_cell(sf, sw, sk(x, last), sk(x, last - 1));
            }
            __syncthreads();
            n = n1 + 1;
        }
    }
    // back to the row-major level
    for (size_t i = tid; i < total; i += FT) {
        const int y = static_cast<int>(i / nx), x = static_cast<int>(i - static_cast<size_t>(y) * nx);
        f[i] = sf[sk(x, y)];
    }
}

// ---- creepfill (src/interpolation.c:680-810) ------------------------------------------------------------------------------------
// Same wavefront as fill2d (a cell sees the west / north neighbours of this sweep, the east / south ones of the previous sweep;
// here that includes the neighbours' WEIGHTS, which a fill switches from 0 to 1).  The reference stops after the first sweep that
// changes nothing; such a sweep leaves a fixed point, so running the rest of a group of sweeps is harmless, and the group's
// last sweep tells whether to go on.  Border cells are not touched by the sweeps.
struct CreepParams {
    float* field;       // [nz][ny][nx], in place
    float* sf;          // diagonal-major scratch, see Fill2dParams
    signed char* sw;    // weights: 0 undefined, setWeight defined, 1 filled
    unsigned short* sr; // how often a cell was filled (defined cells start at `repeat`)
    int nx, ny;
    size_t level, slevel;
    int use_mean;       // creepfill2d: first guess = field mean; creepfillval2d: defaultVal
    float defaultVal;
    unsigned short repeat;
    signed char setWeight;
    unsigned long long* nChanged;
};

// field[i] += field[j]*w[j]; field[i] /= (1 + w[j])  (:758-779), one application
__device__ __forceinline__ float creep_border_step(float fi, float fj, int wj)
{
    return __fdiv_rn(__fadd_rn(fi, __fmul_rn(fj, static_cast<float>(wj))), static_cast<float>(1 + wj));
}

__global__ void __launch_bounds__(FT) creepfill_kernel(const CreepParams P)
{
    __shared__ double stage[FCHUNK];
    __shared__ float s_default;
    __shared__ unsigned long long s_nchg;
    const int nx = P.nx, ny = P.ny, tid = threadIdx.x;
    const size_t total = static_cast<size_t>(nx) * ny;
    float* f = P.field + blockIdx.x * P.level;
    float* sf = P.sf + blockIdx.x * P.slevel;
    signed char* sw = P.sw + blockIdx.x * P.slevel;
    unsigned short* sr = P.sr + blockIdx.x * P.slevel;
    auto sk = [nx](int x, int y) { return static_cast<size_t>(x + y) * nx + x; };
    const int repeat = P.repeat;

    unsigned long long mine = 0;
    for (size_t i = tid; i < total; i += FT)
        mine += isnan(f[i]) ? 1u : 0u;
    if (tid == 0)
        s_nchg = 0;
    __syncthreads();
    if (mine)
        atomicAdd(&s_nchg, mine);
    __syncthreads();
    const unsigned long long nChanged = s_nchg, nUnchanged = total - nChanged;
    if (tid == 0)
        P.nChanged[blockIdx.x] = nChanged;
    if (nUnchanged == 0 || nChanged == 0)
        return; // :687-690, :791
    if (P.use_mean) { // float average = sum / nUnchanged with the f64 sum in index order (:783-794)
        const double sum = sequential_sum(total, stage, [&](size_t i) {
            const float v = f[i];
            return isnan(v) ? 0. : static_cast<double>(v);
        });
        if (tid == 0)
            s_default = __double2float_rn(__ddiv_rn(sum, static_cast<double>(nUnchanged)));
    } else if (tid == 0) {
        s_default = P.defaultVal;
    }
    __syncthreads();
    const float first_guess = s_default;
    for (size_t i = tid; i < total; i += FT) { // :703-715
        const int y = static_cast<int>(i / nx), x = static_cast<int>(i - static_cast<size_t>(y) * nx);
        const float v = f[i];
        const bool undef = isnan(v);
        sf[sk(x, y)] = undef ? first_guess : v;
        sw[sk(x, y)] = undef ? static_cast<signed char>(0) : P.setWeight;
        sr[sk(x, y)] = undef ? static_cast<unsigned short>(0) : P.repeat;
    }
    __syncthreads();

    // ---- sweeps (:724-754) ----
    unsigned long long l = 0;
    while (l < nUnchanged) {
        const int G = static_cast<int>(min(static_cast<unsigned long long>(GMAX), nUnchanged - l));
        int changed = 0;
        const int tmax = (nx + ny - 4) + 2 * (G - 1);
        for (int t = 2; t <= tmax; ++t) {
            for (int x0 = tid; x0 < nx; x0 += FT) {
#pragma unroll
                for (int g = 0; g < GMAX; ++g) {
                    const int d = t - 2 * g;
                    const int x = max(1, d - (ny - 2)) + x0;
                    if (g < G && d >= 2 && x <= min(nx - 2, d - 1)) {
                        const size_t i = static_cast<size_t>(d) * nx + x;
                        if (sr[i] < repeat) {
                            const int wE = sw[i + nx + 1], wW = sw[i - nx - 1], wS = sw[i + nx], wN = sw[i - nx];
                            const size_t wsum = static_cast<size_t>(static_cast<long long>(wE + wW + wS + wN)); // int sum -> size_t, as in C
                            if (wsum != 0) {
                                const float nb = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(static_cast<float>(wE), sf[i + nx + 1]),
                                                                               __fmul_rn(static_cast<float>(wW), sf[i - nx - 1])),
                                                                     __fmul_rn(static_cast<float>(wS), sf[i + nx])),
                                                           __fmul_rn(static_cast<float>(wN), sf[i - nx]));
                                sf[i] = __fdiv_rn(__fadd_rn(sf[i], nb), static_cast<float>(1 + wsum));
                                sw[i] = 1;
                                sr[i] = static_cast<unsigned short>(sr[i] + 1);
                                if (g == G - 1)
                                    changed = 1;
                            }
                        }
                    }
                }
            }
            __syncthreads();
        }
        l += G;
        if (!__syncthreads_or(changed))
            break;
    }

    // ---- borders (:756-781): `repeat` rounds of columns, then rows; an undefined border cell (r == 0 < repeat) is pulled towards
    // its inner neighbour every round.  The corners read a border cell of the same round, so their thread replays that cell.
    if (repeat > 0) {
        const int last = ny - 1;
        float corner = 0.f;
        bool corner_set = false;
        if (tid < 4) {
            const int cx = (tid & 1) ? nx - 1 : 0, cy = (tid & 2) ? last : 0; // the corner
            const int by = (tid & 2) ? last - 1 : 1;                            // its column neighbour (cx, by): a west / east border cell
            const int ix = (tid & 1) ? nx - 2 : 1;                              // that cell's inner neighbour (ix, by)
            if (sr[sk(cx, cy)] < repeat) {
                float fb = sf[sk(cx, by)];
                int wb = sw[sk(cx, by)];
                const bool b_undef = sr[sk(cx, by)] < repeat;
                const float fi = sf[sk(ix, by)];
                const int wi = sw[sk(ix, by)];
                corner = sf[sk(cx, cy)];
                for (int r = 0; r < repeat; ++r) {
                    if (b_undef) { // the column loop of this round comes first (:757-768)
                        fb = creep_border_step(fb, fi, wi);
                        wb = 1;
                    }
                    corner = creep_border_step(corner, fb, wb);
                }
                corner_set = true;
            }
        }
        __syncthreads(); // the corners have read their neighbours' first-guess values
        if (corner_set)
            sf[sk((tid & 1) ? nx - 1 : 0, (tid & 2) ? last : 0)] = corner;
        for (int y = 1 + tid; y < ny - 1; y += FT) {
            if (sr[sk(0, y)] < repeat) {
                float v = sf[sk(0, y)];
                const float fi = sf[sk(1, y)];
                const int wi = sw[sk(1, y)];
                for (int r = 0; r < repeat; ++r)
                    v = creep_border_step(v, fi, wi);
                sf[sk(0, y)] = v;
            }
            if (sr[sk(nx - 1, y)] < repeat) {
                float v = sf[sk(nx - 1, y)];
                const float fi = sf[sk(nx - 2, y)];
                const int wi = sw[sk(nx - 2, y)];
                for (int r = 0; r < repeat; ++r)
                    v = creep_border_step(v, fi, wi);
                sf[sk(nx - 1, y)] = v;
            }
        }
        for (int x = 1 + tid; x < nx - 1; x += FT) {
            if (sr[sk(x, 0)] < repeat) {
                float v = sf[sk(x, 0)];
                const float fi = sf[sk(x, 1)];
                const int wi = sw[sk(x, 1)];
                for (int r = 0; r < repeat; ++r)
                    v = creep_border_step(v, fi, wi);
                sf[sk(x, 0)] = v;
            }
            if (sr[sk(x, last)] < repeat) {
                float v = sf[sk(x, last)];
                const float fi = sf[sk(x, last - 1)];
                const int wi = sw[sk(x, last - 1)];
                for (int r = 0; r < repeat; ++r)
                    v = creep_border_step(v, fi, wi);
                sf[sk(x, last)] = v;
            }
        }
        __syncthreads();
    }
    for (size_t i = tid; i < total; i += FT) {
        const int y = static_cast<int>(i / nx), x = static_cast<int>(i - static_cast<size_t>(y) * nx);
        f[i] = sf[sk(x, y)];
    }
}

} // namespace fxcu

using namespace fxcu;

extern "C" {

int fimexcu_fill2d_dev_f(size_t nx, size_t ny, size_t nz, float* d_field, float relaxCrit, float corrEff, size_t maxLoop, size_t* nChanged_host,
                         void* stream)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    if (nx * ny == 0 || nz == 0)
        return FIMEXCU_OK;
    if (nx < 3 || ny < 3 || nx > 0x3fffffff || ny > 0x3fffffff || nx * ny > 0x7fffffffull)
        FXCU_FAIL("fimexcu_fill2d: the grid must be at least 3 x 3 and smaller than 2^31 cells");
    if (nz > 0x7fffffffull)
        FXCU_FAIL("fimexcu_fill2d: too many levels");
    if (!d_field)
        FXCU_FAIL("fimexcu_fill2d: null field");
    cudaStream_t s = as_stream(stream);
    DevBuf sf, sw, cnt;
    const size_t slevel = (nx + ny - 1) * nx;
    if (!sf.alloc(slevel * nz * sizeof(float), s) || !sw.alloc(slevel * nz * sizeof(float), s) || !cnt.alloc(nz * sizeof(unsigned long long), s))
        return FIMEXCU_ERROR;
    Fill2dParams P{};
    P.field = d_field;
    P.sf = sf.as<float>();
    P.sw = sw.as<float>();
    P.slevel = slevel;
    P.nx = static_cast<int>(nx);
    P.ny = static_cast<int>(ny);
    P.level = nx * ny;
    P.relaxCrit = relaxCrit;
    P.corrEff = corrEff;
    P.maxLoop = maxLoop;
    P.nChanged = cnt.as<unsigned long long>();
    fill2d_kernel<<<static_cast<unsigned>(nz), FT, 0, s>>>(P);
    FXCU_CUDA(cudaGetLastError());
    if (nChanged_host) {
        std::vector<unsigned long long> h(nz);
        FXCU_CUDA(cudaMemcpyAsync(h.data(), cnt.p, nz * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
        FXCU_CUDA(cudaStreamSynchronize(s));
        for (size_t k = 0; k < nz; ++k)
            nChanged_host[k] = static_cast<size_t>(h[k]);
    }
    return FIMEXCU_OK;
}

int fimexcu_fill2d_f(size_t nx, size_t ny, float* field, float relaxCrit, float corrEff, size_t maxLoop, size_t* nChanged)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    if (nx * ny == 0)
        return FIMEXCU_OK;
    const size_t bytes = nx * ny * sizeof(float);
    DevBuf d;
    if (!d.alloc(bytes, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(d.p, field, bytes, cudaMemcpyHostToDevice, nullptr));
    size_t changed = 0;
    if (fimexcu_fill2d_dev_f(nx, ny, 1, d.as<float>(), relaxCrit, corrEff, maxLoop, &changed, nullptr) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(field, d.p, bytes, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    if (nChanged)
        *nChanged = changed;
    return FIMEXCU_OK;
}

int fimexcu_creepfill2d_dev_f(size_t nx, size_t ny, size_t nz, float* d_field, int use_mean, float defaultVal, unsigned short repeat, char setWeight,
                              size_t* nChanged_host, void* stream)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    if (nx * ny == 0 || nz == 0)
        return FIMEXCU_OK;
    if (nx < 3 || ny < 3 || nx > 0x3fffffff || ny > 0x3fffffff || nx * ny > 0x7fffffffull)
        FXCU_FAIL("fimexcu_creepfill2d: the grid must be at least 3 x 3 and smaller than 2^31 cells");
    if (nz > 0x7fffffffull)
        FXCU_FAIL("fimexcu_creepfill2d: too many levels");
    if (!d_field)
        FXCU_FAIL("fimexcu_creepfill2d: null field");
    cudaStream_t s = as_stream(stream);
    DevBuf sf, sw, sr, cnt;
    const size_t slevel = (nx + ny - 1) * nx;
    if (!sf.alloc(slevel * nz * sizeof(float), s) || !sw.alloc(slevel * nz, s) || !sr.alloc(slevel * nz * sizeof(unsigned short), s) ||
        !cnt.alloc(nz * sizeof(unsigned long long), s))
        return FIMEXCU_ERROR;
    CreepParams P{};
    P.field = d_field;
    P.sf = sf.as<float>();
    P.sw = sw.as<signed char>();
    P.sr = sr.as<unsigned short>();
    P.nx = static_cast<int>(nx);
    P.ny = static_cast<int>(ny);
    P.level = nx * ny;
    P.slevel = slevel;
    P.use_mean = use_mean ? 1 : 0;
    P.defaultVal = defaultVal;
    P.repeat = repeat;
    P.setWeight = static_cast<signed char>(setWeight);
    P.nChanged = cnt.as<unsigned long long>();
    creepfill_kernel<<<static_cast<unsigned>(nz), FT, 0, s>>>(P);
    FXCU_CUDA(cudaGetLastError());
    if (nChanged_host) {
        std::vector<unsigned long long> h(nz);
        FXCU_CUDA(cudaMemcpyAsync(h.data(), cnt.p, nz * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
        FXCU_CUDA(cudaStreamSynchronize(s));
        for (size_t k = 0; k < nz; ++k)
            nChanged_host[k] = static_cast<size_t>(h[k]);
    }
    return FIMEXCU_OK;
}

static int creepfill_host(size_t nx, size_t ny, float* field, int use_mean, float defaultVal, unsigned short repeat, char setWeight, size_t* nChanged)
{
    if (!require_device())
        return FIMEXCU_ERROR;
    if (nx * ny == 0)
        return FIMEXCU_OK;
    const size_t bytes = nx * ny * sizeof(float);
    DevBuf d;
    if (!d.alloc(bytes, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(d.p, field, bytes, cudaMemcpyHostToDevice, nullptr));
    size_t changed = 0;
    if (fimexcu_creepfill2d_dev_f(nx, ny, 1, d.as<float>(), use_mean, defaultVal, repeat, setWeight, &changed, nullptr) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(field, d.p, bytes, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    if (nChanged)
        *nChanged = changed;
    return FIMEXCU_OK;
}

int fimexcu_creepfill2d_f(size_t nx, size_t ny, float* field, unsigned short repeat, char setWeight, size_t* nChanged)
{
    clear_error();
    return creepfill_host(nx, ny, field, 1, 0.f, repeat, setWeight, nChanged);
}

int fimexcu_creepfillval2d_f(size_t nx, size_t ny, float* field, float defaultVal, unsigned short repeat, char setWeight, size_t* nChanged)
{
    clear_error();
    return creepfill_host(nx, ny, field, 0, defaultVal, repeat, setWeight, nChanged);
}

} // extern "C"
