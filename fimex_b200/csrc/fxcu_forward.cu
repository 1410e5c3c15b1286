// fxcu_forward.cu -- forward (scatter / aggregate) interpolation: K11 index build, K12 sum/mean/min/max, K13 median.
//
// Reference behaviour restated here (fimex 2.4.1):
//   CachedForwardInterpolation ctor                 src/CachedForwardInterpolation.cc:133-201
//   CachedForwardInterpolation::interpolateValues   src/CachedForwardInterpolation.cc:205-224
//   AggSum / AggMean / AggMin / AggMax / AggMedian   src/CachedForwardInterpolation.cc:63-130
//   RoundAndClamp                                   src/MathUtils.cc:30-46
//
// The reference keeps one std::vector<size_t> of source indices per target cell, filled while walking the source row-major,
// so every list is ascending in source index and the f32 running sums are taken in that order.  Here the lists are a CSR:
// a stable radix sort of (target cell, source index) pairs (one-off, CUB) gives the same order; the apply kernel gives one
// thread one target cell and walks its segment sequentially, so sum and mean are bit-identical to the reference
// (no parallel re-association), min/max/median trivially so.
//
// HBM layout: source [nz][inY][inX] f32 (reduced window), target [nz][outY][outX] f32, CSR segStart int32 [nout+1],
// segIdx int32 [nvalid].  Algorithmic bytes per source point and level: 4 (value) + 4*N_out/N_in (result) + 8/nz (CSR).
#include "fxcu_common.h"

#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <memory>
#include <vector>

struct fimexcu_forward {
    int method = 0;
    int kind = 0; // 0 sum, 1 mean, 2 median, 3 max, 4 min
    bool undef = false;
    size_t inX = 0, inY = 0, outX = 0, outY = 0;
    bool reduced = false;
    size_t xMin = 0, yMin = 0;
    size_t maxPointsInIn = 0;
    size_t nvalid = 0;
    int device = 0;        // where the tables live
    fxcu::DevBuf segStart; // int32 [nout+1]
    fxcu::DevBuf segIdx;   // int32 [nvalid]
};

namespace fxcu {

__device__ __forceinline__ long long fwd_round_and_clamp(double d, long long maxi)
{
    const double r = round(d);
    if (!(r >= 0.0) || r > static_cast<double>(maxi))
        return -1;
    return static_cast<long long>(r);
}

// :141-160 first half: target cell of every source cell (key = nout for "no target"), used source range
__global__ void __launch_bounds__(256) fwd_target_kernel(const double* __restrict__ px, const double* __restrict__ py, int inX, size_t nin, int outX,
                                                         int outY, unsigned* __restrict__ key, unsigned* __restrict__ val, int* __restrict__ range)
{
    const size_t i = blockIdx.x * 256ull + threadIdx.x;
    int lox = 0x7fffffff, hix = -1, loy = 0x7fffffff, hiy = -1;
    if (i < nin) {
        const long long tx = fwd_round_and_clamp(px[i], outX - 1), ty = fwd_round_and_clamp(py[i], outY - 1);
        unsigned k = static_cast<unsigned>(outX) * static_cast<unsigned>(outY);
        if (tx >= 0 && ty >= 0) {
            k = static_cast<unsigned>(ty) * outX + static_cast<unsigned>(tx);
            const int iy = static_cast<int>(i / inX), ix = static_cast<int>(i - static_cast<size_t>(iy) * inX);
            lox = hix = ix;
            loy = hiy = iy;
        }
        key[i] = k;
        val[i] = static_cast<unsigned>(i);
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        lox = min(lox, __shfl_xor_sync(0xffffffffu, lox, s));
        hix = max(hix, __shfl_xor_sync(0xffffffffu, hix, s));
        loy = min(loy, __shfl_xor_sync(0xffffffffu, loy, s));
        hiy = max(hiy, __shfl_xor_sync(0xffffffffu, hiy, s));
    }
    if ((threadIdx.x & 31) == 0 && hix >= 0) {
        atomicMin(&range[0], lox);
        atomicMax(&range[1], hix);
        atomicMin(&range[2], loy);
        atomicMax(&range[3], hiy);
    }
}

// segStart[o] = first sorted position whose key >= o (o = 0..nout); also the longest segment
__global__ void __launch_bounds__(256) fwd_segments_kernel(const unsigned* __restrict__ sortedKey, size_t n, unsigned nout, int* __restrict__ segStart,
                                                           int* __restrict__ maxLen)
{
    const size_t o = blockIdx.x * 256ull + threadIdx.x;
    if (o > nout)
        return;
    size_t lo = 0, hi = n;
    while (lo < hi) {
        const size_t mid = (lo + hi) / 2;
        if (sortedKey[mid] < o)
            lo = mid + 1;
        else
            hi = mid;
    }
    segStart[o] = static_cast<int>(lo);
    if (o < nout) {
        size_t e = lo;
        while (e < n && sortedKey[e] == o)
            ++e;
        atomicMax(maxLen, static_cast<int>(e - lo));
    }
}

// :177-183: source indices re-expressed in the reduced window
__global__ void __launch_bounds__(256) fwd_rebase_kernel(const unsigned* __restrict__ sortedVal, size_t nvalid, int inX, int minX, int minY, int redX,
                                                         int* __restrict__ segIdx)
{
    const size_t k = blockIdx.x * 256ull + threadIdx.x;
    if (k >= nvalid)
        return;
    const int i = static_cast<int>(sortedVal[k]);
    const int iy = i / inX - minY, ix = i % inX - minX;
    segIdx[k] = iy * redX + ix;
}

struct FwdParams {
    const int* segStart;
    const int* segIdx;
    const float* in;
    float* out;
    size_t nout, inLayer;
    int nz, zchunk;
    int kind;
    bool undef, level_compat;
    float bad, fill;
};

__device__ __forceinline__ float fwd_load(const float* p, float bad, bool has_bad)
{
    float v = __ldg(p);
    if (has_bad && v == bad)
        v = nan_f();
    return v;
}

// K12/K13: one thread per target cell, levels z0..z1; the segment is walked in ascending source order
__global__ void __launch_bounds__(256) fwd_apply_kernel(const FwdParams P)
{
    const size_t o = blockIdx.x * 256ull + threadIdx.x;
    if (o >= P.nout)
        return;
    const int z0 = blockIdx.y * P.zchunk, z1 = min(P.nz, z0 + P.zchunk);
    const int s0 = P.segStart[o], s1 = P.segStart[o + 1];
    const bool has_bad = !isnan(P.bad), has_fill = !isnan(P.fill);
    float* po = P.out + static_cast<size_t>(z0) * P.nout + o;
    float carried = nan_f();
    for (int z = z0; z < z1; ++z, po += P.nout) {
        float r;
        if (P.level_compat && z > z0) {
            r = carried; // the reference reads input level 0 for every output level (CachedForwardInterpolation.cc:212-221)
        } else {
            const float* src = P.in + (P.level_compat ? 0 : static_cast<size_t>(z) * P.inLayer);
            int count = 0;
            float val = 0.f;
            if (P.kind != 2) {
                for (int k = s0; k < s1; ++k) {
                    const float f = fwd_load(src + P.segIdx[k], P.bad, has_bad);
                    if (!P.undef && isnan(f))
                        continue;
                    if (P.kind <= 1)
                        val = __fadd_rn(val, f); // AggSum::push, f32 running sum
                    else if (P.kind == 3) {
                        if (count == 0 || f > val)
                            val = f; // AggMax::push
                    } else {
                        if (count == 0 || f < val)
                            val = f; // AggMin::push
                    }
                    ++count;
                }
                if (count == 0)
                    r = nan_f();
                else if (P.kind == 1)
                    r = __fdiv_rn(val, static_cast<float>(count)); // val / count, count converted to float (:90)
                else
                    r = val;
            } else {
                // AggMedian: nth_element at n/2 == element of rank n/2 among the pushed values.  Rank selection without
                // storage: v has rank k iff #(values < v) <= k < #(values < v) + #(values == v).
                int n = 0;
                bool any_nan = false;
                for (int k = s0; k < s1; ++k) {
                    const float f = fwd_load(src + P.segIdx[k], P.bad, has_bad);
                    if (isnan(f)) {
                        if (P.undef) {
                            any_nan = true;
                            ++n;
                        }
                    } else {
                        ++n;
                    }
                }
                r = nan_f();
                if (n > 0 && !any_nan) { // NaN among the values: unspecified order in the reference, defined as NaN
                    const int rank = n / 2;
                    for (int k = s0; k < s1; ++k) {
                        const float v = fwd_load(src + P.segIdx[k], P.bad, has_bad);
                        if (isnan(v))
                            continue;
                        int less = 0, equal = 0;
                        for (int j = s0; j < s1; ++j) {
                            const float w = fwd_load(src + P.segIdx[j], P.bad, has_bad);
                            less += (w < v) ? 1 : 0;
                            equal += (w == v) ? 1 : 0;
                        }
                        if (less <= rank && rank < less + equal) {
                            r = v;
                            break;
                        }
                    }
                }
            }
            carried = r;
        }
        float w = r;
        if (has_fill && isnan(w))
            w = P.fill;
        __stcs(po, w);
    }
}

// ---- round-2 forms of the apply -----------------------------------------------------------------------------------------------
// One block of NK <= ZB levels of one target cell.  FULL: NK == ZB is known at compile time (no per-level tests).  BAD: the input
// carries a bad value that counts as NaN.  Source offsets are 32-bit (the launch guarantees ZB * inLayer < 2^31).
template <int KIND, bool UNDEF, bool BAD, int ZB, bool FULL, int KU = 2>
__device__ __forceinline__ void fwd_reduce_block(const FwdParams& P, const float* __restrict__ src, int s0, int s1, int nk, float* __restrict__ dst, bool has_fill)
{
    const unsigned lay = static_cast<unsigned>(P.inLayer);
    float val[ZB];
    int cnt[ZB];
#pragma unroll
    for (int q = 0; q < ZB; ++q) {
        val[q] = 0.f;
        cnt[q] = 0;
    }
    auto load = [&](unsigned idx, int q) {
        float v = (FULL || q < nk) ? __ldg(src + (idx + static_cast<unsigned>(q) * lay)) : nan_f();
        if (BAD && v == P.bad)
            v = nan_f();
        return v;
    };
    auto push = [&](const float (&f)[ZB]) {
#pragma unroll
        for (int q = 0; q < ZB; ++q) {
            const bool take = UNDEF || !isnan(f[q]);
            if (KIND <= 1) {
                if (take)
                    val[q] = __fadd_rn(val[q], f[q]); // AggSum::push
            } else if (KIND == 3) {
                if (take && (cnt[q] == 0 || f[q] > val[q]))
                    val[q] = f[q]; // AggMax::push
            } else {
                if (take && (cnt[q] == 0 || f[q] < val[q]))
                    val[q] = f[q]; // AggMin::push
            }
            if (take)
                ++cnt[q];
        }
    };
    int k = s0;
    if (KU == 2) {
        for (; k + 2 <= s1; k += 2) { // two sources at a time: their index loads, then all their gathers, are in flight together
            const unsigned i0 = static_cast<unsigned>(__ldg(P.segIdx + k)), i1 = static_cast<unsigned>(__ldg(P.segIdx + k + 1));
            float f0[ZB], f1[ZB];
#pragma unroll
            for (int q = 0; q < ZB; ++q) {
                f0[q] = load(i0, q);
                f1[q] = load(i1, q);
            }
            push(f0);
            push(f1);
        }
    }
    for (; KU != 2 && k + KU <= s1; k += KU) { // KU sources at a time
        unsigned ix[KU];
#pragma unroll
        for (int u = 0; u < KU; ++u)
            ix[u] = static_cast<unsigned>(__ldg(P.segIdx + k + u));
        float f[KU][ZB];
#pragma unroll
        for (int q = 0; q < ZB; ++q)
#pragma unroll
            for (int u = 0; u < KU; ++u)
                f[u][q] = load(ix[u], q);
#pragma unroll
        for (int u = 0; u < KU; ++u)
            push(f[u]); // ascending source order
    }
    if (KU == 2) { // at most one source left
        if (k < s1) {
            const unsigned i0 = static_cast<unsigned>(__ldg(P.segIdx + k));
            float f[ZB];
#pragma unroll
            for (int q = 0; q < ZB; ++q)
                f[q] = load(i0, q);
            push(f);
        }
    } else {
        for (; k < s1; ++k) {
            const unsigned i0 = static_cast<unsigned>(__ldg(P.segIdx + k));
            float f[ZB];
#pragma unroll
            for (int q = 0; q < ZB; ++q)
                f[q] = load(i0, q);
            push(f);
        }
    }
    const unsigned nout = static_cast<unsigned>(P.nout);
#pragma unroll
    for (int q = 0; q < ZB; ++q) {
        if (!FULL && q >= nk)
            break;
        float r = nan_f();
        if (cnt[q] != 0)
            r = KIND == 1 ? __fdiv_rn(val[q], static_cast<float>(cnt[q])) : val[q]; // mean: val / count (:90)
        if (has_fill && isnan(r))
            r = P.fill;
        __stcs(dst + static_cast<unsigned>(q) * nout, r);
    }
}

// Sum / mean / min / max, thread per target cell, the sources in the OUTER loop and a block of ZB levels in registers: the segment's
// source indices are read once per level block instead of once per level, and ZB independent gathers are in flight per source.
// Every level's values are still pushed in ascending source order, so the f32 running sums are the reference's bit for bit
// (CachedForwardInterpolation.cc:79-86, 143-158).  The launch guarantees ZB * inLayer and ZB * nout < 2^31.
template <int KIND, bool UNDEF, bool BAD, int ZB>
__global__ void __launch_bounds__(256) fwd_reduce_kernel(const FwdParams P)
{
    const size_t o = blockIdx.x * 256ull + threadIdx.x;
    if (o >= P.nout)
        return;
    const int z0 = blockIdx.y * P.zchunk, z1 = min(P.nz, z0 + P.zchunk);
    const int s0 = P.segStart[o], s1 = P.segStart[o + 1];
    const bool has_fill = !isnan(P.fill);
    const float* src = P.in + static_cast<size_t>(z0) * P.inLayer;
    float* dst = P.out + static_cast<size_t>(z0) * P.nout + o;
    int zb = z0;
    constexpr int KU = ZB >= 8 ? 2 : (ZB >= 4 ? 4 : 8); // sixteen gathers in flight per thread whatever the level block
    for (; zb + ZB <= z1; zb += ZB, src += static_cast<size_t>(ZB) * P.inLayer, dst += static_cast<size_t>(ZB) * P.nout)
        fwd_reduce_block<KIND, UNDEF, BAD, ZB, true, KU>(P, src, s0, s1, ZB, dst, has_fill);
    if (zb < z1)
        fwd_reduce_block<KIND, UNDEF, BAD, ZB, false, KU>(P, src, s0, s1, z1 - zb, dst, has_fill);
}

// Min / max without the undef flavour do not depend on the order of the values: long segments (coarse targets: hundreds of sources
// per cell) are taken by a whole warp -- lanes stride over the segment, so a warp reads runs of consecutive source cells instead of
// 32 cells that lie a target-cell width apart -- and combined with shuffles.  One warp per target cell, ZB levels per pass.
template <bool MAX, int ZB>
__global__ void __launch_bounds__(256) fwd_minmax_warp_kernel(const FwdParams P)
{
    const size_t o = (blockIdx.x * 256ull + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (o >= P.nout)
        return;
    const int z0 = blockIdx.y * P.zchunk, z1 = min(P.nz, z0 + P.zchunk);
    const int s0 = P.segStart[o], s1 = P.segStart[o + 1];
    const bool has_bad = !isnan(P.bad), has_fill = !isnan(P.fill);
    for (int zb = z0; zb < z1; zb += ZB) {
        const int nk = min(ZB, z1 - zb);
        const float* src = P.in + static_cast<size_t>(zb) * P.inLayer;
        float val[ZB];
#pragma unroll
        for (int q = 0; q < ZB; ++q)
            val[q] = nan_f(); // NaN = nothing pushed yet
        for (int k = s0 + lane; k < s1; k += 32) {
            const float* p = src + P.segIdx[k];
#pragma unroll
            for (int q = 0; q < ZB; ++q) {
                if (q < nk) {
                    const float f = fwd_load(p + static_cast<size_t>(q) * P.inLayer, P.bad, has_bad);
                    if (!isnan(f) && (isnan(val[q]) || (MAX ? f > val[q] : f < val[q])))
                        val[q] = f;
                }
            }
        }
#pragma unroll
        for (int q = 0; q < ZB; ++q) {
#pragma unroll
            for (int sft = 16; sft > 0; sft >>= 1) {
                const float other = __shfl_xor_sync(0xffffffffu, val[q], sft);
                if (!isnan(other) && (isnan(val[q]) || (MAX ? other > val[q] : other < val[q])))
                    val[q] = other;
            }
            if (lane == 0 && q < nk) {
                float r = val[q];
                if (has_fill && isnan(r))
                    r = P.fill;
                __stcs(P.out + static_cast<size_t>(zb + q) * P.nout + o, r);
            }
        }
    }
}

// Median (AggMedian: nth_element at n/2, CachedForwardInterpolation.cc:113-130): the element of rank n/2 among the pushed values.
// The segment's values of one level are loaded ONCE: up to MED_LOCAL of them into a per-thread array, where the rank is found by
// counting (v has rank r iff #less <= r < #less + #equal); longer segments by a radix selection on the ordered bit pattern of the
// floats, most significant bit first (32 sweeps over the segment, each a sequential re-read that stays in L1/L2) -- O(32 n)
// instead of the O(n^2) loads of round 1.
constexpr int MED_LOCAL = 32;
__device__ __forceinline__ unsigned ordered_bits(float v)
{
    const unsigned b = __float_as_uint(v);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
// Segments of at most NMAX <= 8 values (a fine target: two to six sources per cell) never leave the registers: the source offsets are
// read once for all levels, a level's values are sorted by a fixed network of compare-exchanges (missing and skipped values count
// as +inf behind the real ones) and the element of rank n/2 is picked.  Same result as the counting form below.
template <int N> __device__ __forceinline__ void sort_network(float (&v)[N])
{
    auto ce = [&](int i, int j) {
        const float lo = fminf(v[i], v[j]), hi = fmaxf(v[i], v[j]);
        v[i] = lo;
        v[j] = hi;
    };
    if (N == 2) {
        ce(0, 1);
    } else if (N == 4) {
        ce(0, 1); ce(2, 3); ce(0, 2); ce(1, 3); ce(1, 2);
    } else { // 8 inputs, 19 compare-exchanges
        ce(0, 1); ce(2, 3); ce(4, 5); ce(6, 7);
        ce(0, 2); ce(1, 3); ce(4, 6); ce(5, 7);
        ce(1, 2); ce(5, 6); ce(0, 4); ce(3, 7);
        ce(1, 5); ce(2, 6);
        ce(1, 4); ce(3, 6);
        ce(2, 4); ce(3, 5);
        ce(3, 4);
    }
}
template <bool UNDEF, int NMAX>
__device__ __forceinline__ void median_small(const FwdParams& P, size_t o, int s0, int nseg, int z0, int z1, bool has_bad, bool has_fill)
{
    static_assert(NMAX == 2 || NMAX == 4 || NMAX == 8, "sorting networks for 2, 4 and 8 values");
    unsigned idx[NMAX];
#pragma unroll
    for (int j = 0; j < NMAX; ++j)
        idx[j] = j < nseg ? static_cast<unsigned>(__ldg(P.segIdx + s0 + j)) : 0u;
    const float* src = P.in + static_cast<size_t>(z0) * P.inLayer;
    float* dst = P.out + static_cast<size_t>(z0) * P.nout + o;
    for (int z = z0; z < z1; ++z, src += P.inLayer, dst += P.nout) {
        float v[NMAX];
        int n = 0;
        bool any_nan = false;
#pragma unroll
        for (int j = 0; j < NMAX; ++j) {
            float f = INFINITY;
            if (j < nseg) {
                f = __ldg(src + idx[j]);
                if (has_bad && f == P.bad)
                    f = nan_f();
                if (isnan(f)) {
                    any_nan = any_nan || UNDEF;
                    f = INFINITY;
                } else {
                    ++n;
                }
            }
            v[j] = f;
        }
        sort_network<NMAX>(v);
        float r = nan_f();
        if (n > 0 && !any_nan) { // NaN among the values: unspecified order in the reference, defined as NaN
            const int rank = n / 2;
            r = v[0];
#pragma unroll
            for (int j = 1; j < NMAX; ++j)
                r = rank == j ? v[j] : r;
            if (r == 0.f) { // one of the pushed zeros: -0.0 only if no +0.0 was pushed (as the counting form decides)
                bool pos = false;
#pragma unroll
                for (int j = 0; j < NMAX; ++j)
                    pos = pos || (v[j] == 0.f && !signbit(v[j]));
                r = pos ? 0.f : -0.f;
            }
        }
        if (has_fill && isnan(r))
            r = P.fill;
        __stcs(dst, r);
    }
}

template <bool UNDEF>
__global__ void __launch_bounds__(128) fwd_median_kernel(const FwdParams P)
{
    const size_t o = blockIdx.x * 128ull + threadIdx.x;
    const bool live = o < P.nout;
    const int z0 = blockIdx.y * P.zchunk, z1 = min(P.nz, z0 + P.zchunk);
    const int s0 = live ? P.segStart[o] : 0, s1 = live ? P.segStart[o + 1] : 0;
    const bool has_bad = !isnan(P.bad), has_fill = !isnan(P.fill);
    // the longest segment of the warp picks the form (one branch for the whole warp); offsets must fit 32 bits for the small forms
    const int nmax = __reduce_max_sync(0xffffffffu, s1 - s0);
    if (nmax <= 8 && static_cast<size_t>(P.inLayer) < (size_t(1) << 31)) {
        if (!live)
            return;
        if (nmax <= 2)
            median_small<UNDEF, 2>(P, o, s0, s1 - s0, z0, z1, has_bad, has_fill);
        else if (nmax <= 4)
            median_small<UNDEF, 4>(P, o, s0, s1 - s0, z0, z1, has_bad, has_fill);
        else
            median_small<UNDEF, 8>(P, o, s0, s1 - s0, z0, z1, has_bad, has_fill);
        return;
    }
    if (!live)
        return;
    float v[MED_LOCAL];
    for (int z = z0; z < z1; ++z) {
        const float* src = P.in + static_cast<size_t>(z) * P.inLayer;
        float r = nan_f();
        int n = 0;
        bool any_nan = false;
        if (s1 - s0 <= MED_LOCAL) {
            for (int k = s0; k < s1; ++k) {
                const float f = fwd_load(src + P.segIdx[k], P.bad, has_bad);
                if (isnan(f)) {
                    any_nan = any_nan || UNDEF;
                    n += UNDEF ? 1 : 0;
                } else {
                    v[n++] = f;
                }
            }
            if (n > 0 && !any_nan) { // NaN among the values: unspecified order in the reference, defined as NaN
                const int rank = n / 2;
                for (int i = 0; i < n; ++i) {
                    int less = 0, equal = 0;
                    for (int j = 0; j < n; ++j) {
                        less += (v[j] < v[i]) ? 1 : 0;
                        equal += (v[j] == v[i]) ? 1 : 0;
                    }
                    if (less <= rank && rank < less + equal) {
                        r = v[i];
                        break;
                    }
                }
            }
        } else {
            for (int k = s0; k < s1; ++k) {
                const float f = fwd_load(src + P.segIdx[k], P.bad, has_bad);
                if (isnan(f))
                    any_nan = any_nan || UNDEF;
                else
                    ++n;
            }
            if (n > 0 && !any_nan) {
                // -0.0 and +0.0 compare equal in nth_element but differ as bit patterns: they are made one key (+0.0's)
                int rank = n / 2;
                unsigned prefix = 0, mask = 0;
                for (int bit = 31; bit >= 0; --bit) {
                    const unsigned m2 = mask | (1u << bit);
                    int zeros = 0;
                    for (int k = s0; k < s1; ++k) {
                        float f = fwd_load(src + P.segIdx[k], P.bad, has_bad);
                        if (isnan(f))
                            continue;
                        if (f == 0.f)
                            f = 0.f;
                        const unsigned key = ordered_bits(f);
                        zeros += ((key & m2) == prefix) ? 1 : 0; // matches the prefix so far and has a 0 in this bit
                    }
                    if (rank >= zeros) {
                        rank -= zeros;
                        prefix |= 1u << bit;
                    }
                    mask = m2;
                }
                const unsigned b = (prefix & 0x80000000u) ? (prefix & 0x7fffffffu) : ~prefix;
                r = __uint_as_float(b);
                // the reference returns one of the pushed values: give -0.0 back if the rank falls on a -0.0 and no +0.0 was pushed
                if (r == 0.f) {
                    bool pos = false, neg = false;
                    for (int k = s0; k < s1; ++k) {
                        const float f = fwd_load(src + P.segIdx[k], P.bad, has_bad);
                        if (f == 0.f) {
                            pos = pos || !signbit(f);
                            neg = neg || signbit(f);
                        }
                    }
                    r = (neg && !pos) ? -0.f : 0.f;
                }
            }
        }
        if (has_fill && isnan(r))
            r = P.fill;
        __stcs(P.out + static_cast<size_t>(z) * P.nout + o, r);
    }
}

// Median of long segments: one warp per target cell.  The level's values are gathered once, lanes striding over the segment, into
// the warp's slice of shared memory (segments beyond MED_WARP_CAP values are swept from global memory instead); the element of rank
// n/2 is found by the same most-significant-bit-first radix selection as above, each of the 32 passes counting with all lanes and
// one warp reduction.
constexpr int MED_WARP_CAP = 1024;
template <bool UNDEF>
__global__ void __launch_bounds__(256) fwd_median_warp_kernel(const FwdParams P)
{
    __shared__ unsigned buf[8][MED_WARP_CAP]; // the level's values as ordered keys (NaN: all ones, which no number has)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t o = blockIdx.x * 8ull + warp;
    if (o >= P.nout)
        return;
    const int z0 = blockIdx.y * P.zchunk, z1 = min(P.nz, z0 + P.zchunk);
    const int s0 = P.segStart[o], s1 = P.segStart[o + 1];
    const int len = s1 - s0;
    const bool staged = len <= MED_WARP_CAP;
    const bool has_bad = !isnan(P.bad), has_fill = !isnan(P.fill);
    unsigned* mine = buf[warp];
    auto key_of = [&](const float* src, int k) { // -0.0 and +0.0 are one key (they compare equal in nth_element)
        float f = fwd_load(src + P.segIdx[s0 + k], P.bad, has_bad);
        if (isnan(f))
            return 0xffffffffu;
        if (f == 0.f)
            f = 0.f;
        return ordered_bits(f);
    };
    for (int z = z0; z < z1; ++z) {
        const float* src = P.in + static_cast<size_t>(z) * P.inLayer;
        int n = 0, nans = 0, negz = 0, posz = 0;
        for (int k = lane; k < len; k += 32) {
            const float f = fwd_load(src + P.segIdx[s0 + k], P.bad, has_bad);
            unsigned key = 0xffffffffu;
            if (isnan(f)) {
                ++nans;
            } else {
                ++n;
                if (f == 0.f) {
                    negz += signbit(f) ? 1 : 0;
                    posz += signbit(f) ? 0 : 1;
                }
                key = ordered_bits(f == 0.f ? 0.f : f);
            }
            if (staged)
                mine[k] = key;
        }
#pragma unroll
        for (int sft = 16; sft > 0; sft >>= 1) {
            n += __shfl_xor_sync(0xffffffffu, n, sft);
            nans += __shfl_xor_sync(0xffffffffu, nans, sft);
            negz += __shfl_xor_sync(0xffffffffu, negz, sft);
            posz += __shfl_xor_sync(0xffffffffu, posz, sft);
        }
        __syncwarp();
        float r = nan_f();
        if (n > 0 && !(UNDEF && nans > 0)) { // NaN among the values (undef flavour): unspecified order in the reference, NaN here
            // most-significant-bit-first radix selection of the key of rank n/2; `cand` keys still share the prefix, and once a
            // single one is left its remaining bits need no further passes (about log2(n) passes for distinct values, not 32)
            int rank = n / 2, cand = n;
            unsigned prefix = 0, mask = 0;
            for (int bit = 31; bit >= 0 && cand > 1; --bit) {
                const unsigned m2 = mask | (1u << bit);
                int zeros = 0;
                for (int k = lane; k < len; k += 32) {
                    const unsigned key = staged ? mine[k] : key_of(src, k);
                    zeros += ((key & m2) == prefix) ? 1 : 0; // shares the prefix and has a 0 here (the NaN key never does)
                }
#pragma unroll
                for (int sft = 16; sft > 0; sft >>= 1)
                    zeros += __shfl_xor_sync(0xffffffffu, zeros, sft);
                if (rank >= zeros) {
                    rank -= zeros;
                    prefix |= 1u << bit;
                    cand -= zeros;
                } else {
                    cand = zeros;
                }
                mask = m2;
            }
            if (mask != 0xffffffffu) { // stopped early: fetch the one key that carries the prefix
                unsigned found = 0;
                for (int k = lane; k < len; k += 32) {
                    const unsigned key = staged ? mine[k] : key_of(src, k);
                    if ((key & mask) == prefix && key != 0xffffffffu)
                        found = key;
                }
#pragma unroll
                for (int sft = 16; sft > 0; sft >>= 1)
                    found |= __shfl_xor_sync(0xffffffffu, found, sft); // exactly one lane holds a non-zero key... or the key is 0
                prefix = found;
            }
            r = __uint_as_float((prefix & 0x80000000u) ? (prefix & 0x7fffffffu) : ~prefix);
            if (r == 0.f)
                r = (negz > 0 && posz == 0) ? -0.f : 0.f;
        }
        __syncwarp();
        if (lane == 0) {
            if (has_fill && isnan(r))
                r = P.fill;
            __stcs(P.out + static_cast<size_t>(z) * P.nout + o, r);
        }
    }
}

static int fwd_zchunk(size_t nout, int nz, bool compat)
{
    if (compat)
        return nz; // one computation, nz stores
    const size_t blocks = (nout + 255) / 256;
    const size_t want = static_cast<size_t>(sm_count()) * 16;
    if (blocks >= want || nz <= 1)
        return nz;
    // few target cells: split the levels over blockIdx.y until the machine is full (level blocks of the kernels shrink with it)
    const size_t chunks = std::min<size_t>((want + blocks - 1) / blocks, static_cast<size_t>(nz));
    return static_cast<int>((nz + chunks - 1) / chunks);
}

} // namespace fxcu

using namespace fxcu;

extern "C" {

fimexcu_forward* fimexcu_forward_create(int method, const double* pOnX, const double* pOnY, size_t inX, size_t inY, size_t outX, size_t outY)
{
    clear_error();
    if (!require_device())
        return nullptr;
    if (method < FIMEXCU_INTERPOL_FORWARD_SUM || method > FIMEXCU_INTERPOL_FORWARD_UNDEF_MIN) {
        set_error("unknown forward interpolation method: " + std::to_string(method)); // CachedForwardInterpolation.cc:196
        return nullptr;
    }
    const size_t nin = inX * inY, nout = outX * outY;
    if (nin == 0 || nout == 0 || nin >= 2147483647ull || nout >= 2147483647ull) {
        set_error("fimexcu_forward_create: grid sizes out of range (need 0 < size < 2^31)");
        return nullptr;
    }
    std::unique_ptr<fimexcu_forward> f(new fimexcu_forward());
    f->method = method;
    FXCU_CUDA_NULL(cudaGetDevice(&f->device));
    f->undef = method >= FIMEXCU_INTERPOL_FORWARD_UNDEF_SUM;
    switch (method) {
    case FIMEXCU_INTERPOL_FORWARD_SUM: case FIMEXCU_INTERPOL_FORWARD_UNDEF_SUM: f->kind = 0; break;
    case FIMEXCU_INTERPOL_FORWARD_MEAN: case FIMEXCU_INTERPOL_FORWARD_UNDEF_MEAN: f->kind = 1; break;
    case FIMEXCU_INTERPOL_FORWARD_MEDIAN: case FIMEXCU_INTERPOL_FORWARD_UNDEF_MEDIAN: f->kind = 2; break;
    case FIMEXCU_INTERPOL_FORWARD_MAX: case FIMEXCU_INTERPOL_FORWARD_UNDEF_MAX: f->kind = 3; break;
    default: f->kind = 4; break;
    }
    f->inX = inX;
    f->inY = inY;
    f->outX = outX;
    f->outY = outY;
    cudaStream_t s = nullptr;
    DevBuf dx, dy, key, val, keyS, valS, range, maxLen, tmp;
    if (!dx.alloc(nin * 8, s) || !dy.alloc(nin * 8, s) || !key.alloc(nin * 4, s) || !val.alloc(nin * 4, s) || !keyS.alloc(nin * 4, s) ||
        !valS.alloc(nin * 4, s) || !range.alloc(4 * sizeof(int), s) || !maxLen.alloc(sizeof(int), s))
        return nullptr;
    FXCU_CUDA_NULL(cudaMemcpyAsync(dx.p, pOnX, nin * 8, cudaMemcpyHostToDevice, s));
    FXCU_CUDA_NULL(cudaMemcpyAsync(dy.p, pOnY, nin * 8, cudaMemcpyHostToDevice, s));
    const int init[4] = {0x7fffffff, -1, 0x7fffffff, -1};
    FXCU_CUDA_NULL(cudaMemcpyAsync(range.p, init, sizeof(init), cudaMemcpyHostToDevice, s));
    FXCU_CUDA_NULL(cudaMemsetAsync(maxLen.p, 0, sizeof(int), s));
    fwd_target_kernel<<<blocks_for(nin, 256), 256, 0, s>>>(dx.as<double>(), dy.as<double>(), static_cast<int>(inX), nin, static_cast<int>(outX),
                                                           static_cast<int>(outY), key.as<unsigned>(), val.as<unsigned>(), range.as<int>());
    FXCU_CUDA_NULL(cudaGetLastError());
    // stable LSD radix sort by target cell: equal keys keep their ascending source order
    size_t tmpBytes = 0;
    int bits = 1;
    while ((1ull << bits) <= nout)
        ++bits;
    FXCU_CUDA_NULL(cub::DeviceRadixSort::SortPairs(nullptr, tmpBytes, key.as<unsigned>(), keyS.as<unsigned>(), val.as<unsigned>(), valS.as<unsigned>(),
                                                   static_cast<int>(nin), 0, bits, s));
    if (!tmp.alloc(tmpBytes, s))
        return nullptr;
    FXCU_CUDA_NULL(cub::DeviceRadixSort::SortPairs(tmp.p, tmpBytes, key.as<unsigned>(), keyS.as<unsigned>(), val.as<unsigned>(), valS.as<unsigned>(),
                                                   static_cast<int>(nin), 0, bits, s));
    if (!f->segStart.alloc_sync((nout + 1) * sizeof(int)))
        return nullptr;
    fwd_segments_kernel<<<blocks_for(nout + 1, 256), 256, 0, s>>>(keyS.as<unsigned>(), nin, static_cast<unsigned>(nout), f->segStart.as<int>(),
                                                                  maxLen.as<int>());
    FXCU_CUDA_NULL(cudaGetLastError());
    int r[4], ml = 0, nvalid = 0;
    FXCU_CUDA_NULL(cudaMemcpyAsync(r, range.p, sizeof(r), cudaMemcpyDeviceToHost, s));
    FXCU_CUDA_NULL(cudaMemcpyAsync(&ml, maxLen.p, sizeof(int), cudaMemcpyDeviceToHost, s));
    FXCU_CUDA_NULL(cudaMemcpyAsync(&nvalid, f->segStart.as<int>() + nout, sizeof(int), cudaMemcpyDeviceToHost, s));
    FXCU_CUDA_NULL(cudaStreamSynchronize(s));
    f->maxPointsInIn = static_cast<size_t>(ml);
    f->nvalid = static_cast<size_t>(nvalid);
    // CachedForwardInterpolation.cc:165-186
    size_t minInX = 0, maxInX = 0, minInY = 0, maxInY = 0;
    if (nvalid > 0) {
        minInX = static_cast<size_t>(r[0]);
        maxInX = static_cast<size_t>(r[1]);
        minInY = static_cast<size_t>(r[2]);
        maxInY = static_cast<size_t>(r[3]);
    }
    const size_t EXT = 2;
    minInX = (minInX > EXT) ? minInX - EXT : 0;
    minInY = (minInY > EXT) ? minInY - EXT : 0;
    maxInX = std::min(inX - 1, maxInX + EXT);
    maxInY = std::min(inY - 1, maxInY + EXT);
    int rebX = 0, rebY = 0;
    size_t redX = inX;
    if ((minInX > 0 || minInY > 0 || maxInX < inX - 1 || maxInY < inY - 1) && (minInX + 2 * EXT <= maxInX) && (minInY + 2 * EXT <= maxInY)) {
        redX = maxInX - minInX + 1;
        f->reduced = true;
        f->xMin = minInX;
        f->yMin = minInY;
        f->inX = redX;
        f->inY = maxInY - minInY + 1;
        rebX = static_cast<int>(minInX);
        rebY = static_cast<int>(minInY);
    }
    if (!f->segIdx.alloc_sync(std::max<size_t>(f->nvalid, 1) * sizeof(int)))
        return nullptr;
    if (f->nvalid) {
        fwd_rebase_kernel<<<blocks_for(f->nvalid, 256), 256, 0, s>>>(valS.as<unsigned>(), f->nvalid, static_cast<int>(inX), rebX, rebY,
                                                                     static_cast<int>(redX), f->segIdx.as<int>());
        FXCU_CUDA_NULL(cudaGetLastError());
    }
    FXCU_CUDA_NULL(cudaStreamSynchronize(s));
    return f.release();
}

int fimexcu_forward_info(const fimexcu_forward* f, size_t* inX, size_t* inY, size_t* outX, size_t* outY, int* reduced, size_t* xMin,
                         size_t* yMin, size_t* maxPointsInIn)
{
    if (!f)
        FXCU_FAIL("fimexcu_forward_info: null handle");
    if (inX) *inX = f->inX;
    if (inY) *inY = f->inY;
    if (outX) *outX = f->outX;
    if (outY) *outY = f->outY;
    if (reduced) *reduced = f->reduced ? 1 : 0;
    if (xMin) *xMin = f->xMin;
    if (yMin) *yMin = f->yMin;
    if (maxPointsInIn) *maxPointsInIn = f->maxPointsInIn;
    return FIMEXCU_OK;
}

int fimexcu_forward_apply_dev_f(const fimexcu_forward* f, const float* d_in, size_t nz, float* d_out, float badIn, float fillOut,
                                int level_compat, void* stream)
{
    clear_error();
    if (!f)
        FXCU_FAIL("fimexcu_forward_apply: null handle");
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev != f->device) // the segment tables live in that device's memory
        FXCU_FAIL("fimexcu_forward_apply: the handle was built on device " + std::to_string(f->device) + ", the current device is " +
                  std::to_string(dev) + " (call fimexcu_set_device first)");
    const size_t nout = f->outX * f->outY;
    if (nz == 0 || nout == 0)
        return FIMEXCU_OK;
    if (nz > 0x3fffffffull)
        FXCU_FAIL("fimexcu_forward_apply: too many levels");
    FwdParams P;
    P.segStart = f->segStart.as<int>();
    P.segIdx = f->segIdx.as<int>();
    P.in = d_in;
    P.out = d_out;
    P.nout = nout;
    P.inLayer = f->inX * f->inY;
    P.nz = static_cast<int>(nz);
    P.zchunk = fwd_zchunk(nout, P.nz, level_compat != 0);
    P.kind = f->kind;
    P.undef = f->undef;
    P.level_compat = level_compat != 0;
    P.bad = badIn;
    P.fill = fillOut;
    cudaStream_t st = as_stream(stream);
    if (P.level_compat) {
        // the reference's level-0 quirk: one aggregation, nz stores (fwd_apply_kernel carries the value from level to level)
        const dim3 grid(blocks_for(nout, 256), 1);
        fwd_apply_kernel<<<grid, 256, 0, st>>>(P);
        FXCU_CUDA(cudaGetLastError());
        return FIMEXCU_OK;
    }
    constexpr int ZB = 8;
    const dim3 grid(blocks_for(nout, 256), static_cast<unsigned>((P.nz + P.zchunk - 1) / P.zchunk));
    const double mean_len = static_cast<double>(f->nvalid) / static_cast<double>(nout);
    if (f->kind == 2 && mean_len >= 24.) {
        const dim3 gw(blocks_for(nout, 8), grid.y);
        if (f->undef)
            fwd_median_warp_kernel<true><<<gw, 256, 0, st>>>(P);
        else
            fwd_median_warp_kernel<false><<<gw, 256, 0, st>>>(P);
    } else if (f->kind == 2) {
        const dim3 g2(blocks_for(nout, 128), grid.y);
        if (f->undef)
            fwd_median_kernel<true><<<g2, 128, 0, st>>>(P);
        else
            fwd_median_kernel<false><<<g2, 128, 0, st>>>(P);
    } else if ((f->kind == 3 || f->kind == 4) && !f->undef && mean_len >= 48.) {
        const dim3 gw(blocks_for(nout * 32, 256), grid.y);
        if (f->kind == 3)
            fwd_minmax_warp_kernel<true, 4><<<gw, 256, 0, st>>>(P);
        else
            fwd_minmax_warp_kernel<false, 4><<<gw, 256, 0, st>>>(P);
    } else if (static_cast<size_t>(ZB) * P.inLayer >= (size_t(1) << 31) || static_cast<size_t>(ZB) * nout >= (size_t(1) << 31)) {
        fwd_apply_kernel<<<grid, 256, 0, st>>>(P); // grids beyond the 32-bit offsets of the level-blocked kernel: the general form
    } else {
        // (f->kind: 0 sum, 1 mean, 3 max, 4 min) x undef x "the input has a bad value"
        const bool bad = !std::isnan(P.bad);
        // few target cells: the levels are split over blockIdx.y (fwd_zchunk) and the level block shrinks with the chunk, so that
        // no thread issues gathers for levels it does not own
#define FXCU_FWD_ZB(K, U, B) \
    if (P.zchunk >= 8) \
        fwd_reduce_kernel<K, U, B, 8><<<grid, 256, 0, st>>>(P); \
    else if (P.zchunk >= 4) \
        fwd_reduce_kernel<K, U, B, 4><<<grid, 256, 0, st>>>(P); \
    else \
        fwd_reduce_kernel<K, U, B, 1><<<grid, 256, 0, st>>>(P);
#define FXCU_FWD_CASE(K, U) \
    if (bad) { \
        FXCU_FWD_ZB(K, U, true) \
    } else { \
        FXCU_FWD_ZB(K, U, false) \
    } \
    break;
        switch (f->kind * 2 + (f->undef ? 1 : 0)) {
        case 0: FXCU_FWD_CASE(0, false)
        case 1: FXCU_FWD_CASE(0, true)
        case 2: FXCU_FWD_CASE(1, false)
        case 3: FXCU_FWD_CASE(1, true)
        case 6: FXCU_FWD_CASE(3, false)
        case 7: FXCU_FWD_CASE(3, true)
        case 8: FXCU_FWD_CASE(4, false)
        default: FXCU_FWD_CASE(4, true)
        }
#undef FXCU_FWD_CASE
#undef FXCU_FWD_ZB
    }
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}

int fimexcu_forward_apply_f(const fimexcu_forward* f, const float* in, size_t size, float* out, size_t* newSize, float badIn, float fillOut,
                            int level_compat)
{
    clear_error();
    if (!f)
        FXCU_FAIL("fimexcu_forward_apply_f: null handle");
    if (!require_device())
        return FIMEXCU_ERROR;
    const size_t inLayer = f->inX * f->inY, outLayer = f->outX * f->outY;
    const size_t nz = size / inLayer; // CachedForwardInterpolation.cc:209
    if (newSize)
        *newSize = nz * outLayer;
    if (nz == 0 || outLayer == 0)
        return FIMEXCU_OK;
    DevBuf din, dout;
    if (!din.alloc(nz * inLayer * 4, nullptr) || !dout.alloc(nz * outLayer * 4, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(din.p, in, nz * inLayer * 4, cudaMemcpyHostToDevice, nullptr));
    if (fimexcu_forward_apply_dev_f(f, din.as<float>(), nz, dout.as<float>(), badIn, fillOut, level_compat, nullptr) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(out, dout.p, nz * outLayer * 4, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

void fimexcu_forward_destroy(fimexcu_forward* f) { delete f; }

} // extern "C"
