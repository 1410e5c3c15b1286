// fxcu_nn.cu -- coordinate-based nearest neighbour for curvilinear sources: K9 (coord_kdtree) and K10 (coord_nearestneighbor).
//
// Reference behaviour restated here (fimex 2.4.1):
//   flannTranslatePointsToClosestInputCell        src/CDMInterpolator.cc:953-1028   (nanoflann radiusSearch, first sorted match)
//   getGridDistance                               src/CDMInterpolator.cc:1030-1090
//   fastTranslatePointsToClosestInputCell         src/CDMInterpolator.cc:1112-1190
//   nanoflann L2_Simple_Adaptor / RadiusResultSet include/nanoflann/nanoflann.hpp:150-154, 179-186, 323-333
//
// What the reference computes, stated without its data structure: for every target point the source point with the smallest
// squared chord distance on the unit sphere, strictly inside the search radius (kd-tree), resp. with the largest cosine of the
// great-circle angle, strictly above cos(ROI) (coord_nearestneighbor).  The GPU gets there with a uniform 3-D hash grid over
// the unit sphere instead of a kd-tree: cell edge >= search radius, source points sorted by cell key (radix sort), each query
// scans the 27 cells around it.  Distances use the reference's operation order with separately rounded f64 operations.
// K9 through this file is the nanoflann_compat = 0 form only (true nearest point, NaN points skipped, lower source index on exact
// ties); the reference-identical form walks a clone of nanoflann's tree (fxcu_kdtree.cu).  K10 here is reference-identical: the
// device search is final only where its best candidate is clear of the runner-up and of the ROI threshold by more than the
// difference between the device's and glibc's sin/cos; the remaining targets (near ties, exact ties between duplicated
// positions, targets at the edge of the ROI) are decided by the reference's own latitude-sorted scan on the host (fxcu_llscan.h).
#include "fxcu_common.h"
#include "fxcu_llscan.h"

#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <thread>
#include <vector>

namespace fxcu {

constexpr double NN_D2R = 0.017453292519943296; // deg_to_rad<double>, include/fimex/MathUtils.h:158-162
constexpr double NN_EARTH_RADIUS_M = 6371000.;  // MIFI_EARTH_RADIUS_M

struct SphereGrid {
    int nc;     // cells per dimension
    double inv; // 1 / cell edge
};

__device__ __forceinline__ int grid_coord(double v, const SphereGrid g)
{
    int c = static_cast<int>(floor((v + 1.0) * g.inv));
    return max(0, min(g.nc - 1, c));
}
__device__ __forceinline__ unsigned long long grid_key(int cx, int cy, int cz, const SphereGrid g)
{
    return (static_cast<unsigned long long>(cx) * g.nc + cy) * g.nc + cz;
}

// CDMInterpolator.cc:969-988: lon/lat degrees -> unit-sphere xyz; NaN coordinates stay in the cloud as NaN (never matched)
__global__ void __launch_bounds__(256) nn_cloud_kernel(const double* __restrict__ lon, const double* __restrict__ lat, size_t n, SphereGrid g,
                                                       double* __restrict__ x, double* __restrict__ y, double* __restrict__ z, double* __restrict__ latr,
                                                       double* __restrict__ lonr, unsigned long long* __restrict__ key, unsigned* __restrict__ val)
{
    const size_t i = blockIdx.x * 256ull + threadIdx.x;
    if (i >= n)
        return;
    const double la = lat[i], lo = lon[i];
    unsigned long long k = ~0ull;
    double px = nan(""), py = px, pz = px, lar = px, lor = px;
    if (!(isnan(la) || isnan(lo))) {
        lar = __dmul_rn(la, NN_D2R);
        lor = __dmul_rn(lo, NN_D2R);
        const double sla = sin(lar), cla = cos(lar), slo = sin(lor), clo = cos(lor);
        px = __dmul_rn(cla, clo);
        py = __dmul_rn(cla, slo);
        pz = sla;
        if (isfinite(px) && isfinite(py) && isfinite(pz))
            k = grid_key(grid_coord(px, g), grid_coord(py, g), grid_coord(pz, g), g);
    }
    x[i] = px;
    y[i] = py;
    z[i] = pz;
    latr[i] = lar;
    lonr[i] = lor;
    key[i] = k;
    val[i] = static_cast<unsigned>(i);
}

__device__ __forceinline__ size_t lower_bound_key(const unsigned long long* __restrict__ keys, size_t n, unsigned long long k)
{
    size_t lo = 0, hi = n;
    while (lo < hi) {
        const size_t mid = (lo + hi) >> 1;
        if (keys[mid] < k)
            lo = mid + 1;
        else
            hi = mid;
    }
    return lo;
}

// K9: radiusSearch(q, r2, sorted) -> matches.at(0): smallest squared chord distance with dist < r2 (nanoflann.hpp:150-154)
__global__ void __launch_bounds__(128) nn_kdtree_kernel(double* __restrict__ qlon, double* __restrict__ qlat, size_t nq, const double* __restrict__ x,
                                                        const double* __restrict__ y, const double* __restrict__ z,
                                                        const unsigned long long* __restrict__ keys, const unsigned* __restrict__ vals, size_t n,
                                                        SphereGrid g, double r2, int nx)
{
    const size_t i = blockIdx.x * 128ull + threadIdx.x;
    if (i >= nq)
        return;
    const double lar = __dmul_rn(qlat[i], NN_D2R), lor = __dmul_rn(qlon[i], NN_D2R);
    const double sla = sin(lar), cla = cos(lar), slo = sin(lor), clo = cos(lor);
    const double q0 = __dmul_rn(cla, clo), q1 = __dmul_rn(cla, slo), q2 = sla;
    double best = INFINITY;
    unsigned best_idx = 0xffffffffu;
    if (isfinite(q0) && isfinite(q1) && isfinite(q2)) {
        const int cx = grid_coord(q0, g), cy = grid_coord(q1, g), cz = grid_coord(q2, g);
        for (int ax = max(cx - 1, 0); ax <= min(cx + 1, g.nc - 1); ++ax)
            for (int ay = max(cy - 1, 0); ay <= min(cy + 1, g.nc - 1); ++ay) {
                // the three z-neighbours are consecutive keys
                const unsigned long long k0 = grid_key(ax, ay, max(cz - 1, 0), g), k1 = grid_key(ax, ay, min(cz + 1, g.nc - 1), g);
                for (size_t p = lower_bound_key(keys, n, k0); p < n && keys[p] <= k1; ++p) {
                    const unsigned s = vals[p];
                    const double d0 = __dsub_rn(q0, x[s]), d1 = __dsub_rn(q1, y[s]), d2 = __dsub_rn(q2, z[s]);
                    // L2_Simple_Adaptor::evalMetric: result += diff*diff, dimension by dimension
                    const double d = __dadd_rn(__dadd_rn(__dmul_rn(d0, d0), __dmul_rn(d1, d1)), __dmul_rn(d2, d2));
                    if (d < r2 && (d < best || (d == best && s < best_idx))) {
                        best = d;
                        best_idx = s;
                    }
                }
            }
    }
    if (best_idx != 0xffffffffu) {
        qlon[i] = static_cast<double>(best_idx % nx); // pos = ix + nx*iy (CDMInterpolator.cc:1015-1020)
        qlat[i] = static_cast<double>(best_idx / nx);
    } else {
        qlon[i] = -1000.;
        qlat[i] = -1000.;
    }
}

// getGridDistance (CDMInterpolator.cc:1030-1090): per sample, the largest cosine to any other valid point -- in device
// arithmetic; the second pass lists every point within NOISE of that maximum so that the host can take the maximum of the
// glibc-evaluated cosines over them (llscan::exact_sample_cos), which is the reference's number
constexpr int SAMPLE_CAND_CAP = 256;

__device__ __forceinline__ double sample_cos(double c0, double s0, double lon0, double lat1, double lon1)
{
    const double dlon = __dsub_rn(lon0, lon1);
    return __dadd_rn(__dmul_rn(__dmul_rn(c0, cos(lat1)), cos(dlon)), __dmul_rn(s0, sin(lat1)));
}

__global__ void __launch_bounds__(256) nn_sample_kernel(const double* __restrict__ latr, const double* __restrict__ lonr, size_t n, size_t stepSize,
                                                        double* __restrict__ sample_cos_out)
{
    const size_t samplePos = blockIdx.x * stepSize;
    const double lat0 = latr[samplePos], lon0 = lonr[samplePos];
    double best = -2.;
    if (!(isnan(lat0) || isnan(lon0))) {
        const double c0 = cos(lat0), s0 = sin(lat0);
        for (size_t p = threadIdx.x; p < n; p += 256) {
            if (p == samplePos)
                continue;
            const double lat1 = latr[p], lon1 = lonr[p];
            if (isnan(lat1) || isnan(lon1))
                continue;
            const double cd = sample_cos(c0, s0, lon0, lat1, lon1);
            if (cd > best)
                best = cd;
        }
    } else {
        best = nan(""); // sample skipped (:1053)
    }
    __shared__ double red[256];
    red[threadIdx.x] = best;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) {
            const double o = red[threadIdx.x + s];
            if (o > red[threadIdx.x])
                red[threadIdx.x] = o;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0)
        sample_cos_out[blockIdx.x] = isnan(best) ? nan("") : red[0];
}

__global__ void __launch_bounds__(256) nn_sample_collect_kernel(const double* __restrict__ latr, const double* __restrict__ lonr, size_t n,
                                                                size_t stepSize, const double* __restrict__ sample_max, double noise,
                                                                unsigned* __restrict__ cand, unsigned* __restrict__ ncand)
{
    const size_t samplePos = blockIdx.x * stepSize;
    const double lat0 = latr[samplePos], lon0 = lonr[samplePos];
    const double top = sample_max[blockIdx.x];
    if (isnan(lat0) || isnan(lon0) || isnan(top))
        return;
    const double c0 = cos(lat0), s0 = sin(lat0);
    for (size_t p = threadIdx.x; p < n; p += 256) {
        if (p == samplePos)
            continue;
        const double lat1 = latr[p], lon1 = lonr[p];
        if (isnan(lat1) || isnan(lon1))
            continue;
        if (sample_cos(c0, s0, lon0, lat1, lon1) >= top - noise) {
            const unsigned slot = atomicAdd(&ncand[blockIdx.x], 1u);
            if (slot < SAMPLE_CAND_CAP)
                cand[blockIdx.x * SAMPLE_CAND_CAP + slot] = static_cast<unsigned>(p);
        }
    }
}

// K10: fastTranslatePointsToClosestInputCell (:1137-1186) in device arithmetic: the largest cos_d, its runner-up, and the
// verdict -- a clear winner above the ROI threshold, a clear miss, or "let the host scan decide" (see the file header)
__global__ void __launch_bounds__(128) nn_coord_kernel(double* __restrict__ qlon, double* __restrict__ qlat, size_t nq, const double* __restrict__ latr,
                                                       const double* __restrict__ lonr, const unsigned long long* __restrict__ keys,
                                                       const unsigned* __restrict__ vals, size_t n, SphereGrid g, double min_grid_cos_d, double noise,
                                                       int nx, unsigned* __restrict__ redo, unsigned* __restrict__ nredo)
{
    const size_t i = blockIdx.x * 128ull + threadIdx.x;
    if (i >= nq)
        return;
    const double plat = __dmul_rn(qlat[i], NN_D2R), plon = __dmul_rn(qlon[i], NN_D2R);
    const double cpl = cos(plat), spl = sin(plat);
    const double q0 = __dmul_rn(cpl, cos(plon)), q1 = __dmul_rn(cpl, sin(plon)), q2 = spl;
    double best = -INFINITY, second = -INFINITY;
    unsigned best_idx = 0xffffffffu;
    if (isfinite(q0) && isfinite(q1) && isfinite(q2)) {
        const int cx = grid_coord(q0, g), cy = grid_coord(q1, g), cz = grid_coord(q2, g);
        for (int ax = max(cx - 1, 0); ax <= min(cx + 1, g.nc - 1); ++ax)
            for (int ay = max(cy - 1, 0); ay <= min(cy + 1, g.nc - 1); ++ay) {
                const unsigned long long k0 = grid_key(ax, ay, max(cz - 1, 0), g), k1 = grid_key(ax, ay, min(cz + 1, g.nc - 1), g);
                for (size_t p = lower_bound_key(keys, n, k0); p < n && keys[p] <= k1; ++p) {
                    const unsigned s = vals[p];
                    const double slat = latr[s];
                    const double dlon = __dsub_rn(lonr[s], plon);
                    const double cd = __dadd_rn(__dmul_rn(__dmul_rn(cos(slat), cpl), cos(dlon)), __dmul_rn(sin(slat), spl));
                    if (cd > best) {
                        second = best;
                        best = cd;
                        best_idx = s;
                    } else if (cd > second) {
                        second = cd;
                    }
                }
            }
    }
    if (best_idx == 0xffffffffu || best < min_grid_cos_d - noise) {
        qlon[i] = -1.;
        qlat[i] = -1.;
    } else if (best > min_grid_cos_d + noise && best - second > noise) {
        qlon[i] = static_cast<double>(best_idx % nx);
        qlat[i] = static_cast<double>(best_idx / nx);
    } else {
        redo[atomicAdd(nredo, 1u)] = static_cast<unsigned>(i); // qlon/qlat keep the target's coordinates for the host
    }
}

struct Cloud {
    DevBuf x, y, z, latr, lonr, key, val, keyS, valS;
    SphereGrid g;
    size_t n = 0;
};

// chord: the largest chord length a match can have; the cell edge is at least that, so 27 cells cover every candidate
static int build_cloud(Cloud& c, const double* lon_host, const double* lat_host, size_t n, double chord)
{
    c.n = n;
    double h = std::max(chord, 2.0 / 1048575.0) * (1.0 + 1e-9);
    int nc = static_cast<int>(std::floor(2.0 / h)) + 1;
    nc = std::max(1, std::min(nc, 1 << 20));
    c.g.nc = nc;
    c.g.inv = 1.0 / h;
    DevBuf dlon, dlat, tmp;
    if (!dlon.alloc(n * 8, nullptr) || !dlat.alloc(n * 8, nullptr) || !c.x.alloc(n * 8, nullptr) || !c.y.alloc(n * 8, nullptr) || !c.z.alloc(n * 8, nullptr) ||
        !c.latr.alloc(n * 8, nullptr) || !c.lonr.alloc(n * 8, nullptr) || !c.key.alloc(n * 8, nullptr) || !c.val.alloc(n * 4, nullptr) ||
        !c.keyS.alloc(n * 8, nullptr) || !c.valS.alloc(n * 4, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(dlon.p, lon_host, n * 8, cudaMemcpyHostToDevice, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(dlat.p, lat_host, n * 8, cudaMemcpyHostToDevice, nullptr));
    nn_cloud_kernel<<<blocks_for(n, 256), 256>>>(dlon.as<double>(), dlat.as<double>(), n, c.g, c.x.as<double>(), c.y.as<double>(), c.z.as<double>(),
                                                 c.latr.as<double>(), c.lonr.as<double>(), c.key.as<unsigned long long>(), c.val.as<unsigned>());
    FXCU_CUDA(cudaGetLastError());
    size_t tb = 0;
    FXCU_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, c.key.as<unsigned long long>(), c.keyS.as<unsigned long long>(), c.val.as<unsigned>(),
                                              c.valS.as<unsigned>(), static_cast<int>(n)));
    if (!tmp.alloc(tb, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, tb, c.key.as<unsigned long long>(), c.keyS.as<unsigned long long>(), c.val.as<unsigned>(),
                                              c.valS.as<unsigned>(), static_cast<int>(n)));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

} // namespace fxcu

namespace fxcu {

// K9 without the tree (nanoflann_compat = 0 of fimexcu_kdtree_translate_mode; argument checks are the caller's)
int kdtree_translate_grid(double maxDist, double* pointsOnXAxis, double* pointsOnYAxis, size_t pointsSize, const double* lonVals, const double* latVals,
                          size_t orgXDimSize, size_t orgYDimSize)
{
    const size_t n = orgXDimSize * orgYDimSize;
    if (pointsSize == 0)
        return FIMEXCU_OK;
    const double r = maxDist / NN_EARTH_RADIUS_M; // all calculations on the unit sphere (:964)
    const double r2 = r * r;                      // squared, "since distance is not sqrt" (:998)
    Cloud c;
    if (build_cloud(c, lonVals, latVals, n, r) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    DevBuf qx, qy;
    if (!qx.alloc(pointsSize * 8, nullptr) || !qy.alloc(pointsSize * 8, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(qx.p, pointsOnXAxis, pointsSize * 8, cudaMemcpyHostToDevice, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(qy.p, pointsOnYAxis, pointsSize * 8, cudaMemcpyHostToDevice, nullptr));
    nn_kdtree_kernel<<<blocks_for(pointsSize, 128), 128>>>(qx.as<double>(), qy.as<double>(), pointsSize, c.x.as<double>(), c.y.as<double>(),
                                                           c.z.as<double>(), c.keyS.as<unsigned long long>(), c.valS.as<unsigned>(), n, c.g, r2,
                                                           static_cast<int>(orgXDimSize));
    FXCU_CUDA(cudaGetLastError());
    FXCU_CUDA(cudaMemcpyAsync(pointsOnXAxis, qx.p, pointsSize * 8, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(pointsOnYAxis, qy.p, pointsSize * 8, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

} // namespace fxcu

using namespace fxcu;

extern "C" {

int fimexcu_coordnn_translate(double* pointsOnXAxis, double* pointsOnYAxis, size_t pointsSize, const double* lonVals, const double* latVals,
                              size_t orgXDimSize, size_t orgYDimSize)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    const size_t n = orgXDimSize * orgYDimSize;
    if (n == 0 || n >= 2147483647ull)
        FXCU_FAIL("fimexcu_coordnn_translate: source grid size out of range");
    if (pointsSize >= 0xffffffffull)
        FXCU_FAIL("fimexcu_coordnn_translate: too many target points");
    try {
        // the reference's latitude-sorted list (std::sort + glibc sin/cos) is prepared on the host cores while the GPU searches
        llscan::SortedCloud sorted;
        struct Joiner {
            std::thread t;
            ~Joiner()
            {
                if (t.joinable())
                    t.join();
            }
        } sorter{std::thread([&]() { llscan::build(sorted, lonVals, latVals, orgXDimSize, orgYDimSize); })};

        // 1. ROI of the input data (getGridDistance): the cloud for that step needs no grid yet (chord 2 -> one cell)
        Cloud c0;
        if (build_cloud(c0, lonVals, latVals, n, 2.0) != FIMEXCU_OK)
            return FIMEXCU_ERROR;
        int steps;
        size_t stepSize;
        if (n > 1000) {
            steps = 53; // "unusual grid-dimension" (:1040)
            stepSize = n / steps;
        } else {
            stepSize = 1;
            steps = static_cast<int>(n);
        }
        DevBuf sc, cand, ncand;
        if (!sc.alloc(steps * sizeof(double), nullptr) || !cand.alloc(size_t(steps) * SAMPLE_CAND_CAP * 4, nullptr) || !ncand.alloc(size_t(steps) * 4, nullptr))
            return FIMEXCU_ERROR;
        FXCU_CUDA(cudaMemsetAsync(ncand.p, 0, size_t(steps) * 4, nullptr));
        nn_sample_kernel<<<steps, 256>>>(c0.latr.as<double>(), c0.lonr.as<double>(), n, stepSize, sc.as<double>());
        FXCU_CUDA(cudaGetLastError());
        nn_sample_collect_kernel<<<steps, 256>>>(c0.latr.as<double>(), c0.lonr.as<double>(), n, stepSize, sc.as<double>(), llscan::NOISE,
                                                 cand.as<unsigned>(), ncand.as<unsigned>());
        FXCU_CUDA(cudaGetLastError());
        std::vector<double> devMax(steps);
        std::vector<unsigned> hCand(size_t(steps) * SAMPLE_CAND_CAP), hCount(steps);
        FXCU_CUDA(cudaMemcpy(devMax.data(), sc.p, steps * sizeof(double), cudaMemcpyDeviceToHost));
        FXCU_CUDA(cudaMemcpy(hCand.data(), cand.p, hCand.size() * 4, cudaMemcpyDeviceToHost));
        FXCU_CUDA(cudaMemcpy(hCount.data(), ncand.p, size_t(steps) * 4, cudaMemcpyDeviceToHost));
        // the samples in glibc arithmetic: largest cosine over the listed candidates (or over everything, should a sample sit in
        // a heap of more than SAMPLE_CAND_CAP near-equidistant points)
        double min_cos = 0.;
        bool any = false;
        for (int ik = 0; ik < steps; ++ik) {
            const size_t samplePos = ik * stepSize;
            if (std::isnan(lonVals[samplePos] * llscan::D2R) || std::isnan(latVals[samplePos] * llscan::D2R))
                continue; // :1053
            double v;
            if (hCount[ik] <= SAMPLE_CAND_CAP) {
                v = llscan::exact_sample_cos(latVals[samplePos], lonVals[samplePos], lonVals, latVals, &hCand[size_t(ik) * SAMPLE_CAND_CAP], hCount[ik]);
            } else {
                std::vector<unsigned> all;
                all.reserve(n);
                for (size_t ix = 0; ix < orgXDimSize; ++ix)
                    for (size_t iy = 0; iy < orgYDimSize; ++iy) {
                        const size_t pos = ix + iy * orgXDimSize;
                        if (pos != samplePos && !(std::isnan(lonVals[pos]) || std::isnan(latVals[pos])))
                            all.push_back(static_cast<unsigned>(pos));
                    }
                v = llscan::exact_sample_cos(latVals[samplePos], lonVals[samplePos], lonVals, latVals, all.data(), all.size());
            }
            if (!any || v < min_cos)
                min_cos = v;
            any = true;
        }
        if (!any)
            FXCU_FAIL("fimexcu_coordnn_translate: no valid sample point to estimate the grid distance"); // the reference dereferences end() here
        const double MIFI_PI_D = 3.1415926535897932384626433832795;
        const double grid_d = std::acos(min_cos) * 1.414;                                  // :1088
        const double max_grid_d = (MIFI_PI_D < grid_d) ? MIFI_PI_D : grid_d;              // std::min(max_grid_d, MIFI_PI), :1089
        const double min_grid_cos_d = std::cos(max_grid_d);
        if (pointsSize == 0)
            return FIMEXCU_OK;
        if (std::isnan(min_grid_cos_d)) {
            // a lone source point (or coincident sample pairs whose cosine rounds above 1) gives acos() = NaN in the reference:
            // no cosine is greater than NaN, every target stays at its initial (-1, -1)
            for (size_t i = 0; i < pointsSize; ++i)
                pointsOnXAxis[i] = pointsOnYAxis[i] = -1.;
            return FIMEXCU_OK;
        }
        // 2. the search, on a grid whose cells are as wide as the largest admissible chord
        Cloud c;
        if (build_cloud(c, lonVals, latVals, n, 2.0 * std::sin(max_grid_d / 2.0)) != FIMEXCU_OK)
            return FIMEXCU_ERROR;
        DevBuf qx, qy, redo, nredo;
        if (!qx.alloc(pointsSize * 8, nullptr) || !qy.alloc(pointsSize * 8, nullptr) || !redo.alloc(pointsSize * 4, nullptr) || !nredo.alloc(4, nullptr))
            return FIMEXCU_ERROR;
        FXCU_CUDA(cudaMemcpyAsync(qx.p, pointsOnXAxis, pointsSize * 8, cudaMemcpyHostToDevice, nullptr));
        FXCU_CUDA(cudaMemcpyAsync(qy.p, pointsOnYAxis, pointsSize * 8, cudaMemcpyHostToDevice, nullptr));
        FXCU_CUDA(cudaMemsetAsync(nredo.p, 0, 4, nullptr));
        nn_coord_kernel<<<blocks_for(pointsSize, 128), 128>>>(qx.as<double>(), qy.as<double>(), pointsSize, c.latr.as<double>(), c.lonr.as<double>(),
                                                              c.keyS.as<unsigned long long>(), c.valS.as<unsigned>(), n, c.g, min_grid_cos_d,
                                                              llscan::NOISE, static_cast<int>(orgXDimSize), redo.as<unsigned>(), nredo.as<unsigned>());
        FXCU_CUDA(cudaGetLastError());
        unsigned hRedo = 0;
        FXCU_CUDA(cudaMemcpyAsync(&hRedo, nredo.p, 4, cudaMemcpyDeviceToHost, nullptr));
        FXCU_CUDA(cudaMemcpyAsync(pointsOnXAxis, qx.p, pointsSize * 8, cudaMemcpyDeviceToHost, nullptr));
        FXCU_CUDA(cudaMemcpyAsync(pointsOnYAxis, qy.p, pointsSize * 8, cudaMemcpyDeviceToHost, nullptr));
        FXCU_CUDA(cudaStreamSynchronize(nullptr));
        sorter.t.join();
        if (hRedo) {
            std::vector<unsigned> list(hRedo);
            FXCU_CUDA(cudaMemcpy(list.data(), redo.p, size_t(hRedo) * 4, cudaMemcpyDeviceToHost));
            parallel_ranges(hRedo, 16, [&](size_t a, size_t b) {
                for (size_t k = a; k < b; ++k) {
                    const size_t i = list[k]; // pointsOn?Axis[i] still hold the target's lon / lat
                    const long pos = llscan::scan(sorted, pointsOnXAxis[i], pointsOnYAxis[i], min_grid_cos_d);
                    if (pos >= 0) {
                        pointsOnXAxis[i] = static_cast<double>(size_t(pos) % orgXDimSize);
                        pointsOnYAxis[i] = static_cast<double>(size_t(pos) / orgXDimSize);
                    } else {
                        pointsOnXAxis[i] = pointsOnYAxis[i] = -1.;
                    }
                }
            });
        }
        return FIMEXCU_OK;
    } catch (const std::exception& e) {
        FXCU_FAIL(std::string("fimexcu_coordnn_translate: ") + e.what());
    }
}

} // extern "C"
