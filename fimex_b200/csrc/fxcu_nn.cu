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
// Exact ties (duplicated source positions, e.g. wrap columns) are decided by the lower source index for K9 -- the reference's
// choice depends on nanoflann's tree layout -- and by the reference's scan order (larger-or-equal latitude first) for K10.
#include "fxcu_common.h"

#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
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

// getGridDistance (CDMInterpolator.cc:1030-1090): per sample, the largest cosine to any other valid point
__global__ void __launch_bounds__(256) nn_sample_kernel(const double* __restrict__ latr, const double* __restrict__ lonr, size_t n, size_t stepSize,
                                                        double* __restrict__ sample_cos)
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
            const double dlon = __dsub_rn(lon0, lon1);
            const double cd = __dadd_rn(__dmul_rn(__dmul_rn(c0, cos(lat1)), cos(dlon)), __dmul_rn(s0, sin(lat1)));
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
        sample_cos[blockIdx.x] = isnan(best) ? nan("") : red[0];
}

// K10: fastTranslatePointsToClosestInputCell (:1137-1186): largest cos_d strictly above cos(ROI); on exact ties the scan order
// of the reference decides: entries with lat >= p.lat in ascending latitude first, then those below in descending latitude.
__global__ void __launch_bounds__(128) nn_coord_kernel(double* __restrict__ qlon, double* __restrict__ qlat, size_t nq, const double* __restrict__ latr,
                                                       const double* __restrict__ lonr, const unsigned long long* __restrict__ keys,
                                                       const unsigned* __restrict__ vals, size_t n, SphereGrid g, double min_grid_cos_d, int nx)
{
    const size_t i = blockIdx.x * 128ull + threadIdx.x;
    if (i >= nq)
        return;
    const double plat = __dmul_rn(qlat[i], NN_D2R), plon = __dmul_rn(qlon[i], NN_D2R);
    const double cpl = cos(plat), spl = sin(plat);
    const double q0 = __dmul_rn(cpl, cos(plon)), q1 = __dmul_rn(cpl, sin(plon)), q2 = spl;
    double best = min_grid_cos_d, best_lat = 0.;
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
                    bool take = cd > best;
                    if (!take && cd == best && best_idx != 0xffffffffu) {
                        const bool s_up = slat >= plat, b_up = best_lat >= plat;
                        if (s_up != b_up)
                            take = s_up;
                        else if (slat != best_lat)
                            take = s_up ? (slat < best_lat) : (slat > best_lat);
                        else
                            take = s < best_idx;
                    }
                    if (take) {
                        best = cd;
                        best_idx = s;
                        best_lat = slat;
                    }
                }
            }
    }
    if (best_idx != 0xffffffffu) {
        qlon[i] = static_cast<double>(best_idx % nx);
        qlat[i] = static_cast<double>(best_idx / nx);
    } else {
        qlon[i] = -1.;
        qlat[i] = -1.;
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

using namespace fxcu;

extern "C" {

int fimexcu_kdtree_translate(double maxDist, double* pointsOnXAxis, double* pointsOnYAxis, size_t pointsSize, const double* lonVals,
                             const double* latVals, size_t orgXDimSize, size_t orgYDimSize)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    const size_t n = orgXDimSize * orgYDimSize;
    if (n == 0 || n >= 2147483647ull)
        FXCU_FAIL("fimexcu_kdtree_translate: source grid size out of range");
    if (!(maxDist > 0))
        FXCU_FAIL("fimexcu_kdtree_translate: maxDist must be positive"); // assert(maxDist != 0), CDMInterpolator.cc:961
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

int fimexcu_coordnn_translate(double* pointsOnXAxis, double* pointsOnYAxis, size_t pointsSize, const double* lonVals, const double* latVals,
                              size_t orgXDimSize, size_t orgYDimSize)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    const size_t n = orgXDimSize * orgYDimSize;
    if (n == 0 || n >= 2147483647ull)
        FXCU_FAIL("fimexcu_coordnn_translate: source grid size out of range");
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
    DevBuf sc;
    if (!sc.alloc(steps * sizeof(double), nullptr))
        return FIMEXCU_ERROR;
    nn_sample_kernel<<<steps, 256>>>(c0.latr.as<double>(), c0.lonr.as<double>(), n, stepSize, sc.as<double>());
    FXCU_CUDA(cudaGetLastError());
    std::vector<double> samples(steps);
    FXCU_CUDA(cudaMemcpy(samples.data(), sc.p, steps * sizeof(double), cudaMemcpyDeviceToHost));
    double min_cos = 2.;
    bool any = false;
    for (double v : samples)
        if (!std::isnan(v)) {
            any = true;
            min_cos = std::min(min_cos, v);
        }
    if (!any)
        FXCU_FAIL("fimexcu_coordnn_translate: no valid sample point to estimate the grid distance");
    const double PI = 3.1415926535897932384626433832795;
    const double max_grid_d = std::min(std::acos(std::max(-1.0, std::min(1.0, min_cos))) * 1.414, PI); // :1088-1089
    const double min_grid_cos_d = std::cos(max_grid_d);
    if (pointsSize == 0)
        return FIMEXCU_OK;
    // 2. the search, on a grid whose cells are as wide as the largest admissible chord
    Cloud c;
    if (build_cloud(c, lonVals, latVals, n, 2.0 * std::sin(max_grid_d / 2.0)) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    DevBuf qx, qy;
    if (!qx.alloc(pointsSize * 8, nullptr) || !qy.alloc(pointsSize * 8, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(qx.p, pointsOnXAxis, pointsSize * 8, cudaMemcpyHostToDevice, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(qy.p, pointsOnYAxis, pointsSize * 8, cudaMemcpyHostToDevice, nullptr));
    nn_coord_kernel<<<blocks_for(pointsSize, 128), 128>>>(qx.as<double>(), qy.as<double>(), pointsSize, c.latr.as<double>(), c.lonr.as<double>(),
                                                          c.keyS.as<unsigned long long>(), c.valS.as<unsigned>(), n, c.g, min_grid_cos_d,
                                                          static_cast<int>(orgXDimSize));
    FXCU_CUDA(cudaGetLastError());
    FXCU_CUDA(cudaMemcpyAsync(pointsOnXAxis, qx.p, pointsSize * 8, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(pointsOnYAxis, qy.p, pointsSize * 8, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

} // extern "C"
