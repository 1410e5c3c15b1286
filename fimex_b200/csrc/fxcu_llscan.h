// fxcu_llscan.h -- host side of K10 (coord_nearestneighbor): the reference's latitude-sorted scan, used to decide the targets
// the GPU search cannot decide from device arithmetic alone.
//
// Reference: fastTranslatePointsToClosestInputCell + getGridDistance, src/CDMInterpolator.cc:1030-1190.  Its answer for a target
// is the first point, in ITS scan order, whose cos(great-circle angle) is strictly greater than everything before it, where
//   * every cosine is cos(lat_i)*cos(lat_p)*cos(lon_i - lon_p) + sin(lat_i)*sin(lat_p) evaluated with glibc's sin/cos;
//   * the scan order is: the latitude-sorted list from lower_bound(lat_p) upwards, then from there downwards, each direction cut
//     off at the first entry with |dlat| > acos(best so far);
//   * the list is sorted with std::sort on the latitude alone -- not a stable sort, so the order of points with EQUAL latitude
//     (duplicated wrap columns) is whatever libstdc++'s introsort leaves, given the insertion order ix-outer / iy-inner (:1129-1136).
// The GPU search (fxcu_nn.cu) uses the device's sin/cos, whose last bit differs from glibc's: it is final only where the best
// candidate is clear of the runner-up and of the ROI threshold by more than that noise.  Everything else comes here and is
// decided by the scan below, on the same sorted list (the real std::sort of this libstdc++), with the same libm calls.
// The only liberty: cos(lon_i - lon_p) is not evaluated for an entry whose unit-vector dot product with the target is below the
// best so far by more than 1e-13 (2^10 times the rounding error of either formula) -- such an entry cannot change the state.
//
// Host-only C++.
#pragma once

#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <vector>

#include "fxcu_hostpar.h"

namespace fxcu {
namespace llscan {

constexpr double D2R = 0.017453292519943296; // deg_to_rad, include/fimex/MathUtils.h:158-162
constexpr double NOISE = 1e-13;              // >> |device cos_d - glibc cos_d| (a few 1e-16) and << any real gap

struct Key {
    double lat; // radians
    uint32_t src;
    bool operator<(const Key& rhs) const { return lat < rhs.lat; } // LL_POINT::operator<, :1105
};

struct SortedCloud {
    std::vector<Key> keys;                        // after std::sort
    std::vector<double> lon, coslat, sinlat, x, y; // aligned with keys; (x, y, sinlat) is the unit vector
    size_t nx = 1;
};

// :1127-1138
inline void build(SortedCloud& c, const double* lonDeg, const double* latDeg, size_t nx, size_t ny)
{
    c.nx = nx;
    c.keys.clear();
    c.keys.reserve(nx * ny);
    for (size_t ix = 0; ix < nx; ++ix)
        for (size_t iy = 0; iy < ny; ++iy) {
            const size_t pos = ix + iy * nx;
            if (!(std::isnan(lonDeg[pos]) || std::isnan(latDeg[pos])))
                c.keys.push_back(Key{latDeg[pos] * D2R, static_cast<uint32_t>(pos)});
        }
    std::sort(c.keys.begin(), c.keys.end());
    const size_t m = c.keys.size();
    c.lon.resize(m);
    c.coslat.resize(m);
    c.sinlat.resize(m);
    c.x.resize(m);
    c.y.resize(m);
    parallel_ranges(m, 1 << 15, [&](size_t a, size_t b) {
        for (size_t i = a; i < b; ++i) {
            const double lo = lonDeg[c.keys[i].src] * D2R;
            c.lon[i] = lo;
            c.coslat[i] = std::cos(c.keys[i].lat);
            c.sinlat[i] = std::sin(c.keys[i].lat);
            c.x[i] = c.coslat[i] * std::cos(lo);
            c.y[i] = c.coslat[i] * std::sin(lo);
        }
    });
}

// one target, :1143-1186; returns the source index or -1
inline long scan(const SortedCloud& c, double lonDeg, double latDeg, double min_grid_cos_d)
{
    const double plat = latDeg * D2R, plon = lonDeg * D2R;
    const double cpl = std::cos(plat), spl = std::sin(plat);
    const double qx = cpl * std::cos(plon), qy = cpl * std::sin(plon);
    double min_cos_d = min_grid_cos_d;
    double min_d = std::acos(min_cos_d);
    long found = -1;
    const size_t m = c.keys.size();
    const size_t start = static_cast<size_t>(std::lower_bound(c.keys.begin(), c.keys.end(), Key{plat, 0}) - c.keys.begin());
    auto visit = [&](size_t i) -> bool { // false: this direction is finished
        if (std::fabs(c.keys[i].lat - plat) > min_d)
            return false;
        const double dot = c.x[i] * qx + c.y[i] * qy + c.sinlat[i] * spl;
        if (!(dot < min_cos_d - NOISE)) {
            const double dlon = c.lon[i] - plon;
            const double cos_d = c.coslat[i] * cpl * std::cos(dlon) + c.sinlat[i] * spl;
            if (cos_d > min_cos_d) {
                min_cos_d = cos_d;
                min_d = std::acos(min_cos_d);
                found = static_cast<long>(c.keys[i].src);
            }
        }
        return true;
    };
    for (size_t i = start; i < m; ++i)
        if (!visit(i))
            break;
    for (size_t i = start; i-- > 0;)
        if (!visit(i))
            break;
    return found;
}

// getGridDistance, :1030-1090, given for every sample the candidates that can hold its largest cosine (source indices; the GPU
// pass keeps everything within NOISE of the device maximum).  sampleLat/sampleLon in degrees.  Returns max_grid_d (radians).
inline double exact_sample_cos(double lat0Deg, double lon0Deg, const double* lonDeg, const double* latDeg, const uint32_t* cand, size_t ncand)
{
    const double lon0 = lon0Deg * D2R, lat0 = lat0Deg * D2R;
    double best = -2;
    for (size_t k = 0; k < ncand; ++k) {
        const double lon1 = lonDeg[cand[k]] * D2R, lat1 = latDeg[cand[k]] * D2R;
        const double dlon = lon0 - lon1;
        const double cos_d = std::cos(lat0) * std::cos(lat1) * std::cos(dlon) + std::sin(lat0) * std::sin(lat1);
        if (cos_d > best)
            best = cos_d;
    }
    return best;
}

} // namespace llscan
} // namespace fxcu
