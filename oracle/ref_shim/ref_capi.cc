// C entry points over the UNMODIFIED reference objects compiled by oracle/Makefile from
// /root/reference (CachedInterpolation.cc, CachedForwardInterpolation.cc, interpolation.c, ...).
// TEST INFRASTRUCTURE: used to pin oracle/fimex_oracle.c and as the "reference" CPU baseline.
// The mifi_* C functions of src/interpolation.c are exported by the same library as they are.
#include "CachedForwardInterpolation.h"
#include "fimex/CachedInterpolation.h"
#include "fimex/DataUtils.h"
#include "fimex/FindNeighborElements.h"
#include "fimex/MathUtils.h"
#include "fimex/SharedArray.h"
#include "fimex/interpolation.h"
#include "fimex/reproject.h"
#include "leap_iterator.h"
#include "nanoflann/nanoflann.hpp"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <memory>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace MetNoFimex;

namespace {
struct RefCache {
    CachedInterpolationInterface_p ci;
};

shared_array<double> clone_table(const double* p, size_t n)
{
    auto a = make_shared_array<double>(n ? n : 1);
    std::memcpy(a.get(), p, n * sizeof(double));
    return a;
}

shared_array<float> borrow(const float* p)
{
    // interpolateValues() only reads its input; a non-owning handle avoids timing a copy
    return shared_array<float>(const_cast<float*>(p), [](float*) {});
}
} // namespace

// The call sequence of flannTranslatePointsToClosestInputCell (CDMInterpolator.cc:953-1028) on the
// reference's vendored nanoflann (that function sits in an anonymous namespace of a translation unit
// that cannot be linked here, so the few lines that drive nanoflann are re-stated; the kd-tree, its
// traversal order and the result sort are the reference's own code).
namespace {
struct Cloud {
    struct P { double x, y, z; };
    std::vector<P> pts;
    size_t kdtree_get_point_count() const { return pts.size(); }
    double kdtree_distance(const double* p1, const size_t idx, size_t) const
    {
        const double d0 = p1[0] - pts[idx].x, d1 = p1[1] - pts[idx].y, d2 = p1[2] - pts[idx].z;
        return d0 * d0 + d1 * d1 + d2 * d2;
    }
    double kdtree_get_pt(const size_t idx, int dim) const { return dim == 0 ? pts[idx].x : (dim == 1 ? pts[idx].y : pts[idx].z); }
    template <class BBOX> bool kdtree_get_bbox(BBOX&) const { return false; }
};
} // namespace

// generic over the CDM integer types; dtype as enum fimexcu_dtype (CDM_CHAR is `char`, include/fimex/CDMDataType.h:34-46)
namespace {
template <class T> void ref_unpack_t(const void* in, size_t n, double bad, float* out)
{
    const T* p = static_cast<const T*>(in);
    std::transform(p, p + n, out, MetNoFimex::data_caster<float, T>());
    mifi_bad2nanf(out, out + n, bad);
}
template <class T> void ref_unpack_scaled_t(const void* in, size_t n, double fill, double scale, double offset, float* out)
{
    const T* p = static_cast<const T*>(in);
    MetNoFimex::ScaleValue<T, double> sv(fill, scale, offset, MIFI_UNDEFINED_D, 1., 0.);
    for (size_t i = 0; i < n; ++i)
        out[i] = MetNoFimex::data_caster<float, double>()(sv(p[i]));
}
template <class T> void ref_pack_t(const float* in, size_t n, double bad, void* out)
{
    MetNoFimex::ScaleValue<float, T> sv(MIFI_UNDEFINED_F, 1., 0., bad, 1., 0.);
    std::transform(in, in + n, static_cast<T*>(out), sv);
}
} // namespace


extern "C" {

int ref_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void ref_set_num_threads(int n)
{
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}

// createCachedInterpolation (CachedInterpolation.cc:132-147) or the forward cache ctor for method >= FORWARD_SUM
void* ref_cache_create(int method, const double* px, const double* py, size_t inX, size_t inY, size_t outX, size_t outY)
{
    try {
        auto c = std::make_unique<RefCache>();
        if (method >= MIFI_INTERPOL_FORWARD_SUM) {
            c->ci = std::make_shared<CachedForwardInterpolation>("x", "y", method, clone_table(px, inX * inY), clone_table(py, inX * inY), inX, inY, outX, outY);
        } else {
            c->ci = createCachedInterpolation("x", "y", method, clone_table(px, outX * outY), clone_table(py, outX * outY), inX, inY, outX, outY);
        }
        return c.release();
    } catch (...) {
        return nullptr;
    }
}

void ref_cache_info(void* h, size_t* inX, size_t* inY, size_t* outX, size_t* outY, int* reduced, size_t* xMin, size_t* yMin)
{
    auto* c = static_cast<RefCache*>(h);
    *inX = c->ci->getInX();
    *inY = c->ci->getInY();
    *outX = c->ci->getOutX();
    *outY = c->ci->getOutY();
    auto rd = c->ci->reducedDomain();
    *reduced = rd ? 1 : 0;
    *xMin = rd ? rd->xMin : 0;
    *yMin = rd ? rd->yMin : 0;
}

int ref_cache_apply(void* h, const float* in, size_t size, float* out)
{
    auto* c = static_cast<RefCache*>(h);
    try {
        size_t newSize = 0;
        auto r = c->ci->interpolateValues(borrow(in), size, newSize);
        std::memcpy(out, r.get(), newSize * sizeof(float));
        return 0;
    } catch (...) {
        return -1;
    }
}

// time `reps` calls of interpolateValues (allocation of the result included, as in the reference; no copy-out)
double ref_cache_apply_seconds(void* h, const float* in, size_t size, int reps)
{
    auto* c = static_cast<RefCache*>(h);
    auto input = borrow(in);
    size_t newSize = 0;
    const auto t0 = std::chrono::steady_clock::now();
    for (int r = 0; r < reps; ++r) {
        auto res = c->ci->interpolateValues(input, size, newSize);
        if (!res)
            return -1;
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

void ref_cache_destroy(void* h) { delete static_cast<RefCache*>(h); }

size_t ref_round_and_clamp(double d, size_t mini, size_t maxi, size_t invalid) { return RoundAndClamp(mini, maxi, invalid)(d); }

// FindNeighborElements.h instantiated exactly as CDMVerticalInterpolator.cc:529-533 does
void ref_find_closest_neighbor_distinct_elements(const float* col, size_t n, long stride, double x, long* first, long* second)
{
    const leap_iterator<const float*> b(col, stride);
    const leap_iterator<const float*> e = b + n;
    const auto pos = find_closest_neighbor_distinct_elements(b, e, x);
    *first = pos.first;
    *second = pos.second;
}

void ref_find_closest_distinct_elements(const float* col, size_t n, long stride, double x, long* first, long* second)
{
    const leap_iterator<const float*> b(col, stride);
    const leap_iterator<const float*> e = b + n;
    const auto pos = find_closest_distinct_elements(b, e, x);
    *first = pos.first;
    *second = pos.second;
}

int ref_kdtree_translate(double maxDist_m, double* pointsOnX_deg, double* pointsOnY_deg, size_t nPoints, const double* lon_deg, const double* lat_deg,
                         size_t nx, size_t ny)
{
    using namespace nanoflann;
    double maxDist = maxDist_m / 6371000.;
    Cloud cloud;
    cloud.pts.resize(nx * ny);
    for (size_t pos = 0; pos < nx * ny; ++pos) {
        if (!(std::isnan(lat_deg[pos]) || std::isnan(lon_deg[pos]))) {
            const double la = deg_to_rad(lat_deg[pos]), lo = deg_to_rad(lon_deg[pos]);
            cloud.pts[pos] = {std::cos(la) * std::cos(lo), std::cos(la) * std::sin(lo), std::sin(la)};
        } else {
            const double nan = std::numeric_limits<double>::quiet_NaN();
            cloud.pts[pos] = {nan, nan, nan};
        }
    }
    typedef KDTreeSingleIndexAdaptor<L2_Simple_Adaptor<double, Cloud>, Cloud, 3> tree_t;
    tree_t index(3, cloud, KDTreeSingleIndexAdaptorParams(12));
    index.buildIndex();
    const double r2 = maxDist * maxDist;
    SearchParams params;
    params.sorted = true;
    for (size_t i = 0; i < nPoints; ++i) {
        const double la = deg_to_rad(pointsOnY_deg[i]), lo = deg_to_rad(pointsOnX_deg[i]);
        const double q[3] = {std::cos(la) * std::cos(lo), std::cos(la) * std::sin(lo), std::sin(la)};
        std::vector<std::pair<size_t, double>> matches;
        const size_t n = index.radiusSearch(q, r2, matches, params);
        if (n > 0) {
            const size_t pos = matches.at(0).first;
            pointsOnX_deg[i] = pos % nx;
            pointsOnY_deg[i] = pos / nx;
        } else {
            pointsOnX_deg[i] = -1000;
            pointsOnY_deg[i] = -1000;
        }
    }
    return 0;
}

// coord_nearestneighbor: fastTranslatePointsToClosestInputCell + getGridDistance (CDMInterpolator.cc:1030-1190) also sit in the
// anonymous namespace of the translation unit that cannot be linked here.  Restated on the SAME library calls the reference
// makes -- std::sort on a 4-double record compared by latitude only (not stable: the order of equal latitudes is libstdc++'s
// introsort's), std::lower_bound, std::min_element, glibc sin/cos/acos -- so that the tie order is the reference's.
namespace {
struct LatLonRecord {
    double lat, lon, x, y; // radians, radians, source ix, source iy
    bool operator<(const LatLonRecord& o) const { return lat < o.lat; }
};
} // namespace

double ref_grid_distance(const double* lon_deg, const double* lat_deg, size_t nx, size_t ny)
{
    std::vector<double> samples;
    int steps;
    size_t stepSize;
    if (nx * ny > 1000) {
        steps = 53;
        stepSize = nx * ny / steps;
    } else {
        stepSize = 1;
        steps = static_cast<int>(nx * ny);
    }
    for (int ik = 0; ik < steps; ++ik) {
        const size_t samplePos = ik * stepSize;
        const double lon0 = deg_to_rad(lon_deg[samplePos]), lat0 = deg_to_rad(lat_deg[samplePos]);
        if (std::isnan(lon0) || std::isnan(lat0))
            continue;
        double largest = -2;
        for (size_t ix = 0; ix < nx; ++ix)
            for (size_t iy = 0; iy < ny; ++iy) {
                const size_t pos = ix + iy * nx;
                if (pos == samplePos)
                    continue;
                const double lon1 = deg_to_rad(lon_deg[pos]), lat1 = deg_to_rad(lat_deg[pos]);
                if (std::isnan(lon1) || std::isnan(lat1))
                    continue;
                const double dlon = lon0 - lon1;
                const double cos_d = cos(lat0) * cos(lat1) * cos(dlon) + sin(lat0) * sin(lat1);
                if (cos_d > largest)
                    largest = cos_d;
            }
        samples.push_back(largest);
    }
    const double d = acos(*std::min_element(samples.begin(), samples.end())) * 1.414;
    return std::min(d, MIFI_PI);
}

int ref_coordnn_translate(double* pointsOnX_deg, double* pointsOnY_deg, size_t nPoints, const double* lon_deg, const double* lat_deg, size_t nx, size_t ny)
{
    const double roi = ref_grid_distance(lon_deg, lat_deg, nx, ny);
    const double roi_cos = cos(roi);
    std::vector<LatLonRecord> recs;
    recs.reserve(nx * ny);
    for (size_t ix = 0; ix < nx; ++ix)
        for (size_t iy = 0; iy < ny; ++iy) {
            const size_t pos = ix + iy * nx;
            if (!(std::isnan(lon_deg[pos]) || std::isnan(lat_deg[pos])))
                recs.push_back(LatLonRecord{deg_to_rad(lat_deg[pos]), deg_to_rad(lon_deg[pos]), double(ix), double(iy)});
        }
    std::sort(recs.begin(), recs.end());
#pragma omp parallel for
    for (long i = 0; i < static_cast<long>(nPoints); ++i) {
        LatLonRecord p{deg_to_rad(pointsOnY_deg[i]), deg_to_rad(pointsOnX_deg[i]), -1., -1.};
        double best_cos = roi_cos, best_d = acos(best_cos);
        auto up = std::lower_bound(recs.begin(), recs.end(), p);
        auto down = up;
        while (up != recs.end()) {
            if (fabs(up->lat - p.lat) > best_d)
                break;
            const double dlon = up->lon - p.lon;
            const double cos_d = cos(up->lat) * cos(p.lat) * cos(dlon) + sin(up->lat) * sin(p.lat);
            if (cos_d > best_cos) {
                best_cos = cos_d;
                best_d = acos(best_cos);
                p.x = up->x;
                p.y = up->y;
            }
            ++up;
        }
        while (down != recs.begin()) {
            --down;
            if (fabs(down->lat - p.lat) > best_d)
                break;
            const double dlon = down->lon - p.lon;
            const double cos_d = cos(down->lat) * cos(p.lat) * cos(dlon) + sin(down->lat) * sin(p.lat);
            if (cos_d > best_cos) {
                best_cos = cos_d;
                best_d = acos(best_cos);
                p.x = down->x;
                p.y = down->y;
            }
        }
        pointsOnY_deg[i] = p.y;
        pointsOnX_deg[i] = p.x;
    }
    return 0;
}

// ---- src/reproject.cc driven through its public functions (PROJ replaced by ref_shim/proj.h) ----
int ref_reproject_axes(const char* proj_in, const char* proj_out, const double* in_x, const double* in_y, int ix, int iy, double* out_x, double* out_y)
{
    try {
        reproject::reproject_axes(proj_in, proj_out, in_x, in_y, ix, iy, out_x, out_y);
        return 0;
    } catch (...) {
        return -1;
    }
}

int ref_get_vector_reproject_matrix(const char* proj_in, const char* proj_out, const double* out_x_axis, const double* out_y_axis, int ox, int oy,
                                    double* matrix3)
{
    try {
        auto m = reproject::get_vector_reproject_matrix(proj_in, proj_out, out_x_axis, out_y_axis, MIFI_PROJ_AXIS, MIFI_PROJ_AXIS, ox, oy);
        std::memcpy(matrix3, m->mtx(), sizeof(double) * 3 * ox * oy);
        return 0;
    } catch (...) {
        return -1;
    }
}

int ref_get_vector_reproject_matrix_field(const char* proj_in, const char* proj_out, const double* in_x_field, const double* in_y_field, int ox, int oy,
                                          double* matrix3)
{
    try {
        auto m = reproject::get_vector_reproject_matrix_field(proj_in, proj_out, in_x_field, in_y_field, ox, oy);
        std::memcpy(matrix3, m->mtx(), sizeof(double) * 3 * ox * oy);
        return 0;
    } catch (...) {
        return -1;
    }
}

// reproject::vector_reproject_values_f (reproject.cc:532-540): matrix from the axes + rotation, as the reference's tests call it
int ref_vector_reproject_values_f(const char* proj_in, const char* proj_out, float* u, float* v, const double* out_x_axis, const double* out_y_axis,
                                  int out_x_axis_type, int out_y_axis_type, int ox, int oy, int oz)
{
    try {
        reproject::vector_reproject_values_f(proj_in, proj_out, u, v, out_x_axis, out_y_axis, out_x_axis_type, out_y_axis_type, ox, oy, oz);
        return 0;
    } catch (...) {
        return -1;
    }
}

namespace {
reproject::Matrix_cp borrow_matrix(const double* matrix3, int ox, int oy)
{
    auto m = std::make_shared<reproject::Matrix>(ox, oy);
    std::memcpy(m->mtx(), matrix3, sizeof(double) * 3 * ox * oy);
    return m;
}
} // namespace

void ref_vector_reproject_values_by_matrix_f(const double* matrix3, int ox, int oy, float* u, float* v, int oz)
{
    reproject::vector_reproject_values_by_matrix_f(borrow_matrix(matrix3, ox, oy), u, v, oz);
}

void ref_vector_reproject_direction_by_matrix_f(const double* matrix3, int ox, int oy, float* angle, int oz)
{
    reproject::vector_reproject_direction_by_matrix_f(borrow_matrix(matrix3, ox, oy), angle, oz);
}

int ref_reproject_f(int method, const char* proj_in, const float* infield, const double* in_x_axis, const double* in_y_axis, int in_x_type, int in_y_type,
                    int ix, int iy, int iz, const char* proj_out, float* outfield, const double* out_x_axis, const double* out_y_axis, int out_x_type,
                    int out_y_type, int ox, int oy)
{
    try {
        reproject::reproject_f(method, proj_in, infield, in_x_axis, in_y_axis, in_x_type, in_y_type, ix, iy, iz, proj_out, outfield, out_x_axis,
                               out_y_axis, out_x_type, out_y_type, ox, oy);
        return 0;
    } catch (...) {
        return -1;
    }
}

// The reference's own conversion templates for a CDM_SHORT variable around the interpolation, instantiated exactly as
// CDMInterpolator.cc:140-151 / DataImpl.h:296-307 / CDMReader.cc:177-192 instantiate them (header-only code of the reference).
void ref_i16_to_interpolation_array(const short* in, size_t n, double badValue, float* out)
{
    std::transform(in, in + n, out, MetNoFimex::data_caster<float, short>()); // Data::asFloat -> ArrayTypeConverter<float,short>
    mifi_bad2nanf(out, out + n, badValue);
}

void ref_interpolation_array_to_i16(const float* in, size_t n, double badValue, short* out)
{
    MetNoFimex::ScaleValue<float, short> sv(MIFI_UNDEFINED_F, 1., 0., badValue, 1., 0.);
    std::transform(in, in + n, out, sv);
}

void ref_i16_scaled_to_float(const short* in, size_t n, double fill, double scale, double offset, float* out)
{
    MetNoFimex::ScaleValue<short, double> sv(fill, scale, offset, MIFI_UNDEFINED_D, 1., 0.);
    for (size_t i = 0; i < n; ++i)
        out[i] = MetNoFimex::data_caster<float, double>()(sv(in[i]));
}

int ref_unpack_f(int dtype, const void* in, size_t n, double bad, float* out)
{
    switch (dtype) {
    case 0: ref_unpack_t<char>(in, n, bad, out); return 1;
    case 1: ref_unpack_t<short>(in, n, bad, out); return 1;
    case 2: ref_unpack_t<int>(in, n, bad, out); return 1;
    case 3: ref_unpack_t<unsigned char>(in, n, bad, out); return 1;
    case 4: ref_unpack_t<unsigned short>(in, n, bad, out); return 1;
    case 5: ref_unpack_t<unsigned int>(in, n, bad, out); return 1;
    default: return 0;
    }
}

int ref_unpack_scaled_f(int dtype, const void* in, size_t n, double fill, double scale, double offset, float* out)
{
    switch (dtype) {
    case 0: ref_unpack_scaled_t<char>(in, n, fill, scale, offset, out); return 1;
    case 1: ref_unpack_scaled_t<short>(in, n, fill, scale, offset, out); return 1;
    case 2: ref_unpack_scaled_t<int>(in, n, fill, scale, offset, out); return 1;
    case 3: ref_unpack_scaled_t<unsigned char>(in, n, fill, scale, offset, out); return 1;
    case 4: ref_unpack_scaled_t<unsigned short>(in, n, fill, scale, offset, out); return 1;
    case 5: ref_unpack_scaled_t<unsigned int>(in, n, fill, scale, offset, out); return 1;
    default: return 0;
    }
}

int ref_pack_f(int dtype, const float* in, size_t n, double bad, void* out)
{
    switch (dtype) {
    case 0: ref_pack_t<char>(in, n, bad, out); return 1;
    case 1: ref_pack_t<short>(in, n, bad, out); return 1;
    case 2: ref_pack_t<int>(in, n, bad, out); return 1;
    case 3: ref_pack_t<unsigned char>(in, n, bad, out); return 1;
    case 4: ref_pack_t<unsigned short>(in, n, bad, out); return 1;
    case 5: ref_pack_t<unsigned int>(in, n, bad, out); return 1;
    default: return 0;
    }
}

} // extern "C"
