// Test harness (CPU): drives the HOST part of the product's coord_kdtree path -- fimex_b200/csrc/fxcu_kdclone.h: libm xyz,
// tree build, host walk + std::sort -- for every target, so that its agreement with the reference's nanoflann
// (oracle/_ref: ref_kdtree_translate) can be checked without a GPU.  Not part of libfimexcu.so.
#include "../../fimex_b200/csrc/fxcu_kdclone.h"

#include <cstddef>
#include <cstring>

extern "C" int kdclone_translate_host(double maxDist_m, double* px, double* py, size_t nq, const double* lon, const double* lat, size_t nx, size_t ny,
                                      unsigned threads, char* err, size_t errlen)
{
    using namespace fxcu::kdclone;
    const size_t n = nx * ny;
    std::vector<double> xyz(3 * n);
    sphere_xyz(lon, lat, 0, n, true, xyz.data());
    Tree t;
    build(t, xyz.data(), n, threads);
    if (!t.error.empty()) {
        std::strncpy(err, t.error.c_str(), errlen - 1);
        err[errlen - 1] = 0;
        return -1;
    }
    const double r = maxDist_m / 6371000.;
    const double r2 = r * r;
    for (size_t i = 0; i < nq; ++i) {
        double q[3];
        sphere_xyz(px, py, i, i + 1, false, q - 3 * i);
        const long pos = search_first(t, q, r2);
        if (pos >= 0) {
            px[i] = static_cast<double>(size_t(pos) % nx);
            py[i] = static_cast<double>(size_t(pos) / nx);
        } else {
            px[i] = py[i] = -1000;
        }
    }
    return 0;
}

extern "C" size_t kdclone_node_count(const double* lon, const double* lat, size_t n, unsigned threads)
{
    using namespace fxcu::kdclone;
    std::vector<double> xyz(3 * n);
    sphere_xyz(lon, lat, 0, n, true, xyz.data());
    Tree t;
    build(t, xyz.data(), n, threads);
    return t.nodes.size();
}
