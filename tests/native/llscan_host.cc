// Test harness (CPU): drives the HOST part of the product's coord_nearestneighbor path -- fimex_b200/csrc/fxcu_llscan.h: the
// std::sort'ed latitude list and the reference's scan with its dot-product shortcut -- for every target, so that it can be
// compared with the reference restatement in oracle/_ref (ref_coordnn_translate) without a GPU.  Not part of libfimexcu.so.
#include "../../fimex_b200/csrc/fxcu_llscan.h"

extern "C" int llscan_translate_host(double* px, double* py, size_t nq, const double* lon, const double* lat, size_t nx, size_t ny, double roi_rad)
{
    using namespace fxcu::llscan;
    SortedCloud c;
    build(c, lon, lat, nx, ny);
    const double min_grid_cos_d = std::cos(roi_rad);
    fxcu::parallel_ranges(nq, 64, [&](size_t a, size_t b) {
        for (size_t i = a; i < b; ++i) {
            const long pos = scan(c, px[i], py[i], min_grid_cos_d);
            if (pos >= 0) {
                px[i] = static_cast<double>(size_t(pos) % nx);
                py[i] = static_cast<double>(size_t(pos) / nx);
            } else {
                px[i] = py[i] = -1;
            }
        }
    });
    return 0;
}
