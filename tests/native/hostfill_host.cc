// Host-side logic of libfimexcu's host entry points, driven without a GPU (tests/test_host_helpers.py): the streaming fill and copy
// (fimex_b200/csrc/fxcu_hostfill.h), the row bands, and the persistent worker pool (fxcu_hostpar.h).
#include "../../fimex_b200/csrc/fxcu_hostfill.h"
#include "../../fimex_b200/csrc/fxcu_hostpar.h"

#include <atomic>

using namespace fxcu;

extern "C" {

void hostfill_fill_f(float* p, size_t n, float v) { hostfill::fill_lines(p, n, v); _mm_sfence(); }
void hostfill_fill_i16(short* p, size_t n, short v) { hostfill::fill_lines(p, n, v); _mm_sfence(); }
void hostfill_copy(void* dst, const void* src, size_t bytes) { hostfill::copy_lines(dst, src, bytes); _mm_sfence(); }

// spans2: [n][2] = lo, hi per row; bands4: room for max_bands quadruples; returns the number of bands
int hostfill_merge(const int* spans2, size_t n, int outX, size_t max_bands, int* bands4, double* fraction)
{
    std::vector<hostfill::Span> rows(n);
    for (size_t y = 0; y < n; ++y)
        rows[y] = hostfill::Span{spans2[2 * y], spans2[2 * y + 1]};
    const std::vector<hostfill::Band> b = hostfill::merge_row_spans(rows, outX, max_bands, *fraction);
    for (size_t k = 0; k < b.size(); ++k) {
        bands4[4 * k] = b[k].y0;
        bands4[4 * k + 1] = b[k].y1;
        bands4[4 * k + 2] = b[k].x0;
        bands4[4 * k + 3] = b[k].x1;
    }
    return static_cast<int>(b.size());
}

// every index of [0, n) visited exactly once, `rounds` times over, with nested calls from inside the ranges; returns the number of
// indices whose count is wrong
size_t hostpar_cover(size_t n, size_t grain, size_t max_parts, int rounds, int nested)
{
    std::vector<std::atomic<int>> hits(n);
    for (auto& h : hits)
        h.store(0);
    for (int r = 0; r < rounds; ++r)
        parallel_ranges(n, grain, [&](size_t a, size_t b) {
            if (nested)
                parallel_ranges(b - a, 1, [&](size_t c, size_t d) {
                    for (size_t i = a + c; i < a + d; ++i)
                        hits[i].fetch_add(1);
                });
            else
                for (size_t i = a; i < b; ++i)
                    hits[i].fetch_add(1);
        }, max_parts);
    size_t bad = 0;
    for (auto& h : hits)
        bad += h.load() != rounds;
    return bad;
}
}
