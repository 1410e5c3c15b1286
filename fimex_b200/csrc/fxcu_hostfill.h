// fxcu_hostfill.h -- host-only pieces of the host entry points (fxcu_apply.cu): the row bands of a result that can hold data, and
// the streaming fill / copy of the caller's array around them.  No CUDA in here, so tests/native/hostfill_host.cc can drive it
// without a GPU.
#pragma once

#include <algorithm>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <vector>

#include <emmintrin.h>

namespace fxcu {
namespace hostfill {

struct Span { // a target row: first and one-past-last column that can hold data (lo >= hi: none)
    int lo, hi;
};
struct Band { // rows [y0, y1) with the common column span [x0, x1) (x0 >= x1: no row of the band holds data)
    int y0, y1, x0, x1;
};

// Rows -> at most max_bands bands covering [0, rows.size()) in order: blocks of rows first (at most 512 of them), then always the
// two neighbours joined whose common rectangle adds the least area.  fraction = area of the rectangles / area of the target.
inline std::vector<Band> merge_row_spans(const std::vector<Span>& rows, int outX, size_t max_bands, double& fraction)
{
    auto join = [](const Band& a, const Band& b) {
        Band r{a.y0, b.y1, 0, 0};
        if (a.x0 >= a.x1) {
            r.x0 = b.x0;
            r.x1 = b.x1;
        } else if (b.x0 >= b.x1) {
            r.x0 = a.x0;
            r.x1 = a.x1;
        } else {
            r.x0 = std::min(a.x0, b.x0);
            r.x1 = std::max(a.x1, b.x1);
        }
        return r;
    };
    auto area = [](const Band& b) { return static_cast<long long>(b.y1 - b.y0) * std::max(0, b.x1 - b.x0); };
    const size_t outY = rows.size();
    std::vector<Band> bands;
    fraction = 1.0;
    if (outY == 0 || outX <= 0)
        return bands;
    const size_t block = (outY + 511) / 512;
    for (size_t y0 = 0; y0 < outY; y0 += block) {
        Band b{static_cast<int>(y0), static_cast<int>(y0), 0, 0};
        for (size_t y = y0; y < std::min(outY, y0 + block); ++y)
            b = join(b, Band{static_cast<int>(y), static_cast<int>(y + 1), rows[y].lo, std::max(rows[y].lo, rows[y].hi)});
        bands.push_back(b);
    }
    while (bands.size() > std::max<size_t>(max_bands, 1)) {
        size_t best = 0;
        long long best_cost = -1;
        for (size_t i = 0; i + 1 < bands.size(); ++i) {
            const long long cost = area(join(bands[i], bands[i + 1])) - area(bands[i]) - area(bands[i + 1]);
            if (best_cost < 0 || cost < best_cost) {
                best_cost = cost;
                best = i;
            }
        }
        bands[best] = join(bands[best], bands[best + 1]);
        bands.erase(bands.begin() + static_cast<long>(best) + 1);
    }
    long long total = 0;
    for (const Band& b : bands)
        total += area(b);
    fraction = static_cast<double>(total) / (static_cast<double>(outX) * static_cast<double>(outY));
    return bands;
}

// n copies of v at p: ordinary stores up to the first 64-byte boundary and after the last one, streaming stores for the whole lines
// in between (written once, not read by this library: an ordinary store would first pull every line into the cache).  Streaming
// stores to PART of a line, or mixed with ordinary stores to the same line, run four times slower (profiles/exp/hostcopy_probe.cu).
template <class T> inline void fill_lines(T* p, size_t n, T v)
{
    static_assert(sizeof(T) == 4 || sizeof(T) == 2, "float or short");
    while (n && (reinterpret_cast<uintptr_t>(p) & 63u)) {
        *p++ = v;
        --n;
    }
    constexpr size_t PER = 64 / sizeof(T);
    if (n >= PER) {
        alignas(16) T pat[16 / sizeof(T)];
        for (T& e : pat)
            e = v;
        const __m128i q = _mm_load_si128(reinterpret_cast<const __m128i*>(pat));
        for (; n >= PER; n -= PER, p += PER) {
            __m128i* line = reinterpret_cast<__m128i*>(p);
            _mm_stream_si128(line, q);
            _mm_stream_si128(line + 1, q);
            _mm_stream_si128(line + 2, q);
            _mm_stream_si128(line + 3, q);
        }
    }
    while (n) {
        *p++ = v;
        --n;
    }
}

// the same for a copy: a result span leaves the pinned bounce buffer for the caller's array, whole destination lines streamed
inline void copy_lines(void* dst_, const void* src_, size_t bytes)
{
    char* dst = static_cast<char*>(dst_);
    const char* src = static_cast<const char*>(src_);
    const size_t head = std::min<size_t>(bytes, (64 - (reinterpret_cast<uintptr_t>(dst) & 63u)) & 63u);
    std::memcpy(dst, src, head);
    dst += head;
    src += head;
    bytes -= head;
    for (; bytes >= 64; bytes -= 64, dst += 64, src += 64) {
        const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src)), b = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src) + 1);
        const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src) + 2), d = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src) + 3);
        _mm_stream_si128(reinterpret_cast<__m128i*>(dst), a);
        _mm_stream_si128(reinterpret_cast<__m128i*>(dst) + 1, b);
        _mm_stream_si128(reinterpret_cast<__m128i*>(dst) + 2, c);
        _mm_stream_si128(reinterpret_cast<__m128i*>(dst) + 3, d);
    }
    std::memcpy(dst, src, bytes);
}

} // namespace hostfill
} // namespace fxcu
