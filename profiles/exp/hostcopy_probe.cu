// Host-side rates behind the host entry points: pageable -> pinned memcpy and streaming fill, against the thread count.
// Build: nvcc -O3 -std=c++17 -o hostcopy_probe hostcopy_probe.cu
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include <emmintrin.h>
#include <cuda_runtime.h>

static double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

template <class F> static void par(size_t n, int threads, F fn)
{
    std::vector<std::thread> th;
    const size_t step = (n + threads - 1) / threads;
    for (int k = 1; k < threads; ++k)
        th.emplace_back([=]() { fn(std::min(n, k * step), std::min(n, (k + 1) * step)); });
    fn(0, std::min(n, step));
    for (auto& t : th)
        t.join();
}

int main()
{
    const size_t total = 264ull << 20, chunk = 33ull << 20;
    char* pageable = static_cast<char*>(std::malloc(total));
    std::memset(pageable, 1, total);
    char* pinned;
    if (cudaMallocHost(&pinned, total) != cudaSuccess)
        return 1;
    std::memset(pinned, 2, total);
    std::printf("hardware threads: %u\n", std::thread::hardware_concurrency());
    for (int threads : {1, 2, 4, 8, 16, 32}) {
        double best_all = 1e30, best_chunks = 1e30, best_fill = 1e30;
        for (int r = 0; r < 3; ++r) {
            double t0 = now_ms();
            par(total, threads, [&](size_t a, size_t b) { std::memcpy(pinned + a, pageable + a, b - a); });
            best_all = std::min(best_all, now_ms() - t0);
            t0 = now_ms();
            for (size_t o = 0; o < total; o += chunk)
                par(chunk, threads, [&](size_t a, size_t b) { std::memcpy(pinned + o + a, pageable + o + a, b - a); });
            best_chunks = std::min(best_chunks, now_ms() - t0);
            t0 = now_ms();
            par(total, threads, [&](size_t a, size_t b) {
                const __m128i q = _mm_set1_epi32(0x7fc00000);
                for (size_t i = a; i + 16 <= b; i += 16)
                    _mm_stream_si128(reinterpret_cast<__m128i*>(pinned + i), q);
                _mm_sfence();
            });
            best_fill = std::min(best_fill, now_ms() - t0);
        }
        std::printf("%2d threads: memcpy 264 MB at once %.2f ms (%.1f GB/s), in 8 chunks %.2f ms (%.1f GB/s), streaming fill %.2f ms (%.1f GB/s)\n", threads, best_all,
                    total / best_all * 1e-6, best_chunks, total / best_chunks * 1e-6, best_fill, total / best_fill * 1e-6);
    }
    // the undefined margins of a banded S->T1 result (rows of 2891 floats, data in [500, 2400)), written the way the library does,
    // alone and while the DMA engine writes a result into the same array's spans
    {
        const size_t outX = 2891, outY = 1021, nz = 65, full = outX * outY * nz * 4;
        float *res, *dev;
        if (cudaMallocHost(&res, full) != cudaSuccess || cudaMalloc(&dev, full) != cudaSuccess)
            return 1;
        std::memset(res, 0, full);
        cudaStream_t st;
        cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
        auto fill_stream = [](float* p, size_t n, float v) {
            while (n && (reinterpret_cast<uintptr_t>(p) & 15u)) {
                *p++ = v;
                --n;
            }
            const __m128i q = _mm_set1_epi32(0x7fc00000);
            for (; n >= 4; n -= 4, p += 4)
                _mm_stream_si128(reinterpret_cast<__m128i*>(p), q);
            while (n) {
                *p++ = v;
                --n;
            }
        };
        // B: the right margin of a row and the left margin of the next one as ONE run; ordinary stores up to the first 64-byte
        //    boundary and after the last one, whole lines with streaming stores
        auto fill_lines = [](float* p, size_t n, float v) {
            while (n && (reinterpret_cast<uintptr_t>(p) & 63u)) {
                *p++ = v;
                --n;
            }
            const __m128i q = _mm_set1_epi32(0x7fc00000);
            for (; n >= 16; n -= 16, p += 16) {
                _mm_stream_si128(reinterpret_cast<__m128i*>(p), q);
                _mm_stream_si128(reinterpret_cast<__m128i*>(p) + 1, q);
                _mm_stream_si128(reinterpret_cast<__m128i*>(p) + 2, q);
                _mm_stream_si128(reinterpret_cast<__m128i*>(p) + 3, q);
            }
            while (n) {
                *p++ = v;
                --n;
            }
        };
        auto fill_stos = [](float* p, size_t n, float v) { // E: the string instruction (fast-string microcode writes whole lines without reading them)
            unsigned bits;
            std::memcpy(&bits, &v, 4);
            asm volatile("rep stosl" : "+D"(p), "+c"(n) : "a"(bits) : "memory");
        };
        auto fill_std = [](float* p, size_t n, float v) { std::fill(p, p + n, v); };
        const float nanv = __builtin_nanf("");
        struct Variant {
            const char* name;
            int kind;
        } variants[] = {{"A 16-byte streaming stores per piece", 0}, {"B merged runs, whole lines streamed", 1}, {"C std::fill per piece", 2}, {"E rep stosl, merged runs", 3},
                        {"F rep stosl per piece", 4}};
        for (auto& v : variants)
            for (int dma : {0, 1})
                for (int threads : {4, 8}) {
                    double best = 1e30;
                    for (int r = 0; r < 3; ++r) {
                        if (dma)
                            cudaMemcpy2DAsync(res + 500, outX * 4, dev + 500, outX * 4, 1900 * 4, outY * nz, cudaMemcpyDeviceToHost, st);
                        const double t0 = now_ms();
                        par(nz * outY, threads, [&](size_t a, size_t b) {
                            if (v.kind == 1 || v.kind == 3) {
                                float* first = res + a * outX;
                                (v.kind == 1 ? (void)fill_lines(first, 500, nanv) : (void)fill_stos(first, 500, nanv));
                                for (size_t k = a; k < b; ++k) {
                                    float* run = res + k * outX + 2400;
                                    const size_t n = (outX - 2400) + (k + 1 < b ? 500 : 0);
                                    if (v.kind == 1)
                                        fill_lines(run, n, nanv);
                                    else
                                        fill_stos(run, n, nanv);
                                }
                            } else
                                for (size_t k = a; k < b; ++k) {
                                    float* row = res + k * outX;
                                    if (v.kind == 0) {
                                        fill_stream(row, 500, nanv);
                                        fill_stream(row + 2400, outX - 2400, nanv);
                                    } else if (v.kind == 2) {
                                        fill_std(row, 500, nanv);
                                        fill_std(row + 2400, outX - 2400, nanv);
                                    } else {
                                        fill_stos(row, 500, nanv);
                                        fill_stos(row + 2400, outX - 2400, nanv);
                                    }
                                }
                            _mm_sfence();
                        });
                        best = std::min(best, now_ms() - t0);
                        cudaStreamSynchronize(st);
                    }
                    std::printf("%-40s %d threads, DMA %s: %6.2f ms (%.1f GB/s)\n", v.name, threads, dma ? "running" : "idle   ", best,
                                (outX - 1900) * 4.0 * outY * nz / best * 1e-6);
                }
    }
    return 0;
}
