// How fast are pitched device->host copies against plain ones?  (Decides how the host entry points bring back only the part of
// a result row that can hold data: 36.6 % of the S->T1 target lies outside the source domain.)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o d2h_probe d2h_probe.cu
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>

#define CK(x)                                                                          \
    do {                                                                               \
        cudaError_t e_ = (x);                                                          \
        if (e_ != cudaSuccess) {                                                       \
            std::printf("%s: %s\n", #x, cudaGetErrorString(e_));                      \
            std::exit(1);                                                              \
        }                                                                              \
    } while (0)

static double now_ms()
{
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int main()
{
    const size_t outX = 2891, outY = 1021, nz = 65, pitch = outX * 4;
    const size_t full = outX * outY * nz * 4;
    char *d, *h;
    CK(cudaMalloc(&d, full));
    CK(cudaMemset(d, 1, full));
    CK(cudaMallocHost(&h, full));
    std::memset(h, 0, full);
    cudaStream_t st;
    CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    auto timed = [&](const char* name, size_t bytes, auto&& issue) {
        double best = 1e30, issue_ms = 0;
        for (int r = 0; r < 4; ++r) {
            CK(cudaStreamSynchronize(st));
            const double t0 = now_ms();
            issue();
            const double t1 = now_ms();
            CK(cudaStreamSynchronize(st));
            const double t2 = now_ms();
            if (t2 - t0 < best) {
                best = t2 - t0;
                issue_ms = t1 - t0;
            }
        }
        std::printf("%-72s %8.3f ms  %6.1f GB/s  (issue %.3f ms)\n", name, best, bytes / best * 1e-6, issue_ms);
    };
    timed("1-D, whole slab (767 MB)", full, [&] { CK(cudaMemcpyAsync(h, d, full, cudaMemcpyDeviceToHost, st)); });
    timed("1-D, 66 % of the slab", full * 66 / 100, [&] { CK(cudaMemcpyAsync(h, d, full * 66 / 100, cudaMemcpyDeviceToHost, st)); });
    for (size_t w : {1900ul, 1024ul, 512ul}) {
        char name[128];
        std::snprintf(name, sizeof name, "2-D, %zu of 2891 floats per row, one call (66365 rows)", w);
        timed(name, w * 4 * outY * nz, [&] { CK(cudaMemcpy2DAsync(h + 400 * 4, pitch, d + 400 * 4, pitch, w * 4, outY * nz, cudaMemcpyDeviceToHost, st)); });
    }
    for (size_t rb : {16ul, 64ul}) {
        char name[128];
        std::snprintf(name, sizeof name, "3-D, 1900 floats x %zu rows x 8 levels per call, all of the slab", rb);
        timed(name, 1900 * 4 * outY * nz, [&] {
            for (size_t z0 = 0; z0 < nz; z0 += 8)
                for (size_t y0 = 0; y0 < outY; y0 += rb) {
                    cudaMemcpy3DParms p = {};
                    p.srcPtr = make_cudaPitchedPtr(d, pitch, outX, outY);
                    p.dstPtr = make_cudaPitchedPtr(h, pitch, outX, outY);
                    p.srcPos = make_cudaPos(400 * 4, y0, z0);
                    p.dstPos = p.srcPos;
                    p.extent = make_cudaExtent(1900 * 4, (y0 + rb <= outY ? rb : outY - y0), (z0 + 8 <= nz ? 8 : nz - z0));
                    p.kind = cudaMemcpyDeviceToHost;
                    CK(cudaMemcpy3DAsync(&p, st));
                }
        });
    }
    // the actual row bands of the S->T1 cache (fimexcu_cache_row_bands), level chunks of 9 as the host entry point issues them
    {
        const int bands[24][4] = {{0, 22, 1921, 2051},    {22, 82, 710, 2079},    {82, 134, 689, 2106},   {134, 204, 658, 2144},  {204, 292, 615, 2198},
                                  {292, 350, 584, 2238},  {350, 428, 537, 2298},  {428, 482, 501, 2343},  {482, 522, 472, 2380},  {522, 580, 426, 2439},
                                  {580, 628, 384, 2493},  {628, 664, 349, 2537},  {664, 710, 301, 2599},  {710, 770, 231, 2691},  {770, 828, 152, 2795},
                                  {828, 864, 97, 2868},   {864, 896, 43, 2890},   {896, 930, 0, 2780},    {930, 946, 54, 2587},   {946, 962, 162, 2479},
                                  {962, 978, 287, 2354},  {978, 994, 435, 2206},  {994, 1010, 621, 2020}, {1010, 1021, 894, 1747}};
        size_t cells = 0;
        for (auto& b : bands)
            cells += static_cast<size_t>(b[1] - b[0]) * (b[3] - b[2]);
        cudaStream_t st3[3];
        for (auto& q : st3)
            CK(cudaStreamCreateWithFlags(&q, cudaStreamNonBlocking));
        for (int align : {1, 16, 32})
            for (size_t zc : {9ul, 65ul}) {
                char name[128];
                std::snprintf(name, sizeof name, "3-D, the 24 S->T1 row bands, %zu levels per call, spans widened to %d floats", zc, align);
                size_t moved = 0;
                timed(name, cells * 4 * nz, [&] {
                    moved = 0;
                    for (size_t z0 = 0; z0 < nz; z0 += zc)
                        for (auto& b : bands) {
                            const size_t x0 = static_cast<size_t>(b[2]) / align * align, x1 = std::min<size_t>(outX, (static_cast<size_t>(b[3]) + align - 1) / align * align);
                            cudaMemcpy3DParms p = {};
                            p.srcPtr = make_cudaPitchedPtr(d, pitch, outX, outY);
                            p.dstPtr = make_cudaPitchedPtr(h, pitch, outX, outY);
                            p.srcPos = make_cudaPos(x0 * 4, b[0], z0);
                            p.dstPos = p.srcPos;
                            p.extent = make_cudaExtent((x1 - x0) * 4, b[1] - b[0], (z0 + zc <= nz ? zc : nz - z0));
                            p.kind = cudaMemcpyDeviceToHost;
                            CK(cudaMemcpy3DAsync(&p, st));
                            moved += (x1 - x0) * 4 * (b[1] - b[0]) * p.extent.depth;
                        }
                });
                std::printf("      (bytes actually moved: %zu)\n", moved);
            }
        // one 2-D call per band and level (what a 3-D copy may be split into anyway)
        timed("2-D per band and level (1560 calls)", cells * 4 * nz, [&] {
            for (size_t z = 0; z < nz; ++z)
                for (auto& b : bands)
                    CK(cudaMemcpy2DAsync(h + (z * outY + b[0]) * pitch + b[2] * 4, pitch, d + (z * outY + b[0]) * pitch + b[2] * 4, pitch, (b[3] - b[2]) * 4ul, b[1] - b[0],
                                         cudaMemcpyDeviceToHost, st));
        });
    }
    // compact on the device side, pitched on the host side (what a span-compacted device result would need for a pinned destination)
    timed("2-D, compact device rows (pitch = width) -> pitched host rows, 1900 floats", 1900 * 4 * outY * nz,
          [&] { CK(cudaMemcpy2DAsync(h + 400 * 4, pitch, d, 1900 * 4, 1900 * 4, outY * nz, cudaMemcpyDeviceToHost, st)); });
    // host -> device for comparison
    const size_t inX = 949, inY = 1069, inFull = inX * inY * nz * 4;
    timed("H2D 1-D, source slab (264 MB)", inFull, [&] { CK(cudaMemcpyAsync(d, h, inFull, cudaMemcpyHostToDevice, st)); });
    timed("H2D 2-D, 949-float rows into a 952-float pitch", inFull, [&] { CK(cudaMemcpy2DAsync(d, 952 * 4, h, inX * 4, inX * 4, inY * nz, cudaMemcpyHostToDevice, st)); });
    // host-side fill and scatter rates (what the CPU cores add): NaN margins and row scatter from a compact bounce buffer
    char* pageable = static_cast<char*>(std::malloc(full));
    std::memset(pageable, 0, full);
    double t0 = now_ms();
    std::memcpy(pageable, h, full);
    double t1 = now_ms();
    std::printf("host memcpy pinned -> pageable, one thread, 767 MB: %.2f ms  %.1f GB/s\n", t1 - t0, full / (t1 - t0) * 1e-6);
    t0 = now_ms();
    std::memset(pageable, 0xff, full / 3);
    t1 = now_ms();
    std::printf("host memset, one thread, 256 MB: %.2f ms  %.1f GB/s\n", t1 - t0, full / 3 / (t1 - t0) * 1e-6);
    return 0;
}
