// What does pinning a caller's pageable slab in place cost (cudaHostRegister + cudaHostUnregister), against copying it through
// pinned bounce buffers?  264 MB = one S->T1 input slab.  Build: nvcc -O3 -std=c++17 -o hostregister_probe hostregister_probe.cu
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
static double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main()
{
    const size_t bytes = 264ull << 20;
    char* p = static_cast<char*>(std::malloc(bytes + 4096));
    std::memset(p, 1, bytes + 4096);
    char* dev;
    cudaMalloc(&dev, bytes);
    cudaFree(0);
    for (int r = 0; r < 4; ++r) {
        double t0 = now_ms();
        cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterDefault);
        double t1 = now_ms();
        if (e != cudaSuccess) { std::printf("register: %s\n", cudaGetErrorString(e)); return 1; }
        cudaMemcpy(dev, p, bytes, cudaMemcpyHostToDevice);
        double t2 = now_ms();
        cudaHostUnregister(p);
        double t3 = now_ms();
        std::printf("register %.2f ms, H2D from the registered slab %.2f ms (%.1f GB/s), unregister %.2f ms\n", t1 - t0, t2 - t1, bytes / (t2 - t1) * 1e-6, t3 - t2);
    }
    // in 8 pieces (what a chunked pipeline would do)
    double t0 = now_ms();
    for (int k = 0; k < 8; ++k)
        cudaHostRegister(p + k * (bytes / 8), bytes / 8, cudaHostRegisterDefault);
    double t1 = now_ms();
    for (int k = 0; k < 8; ++k)
        cudaHostUnregister(p + k * (bytes / 8));
    double t2 = now_ms();
    std::printf("8 pieces: register %.2f ms, unregister %.2f ms\n", t1 - t0, t2 - t1);
    // pageable H2D straight (the driver's own staging)
    t0 = now_ms();
    cudaMemcpy(dev, p, bytes, cudaMemcpyHostToDevice);
    t1 = now_ms();
    std::printf("cudaMemcpy from pageable memory: %.2f ms (%.1f GB/s)\n", t1 - t0, bytes / (t1 - t0) * 1e-6);
    return 0;
}
