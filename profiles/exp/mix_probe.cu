// What does HBM deliver for the headline kernel's traffic MIX?  K4 (S->T1) reads 303 MB and writes 712 MB per launch; the roofline's
// denominator (MEASURED_PEAKS.json) is a copy, i.e. 1 read : 1 write.  This probe streams the same byte counts with no arithmetic
// at all (grid-stride float4 loads folded into a never-taken store, float4 stores of a constant) and, for calibration, a plain copy
// and a pure store of the same total.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mix_probe mix_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x)                                                                          \
    do {                                                                               \
        cudaError_t e_ = (x);                                                          \
        if (e_ != cudaSuccess) {                                                       \
            std::printf("%s: %s\n", #x, cudaGetErrorString(e_));                      \
            std::exit(1);                                                              \
        }                                                                              \
    } while (0)

// every block handles a contiguous share of the reads and of the writes, interleaved 1 read : `ratio` writes like the real kernel
__global__ void __launch_bounds__(256) mix_kernel(const float4* __restrict__ src, size_t n_read, float4* __restrict__ dst, size_t n_write, float4* sink)
{
    const size_t tid = blockIdx.x * 256ull + threadIdx.x, nth = gridDim.x * 256ull;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const size_t steps = (max(n_read, n_write) + nth - 1) / nth;
    const double rs = static_cast<double>(n_read) / static_cast<double>(max(n_read, n_write));
    const double ws = static_cast<double>(n_write) / static_cast<double>(max(n_read, n_write));
    size_t r_done = 0, w_done = 0;
    for (size_t s = 0; s < steps; ++s) {
        const size_t r_to = static_cast<size_t>((s + 1) * rs), w_to = static_cast<size_t>((s + 1) * ws);
        for (; r_done < r_to; ++r_done) {
            const size_t i = r_done * nth + tid;
            if (i < n_read) {
                const float4 v = __ldg(src + i);
                acc.x += v.x;
                acc.y += v.y;
                acc.z += v.z;
                acc.w += v.w;
            }
        }
        for (; w_done < w_to; ++w_done) {
            const size_t i = w_done * nth + tid;
            if (i < n_write)
                dst[i] = make_float4(1.f, 2.f, 3.f, static_cast<float>(s));
        }
    }
    if (acc.x == 12345.678f)
        *sink = acc;
}

static float time_ms(void (*launch)(void*), void* ctx, int reps)
{
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    for (int i = 0; i < 3; ++i)
        launch(ctx);
    float best = 1e30f;
    for (int i = 0; i < reps; ++i) {
        CK(cudaEventRecord(a));
        launch(ctx);
        CK(cudaEventRecord(b));
        CK(cudaEventSynchronize(b));
        float ms;
        CK(cudaEventElapsedTime(&ms, a, b));
        best = ms < best ? ms : best;
    }
    return best;
}

struct Ctx {
    const float4* src;
    float4* dst;
    float4* sink;
    size_t n_read, n_write;
    int blocks;
};
static void launch_mix(void* p)
{
    Ctx* c = static_cast<Ctx*>(p);
    mix_kernel<<<c->blocks, 256>>>(c->src, c->n_read, c->dst, c->n_write, c->sink);
}
static void launch_copy(void* p)
{
    Ctx* c = static_cast<Ctx*>(p);
    CK(cudaMemcpyAsync(c->dst, c->src, c->n_write * 16, cudaMemcpyDeviceToDevice));
}
static void launch_set(void* p)
{
    Ctx* c = static_cast<Ctx*>(p);
    CK(cudaMemsetAsync(c->dst, 1, c->n_write * 16));
}

int main()
{
    const size_t total = 1015309824ull; // K4's measured DRAM bytes per launch
    float4 *src, *dst, *sink;
    CK(cudaMalloc(&src, total));
    CK(cudaMalloc(&dst, total));
    CK(cudaMalloc(&sink, 16));
    CK(cudaMemset(src, 0, total));
    CK(cudaMemset(dst, 0, total));
    struct {
        const char* name;
        size_t rd, wr;
    } mixes[] = {{"K4 S->T1 mix (303 MB read, 712 MB written)", 303491328ull, 711818496ull},
                 {"half and half", total / 2, total / 2},
                 {"reads only", total, 0},
                 {"writes only", 0, total},
                 {"T2 mix (84 MB read, 117 MB written)", 84294600ull, 116833860ull}};
    for (auto& m : mixes)
        for (int per_sm : {8, 16, 32}) {
            Ctx c{src, dst, sink, m.rd / 16, m.wr / 16, 148 * per_sm};
            const float ms = time_ms(launch_mix, &c, 20);
            std::printf("%-46s %2d blocks/SM: %.4f ms  %.1f GB/s\n", m.name, per_sm, ms, (m.rd + m.wr) / ms * 1e-6);
        }
    Ctx c{src, dst, sink, total / 32, total / 32, 0};
    float ms = time_ms(launch_copy, &c, 20);
    std::printf("cudaMemcpy D2D of %zu bytes (read + write = %zu): %.4f ms  %.1f GB/s\n", total / 2, total, ms, total / ms * 1e-6);
    c.n_write = total / 16;
    ms = time_ms(launch_set, &c, 20);
    std::printf("cudaMemset of %zu bytes: %.4f ms  %.1f GB/s\n", total, ms, total / ms * 1e-6);
    return 0;
}
