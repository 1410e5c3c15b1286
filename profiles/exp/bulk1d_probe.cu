// Can a ring of shared-memory stages be fed with many small 1-D bulk copies (cp.async.bulk, one per box row) fast enough to
// replace the single 3-D tensor copy per stage where the source rows are not 16-byte aligned (the reference's dense layout)?
// This probe keeps the headline kernel's structure (32 x 8 target tiles, 16 rows x 4 levels of 36 floats per stage, 4 stages,
// 8 consumer warps + producer warp, 6 CTAs per SM, 17 stages per tile, 11648 tiles) but does no arithmetic: consumers read one
// value per stage and store one line per level, the producer warp issues the 64 row copies of a stage from its 32 lanes.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bulk1d_probe bulk1d_probe.cu
#include <cstdint>
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

constexpr int ROWF = 36;  // floats per row copy (32 + up to 3 of misalignment, rounded to 16 B)
constexpr int BH = 16, ZB = 4, STAGES = 4;
constexpr int STAGE_FLOATS = ROWF * BH * ZB;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity)
{
    uint32_t ok = 0;
    const uint32_t addr = smem_u32(bar);
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_1d(void* dst, const void* src, unsigned bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes),
                 "r"(smem_u32(bar))
                 : "memory");
}

// src: [nz][inY][pitch] floats, pitch a multiple of 4 here so that every row start is 16-byte aligned (the probe measures the copy
// machinery; in the real kernel the start is rounded down to the boundary)
__global__ void __launch_bounds__(288, 6) probe(const float* __restrict__ src, int pitch, size_t level, int inY, int nz, float* __restrict__ out, int outX, int outY,
                                                 int tiles_x, int lanes_issuing)
{
    __shared__ __align__(128) float tile[STAGES][STAGE_FLOATS];
    __shared__ __align__(8) uint64_t full[STAGES], empty[STAGES];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 8);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
    // source box of this tile: ~18 columns x 9 rows, as on S->T1
    const int bx0 = ((tx * 18) % (pitch - 64)) & ~3, by0 = (ty * 8) % (inY - BH);
    const int niter = (nz + ZB - 1) / ZB;
    if (warp == 8) { // producer warp
        for (int it = 0; it < niter; ++it) {
            const int s = it % STAGES;
            if (lane == 0) {
                mbar_wait(&empty[s], ((it / STAGES) & 1) ^ 1);
                mbar_expect_tx(&full[s], STAGE_FLOATS * 4);
            }
            __syncwarp();
            for (int r = lane; r < BH * ZB; r += lanes_issuing) {
                if (lane >= lanes_issuing)
                    break;
                const int z = it * ZB + r / BH, y = by0 + r % BH;
                const float* p = src + static_cast<size_t>(min(z, nz - 1)) * level + static_cast<size_t>(y) * pitch + bx0;
                bulk_1d(&tile[s][r * ROWF], p, ROWF * 4, &full[s]);
            }
        }
        return;
    }
    const int ox = tx * 32 + lane, oy = ty * 8 + warp;
    float* po = out + static_cast<size_t>(oy) * outX + ox;
    const size_t olevel = static_cast<size_t>(outX) * outY;
    for (int it = 0; it < niter; ++it) {
        const int s = it % STAGES;
        mbar_wait(&full[s], (it / STAGES) & 1);
#pragma unroll
        for (int k = 0; k < ZB; ++k) {
            const float* t = &tile[s][k * BH * ROWF + (warp + 1) * ROWF + (lane >> 1)];
            const float v = t[0] + t[1] + t[ROWF] + t[ROWF + 1];
            if (it * ZB + k < nz && ox < outX && oy < outY)
                __stcs(po + k * olevel, v);
        }
        po += ZB * olevel;
        __syncwarp();
        if (lane == 0)
            mbar_arrive(&empty[s]);
    }
}

int main()
{
    const int inX = 949, inY = 1069, nz = 65, pitch = 952, outX = 2912, outY = 1021; // padded pitches: aligned rows, full store lines
    const size_t level = static_cast<size_t>(pitch) * inY;
    float *src, *out;
    CK(cudaMalloc(&src, level * nz * 4));
    CK(cudaMalloc(&out, static_cast<size_t>(outX) * outY * nz * 4));
    CK(cudaMemset(src, 0, level * nz * 4));
    const int tiles_x = outX / 32, tiles_y = (outY + 7) / 8;
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    (void)inX;
    for (int lanes : {32, 16, 8, 1}) {
        float best = 1e9f;
        for (int r = 0; r < 6; ++r) {
            CK(cudaEventRecord(a));
            probe<<<tiles_x * tiles_y, 288>>>(src, pitch, level, inY, nz, out, outX, outY, tiles_x, lanes);
            CK(cudaEventRecord(b));
            CK(cudaEventSynchronize(b));
            float ms;
            CK(cudaEventElapsedTime(&ms, a, b));
            best = ms < best ? ms : best;
        }
        CK(cudaGetLastError());
        std::printf("64 row copies of %d B per stage issued by %2d lanes: %.4f ms per slab-sized launch (%d tiles x %d stages)\n", ROWF * 4, lanes, best,
                    tiles_x * tiles_y, (nz + ZB - 1) / ZB);
    }
    return 0;
}
