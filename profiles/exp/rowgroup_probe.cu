// Probe: a 2-D tensor map {4 * pitch elements, groups of four rows} over a dense slab with an odd row pitch, box {32, 4}.
// nvcc -gencode arch=compute_100a,code=sm_100a -o rowgroup_probe rowgroup_probe.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__global__ void probe(const __grid_constant__ CUtensorMap map, int x, int y, float* out)
{
    __shared__ __align__(128) float tile[128];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(512) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(tile)),
                     "l"(&map), "r"(smem_u32(&bar)), "r"(x), "r"(y)
                     : "memory");
    }
    uint32_t ok = 0;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(smem_u32(&bar)), "r"(0) : "memory");
    } while (!ok);
    out[threadIdx.x] = tile[threadIdx.x];
}

int main(int argc, char** argv)
{
    const int pitch = argc > 1 ? atoi(argv[1]) : 949, rows = 64;
    std::vector<float> h(size_t(pitch) * rows);
    for (size_t i = 0; i < h.size(); ++i)
        h[i] = float(i / pitch) * 1000.f + float(i % pitch);
    float *d, *o;
    cudaMalloc(&d, h.size() * 4);
    cudaMalloc(&o, 128 * 4);
    cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    typedef CUresult (*enc_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                              CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q);
    CUtensorMap map;
    cuuint64_t dims[2] = {cuuint64_t(pitch) * 4, cuuint64_t(rows / 4)};
    cuuint64_t strides[1] = {cuuint64_t(pitch) * 16};
    cuuint32_t box[2] = {32, 4}, es[2] = {1, 1};
    CUresult rc = ((enc_t)fp)(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc %d (pitch %d)\n", int(rc), pitch);
    const int tests[][2] = {{0, 0}, {8, 1}, {949 + 5, 2}, {2 * 949 + 7, 0}, {3 * 949 + 930, 3}, {13, 100}};
    for (auto& t : tests) {
        probe<<<1, 128>>>(map, t[0] / 949 * pitch + t[0] % 949, t[1], o);
        cudaError_t e = cudaDeviceSynchronize();
        float r[128];
        cudaMemcpy(r, o, sizeof(r), cudaMemcpyDeviceToHost);
        printf("x %d y %d -> %s : row0 %.0f %.0f .. %.0f | row1 %.0f | row3 %.0f\n", t[0], t[1], cudaGetErrorString(e), r[0], r[1], r[31], r[32], r[96]);
        if (e != cudaSuccess)
            return 1;
    }
    return 0;
}
