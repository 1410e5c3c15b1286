// fxcu_runtime.cu -- error state, device selection and memory helpers of the C ABI (include/fimexcu.h, "runtime").
#include "fxcu_common.h"

#include <mutex>

namespace fxcu {

namespace {
thread_local std::string g_last_error;
std::once_flag g_pool_once[64];
}

void set_error(const std::string& msg) { g_last_error = msg; }
void clear_error() { g_last_error.clear(); }
const char* last_error_cstr() { return g_last_error.c_str(); }

bool cuda_ok(cudaError_t e, const char* what, const char* file, int line)
{
    if (e == cudaSuccess)
        return true;
    char buf[512];
    std::snprintf(buf, sizeof(buf), "CUDA error %d (%s) in %s at %s:%d", static_cast<int>(e), cudaGetErrorString(e), what, file, line);
    set_error(buf);
    cudaGetLastError(); // clear the sticky-less error state
    return false;
}

bool require_device()
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        set_error(std::string("no usable CUDA device: ") + cudaGetErrorString(e) + " (libfimexcu has no CPU fallback)");
        cudaGetLastError();
        return false;
    }
    int major = 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess || major != 10) {
        set_error("libfimexcu is built for sm_100a only; the current device is not a compute-capability 10.x GPU");
        cudaGetLastError();
        return false;
    }
    if (dev >= 0 && dev < 64) {
        std::call_once(g_pool_once[dev], [dev]() {
            cudaMemPool_t pool;
            if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
                unsigned long long keep = ~0ull;
                cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
            }
            cudaGetLastError();
        });
    }
    return true;
}

int sm_count()
{
    int dev = 0, n = 148;
    if (cudaGetDevice(&dev) == cudaSuccess)
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    return n > 0 ? n : 148;
}

bool DevBuf::alloc(size_t bytes, cudaStream_t s)
{
    release();
    if (bytes == 0)
        bytes = 16;
    stream = s;
    async = true;
    return cuda_ok(cudaMallocAsync(&p, bytes, s), "cudaMallocAsync", __FILE__, __LINE__);
}

bool DevBuf::alloc_sync(size_t bytes)
{
    release();
    if (bytes == 0)
        bytes = 16;
    async = false;
    return cuda_ok(cudaMalloc(&p, bytes), "cudaMalloc", __FILE__, __LINE__);
}

void DevBuf::release()
{
    if (!p)
        return;
    if (async)
        cudaFreeAsync(p, stream);
    else
        cudaFree(p);
    p = nullptr;
}

} // namespace fxcu

extern "C" {

const char* fimexcu_last_error(void) { return fxcu::last_error_cstr(); }

int fimexcu_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    int usable = 0;
    for (int d = 0; d < n; ++d) {
        int major = 0;
        if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, d) == cudaSuccess && major == 10)
            ++usable;
    }
    return usable;
}

int fimexcu_set_device(int ordinal)
{
    FXCU_CUDA(cudaSetDevice(ordinal));
    return fxcu::require_device() ? FIMEXCU_OK : FIMEXCU_ERROR;
}

void* fimexcu_host_alloc(size_t bytes)
{
    void* p = nullptr;
    if (!fxcu::require_device())
        return nullptr;
    FXCU_CUDA_NULL(cudaMallocHost(&p, bytes ? bytes : 1));
    return p;
}

void fimexcu_host_free(void* p)
{
    if (p)
        cudaFreeHost(p);
}

void* fimexcu_dev_alloc(size_t bytes)
{
    void* p = nullptr;
    if (!fxcu::require_device())
        return nullptr;
    FXCU_CUDA_NULL(cudaMalloc(&p, bytes ? bytes : 1));
    return p;
}

void fimexcu_dev_free(void* p)
{
    if (p)
        cudaFree(p);
}

int fimexcu_memcpy_h2d(void* dst_dev, const void* src_host, size_t bytes)
{
    FXCU_CUDA(cudaMemcpy(dst_dev, src_host, bytes, cudaMemcpyHostToDevice));
    return FIMEXCU_OK;
}

int fimexcu_memcpy_d2h(void* dst_host, const void* src_dev, size_t bytes)
{
    FXCU_CUDA(cudaMemcpy(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost));
    return FIMEXCU_OK;
}

int fimexcu_synchronize(void)
{
    FXCU_CUDA(cudaDeviceSynchronize());
    return FIMEXCU_OK;
}

} // extern "C"
