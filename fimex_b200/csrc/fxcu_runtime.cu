// fxcu_runtime.cu -- error state, device selection and memory helpers of the C ABI (include/fimexcu.h, "runtime").
#include "fxcu_common.h"

#include <map>
#include <mutex>
#include <unordered_map>

namespace fxcu {

namespace {
thread_local std::string g_last_error;
constexpr int MAX_DEVICES = 64;
std::once_flag g_pool_once[MAX_DEVICES];
cudaMemPool_t g_pool[MAX_DEVICES] = {};

// the library's own stream-ordered pool on `dev` (nullptr: creation failed, callers fall back to the default pool)
cudaMemPool_t device_pool(int dev)
{
    if (dev < 0 || dev >= MAX_DEVICES)
        return nullptr;
    std::call_once(g_pool_once[dev], [dev]() {
        cudaMemPoolProps props = {};
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = dev;
        cudaMemPool_t pool = nullptr;
        if (cudaMemPoolCreate(&pool, &props) == cudaSuccess) {
            const char* e = getenv("FIMEXCU_DEVICE_POOL_KEEP_MB");
            unsigned long long keep = (e ? strtoull(e, nullptr, 10) : 2048ull) << 20;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
            g_pool[dev] = pool;
        }
        cudaGetLastError();
    });
    return g_pool[dev];
}

// ---- pinned host pool: size classes, free lists; cudaFreeHost never runs on the hot path ----
struct HostPool {
    std::mutex m;
    std::unordered_map<void*, size_t> live;  // handed out: pointer -> class size
    std::multimap<size_t, void*> idle;       // class size -> pointer
    size_t idle_bytes = 0;
    size_t max_idle = []() {
        const char* e = getenv("FIMEXCU_HOST_POOL_KEEP_MB");
        return static_cast<size_t>(e ? strtoull(e, nullptr, 10) : 8192ull) << 20;
    }();
};
HostPool& host_pool()
{
    static HostPool* p = new HostPool; // never destroyed: the CUDA context may be gone at static destruction time
    return *p;
}
size_t host_class(size_t bytes)
{
    if (bytes <= (1u << 20)) {
        size_t c = 4096;
        while (c < bytes)
            c <<= 1;
        return c;
    }
    const size_t step = 2u << 20; // 2 MiB steps above 1 MiB: slabs of one variable land in the same class call after call
    return (bytes + step - 1) / step * step;
}
} // namespace

void* host_pool_alloc(size_t bytes)
{
    HostPool& hp = host_pool();
    const size_t cls = host_class(bytes ? bytes : 1);
    {
        std::lock_guard<std::mutex> g(hp.m);
        auto it = hp.idle.lower_bound(cls);
        if (it != hp.idle.end() && it->first <= cls + cls / 8) {
            void* p = it->second;
            hp.live[p] = it->first;
            hp.idle_bytes -= it->first;
            hp.idle.erase(it);
            return p;
        }
    }
    void* p = nullptr;
    // portable: one process may drive several devices from its host threads
    if (!cuda_ok(cudaHostAlloc(&p, cls, cudaHostAllocPortable), "cudaHostAlloc", __FILE__, __LINE__))
        return nullptr;
    std::lock_guard<std::mutex> g(hp.m);
    hp.live[p] = cls;
    return p;
}

void host_pool_free(void* p)
{
    if (!p)
        return;
    HostPool& hp = host_pool();
    size_t cls = 0;
    {
        std::lock_guard<std::mutex> g(hp.m);
        auto it = hp.live.find(p);
        if (it == hp.live.end())
            return; // not ours (or freed twice): leave it alone
        cls = it->second;
        hp.live.erase(it);
        if (hp.idle_bytes + cls <= hp.max_idle) {
            hp.idle.emplace(cls, p);
            hp.idle_bytes += cls;
            return;
        }
    }
    cudaFreeHost(p);
}

static void host_pool_trim()
{
    HostPool& hp = host_pool();
    std::multimap<size_t, void*> drop;
    {
        std::lock_guard<std::mutex> g(hp.m);
        drop.swap(hp.idle);
        hp.idle_bytes = 0;
    }
    for (auto& kv : drop)
        cudaFreeHost(kv.second);
}

bool is_pageable_host(const void* p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return true;
    }
    return a.type == cudaMemoryTypeUnregistered;
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
    device_pool(dev);
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
    this->bytes = bytes;
    int dev = -1;
    cudaGetDevice(&dev);
    cudaMemPool_t pool = device_pool(dev);
    if (pool)
        return cuda_ok(cudaMallocFromPoolAsync(&p, bytes, pool, s), "cudaMallocFromPoolAsync", __FILE__, __LINE__);
    return cuda_ok(cudaMallocAsync(&p, bytes, s), "cudaMallocAsync", __FILE__, __LINE__);
}

bool DevBuf::alloc_sync(size_t bytes)
{
    release();
    if (bytes == 0)
        bytes = 16;
    async = false;
    this->bytes = bytes;
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
    bytes = 0;
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
    if (!fxcu::require_device())
        return nullptr;
    return fxcu::host_pool_alloc(bytes);
}

void fimexcu_host_free(void* p) { fxcu::host_pool_free(p); }

void fimexcu_host_pool_trim(void) { fxcu::host_pool_trim(); }

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

int fimexcu_memcpy_d2d(void* dst_dev, const void* src_dev, size_t bytes)
{
    FXCU_CUDA(cudaMemcpy(dst_dev, src_dev, bytes, cudaMemcpyDeviceToDevice));
    return FIMEXCU_OK;
}

int fimexcu_memcpy_d2h(void* dst_host, const void* src_dev, size_t bytes)
{
    FXCU_CUDA(cudaMemcpy(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost));
    return FIMEXCU_OK;
}

// dense host [nz][ny][row] <-> pitched device slab: one 2-D copy when the levels follow each other at the row pitch, else one per level
static int copy_slab(void* dst, size_t dpitch, size_t dlevel, const void* src, size_t spitch, size_t slevel, size_t width, size_t ny, size_t nz,
                     cudaMemcpyKind kind)
{
    if (width == 0 || ny == 0 || nz == 0)
        return FIMEXCU_OK;
    if (dlevel == dpitch * ny && slevel == spitch * ny) {
        FXCU_CUDA(cudaMemcpy2D(dst, dpitch, src, spitch, width, ny * nz, kind));
        return FIMEXCU_OK;
    }
    for (size_t z = 0; z < nz; ++z)
        FXCU_CUDA(cudaMemcpy2D(static_cast<char*>(dst) + z * dlevel, dpitch, static_cast<const char*>(src) + z * slevel, spitch, width, ny, kind));
    return FIMEXCU_OK;
}

int fimexcu_memcpy2d_h2d(void* dst_dev, size_t dst_pitch_bytes, const void* src_host, size_t src_pitch_bytes, size_t width_bytes, size_t ny, size_t nz,
                         size_t dst_level_bytes)
{
    return copy_slab(dst_dev, dst_pitch_bytes, dst_level_bytes, src_host, src_pitch_bytes, src_pitch_bytes * ny, width_bytes, ny, nz, cudaMemcpyHostToDevice);
}

int fimexcu_memcpy2d_d2h(void* dst_host, size_t dst_pitch_bytes, const void* src_dev, size_t src_pitch_bytes, size_t width_bytes, size_t ny, size_t nz,
                         size_t src_level_bytes)
{
    return copy_slab(dst_host, dst_pitch_bytes, dst_pitch_bytes * ny, src_dev, src_pitch_bytes, src_level_bytes, width_bytes, ny, nz, cudaMemcpyDeviceToHost);
}

int fimexcu_synchronize(void)
{
    FXCU_CUDA(cudaDeviceSynchronize());
    return FIMEXCU_OK;
}

} // extern "C"
