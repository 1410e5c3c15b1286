// fxcu_common.h -- shared plumbing of libfimexcu.so: error reporting, RAII device buffers, launch helpers.
// Product code: never includes or links anything from oracle/.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <limits>
#include <string>

#include "../../include/fimexcu.h"

namespace fxcu {

// ---- per-thread error message (fimexcu_last_error) ----
void set_error(const std::string& msg);
void clear_error();
const char* last_error_cstr();

// returns false (and records the message) when a CUDA call failed
bool cuda_ok(cudaError_t e, const char* what, const char* file, int line);
#define FXCU_CUDA(call)                                              \
    do {                                                             \
        if (!::fxcu::cuda_ok((call), #call, __FILE__, __LINE__))     \
            return FIMEXCU_ERROR;                                    \
    } while (0)
#define FXCU_CUDA_NULL(call)                                         \
    do {                                                             \
        if (!::fxcu::cuda_ok((call), #call, __FILE__, __LINE__))     \
            return nullptr;                                          \
    } while (0)
#define FXCU_FAIL(msg)                \
    do {                              \
        ::fxcu::set_error(msg);       \
        return FIMEXCU_ERROR;         \
    } while (0)

// true when the current device can run this library (sm_100); records an error otherwise.
bool require_device();
int sm_count();

inline cudaStream_t as_stream(void* s) { return static_cast<cudaStream_t>(s); }

// stream-ordered scratch allocation from a pool that belongs to this library (one per device, created on first use; it keeps
// up to FIMEXCU_DEVICE_POOL_KEEP_MB, default 2048, of freed memory for the next call -- the device's default pool and its
// settings, which the host application may share, are left alone)
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    cudaStream_t stream = nullptr;
    bool async = false;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { release(); }
    bool alloc(size_t bytes, cudaStream_t s);        // stream ordered
    bool alloc_sync(size_t bytes);                   // plain cudaMalloc (for long-lived tables)
    void release();
    template <typename T> T* as() const { return static_cast<T*>(p); }
};

// pinned host memory from the library's pool (fimexcu_host_alloc / fimexcu_host_free)
void* host_pool_alloc(size_t bytes);
void host_pool_free(void* p);
// true when p is ordinary pageable host memory (not pinned / registered / managed / device)
bool is_pageable_host(const void* p);

// grid size for a grid-stride / one-shot kernel over n items with `threads` per block
inline unsigned blocks_for(size_t n, unsigned threads) { return static_cast<unsigned>((n + threads - 1) / threads); }

__device__ __forceinline__ float nan_f() { return __int_as_float(0x7fc00000); }

} // namespace fxcu
