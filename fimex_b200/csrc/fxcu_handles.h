// fxcu_handles.h -- the opaque handle types of include/fimexcu.h.
#pragma once
#include "fxcu_common.h"

#include <vector>

namespace fxcu {
enum CacheKind { KIND_BILINEAR = 0, KIND_BICUBIC = 1, KIND_NN = 2 };
// what the staged kernels need to know about a target tile before any of its table entries has arrived
enum TileFlags { TILE_HAS_DATA = 1, TILE_FITS = 2, TILE_ALL_INTERIOR = 4 };
}

struct fimexcu_cache {
    int method = 0;
    int kind = 0;
    int device = 0;
    size_t inX = 0, inY = 0, outX = 0, outY = 0; // inX/inY: the (possibly reduced) source window
    bool reduced = false;
    size_t xMin = 0, yMin = 0;
    bool fast_bicubic = false; // fimexcu_cache_set_fast_bicubic: f32 FMA arithmetic in the staged bicubic kernel (1e-5 contract) instead of the reference's rounding chain
    fxcu::DevBuf off;    // int32 [outX*outY]
    fxcu::DevBuf wx, wy; // f32 (bilinear) or f64 (bicubic) [outX*outY]
    unsigned tiles_with_data = 0, tiles_fitting = 0; // 32 x 8 target tiles: with defined cells / whose footprint fits the TMA box
    fxcu::DevBuf tile_meta; // int4 per 32 x 8 target tile (row-major over tiles): {box x0 (sector aligned), box y0, TILE_* flags, 0}
    // Where a result can hold data at all: target rows grouped into a few bands, each with the column span [x0, x1) outside which
    // every row of the band is undefined whatever the input (table entry -1).  The host entry points bring back only these
    // rectangles and write the undefined value into the rest on the host (S->T1: 36.6 % of the target lies outside the source).
    struct RowBand {
        int y0, y1, x0, x1; // rows [y0, y1), columns [x0, x1); x0 == x1: no row of the band holds data
    };
    std::vector<RowBand> bands;  // covers [0, outY) in order; empty: not analysed (treat everything as possibly defined)
    double band_fraction = 1.0;  // area of the bands' rectangles / area of the target
    // copies of the tables on other devices (fimexcu_cache_replicate), indexed by device ordinal; owned by this handle
    std::vector<fimexcu_cache*> replicas;
    ~fimexcu_cache()
    {
        for (fimexcu_cache* r : replicas)
            delete r;
    }
};

struct fimexcu_vecrot {
    size_t ox = 0, oy = 0;
    int device = 0;
    fxcu::DevBuf c, s, deg; // f64 [ox*oy] each: cos, sin, angle in degree
    std::vector<fimexcu_vecrot*> replicas;
    ~fimexcu_vecrot()
    {
        for (fimexcu_vecrot* r : replicas)
            delete r;
    }
};

namespace fxcu {
// fxcu_apply.cu: cache construction from DEVICE index tables (K2/K3 tail: reduced domain, table entries)
fimexcu_cache* cache_build(int method, const double* d_px, const double* d_py, size_t inX, size_t inY, size_t outX, size_t outY, cudaStream_t s);
// fxcu_axes.cu: mifi_points2position on device points with the axis already on the device
int points2position_dev(double* d_points, size_t n, const double* d_axis, const double* axis_host, int num, int axis_type, cudaStream_t s);

// the table set of a handle that lives on the calling thread's current device: the handle itself, one of its replicas, or
// nullptr with the error message set (template over fimexcu_cache / fimexcu_vecrot)
template <class H> const H* on_current_device(const H* h, const char* who)
{
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        cudaGetLastError();
        set_error(std::string(who) + ": no current CUDA device");
        return nullptr;
    }
    if (dev == h->device)
        return h;
    if (dev >= 0 && static_cast<size_t>(dev) < h->replicas.size() && h->replicas[dev])
        return h->replicas[dev];
    set_error(std::string(who) + ": the handle's tables live on device " + std::to_string(h->device) + ", the current device is " + std::to_string(dev) +
              " (call fimexcu_set_device, or replicate the handle onto this device first)");
    return nullptr;
}

// copy one device buffer of a handle onto the current device (peer copy; no peer mapping is set up or needed)
inline bool replicate_buf(DevBuf& dst, int dstDev, const DevBuf& src, int srcDev, cudaStream_t s)
{
    if (!src.p)
        return true;
    if (!dst.alloc_sync(src.bytes))
        return false;
    return cuda_ok(cudaMemcpyPeerAsync(dst.p, dstDev, src.p, srcDev, src.bytes, s), "cudaMemcpyPeerAsync", __FILE__, __LINE__);
}
}
