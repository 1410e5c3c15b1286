// CudaCachedInterpolation.h -- the reference-side glue: fimex's own interpolation-cache classes implemented on libfimexcu.so.
//
// Drop-in for (fimex 2.4.1):
//   createCachedInterpolation                         src/CachedInterpolation.cc:132-147
//   CachedInterpolation / CachedNNInterpolation       include/fimex/CachedInterpolation.h:124-189
//   CachedForwardInterpolation                        src/CachedForwardInterpolation.h:36-50
//   CachedVectorReprojection                          include/fimex/CachedVectorReprojection.h:38-63
// Same constructor arguments, same virtual interface, same exceptions (CDMException); CDMInterpolator needs no other change
// than calling createCudaCachedInterpolation() from its factory.  Header only; link with -lfimexcu.
#pragma once

#include "../../include/fimexcu.h"
#include "fimex_stand_in.h"

#include <atomic>
#include <cmath>
#include <cstdlib>
#include <limits>
#include <string>
#include <vector>

namespace MetNoFimex {

// Which GPUs the drop-in uses, and which of them serves the calling host thread.
// fimex is ONE process; its NetCDF writer pulls the time steps through `#pragma omp parallel for`
// (src/io/netcdf/NetCDF_CDMWriter.cc:599-618), so getDataSlice -- and with it interpolateValues -- is entered concurrently from
// the OpenMP threads, all on the same cache object.  Each host thread is bound to one device the first time it gets here
// (round robin over FIMEX_CUDA_DEVICES = "0,2,5", default: every sm_100 device), the caches keep a copy of their tables on
// each of those devices (fimexcu_cache_replicate), and the slabs of one variable are thereby dealt over the GPUs by the
// reference's own loop -- no new option, no MPI launch.
namespace cuda_glue {
inline const std::vector<int>& devices()
{
    static const std::vector<int> list = []() {
        std::vector<int> v;
        const int n = fimexcu_device_count();
        if (const char* e = std::getenv("FIMEX_CUDA_DEVICES")) {
            const char* p = e;
            while (*p) {
                char* end = nullptr;
                const long d = std::strtol(p, &end, 10);
                if (end == p)
                    break;
                if (d >= 0 && d < n)
                    v.push_back(static_cast<int>(d));
                p = (*end == ',') ? end + 1 : end;
            }
        }
        if (v.empty())
            for (int d = 0; d < n; ++d)
                v.push_back(d);
        return v;
    }();
    return list;
}
// binds the calling thread to its device (once) and returns the ordinal; throws when there is no usable GPU
inline int bind_thread()
{
    static std::atomic<unsigned> next{0};
    thread_local int mine = -1;
    if (mine < 0) {
        const std::vector<int>& d = devices();
        if (d.empty())
            throw CDMException("no usable sm_100 device (libfimexcu has no CPU fallback)");
        mine = d[next.fetch_add(1) % d.size()];
    }
    if (fimexcu_set_device(mine) != FIMEXCU_OK)
        throw CDMException(fimexcu_last_error());
    return mine;
}
} // namespace cuda_glue

class CudaCachedInterpolation : public CachedInterpolationInterface {
public:
    // px/py as handed to createCachedInterpolation: fractional source indices of every target cell.  Like the reference's
    // constructors (CachedInterpolation.cc:243-246) the shared tables are rebased in place when the domain is reduced.
    CudaCachedInterpolation(const std::string& xDimName, const std::string& yDimName, int method, shared_array<double> pointsOnXAxis,
                            shared_array<double> pointsOnYAxis, size_t inx, size_t iny, size_t outx, size_t outy, float badValue = NAN,
                            float fillValue = NAN)
        : CachedInterpolationInterface(xDimName, yDimName, inx, iny, outx, outy)
        , h_((cuda_glue::bind_thread(), fimexcu_cache_create(method, pointsOnXAxis.get(), pointsOnYAxis.get(), inx, iny, outx, outy)))
        , bad_(badValue)
        , fill_(fillValue)
    {
        if (!h_)
            throw CDMException(fimexcu_last_error());
        const std::vector<int>& devs = cuda_glue::devices(); // the tables on every device the writer threads will use
        if (fimexcu_cache_replicate(h_, static_cast<int>(devs.size()), devs.data()) != FIMEXCU_OK) {
            const std::string msg = fimexcu_last_error();
            fimexcu_cache_destroy(h_);
            throw CDMException(msg);
        }
        int reduced = 0;
        size_t xMin = 0, yMin = 0;
        fimexcu_cache_info(h_, &inX, &inY, nullptr, nullptr, &reduced, &xMin, &yMin);
        if (reduced) {
            reducedDomain_ = std::make_shared<ReducedInterpolationDomain>(xDimName, yDimName, xMin, yMin);
            if (method == FIMEXCU_INTERPOL_BILINEAR || method == FIMEXCU_INTERPOL_BICUBIC)
                for (size_t i = 0; i < outx * outy; ++i) {
                    pointsOnXAxis[i] -= xMin;
                    pointsOnYAxis[i] -= yMin;
                }
        }
    }
    ~CudaCachedInterpolation() override { fimexcu_cache_destroy(h_); }
    CudaCachedInterpolation(const CudaCachedInterpolation&) = delete;
    CudaCachedInterpolation& operator=(const CudaCachedInterpolation&) = delete;

    // CachedInterpolation.cc:173-209.  The result lives in pinned host memory (full-speed device-to-host copy) taken from the
    // library's pool -- a fresh cudaHostAlloc of a 767 MB result per call would cost more than the regridding -- and the
    // shared_array's deleter parks it there again, so callers see an ordinary shared_array<float>.  inData is whatever the
    // upstream reader allocated (pageable `new float[]`): the library moves it through pinned bounce buffers.
    // Re-entrant: called concurrently from the writer's OpenMP threads, each bound to its own GPU.
    shared_array<float> interpolateValues(shared_array<float> inData, size_t size, size_t& newSize) const override
    {
        cuda_glue::bind_thread();
        const size_t nz = (inX * inY != 0) ? size / (inX * inY) : 0;
        newSize = nz * outX * outY;
        float* raw = static_cast<float*>(fimexcu_host_alloc((newSize ? newSize : 1) * sizeof(float)));
        if (!raw)
            throw CDMException(fimexcu_last_error());
        shared_array<float> out(raw, [](float* p) { fimexcu_host_free(p); });
        if (fimexcu_cache_apply_f(h_, inData.get(), size, out.get(), &newSize, bad_, fill_) != FIMEXCU_OK)
            throw CDMException(std::string("error during interpolation: ") + fimexcu_last_error()); // CachedInterpolation.cc:202
        return out;
    }
    const fimexcu_cache* handle() const { return h_; }

private:
    fimexcu_cache* h_;
    float bad_, fill_;
};

class CudaCachedForwardInterpolation : public CachedInterpolationInterface {
public:
    CudaCachedForwardInterpolation(const std::string& xDimName, const std::string& yDimName, int funcType, shared_array<double> pointsOnXAxis,
                                   shared_array<double> pointsOnYAxis, size_t inx, size_t iny, size_t outx, size_t outy, bool levelCompat = true)
        : CachedInterpolationInterface(xDimName, yDimName, inx, iny, outx, outy)
        , device_(cuda_glue::bind_thread())
        , h_(fimexcu_forward_create(funcType, pointsOnXAxis.get(), pointsOnYAxis.get(), inx, iny, outx, outy))
        , compat_(levelCompat)
    {
        if (!h_)
            throw CDMException(fimexcu_last_error());
        int reduced = 0;
        size_t xMin = 0, yMin = 0;
        fimexcu_forward_info(h_, &inX, &inY, nullptr, nullptr, &reduced, &xMin, &yMin, nullptr);
        if (reduced)
            reducedDomain_ = std::make_shared<ReducedInterpolationDomain>(xDimName, yDimName, xMin, yMin);
    }
    ~CudaCachedForwardInterpolation() override { fimexcu_forward_destroy(h_); }
    shared_array<float> interpolateValues(shared_array<float> inData, size_t size, size_t& newSize) const override
    {
        const size_t nz = (inX * inY != 0) ? size / (inX * inY) : 0;
        newSize = nz * outX * outY;
        auto out = make_shared_array<float>(newSize ? newSize : 1);
        if (fimexcu_set_device(device_) != FIMEXCU_OK) // the segment lists live on the device the cache was built on
            throw CDMException(fimexcu_last_error());
        if (fimexcu_forward_apply_f(h_, inData.get(), size, out.get(), &newSize, NAN, NAN, compat_ ? 1 : 0) != FIMEXCU_OK)
            throw CDMException(std::string("error during interpolation: ") + fimexcu_last_error());
        return out;
    }

private:
    int device_;
    fimexcu_forward* h_;
    bool compat_;
};

// src/CachedInterpolation.cc:132-147 (+ the forward methods, which CDMInterpolator constructs directly, CDMInterpolator.cc:1271)
inline CachedInterpolationInterface_p createCudaCachedInterpolation(const std::string& xDimName, const std::string& yDimName, int method,
                                                                    shared_array<double> pointsOnXAxis, shared_array<double> pointsOnYAxis, size_t inX,
                                                                    size_t inY, size_t outX, size_t outY)
{
    switch (method) {
    case FIMEXCU_INTERPOL_BILINEAR:
    case FIMEXCU_INTERPOL_BICUBIC:
    case FIMEXCU_INTERPOL_NEAREST_NEIGHBOR:
    case FIMEXCU_INTERPOL_COORD_NN:
    case FIMEXCU_INTERPOL_COORD_NN_KD:
        return std::make_shared<CudaCachedInterpolation>(xDimName, yDimName, method, pointsOnXAxis, pointsOnYAxis, inX, inY, outX, outY);
    default:
        if (method >= FIMEXCU_INTERPOL_FORWARD_SUM && method <= FIMEXCU_INTERPOL_FORWARD_UNDEF_MIN)
            return std::make_shared<CudaCachedForwardInterpolation>(xDimName, yDimName, method, pointsOnXAxis, pointsOnYAxis, inX, inY, outX, outY);
        throw CDMException("unknown interpolation function: " + std::to_string(method));
    }
}

class CudaCachedVectorReprojection {
public:
    explicit CudaCachedVectorReprojection(reproject::Matrix_cp matrix)
        : matrix_(matrix), ox_(matrix->size_x), oy_(matrix->size_y)
        , h_((cuda_glue::bind_thread(), fimexcu_vecrot_create(matrix->mtx(), matrix->size_x, matrix->size_y)))
    {
        if (!h_)
            throw CDMException(fimexcu_last_error());
        const std::vector<int>& devs = cuda_glue::devices();
        if (fimexcu_vecrot_replicate(h_, static_cast<int>(devs.size()), devs.data()) != FIMEXCU_OK) {
            const std::string msg = fimexcu_last_error();
            fimexcu_vecrot_destroy(h_);
            throw CDMException(msg);
        }
    }
    ~CudaCachedVectorReprojection() { fimexcu_vecrot_destroy(h_); }
    CudaCachedVectorReprojection(const CudaCachedVectorReprojection&) = delete;
    // src/CachedVectorReprojection.cc:44-48
    void reprojectValues(shared_array<float>& uValues, shared_array<float>& vValues, size_t size) const
    {
        cuda_glue::bind_thread();
        if (fimexcu_vecrot_apply_f(h_, uValues.get(), vValues.get(), size) != FIMEXCU_OK)
            throw CDMException(fimexcu_last_error());
    }
    // src/CachedVectorReprojection.cc:50-54
    void reprojectDirectionValues(shared_array<float>& angles, size_t size) const
    {
        cuda_glue::bind_thread();
        if (fimexcu_vecrot_direction_f(h_, angles.get(), size) != FIMEXCU_OK)
            throw CDMException(fimexcu_last_error());
    }
    size_t getXSize() const { return ox_; }
    size_t getYSize() const { return oy_; }
    const fimexcu_vecrot* handle() const { return h_; }

private:
    reproject::Matrix_cp matrix_;
    size_t ox_, oy_;
    fimexcu_vecrot* h_;
};

} // namespace MetNoFimex
