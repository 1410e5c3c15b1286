// CudaInterpolatorProcess2d.h -- the pre/post-process functors of CDMInterpolator over libfimexcu.so: drop-in subclasses of
// MetNoFimex::InterpolatorProcess2d (include/fimex/CDMInterpolator.h:45-84) with the constructors of InterpolatorFill2d,
// InterpolatorCreepFill2d and InterpolatorCreepFillVal2d, whose operator() bodies (src/CDMInterpolator.cc:76-92) call
// mifi_fill2d_f / mifi_creepfill2d_f / mifi_creepfillval2d_f.  Results are bit-identical; errors become CDMException.
//
// operator()(array, nx, ny) works on ONE host level, like the reference's (processArray_, src/CDMInterpolator.cc:163-186, calls it
// once per level inside an OpenMP loop).  processSlab(array, size, nx, ny) is the whole of processArray_ for one functor: all
// levels go to the device in one copy and are filled concurrently, one CTA per level.
#pragma once
#include "fimex_stand_in.h"
#include "../../include/fimexcu.h"

#include <vector>

namespace MetNoFimex {

class CudaInterpolatorProcess2d : public InterpolatorProcess2d {
public:
    // processArray_ for this functor alone: array holds size / (nx*ny) levels
    void processSlab(float* array, size_t size, size_t nx, size_t ny)
    {
        const size_t layer = nx * ny;
        if (layer == 0 || size == 0)
            return;
        if (size % layer != 0)
            throw CDMException("processSlab: size is not a multiple of nx*ny");
        const size_t nz = size / layer, bytes = size * sizeof(float);
        float* d = static_cast<float*>(fimexcu_dev_alloc(bytes));
        if (!d)
            throw CDMException(fimexcu_last_error());
        int rc = fimexcu_memcpy_h2d(d, array, bytes);
        if (rc == FIMEXCU_OK)
            rc = runDevice(d, nx, ny, nz);
        if (rc == FIMEXCU_OK)
            rc = fimexcu_memcpy_d2h(array, d, bytes);
        fimexcu_dev_free(d);
        if (rc != FIMEXCU_OK)
            throw CDMException(fimexcu_last_error());
    }

protected:
    virtual int runDevice(float* d_array, size_t nx, size_t ny, size_t nz) = 0;
    static void check(int rc)
    {
        if (rc != FIMEXCU_OK)
            throw CDMException(fimexcu_last_error());
    }
};

// include/fimex/CDMInterpolator.h:53-62, src/CDMInterpolator.cc:76-80
class CudaInterpolatorFill2d : public CudaInterpolatorProcess2d {
public:
    CudaInterpolatorFill2d(float relaxCrit, float corrEff, size_t maxLoop) : relaxCrit_(relaxCrit), corrEff_(corrEff), maxLoop_(maxLoop) {}
    void operator()(float* array, size_t nx, size_t ny) override
    {
        size_t nChanged;
        check(fimexcu_fill2d_f(nx, ny, array, relaxCrit_, corrEff_, maxLoop_, &nChanged));
    }

protected:
    int runDevice(float* d, size_t nx, size_t ny, size_t nz) override
    {
        return fimexcu_fill2d_dev_f(nx, ny, nz, d, relaxCrit_, corrEff_, maxLoop_, nullptr, nullptr);
    }

private:
    float relaxCrit_, corrEff_;
    size_t maxLoop_;
};

// include/fimex/CDMInterpolator.h:64-72, src/CDMInterpolator.cc:82-86
class CudaInterpolatorCreepFill2d : public CudaInterpolatorProcess2d {
public:
    CudaInterpolatorCreepFill2d(unsigned short repeat, char setWeight) : repeat_(repeat), setWeight_(setWeight) {}
    void operator()(float* array, size_t nx, size_t ny) override
    {
        size_t nChanged;
        check(fimexcu_creepfill2d_f(nx, ny, array, repeat_, setWeight_, &nChanged));
    }

protected:
    int runDevice(float* d, size_t nx, size_t ny, size_t nz) override
    {
        return fimexcu_creepfill2d_dev_f(nx, ny, nz, d, 1, 0.f, repeat_, setWeight_, nullptr, nullptr);
    }

private:
    unsigned short repeat_;
    char setWeight_;
};

// include/fimex/CDMInterpolator.h:74-83, src/CDMInterpolator.cc:88-92
class CudaInterpolatorCreepFillVal2d : public CudaInterpolatorProcess2d {
public:
    CudaInterpolatorCreepFillVal2d(unsigned short repeat, char setWeight, float defaultValue)
        : repeat_(repeat), setWeight_(setWeight), defVal_(defaultValue) {}
    void operator()(float* array, size_t nx, size_t ny) override
    {
        size_t nChanged;
        check(fimexcu_creepfillval2d_f(nx, ny, array, defVal_, repeat_, setWeight_, &nChanged));
    }

protected:
    int runDevice(float* d, size_t nx, size_t ny, size_t nz) override
    {
        return fimexcu_creepfill2d_dev_f(nx, ny, nz, d, 0, defVal_, repeat_, setWeight_, nullptr, nullptr);
    }

private:
    unsigned short repeat_;
    char setWeight_;
    float defVal_;
};

} // namespace MetNoFimex
