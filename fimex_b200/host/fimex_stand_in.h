// fimex_stand_in.h -- the handful of reference declarations CudaCachedInterpolation.h needs, for building it WITHOUT the
// reference tree (the GPU box has no /root/reference, and the reference cannot be linked here anyway: PROJ, NetCDF, libxml2
// and udunits2 are absent).  Inside fimex these come from the real headers; the shapes below follow
//   include/fimex/SharedArray.h:39-48           shared_array<T> = std::shared_ptr<T[]>, make_shared_array
//   include/fimex/CDMException.h:36-41          CDMException : std::runtime_error
//   include/fimex/CachedInterpolation.h:42-112  ReducedInterpolationDomain, CachedInterpolationInterface (without the CDM slicing helpers)
//   include/fimex/reproject.h:34-47             reproject::Matrix
//   include/fimex/CDMInterpolator.h:45-52       InterpolatorProcess2d
// Define FIMEXCU_WITH_FIMEX_HEADERS to compile against the real headers instead.
#pragma once
#ifdef FIMEXCU_WITH_FIMEX_HEADERS
#include "fimex/CDMException.h"
#include "fimex/CDMInterpolator.h"
#include "fimex/CachedInterpolation.h"
#include "fimex/CachedVectorReprojection.h"
#include "fimex/SharedArray.h"
#include "fimex/reproject.h"
#else
#include <cstddef>
#include <memory>
#include <stdexcept>
#include <string>

namespace MetNoFimex {

template <typename T> using shared_array = std::shared_ptr<T[]>;
template <typename T> shared_array<T> make_shared_array(size_t n) { return shared_array<T>(new T[n]); }

class CDMException : public std::runtime_error {
public:
    explicit CDMException(const std::string& msg) : std::runtime_error("CDMException: " + msg) {}
};

struct ReducedInterpolationDomain {
    std::string xDim, yDim;
    size_t xMin, yMin;
    ReducedInterpolationDomain(const std::string& xdim, const std::string& ydim, size_t xmin, size_t ymin) : xDim(xdim), yDim(ydim), xMin(xmin), yMin(ymin) {}
};
typedef std::shared_ptr<ReducedInterpolationDomain> ReducedInterpolationDomain_p;

class CachedInterpolationInterface {
public:
    CachedInterpolationInterface(const std::string& xDimName, const std::string& yDimName, size_t inX, size_t inY, size_t outX, size_t outY)
        : _xDimName(xDimName), _yDimName(yDimName), inX(inX), inY(inY), outX(outX), outY(outY) {}
    virtual ~CachedInterpolationInterface() {}
    virtual shared_array<float> interpolateValues(shared_array<float> inData, size_t size, size_t& newSize) const = 0;
    size_t getInX() const { return inX; }
    size_t getInY() const { return inY; }
    size_t getOutX() const { return outX; }
    size_t getOutY() const { return outY; }
    ReducedInterpolationDomain_p reducedDomain() const { return reducedDomain_; }

private:
    std::string _xDimName, _yDimName;

protected:
    size_t inX, inY, outX, outY;
    ReducedInterpolationDomain_p reducedDomain_;
};
typedef std::shared_ptr<CachedInterpolationInterface> CachedInterpolationInterface_p;

class InterpolatorProcess2d {
public:
    virtual void operator()(float* array, size_t nx, size_t ny) = 0;
    virtual ~InterpolatorProcess2d() {}
};

namespace reproject {
struct Matrix {
    enum { stride = 3 };
    Matrix(int sizex, int sizey) : size_x(sizex), size_y(sizey), matrix(make_shared_array<double>(static_cast<size_t>(sizex) * sizey * stride)) {}
    const double* mtx() const { return matrix.get(); }
    double* mtx() { return matrix.get(); }
    int size_x, size_y;
    shared_array<double> matrix;
};
typedef std::shared_ptr<const Matrix> Matrix_cp;
} // namespace reproject
} // namespace MetNoFimex
#endif
