// CudaSlab.h -- a slab that stays on the GPU between chained readers: the device-side analogue of fimex's Data
// (include/fimex/Data.h:48-166) for the hot path.
//
// In the reference every CDMReader hands its neighbour a DataPtr in host memory: the interpolator returns a host array
// (src/CDMInterpolator.cc:219-263), the vertical interpolator wrapped around it pulls the regridded field AND the regridded surface
// pressure through getDataSlice() (src/CDMVerticalInterpolator.cc:363-364, 478-486; fimex.cc:862-863 fixes the order horizontal ->
// vertical), and the writer finally takes the bytes.  With the kernels at ~0.2 ms per slab and PCIe at ~15 ms, a chain that went
// through host memory at every link would spend all its time on the bus.  A CudaSlab carries what the next link needs instead:
// the device pointer, the element type, the logical shape [nz][ny][nx] and the layout (row pitch and level stride in elements,
// so the padded rows of the staged kernels pass through unchanged), the fill value, and the device it lives on.  It converts to
// the reference's types only at the ends of the chain:
//     CudaSlab::upload(shared_array<float>|shared_array<short>, ...)      reader side (one H2D, pageable or pinned source)
//     slab->asFloat() / slab->asShort()                                    writer side (one D2H into pooled pinned memory)
// and in between
//     CudaCachedInterpolation::interpolateSlab(slab)                       K4/K5/K6 (+ fused unpack for short slabs)
//     CudaCachedVectorReprojection::reprojectSlabs(u, v)                   K7
//     cudaVerticalToFixedLevels(...)                                       K14+K15 with the pressure formed from a ps SLAB
// all enqueue on one stream per slab chain and never synchronise.  Header only; link with -lfimexcu.
#pragma once

#include "CudaCachedInterpolation.h"

#include <memory>
#include <vector>

namespace MetNoFimex {

class CudaSlab;
typedef std::shared_ptr<CudaSlab> CudaSlab_p;

class CudaSlab {
public:
    enum Type { FLOAT = 0, SHORT = 1 }; // CDM_FLOAT / CDM_SHORT (include/fimex/CDMDataType.h:34-46)

    // an uninitialised slab of the given shape and layout on the calling thread's device
    static CudaSlab_p allocate(Type type, size_t nz, size_t ny, size_t nx, size_t rowPitch = 0, size_t levelStride = 0)
    {
        CudaSlab_p s(new CudaSlab());
        s->type_ = type;
        s->nz_ = nz;
        s->ny_ = ny;
        s->nx_ = nx;
        s->pitch_ = rowPitch ? rowPitch : nx;
        s->level_ = levelStride ? levelStride : s->pitch_ * ny;
        s->device_ = cuda_glue::bind_thread();
        s->ptr_ = fimexcu_dev_alloc(std::max<size_t>(1, s->level_ * nz) * s->elementSize());
        if (!s->ptr_)
            throw CDMException(fimexcu_last_error());
        return s;
    }
    // reader side: dense host array [nz][ny][nx] -> device slab with the given layout (0 = dense)
    template <class T> static CudaSlab_p upload(const shared_array<T>& host, size_t nz, size_t ny, size_t nx, size_t rowPitch = 0, size_t levelStride = 0)
    {
        CudaSlab_p s = allocate(sizeof(T) == 4 ? FLOAT : SHORT, nz, ny, nx, rowPitch, levelStride);
        if (s->dense()) {
            if (fimexcu_memcpy_h2d(s->ptr_, host.get(), nz * ny * nx * sizeof(T)) != FIMEXCU_OK)
                throw CDMException(fimexcu_last_error());
        } else {
            if (fimexcu_memcpy2d_h2d(s->ptr_, s->pitch_ * sizeof(T), host.get(), nx * sizeof(T), nx * sizeof(T), ny, nz, s->level_ * sizeof(T)) != FIMEXCU_OK)
                throw CDMException(fimexcu_last_error());
        }
        return s;
    }
    ~CudaSlab() { fimexcu_dev_free(ptr_); }
    CudaSlab(const CudaSlab&) = delete;
    CudaSlab& operator=(const CudaSlab&) = delete;

    Type type() const { return type_; }
    size_t elementSize() const { return type_ == FLOAT ? 4 : 2; }
    size_t size() const { return nz_ * ny_ * nx_; } // Data::size(): logical number of values
    size_t nz() const { return nz_; }
    size_t ny() const { return ny_; }
    size_t nx() const { return nx_; }
    size_t rowPitch() const { return pitch_; }
    size_t levelStride() const { return level_; }
    bool dense() const { return pitch_ == nx_ && level_ == nx_ * ny_; }
    int device() const { return device_; }
    void* devicePtr() const { return ptr_; }
    float* floatPtr() const { return static_cast<float*>(ptr_); }

    // writer side: Data::asFloat() / asShort() -- dense host copy in pooled pinned memory
    shared_array<float> asFloat() const
    {
        if (type_ != FLOAT)
            throw CDMException("CudaSlab::asFloat on a packed slab: unpack it first");
        return download<float>();
    }
    shared_array<short> asShort() const
    {
        if (type_ != SHORT)
            throw CDMException("CudaSlab::asShort on a float slab: pack it first");
        return download<short>();
    }

private:
    CudaSlab() = default;
    template <class T> shared_array<T> download() const
    {
        if (fimexcu_set_device(device_) != FIMEXCU_OK)
            throw CDMException(fimexcu_last_error());
        T* raw = static_cast<T*>(fimexcu_host_alloc(std::max<size_t>(1, size()) * sizeof(T)));
        if (!raw)
            throw CDMException(fimexcu_last_error());
        shared_array<T> out(raw, [](T* p) { fimexcu_host_free(p); });
        const int rc = dense() ? fimexcu_memcpy_d2h(raw, ptr_, size() * sizeof(T))
                               : fimexcu_memcpy2d_d2h(raw, nx_ * sizeof(T), ptr_, pitch_ * sizeof(T), nx_ * sizeof(T), ny_, nz_, level_ * sizeof(T));
        if (rc != FIMEXCU_OK)
            throw CDMException(fimexcu_last_error());
        return out;
    }
    Type type_ = FLOAT;
    size_t nz_ = 0, ny_ = 0, nx_ = 0, pitch_ = 0, level_ = 0;
    int device_ = 0;
    void* ptr_ = nullptr;
};

// ---- the links of the chain ------------------------------------------------------------------------------------------------------

// getInputDataSlice + data2InterpolationArray, reader side: the (already reduced) input window goes up in the padded layout the
// staged kernels want; a short slab goes up as it is (2 bytes per value over PCIe) and is unpacked on the device
inline CudaSlab_p uploadInterpolationInput(const CudaCachedInterpolation& ci, const shared_array<float>& in, size_t size)
{
    size_t ip, il, op, ol;
    if (fimexcu_cache_padded_layout(ci.handle(), &ip, &il, &op, &ol) != FIMEXCU_OK)
        throw CDMException(fimexcu_last_error());
    const size_t nz = size / (ci.getInX() * ci.getInY());
    return CudaSlab::upload(in, nz, ci.getInY(), ci.getInX(), ip, il);
}

// interpolateValues on a device slab (float, any layout).  badIn / fillOut fuse data2InterpolationArray / interpolationArray2Data
// (src/CDMInterpolator.cc:140-151); NaN = off.  The result has the padded target layout.
inline CudaSlab_p interpolateSlab(const CudaCachedInterpolation& ci, const CudaSlab& in, float badIn = NAN, float fillOut = NAN)
{
    if (in.type() != CudaSlab::FLOAT || in.nx() != ci.getInX() || in.ny() != ci.getInY())
        throw CDMException("interpolateSlab: the slab is not the cache's input window");
    cuda_glue::bind_thread();
    size_t ip, il, op, ol;
    if (fimexcu_cache_padded_layout(ci.handle(), &ip, &il, &op, &ol) != FIMEXCU_OK)
        throw CDMException(fimexcu_last_error());
    CudaSlab_p out = CudaSlab::allocate(CudaSlab::FLOAT, in.nz(), ci.getOutY(), ci.getOutX(), op, ol);
    if (fimexcu_cache_apply_dev_strided_f(ci.handle(), in.floatPtr(), in.rowPitch(), in.levelStride(), in.nz(), out->floatPtr(), op, ol, badIn, fillOut,
                                          nullptr) != FIMEXCU_OK)
        throw CDMException(std::string("error during interpolation: ") + fimexcu_last_error());
    return out;
}

// the vector branch of getDataSlice (src/CDMInterpolator.cc:232-259) on device slabs: both components interpolated and rotated in one pass
inline std::pair<CudaSlab_p, CudaSlab_p> interpolateVectorSlabs(const CudaCachedInterpolation& ci, const CudaCachedVectorReprojection& cvr, const CudaSlab& u,
                                                              const CudaSlab& v)
{
    if (u.rowPitch() != v.rowPitch() || u.levelStride() != v.levelStride() || u.nz() != v.nz())
        throw CDMException("interpolateVectorSlabs: u and v differ in shape or layout");
    cuda_glue::bind_thread();
    size_t ip, il, op, ol;
    if (fimexcu_cache_padded_layout(ci.handle(), &ip, &il, &op, &ol) != FIMEXCU_OK)
        throw CDMException(fimexcu_last_error());
    CudaSlab_p uo = CudaSlab::allocate(CudaSlab::FLOAT, u.nz(), ci.getOutY(), ci.getOutX(), op, ol);
    CudaSlab_p vo = CudaSlab::allocate(CudaSlab::FLOAT, u.nz(), ci.getOutY(), ci.getOutX(), op, ol);
    if (fimexcu_cache_apply_uv_dev_strided_f(ci.handle(), cvr.handle(), u.floatPtr(), v.floatPtr(), u.rowPitch(), u.levelStride(), u.nz(), uo->floatPtr(),
                                             vo->floatPtr(), op, ol, NAN, NAN, NAN, NAN, nullptr) != FIMEXCU_OK)
        throw CDMException(std::string("error during interpolation: ") + fimexcu_last_error());
    return {uo, vo};
}

// CDMProcessor's rotation of a vector to another projection's axes (src/CDMProcessor.cc:114-157 -> CachedVectorReprojection::
// reprojectValues) on device slabs of whatever row pitch, in place; the pad elements of padded rows are not touched
inline void rotateVectorSlabs(const CudaCachedVectorReprojection& cvr, CudaSlab& u, CudaSlab& v)
{
    if (u.rowPitch() != v.rowPitch() || u.levelStride() != v.levelStride() || u.nz() != v.nz() || u.type() != CudaSlab::FLOAT || v.type() != CudaSlab::FLOAT)
        throw CDMException("rotateVectorSlabs: u and v differ in shape or layout, or are not float slabs");
    cuda_glue::bind_thread();
    if (fimexcu_vecrot_apply_dev_strided_f(cvr.handle(), u.floatPtr(), v.floatPtr(), u.rowPitch(), u.levelStride(), u.nz(), nullptr) != FIMEXCU_OK)
        throw CDMException(std::string("error during vector rotation: ") + fimexcu_last_error());
}

// CDMVerticalInterpolator::getLevelDataSlice for a hybrid-sigma source and fixed target levels (src/CDMVerticalInterpolator.cc:
// 356-565 with HybridSigmaApToPressureConverter.cc:71-100): field [nLevels x nOther][ny][nx] laid out level-major (hybrid, member,
// y, x) as CDM orders it, surface pressure [nOther][ny][nx] in the SAME layout -- both typically straight out of interpolateSlab().
// The pad elements of padded rows are skipped (neither read nor written).  ap in the unit of the target levels (hPa), ps * psScale too.
inline CudaSlab_p cudaVerticalToFixedLevels(int method, const CudaSlab& field, size_t nLevels, const std::vector<double>& ap, const std::vector<double>& b,
                                            const CudaSlab& ps, double psScale, const std::vector<double>& levels, double validMin = NAN, double validMax = NAN)
{
    if (field.nz() % nLevels != 0 || ap.size() != nLevels || b.size() != nLevels)
        throw CDMException("cudaVerticalToFixedLevels: level count and coefficients disagree");
    const size_t nOther = field.nz() / nLevels;
    if (ps.nz() != nOther || ps.rowPitch() != field.rowPitch() || ps.levelStride() != field.levelStride() || ps.ny() != field.ny())
        throw CDMException("cudaVerticalToFixedLevels: the surface pressure slab must share the field's horizontal layout");
    cuda_glue::bind_thread();
    const size_t nInner = nOther * field.levelStride(); // (member, y, padded x) is one inner index: column = everything but the level
    CudaSlab_p out = CudaSlab::allocate(CudaSlab::FLOAT, levels.size() * nOther, field.ny(), field.nx(), field.rowPitch(), field.levelStride());
    // rows: only where the level stride is a whole number of rows (the padded layout's is) and the slab is small enough for the form
    const bool rows = field.rowPitch() > field.nx() && field.levelStride() % field.rowPitch() == 0 && nInner < (size_t(1) << 32);
    if (fimexcu_vertical_hybrid_pressure_rows_dev_f(method, field.floatPtr(), 1, nLevels, nInner, ap.data(), b.data(), ps.floatPtr(), 1, psScale, levels.data(),
                                                    levels.size(), validMin, validMax, NAN, NAN, out->floatPtr(), rows ? field.nx() : 0,
                                                    rows ? field.rowPitch() : 0, nullptr) != FIMEXCU_OK)
        throw CDMException(fimexcu_last_error());
    return out;
}

// packed slabs at the ends of a chain: Data::asFloat + mifi_bad2nanf / ScaleValue<float,short> on the device
inline CudaSlab_p unpackSlab(const CudaSlab& raw, double fillValue)
{
    if (raw.type() != CudaSlab::SHORT)
        throw CDMException("unpackSlab: not a packed slab");
    cuda_glue::bind_thread();
    CudaSlab_p out = CudaSlab::allocate(CudaSlab::FLOAT, raw.nz(), raw.ny(), raw.nx(), raw.rowPitch(), raw.levelStride());
    if (fimexcu_unpack_dev_f(FIMEXCU_DTYPE_I16, raw.devicePtr(), raw.levelStride() * raw.nz(), fillValue, out->floatPtr(), nullptr) != FIMEXCU_OK)
        throw CDMException(fimexcu_last_error());
    return out;
}
inline CudaSlab_p packSlab(const CudaSlab& f, double fillValue)
{
    if (f.type() != CudaSlab::FLOAT)
        throw CDMException("packSlab: not a float slab");
    cuda_glue::bind_thread();
    CudaSlab_p out = CudaSlab::allocate(CudaSlab::SHORT, f.nz(), f.ny(), f.nx(), f.rowPitch(), f.levelStride());
    if (fimexcu_pack_dev_f(FIMEXCU_DTYPE_I16, f.floatPtr(), f.levelStride() * f.nz(), fillValue, out->devicePtr(), nullptr) != FIMEXCU_OK)
        throw CDMException(fimexcu_last_error());
    return out;
}

} // namespace MetNoFimex
