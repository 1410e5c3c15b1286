// fxcu_pack.cu -- slab feed / sink for packed integer variables: the conversions the reference wraps
// around every operator of the path, as device passes, so that packed slabs cross PCIe in their file type and every operator
// (horizontal, vertical, rotation) can be chained on the device in f32.
//
// Reference behaviour restated here (fimex 2.4.1):
//   data2InterpolationArray      src/CDMInterpolator.cc:140-145: Data::asFloat() (data_caster<float,T>, MathUtils.h:60-96) of the RAW
//                                values, then mifi_bad2nanf(float(badValue)) (interpolation.c:1064-1072)
//   interpolationArray2Data      src/CDMInterpolator.cc:147-151: convertDataType(NaN,1,0,type,badValue,1,0) -> ScaleValue<float,T>
//                                (DataUtils.h:38-60): NaN -> T(badValue), else static_cast<T>(int(lround(double(v))))
//   getScaledDataSlice           src/CDMReader.cc:177-192: ScaleValue<T,double>(fill, scale, offset, NaN, 1, 0), narrowed with
//                                Data::asFloat() by the vertical converters / interpolator
// One thread per element, grid-stride, 16 B of input per thread where the type allows: pure streaming kernels.
#include "fxcu_common.h"

#include <cstdint>

namespace fxcu {

template <class T>
__global__ void __launch_bounds__(256) unpack_kernel(const T* __restrict__ in, size_t n, float bad, float* __restrict__ out)
{
    for (size_t i = blockIdx.x * 256ull + threadIdx.x; i < n; i += 256ull * gridDim.x) {
        float v = static_cast<float>(in[i]);
        if (v == bad)
            v = nan_f();
        out[i] = v;
    }
}

template <class T>
__global__ void __launch_bounds__(256) unpack_scaled_kernel(const T* __restrict__ in, size_t n, T fill, double scale, double offset,
                                                            float* __restrict__ out)
{
    for (size_t i = blockIdx.x * 256ull + threadIdx.x; i < n; i += 256ull * gridDim.x) {
        const T v = in[i];
        out[i] = (v == fill) ? nan_f() : __double2float_rn(__dadd_rn(__dmul_rn(scale, static_cast<double>(v)), offset));
    }
}

template <class T>
__global__ void __launch_bounds__(256) pack_kernel(const float* __restrict__ in, size_t n, T fill, T* __restrict__ out)
{
    for (size_t i = blockIdx.x * 256ull + threadIdx.x; i < n; i += 256ull * gridDim.x) {
        const float v = in[i];
        T r = fill;
        if (!isnan(v)) {
            float t = truncf(v);
            if (fabsf(__fsub_rn(v, t)) >= 0.5f) // lround: half away from zero; v - trunc(v) is exact
                t = __fadd_rn(t, copysignf(1.f, v));
            r = static_cast<T>(static_cast<int>(__float2ll_rz(t))); // int round(double) { return ::lround(num); } then static_cast<T>
        }
        out[i] = r;
    }
}

static unsigned stream_blocks(size_t n) { return static_cast<unsigned>(std::min<size_t>((n + 255) / 256, static_cast<size_t>(sm_count()) * 16)); }

template <class T>
static int unpack_t(const void* d_in, size_t n, double bad, float* d_out, cudaStream_t s)
{
    unpack_kernel<T><<<stream_blocks(n), 256, 0, s>>>(static_cast<const T*>(d_in), n, static_cast<float>(bad), d_out);
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}
template <class T>
static int unpack_scaled_t(const void* d_in, size_t n, double fill, double scale, double offset, float* d_out, cudaStream_t s)
{
    // ScaleValue's constructor: oldFill_(static_cast<IN>(oldFill)), oldScale/newScale with newScale = 1, (oldOffset - 0)/1
    unpack_scaled_kernel<T><<<stream_blocks(n), 256, 0, s>>>(static_cast<const T*>(d_in), n, static_cast<T>(fill), scale / 1.0, (offset - 0.0) / 1.0, d_out);
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}
template <class T>
static int pack_t(const float* d_in, size_t n, double bad, void* d_out, cudaStream_t s)
{
    pack_kernel<T><<<stream_blocks(n), 256, 0, s>>>(d_in, n, static_cast<T>(bad), static_cast<T*>(d_out));
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}

} // namespace fxcu

using namespace fxcu;

#define FXCU_DTYPE_SWITCH(dtype, CALL)                          \
    switch (dtype) {                                            \
    case FIMEXCU_DTYPE_I8: return CALL(signed char);            \
    case FIMEXCU_DTYPE_I16: return CALL(short);                 \
    case FIMEXCU_DTYPE_I32: return CALL(int);                   \
    case FIMEXCU_DTYPE_U8: return CALL(unsigned char);          \
    case FIMEXCU_DTYPE_U16: return CALL(unsigned short);        \
    case FIMEXCU_DTYPE_U32: return CALL(unsigned int);          \
    default: FXCU_FAIL("fimexcu_pack/unpack: unknown data type"); \
    }

extern "C" {

int fimexcu_unpack_dev_f(int dtype, const void* d_in, size_t n, double badValue, float* d_out, void* stream)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    if (n == 0)
        return FIMEXCU_OK;
    if (!d_in || !d_out)
        FXCU_FAIL("fimexcu_unpack: null argument");
#define CALL(T) unpack_t<T>(d_in, n, badValue, d_out, as_stream(stream))
    FXCU_DTYPE_SWITCH(dtype, CALL)
#undef CALL
}

int fimexcu_unpack_scaled_dev_f(int dtype, const void* d_in, size_t n, double fill, double scale, double offset, float* d_out, void* stream)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    if (n == 0)
        return FIMEXCU_OK;
    if (!d_in || !d_out)
        FXCU_FAIL("fimexcu_unpack_scaled: null argument");
#define CALL(T) unpack_scaled_t<T>(d_in, n, fill, scale, offset, d_out, as_stream(stream))
    FXCU_DTYPE_SWITCH(dtype, CALL)
#undef CALL
}

int fimexcu_pack_dev_f(int dtype, const float* d_in, size_t n, double badValue, void* d_out, void* stream)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    if (n == 0)
        return FIMEXCU_OK;
    if (!d_in || !d_out)
        FXCU_FAIL("fimexcu_pack: null argument");
#define CALL(T) pack_t<T>(d_in, n, badValue, d_out, as_stream(stream))
    FXCU_DTYPE_SWITCH(dtype, CALL)
#undef CALL
}

} // extern "C"
