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

// element conversions (the arithmetic of the three passes)
template <class T>
struct UnpackRaw { // Data::asFloat + mifi_bad2nanf
    float bad;
    __device__ __forceinline__ float operator()(T e) const
    {
        const float v = static_cast<float>(e);
        return v == bad ? nan_f() : v;
    }
};
template <class T>
struct UnpackScaled { // ScaleValue<T,double>(fill, scale, offset, NaN, 1, 0) narrowed to f32
    T fill;
    double scale, offset;
    __device__ __forceinline__ float operator()(T e) const
    {
        return (e == fill) ? nan_f() : __double2float_rn(__dadd_rn(__dmul_rn(scale, static_cast<double>(e)), offset));
    }
};
template <class T>
struct PackRound { // ScaleValue<float,T>(NaN, 1, 0, fill, 1, 0)
    T fill;
    __device__ __forceinline__ T operator()(float v) const
    {
        T r = fill;
        if (!isnan(v)) {
            float t = truncf(v);
            if (fabsf(__fsub_rn(v, t)) >= 0.5f) // lround: half away from zero; v - trunc(v) is exact
                t = __fadd_rn(t, copysignf(1.f, v));
            r = static_cast<T>(static_cast<int>(__float2ll_rz(t))); // int round(double) { return ::lround(num); } then static_cast<T>
        }
        return r;
    }
};

// four elements as one word of 4 * sizeof(T) bytes
template <int BYTES> struct Word;
template <> struct Word<4> { typedef unsigned type; };
template <> struct Word<8> { typedef uint2 type; };
template <> struct Word<16> { typedef uint4 type; };

// integer slab -> f32 slab.  VEC (both bases on 16 bytes): a thread takes FOUR elements -- one 4-, 8- or 16-byte load and one
// 16-byte store, so that a warp's store is 512 contiguous bytes; the elements beyond the last whole group, or everything when a
// base is not aligned, go one by one.  (16-byte accesses stream at ~1.5x the rate of 4-byte ones on this GPU, 2-byte ones are worse.)
template <class T, bool VEC, class F>
__global__ void __launch_bounds__(256) unpack_kernel(const T* __restrict__ in, size_t n, F f, float* __restrict__ out)
{
    typedef typename Word<4 * sizeof(T)>::type W;
    const size_t nv = VEC ? n / 4 : 0;
    if constexpr (VEC)
        for (size_t k = blockIdx.x * 256ull + threadIdx.x; k < nv; k += 256ull * gridDim.x) {
            union {
                W q;
                T e[4];
            } a;
            a.q = __ldcs(reinterpret_cast<const W*>(in) + k);
            reinterpret_cast<float4*>(out)[k] = make_float4(f(a.e[0]), f(a.e[1]), f(a.e[2]), f(a.e[3]));
        }
    for (size_t i = nv * 4 + blockIdx.x * 256ull + threadIdx.x; i < n; i += 256ull * gridDim.x)
        out[i] = f(in[i]);
}

// f32 slab -> integer slab, the same way round
template <class T, bool VEC, class F>
__global__ void __launch_bounds__(256) pack_kernel(const float* __restrict__ in, size_t n, F f, T* __restrict__ out)
{
    typedef typename Word<4 * sizeof(T)>::type W;
    const size_t nv = VEC ? n / 4 : 0;
    if constexpr (VEC)
        for (size_t k = blockIdx.x * 256ull + threadIdx.x; k < nv; k += 256ull * gridDim.x) {
            union {
                W q;
                T e[4];
            } a;
            const float4 v = __ldcs(reinterpret_cast<const float4*>(in) + k);
            a.e[0] = f(v.x);
            a.e[1] = f(v.y);
            a.e[2] = f(v.z);
            a.e[3] = f(v.w);
            reinterpret_cast<W*>(out)[k] = a.q;
        }
    for (size_t i = nv * 4 + blockIdx.x * 256ull + threadIdx.x; i < n; i += 256ull * gridDim.x)
        out[i] = f(in[i]);
}

static bool on16(const void* a, const void* b) { return ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15u) == 0; }

static unsigned stream_blocks(size_t n, size_t per_thread = 1)
{
    return static_cast<unsigned>(std::max<size_t>(1, std::min<size_t>((n / per_thread + 255) / 256, static_cast<size_t>(sm_count()) * 16)));
}

template <class T>
static int unpack_t(const void* d_in, size_t n, double bad, float* d_out, cudaStream_t s)
{
    const UnpackRaw<T> f{static_cast<float>(bad)};
    if (on16(d_in, d_out))
        unpack_kernel<T, true><<<stream_blocks(n, 4), 256, 0, s>>>(static_cast<const T*>(d_in), n, f, d_out);
    else
        unpack_kernel<T, false><<<stream_blocks(n), 256, 0, s>>>(static_cast<const T*>(d_in), n, f, d_out);
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}
template <class T>
static int unpack_scaled_t(const void* d_in, size_t n, double fill, double scale, double offset, float* d_out, cudaStream_t s)
{
    // ScaleValue's constructor: oldFill_(static_cast<IN>(oldFill)), oldScale/newScale with newScale = 1, (oldOffset - 0)/1
    const UnpackScaled<T> f{static_cast<T>(fill), scale / 1.0, (offset - 0.0) / 1.0};
    if (on16(d_in, d_out))
        unpack_kernel<T, true><<<stream_blocks(n, 4), 256, 0, s>>>(static_cast<const T*>(d_in), n, f, d_out);
    else
        unpack_kernel<T, false><<<stream_blocks(n), 256, 0, s>>>(static_cast<const T*>(d_in), n, f, d_out);
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}
template <class T>
static int pack_t(const float* d_in, size_t n, double bad, void* d_out, cudaStream_t s)
{
    const PackRound<T> f{static_cast<T>(bad)};
    if (on16(d_in, d_out))
        pack_kernel<T, true><<<stream_blocks(n, 4), 256, 0, s>>>(d_in, n, f, static_cast<T*>(d_out));
    else
        pack_kernel<T, false><<<stream_blocks(n), 256, 0, s>>>(d_in, n, f, static_cast<T*>(d_out));
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
