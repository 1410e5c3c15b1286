// fxcu_axes.cu -- coordinate -> fractional axis index (K2).
//
// Reference behaviour restated here (fimex 2.4.1): mifi_points2position, src/interpolation.c:154-227, with its
// helper bsearchDoubleIndex (:129-152): longitude normalisation to the axis convention, circular-axis detection,
// exact hits, linear spline on the bracketing interval with end-interval extrapolation, non-finite -> -999.
// One thread per point; the axis (a few thousand f64) stays in L1/L2.
#include "fxcu_common.h"

#include <mutex>
#include <vector>

namespace fxcu {

__device__ __forceinline__ int axis_search(double key, const double* __restrict__ base, int num, bool ascending)
{
    int lo = 0, hi = num - 1, mid = 0, cmp = 0;
    while (lo <= hi) {
        mid = (lo + hi) / 2;
        const double b = __ldg(base + mid);
        cmp = (key == b) ? 0 : ((key > b) ? 1 : -1);
        if (!ascending)
            cmp = -cmp;
        if (cmp == 0)
            return mid;
        if (cmp > 0)
            lo = mid + 1;
        else
            hi = mid - 1;
    }
    return (cmp > 0) ? -(mid + 1) - 1 : -mid - 1;
}

// lon_mode: 0 none, 1 axis is -180..180 (points > 180 lose 360), 2 axis is 0..360 (negative points gain 360)
__device__ __forceinline__ double point_to_position(double p, const double* __restrict__ axis, int num, bool ascending, int lon_mode, bool circular)
{
    if (lon_mode == 1) {
        if (p > 180.)
            p -= 360.;
    } else if (lon_mode == 2) {
        if (p < 0)
            p += 360.;
    }
    if (!isfinite(p))
        return -999.;
    const int pos = axis_search(p, axis, num, ascending);
    if (pos >= 0)
        return static_cast<double>(pos);
    int k = -(pos + 1);
    if (k == num)
        k--;
    else if (k == 0)
        k++;
    const double ak = __ldg(axis + k), akm = __ldg(axis + k - 1);
    const double slope = __dsub_rn(ak, akm);
    const double offset = __dsub_rn(ak, __dmul_rn(slope, static_cast<double>(k)));
    double r = __ddiv_rn(__dsub_rn(p, offset), slope);
    if (circular && r <= -0.5)
        r = __dadd_rn(r, static_cast<double>(num));
    if (circular && r > (num - 0.5))
        r = __dsub_rn(r, static_cast<double>(num));
    return r;
}

__global__ void __launch_bounds__(256) points2position_kernel(double* __restrict__ pts, size_t n, const double* __restrict__ axis, int num, bool ascending,
                                                              int lon_mode, bool circular)
{
    const size_t i = blockIdx.x * 256ull + threadIdx.x;
    if (i >= n)
        return;
    pts[i] = point_to_position(pts[i], axis, num, ascending, lon_mode, circular);
}

// host part of interpolation.c:156-190: comparator choice, longitude convention, circularity
void axis_traits(const double* axis, int num, int axis_type, bool& ascending, int& lon_mode, bool& circular)
{
    ascending = axis[0] < axis[num - 1];
    lon_mode = 0;
    circular = false;
    if (axis_type == FIMEXCU_LONGITUDE) {
        lon_mode = (axis[0] < 0 || axis[num - 1] < 0) ? 1 : 2;
        double next = axis[num - 1] + (axis[1] - axis[0]) * 1.01;
        if (ascending) {
            next -= 360.;
            circular = next >= axis[0];
        } else {
            next += 360.;
            circular = next <= axis[0];
        }
    }
}

int points2position_dev(double* d_points, size_t n, const double* d_axis, const double* axis_host, int num, int axis_type, cudaStream_t s)
{
    if (num < 2)
        FXCU_FAIL("fimexcu_points2position: axis needs at least 2 values");
    bool ascending, circular;
    int lon_mode;
    axis_traits(axis_host, num, axis_type, ascending, lon_mode, circular);
    if (n == 0)
        return FIMEXCU_OK;
    points2position_kernel<<<blocks_for(n, 256), 256, 0, s>>>(d_points, n, d_axis, num, ascending, lon_mode, circular);
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}

// ---- scratch for fimexcu_points2position_dev_noalloc: the device copy of the axis lives in a fixed array that is
// part of the loaded module (one instance per GPU that loads the library, held until the process exits) ----
constexpr int NOALLOC_MAX_AXIS = FIMEXCU_NOALLOC_MAX_AXIS;
static __device__ __align__(256) double g_noalloc_axis[NOALLOC_MAX_AXIS];

// In host code `g_noalloc_axis` is only a stand-in for the symbol, not a device address: the address the kernel
// needs is looked up once per device, with that device current.
static double* noalloc_axis_address()
{
    static std::mutex lock;
    static double* address[64] = {nullptr};
    int dev = -1;
    if (!cuda_ok(cudaGetDevice(&dev), "cudaGetDevice", __FILE__, __LINE__))
        return nullptr;
    if (dev < 0 || dev >= 64) {
        set_error("fimexcu_points2position_dev_noalloc: device ordinal out of range");
        return nullptr;
    }
    std::lock_guard<std::mutex> guard(lock);
    if (!address[dev]) {
        double* p = nullptr;
        if (!cuda_ok(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_axis), "cudaGetSymbolAddress(g_noalloc_axis)", __FILE__, __LINE__))
            return nullptr;
        address[dev] = p;
    }
    return address[dev];
}

} // namespace fxcu

using namespace fxcu;

extern "C" {

int fimexcu_points2position_dev(double* d_points, size_t n, const double* axis_host, int num, int axis_type, void* stream)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    cudaStream_t s = as_stream(stream);
    DevBuf ax;
    if (!ax.alloc(static_cast<size_t>(num > 0 ? num : 1) * sizeof(double), s))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(ax.p, axis_host, static_cast<size_t>(num) * sizeof(double), cudaMemcpyHostToDevice, s));
    const int rc = points2position_dev(d_points, n, ax.as<double>(), axis_host, num, axis_type, s);
    FXCU_CUDA(cudaStreamSynchronize(s)); // axis_host may be a temporary of the caller
    return rc;
}

int fimexcu_points2position_dev_noalloc(double* d_points, size_t n, const double* axis_host, int num, int axis_type, void* stream)
{
    clear_error();
    if (num > NOALLOC_MAX_AXIS) {
        char msg[160];
        std::snprintf(msg, sizeof(msg), "fimexcu_points2position_dev_noalloc: axis has %d values, the fixed scratch holds at most %d "
                                        "(use fimexcu_points2position_dev, which allocates)", num, NOALLOC_MAX_AXIS);
        FXCU_FAIL(msg);
    }
    if (num < 2)
        FXCU_FAIL("fimexcu_points2position: axis needs at least 2 values");
    int dev = 0, major = 0;
    FXCU_CUDA(cudaGetDevice(&dev));
    FXCU_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    if (major != 10)
        FXCU_FAIL("libfimexcu is built for sm_100a only; the current device is not a compute-capability 10.x GPU");
    double* d_axis = noalloc_axis_address(); // first call on a device: symbol look-up
    if (!d_axis)
        return FIMEXCU_ERROR;
    cudaStream_t s = as_stream(stream);
    // every call overwrites the first `num` entries it reads; nothing in the scratch has to start at zero
    FXCU_CUDA(cudaMemcpyAsync(d_axis, axis_host, static_cast<size_t>(num) * sizeof(double), cudaMemcpyHostToDevice, s));
    const int rc = points2position_dev(d_points, n, d_axis, axis_host, num, axis_type, s);
    FXCU_CUDA(cudaStreamSynchronize(s)); // axis_host may be a temporary of the caller; also ends this call's use of the scratch
    return rc;
}

int fimexcu_points2position(double* points, size_t n, const double* axis, int num, int axis_type)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    DevBuf d;
    if (!d.alloc(n * sizeof(double), nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(d.p, points, n * sizeof(double), cudaMemcpyHostToDevice, nullptr));
    if (fimexcu_points2position_dev(d.as<double>(), n, axis, num, axis_type, nullptr) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(points, d.p, n * sizeof(double), cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

} // extern "C"
