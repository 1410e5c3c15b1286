// fxcu_proj.cu -- cache build on the device: lattice reprojection (K1), rotation-matrix build (K8) and the one-call
// "projection parameters -> cache" path.
//
// Reference behaviour restated here (fimex 2.4.1):
//   reproject::reproject_axes / reproject_values                  src/reproject.cc:567-587
//   reproject::get_vector_reproject_matrix (+ _proj, _points_proj_delta, _points)   src/reproject.cc:310-397, 432-486
//   mifi_bearing, store_rotation                                   src/reproject.cc:241-264, 297-302
//   CDMInterpolator::changeProjectionByProjectionParameters        src/CDMInterpolator.cc:1389-1416
// The projection arithmetic itself (PROJ in the reference) is fxcu_proj.h.  One thread per lattice point, f64.
#include "fxcu_handles.h"
#include "fxcu_proj.h"

#include <algorithm>
#include <cstring>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "fxcu_proj_setup.cpp.inc"

namespace fxcu {

using fxproj::Crs;

__global__ void __launch_bounds__(256) transform_kernel(const Crs src, const Crs dst, size_t n, double* __restrict__ x, double* __restrict__ y)
{
    const size_t i = blockIdx.x * 256ull + threadIdx.x;
    if (i >= n)
        return;
    double px = x[i], py = y[i];
    fxproj::crs_transform(src, dst, px, py);
    x[i] = px;
    y[i] = py;
}

// reproject_axes: broadcast the 1-D axes to the lattice and transform (reproject.cc:573-587)
__global__ void __launch_bounds__(256) lattice_kernel(const Crs src, const Crs dst, const double* __restrict__ ax, const double* __restrict__ ay, int ix,
                                                      int iy, double* __restrict__ ox, double* __restrict__ oy)
{
    const size_t i = blockIdx.x * 256ull + threadIdx.x;
    if (i >= static_cast<size_t>(ix) * iy)
        return;
    const int yy = static_cast<int>(i / ix), xx = static_cast<int>(i - static_cast<size_t>(yy) * ix);
    double px = ax[xx], py = ay[yy];
    fxproj::crs_transform(src, dst, px, py);
    ox[i] = px;
    oy[i] = py;
}

__device__ __forceinline__ double bearing_deg_args(double lat0, double lon0, double lat1, double lon1)
{
    // mifi_bearing (reproject.cc:241-264): DEG_TO_RAD there is (M_PI / 180)
    const double d2r = fxproj::PI / 180;
    lat0 *= d2r;
    lon0 *= d2r;
    lat1 *= d2r;
    lon1 *= d2r;
    const double dlon = lon0 - lon1;
    return atan2(sin(dlon) * cos(lat1), cos(lat0) * sin(lat1) - sin(lat0) * cos(lat1) * cos(dlon));
}

// get_vector_reproject_matrix_points_proj_delta (reproject.cc:310-357): src -> dst is c2c.fwd
__global__ void __launch_bounds__(256) matrix_kernel(const Crs src, const Crs dst, const double* __restrict__ in_x, const double* __restrict__ in_y,
                                                     const double* __restrict__ out_x, const double* __restrict__ out_y, double dX, double dY, size_t n,
                                                     double* __restrict__ mtx)
{
    const size_t i = blockIdx.x * 256ull + threadIdx.x;
    if (i >= n)
        return;
    const double ix = in_x[i], iy = in_y[i], oxv = out_x[i], oyv = out_y[i];
    double ty_x = ix, ty_y = iy + dY; // step along y
    fxproj::crs_transform(src, dst, ty_x, ty_y);
    const double add_y = dY > 0 ? 0 : fxproj::PI;
    double phi0;
    if (dst.geographic) {
        phi0 = bearing_deg_args(oyv, oxv, ty_y, ty_x) + add_y; // true-north bearing only
    } else {
        double tx_x = ix + dX, tx_y = iy; // step along x
        fxproj::crs_transform(src, dst, tx_x, tx_y);
        const double add_x = dX > 0 ? 0 : fxproj::PI;
        const double phi_x = atan2(tx_y - oyv, tx_x - oxv) + add_x;
        const double phi_y = -atan2(ty_x - oxv, ty_y - oyv) + add_y;
        phi0 = (phi_y + phi_x) / 2; // average, for non-conformal cases
    }
    mtx[3 * i + 0] = cos(phi0);
    mtx[3 * i + 1] = sin(phi0);
    mtx[3 * i + 2] = (180 / fxproj::PI) * phi0;
}

static bool parse_pair(const char* a, const char* b, Crs& ca, Crs& cb)
{
    std::string err;
    if (!parse_crs(a, ca, err) || !parse_crs(b, cb, err)) {
        set_error(err);
        return false;
    }
    return true;
}

// lattice of the 1-D axes transformed src -> dst into device arrays
static int lattice_dev(const Crs& src, const Crs& dst, const double* ax_host, const double* ay_host, int ix, int iy, double* d_ox, double* d_oy,
                       cudaStream_t s)
{
    if (ix <= 0 || iy <= 0)
        return FIMEXCU_OK;
    DevBuf axes;
    if (!axes.alloc(static_cast<size_t>(ix + iy) * sizeof(double), s))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(axes.p, ax_host, ix * sizeof(double), cudaMemcpyHostToDevice, s));
    FXCU_CUDA(cudaMemcpyAsync(axes.as<double>() + ix, ay_host, iy * sizeof(double), cudaMemcpyHostToDevice, s));
    const size_t n = static_cast<size_t>(ix) * iy;
    lattice_kernel<<<blocks_for(n, 256), 256, 0, s>>>(src, dst, axes.as<double>(), axes.as<double>() + ix, ix, iy, d_ox, d_oy);
    FXCU_CUDA(cudaGetLastError());
    FXCU_CUDA(cudaStreamSynchronize(s)); // the axes are temporaries of the caller
    return FIMEXCU_OK;
}

static int matrix_from_dev_points(const Crs& src, const Crs& dst, const double* d_in_x, const double* d_in_y, const double* d_out_x,
                                  const double* d_out_y, double dX, double dY, size_t n, double* matrix3_host)
{
    DevBuf m;
    if (!m.alloc(std::max<size_t>(n, 1) * 3 * sizeof(double), nullptr))
        return FIMEXCU_ERROR;
    if (n) {
        matrix_kernel<<<blocks_for(n, 256), 256>>>(src, dst, d_in_x, d_in_y, d_out_x, d_out_y, dX, dY, n, m.as<double>());
        FXCU_CUDA(cudaGetLastError());
        FXCU_CUDA(cudaMemcpyAsync(matrix3_host, m.p, n * 3 * sizeof(double), cudaMemcpyDeviceToHost, nullptr));
    }
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

// get_vector_reproject_matrix_proj (reproject.cc:363-397): delta = 0.1 % of the source spacing, sampled at two places, then the
// per-point angles (:310-357).  All four arrays on the device, ox*oy entries, row-major.
static int matrix_proj(const Crs& src, const Crs& dst, const double* d_in_x, const double* d_in_y, const double* d_out_x, const double* d_out_y, int ox,
                       int oy, double* matrix3_host)
{
    const size_t n = static_cast<size_t>(ox) * oy;
    auto at = [&](size_t idx) {
        double v = 0;
        cudaMemcpy(&v, d_in_x + idx, sizeof(double), cudaMemcpyDeviceToHost);
        return v;
    };
    const double dflt = 1e-3;
    double delta;
    if (ox > 1) {
        if (oy > 1) {
            delta = dflt * (at(static_cast<size_t>(ox) + 1) - at(0));
            const size_t ox2 = ox / 2, oy2 = oy / 2;
            const double d2 = dflt * (at((oy2 + 1) * ox + (ox2 + 1)) - at(oy2 * ox + ox2));
            delta += d2;
            delta /= 2;
        } else {
            delta = dflt * (at(1) - at(0));
        }
    } else {
        if (oy > 1) {
            delta = dflt * (at(static_cast<size_t>(ox)) - at(0));
        } else {
            const double v0 = at(0);
            delta = (v0 > 1) ? (v0 * dflt) : dflt;
        }
    }
    if (std::fabs(delta) < 1e-9)
        delta = dflt;
    return matrix_from_dev_points(src, dst, d_in_x, d_in_y, d_out_x, d_out_y, delta, delta, n, matrix3_host);
}

} // namespace fxcu

using namespace fxcu;

extern "C" {

int fimexcu_reproject_values_dev(const char* proj_input, const char* proj_output, double* d_x, double* d_y, size_t num, void* stream)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    Crs src, dst;
    if (!parse_pair(proj_input, proj_output, src, dst))
        return FIMEXCU_ERROR;
    if (num) {
        transform_kernel<<<blocks_for(num, 256), 256, 0, as_stream(stream)>>>(src, dst, num, d_x, d_y);
        FXCU_CUDA(cudaGetLastError());
    }
    return FIMEXCU_OK;
}

int fimexcu_reproject_values(const char* proj_input, const char* proj_output, double* in_out_x_vals, double* in_out_y_vals, int num)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    if (num <= 0)
        return FIMEXCU_OK;
    const size_t bytes = static_cast<size_t>(num) * sizeof(double);
    DevBuf dx, dy;
    if (!dx.alloc(bytes, nullptr) || !dy.alloc(bytes, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(dx.p, in_out_x_vals, bytes, cudaMemcpyHostToDevice, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(dy.p, in_out_y_vals, bytes, cudaMemcpyHostToDevice, nullptr));
    if (fimexcu_reproject_values_dev(proj_input, proj_output, dx.as<double>(), dy.as<double>(), static_cast<size_t>(num), nullptr) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(in_out_x_vals, dx.p, bytes, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(in_out_y_vals, dy.p, bytes, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

int fimexcu_reproject_axes_dev(const char* proj_input, const char* proj_output, const double* in_x_axis_host, const double* in_y_axis_host, int ix,
                               int iy, double* d_out_xproj_axis, double* d_out_yproj_axis, void* stream)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    Crs src, dst;
    if (!parse_pair(proj_input, proj_output, src, dst))
        return FIMEXCU_ERROR;
    return lattice_dev(src, dst, in_x_axis_host, in_y_axis_host, ix, iy, d_out_xproj_axis, d_out_yproj_axis, as_stream(stream));
}

int fimexcu_reproject_axes(const char* proj_input, const char* proj_output, const double* in_x_axis, const double* in_y_axis, int ix, int iy,
                           double* out_xproj_axis, double* out_yproj_axis)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    if (ix <= 0 || iy <= 0)
        return FIMEXCU_OK;
    const size_t bytes = static_cast<size_t>(ix) * iy * sizeof(double);
    DevBuf dx, dy;
    if (!dx.alloc(bytes, nullptr) || !dy.alloc(bytes, nullptr))
        return FIMEXCU_ERROR;
    if (fimexcu_reproject_axes_dev(proj_input, proj_output, in_x_axis, in_y_axis, ix, iy, dx.as<double>(), dy.as<double>(), nullptr) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(out_xproj_axis, dx.p, bytes, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(out_yproj_axis, dy.p, bytes, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

int fimexcu_get_vector_reproject_matrix(const char* proj_input, const char* proj_output, const double* out_x_axis, const double* out_y_axis, int ox,
                                        int oy, double* matrix3)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    Crs src, dst;
    if (!parse_pair(proj_input, proj_output, src, dst))
        return FIMEXCU_ERROR;
    if (ox <= 0 || oy <= 0)
        return FIMEXCU_OK;
    const size_t n = static_cast<size_t>(ox) * oy;
    DevBuf inx, iny, outx, outy;
    if (!inx.alloc(n * 8, nullptr) || !iny.alloc(n * 8, nullptr) || !outx.alloc(n * 8, nullptr) || !outy.alloc(n * 8, nullptr))
        return FIMEXCU_ERROR;
    // positions of the target lattice in the target CRS (identity transform) and in the source CRS (c2c.inv)
    if (lattice_dev(dst, src, out_x_axis, out_y_axis, ox, oy, inx.as<double>(), iny.as<double>(), nullptr) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    {
        std::vector<double> hx(n), hy(n);
        for (int y = 0; y < oy; ++y)
            for (int x = 0; x < ox; ++x) {
                hx[static_cast<size_t>(y) * ox + x] = out_x_axis[x];
                hy[static_cast<size_t>(y) * ox + x] = out_y_axis[y];
            }
        FXCU_CUDA(cudaMemcpy(outx.p, hx.data(), n * 8, cudaMemcpyHostToDevice));
        FXCU_CUDA(cudaMemcpy(outy.p, hy.data(), n * 8, cudaMemcpyHostToDevice));
    }
    return matrix_proj(src, dst, inx.as<double>(), iny.as<double>(), outx.as<double>(), outy.as<double>(), ox, oy, matrix3);
}

int fimexcu_get_vector_reproject_matrix_field(const char* proj_input, const char* proj_output, const double* in_x_field, const double* in_y_field, int ox,
                                              int oy, double* matrix3)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    Crs src, dst;
    if (!parse_pair(proj_input, proj_output, src, dst))
        return FIMEXCU_ERROR;
    if (ox <= 0 || oy <= 0)
        return FIMEXCU_OK;
    const size_t n = static_cast<size_t>(ox) * oy;
    DevBuf inx, iny, outx, outy;
    if (!inx.alloc(n * 8, nullptr) || !iny.alloc(n * 8, nullptr) || !outx.alloc(n * 8, nullptr) || !outy.alloc(n * 8, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpy(inx.p, in_x_field, n * 8, cudaMemcpyHostToDevice));
    FXCU_CUDA(cudaMemcpy(iny.p, in_y_field, n * 8, cudaMemcpyHostToDevice));
    FXCU_CUDA(cudaMemcpy(outx.p, inx.p, n * 8, cudaMemcpyDeviceToDevice));
    FXCU_CUDA(cudaMemcpy(outy.p, iny.p, n * 8, cudaMemcpyDeviceToDevice));
    transform_kernel<<<blocks_for(n, 256), 256>>>(src, dst, n, outx.as<double>(), outy.as<double>()); // c2c.fwd (reproject.cc:459)
    FXCU_CUDA(cudaGetLastError());
    return matrix_proj(src, dst, inx.as<double>(), iny.as<double>(), outx.as<double>(), outy.as<double>(), ox, oy, matrix3);
}

int fimexcu_get_vector_reproject_matrix_points(const char* proj_input, const char* proj_output, int inputIsMetric, const double* out_x_points,
                                               const double* out_y_points, int ox, int oy, double* matrix3)
{
    clear_error();
    if (!require_device())
        return FIMEXCU_ERROR;
    Crs src, dst;
    if (!parse_pair(proj_input, proj_output, src, dst))
        return FIMEXCU_ERROR;
    if (ox <= 0 || oy <= 0)
        return FIMEXCU_OK;
    const size_t n = static_cast<size_t>(ox) * oy;
    DevBuf inx, iny, outx, outy;
    if (!inx.alloc(n * 8, nullptr) || !iny.alloc(n * 8, nullptr) || !outx.alloc(n * 8, nullptr) || !outy.alloc(n * 8, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpy(outx.p, out_x_points, n * 8, cudaMemcpyHostToDevice));
    FXCU_CUDA(cudaMemcpy(outy.p, out_y_points, n * 8, cudaMemcpyHostToDevice));
    FXCU_CUDA(cudaMemcpy(inx.p, outx.p, n * 8, cudaMemcpyDeviceToDevice));
    FXCU_CUDA(cudaMemcpy(iny.p, outy.p, n * 8, cudaMemcpyDeviceToDevice));
    transform_kernel<<<blocks_for(n, 256), 256>>>(dst, src, n, inx.as<double>(), iny.as<double>()); // c2c.inv
    FXCU_CUDA(cudaGetLastError());
    const double delta = inputIsMetric ? 100 : 0.00001; // reproject.cc:445
    return matrix_from_dev_points(src, dst, inx.as<double>(), iny.as<double>(), outx.as<double>(), outy.as<double>(), delta, delta, n, matrix3);
}

fimexcu_cache* fimexcu_cache_create_from_projection(int method, const char* proj_input, const char* proj_org, const double* out_x_axis,
                                                    const double* out_y_axis, size_t outX, size_t outY, const double* org_x_axis,
                                                    const double* org_y_axis, size_t inX, size_t inY, int org_x_axis_type, int org_y_axis_type)
{
    clear_error();
    if (!require_device())
        return nullptr;
    Crs target, source;
    if (!parse_pair(proj_input, proj_org, target, source))
        return nullptr;
    if (outX == 0 || outY == 0 || inX < 2 || inY < 2 || outX > 0x3fffffff || outY > 0x3fffffff || inX > 0x3fffffff || inY > 0x3fffffff) {
        set_error("fimexcu_cache_create_from_projection: axis sizes out of range");
        return nullptr;
    }
    const size_t n = outX * outY;
    DevBuf px, py, axes;
    if (!px.alloc(n * 8, nullptr) || !py.alloc(n * 8, nullptr) || !axes.alloc((inX + inY) * 8, nullptr))
        return nullptr;
    // CDMInterpolator.cc:1389: target lattice expressed in the source CRS
    if (lattice_dev(target, source, out_x_axis, out_y_axis, static_cast<int>(outX), static_cast<int>(outY), px.as<double>(), py.as<double>(), nullptr) !=
        FIMEXCU_OK)
        return nullptr;
    // :1398-1399: coordinates -> fractional indices on the source axes
    FXCU_CUDA_NULL(cudaMemcpy(axes.p, org_x_axis, inX * 8, cudaMemcpyHostToDevice));
    FXCU_CUDA_NULL(cudaMemcpy(axes.as<double>() + inX, org_y_axis, inY * 8, cudaMemcpyHostToDevice));
    if (points2position_dev(px.as<double>(), n, axes.as<double>(), org_x_axis, static_cast<int>(inX), org_x_axis_type, nullptr) != FIMEXCU_OK ||
        points2position_dev(py.as<double>(), n, axes.as<double>() + inX, org_y_axis, static_cast<int>(inY), org_y_axis_type, nullptr) != FIMEXCU_OK)
        return nullptr;
    // :1404: createCachedInterpolation
    return cache_build(method, px.as<double>(), py.as<double>(), inX, inY, outX, outY, nullptr);
}

fimexcu_cache* fimexcu_cache_create_from_points(int method, const char* proj_points, const char* proj_org, const double* x_points,
                                                const double* y_points, size_t outX, size_t outY, const double* org_x_axis,
                                                const double* org_y_axis, size_t inX, size_t inY, int org_x_axis_type, int org_y_axis_type)
{
    clear_error();
    if (!require_device())
        return nullptr;
    Crs points, source;
    if (!parse_pair(proj_points, proj_org, points, source))
        return nullptr;
    if (outX == 0 || outY == 0 || inX < 2 || inY < 2 || outX > 0x3fffffff || outY > 0x3fffffff || inX > 0x3fffffff || inY > 0x3fffffff) {
        set_error("fimexcu_cache_create_from_points: sizes out of range");
        return nullptr;
    }
    if (!x_points || !y_points || !org_x_axis || !org_y_axis) {
        set_error("fimexcu_cache_create_from_points: null argument");
        return nullptr;
    }
    const size_t n = outX * outY;
    DevBuf px, py, axes;
    if (!px.alloc(n * 8, nullptr) || !py.alloc(n * 8, nullptr) || !axes.alloc((inX + inY) * 8, nullptr))
        return nullptr;
    FXCU_CUDA_NULL(cudaMemcpyAsync(px.p, x_points, n * 8, cudaMemcpyHostToDevice, nullptr));
    FXCU_CUDA_NULL(cudaMemcpyAsync(py.p, y_points, n * 8, cudaMemcpyHostToDevice, nullptr));
    // CDMInterpolator.cc:1647: the template's / list's points expressed in the source CRS
    transform_kernel<<<blocks_for(n, 256), 256>>>(points, source, n, px.as<double>(), py.as<double>());
    FXCU_CUDA_NULL(cudaGetLastError());
    // :1659-1660: coordinates -> fractional indices on the source axes
    FXCU_CUDA_NULL(cudaMemcpy(axes.p, org_x_axis, inX * 8, cudaMemcpyHostToDevice));
    FXCU_CUDA_NULL(cudaMemcpy(axes.as<double>() + inX, org_y_axis, inY * 8, cudaMemcpyHostToDevice));
    if (points2position_dev(px.as<double>(), n, axes.as<double>(), org_x_axis, static_cast<int>(inX), org_x_axis_type, nullptr) != FIMEXCU_OK ||
        points2position_dev(py.as<double>(), n, axes.as<double>() + inX, org_y_axis, static_cast<int>(inY), org_y_axis_type, nullptr) != FIMEXCU_OK)
        return nullptr;
    // :1665: createCachedInterpolation
    return cache_build(method, px.as<double>(), py.as<double>(), inX, inY, outX, outY, nullptr);
}

} // extern "C"
