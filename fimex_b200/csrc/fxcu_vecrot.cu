// fxcu_vecrot.cu -- wind-vector rotation by a cached per-cell matrix (K7) and the fused
// "interpolate u and v, then rotate" pass that replaces the vector branch of CDMInterpolator::getDataSlice.
//
// Reference behaviour restated here (fimex 2.4.1):
//   CachedVectorReprojection                 src/CachedVectorReprojection.cc:37-54
//   vector_reproject_values_by_matrix_f      src/reproject.cc:488-507   (f64 products and sums, rounded one by one)
//   vector_reproject_direction_by_matrix_f   src/reproject.cc:509-530
//   reproject::Matrix (stride 3: cos, sin, angle in degree)   include/fimex/reproject.h:34-46
//
// HBM layout: the matrix is split into three f64 planes (cos, sin, degree) of ox*oy entries so that a warp reads
// 256 contiguous bytes per plane; u, v are [nz][oy][ox] f32 rotated in place.
#include "fxcu_handles.h"

#include <memory>

namespace fxcu {

__global__ void __launch_bounds__(256) split_matrix_kernel(const double* __restrict__ m3, size_t n, double* __restrict__ c, double* __restrict__ s,
                                                           double* __restrict__ deg)
{
    const size_t i = blockIdx.x * 256ull + threadIdx.x;
    if (i >= n)
        return;
    c[i] = m3[3 * i];
    s[i] = m3[3 * i + 1];
    deg[i] = m3[3 * i + 2];
}

__device__ __forceinline__ void rotate_uv(float u, float v, double c, double s, float& ur, float& vr)
{
    const double du = static_cast<double>(u), dv = static_cast<double>(v);
    ur = __double2float_rn(__dsub_rn(__dmul_rn(du, c), __dmul_rn(dv, s)));
    vr = __double2float_rn(__dadd_rn(__dmul_rn(du, s), __dmul_rn(dv, c)));
}

__global__ void __launch_bounds__(256) rotate_kernel(const double* __restrict__ cs, const double* __restrict__ sn, float* __restrict__ u,
                                                     float* __restrict__ v, size_t layer, int nz, int zchunk)
{
    const size_t i = blockIdx.x * 256ull + threadIdx.x;
    if (i >= layer)
        return;
    const int z0 = blockIdx.y * zchunk, z1 = min(nz, z0 + zchunk);
    const double c = cs[i], s = sn[i];
    float* pu = u + static_cast<size_t>(z0) * layer + i;
    float* pv = v + static_cast<size_t>(z0) * layer + i;
#pragma unroll 4
    for (int z = z0; z < z1; ++z, pu += layer, pv += layer) {
        float ur, vr;
        rotate_uv(*pu, *pv, c, s, ur, vr);
        __stcs(pu, ur);
        __stcs(pv, vr);
    }
}

// The same rotation on slabs with a row pitch (the padded rows of a device-resident chain): a thread takes four neighbouring cells
// of a row -- one 16-byte load and store per component and level where the layout allows (VEC: pitch and level stride multiples
// of four, bases on 16 bytes), which is what the memory system wants (16-byte accesses stream at 6.7 TB/s on this GPU, 4-byte
// ones at 4.4: profiles/exp/inplace_probe.py).  The pad elements of a row are not touched.
template <bool VEC>
__global__ void __launch_bounds__(256) rotate_rows_kernel(const double* __restrict__ cs, const double* __restrict__ sn, float* __restrict__ u,
                                                          float* __restrict__ v, int ox, int oy, size_t pitch, size_t level, int nz, int zchunk)
{
    const int x = (blockIdx.x * 64 + (threadIdx.x & 63)) * 4, y = blockIdx.y * 4 + (threadIdx.x >> 6);
    if (x >= ox || y >= oy)
        return;
    const int z0 = blockIdx.z * zchunk, z1 = min(nz, z0 + zchunk);
    const int n = min(4, ox - x); // cells of this thread that exist
    double c[4], s[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const size_t m = static_cast<size_t>(y) * ox + x + min(i, n - 1);
        c[i] = cs[m];
        s[i] = sn[m];
    }
    const size_t o = static_cast<size_t>(z0) * level + static_cast<size_t>(y) * pitch + x;
    float* pu = u + o;
    float* pv = v + o;
    for (int z = z0; z < z1; ++z, pu += level, pv += level) {
        if (VEC && n == 4) {
            const float4 a = *reinterpret_cast<const float4*>(pu), b = *reinterpret_cast<const float4*>(pv);
            float4 ra, rb;
            rotate_uv(a.x, b.x, c[0], s[0], ra.x, rb.x);
            rotate_uv(a.y, b.y, c[1], s[1], ra.y, rb.y);
            rotate_uv(a.z, b.z, c[2], s[2], ra.z, rb.z);
            rotate_uv(a.w, b.w, c[3], s[3], ra.w, rb.w);
            __stcs(reinterpret_cast<float4*>(pu), ra);
            __stcs(reinterpret_cast<float4*>(pv), rb);
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if (i < n) {
                    float ur, vr;
                    rotate_uv(pu[i], pv[i], c[i], s[i], ur, vr);
                    __stcs(pu + i, ur);
                    __stcs(pv + i, vr);
                }
        }
    }
}

__global__ void __launch_bounds__(256) direction_kernel(const double* __restrict__ deg, float* __restrict__ ang, size_t layer, int nz, int zchunk)
{
    const size_t i = blockIdx.x * 256ull + threadIdx.x;
    if (i >= layer)
        return;
    const int z0 = blockIdx.y * zchunk, z1 = min(nz, z0 + zchunk);
    const double rot = deg[i];
    float* p = ang + static_cast<size_t>(z0) * layer + i;
#pragma unroll 4
    for (int z = z0; z < z1; ++z, p += layer) {
        double a = __dsub_rn(static_cast<double>(*p), rot);
        if (a < 0)
            a = __dadd_rn(a, 360.);
        if (a > 360)
            a = __dsub_rn(a, 360.);
        __stcs(p, __double2float_rn(a));
    }
}

static int zchunk_for(size_t layer, int nz)
{

    const size_t blocks_xy = (layer + 255) / 256;
    const size_t want = static_cast<size_t>(sm_count()) * 16;
    if (blocks_xy >= want || nz <= 1)
        return std::min(nz, 8); // 8 levels per block: 0.72 against 0.75 ms for [65][1021][2891] with all levels in one block (the tables are re-read per block)
    size_t chunks = std::min<size_t>((want + blocks_xy - 1) / blocks_xy, static_cast<size_t>(nz));
    return static_cast<int>((nz + chunks - 1) / chunks);
}

static fimexcu_vecrot* vecrot_build(const double* d_m3, size_t ox, size_t oy, cudaStream_t st)
{
    std::unique_ptr<fimexcu_vecrot> r(new fimexcu_vecrot());
    r->ox = ox;
    r->oy = oy;
    FXCU_CUDA_NULL(cudaGetDevice(&r->device));
    const size_t n = ox * oy;
    if (!r->c.alloc_sync(n * sizeof(double)) || !r->s.alloc_sync(n * sizeof(double)) || !r->deg.alloc_sync(n * sizeof(double)))
        return nullptr;
    if (n) {
        split_matrix_kernel<<<blocks_for(n, 256), 256, 0, st>>>(d_m3, n, r->c.as<double>(), r->s.as<double>(), r->deg.as<double>());
        FXCU_CUDA_NULL(cudaGetLastError());
        FXCU_CUDA_NULL(cudaStreamSynchronize(st));
    }
    return r.release();
}

} // namespace fxcu

using namespace fxcu;

extern "C" {

fimexcu_vecrot* fimexcu_vecrot_create_dev(const double* d_matrix3, size_t ox, size_t oy, void* stream)
{
    clear_error();
    if (!require_device())
        return nullptr;
    return vecrot_build(d_matrix3, ox, oy, as_stream(stream));
}

fimexcu_vecrot* fimexcu_vecrot_create(const double* matrix3, size_t ox, size_t oy)
{
    clear_error();
    if (!require_device())
        return nullptr;
    DevBuf m;
    const size_t bytes = ox * oy * 3 * sizeof(double);
    if (!m.alloc_sync(bytes))
        return nullptr;
    FXCU_CUDA_NULL(cudaMemcpy(m.p, matrix3, bytes, cudaMemcpyHostToDevice));
    return vecrot_build(m.as<double>(), ox, oy, nullptr);
}

int fimexcu_vecrot_apply_dev_f(const fimexcu_vecrot* r, float* d_u, float* d_v, size_t nz, void* stream)
{
    clear_error();
    if (!r)
        FXCU_FAIL("fimexcu_vecrot_apply: null handle");
    if (!(r = on_current_device(r, "fimexcu_vecrot_apply")))
        return FIMEXCU_ERROR;
    const size_t layer = r->ox * r->oy;
    if (layer == 0 || nz == 0)
        return FIMEXCU_OK;
    const int inz = static_cast<int>(nz);
    const int zc = zchunk_for(layer, inz);
    const dim3 grid(blocks_for(layer, 256), (inz + zc - 1) / zc);
    rotate_kernel<<<grid, 256, 0, as_stream(stream)>>>(r->c.as<double>(), r->s.as<double>(), d_u, d_v, layer, inz, zc);
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}

int fimexcu_vecrot_apply_dev_strided_f(const fimexcu_vecrot* r, float* d_u, float* d_v, size_t row_pitch, size_t level_stride, size_t nz, void* stream)
{
    clear_error();
    if (!r)
        FXCU_FAIL("fimexcu_vecrot_apply: null handle");
    if (!(r = on_current_device(r, "fimexcu_vecrot_apply")))
        return FIMEXCU_ERROR;
    if (r->ox == 0 || r->oy == 0 || nz == 0)
        return FIMEXCU_OK;
    if (row_pitch < r->ox || level_stride < row_pitch * (r->oy - 1) + r->ox)
        FXCU_FAIL("fimexcu_vecrot_apply: row pitch / level stride smaller than the grid");
    if (r->ox > 0x7fffffff || r->oy > 0x7fffffff || nz > 0x7fffffff)
        FXCU_FAIL("fimexcu_vecrot_apply: grid too large");
    const bool vec = row_pitch % 4 == 0 && level_stride % 4 == 0 && ((reinterpret_cast<uintptr_t>(d_u) | reinterpret_cast<uintptr_t>(d_v)) & 15u) == 0;
    const int inz = static_cast<int>(nz);
    const size_t blocks_xy = ((r->ox + 255) / 256) * ((r->oy + 3) / 4);
    const int zc = zchunk_for(blocks_xy * 256, inz);
    const dim3 grid(static_cast<unsigned>((r->ox + 255) / 256), static_cast<unsigned>((r->oy + 3) / 4), static_cast<unsigned>((inz + zc - 1) / zc));
    if (grid.y > 65535 * 1u || grid.z > 65535u)
        FXCU_FAIL("fimexcu_vecrot_apply: grid too large");
    if (vec)
        rotate_rows_kernel<true><<<grid, 256, 0, as_stream(stream)>>>(r->c.as<double>(), r->s.as<double>(), d_u, d_v, static_cast<int>(r->ox), static_cast<int>(r->oy),
                                                                      row_pitch, level_stride, inz, zc);
    else
        rotate_rows_kernel<false><<<grid, 256, 0, as_stream(stream)>>>(r->c.as<double>(), r->s.as<double>(), d_u, d_v, static_cast<int>(r->ox), static_cast<int>(r->oy),
                                                                       row_pitch, level_stride, inz, zc);
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}

int fimexcu_vecrot_apply_f(const fimexcu_vecrot* r, float* u, float* v, size_t size)
{
    clear_error();
    if (!r)
        FXCU_FAIL("fimexcu_vecrot_apply: null handle");
    if (!require_device())
        return FIMEXCU_ERROR;
    const size_t layer = r->ox * r->oy;
    if (layer == 0)
        return FIMEXCU_OK;
    const size_t nz = size / layer; // CachedVectorReprojection.cc:46
    if (nz == 0)
        return FIMEXCU_OK;
    const size_t bytes = nz * layer * sizeof(float);
    DevBuf du, dv;
    if (!du.alloc(bytes, nullptr) || !dv.alloc(bytes, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(du.p, u, bytes, cudaMemcpyHostToDevice, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(dv.p, v, bytes, cudaMemcpyHostToDevice, nullptr));
    if (fimexcu_vecrot_apply_dev_f(r, du.as<float>(), dv.as<float>(), nz, nullptr) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(u, du.p, bytes, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaMemcpyAsync(v, dv.p, bytes, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

int fimexcu_vecrot_direction_dev_f(const fimexcu_vecrot* r, float* d_angles, size_t nz, void* stream)
{
    clear_error();
    if (!r)
        FXCU_FAIL("fimexcu_vecrot_direction: null handle");
    if (!(r = on_current_device(r, "fimexcu_vecrot_direction")))
        return FIMEXCU_ERROR;
    const size_t layer = r->ox * r->oy;
    if (layer == 0 || nz == 0)
        return FIMEXCU_OK;
    const int inz = static_cast<int>(nz);
    const int zc = zchunk_for(layer, inz);
    const dim3 grid(blocks_for(layer, 256), (inz + zc - 1) / zc);
    direction_kernel<<<grid, 256, 0, as_stream(stream)>>>(r->deg.as<double>(), d_angles, layer, inz, zc);
    FXCU_CUDA(cudaGetLastError());
    return FIMEXCU_OK;
}

int fimexcu_vecrot_direction_f(const fimexcu_vecrot* r, float* angles, size_t size)
{
    clear_error();
    if (!r)
        FXCU_FAIL("fimexcu_vecrot_direction: null handle");
    if (!require_device())
        return FIMEXCU_ERROR;
    const size_t layer = r->ox * r->oy;
    if (layer == 0)
        return FIMEXCU_OK;
    const size_t nz = size / layer;
    if (nz == 0)
        return FIMEXCU_OK;
    const size_t bytes = nz * layer * sizeof(float);
    DevBuf d;
    if (!d.alloc(bytes, nullptr))
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(d.p, angles, bytes, cudaMemcpyHostToDevice, nullptr));
    if (fimexcu_vecrot_direction_dev_f(r, d.as<float>(), nz, nullptr) != FIMEXCU_OK)
        return FIMEXCU_ERROR;
    FXCU_CUDA(cudaMemcpyAsync(angles, d.p, bytes, cudaMemcpyDeviceToHost, nullptr));
    FXCU_CUDA(cudaStreamSynchronize(nullptr));
    return FIMEXCU_OK;
}

void fimexcu_vecrot_destroy(fimexcu_vecrot* r) { delete r; }

int fimexcu_vecrot_replicate(fimexcu_vecrot* v, int n_devices, const int* device_ids)
{
    clear_error();
    if (!v)
        FXCU_FAIL("fimexcu_vecrot_replicate: null handle");
    if (n_devices < 0 || (n_devices > 0 && !device_ids))
        FXCU_FAIL("fimexcu_vecrot_replicate: bad device list");
    int before = -1;
    FXCU_CUDA(cudaGetDevice(&before));
    int rc = FIMEXCU_OK;
    for (int k = 0; k < n_devices && rc == FIMEXCU_OK; ++k) {
        const int d = device_ids[k];
        if (d < 0 || d >= 64) {
            set_error("fimexcu_vecrot_replicate: device ordinal out of range");
            rc = FIMEXCU_ERROR;
            break;
        }
        if (d == v->device || (static_cast<size_t>(d) < v->replicas.size() && v->replicas[d]))
            continue;
        if (!cuda_ok(cudaSetDevice(d), "cudaSetDevice", __FILE__, __LINE__) || !require_device()) {
            rc = FIMEXCU_ERROR;
            break;
        }
        std::unique_ptr<fimexcu_vecrot> r(new fimexcu_vecrot());
        r->ox = v->ox;
        r->oy = v->oy;
        r->device = d;
        if (!replicate_buf(r->c, d, v->c, v->device, nullptr) || !replicate_buf(r->s, d, v->s, v->device, nullptr) ||
            !replicate_buf(r->deg, d, v->deg, v->device, nullptr) || !cuda_ok(cudaStreamSynchronize(nullptr), "cudaStreamSynchronize", __FILE__, __LINE__)) {
            rc = FIMEXCU_ERROR;
            break;
        }
        if (v->replicas.size() <= static_cast<size_t>(d))
            v->replicas.resize(d + 1, nullptr);
        v->replicas[d] = r.release();
    }
    cudaSetDevice(before);
    return rc;
}

} // extern "C"
