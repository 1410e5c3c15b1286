/*
 * fimexcu.h -- C ABI of libfimexcu.so: the fimex regridding hot path on B200 (sm_100a).
 *
 * This is the drop-in boundary: plain pointers and sizes, opaque handles, int status codes; no C++,
 * torch or CUDA types.  Every entry point names the reference interface it replaces
 * (fimex 2.4.1, paths relative to the reference tree).  INTEGRATION.md shows the few lines a fimex
 * maintainer adds on the reference side to bind them.
 *
 * Conventions
 *   - status: FIMEXCU_OK (1, == MIFI_OK) or FIMEXCU_ERROR (-1, == MIFI_ERROR, mifi_constants.h:259-261);
 *     creators return NULL on error.  fimexcu_last_error() gives the message for the calling thread.
 *   - fields are C-order [z][y][x], x fastest (mifi_3d_array_pos, interpolation.c:229-233), raw f32
 *     values with NaN = undefined; z = product of all non-horizontal dimensions of the slab.
 *   - "_dev" entry points take DEVICE pointers of the current CUDA device and an optional stream
 *     (a cudaStream_t passed as void*, NULL = default stream); they enqueue and return.
 *     The others take HOST pointers, copy in and out, and return when the result is in the output.
 *   - handles are immutable after creation and may be used from several host threads at once.
 *   - there is NO CPU fallback: every call fails with FIMEXCU_ERROR when no sm_100 device is usable.
 */
#ifndef FIMEXCU_H
#define FIMEXCU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FIMEXCU_OK 1
#define FIMEXCU_ERROR (-1)

/* interpolation methods: enum mifi_interpol_method, include/fimex/mifi_constants.h:52-138 */
enum fimexcu_interpol_method {
    FIMEXCU_INTERPOL_NEAREST_NEIGHBOR = 0,
    FIMEXCU_INTERPOL_BILINEAR = 1,
    FIMEXCU_INTERPOL_BICUBIC = 2,
    FIMEXCU_INTERPOL_COORD_NN = 3,
    FIMEXCU_INTERPOL_COORD_NN_KD = 4,
    FIMEXCU_INTERPOL_FORWARD_SUM = 5,
    FIMEXCU_INTERPOL_FORWARD_MEAN = 6,
    FIMEXCU_INTERPOL_FORWARD_MEDIAN = 7,
    FIMEXCU_INTERPOL_FORWARD_MAX = 8,
    FIMEXCU_INTERPOL_FORWARD_MIN = 9,
    FIMEXCU_INTERPOL_FORWARD_UNDEF_SUM = 10,
    FIMEXCU_INTERPOL_FORWARD_UNDEF_MEAN = 11,
    FIMEXCU_INTERPOL_FORWARD_UNDEF_MEDIAN = 12,
    FIMEXCU_INTERPOL_FORWARD_UNDEF_MAX = 13,
    FIMEXCU_INTERPOL_FORWARD_UNDEF_MIN = 14
};
/* vertical methods: enum mifi_vertical_interpol_method, mifi_constants.h:206-231 */
enum fimexcu_vint_method {
    FIMEXCU_VINT_METHOD_LIN = 0,
    FIMEXCU_VINT_METHOD_LOG = 1,
    FIMEXCU_VINT_METHOD_LOGLOG = 2,
    FIMEXCU_VINT_METHOD_NN = 3,
    FIMEXCU_VINT_METHOD_LIN_WEAK_EXTRA = 4,
    FIMEXCU_VINT_METHOD_LIN_NO_EXTRA = 5,
    FIMEXCU_VINT_METHOD_LIN_CONST_EXTRA = 6
};
/* axis types: mifi_constants.h:264-268 */
enum fimexcu_axis_type { FIMEXCU_PROJ_AXIS = 0, FIMEXCU_LONGITUDE = 1, FIMEXCU_LATITUDE = 2 };

/* ------------------------------------------------------------------------------------------------
 * runtime
 * ---------------------------------------------------------------------------------------------- */
/* message of the last failed call on this thread ("" if none).  Replaces the CDMException text thrown at
 * src/CachedInterpolation.cc:197-203 / src/reproject.cc:65-70. */
const char* fimexcu_last_error(void);
/* number of usable sm_100 devices; replaces mifi_setNumThreads()'s role of sizing the worker pool
 * (src/ThreadPool.c:33-47): slabs are dealt to devices instead of OpenMP threads. */
int fimexcu_device_count(void);
/* make device `ordinal` current for the calling thread (one process or thread per GPU). */
int fimexcu_set_device(int ordinal);
/* pinned host memory for slabs (so the host<->device copies of the *_f entry points run at full PCIe
 * speed); plays the part of make_shared_array<float>() in include/fimex/SharedArray.h:42-48.  The reference allocates a fresh
 * result array in every interpolateValues() (src/CachedInterpolation.cc:177); pinning 767 MB per call would cost more than
 * the regridding, so the memory comes from a pool: fimexcu_host_free() parks the block in a size-class free list (up to
 * FIMEXCU_HOST_POOL_KEEP_MB, default 8192, stay parked) and the next fimexcu_host_alloc() of that class takes it back without
 * a driver call.  Blocks are portable (usable with every device of the process).  Thread safe. */
void* fimexcu_host_alloc(size_t bytes);
void fimexcu_host_free(void* p);
/* give the parked blocks back to the operating system */
void fimexcu_host_pool_trim(void);
/* plain device memory for callers that keep slabs resident between chained operators */
void* fimexcu_dev_alloc(size_t bytes);
void fimexcu_dev_free(void* p);
int fimexcu_memcpy_h2d(void* dst_dev, const void* src_host, size_t bytes);
int fimexcu_memcpy_d2h(void* dst_host, const void* src_dev, size_t bytes);
int fimexcu_memcpy_d2d(void* dst_dev, const void* src_dev, size_t bytes);
/* dense host slab [nz][ny][width] <-> device slab with a row pitch and a level stride (bytes): how a slab enters / leaves the padded
 * device layout of fimexcu_cache_padded_layout at the ends of a device-resident reader chain (fimex_b200/host/CudaSlab.h) */
int fimexcu_memcpy2d_h2d(void* dst_dev, size_t dst_pitch_bytes, const void* src_host, size_t src_pitch_bytes, size_t width_bytes, size_t ny, size_t nz,
                         size_t dst_level_bytes);
int fimexcu_memcpy2d_d2h(void* dst_host, size_t dst_pitch_bytes, const void* src_dev, size_t src_pitch_bytes, size_t width_bytes, size_t ny, size_t nz,
                         size_t src_level_bytes);
int fimexcu_synchronize(void);

/* ------------------------------------------------------------------------------------------------
 * horizontal interpolation cache (bilinear / bicubic / nearest / coord_nn / coord_kdtree)
 * ---------------------------------------------------------------------------------------------- */
typedef struct fimexcu_cache fimexcu_cache;

/* Replaces createCachedInterpolation() (src/CachedInterpolation.cc:132-147) with the constructors of
 * CachedInterpolation (:149-171, reduced domain :222-253) and CachedNNInterpolation (:262-315).
 * pointsOnXAxis/pointsOnYAxis: outX*outY fractional source indices (HOST, not modified; the reference
 * rebases its arrays in place -- read the offsets with fimexcu_cache_info()). */
fimexcu_cache* fimexcu_cache_create(int method, const double* pointsOnXAxis, const double* pointsOnYAxis, size_t inX, size_t inY,
                                    size_t outX, size_t outY);
/* same with DEVICE tables (e.g. straight from fimexcu_project_axes_dev + fimexcu_points2position_dev) */
fimexcu_cache* fimexcu_cache_create_dev(int method, const double* d_pointsOnXAxis, const double* d_pointsOnYAxis, size_t inX, size_t inY,
                                        size_t outX, size_t outY, void* stream);
/* getInX/getInY/getOutX/getOutY and reducedDomain() of CachedInterpolationInterface
 * (include/fimex/CachedInterpolation.h:63-103): inX/inY are the sizes of the (possibly reduced) input
 * window the apply calls expect, xMin/yMin its offset in the full source grid. */
int fimexcu_cache_info(const fimexcu_cache* c, size_t* inX, size_t* inY, size_t* outX, size_t* outY, int* reduced, size_t* xMin, size_t* yMin);
/* Replaces CachedInterpolationInterface::interpolateValues(inData, size, newSize)
 * (CachedInterpolation.cc:173-209 bilinear/bicubic via interpolation.c:267-414; :317-337 nearest) with
 * mifi_bad2nanf (interpolation.c:1064-1072, caller: CDMInterpolator.cc:140-145) fused on load and
 * mifi_nanf2bad (:1074-1082; caller interpolationArray2Data, CDMInterpolator.cc:147-151) fused on store.
 *   in   HOST [nz][inY][inX] f32, size = nz*inX*inY;  out HOST [nz][outY][outX] f32;  *newSize = nz*outX*outY
 *   badIn   values equal to it are read as NaN (pass NaN for "none")
 *   fillOut NaN results are written as this value (pass NaN to keep NaN) */
int fimexcu_cache_apply_f(const fimexcu_cache* c, const float* in, size_t size, float* out, size_t* newSize, float badIn, float fillOut);
/* `in` / `out` of the host entry points may be ordinary pageable memory (the reference hands `new float[]` arrays,
 * include/fimex/SharedArray.h:42-48): such a slab is moved through pinned bounce buffers of the pool above by the host cores,
 * chunk by chunk, overlapped with the DMA of the neighbouring chunk.  Pinned memory (fimexcu_host_alloc) skips that step. */
/* interpolateValues() followed by getOutputDataSlice() (src/CachedInterpolation.cc:113-130) for callers that asked for a
 * sub-window of the target grid (cross-sections, point lists, tiled writers): only the cells [yStart, yStart+ySize) x
 * [xStart, xStart+xSize) are computed and copied back; out is [nz][ySize][xSize], *newSize = nz*xSize*ySize. */
int fimexcu_cache_apply_window_f(const fimexcu_cache* c, const float* in, size_t size, size_t xStart, size_t xSize, size_t yStart, size_t ySize,
                                 float* out, size_t* newSize, float badIn, float fillOut);
int fimexcu_cache_apply_window_dev_f(const fimexcu_cache* c, const float* d_in, size_t nz, size_t xStart, size_t xSize, size_t yStart, size_t ySize,
                                     float* d_out, float badIn, float fillOut, void* stream);
/* same on DEVICE-resident slabs in the reference's dense layout; enqueues on `stream` */
int fimexcu_cache_apply_dev_f(const fimexcu_cache* c, const float* d_in, size_t nz, float* d_out, float badIn, float fillOut, void* stream);
/* same on DEVICE-resident slabs with explicit element strides: element (z, y, x) of the source lives at
 * d_in[z*in_level_stride + y*in_row_pitch + x], of the target at d_out[z*out_level_stride + y*out_row_pitch + x].
 * Any strides >= the grid sizes are accepted.  The fast (TMA-staged) bilinear kernel needs d_in on a 16 B boundary and
 * in_row_pitch, in_level_stride multiples of 4; target rows on 128 B boundaries make every warp store one full line.
 * fimexcu_cache_padded_layout() returns strides that satisfy both (what slabs that stay on the device between chained
 * operators should use).  The host entry points (*_f, *_i16) do NOT go through this layout: they are bound by the PCIe copies,
 * plain 1-D copies of the dense host layout run at full DMA speed and strided copies into padded rows do not, so their level
 * chunks keep the dense layout on the device. */
int fimexcu_cache_apply_dev_strided_f(const fimexcu_cache* c, const float* d_in, size_t in_row_pitch, size_t in_level_stride, size_t nz,
                                      float* d_out, size_t out_row_pitch, size_t out_level_stride, float badIn, float fillOut, void* stream);
int fimexcu_cache_padded_layout(const fimexcu_cache* c, size_t* in_row_pitch, size_t* in_level_stride, size_t* out_row_pitch,
                                size_t* out_level_stride);
/* Where a result of this cache can hold data at all.  A target cell whose table entry has no source cell is undefined whatever the
 * input (CachedInterpolation.cc:176-209 writes NaN there on every call).  The target rows are grouped into at most 24 bands, band k
 * = rows [y0, y1) with the column span [x0, x1) outside which all of its rows are undefined (x0 == x1: the whole band is).  The host
 * entry points (*_f, *_i16) copy only these rectangles back from the device and write the undefined value (NaN / fillOut / the
 * packed fill) into the rest of the caller's array on the host cores meanwhile -- for a lat/lon target around a Lambert source
 * (BASELINE configs[1]: 36.6 % of the target outside the source) a third of the result never crosses PCIe.  Results are
 * identical; FIMEXCU_E2E_BANDS=0 switches it off.  bands4 receives up to max_bands quadruples {y0, y1, x0, x1}; returns the
 * number of bands (0: not analysed), *fraction = area of the rectangles / area of the target. */
int fimexcu_cache_row_bands(const fimexcu_cache* c, int* bands4, int max_bands, double* fraction);
/* Packed 16-bit slabs (slab feed / sink with fused unpack and pack).  Replaces, for a CDM_SHORT variable,
 * the whole of CDMInterpolator::getDataSlice around the cached interpolation (src/CDMInterpolator.cc:140-151, 188-260):
 *   data2InterpolationArray   Data::asFloat() of the RAW packed values, mifi_bad2nanf(badValue)
 *   interpolateValues         as fimexcu_cache_apply_f
 *   interpolationArray2Data   convertDataType(NaN,1,0, CDM_SHORT, badValue,1,0) = ScaleValue<float,short>
 *                             (include/fimex/DataUtils.h:38-60): NaN -> (short)badValue, else
 *                             static_cast<short>(int(lround(double(v)))) (include/fimex/MathUtils.h:40-96)
 * so only 2 bytes per value cross PCIe and HBM in either direction.  badValue = the variable's _FillValue
 * (CDM::getFillValue), used for both sides like the reference.  Layouts as fimexcu_cache_apply_f / _dev_f with int16 elements. */
int fimexcu_cache_apply_i16(const fimexcu_cache* c, const int16_t* in, size_t size, int16_t* out, size_t* newSize, double badValue);
int fimexcu_cache_apply_dev_i16(const fimexcu_cache* c, const int16_t* d_in, size_t nz, int16_t* d_out, double badValue, void* stream);
void fimexcu_cache_destroy(fimexcu_cache* c);
/* One cache for all GPUs of the process.  The reference is ONE process whose writer threads call getDataSlice concurrently
 * (#pragma omp parallel for over the unlimited dimension, src/io/netcdf/NetCDF_CDMWriter.cc:599-618), all through the same
 * CachedInterpolation object.  fimexcu_cache_replicate() copies the handle's tables (and tile metadata) from the device they
 * were built on to every listed device (device-to-device copies; about 35 MB for the 2.95 M-cell bilinear T1 tables); from then on
 * every apply entry point uses the copy that lives on the CALLING THREAD's current device (fimexcu_set_device), so host threads
 * bound to different GPUs share one handle.  Devices already covered are skipped; the copies die with the handle.  Not thread
 * safe against concurrent applies on the same handle: replicate once, right after the cache is built. */
int fimexcu_cache_replicate(fimexcu_cache* c, int n_devices, const int* device_ids);
/* Opt-in arithmetic for bicubic caches on device-resident padded slabs (the staged kernel): on = 1 computes the 4 x 4 convolution
 * with the weights narrowed to f32 and 20 fused multiply-adds per value instead of the reference's chain of 40 separately rounded
 * f64 operations with an f32 accumulator (src/interpolation.c:388-406), which the FP64 pipe bounds at 40 % of the HBM roofline
 * (the bit-exact form runs at 23 %, this one at ~40 %).
 * Results then agree with the reference within BASELINE.json's 1e-5 (relative to the largest tap) instead of bit for bit;
 * undefined masks stay identical.  Default 0 (bit-exact).  No counterpart in the reference. */
int fimexcu_cache_set_fast_bicubic(fimexcu_cache* c, int on);
/* A/B switch between the staged (TMA) and the direct-gather apply kernels, for measurements and tests: 1 = staged kernels where they
 * apply (default), 0 = direct kernels only, -1 = back to the default (which the environment variable FIMEXCU_NO_TMA, read once per
 * process, can set to 0).  No counterpart in the reference. */
void fimexcu_set_staged_kernels(int mode);

/* Slab feed / sink for packed integer variables as stand-alone device passes, so that packed slabs cross
 * PCIe in their file type and any operator of the path can be chained on the device in f32:
 *   fimexcu_unpack_dev_f         data2InterpolationArray (src/CDMInterpolator.cc:140-145): Data::asFloat() of the raw values, fill -> NaN
 *   fimexcu_unpack_scaled_dev_f  getScaledDataSlice (src/CDMReader.cc:177-192): fill -> NaN, scale*v + offset in f64, narrowed like Data::asFloat()
 *   fimexcu_pack_dev_f           interpolationArray2Data (src/CDMInterpolator.cc:147-151): NaN -> fill, lround, the C casts of ScaleValue<float,T> */
enum fimexcu_dtype {
    FIMEXCU_DTYPE_I8 = 0,  /* CDM_CHAR (NC_BYTE) */
    FIMEXCU_DTYPE_I16 = 1, /* CDM_SHORT */
    FIMEXCU_DTYPE_I32 = 2, /* CDM_INT */
    FIMEXCU_DTYPE_U8 = 3,  /* CDM_UCHAR */
    FIMEXCU_DTYPE_U16 = 4, /* CDM_USHORT */
    FIMEXCU_DTYPE_U32 = 5  /* CDM_UINT */
};
int fimexcu_unpack_dev_f(int dtype, const void* d_in, size_t n, double badValue, float* d_out, void* stream);
int fimexcu_unpack_scaled_dev_f(int dtype, const void* d_in, size_t n, double fill, double scale, double offset, float* d_out, void* stream);
int fimexcu_pack_dev_f(int dtype, const float* d_in, size_t n, double badValue, void* d_out, void* stream);

/* Drop-in for the per-point C kernels with the reference signatures (include/fimex/interpolation.h:95-137,
 * src/interpolation.c:251-414): one (x, y) position over iz levels.  infield/outvalues are HOST pointers.
 * Provided for callers of the uncached path (reproject::reproject_f, src/reproject.cc:266-295); one launch
 * per call -- use the cache for anything but single points. */
int fimexcu_get_values_f(const float* infield, float* outvalues, double x, double y, int ix, int iy, int iz);
int fimexcu_get_values_bilinear_f(const float* infield, float* outvalues, double x, double y, int ix, int iy, int iz);
int fimexcu_get_values_bicubic_f(const float* infield, float* outvalues, double x, double y, int ix, int iy, int iz);

/* ------------------------------------------------------------------------------------------------
 * vector (wind) rotation
 * ---------------------------------------------------------------------------------------------- */
typedef struct fimexcu_vecrot fimexcu_vecrot;

/* Replaces CachedVectorReprojection(reproject::Matrix_cp) (src/CachedVectorReprojection.cc:37-42).
 * matrix3: HOST ox*oy*3 doubles (cos, sin, angle in degree), the layout of reproject::Matrix
 * (include/fimex/reproject.h:34-46). */
fimexcu_vecrot* fimexcu_vecrot_create(const double* matrix3, size_t ox, size_t oy);
fimexcu_vecrot* fimexcu_vecrot_create_dev(const double* d_matrix3, size_t ox, size_t oy, void* stream);
/* Replaces CachedVectorReprojection::reprojectValues(u, v, size) -> reproject::vector_reproject_values_by_matrix_f
 * (CachedVectorReprojection.cc:44-48, reproject.cc:488-507): in place, size = nz*ox*oy. */
int fimexcu_vecrot_apply_f(const fimexcu_vecrot* r, float* u, float* v, size_t size);
int fimexcu_vecrot_apply_dev_f(const fimexcu_vecrot* r, float* d_u, float* d_v, size_t nz, void* stream);
/* the same in place on device slabs with a row pitch and a level stride in elements (the padded rows of fimexcu_cache_padded_layout,
 * as a CDMProcessor-style rotation behind a device-resident horizontal step meets them: src/CDMProcessor.cc:114-157); the pad
 * elements of the rows are not touched */
int fimexcu_vecrot_apply_dev_strided_f(const fimexcu_vecrot* r, float* d_u, float* d_v, size_t row_pitch, size_t level_stride, size_t nz, void* stream);
/* Replaces CachedVectorReprojection::reprojectDirectionValues -> vector_reproject_direction_by_matrix_f
 * (CachedVectorReprojection.cc:50-54, reproject.cc:509-530). */
int fimexcu_vecrot_direction_f(const fimexcu_vecrot* r, float* angles, size_t size);
int fimexcu_vecrot_direction_dev_f(const fimexcu_vecrot* r, float* d_angles, size_t nz, void* stream);
void fimexcu_vecrot_destroy(fimexcu_vecrot* r);
/* the rotation matrix on every listed device, like fimexcu_cache_replicate (CachedVectorReprojection is shared by the writer threads too) */
int fimexcu_vecrot_replicate(fimexcu_vecrot* r, int n_devices, const int* device_ids);
/* The vector branch of CDMInterpolator::getDataSlice (src/CDMInterpolator.cc:232-259) as ONE pass:
 * interpolate u and v with the same cache and rotate them, instead of the reference's two
 * interpolateValues() calls per component plus reprojectValues().  r may be NULL (no rotation).  The host form is pipelined like
 * fimexcu_cache_apply_f: level chunks of both components on rotating streams, pageable arrays through pinned bounce chunks, only
 * the row bands of the target that can hold data copied back (fimexcu_cache_row_bands). */
int fimexcu_cache_apply_uv_f(const fimexcu_cache* c, const fimexcu_vecrot* r, const float* u_in, const float* v_in, size_t size, float* u_out,
                             float* v_out, size_t* newSize, float badU, float badV, float fillU, float fillV);
int fimexcu_cache_apply_uv_dev_f(const fimexcu_cache* c, const fimexcu_vecrot* r, const float* d_u_in, const float* d_v_in, size_t nz,
                                 float* d_u_out, float* d_v_out, float badU, float badV, float fillU, float fillV, void* stream);
/* same with explicit element strides for both sources and both targets (see fimexcu_cache_apply_dev_strided_f); with the
 * padded layout the bilinear case runs as one TMA-staged kernel. */
int fimexcu_cache_apply_uv_dev_strided_f(const fimexcu_cache* c, const fimexcu_vecrot* r, const float* d_u_in, const float* d_v_in,
                                         size_t in_row_pitch, size_t in_level_stride, size_t nz, float* d_u_out, float* d_v_out,
                                         size_t out_row_pitch, size_t out_level_stride, float badU, float badV, float fillU, float fillV,
                                         void* stream);

/* ------------------------------------------------------------------------------------------------
 * cache build: axes, projections, rotation matrix
 * ---------------------------------------------------------------------------------------------- */
/* Replaces mifi_points2position (include/fimex/interpolation.h:76-93, src/interpolation.c:154-227):
 * coordinates -> fractional axis indices, in place.  HOST pointers. */
int fimexcu_points2position(double* points, size_t n, const double* axis, int num, int axis_type);
int fimexcu_points2position_dev(double* d_points, size_t n, const double* axis_host, int num, int axis_type, void* stream);
/* No counterpart in the reference.  Same result as fimexcu_points2position_dev, for callers that must not allocate GPU
 * memory while they run: a reader chain embedded in a host application that has handed the whole device to its own
 * memory pool at start-up (an inference or visualisation server with a fixed budget), where cudaMalloc fails, or a
 * latency-sensitive thread that cannot afford the device-wide synchronisation a cudaMalloc/cudaFree may imply.  The
 * device copy of the axis, which the other entry point keeps in a freshly allocated buffer, lives here in a fixed
 * array inside the library image: FIMEXCU_NOALLOC_MAX_AXIS doubles (128 KiB) on every GPU that loads the library,
 * held until the process exits.  The call allocates nothing (no cudaMalloc/cudaFree, no pool, no torch); the caller
 * still owns d_points (in/out, device) and axis_host (host).  Longer axes are rejected with FIMEXCU_ERROR.
 * The first call on each device looks up the array's device address; make that first call before any stream capture
 * begins.  All calls share the one array and it keeps the previous call's axis: a call overwrites the part it uses
 * before reading it, and two calls whose GPU work overlaps (different streams or threads) race on it -- serialise
 * them.  The call returns after `stream` has drained, like fimexcu_points2position_dev. */
#define FIMEXCU_NOALLOC_MAX_AXIS 16384
int fimexcu_points2position_dev_noalloc(double* d_points, size_t n, const double* axis_host, int num, int axis_type, void* stream);
/* Replaces reproject::reproject_axes (include/fimex/reproject.h:211-227, src/reproject.cc:573-587): broadcast the
 * 1-D axes to an ix*iy lattice and transform proj_input -> proj_output (proj4 strings; degrees for geographic
 * CRS as with PROJ >= 6).  Points that cannot be projected become +inf.
 * Understood: every projection fimex's Projection classes emit (src/coordSys/ ...Projection.cc) -- longlat, lcc, stere / ups,
 * ob_tran over longlat, merc, tmerc / utm, laea, aea, aeqd, cea, ortho, sinu, nsper, geos, omerc -- on a sphere or an
 * ellipsoid (+R | +a +b/+rf/+es | +ellps | +datum=WGS84), without datum shifts.  A string that asks for anything else
 * (+towgs84 other than zero, +nadgrids, +pm, +axis, another projection, oblique aeqd or ortho on an ellipsoid, the two-point
 * omerc) is an error ("Proj error: ..."), never an approximation. */
int fimexcu_reproject_axes(const char* proj_input, const char* proj_output, const double* in_x_axis, const double* in_y_axis, int ix, int iy,
                           double* out_xproj_axis, double* out_yproj_axis);
int fimexcu_reproject_axes_dev(const char* proj_input, const char* proj_output, const double* in_x_axis_host, const double* in_y_axis_host, int ix,
                               int iy, double* d_out_xproj_axis, double* d_out_yproj_axis, void* stream);
/* Replaces reproject::reproject_values (reproject.h:195-209, reproject.cc:567-571): n points in place. */
int fimexcu_reproject_values(const char* proj_input, const char* proj_output, double* in_out_x_vals, double* in_out_y_vals, int num);
int fimexcu_reproject_values_dev(const char* proj_input, const char* proj_output, double* d_x, double* d_y, size_t num, void* stream);
/* Replaces reproject::get_vector_reproject_matrix (reproject.h:146-160, reproject.cc:463-486 with :363-397 and
 * :310-357).  matrix3: ox*oy*3 doubles out. */
int fimexcu_get_vector_reproject_matrix(const char* proj_input, const char* proj_output, const double* out_x_axis, const double* out_y_axis, int ox,
                                        int oy, double* matrix3);
/* Replaces reproject::get_vector_reproject_matrix_field (reproject.h:165-178, reproject.cc:449-461): the grid keeps its cells, only
 * the direction of the vectors changes (rotation of model winds to true north in CDMProcessor, src/CDMProcessor.cc:149);
 * in_x_field / in_y_field: ox*oy positions in the input projection (HOST). */
int fimexcu_get_vector_reproject_matrix_field(const char* proj_input, const char* proj_output, const double* in_x_field, const double* in_y_field, int ox,
                                              int oy, double* matrix3);
/* Replaces reproject::get_vector_reproject_matrix_points (reproject.h:177-193, reproject.cc:432-447). */
int fimexcu_get_vector_reproject_matrix_points(const char* proj_input, const char* proj_output, int inputIsMetric, const double* out_x_points,
                                               const double* out_y_points, int ox, int oy, double* matrix3);
/* One call for the whole of CDMInterpolator::changeProjectionByProjectionParameters' numeric part
 * (src/CDMInterpolator.cc:1389-1404): reproject_axes(out -> in) + 2 x points2position + cache construction,
 * all on the device; the index tables never visit the host. */
fimexcu_cache* fimexcu_cache_create_from_projection(int method, const char* proj_input /* target CRS */, const char* proj_org /* source CRS */,
                                                    const double* out_x_axis, const double* out_y_axis, size_t outX, size_t outY,
                                                    const double* org_x_axis, const double* org_y_axis, size_t inX, size_t inY,
                                                    int org_x_axis_type, int org_y_axis_type);

/* The numeric part of CDMInterpolator::changeProjectionByProjectionParametersToLatLonTemplate (src/CDMInterpolator.cc:1636-1666):
 * the target is a LIST of points (outX*outY of them: a 2-D lat/lon template, a point list with outY = 1, or the sampled lines
 * of changeProjectionToCrossSections, :492-613) given in proj_points (the reference passes "+proj=latlong +datum=WGS84" and
 * degrees): reproject_values(points -> source CRS) + 2 x points2position + cache construction, all on the device
 * (SURVEY 8(f) rank 4). */
fimexcu_cache* fimexcu_cache_create_from_points(int method, const char* proj_points, const char* proj_org, const double* x_points,
                                                const double* y_points, size_t outX, size_t outY, const double* org_x_axis,
                                                const double* org_y_axis, size_t inX, size_t inY, int org_x_axis_type, int org_y_axis_type);

/* ------------------------------------------------------------------------------------------------
 * coordinate-based nearest neighbour (curvilinear sources)
 * ---------------------------------------------------------------------------------------------- */
/* Replaces flannTranslatePointsToClosestInputCell (src/CDMInterpolator.cc:953-1028; nanoflann radiusSearch, first
 * sorted match).  pointsOnXAxis/pointsOnYAxis: in = target lon/lat in degrees, out = source (ix, iy) as doubles,
 * or -1000 when nothing lies within maxDist (metres on the R=6371000 sphere).  HOST pointers.
 * Result identical to the reference's, including what follows from nanoflann's tree layout (include/nanoflann/nanoflann.hpp
 * :1053-1237, :1263-1310) and libstdc++'s std::sort (:972): which of several exactly equidistant source points wins, and which
 * matches are lost when the cloud holds NaN coordinates (the reference keeps them in the tree, CDMInterpolator.cc:982-986).
 * Unit-sphere coordinates come from the host's libm, the tree is built on the host cores, the walk runs on the GPU. */
int fimexcu_kdtree_translate(double maxDist, double* pointsOnXAxis, double* pointsOnYAxis, size_t pointsSize, const double* lonVals,
                             const double* latVals, size_t orgXDimSize, size_t orgYDimSize);
/* Same with a switch, like level_compat of the forward cache: nanoflann_compat = 1 is fimexcu_kdtree_translate;
 * nanoflann_compat = 0 is the corrected behaviour -- source points with NaN coordinates are left out, so that every target gets
 * its true nearest source point within maxDist, and exact ties go to the lower source index (tree-free search on the GPU). */
int fimexcu_kdtree_translate_mode(double maxDist, double* pointsOnXAxis, double* pointsOnYAxis, size_t pointsSize, const double* lonVals,
                                  const double* latVals, size_t orgXDimSize, size_t orgYDimSize, int nanoflann_compat);
/* Replaces fastTranslatePointsToClosestInputCell + getGridDistance (CDMInterpolator.cc:1030-1190): misses give -1.
 * Result identical to the reference's: the GPU search decides every target whose best candidate is clear of the runner-up and of
 * the ROI threshold by more than the device/glibc sin-cos noise (1e-13 in the cosine); the others (near and exact ties, ROI edge)
 * are decided by the reference's own latitude-sorted scan on the host, on a list ordered by this libstdc++'s std::sort. */
int fimexcu_coordnn_translate(double* pointsOnXAxis, double* pointsOnYAxis, size_t pointsSize, const double* lonVals, const double* latVals,
                              size_t orgXDimSize, size_t orgYDimSize);

/* ------------------------------------------------------------------------------------------------
 * forward (scatter / aggregate) interpolation
 * ---------------------------------------------------------------------------------------------- */
typedef struct fimexcu_forward fimexcu_forward;
/* Replaces CachedForwardInterpolation's constructor (src/CachedForwardInterpolation.cc:133-201).  pOnX/pOnY:
 * inX*inY target positions of every SOURCE cell (HOST). */
fimexcu_forward* fimexcu_forward_create(int method, const double* pOnX, const double* pOnY, size_t inX, size_t inY, size_t outX, size_t outY);
int fimexcu_forward_info(const fimexcu_forward* f, size_t* inX, size_t* inY, size_t* outX, size_t* outY, int* reduced, size_t* xMin,
                         size_t* yMin, size_t* maxPointsInIn);
/* Replaces CachedForwardInterpolation::interpolateValues (:205-224).  level_compat != 0 reproduces the reference
 * literally (it reads inData[ii] without the level offset, so every output level aggregates input level 0);
 * 0 aggregates each level from its own input level. */
int fimexcu_forward_apply_f(const fimexcu_forward* f, const float* in, size_t size, float* out, size_t* newSize, float badIn, float fillOut,
                            int level_compat);
int fimexcu_forward_apply_dev_f(const fimexcu_forward* f, const float* d_in, size_t nz, float* d_out, float badIn, float fillOut,
                                int level_compat, void* stream);
void fimexcu_forward_destroy(fimexcu_forward* f);

/* ------------------------------------------------------------------------------------------------
 * vertical interpolation
 * ---------------------------------------------------------------------------------------------- */
/* Replaces the hot loop of CDMVerticalInterpolator::getLevelDataSlice (src/CDMVerticalInterpolator.cc:488-561):
 * neighbour search (include/fimex/FindNeighborElements.h:46-125), the two-level kernel selected at :379-388
 * (src/interpolation.c:416-546), the validity range test (:505-528) and the valid_min/valid_max clamp (:551-561).
 * Layout: in [nOuter][nzi][nInner], vcoord_in the same shape (f32, as :483-486), out [nOuter][nzo][nInner];
 * a column is (outer, inner).  Target: levels[nzo] (f64, `level1`) or, when levels == NULL, vcoord_out
 * [nOuter][nzo][nInner] f32 (template / to-axis).  vmin/vmax: f64 per column (NULL = none).  valid_min /
 * valid_max: NaN = none.  HOST pointers. */
int fimexcu_vertical_interpolate_f(int method, const float* in, const float* vcoord_in, size_t nOuter, size_t nzi, size_t nInner,
                                   const double* levels, const float* vcoord_out, size_t nzo, const double* vmin, const double* vmax,
                                   double valid_min, double valid_max, float badIn, float fillOut, float* out);
int fimexcu_vertical_interpolate_dev_f(int method, const float* d_in, const float* d_vcoord_in, size_t nOuter, size_t nzi, size_t nInner,
                                       const double* levels_host, const float* d_vcoord_out, size_t nzo, const double* d_vmin,
                                       const double* d_vmax, double valid_min, double valid_max, float badIn, float fillOut, float* d_out,
                                       void* stream);
/* Same with the hybrid-sigma pressure formed on the fly (never materialised): replaces
 * HybridSigmaApToPressureConverter::getDataSlice (src/coordSys/verticalTransform/HybridSigmaApToPressureConverter.cc:71-100)
 * -> mifi_atmosphere_hybrid_sigma_ap_pressure (src/vertical_coordinate_transformations.c:62-68) feeding the loop above.
 * p[k] = ap[k] + b[k] * (ps * ps_scale) in f64, narrowed to f32 as at CDMVerticalInterpolator.cc:483.
 * ps: f32 [nOuterPs][nInner] with nOuterPs == nOuter or 1 (broadcast over the outer dimension).
 * For hybrid_sigma (a, p0) pass ap[k] = a[k]*p0.
 * The _dev form enqueues on `stream` and may return before the work is done (like the other _dev entry points): it keeps the
 * device copy of the last coefficient sets (ap, b, levels: a few KB per host thread and device) between calls, and only a call
 * that sees a new set waits for its upload. */
int fimexcu_vertical_hybrid_pressure_f(int method, const float* in, size_t nOuter, size_t nzi, size_t nInner, const double* ap, const double* b,
                                       const float* ps, size_t nOuterPs, double ps_scale, const double* levels, size_t nzo, double valid_min,
                                       double valid_max, float badIn, float fillOut, float* out);
int fimexcu_vertical_hybrid_pressure_dev_f(int method, const float* d_in, size_t nOuter, size_t nzi, size_t nInner, const double* ap_host,
                                           const double* b_host, const float* d_ps, size_t nOuterPs, double ps_scale, const double* levels_host,
                                           size_t nzo, double valid_min, double valid_max, float badIn, float fillOut, float* d_out,
                                           void* stream);
/* The two _dev forms above for slabs whose rows are PADDED (fimexcu_cache_padded_layout: the result of a device-resident horizontal
 * step, handed on without repacking): the inner index runs over rows of rowPitch elements of which only the first rowWidth are
 * columns.  The pad elements are neither read nor written, in any of the arrays (in, vcoord_in / ps, vcoord_out, vmin / vmax,
 * out all share the layout) -- whatever they hold (they are never initialised by the horizontal kernels) costs nothing and cannot
 * reach the surface-pressure statistics of the launch.  0 < rowWidth <= rowPitch, rowPitch divides nInner, nInner < 2^32;
 * rowWidth = rowPitch = 0 means "every inner index is a column" (the plain _dev forms).  A reader chain over the reference's
 * Interpolator -> VerticalInterpolator (src/CDMVerticalInterpolator.cc:356-565 pulling from src/CDMInterpolator.cc:219-263)
 * passes rowWidth = outX and rowPitch = out_row_pitch of the horizontal cache. */
int fimexcu_vertical_interpolate_rows_dev_f(int method, const float* d_in, const float* d_vcoord_in, size_t nOuter, size_t nzi, size_t nInner,
                                            const double* levels_host, const float* d_vcoord_out, size_t nzo, const double* d_vmin,
                                            const double* d_vmax, double valid_min, double valid_max, float badIn, float fillOut, float* d_out,
                                            size_t rowWidth, size_t rowPitch, void* stream);
int fimexcu_vertical_hybrid_pressure_rows_dev_f(int method, const float* d_in, size_t nOuter, size_t nzi, size_t nInner, const double* ap_host,
                                                const double* b_host, const float* d_ps, size_t nOuterPs, double ps_scale,
                                                const double* levels_host, size_t nzo, double valid_min, double valid_max, float badIn,
                                                float fillOut, float* d_out, size_t rowWidth, size_t rowPitch, void* stream);
/* Drop-ins with the reference signatures for the two-level kernels (include/fimex/interpolation.h:153-232,
 * src/interpolation.c:416-546); HOST pointers, n values per level. */
int fimexcu_get_values_nearest_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x);
int fimexcu_get_values_linear_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x);
int fimexcu_get_values_linear_weak_extrapol_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x);
int fimexcu_get_values_linear_no_extrapol_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x);
int fimexcu_get_values_linear_const_extrapol_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x);
int fimexcu_get_values_log_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x);
int fimexcu_get_values_log_log_f(const float* infieldA, const float* infieldB, float* outfield, size_t n, double a, double b, double x);
/* The same seven functions on DEVICE-resident slabs, selected by `method` (FIMEXCU_VINT_METHOD_*), enqueued on `stream`: what
 * CDMTimeInterpolator::getDataSlice does between two time slices (src/CDMTimeInterpolator.cc:127:
 * mifi_get_values_linear_weak_extrapol_f(d1, d2, out, size, t1, t2, t)) for slabs that stay on the device.
 * d_out may alias d_A or d_B. */
int fimexcu_get_values_two_level_dev_f(int method, const float* d_A, const float* d_B, float* d_out, size_t n, double a, double b, double x,
                                       void* stream);

/* ---- fill2d pre / post-process -----------------------------------------------------------------------------
 * Drop-in for mifi_fill2d_f (include/fimex/interpolation.h:253, src/interpolation.c:548-678), the `fill2d(relaxCrit, corrEff,
 * maxLoop)` process that CDMInterpolator::getDataSlice runs level by level before / after the interpolation
 * (src/CDMInterpolator.cc:76-80, 163-186): undefined cells are filled by Laplace relaxation.  Bit-exact: the sequential
 * Gauss-Seidel sweeps run as a wavefront over anti-diagonals with several sweeps in flight (same dependency graph).
 * HOST field [ny][nx] in place; *nChanged = number of undefined cells found. */
int fimexcu_fill2d_f(size_t nx, size_t ny, float* field, float relaxCrit, float corrEff, size_t maxLoop, size_t* nChanged);
/* the same for nz DEVICE-resident levels [nz][ny][nx] at once (processArray_'s loop over levels); nChanged_host: [nz] or NULL */
int fimexcu_fill2d_dev_f(size_t nx, size_t ny, size_t nz, float* d_field, float relaxCrit, float corrEff, size_t maxLoop, size_t* nChanged_host,
                         void* stream);

/* Drop-ins for mifi_creepfill2d_f / mifi_creepfillval2d_f (include/fimex/interpolation.h:275, 298, src/interpolation.c:680-810),
 * the `creepfill2d(repeat, setWeight)` / `creepfillval2d(repeat, setWeight, defaultValue)` processes
 * (src/CDMInterpolator.cc:82-92).  Bit-exact, same wavefront form as fill2d.  HOST field [ny][nx] in place. */
int fimexcu_creepfill2d_f(size_t nx, size_t ny, float* field, unsigned short repeat, char setWeight, size_t* nChanged);
int fimexcu_creepfillval2d_f(size_t nx, size_t ny, float* field, float defaultVal, unsigned short repeat, char setWeight, size_t* nChanged);
/* nz DEVICE-resident levels at once; use_mean != 0: creepfill2d (first guess = the level's mean), else creepfillval2d(defaultVal) */
int fimexcu_creepfill2d_dev_f(size_t nx, size_t ny, size_t nz, float* d_field, int use_mean, float defaultVal, unsigned short repeat, char setWeight,
                              size_t* nChanged_host, void* stream);

/* ---- vertical coordinate converters -----------------------------------------------------------------------
 * The fields CDMVerticalInterpolator asks its ToVLevelConverter / VerticalConverter chain for (src/CDMVerticalInterpolator.cc:
 * 380-470), produced on the device so that they feed fimexcu_vertical_interpolate_dev_f as d_vcoord_in / d_vcoord_out without
 * leaving HBM.  The reference's converters hand doubles to one another and the interpolator narrows the last array with
 * Data::asFloat(); here a chain  level formula -> standard-atmosphere step -> +- topography  is ONE pass in f64 with one f32 store.
 * Fields: f32 [nOuter][nz][nInner]; surface fields f32 [nOuterS][nInner], nOuterS = 1 or nOuter; topography f32 [nInner]. */
enum fimexcu_vcoord_kind {
    FIMEXCU_VCOORD_HYBRID_SIGMA_AP = 0, /* p = ap + b*ps          mifi_atmosphere_hybrid_sigma_ap_pressure (vertical_coordinate_transformations.c:64-70), HybridSigmaApToPressureConverter.cc:71-100 */
    FIMEXCU_VCOORD_HYBRID_SIGMA = 1,    /* p = a*p0 + b*ps        mifi_atmosphere_hybrid_sigma_pressure (:56-62), HybridSigmaToPressureConverter.cc:65-90; scalar = p0 */
    FIMEXCU_VCOORD_SIGMA = 2,           /* p = ptop + sigma*(ps - ptop)  mifi_atmosphere_sigma_pressure (:37-44), SigmaToPressureConverter.cc:50-72; a = sigma, scalar = ptop */
    FIMEXCU_VCOORD_LN_PRESSURE = 3,     /* p = p0*exp(-lev)       mifi_atmosphere_ln_pressure (:29-35), LnPressureToPressureConverter.cc:50-72; a = lev, scalar = p0, no surface field */
    FIMEXCU_VCOORD_HYBRID_HEIGHT = 4,   /* h = a + b*orog         mifi_atmosphere_hybrid_height (:72-78) */
    FIMEXCU_VCOORD_OCEAN_S_G1 = 5,      /* depth = -z, mifi_ocean_s_g1_z (:166-174), OceanSCoordinateGToDepthConverter.cc:64-107; a = s, b = C, scalar = depth_c, surf = depth, eta optional */
    FIMEXCU_VCOORD_OCEAN_S_G2 = 6       /* depth = -z, mifi_ocean_s_g2_z (:175-183) */
};
enum fimexcu_vcoord_post {
    FIMEXCU_VPOST_NONE = 0,
    FIMEXCU_VPOST_STANDARD_ALTITUDE = 1, /* pressure [hPa] -> altitude: mifi_barometric_standard_altitude (:110-113), PressureToStandardAltitudeConverter.cc:62-71 */
    FIMEXCU_VPOST_STANDARD_PRESSURE = 2  /* altitude [m] -> pressure: mifi_barometric_standard_pressure (:95-99), AltitudeStandardToPressureConverter.cc:62-71 */
};
/* a_host, b_host: the per-level terms [nz] (HOST); d_surf is multiplied by surf_scale in f64 (the linear unit conversion of
 * getScaledDataSliceInUnit, e.g. Pa -> hPa); d_topo / topo_factor: AltitudeHeightConverter::getDataSlice
 * (AltitudeHeightConverter.cc:110-131): value + topo_factor*topo with topo_factor = +-1 (surface altitude) or +-1/g (geopotential);
 * NULL = no such step. */
int fimexcu_vcoord_levels_dev_f(int kind, const double* a_host, const double* b_host, double scalar, const float* d_surf, size_t nOuterS,
                                double surf_scale, const float* d_eta, int post, const float* d_topo, double topo_factor, size_t nOuter, size_t nz,
                                size_t nInner, float* d_out, void* stream);
/* the standard-atmosphere step alone on a device field (when the converter's input is a data variable) */
int fimexcu_vcoord_transform_dev_f(int post, const float* d_in, size_t n, float* d_out, void* stream);
/* Replaces PressureIntegrationToAltitudeConverter::getDataSlice (PressureIntegrationToAltitudeConverter.cc:114-247): hypsometric
 * integration from the surface upwards, a += mifi_barometric_layer_thickness(p_low, p_high, Tv) with
 * Tv = mifi_virtual_temperature(q, T) when d_shum is given (vertical_coordinate_transformations.c:115-118, 161-164).
 * d_p [hPa], d_airt [K], d_shum [1] are [nOuter][nz][nInner]; d_sap*sap_scale [hPa] and d_sgp*sgp_scale [m^2/s^2] are surface
 * fields.  start_high_p: 1 = level 0 is the lowest level, 0 = the last level is, -1 = decide like the reference from the first
 * column whose first and last pressure are positive (error if there is none).  d_topo / topo_factor as above. */
int fimexcu_vcoord_pressure_integration_dev_f(const float* d_p, const float* d_sap, double sap_scale, const float* d_sgp, double sgp_scale,
                                              const float* d_airt, const float* d_shum, size_t nOuter, size_t nOuterS, size_t nz, size_t nInner,
                                              int start_high_p, const float* d_topo, double topo_factor, float* d_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FIMEXCU_H */
