/*
 * fimex_oracle.h -- CPU restatement of the fimex regridding hot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this library.  The product
 * (fimex_b200/libfimexcu.so) never links, loads or calls anything in this directory.
 *
 * Every function cites the reference (fimex 2.4.1, paths relative to /root/reference) it
 * restates.  The restatement is pinned in tests/test_oracle_*.py against
 *   (1) the golden vectors of the reference's own unit tests, and
 *   (2) the reference itself, compiled from its own sources into oracle/_ref/libfimexref.so
 *       (oracle/Makefile), on seeded random inputs.
 * Projection arithmetic (oracle/miniproj.c) restates the published PROJ / Snyder formulas; PROJ
 * itself is not vendored in the reference; that part is pinned against the PROJ-generated
 * longitude/latitude fields of the reference's fixtures for ob_tran / lcc / stere on the sphere
 * (tests/test_proj_pins.py) and by analytic identities for the rest (see miniproj.h).
 *
 * Arithmetic is written so that gcc on x86-64 (SSE2, no FMA contraction: build with
 * -ffp-contract=off) produces the same bits as the reference built with its default flags.
 */
#ifndef FIMEX_ORACLE_H
#define FIMEX_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* mifi_constants.h:52-138 */
enum {
    ORC_INTERPOL_NEAREST_NEIGHBOR = 0,
    ORC_INTERPOL_BILINEAR = 1,
    ORC_INTERPOL_BICUBIC = 2,
    ORC_INTERPOL_COORD_NN = 3,
    ORC_INTERPOL_COORD_NN_KD = 4,
    ORC_INTERPOL_FORWARD_SUM = 5,
    ORC_INTERPOL_FORWARD_MEAN = 6,
    ORC_INTERPOL_FORWARD_MEDIAN = 7,
    ORC_INTERPOL_FORWARD_MAX = 8,
    ORC_INTERPOL_FORWARD_MIN = 9,
    ORC_INTERPOL_FORWARD_UNDEF_SUM = 10,
    ORC_INTERPOL_FORWARD_UNDEF_MEAN = 11,
    ORC_INTERPOL_FORWARD_UNDEF_MEDIAN = 12,
    ORC_INTERPOL_FORWARD_UNDEF_MAX = 13,
    ORC_INTERPOL_FORWARD_UNDEF_MIN = 14
};
/* mifi_constants.h:206-231 */
enum {
    ORC_VINT_METHOD_LIN = 0,
    ORC_VINT_METHOD_LOG = 1,
    ORC_VINT_METHOD_LOGLOG = 2,
    ORC_VINT_METHOD_NN = 3,
    ORC_VINT_METHOD_LIN_WEAK_EXTRA = 4,
    ORC_VINT_METHOD_LIN_NO_EXTRA = 5,
    ORC_VINT_METHOD_LIN_CONST_EXTRA = 6
};
/* mifi_constants.h:264-268 */
enum { ORC_PROJ_AXIS = 0, ORC_LONGITUDE = 1, ORC_LATITUDE = 2 };

#define ORC_OK 1
#define ORC_ERROR (-1)
#define ORC_INVALID_IDX ((size_t)0xFFFFFFFFu) /* "~0u" widened to size_t, CachedInterpolation.cc:257 */

/* ---- src/interpolation.c ---- */
int orc_string_to_interpolation_method(const char* name);                                     /* :72-107 */
int orc_bsearch_double_index(double key, const double* base, int num, int ascending);         /* :129-152 */
int orc_points2position(double* points, size_t n, const double* axis, int num, int axis_type); /* :154-227 */
int orc_get_values_f(const float* in, float* out, double x, double y, int ix, int iy, int iz);          /* :251-265 */
int orc_get_values_bilinear_f(const float* in, float* out, double x, double y, int ix, int iy, int iz); /* :267-343 */
int orc_get_values_bicubic_f(const float* in, float* out, double x, double y, int ix, int iy, int iz);  /* :345-414 */
int orc_get_values_nearest_f(const float* a, const float* b, float* out, size_t n, double xa, double xb, double x);               /* :416-425 */
int orc_get_values_linear_f(const float* a, const float* b, float* out, size_t n, double xa, double xb, double x);                /* :439-452 */
int orc_get_values_linear_weak_extrapol_f(const float* a, const float* b, float* out, size_t n, double xa, double xb, double x);  /* :496-499 */
int orc_get_values_linear_no_extrapol_f(const float* a, const float* b, float* out, size_t n, double xa, double xb, double x);    /* :500-503 */
int orc_get_values_linear_const_extrapol_f(const float* a, const float* b, float* out, size_t n, double xa, double xb, double x); /* :505-516 */
int orc_get_values_log_f(const float* a, const float* b, float* out, size_t n, double xa, double xb, double x);                   /* :521-532 */
int orc_get_values_log_log_f(const float* a, const float* b, float* out, size_t n, double xa, double xb, double x);               /* :534-546 */
void orc_bad2nanf(float* v, size_t n, float bad); /* :1064-1072 */
void orc_nanf2bad(float* v, size_t n, float bad); /* :1074-1082 */
/* packed CDM_SHORT slabs around the interpolation, src/CDMInterpolator.cc:140-151:
 *   orc_i16_to_interpolation_array  Data::asFloat() (data_caster<float,short>, include/fimex/MathUtils.h:60-96) + mifi_bad2nanf
 *   orc_interpolation_array_to_i16  convertDataType(NaN,1,0,CDM_SHORT,badValue,1,0) -> ScaleValue<float,short>
 *                                   (include/fimex/DataUtils.h:38-60, src/DataImpl.h:296-307): NaN -> (short)badValue, else
 *                                   (short)(int)lround(1.0*v + 0.0) */
void orc_i16_to_interpolation_array(const short* in, size_t n, double badValue, float* out);
void orc_interpolation_array_to_i16(const float* in, size_t n, double badValue, short* out);
/* getScaledDataSlice of a CDM_SHORT variable as the vertical converters see it (src/CDMReader.cc:177-192 ->
 * ScaleValue<short,double>(fill, scale, offset, NaN, 1, 0), then Data::asFloat()) */
void orc_i16_scaled_to_float(const short* in, size_t n, double fill, double scale, double offset, float* out);

/* ---- src/MathUtils.cc:30-46 : RoundAndClamp ---- */
size_t orc_round_and_clamp(double d, size_t mini, size_t maxi, size_t invalid);

/* ---- src/CachedInterpolation.cc : CachedInterpolation (bilinear/bicubic) ---- */
typedef struct {
    int method;
    size_t inX, inY, outX, outY; /* inX/inY after domain reduction */
    int reduced;
    size_t xMin, yMin;
    double* px; /* owned copies, rebased when reduced */
    double* py;
} orc_cached_interp;
orc_cached_interp* orc_cached_interp_create(int method, const double* px, const double* py, size_t inX, size_t inY, size_t outX,
                                            size_t outY);                                  /* :149-171, :222-253 */
int orc_cached_interp_apply(const orc_cached_interp* c, const float* in, size_t size, float* out); /* :173-209 */
void orc_cached_interp_destroy(orc_cached_interp* c);

/* ---- src/CachedInterpolation.cc : CachedNNInterpolation ---- */
typedef struct {
    size_t inX, inY, outX, outY;
    int reduced;
    size_t xMin, yMin;
    size_t* pointsInIn; /* outX*outY entries, ORC_INVALID_IDX when no source */
} orc_cached_nn;
orc_cached_nn* orc_cached_nn_create(const double* px, const double* py, size_t inX, size_t inY, size_t outX, size_t outY); /* :262-315 */
int orc_cached_nn_apply(const orc_cached_nn* c, const float* in, size_t size, float* out);                                 /* :317-337 */
void orc_cached_nn_destroy(orc_cached_nn* c);

/* ---- src/CachedForwardInterpolation.cc ---- */
typedef struct {
    int method;
    int undefAggr;
    size_t inX, inY, outX, outY;
    int reduced;
    size_t xMin, yMin;
    size_t maxPointsInIn;
    size_t* segStart; /* CSR: outX*outY+1 */
    size_t* segIdx;   /* source indices, ascending inside a segment */
} orc_cached_fwd;
orc_cached_fwd* orc_cached_fwd_create(int method, const double* pOnX, const double* pOnY, size_t inX, size_t inY, size_t outX,
                                      size_t outY); /* :133-201 */
/* level_bug != 0 reproduces the reference literally: inData[ii] without the z offset (:212-221),
 * i.e. every output level aggregates input level 0.  level_bug == 0 uses inData[z*inLayer+ii]. */
int orc_cached_fwd_apply(const orc_cached_fwd* c, const float* in, size_t size, float* out, int level_bug); /* :205-224 */
void orc_cached_fwd_destroy(orc_cached_fwd* c);

/* ---- src/reproject.cc ---- */
void orc_vector_reproject_values_by_matrix_f(const double* matrix3, size_t layer, float* u, float* v, size_t nz); /* :488-507 */
void orc_vector_reproject_direction_by_matrix_f(const double* matrix3, size_t layer, float* angle, size_t nz);    /* :509-530 */

/* ---- include/fimex/FindNeighborElements.h (value_type = float, as instantiated by
 *      CDMVerticalInterpolator.cc:531-533 with leap_iterator<const float*>) ---- */
void orc_find_closest_distinct_elements(const float* col, size_t n, ptrdiff_t stride, float x, size_t* first, size_t* second);          /* :46-77 */
void orc_find_closest_neighbor_distinct_elements(const float* col, size_t n, ptrdiff_t stride, float x, size_t* first, size_t* second); /* :91-125 */

/* ---- src/vertical_coordinate_transformations.c:29-77 ---- */
void orc_atmosphere_ln_pressure(size_t n, double p0, const double* lev, double* p);
void orc_atmosphere_sigma_pressure(size_t n, double ptop, double ps, const double* sigma, double* p);
void orc_atmosphere_hybrid_sigma_pressure(size_t n, double p0, double ps, const double* a, const double* b, double* p);
void orc_atmosphere_hybrid_sigma_ap_pressure(size_t n, double ps, const double* ap, const double* b, double* p);
void orc_atmosphere_hybrid_height(size_t n, double orog, const double* a, const double* b, double* h);

/* ---- src/CDMVerticalInterpolator.cc:488-561, restated for the dense layouts the CDM produces:
 *   in      [nzi][ncol]   f32 (NaN = undefined), vcoord_in [nzi][ncol] f32,
 *   out     [nzo][ncol]   f32,
 *   vertical target: levels[nzo] f64 (fixed levels)  or  vcoord_out [nzo][ncol] f32 (template / to-axis).
 *   vmin/vmax: per-column f64 validity range (NULL = none).  valid_min/valid_max: NaN = none.
 * Where the reference leaves the output unwritten (log of a non-positive coordinate, interpolation.c:524,
 * result ignored at CDMVerticalInterpolator.cc:539) this restatement writes NaN. ---- */
int orc_vertical_interpolate_f(int method, const float* in, const float* vcoord_in, size_t nzi, size_t ncol, const double* levels,
                               const float* vcoord_out, size_t nzo, const double* vmin, const double* vmax, double valid_min,
                               double valid_max, float* out);

/* ---- src/CDMInterpolator.cc:953-1028 (coord_kdtree) stated without nanoflann: for each target the source point with the
 * smallest squared chord distance on the unit sphere, strictly below (maxDist/R)^2; misses give -1000.  Brute force, O(N*M):
 * small cases only.  Exact-distance ties go to the lower source index here; nanoflann's own choice (tree order) is available
 * from oracle/_ref (ref_kdtree_translate).  second_best_gap[i] (may be NULL) receives d2(second best) - d2(best), so that tests
 * can tell genuine ties / near-ties from disagreements. ---- */
int orc_kdtree_translate(double maxDist_m, double* pointsOnX_deg, double* pointsOnY_deg, size_t nPoints, const double* lon_deg,
                         const double* lat_deg, size_t nx, size_t ny, double* second_best_gap);
/* ---- src/CDMInterpolator.cc:1030-1190 (coord_nearestneighbor): getGridDistance + fastTranslatePointsToClosestInputCell,
 * restated literally (latitude-sorted list, scan up then down).  Misses give -1. ---- */
double orc_get_grid_distance(const double* lon_deg, const double* lat_deg, size_t nx, size_t ny);
int orc_coordnn_translate(double* pointsOnX_deg, double* pointsOnY_deg, size_t nPoints, const double* lon_deg, const double* lat_deg, size_t nx,
                          size_t ny);

#ifdef __cplusplus
}
#endif
#endif

/* ---- vertical coordinate converters -------------------------------------------------------------------
 * Restated from src/vertical_coordinate_transformations.c:29-183 and the converter classes under
 * src/coordSys/verticalTransform/ (HybridSigmaApToPressureConverter.cc:71-100, HybridSigmaToPressureConverter.cc:65-90,
 * SigmaToPressureConverter.cc:50-72, LnPressureToPressureConverter.cc:50-72, AtmosphereHybridHeight / mifi_atmosphere_hybrid_height,
 * OceanSCoordinateGToDepthConverter.cc:64-107, PressureToStandardAltitudeConverter.cc:62-71,
 * AltitudeStandardToPressureConverter.cc:62-71, AltitudeHeightConverter.cc:110-131,
 * PressureIntegrationToAltitudeConverter.cc:114-247).  The converters hand DOUBLES to one another; the vertical interpolator
 * narrows the final field with Data::asFloat() (CDMVerticalInterpolator.cc:483).  Fields: [nOuter][nz][nInner], surface fields
 * [nOuterS][nInner] with nOuterS = 1 or nOuter. */
enum { ORC_VCOORD_HYBRID_SIGMA_AP = 0, ORC_VCOORD_HYBRID_SIGMA = 1, ORC_VCOORD_SIGMA = 2, ORC_VCOORD_LN_PRESSURE = 3,
       ORC_VCOORD_HYBRID_HEIGHT = 4, ORC_VCOORD_OCEAN_S_G1 = 5, ORC_VCOORD_OCEAN_S_G2 = 6 };
enum { ORC_VPOST_NONE = 0, ORC_VPOST_STANDARD_ALTITUDE = 1, ORC_VPOST_STANDARD_PRESSURE = 2 };
/* a, b: per-level terms (ap|a|sigma|lev|a|s and b|b|-|-|b|C); scalar: p0 | ptop | depth_c; surf: ps | orog | depth (scaled by
 * surf_scale in double, like getScaledDataSliceInUnit); eta: ocean zeta or NULL; post: standard-atmosphere step on the double;
 * topo: AltitudeHeightConverter's term (NULL = none), added as topo_factor * topo in double */
int orc_vcoord_levels_f(int kind, const double* a, const double* b, double scalar, const float* surf, size_t nOuterS, double surf_scale,
                        const float* eta, int post, const float* topo, double topo_factor, size_t nOuter, size_t nz, size_t nInner, float* out);
/* hypsometric integration from the surface upwards; start_high_p: 1 level 0 is the lowest, 0 the last level is, -1 decide
 * like the reference from the first column whose first and last pressure are positive (returns 0 if there is none) */
int orc_vcoord_pressure_integration_f(const float* p, const float* sap, double sap_scale, const float* sgp, double sgp_scale, const float* airt,
                                      const float* shum, size_t nOuter, size_t nOuterS, size_t nz, size_t nInner, int start_high_p,
                                      const float* topo, double topo_factor, float* out);
/* stand-alone standard-atmosphere step on a field (f32 in, double arithmetic, f32 out) */
void orc_vcoord_transform_f(int post, const float* in, size_t n, float* out);
float orc_virtual_temperature(float spec_humidity, float T);                       /* vertical_coordinate_transformations.c:115-118 */
float orc_barometric_layer_thickness(float p_low_alti, float p_high_alti, float T); /* :161-164 */

/* ---- packed integer slabs, any CDM integer type (see orc_i16_* above for the citations); dtype as enum fimexcu_dtype:
 * 0 CDM_CHAR (signed char), 1 CDM_SHORT, 2 CDM_INT, 3 CDM_UCHAR, 4 CDM_USHORT, 5 CDM_UINT */
int orc_unpack_f(int dtype, const void* in, size_t n, double badValue, float* out);
int orc_unpack_scaled_f(int dtype, const void* in, size_t n, double fill, double scale, double offset, float* out);
int orc_pack_f(int dtype, const float* in, size_t n, double badValue, void* out);

/* mifi_fill2d_f (src/interpolation.c:548-678): undefined cells replaced by the solution of a Laplace relaxation (sequential
 * Gauss-Seidel sweeps in row-major order with over-relaxation corrEff, borders extrapolated after every sweep, convergence
 * tested after sweeps 0, 10, 20, ... below maxLoop-5).  In place; *nChanged = number of undefined cells found. */
int orc_fill2d_f(size_t nx, size_t ny, float* field, float relaxCrit, float corrEff, size_t maxLoop, size_t* nChanged);

/* mifi_creepfill2d_f / mifi_creepfillval2d_f (src/interpolation.c:680-810): undefined cells take the weighted mean of their
 * defined 4-neighbours and their own first guess (field mean resp. defaultVal), sweep after sweep in row-major order (a cell
 * filled in this sweep already counts for its east and south neighbours), every cell at most `repeat` times, until a sweep
 * changes nothing; borders afterwards from their inner neighbours. */
int orc_creepfill2d_f(size_t nx, size_t ny, float* field, unsigned short repeat, char setWeight, size_t* nChanged);
int orc_creepfillval2d_f(size_t nx, size_t ny, float* field, float defaultVal, unsigned short repeat, char setWeight, size_t* nChanged);
