/*
 * miniproj.h -- CPU restatement of the cartographic projections fimex can emit, standing in for the
 * third-party PROJ library (TEST INFRASTRUCTURE ONLY; see fimex_oracle.h).
 *
 * PROJ is NOT vendored in /root/reference and not pinned (CMakeLists.txt:56-88 accepts >= 4.6 < 5 or
 * >= 6.2); it is absent from this image.  The reference reaches it only through src/reproject.cc:56-116
 * (proj_create_crs_to_crs, proj_trans_generic, proj_degree_input/output).  The formulas below follow the
 * published algorithms PROJ implements (Snyder, "Map Projections - A Working Manual", USGS PP 1395; PROJ's
 * documented lcc / stere / merc / ob_tran / tmerc(Poder-Engsager) / laea operations; miniproj_more.inc adds sinu, cea,
 * aea, aeqd, ortho, nsper, geos and omerc).
 *
 * PINNED against real PROJ output for ob_tran, lcc (sphere) and polar stere (sphere): the reference's NetCDF
 * fixtures store PROJ-generated longitude/latitude fields for six projected grids; tests/test_proj_pins.py
 * reproduces them to 3e-14 degree in both directions (fixture tests/golden/reference_proj_pins.npz).
 * Every projection, on the sphere and on an ellipsoid, is also pinned against the published worked examples of
 * Snyder's manual (Appendix A; 21 examples, forward within 0.05 m and back: tests/test_more_projections.py).
 * PARITY UNPINNED only below those examples' 0.1 m for the ellipsoidal forms, and for geos / omerc (limiting cases
 * and an independent formulation instead): the reference has no test that fixes reproject_axes output numerically
 * (test/testInterpolation.cc:260-272 only checks "> 89").  Further pins: round trips fwd(inv(x)) <= 1e-9 deg /
 * 1e-6 m, closed-form corner coordinates of the MEPS grid, the EMEP cell test of test/testInterpolation.cc:274-385
 * and the 90/180 degree rotation tests (:387-628), all replayed in tests/test_oracle_proj.py.
 *
 * Transform semantics (crs -> crs): source inverse projection to geographic lon/lat on the source
 * figure, NO datum shift (what proj4's pj_transform does for an unknown datum and PROJ >= 6 does as a
 * "ballpark" transformation for the +R sphere strings fimex writes, ProjectionImpl.cc:157-180), destination
 * forward projection.  Geographic CRS ("longlat", and "ob_tran +o_proj=longlat") take and give degrees,
 * as PROJ >= 6 proj_create_crs_to_crs does.  A point that cannot be projected becomes +inf (HUGE_VAL).
 */
#ifndef MINIPROJ_H
#define MINIPROJ_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { MP_LONGLAT = 0, MP_LCC, MP_STERE, MP_OB_TRAN, MP_MERC, MP_TMERC, MP_LAEA, MP_SINU, MP_CEA, MP_AEA, MP_AEQD, MP_ORTHO, MP_NSPER, MP_GEOS, MP_OMERC };

typedef struct {
    int kind;
    int geographic; /* I/O in degrees */
    int ellips;     /* es != 0 */
    int mode;       /* projection specific */
    int over;
    double a, ra, es, e, one_es, rone_es;
    double lam0, phi0, k0, x0, y0, to_meter, fr_meter;
    double q[40]; /* projection constants, see miniproj.c */
} miniproj_crs;

/* parse a proj4 string ("+proj=lcc +lat_0=63.3 ... +R=6.371e+06"); 0 on success, else message in err */
int miniproj_parse(const char* def, miniproj_crs* crs, char* err, size_t errlen);

/* geographic radians (lam, phi) -> CRS coordinates (metres, or degrees when geographic); 0 ok */
int miniproj_fwd(const miniproj_crs* crs, double lam, double phi, double* x, double* y);
/* CRS coordinates -> geographic radians */
int miniproj_inv(const miniproj_crs* crs, double x, double y, double* lam, double* phi);

/* crs -> crs on n points in place; failed points become +inf.  Returns number of failed points. */
size_t miniproj_transform(const miniproj_crs* src, const miniproj_crs* dst, size_t n, double* x, double* y);

/* convenience: parse both and transform; -1 on parse error */
long miniproj_transform_str(const char* src, const char* dst, size_t n, double* x, double* y, char* err, size_t errlen);

/* ---- restatements of src/reproject.cc on top of the above ---- */
/* :573-587 */
int orc_reproject_axes(const char* proj_in, const char* proj_out, const double* in_x, const double* in_y, int ix, int iy, double* out_x,
                       double* out_y);
/* :463-486 with :363-397 and :310-357 (PROJ >= 6 semantics: "target is latlon" == geographic output) */
int orc_get_vector_reproject_matrix(const char* proj_in, const char* proj_out, const double* out_x_axis, const double* out_y_axis, int ox, int oy,
                                    double* matrix3);
/* :432-447 */
int orc_get_vector_reproject_matrix_points(const char* proj_in, const char* proj_out, int inputIsMetric, const double* out_x_points,
                                           const double* out_y_points, int ox, int oy, double* matrix3);

#ifdef __cplusplus
}
#endif
#endif
