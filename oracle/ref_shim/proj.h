/*
 * proj.h -- stand-in for the PROJ >= 6 C API, limited to the names src/reproject.cc uses
 * (reproject.cc:36-44, 56-116, 542-548), so that the reference translation unit compiles unmodified
 * in oracle/_ref.  The arithmetic behind it is oracle/miniproj.c (see miniproj.h for what that pins).
 * TEST INFRASTRUCTURE ONLY.  Not a copy of PROJ's header: only the handful of declarations needed.
 */
#ifndef MINIPROJ_PROJ_H_SHIM
#define MINIPROJ_PROJ_H_SHIM

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mp_shim_ctx PJ_CONTEXT;
typedef struct mp_shim_pj PJ;
typedef struct mp_shim_area PJ_AREA;

typedef enum { PJ_FWD = 1, PJ_IDENT = 0, PJ_INV = -1 } PJ_DIRECTION;

typedef struct { double e, n, u; } PJ_ENU;
typedef struct { double x, y, z, t; } PJ_XYZT;
typedef union {
    double v[4];
    PJ_XYZT xyzt;
    PJ_ENU enu;
} PJ_COORD;

PJ_CONTEXT* proj_context_create(void);
int proj_context_errno(PJ_CONTEXT* ctx);
const char* proj_errno_string(int err);
int proj_errno(const PJ* P);

PJ* proj_create_crs_to_crs(PJ_CONTEXT* ctx, const char* source_crs, const char* target_crs, PJ_AREA* area);
PJ* proj_destroy(PJ* P);

int proj_degree_input(PJ* P, PJ_DIRECTION dir);
int proj_degree_output(PJ* P, PJ_DIRECTION dir);

size_t proj_trans_generic(PJ* P, PJ_DIRECTION direction, double* x, size_t sx, size_t nx, double* y, size_t sy, size_t ny, double* z, size_t sz,
                          size_t nz, double* t, size_t st, size_t nt);
PJ_COORD proj_coord(double x, double y, double z, double t);
PJ_COORD proj_trans(PJ* P, PJ_DIRECTION direction, PJ_COORD coord);

#ifdef __cplusplus
}
#endif
#endif
