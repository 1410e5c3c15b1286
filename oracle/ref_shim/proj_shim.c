/* proj_shim.c -- the proj.h stand-in of this directory implemented on oracle/miniproj.c.
 * TEST INFRASTRUCTURE ONLY. */
#include "proj.h"

#include "miniproj.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

struct mp_shim_ctx {
    int last_errno;
};
struct mp_shim_pj {
    miniproj_crs src, dst;
    int src_latlon_first; /* "EPSG:4326": axis order lat, lon */
    int err;
};

static struct mp_shim_ctx g_ctx;

PJ_CONTEXT* proj_context_create(void) { return &g_ctx; }
int proj_context_errno(PJ_CONTEXT* ctx) { return ctx ? ctx->last_errno : 0; }
const char* proj_errno_string(int err) { return err ? "miniproj: unsupported or invalid projection definition" : "no error"; }
int proj_errno(const PJ* P) { return P ? P->err : 0; }

PJ* proj_create_crs_to_crs(PJ_CONTEXT* ctx, const char* source_crs, const char* target_crs, PJ_AREA* area)
{
    (void)area;
    PJ* P = (PJ*)calloc(1, sizeof(PJ));
    char err[128];
    if (ctx)
        ctx->last_errno = 0;
    P->src_latlon_first = strncmp(source_crs, "EPSG:4326", 9) == 0;
    if (miniproj_parse(source_crs, &P->src, err, sizeof(err)) || miniproj_parse(target_crs, &P->dst, err, sizeof(err))) {
        if (ctx)
            ctx->last_errno = 1;
        free(P);
        return NULL;
    }
    return P;
}

PJ* proj_destroy(PJ* P)
{
    free(P);
    return NULL;
}

int proj_degree_input(PJ* P, PJ_DIRECTION dir) { return dir == PJ_FWD ? P->src.geographic : P->dst.geographic; }
int proj_degree_output(PJ* P, PJ_DIRECTION dir) { return dir == PJ_FWD ? P->dst.geographic : P->src.geographic; }

size_t proj_trans_generic(PJ* P, PJ_DIRECTION direction, double* x, size_t sx, size_t nx, double* y, size_t sy, size_t ny, double* z, size_t sz,
                          size_t nz, double* t, size_t st, size_t nt)
{
    (void)z; (void)sz; (void)nz; (void)t; (void)st; (void)nt;
    const size_t n = nx < ny ? nx : ny;
    const miniproj_crs* a = direction == PJ_FWD ? &P->src : &P->dst;
    const miniproj_crs* b = direction == PJ_FWD ? &P->dst : &P->src;
    for (size_t i = 0; i < n; ++i) {
        double* px = (double*)((char*)x + i * sx);
        double* py = (double*)((char*)y + i * sy);
        if (miniproj_transform(a, b, 1, px, py))
            P->err = 2;
    }
    return n; /* like PROJ: failed points carry HUGE_VAL, the count is the number processed */
}

PJ_COORD proj_coord(double x, double y, double z, double t)
{
    PJ_COORD c;
    c.v[0] = x; c.v[1] = y; c.v[2] = z; c.v[3] = t;
    return c;
}

PJ_COORD proj_trans(PJ* P, PJ_DIRECTION direction, PJ_COORD c)
{
    double x = c.v[0], y = c.v[1];
    if (direction == PJ_FWD && P->src_latlon_first) {
        x = c.v[1];
        y = c.v[0];
    }
    const miniproj_crs* a = direction == PJ_FWD ? &P->src : &P->dst;
    const miniproj_crs* b = direction == PJ_FWD ? &P->dst : &P->src;
    miniproj_transform(a, b, 1, &x, &y);
    if (direction != PJ_FWD && P->src_latlon_first) {
        const double tmp = x;
        x = y;
        y = tmp;
    }
    c.v[0] = x;
    c.v[1] = y;
    return c;
}
