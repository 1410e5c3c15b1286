"""Generates tests/golden/*.npz from the reference's OWN code (oracle/_ref, compiled from /root/reference by
oracle/Makefile).  Run in the build container:  python tests/golden/make_golden.py
The GPU box has no /root/reference; the committed fixtures carry the reference's answers there."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as ol  # noqa: E402
from oracle_lib import dp, fp  # noqa: E402


def tables(rng, in_x, in_y, n, margin=3.0):
    px = rng.uniform(-margin, in_x - 1 + margin, n)
    py = rng.uniform(-margin, in_y - 1 + margin, n)
    k = n // 8
    px[:k] = np.round(px[:k] * 2) / 2
    py[:k] = np.round(py[:k] * 2) / 2
    px[k:k + 4] = -999.
    py[k + 2:k + 6] = -1000.
    strip = (py >= in_y - 0.5) & (py < in_y + 0.5) & ~((np.floor(px) >= 0) & (np.floor(px) + 1 < in_x))
    py[strip] += 1.0  # the reference's out-of-bounds corner strip (interpolation.c:322)
    return px, py


def main():
    ref = ol.ref()
    assert ref is not None, "build oracle/_ref first (make -C oracle)"
    rng = np.random.default_rng(20261106)
    out = {}
    in_x, in_y, out_x, out_y, nz = 41, 33, 37, 29, 3
    fld = rng.normal(280, 10, size=(nz, in_y, in_x)).astype(np.float32)
    fld[rng.random(fld.shape) < 0.04] = np.nan
    out["h_field"] = fld
    out["h_dims"] = np.array([in_x, in_y, out_x, out_y, nz])
    for tag, scale in (("full", 1.0), ("sub", 0.45)):
        px, py = tables(rng, in_x, in_y, out_x * out_y)
        if scale != 1.0:
            px, py = 9 + px * scale, 7 + py * scale
        out[f"h_{tag}_px"], out[f"h_{tag}_py"] = px, py
        for name, m in (("bilinear", ol.BILINEAR), ("bicubic", ol.BICUBIC), ("nearest", ol.NEAREST)):
            r, info = ol.r_cached_apply(m, px, py, in_x, in_y, out_x, out_y, fld)
            out[f"h_{tag}_{name}_out"] = r
            out[f"h_{tag}_{name}_info"] = np.array([info["inX"], info["inY"], int(info["reduced"]), info["xMin"], info["yMin"]])
    # forward aggregation
    fin_x, fin_y, fout_x, fout_y, fnz = 57, 43, 13, 11, 2
    pox = rng.uniform(-2, fout_x + 1, fin_x * fin_y)
    poy = rng.uniform(-2, fout_y + 1, fin_x * fin_y)
    ffld = rng.normal(size=(fnz, fin_y, fin_x)).astype(np.float32)
    ffld[rng.random(ffld.shape) < 0.1] = np.nan
    out["f_field"], out["f_px"], out["f_py"] = ffld, pox, poy
    out["f_dims"] = np.array([fin_x, fin_y, fout_x, fout_y, fnz])
    for name, m in (("sum", ol.FWD_SUM), ("mean", ol.FWD_MEAN), ("median", ol.FWD_MEDIAN), ("max", ol.FWD_MAX), ("min", ol.FWD_MIN),
                    ("undef_sum", ol.FWD_UNDEF_SUM), ("undef_mean", ol.FWD_UNDEF_MEAN), ("undef_max", ol.FWD_UNDEF_MAX),
                    ("undef_min", ol.FWD_UNDEF_MIN)):
        r, info = ol.r_cached_apply(m, pox, poy, fin_x, fin_y, fout_x, fout_y, ffld)
        out[f"f_{name}_out"] = r
    # vector rotation
    ox, oy, vz = 23, 19, 2
    ang = rng.uniform(-np.pi, np.pi, ox * oy)
    m3 = np.stack([np.cos(ang), np.sin(ang), np.degrees(ang)], axis=1).copy()
    u = rng.normal(size=(vz, oy, ox)).astype(np.float32)
    v = rng.normal(size=(vz, oy, ox)).astype(np.float32)
    u[0, 0, :2] = np.nan
    d = rng.uniform(0, 360, size=(vz, oy, ox)).astype(np.float32)
    out["v_matrix"], out["v_u"], out["v_v"], out["v_dir"] = m3, u.copy(), v.copy(), d.copy()
    ref.ref_vector_reproject_values_by_matrix_f(dp(m3), ox, oy, fp(u), fp(v), vz)
    ref.ref_vector_reproject_direction_by_matrix_f(dp(m3), ox, oy, fp(d), vz)
    out["v_u_out"], out["v_v_out"], out["v_dir_out"] = u, v, d
    # points2position
    axes = {"lcc_x": np.linspace(-1060084., 1309916., 949), "lat_desc": np.linspace(75.25, 49.75, 1021), "lon360": np.arange(0., 360., 0.5),
            "lon180": np.arange(-180., 180., 0.25), "irregular": np.cumsum(rng.uniform(0.5, 2.0, 200))}
    for k, ax in axes.items():
        for t in (0, 1, 2):
            lo, hi = ax.min(), ax.max()
            pts = rng.uniform(lo - 0.3 * (hi - lo), hi + 0.3 * (hi - lo), 600)
            pts[:60] = rng.choice(ax, 60)
            pts[60:70] = [np.nan, np.inf, -np.inf, 360., -360., 180., -180., 0., 720., -0.0]
            res = pts.copy()
            ref.mifi_points2position(dp(res), res.size, dp(ax), ax.size, t)
            out[f"p_{k}_{t}_axis"], out[f"p_{k}_{t}_in"], out[f"p_{k}_{t}_out"] = ax, pts, res
    np.savez_compressed(os.path.join(HERE, "reference_small.npz"), **out)
    print("wrote reference_small.npz with", len(out), "arrays")



def vertical_fixtures():
    """The reference's own expected-output files of its vertical-interpolation shell tests
    (test/testVerticalInterpolatorSigmaHybridToPressure.sh, test/testVerticalInterpolatorEnsembleSigmaHybridToPressure.sh:
    fimex --verticalInterpolate.type=pressure --verticalInterpolate.method=log --verticalInterpolate.level1=1000,850,500,300,100,50),
    NetCDF-3 classic, read with scipy.  Stored: the raw inputs the hot loop sees and the values fimex wrote."""
    from scipy.io import netcdf_file
    d = "/root/reference/test/"
    out = {}
    fi, fo = netcdf_file(d + "verticalSigmaHybrid.nc", mmap=False), netcdf_file(d + "verticalSigmaHybridPressure.nc", mmap=False)
    out["sh_ap_pa"] = fi.variables["ap"].data.astype(np.float64)
    out["sh_b"] = fi.variables["b"].data.astype(np.float64)
    psv = fi.variables["air_pressure"]
    out["sh_ps_raw"] = psv.data.astype(np.float32)                     # packed short, unit hPa after scaling
    out["sh_ps_scale"] = np.float64(psv.scale_factor)
    out["sh_levels"] = fo.variables["pressure"].data.astype(np.float64)
    for n in ("air_potential_temperature", "specific_humidity"):
        out["sh_in_" + n] = fi.variables[n].data.astype(np.float32)    # (time, k, Yc, Xc), _FillValue 9.96921e36
        out["sh_ex_" + n] = fo.variables[n].data.astype(np.float32)    # (time, pressure, Yc, Xc)
    fi, fo = netcdf_file(d + "testdata_vertical_ensemble_in.nc", mmap=False), netcdf_file(d + "testdata_vertical_ensemble_ex.nc", mmap=False)
    out["en_ap_pa"] = fi.variables["ap"].data.astype(np.float64)
    out["en_b"] = fi.variables["b"].data.astype(np.float64)
    out["en_ps_pa"] = fi.variables["surface_air_pressure"].data.astype(np.float32)  # (time, height0, ensemble, y, x)
    out["en_levels"] = fo.variables["pressure"].data.astype(np.float64)
    for n in ("specific_humidity_ml", "x_wind_ml", "y_wind_ml", "upward_air_velocity_ml"):
        out["en_in_" + n] = fi.variables[n].data.astype(np.int32)      # packed int32, _FillValue -2147483647: the loop sees the raw numbers as f32
        out["en_ex_" + n] = fo.variables[n].data.astype(np.int32)      # (time, pressure, ensemble, y, x)
    np.savez_compressed(os.path.join(HERE, "reference_vertical.npz"), **out)
    print("wrote reference_vertical.npz with", len(out), "arrays")



def hirlam12_fixture():
    """BASELINE configs[0]: test/hirlam12.nc (NetCDF-3 classic: time 2, pressure 2, Yc 12, Xc 17; Xc/Yc plain degrees, i.e. a
    lat/lon source) regridded bilinearly as `fimex --interpolate.method=bilinear --interpolate.projString=...` would, by driving
    the reference's own objects: reproject_axes (src/reproject.cc over the proj.h shim) -> mifi_points2position ->
    createCachedInterpolation -> interpolateValues, and for x_wind/y_wind get_vector_reproject_matrix -> reprojectValues."""
    import ctypes as C
    from scipy.io import netcdf_file
    ref = ol.ref()
    f = netcdf_file("/root/reference/test/hirlam12.nc", mmap=False)
    xc, yc = f.variables["Xc"].data.astype(np.float64), f.variables["Yc"].data.astype(np.float64)
    out = {"xc": xc, "yc": yc}
    fields = {n: f.variables[n].data.astype(np.float32) for n in ("geopotential_height", "air_potential_temperature", "x_wind", "y_wind")}
    for n, v in fields.items():
        out["in_" + n] = v  # (time, pressure, Yc, Xc)
    src = "+proj=latlong +R=6371000 +no_defs"
    targets = {
        "latlong": ("+proj=latlong +R=6371000 +no_defs", np.arange(4.9, 6.7001, 0.05), np.arange(61.4, 62.7001, 0.05), 1, 2),
        "stere": ("+proj=stere +lat_0=90 +lon_0=0 +lat_ts=60 +R=6371000 +no_defs", 295000.0 + 2500.0 * np.arange(50), -3100000.0 + 2500.0 * np.arange(60), 0, 0),
    }
    for tag, (proj, ax, ay, xt, yt) in targets.items():
        ax, ay = np.ascontiguousarray(ax), np.ascontiguousarray(ay)
        n = ax.size * ay.size
        px, py = np.empty(n), np.empty(n)
        assert ref.ref_reproject_axes(proj.encode(), src.encode(), dp(ax), dp(ay), ax.size, ay.size, dp(px), dp(py)) == 0
        ref.mifi_points2position(dp(px), n, dp(xc), xc.size, 1)  # MIFI_LONGITUDE
        ref.mifi_points2position(dp(py), n, dp(yc), yc.size, 2)  # MIFI_LATITUDE
        out[f"{tag}_proj"] = np.frombuffer(proj.encode(), dtype=np.uint8)
        out[f"{tag}_ax"], out[f"{tag}_ay"], out[f"{tag}_px"], out[f"{tag}_py"] = ax, ay, px, py
        m = np.empty(n * 3)
        assert ref.ref_get_vector_reproject_matrix(src.encode(), proj.encode(), dp(ax), dp(ay), ax.size, ay.size, dp(m)) == 0
        out[f"{tag}_matrix"] = m
        res = {}
        for name, v in fields.items():
            fld = v.reshape(-1, yc.size, xc.size)
            r, info = ol.r_cached_apply(ol.BILINEAR, px, py, xc.size, yc.size, ax.size, ay.size, fld)
            res[name] = r
            out[f"{tag}_info"] = np.array([info["inX"], info["inY"], int(info["reduced"]), info["xMin"], info["yMin"]])
        u, v = res["x_wind"].copy(), res["y_wind"].copy()
        ref.ref_vector_reproject_values_by_matrix_f(dp(m), ax.size, ay.size, fp(u), fp(v), u.shape[0])
        res["x_wind"], res["y_wind"] = u, v
        for name, r in res.items():
            out[f"{tag}_ex_{name}"] = r
    np.savez_compressed(os.path.join(HERE, "reference_hirlam12.npz"), **out)
    print("wrote reference_hirlam12.npz with", len(out), "arrays")


if __name__ == "__main__":
    main()
    vertical_fixtures()
    hirlam12_fixture()
