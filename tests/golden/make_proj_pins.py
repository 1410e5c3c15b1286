"""Extract PROJ-generated longitude/latitude fields from the reference's own NetCDF fixtures.

Several files under /root/reference/test hold, next to a projected grid (axes + a `proj4` attribute on the grid-mapping variable), the
2-D double `longitude`/`latitude` fields fimex wrote for that grid -- that is, the output of the real PROJ library for
proj4 -> "+proj=longlat" at every grid point (CDMInterpolator / CoordinateSystemSliceBuilder generate them through
reproject::reproject_values, src/reproject.cc:351-415).  They are the only PROJ output available here (PROJ itself is not
installed), so they pin the projection arithmetic of oracle/miniproj.c and of the device projections.

Run in the build container only (reads /root/reference); writes tests/golden/reference_proj_pins.npz, which travels.
Per case: proj string, the sampled axis values x[i], y[i] (projection units: metres, or degrees for ob_tran axes), lon[i], lat[i]
in degrees, and `axis_rounding` (metres; 0 where the file stores the axes in floating point, 1 where the axes are integers that
were rounded AFTER lon/lat had been generated, so only a ~1 m pin is possible).
"""
import os

import numpy as np
from scipy.io import netcdf_file

REF = "/root/reference/test"
CASES = [  # file, grid-mapping variable, x axis, y axis
    ("c11.nc", None, None, None),
    ("testdata_vertical_ensemble_in.nc", None, None, None),
    ("verticalPressure.nc", None, None, None),
    ("verticalSigmaHybrid.nc", None, None, None),
    ("coordTest.nc", None, None, None),
    ("merge_target_top.nc", None, None, None),
]
MAX_POINTS = 6000


def find_grid(nc):
    gm = [n for n, v in nc.variables.items() if "proj4" in v._attributes]
    assert gm, "no grid mapping with a proj4 attribute"   # c11.nc carries two identical ones; the first belongs to longitude/latitude
    lon, lat = nc.variables["longitude"], nc.variables["latitude"]
    assert lon.dimensions == lat.dimensions and len(lon.dimensions) == 2
    ydim, xdim = lon.dimensions
    return gm[0], nc.variables[xdim], nc.variables[ydim], lon, lat


def main():
    out = {}
    names = []
    for fname, *_ in CASES:
        nc = netcdf_file(os.path.join(REF, fname), "r", mmap=False)
        gm, xv, yv, lon, lat = find_grid(nc)
        proj = nc.variables[gm]._attributes["proj4"].decode()
        x = np.asarray(xv.data, dtype=np.float64)
        y = np.asarray(yv.data, dtype=np.float64)
        integer_axes = np.issubdtype(xv.data.dtype, np.integer)
        units = xv._attributes.get("units", b"").decode()
        if units == "km":
            x, y = x * 1000.0, y * 1000.0
        rounding = 1.0 if integer_axes else 0.0   # float32 axes were widened to double BEFORE PROJ saw them: no rounding in between
        X, Y = np.meshgrid(x, y)
        LON = np.asarray(lon.data, dtype=np.float64)
        LAT = np.asarray(lat.data, dtype=np.float64)
        step = max(1, int(np.ceil(X.size / MAX_POINTS)))
        sel = np.arange(0, X.size, step)
        key = fname[:-3]
        names.append(key)
        out[key + "/proj"] = np.array(proj)
        out[key + "/x"] = X.ravel()[sel]
        out[key + "/y"] = Y.ravel()[sel]
        out[key + "/lon"] = LON.ravel()[sel]
        out[key + "/lat"] = LAT.ravel()[sel]
        out[key + "/axis_rounding"] = np.array(rounding)
        out[key + "/axis_dtype"] = np.array(str(xv.data.dtype))
        print(f"{fname}: {proj!r} axes {xv.data.dtype} units {units!r} grid {X.shape} -> {sel.size} points, rounding {rounding}")
    out["cases"] = np.array(names)
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_proj_pins.npz")
    np.savez_compressed(dst, **out)
    print("wrote", dst, os.path.getsize(dst), "bytes")


if __name__ == "__main__":
    main()
