"""Synthetic MEPS-shaped workloads of BASELINE.json / SURVEY.md section 8(d) (inputs only, no regridding code).

S  : Lambert conformal source grid "+proj=lcc +lat_0=63.3 +lon_0=15 +lat_1=63.3 +lat_2=63.3 +no_defs +R=6.371e+06",
     x = -1060084 + 2500 i (949), y = -1332518 + 2500 j (1069), 65 levels, 67 times, float32.
T1 : 0.025 degree lat/lon, lon -18.0 .. 54.25 (2891) x lat 49.75 .. 75.25 (1021)  (36.6 % of targets outside S)
T2 : lon 5 .. 25 (801) x lat 56 .. 70 (561)                                        (all inside)
The index tables produced here are *inputs* of the cached apply (what CDMInterpolator hands to
createCachedInterpolation); they come from the closed-form spherical LCC below so that the generator needs
neither PROJ nor this library.  The library's own cache build (fimexcu_reproject_axes + points2position) is
checked against the same closed form in the tests.
"""
import math

import numpy as np

MEPS_PROJ = "+proj=lcc +lat_0=63.3 +lon_0=15 +lat_1=63.3 +lat_2=63.3 +no_defs +R=6.371e+06"
LATLON_PROJ = "+proj=latlong +R=6.371e+06 +no_defs"
MEPS_NX, MEPS_NY, MEPS_NZ, MEPS_NT = 949, 1069, 65, 67
MEPS_X0, MEPS_Y0, MEPS_DX = -1060084.0, -1332518.0, 2500.0
SEED = 20261106


def dotted_axis(start, second, end):
    """Axis string "a,b,...,c" expanded like tokenizeDotted (include/fimex/TokenizeDotted.h:42-74): repeated
    `cur += (b - a)` while cur is before the end value by more than 1e-5 of a step; the end value is appended."""
    vals = [start, second]
    dist = second - start
    direction = 1.0 if dist > 0 else -1.0
    cur = second + dist
    round_error = direction * dist * -1.0e-5
    while (cur - end) * direction < round_error:
        vals.append(cur)
        cur += dist
    vals.append(end)
    return np.array(vals, dtype=np.float64)


def meps_axes():
    x = MEPS_X0 + MEPS_DX * np.arange(MEPS_NX, dtype=np.float64)
    y = MEPS_Y0 + MEPS_DX * np.arange(MEPS_NY, dtype=np.float64)
    return x, y


def target_axes(name):
    if name == "T1":
        return dotted_axis(-18.0, -17.975, 54.25), dotted_axis(49.75, 49.775, 75.25)
    if name == "T2":
        return dotted_axis(5.0, 5.025, 25.0), dotted_axis(56.0, 56.025, 70.0)
    raise ValueError(name)


def lcc_sphere_forward(lon_deg, lat_deg, lat0=63.3, lon0=15.0, lat1=63.3, R=6.371e6):
    """Snyder (15-1)..(15-4), tangent cone on the sphere."""
    phi = np.radians(lat_deg)
    lam = np.radians(lon_deg - lon0)
    phi1, phi0 = math.radians(lat1), math.radians(lat0)
    n = math.sin(phi1)
    F = math.cos(phi1) * math.tan(math.pi / 4 + phi1 / 2) ** n / n
    rho = R * F / np.tan(np.pi / 4 + phi / 2) ** n
    rho0 = R * F / math.tan(math.pi / 4 + phi0 / 2) ** n
    return rho * np.sin(n * lam), rho0 - rho * np.cos(n * lam)


def meps_index_tables(target="T1"):
    """pointsOnXAxis / pointsOnYAxis (fractional source indices of every target cell, row-major [outY][outX])."""
    lon, lat = target_axes(target)
    lon2, lat2 = np.meshgrid(lon, lat)
    x, y = lcc_sphere_forward(lon2, lat2)
    px = (x - MEPS_X0) / MEPS_DX
    py = (y - MEPS_Y0) / MEPS_DX
    return np.ascontiguousarray(px), np.ascontiguousarray(py), lon.size, lat.size


def meps_slab_numpy(t=0, nz=MEPS_NZ, nx=MEPS_NX, ny=MEPS_NY, seed=SEED, nan_fraction=0.0):
    """One time step of the synthetic field: smooth part + N(0,1) noise."""
    rng = np.random.default_rng(seed + t)
    x = (MEPS_X0 + MEPS_DX * np.arange(nx))[None, None, :]
    y = (MEPS_Y0 + MEPS_DX * np.arange(ny))[None, :, None]
    k = np.arange(nz, dtype=np.float64)[:, None, None]
    v = 280.0 + 20.0 * np.sin(2 * np.pi * x / 1.2e6) * np.cos(2 * np.pi * y / 0.9e6) - 0.08 * k + 0.5 * t
    v = (v + rng.standard_normal(v.shape)).astype(np.float32)
    if nan_fraction > 0:
        v[rng.random(v.shape) < nan_fraction] = np.nan
    return v


def meps_slab_torch(t, device, nz=MEPS_NZ, nx=MEPS_NX, ny=MEPS_NY, seed=SEED):
    """Same field generated on the device (torch is only the allocator/RNG here)."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed + t)
    x = (MEPS_X0 + MEPS_DX * torch.arange(nx, device=device, dtype=torch.float32))[None, None, :]
    y = (MEPS_Y0 + MEPS_DX * torch.arange(ny, device=device, dtype=torch.float32))[None, :, None]
    k = torch.arange(nz, device=device, dtype=torch.float32)[:, None, None]
    v = 280.0 + 20.0 * torch.sin(2 * math.pi * x / 1.2e6) * torch.cos(2 * math.pi * y / 0.9e6) - 0.08 * k + 0.5 * t
    v = v + torch.randn((nz, ny, nx), device=device, dtype=torch.float32, generator=g)
    return v.contiguous()


def bilinear_algorithmic_bytes(n_in_reduced, n_out, nz, table_bytes=16):
    """per output point 4 (store) + 4*N_in'/N_out (every source element of the reduced window once)
    + table_bytes/nz; returned for one slab of nz levels."""
    per_point = 4.0 + 4.0 * n_in_reduced / n_out + table_bytes / nz
    return per_point * n_out * nz


def swath_lonlat(nx=1354, ny=2030, seed=SEED):
    """curvilinear longitude / latitude of a tilted satellite swath, NaN-free: [ny][nx] f64 each"""
    rng = np.random.default_rng(seed)
    u = np.linspace(-1, 1, nx)[None, :]
    v = np.linspace(0, 1, ny)[:, None]
    lat = 55 + 20 * v + 2.5 * u + 0.3 * np.sin(6 * v) * u
    lon = 5 + 18 * u / np.cos(np.radians(lat)) * 0.5 + 9 * v
    lat = lat + rng.normal(0, 1e-4, lat.shape)
    lon = lon + rng.normal(0, 1e-4, lon.shape)
    return np.ascontiguousarray(lon), np.ascontiguousarray(lat)
