"""creepfill2d on MEPS-sized levels: GPU (all levels at once) against the reference's mifi_creepfill2d_f from oracle/_ref (one level)."""
import sys, os, time, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "tests"))
import numpy as np, torch
import oracle_lib as ol
from oracle_lib import fp
from fimex_b200 import interpolation as I
rng = np.random.default_rng(1)
ny, nx, nz = 1069, 949, 65
yy, xx = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
base = (280 + 15 * np.sin(xx / 90.0) * np.cos(yy / 70.0)).astype(np.float32)
land = (np.sin(xx / 37.0) * np.cos(yy / 53.0) > 0.55)
fld = np.repeat(base[None], nz, 0) + rng.normal(0, 1, (nz, ny, nx)).astype(np.float32)
fld[:, land] = np.nan
d = torch.from_numpy(fld).cuda()
I.mifi_creepfill2d_f(d.clone(), 20, 2); torch.cuda.synchronize()
for _ in range(2):
    x = d.clone(); torch.cuda.synchronize(); t0 = time.perf_counter()
    I.mifi_creepfill2d_f(x, 20, 2); torch.cuda.synchronize()
    print("GPU", nz, "levels:", round((time.perf_counter() - t0) * 1e3, 2), "ms")
r = ol.ref()
if r is not None:
    a = fld[0].copy(); n = C.c_size_t(); t0 = time.perf_counter()
    r.mifi_creepfill2d_f(nx, ny, fp(a), 20, bytes([2]), C.byref(n)); dt = time.perf_counter() - t0
    print("reference CPU, 1 level:", round(dt * 1e3, 1), "ms ->", nz, "levels on 1 core:", round(dt * nz, 2), "s")
    print("equal to GPU level 0:", bool(np.array_equal(a, x[0].cpu().numpy())))
