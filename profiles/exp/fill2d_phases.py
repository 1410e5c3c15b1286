"""Where does fill2d's time go?  Set-up only (maxLoop = 0) against 1, 11, 51, 100 sweeps; 65 MEPS-sized levels."""
import sys, os, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import numpy as np, torch
from fimex_b200 import interpolation as I
rng = np.random.default_rng(1)
ny, nx, nz = 1069, 949, 65
yy, xx = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
base = (280 + 15 * np.sin(xx / 90.0) * np.cos(yy / 70.0)).astype(np.float32)
land = (np.sin(xx / 37.0) * np.cos(yy / 53.0) > 0.55)
fld = np.repeat(base[None], nz, 0) + rng.normal(0, 1, (nz, ny, nx)).astype(np.float32)
fld[:, land] = np.nan
d = torch.from_numpy(fld).cuda()
for ml in (0, 1, 11, 51, 100):
    for rep in range(2):
        x = d.clone(); torch.cuda.synchronize(); t0 = time.perf_counter()
        I.mifi_fill2d_f(x, 1e-9, 1.6, ml); torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    print("maxLoop", ml, round(dt * 1e3, 2), "ms")
