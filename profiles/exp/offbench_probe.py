"""Off-benchmark conditions of the main kernels (looking for cliffs): bad / fill conversions on, NaN-heavy inputs, other level counts."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import numpy as np, torch
import fimex_b200
from fimex_b200 import synthetic as S, vertical as V
from fimex_b200.interpolation import interpolateVectorValuesPadded


def timed(fn, reps=6):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


px, py, ox, oy = S.meps_index_tables("T1")
ci = fimex_b200.CachedInterpolation("x", "y", 1, px, py, S.MEPS_NX, S.MEPS_NY, ox, oy)
ip, il, op, ol = ci.paddedLayout()
nz = 65
f = torch.randn((nz, ci.getInY(), ip), dtype=torch.float32, device="cuda")
o = torch.empty((nz, oy, op), dtype=torch.float32, device="cuda")
print("K4 staged plain            %.4f ms" % timed(lambda: ci.interpolateValuesPadded(f, out=o)))
print("K4 staged bad+fill         %.4f ms" % timed(lambda: ci.interpolateValuesPadded(f, out=o, badValue=-999.0, fillValue=1e20)))
fn = f.clone(); fn[torch.rand_like(fn) < 0.3] = float("nan")
print("K4 staged 30 %% NaN input   %.4f ms" % timed(lambda: ci.interpolateValuesPadded(fn, out=o)))
v = torch.randn_like(f)
print("uv staged plain            %.4f ms" % timed(lambda: interpolateVectorValuesPadded(ci, None, f, v)))
print("uv staged bad+fill         %.4f ms" % timed(lambda: interpolateVectorValuesPadded(ci, None, f, v, badU=-999.0, badV=-999.0, fillU=1e20, fillV=1e20)))
del f, o, fn, v
ci.close()

g = np.random.default_rng(2)
for nzi, nzo in ((65, 12), (137, 12), (137, 30), (65, 30), (40, 6)):
    ncol = 6 * 801 * 561
    eta = np.linspace(0.0, 1.0, nzi)
    ap, b = 100.0 * (1 - eta) * 0.2 * eta * 4 + 0.1, eta ** 1.5
    ps = torch.from_numpy((101325 - 12 * g.uniform(0, 2500, ncol) + g.normal(0, 300, ncol)).astype(np.float32)).cuda().reshape(1, ncol)
    vin = torch.randn((1, nzi, ncol), dtype=torch.float32, device="cuda")
    levels = np.geomspace(1000, 50, nzo)
    for m, name in ((0, "lin"), (1, "log")):
        ms = timed(lambda: V.verticalHybridPressure(m, vin, ap, b, ps, levels, ps_scale=0.01), reps=3)
        alg = (4.0 * nzi + 4 + 4 * nzo) * ncol
        print("vertical hybrid %3d -> %2d %s: %.3f ms  (%.0f GB/s algorithmic)" % (nzi, nzo, name, ms, alg / ms / 1e6))
    p_e = (torch.from_numpy(ap).cuda().float()[None, :, None] + torch.from_numpy(b).cuda().float()[None, :, None] * (ps[:, None, :] * 0.01)).contiguous()
    ms = timed(lambda: V.verticalInterpolate(0, vin, p_e, level1=levels), reps=3)
    print("vertical explicit %3d -> %2d lin: %.3f ms  (%.0f GB/s algorithmic)" % (nzi, nzo, ms, (8.0 * nzi + 4 * nzo) * ncol / ms / 1e6))
    if nzi == 65 and nzo == 12:
        p_n = p_e.clone(); p_n[:, :, ::3] = float("nan")
        ms = timed(lambda: V.verticalInterpolate(0, vin, p_n, level1=levels), reps=3)
        print("vertical explicit, every third column all NaN: %.3f ms" % ms)
        p_n = p_e.clone(); p_n[:, :, : ncol // 3] = float("nan")
        ms = timed(lambda: V.verticalInterpolate(0, vin, p_n, level1=levels), reps=3)
        print("vertical explicit, first third of the columns all NaN: %.3f ms" % ms)
        vn = vin.clone(); vn[torch.rand_like(vn) < 0.3] = float("nan")
        ms = timed(lambda: V.verticalHybridPressure(1, vn, ap, b, ps, levels, ps_scale=0.01), reps=3)
        print("vertical hybrid log, 30 %% NaN values: %.3f ms" % ms)
        del p_n, vn
    del vin, p_e, ps
