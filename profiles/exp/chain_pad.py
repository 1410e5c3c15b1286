"""Does the content of the pad columns of a padded slab (pitch 832 for 801 columns) decide the speed of the vertical kernel behind it?"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import numpy as np, torch
import fimex_b200
from fimex_b200 import synthetic as S, vertical as V

eta = np.linspace(0.0, 1.0, 65)
ap, b = 100.0 * (1 - eta) * 0.2 * eta * 4 + 0.1, eta ** 1.5
levels = np.array([1000, 925, 850, 700, 500, 400, 300, 250, 200, 150, 100, 50], dtype=np.float64)
px2, py2, ox2, oy2 = S.meps_index_tables("T2")
c2 = fimex_b200.CachedInterpolation("x", "y", 1, px2, py2, S.MEPS_NX, S.MEPS_NY, ox2, oy2)
ip2, il2, op2, ol2 = c2.paddedLayout()
nmem = 6
f_in = torch.randn((65 * nmem, c2.getInY(), ip2), dtype=torch.float32, device="cuda")
ps_in = 101325.0 + 300.0 * torch.randn((nmem, c2.getInY(), ip2), dtype=torch.float32, device="cuda")
n_inner = nmem * oy2 * op2


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for name, padval in (("zeros", 0.0), ("NaN", float("nan")), ("1e30", 1e30), ("plausible ps (101325) / 0 values", None)):
    f_out = torch.zeros((65 * nmem, oy2, op2), dtype=torch.float32, device="cuda")
    ps_out = torch.zeros((nmem, oy2, op2), dtype=torch.float32, device="cuda")
    if padval is None:
        ps_out.fill_(101325.0)
    else:
        f_out.fill_(padval); ps_out.fill_(padval)
    c2.interpolateValuesPadded(f_in, out=f_out)
    c2.interpolateValuesPadded(ps_in, out=ps_out)
    v = lambda: V.verticalHybridPressure(0, f_out.reshape(1, 65, n_inner), ap, b, ps_out.reshape(1, n_inner), levels, ps_scale=0.01)
    print("pad = %-34s vertical %.3f ms   (valid columns %d of %d; NaN in ps_out valid part: %d)" % (
        name, timed(v), nmem * oy2 * ox2, n_inner, int(torch.isnan(ps_out[:, :, :ox2]).sum())))
vr = lambda: V.verticalHybridPressure(0, f_out.reshape(1, 65, n_inner), ap, b, ps_out.reshape(1, n_inner), levels, ps_scale=0.01, rows=(ox2, op2))
print("rows = (%d, %d): pad skipped, whatever it holds: vertical %.3f ms" % (ox2, op2, timed(vr)))
f_out.zero_(); ps_out.zero_()
c2.interpolateValuesPadded(f_in, out=f_out)
c2.interpolateValuesPadded(ps_in, out=ps_out)
print("rows, pad = zeros: vertical %.3f ms" % timed(vr))
# a third of the columns undefined (a target that reaches beyond the source grid, like T1): NaN surface pressure
ps_nan = ps_out.clone()
ps_nan.reshape(-1)[::3] = float("nan")
print("rows, every third column NaN ps: vertical %.3f ms" % timed(
    lambda: V.verticalHybridPressure(0, f_out.reshape(1, 65, n_inner), ap, b, ps_nan.reshape(1, n_inner), levels, ps_scale=0.01, rows=(ox2, op2))))
ps_nan = ps_out.clone()
ps_nan[:, :, : ox2 // 3] = float("nan")
print("rows, the left third of every row NaN ps: vertical %.3f ms" % timed(
    lambda: V.verticalHybridPressure(0, f_out.reshape(1, 65, n_inner), ap, b, ps_nan.reshape(1, n_inner), levels, ps_scale=0.01, rows=(ox2, op2))))
# the same number of columns, dense
ncol = nmem * oy2 * ox2
f_d = torch.randn((1, 65, ncol), dtype=torch.float32, device="cuda")
ps_d = (101325.0 + 300.0 * torch.randn((1, ncol), dtype=torch.float32, device="cuda"))
print("dense columns, random values: vertical %.3f ms" % timed(lambda: V.verticalHybridPressure(0, f_d, ap, b, ps_d, levels, ps_scale=0.01)))
