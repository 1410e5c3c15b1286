"""Where the 0.77 ms of bench.py's chained_bilinear_T2_then_vertical go: each call alone (CUDA events), the host time each call
takes to enqueue (wall clock, no synchronisation), and the chain at 6 and at 30 members."""
import os, sys, time
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


def host(fn, reps=10):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    dt = (time.perf_counter() - t0) / reps
    torch.cuda.synchronize()
    return dt * 1e3


for nmem in (6, 30):
    f_in = torch.randn((65 * nmem, c2.getInY(), ip2), dtype=torch.float32, device="cuda")
    ps_in = 101325.0 + 300.0 * torch.randn((nmem, c2.getInY(), ip2), dtype=torch.float32, device="cuda")
    f_out = torch.empty((65 * nmem, oy2, op2), dtype=torch.float32, device="cuda")
    ps_out = torch.empty((nmem, oy2, op2), dtype=torch.float32, device="cuda")
    n_inner = nmem * oy2 * op2
    h = lambda: c2.interpolateValuesPadded(f_in, out=f_out)
    p = lambda: c2.interpolateValuesPadded(ps_in, out=ps_out)
    v = lambda: V.verticalHybridPressure(0, f_out.reshape(1, 65, n_inner), ap, b, ps_out.reshape(1, n_inner), levels, ps_scale=0.01)

    def chain():
        h(); p()
        return v()
    chain()
    print("members", nmem)
    for name, fn in (("bilinear 65 x members", h), ("bilinear ps", p), ("vertical", v), ("chain", chain)):
        print("  %-22s device %.3f ms   host enqueue %.3f ms" % (name, timed(fn), host(fn)))
    del f_in, ps_in, f_out, ps_out
