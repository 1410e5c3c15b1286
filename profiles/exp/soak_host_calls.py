"""Soak: the host-pointer apply called a few hundred times (pageable input, result from the pinned pool, f32 and packed int16, plus a
vertical call): resident set size of the process and free device memory must stop growing after the first calls."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import numpy as np, psutil, torch
import fimex_b200
from fimex_b200 import synthetic as S, _lib, vertical as V

lib = fimex_b200.lib
px, py, ox, oy = S.meps_index_tables("T1")
ci = fimex_b200.CachedInterpolation("x", "y", 1, px, py, S.MEPS_NX, S.MEPS_NY, ox, oy)
nz = 16
src = np.random.default_rng(1).normal(size=(nz, ci.getInY(), ci.getInX())).astype(np.float32)
raw = (src * 1000).astype(np.int16)
n_in, n_out = src.size, nz * ox * oy
new_size = C.c_size_t()
proc = psutil.Process()
eta = np.linspace(0.0, 1.0, 65)
ap, b = 100.0 * (1 - eta) * 0.2 * eta * 4 + 0.1, eta ** 1.5
ncol = 200000
ps = (101325 - 12 * np.random.default_rng(2).uniform(0, 2500, ncol)).astype(np.float32).reshape(1, ncol)
vin = np.random.default_rng(3).normal(size=(1, 65, ncol)).astype(np.float32)
levels = np.array([1000, 850, 500, 250, 100], dtype=np.float64)


def once():
    p = lib.fimexcu_host_alloc(n_out * 4)
    _lib.check(lib.fimexcu_cache_apply_f(ci._h, src.ctypes.data, n_in, p, C.byref(new_size), float("nan"), float("nan")), "apply")
    lib.fimexcu_host_free(p)
    q = lib.fimexcu_host_alloc(n_out * 2)
    _lib.check(lib.fimexcu_cache_apply_i16(ci._h, raw.ctypes.data, n_in, q, C.byref(new_size), -32767.0), "apply_i16")
    lib.fimexcu_host_free(q)
    V.verticalHybridPressure(1, vin, ap, b, ps, levels, ps_scale=0.01)


marks = {}
for k in range(1, 301):
    once()
    if k in (20, 100, 200, 300):
        torch.cuda.synchronize()
        free, total = torch.cuda.mem_get_info()
        marks[k] = (proc.memory_info().rss / 2**20, (total - free) / 2**20)
        print("after %3d rounds: RSS %.0f MiB, device memory in use %.0f MiB" % (k, *marks[k]))
grow_rss = marks[300][0] - marks[100][0]
grow_dev = marks[300][1] - marks[100][1]
print("growth between round 100 and 300: RSS %+.1f MiB, device %+.1f MiB -> %s" % (grow_rss, grow_dev, "ok" if grow_rss < 16 and grow_dev < 16 else "LEAK?"))
