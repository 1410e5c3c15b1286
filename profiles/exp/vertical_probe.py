"""Scratch timing of the vertical kernel by method (where do the instructions go?)."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import numpy as np, torch
from fimex_b200 import vertical as V
g = np.random.default_rng(2)
ncol = 30 * 801 * 561
eta = np.linspace(0.0, 1.0, 65)
ap, b = 100.0 * (1 - eta) * 0.2 * eta * 4 + 0.1, eta ** 1.5
ps = torch.from_numpy((101325 - 12 * g.uniform(0, 2500, 801 * 561) + g.normal(0, 300, 801 * 561)).astype(np.float32)).cuda().repeat(30).reshape(1, ncol).contiguous()
vin = torch.randn((1, 65, ncol), dtype=torch.float32, device="cuda")
levels = np.array([1000, 925, 850, 700, 500, 400, 300, 250, 200, 150, 100, 50], dtype=np.float64)
def t(fn, reps=5):
    for _ in range(2): fn()
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / reps
for name, m in (("lin", 0), ("log", 1), ("loglog", 2), ("nn", 3)):
    print(name, "hybrid", round(t(lambda: V.verticalHybridPressure(m, vin, ap, b, ps, levels, ps_scale=0.01)), 3), "ms")
p = (torch.from_numpy(ap).cuda()[None, :, None] + torch.from_numpy(b).cuda()[None, :, None] * (ps.double() * 0.01)[:, None, :]).float().contiguous()
for name, m in (("lin", 0), ("log", 1)):
    print(name, "explicit coordinate", round(t(lambda: V.verticalInterpolate(m, vin, p, level1=levels)), 3), "ms")
print(name, "1 level", round(t(lambda: V.verticalHybridPressure(1, vin, ap, b, ps, levels[:1], ps_scale=0.01)), 3), "ms")
