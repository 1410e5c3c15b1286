"""Explicit-coordinate vertical interpolation (sweep kernel) once, for ncu: python vertical_explicit_one.py <method>"""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import numpy as np, torch
from fimex_b200 import vertical as V
m = int(sys.argv[1]) if len(sys.argv) > 1 else 0
g = np.random.default_rng(2)
ncol = 10 * 801 * 561
eta = np.linspace(0.0, 1.0, 65)
ap, b = 100.0 * (1 - eta) * 0.2 * eta * 4 + 0.1, eta ** 1.5
ps = torch.from_numpy((101325 - 12 * g.uniform(0, 2500, 801 * 561) + g.normal(0, 300, 801 * 561)).astype(np.float32)).cuda().repeat(10).reshape(1, ncol).contiguous()
vin = torch.randn((1, 65, ncol), dtype=torch.float32, device="cuda")
levels = np.array([1000, 925, 850, 700, 500, 400, 300, 250, 200, 150, 100, 50], dtype=np.float64)
p = (torch.from_numpy(ap).cuda()[None, :, None] + torch.from_numpy(b).cuda()[None, :, None] * (ps.double() * 0.01)[:, None, :]).float().contiguous()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(3):
    e0.record(); V.verticalInterpolate(m, vin, p, level1=levels); e1.record(); torch.cuda.synchronize()
    print("explicit method", m, "ms", round(e0.elapsed_time(e1), 3))
