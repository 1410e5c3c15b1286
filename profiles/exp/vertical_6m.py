"""Hybrid -> pressure for 6 members of T2 (dense columns), a few calls: for a launch list of the three kernels of a call."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import numpy as np, torch
from fimex_b200 import vertical as V
g = np.random.default_rng(2)
ncol = 6 * 801 * 561
eta = np.linspace(0.0, 1.0, 65)
ap, b = 100.0 * (1 - eta) * 0.2 * eta * 4 + 0.1, eta ** 1.5
ps = torch.from_numpy((101325 - 12 * g.uniform(0, 2500, ncol) + g.normal(0, 300, ncol)).astype(np.float32)).cuda().reshape(1, ncol)
vin = torch.randn((1, 65, ncol), dtype=torch.float32, device="cuda")
levels = np.array([1000, 925, 850, 700, 500, 400, 300, 250, 200, 150, 100, 50], dtype=np.float64)
m = int(sys.argv[1]) if len(sys.argv) > 1 else 0
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(4):
    e0.record()
    V.verticalHybridPressure(m, vin, ap, b, ps, levels, ps_scale=0.01)
    e1.record()
    torch.cuda.synchronize()
    print("method", m, "ms", round(e0.elapsed_time(e1), 4))
