"""Latency of the host-pointer entry point on small slabs (BASELINE configs[0] size: hirlam12 229x196 -> 300x300, 1..4 levels)."""
import sys, os, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import numpy as np
import fimex_b200
rng = np.random.default_rng(0)
in_x, in_y, out_x, out_y = 229, 196, 300, 300
px = rng.uniform(0, in_x - 1, out_x * out_y); py = rng.uniform(0, in_y - 1, out_x * out_y)
ci = fimex_b200.CachedInterpolation("x", "y", 1, px, py, in_x, in_y, out_x, out_y)
for nz in (1, 4, 16):
    f = rng.normal(size=(nz, ci.getInY(), ci.getInX())).astype(np.float32)
    for _ in range(5):
        ci.interpolateValues(f)
    t0 = time.perf_counter(); n = 200
    for _ in range(n):
        ci.interpolateValues(f)
    print("nz", nz, "host-pointer call:", round((time.perf_counter() - t0) / n * 1e6, 1), "us")
