"""How often does the device LOG weight differ from the oracle's (glibc) result?  Counts exact / 1-ulp / larger differences."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "tests"))
import numpy as np
import oracle_lib as ol
import test_gpu_vertical_forward as T
from fimex_b200.vertical import verticalHybridPressure
g = np.load(os.path.join(T.GOLD, "reference_vertical.npz"))
ap, b = g["en_ap_pa"] * 0.01, g["en_b"]
rng = np.random.default_rng(5)
nI = 200000
levels = np.array([1000, 925, 850, 700, 500, 400, 300, 250, 200, 150, 100, 50], dtype=np.float64)
ps = (101325 - 12 * rng.uniform(0, 2500, nI) + rng.normal(0, 300, nI)).astype(np.float32)
fld = rng.normal(0, 10, (1, 65, nI)).astype(np.float32)
p = (ap[:, None] + b[:, None] * (ps.astype(np.float64) * 0.01)[None, :]).astype(np.float32)[None]
for m in (ol.VINT_LOG, ol.VINT_LOGLOG):
    want = T._oracle_vertical(m, fld, np.ascontiguousarray(p), levels=levels)
    got = verticalHybridPressure(m, fld, ap, b, ps.reshape(1, nI), levels, ps_scale=0.01)
    both = ~np.isnan(want) & ~np.isnan(got)
    assert (np.isnan(want) == np.isnan(got)).all()
    d = np.abs(got[both].view(np.int32).astype(np.int64) - want[both].view(np.int32).astype(np.int64))
    print("method", m, "values", both.sum(), "exact", (d == 0).sum(), "1ulp", (d == 1).sum(), ">1ulp", (d > 1).sum(), "max rel", float(np.max(np.abs(got[both] - want[both]) / np.maximum(np.abs(want[both]), 1e-30))))
