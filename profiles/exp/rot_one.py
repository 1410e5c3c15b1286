"""the stand-alone rotation of u, v [65][1021][2891] for timing: python rot_one.py"""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import numpy as np, torch
import fimex_b200
from fimex_b200 import _lib
lib = fimex_b200.lib
ox, oy, nz = 2891, 1021, 65
rng = np.random.default_rng(1)
ang = rng.uniform(-3, 3, ox * oy)
m3 = np.stack([np.cos(ang), np.sin(ang), np.degrees(ang)], axis=1).copy()
cvr = fimex_b200.CachedVectorReprojection(m3, ox, oy)
u = torch.randn((nz, oy, ox), dtype=torch.float32, device="cuda"); v = torch.randn_like(u)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
best = 1e9
for i in range(6):
    e0.record()
    _lib.check(lib.fimexcu_vecrot_apply_dev_f(cvr._h, u.data_ptr(), v.data_ptr(), nz, _lib.current_stream_ptr()), "rot")
    e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
print("zchunk", os.environ.get("FIMEXCU_ROT_ZCHUNK", "default"), "ms", round(best, 4))
