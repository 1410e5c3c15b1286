"""One bicubic apply on the MEPS S->T1 slab (dense layout), for ncu captures."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import torch
import fimex_b200
from fimex_b200 import synthetic as S
px, py, ox, oy = S.meps_index_tables("T1")
ci = fimex_b200.CachedInterpolation("x", "y", 2, px, py, S.MEPS_NX, S.MEPS_NY, ox, oy)
src = S.meps_slab_torch(t=1, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(3):
    e0.record(); out, _ = ci.interpolateValues(src); e1.record(); torch.cuda.synchronize()
    print("bicubic ms", round(e0.elapsed_time(e1), 3))
