"""The opt-in f32 bicubic (staged kernel) on the padded S->T1 slab: time per launch, NaN-free and with 2 % NaN blobs in the source."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import torch
import fimex_b200
from fimex_b200 import synthetic as S
px, py, ox, oy = S.meps_index_tables("T1")
ci = fimex_b200.CachedInterpolation("x", "y", 2, px, py, S.MEPS_NX, S.MEPS_NY, ox, oy)
ci.setFastBicubic(True)
src = ci.toPadded(ci.reduceInput(S.meps_slab_torch(t=1, device="cuda"), S.MEPS_NX, S.MEPS_NY))
out = torch.empty((65, oy, ci.paddedLayout()[2]), dtype=torch.float32, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for name, s in (("NaN-free", src), ("2 % NaN", torch.where(torch.rand_like(src) < 0.02, torch.full_like(src, float("nan")), src))):
    for i in range(3):
        ci.interpolateValuesPadded(s, out=out)
    torch.cuda.synchronize()
    e0.record()
    for i in range(10):
        ci.interpolateValuesPadded(s, out=out)
    e1.record(); torch.cuda.synchronize()
    print("fast bicubic, %s: %.4f ms" % (name, e0.elapsed_time(e1) / 10))
