"""Staged bilinear S->T2 with many levels per launch: time per level against the level chunk a CTA walks (FIMEXCU_ZCHUNK_MAX, a
probe-only knob)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import torch
import fimex_b200
from fimex_b200 import synthetic as S

tgt = sys.argv[1] if len(sys.argv) > 1 else "T2"
px2, py2, ox2, oy2 = S.meps_index_tables(tgt)
c2 = fimex_b200.CachedInterpolation("x", "y", 1, px2, py2, S.MEPS_NX, S.MEPS_NY, ox2, oy2)
ip2, il2, op2, ol2 = c2.paddedLayout()
for nz in [int(a) for a in sys.argv[2:]] or [65, 390, 1950]:
    f_in = torch.randn((nz, c2.getInY(), ip2), dtype=torch.float32, device="cuda")
    f_out = torch.empty((nz, oy2, op2), dtype=torch.float32, device="cuda")
    for _ in range(3):
        c2.interpolateValuesPadded(f_in, out=f_out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        c2.interpolateValuesPadded(f_in, out=f_out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print("%s zchunk_max=%s levels %5d: %.3f ms, %.3f us per level" % (tgt, os.environ.get("FIMEXCU_ZCHUNK_MAX", "-"), nz, ms, 1e3 * ms / nz))
    del f_in, f_out
