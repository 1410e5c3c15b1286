"""Per-level time of the apply kernels for launches of 65 and 260 levels (S->T1), with the level-chunk cap at its default and
switched off (FIMEXCU_ZCHUNK_MAX=1000000): python zchunk_probe2.py"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import torch
import fimex_b200
from fimex_b200 import synthetic as S
from fimex_b200.interpolation import interpolateVectorValuesPadded

px, py, ox, oy = S.meps_index_tables("T1")


def timed(fn, reps=6):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for method, name in ((1, "bilinear"), (0, "nearest"), (2, "bicubic")):
    if method == 0:
        ci = fimex_b200.CachedNNInterpolation("x", "y", px, py, S.MEPS_NX, S.MEPS_NY, ox, oy)
    else:
        ci = fimex_b200.CachedInterpolation("x", "y", method, px, py, S.MEPS_NX, S.MEPS_NY, ox, oy)
    ip, il, op, ol = ci.paddedLayout()
    for nz in (65, 260):
        f_in = torch.randn((nz, ci.getInY(), ip), dtype=torch.float32, device="cuda")
        f_out = torch.empty((nz, oy, op), dtype=torch.float32, device="cuda")
        d_in = torch.randn((nz, ci.getInY(), ci.getInX()), dtype=torch.float32, device="cuda")
        res = {}
        res["staged/padded"] = timed(lambda: ci.interpolateValuesPadded(f_in, out=f_out))
        res["dense"] = timed(lambda: ci.interpolateValues(d_in))
        if method == 1:
            v_in = torch.randn_like(f_in)
            res["uv staged"] = timed(lambda: interpolateVectorValuesPadded(ci, None, f_in, v_in))
        print("cap=%s %-8s %4d levels: " % (os.environ.get("FIMEXCU_ZCHUNK_MAX", "default"), name, nz) +
              "  ".join("%s %.3f ms (%.2f us/level)" % (k, v, 1e3 * v / nz) for k, v in res.items()))
        del f_in, f_out, d_in
    ci.close()
