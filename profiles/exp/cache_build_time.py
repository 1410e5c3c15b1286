"""Wall time of the cache build from projection strings (T1: 2.95 M targets), repeated: python cache_build_time.py"""
import sys, os, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import torch
import fimex_b200
from fimex_b200 import reproject as RP, synthetic as S
lon, lat = S.target_axes("T1")
mx, my = S.meps_axes()
for i in range(6):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    c = RP.cachedInterpolationFromProjection(1, S.LATLON_PROJ, S.MEPS_PROJ, lon, lat, mx, my)
    torch.cuda.synchronize()
    print("build %d: %.3f ms" % (i, (time.perf_counter() - t0) * 1e3))
    c.close()
