"""One forward-aggregation launch for ncu: python forward_one.py <method 6 mean|8 max|7 median> [levels]"""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import numpy as np, torch
import fimex_b200
from fimex_b200 import _lib, interpolation as I
lib = fimex_b200.lib
method = int(sys.argv[1]) if len(sys.argv) > 1 else 6
nz = int(sys.argv[2]) if len(sys.argv) > 2 else 65
tox = toy = 1000
gx, gy = np.meshgrid(np.arange(1354), np.arange(2030))
fwd = I.CachedForwardInterpolation("x", "y", method, (gx * (tox / 1354.0)).ravel() - 0.3, (gy * (toy / 2030.0)).ravel() - 0.3, 1354, 2030, tox, toy, level_compat=False)
fin = torch.randn((nz, fwd.getInY(), fwd.getInX()), dtype=torch.float32, device="cuda")
fout = torch.empty((nz, toy, tox), dtype=torch.float32, device="cuda")
nan = float("nan")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(3):
    e0.record()
    _lib.check(lib.fimexcu_forward_apply_dev_f(fwd._h, fin.data_ptr(), nz, fout.data_ptr(), nan, nan, 0, _lib.current_stream_ptr()), "fwd")
    e1.record()
    torch.cuda.synchronize()
    print("method", method, "levels", nz, "ms", round(e0.elapsed_time(e1), 4))
