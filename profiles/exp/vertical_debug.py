import sys, os
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
import oracle_lib as ol
import test_gpu_vertical_forward as T
from fimex_b200.vertical import verticalInterpolate
method, kind, order = 0, "ties", "up"
rng = np.random.default_rng(1000 * method + 10 * len(kind) + len(order))
nO, nzi, nI = 2, 65, 389
fld, _ = T._columns(rng, nO, nzi, nI, "pressure")
step = rng.choice([8.0, 16.0, -8.0], (nO, 1, nI))
vc = np.ascontiguousarray((1.0e8 + step * np.arange(nzi)[None, :, None]).astype(np.float32))
levels = T._sorted_level_list(vc, order)
levels = sorted(levels + [1.0e8, 1.0e8 + 16, 1.0e8 + 20, 1.0e8 - 300, 1.0e8 + 8 * 64, 1.0e8 + 8 * 64 + 4, 2.0e8, 1.7e15], reverse=(order == "down"))
want = T._oracle_vertical(method, fld, vc, levels=levels)
got = verticalInterpolate(method, fld, vc, level1=levels)
bad = ~((got == want) | (np.isnan(got) & np.isnan(want)))
print("mismatches", bad.sum(), "of", bad.size)
idx = np.argwhere(bad)
for o, k, i in idx[:12]:
    print("outer", o, "level", k, levels[k], np.float32(levels[k]), "col", i, "step", step[o, 0, i], "got", got[o, k, i], "want", want[o, k, i], "c0", vc[o, 0, i], "cl", vc[o, -1, i])
print("levels with mismatches:", sorted(set(levels[k] for _, k, _ in idx)))
print("steps with mismatches:", sorted(set(step[o, 0, i] for o, _, i in idx)))
