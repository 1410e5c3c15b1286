"""Hypsometric integration: exactness against the oracle (glibc log) and time."""
import sys, os, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "tests"))
import numpy as np, torch
import oracle_lib as ol
from oracle_lib import fp
from fimex_b200 import vertical as V
rng = np.random.default_rng(3)
nz, nI = 65, 40000
eta = np.linspace(0.998, 0.01, nz)
sap = rng.uniform(650, 1040, (1, nI)).astype(np.float32)
p = (eta[None, :, None] * sap[:, None, :]).astype(np.float32)
sgp = rng.uniform(0, 25000, (1, nI)).astype(np.float32)
T = rng.uniform(205, 305, (1, nz, nI)).astype(np.float32)
q = rng.uniform(0, 0.02, (1, nz, nI)).astype(np.float32)
want = np.empty((1, nz, nI), np.float32)
assert ol.oracle().orc_vcoord_pressure_integration_f(fp(p), fp(sap), 1.0, fp(sgp), 1.0, fp(T), fp(q), 1, 1, nz, nI, 1, None, 0.0, fp(want)) == 1
got = V.pressureIntegrationAltitude(p, sap, sgp, T, q, start_high_p=1).cpu().numpy()
d = np.abs(got.view(np.int32).astype(np.int64) - want.view(np.int32).astype(np.int64))
print("values", d.size, "exact", int((d == 0).sum()), "1ulp", int((d == 1).sum()), ">1ulp", int((d > 1).sum()))
ncol = 6 * 801 * 561
p_f = (torch.linspace(0.998, 0.01, 65, device="cuda")[None, :, None] * (950.0 + 60.0 * torch.rand((1, 1, ncol), device="cuda"))).contiguous()
sap_f = (p_f[:, 0, :] / 0.998).contiguous(); sgp_f = 20000.0 * torch.rand((1, ncol), device="cuda")
t_f = 220.0 + 80.0 * torch.rand((1, 65, ncol), device="cuda"); q_f = 0.02 * torch.rand((1, 65, ncol), device="cuda")
for _ in range(3):
    V.pressureIntegrationAltitude(p_f, sap_f, sgp_f, t_f, q_f, start_high_p=1)
torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    V.pressureIntegrationAltitude(p_f, sap_f, sgp_f, t_f, q_f, start_high_p=1)
e1.record(); torch.cuda.synchronize()
print("integration, 2.7 M columns x 65 levels:", round(e0.elapsed_time(e1) / 5, 3), "ms")
