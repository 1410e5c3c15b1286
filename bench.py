#!/usr/bin/env python
"""bench.py -- cached bilinear regridding of the synthetic MEPS field (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--target T1|T2]

One step = one pass of the hot path over one time-step slab: source [65][1069][949] f32 (Lambert, 2.5 km)
-> target T1 [65][1021][2891] f32 (0.025 degree lat/lon), i.e. 191.9 M regridded points, through
fimexcu_cache_apply_dev_f (one kernel launch) with the slab resident in HBM.  With N GPUs (torchrun, one
process per GPU) every rank regrids its own slabs: the time/level/ensemble slabs are independent, so the
path shards without any collective and the scaling is "weak".  Timed with CUDA events on the launching
stream, barrier + synchronize on both sides, max over ranks.

Prints ONE JSON line (rank 0).  Extra objects:
  roofline      algorithmic bytes per launch (4 + 4*N_in'/N_out + 16/nz bytes per point) over the
                mean launch duration, against the measured HBM copy bandwidth of MEASURED_PEAKS.json
  cpu_baseline  the reference's own CPU code (oracle/_ref, OpenMP, all host cores) on a bounded sample
  e2e           same metric through the host-pointer C ABI call (pinned host buffers; H2D and D2H inside)
`--impl reference` times only the reference CPU implementation (oracle/_ref, else the oracle port).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "regridded Gpoints/s (cached bilinear, whole job)"
UNIT = "Gpoints/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--target", default="T1", choices=["T1", "T2"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-job", action="store_true", help="skip the 67-slab job figures")
    ap.add_argument("--full-job", action="store_true", help="run every slab of the job through the host ABI instead of a bounded sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the short measurements of the other kernels of the path")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx = float(parts[1])
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def _load_by_path(name):
    """fimex_b200/<name>.py loaded by file path: the reference arm must not import the package (whose __init__ maps libfimexcu.so)"""
    import importlib.util
    spec = importlib.util.spec_from_file_location("_fimex_b200_" + name, os.path.join(ROOT, "fimex_b200", name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def workload(target):
    S = _load_by_path("synthetic")  # inputs only: axes, index tables, the synthetic field
    px, py, ox, oy = S.meps_index_tables(target)
    return S, px, py, ox, oy


def workload_name(S, target, ox, oy):
    """config.workload, the same string in both arms"""
    pts = S.MEPS_NZ * ox * oy
    return (f"BASELINE configs[1]: synthetic MEPS Lambert 949x1069 x 65 levels f32 -> {target} 0.025deg lat/lon "
            f"({ox}x{oy}), cached bilinear, one time-step slab ({pts / 1e6:.1f} Mpoints) per step per GPU")


# ------------------------------------------------------------------------------------------------ reference arm
def cpu_vertical_port(ncol=60000):
    """configs[4] on the CPU: the oracle's restatement of CDMVerticalInterpolator::getLevelDataSlice (the reference's loop is not
    OpenMP-parallel) on a bounded sample of columns, 65 hybrid levels -> 12 pressure levels, log.  Part of the cpu_baseline leg."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    rng = np.random.default_rng(1)
    g = np.load(os.path.join(ROOT, "tests", "golden", "reference_vertical.npz"))
    ap, b = g["en_ap_pa"] * 0.01, g["en_b"]
    levels = np.array([1000, 925, 850, 700, 500, 400, 300, 250, 200, 150, 100, 50], dtype=np.float64)
    ps = (101325 - 12 * rng.uniform(0, 2500, ncol) + rng.normal(0, 300, ncol)).astype(np.float32)
    fld = rng.normal(0, 10, (65, ncol)).astype(np.float32)
    t0 = time.perf_counter()
    p = np.ascontiguousarray((ap[:, None] + b[:, None] * (ps.astype(np.float64) * 0.01)[None, :]).astype(np.float32))  # the converter's part
    out = np.empty((12, ncol), np.float32)
    rc = ol.oracle().orc_vertical_interpolate_f(ol.VINT_LOG, ol.fp(fld), ol.fp(p), 65, ncol, ol.dp(levels), None, 12, None, None, float("nan"),
                                                float("nan"), ol.fp(out))
    dt = time.perf_counter() - t0
    if rc != 1:
        raise RuntimeError("oracle vertical interpolation failed")
    return {"value": 12 * ncol / dt / 1e9, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"{ncol} columns, 65 hybrid levels -> 12 pressure levels, log (oracle restatement of getLevelDataSlice + hybrid pressure)"}


def cpu_reference(px, py, S, ox, oy, budget_s=12.0, max_levels=None):
    """Times CachedInterpolation::interpolateValues of the reference (oracle/_ref, built with OpenMP = the reference's
    ENABLE_FIMEX_OMP) on a bounded number of levels of the same slab.  Falls back to the scalar oracle port."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    lib = ol.ref(omp=True) or ol.ref(omp=False)
    cores = os.cpu_count() or 1
    if lib is not None:
        kind = "reference"
        lib.ref_set_num_threads(cores)
        threads = lib.ref_num_threads()
        h = lib.ref_cache_create(ol.BILINEAR, ol.dp(px.ravel()), ol.dp(py.ravel()), S.MEPS_NX, S.MEPS_NY, ox, oy)
        import ctypes as C
        v = [C.c_size_t() for _ in range(4)]
        red, xm, ym = C.c_int(), C.c_size_t(), C.c_size_t()
        lib.ref_cache_info(h, *[C.byref(t) for t in v], C.byref(red), C.byref(xm), C.byref(ym))
        inx, iny = v[0].value, v[1].value

        fields = {}

        def run(levels, reps):
            if levels not in fields:
                fields.clear()
                fields[levels] = S.meps_slab_numpy(t=0, nz=levels)[:, ym.value:ym.value + iny, xm.value:xm.value + inx].copy()
            fld = fields[levels]
            return lib.ref_cache_apply_seconds(h, ol.fp(fld), fld.size, reps) / reps
    else:
        kind, threads = "port", 1
        o = ol.oracle()
        h = o.orc_cached_interp_create(ol.BILINEAR, ol.dp(px.ravel()), ol.dp(py.ravel()), S.MEPS_NX, S.MEPS_NY, ox, oy)
        c = h.contents

        fields = {}

        def run(levels, reps):
            if levels not in fields:
                fields.clear()
                fields[levels] = S.meps_slab_numpy(t=0, nz=levels)[:, c.yMin:c.yMin + c.inY, c.xMin:c.xMin + c.inX].copy()
            fld = fields[levels]
            out = np.empty((levels, oy, ox), np.float32)
            t0 = time.perf_counter()
            for _ in range(reps):
                o.orc_cached_interp_apply(h, ol.fp(fld), fld.size, ol.fp(out))
            return (time.perf_counter() - t0) / reps
    t1 = run(1, 1)  # calibration: one level
    levels = int(max(1, min(S.MEPS_NZ if max_levels is None else max_levels, budget_s / 2 / max(t1, 1e-4))))
    return dict(run=run, levels=levels, kind=kind, threads=threads, t_level=t1)


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    S, px, py, ox, oy = workload(args.target)
    ref = cpu_reference(px, py, S, ox, oy, budget_s=2.0)
    levels = ref["levels"]
    pts = levels * ox * oy
    for _ in range(args.warmup):
        ref["run"](levels, 1)
    dt = sum(ref["run"](levels, 1) for _ in range(args.steps)) / args.steps  # seconds inside interpolateValues
    value = pts / dt / 1e9
    sample = f"{levels} of {S.MEPS_NZ} levels of one S->{args.target} slab per step ({pts / 1e6:.1f} Mpoints), {ref['threads']} OpenMP threads"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(S, args.target, ox, oy),
                   "arm": "reference CPU path (CachedInterpolation::interpolateValues compiled from the reference sources) on the host cores",
                   "sample": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": ref["threads"], "kind": ref["kind"], "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0



# ------------------------------------------------------------------------------------------------ the other kernels of the path
def secondary_kernels(S, ci, px, py, ox, oy, peak):
    """Short CUDA-event measurements (3 warm-up + 5 timed launches each) of the remaining SURVEY 8 kernels on BASELINE-shaped
    inputs, each with its algorithmic bytes against the same measured HBM peak.  Reported under "kernels"."""
    import torch
    import fimex_b200
    from fimex_b200 import _lib, interpolation as I, vertical as V, reproject as RP
    lib = fimex_b200.lib
    nz, nan = S.MEPS_NZ, float("nan")
    stream = _lib.current_stream_ptr()
    out = {}

    def timed(fn, reps=5):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    def entry(ms, alg_bytes, points, what):
        gbs = alg_bytes / (ms * 1e-3) / 1e9
        return {"ms": ms, "Gpoints_per_s": points / (ms * 1e-3) / 1e9, "achieved_GBs": gbs, "frac_of_hbm_peak": gbs / peak,
                "algorithmic_bytes": alg_bytes, "what": what}

    inx, iny = ci.getInX(), ci.getInY()
    pts = nz * ox * oy
    alg = S.bilinear_algorithmic_bytes(inx * iny, ox * oy, nz)
    src = S.meps_slab_torch(t=7, device="cuda")
    dst = torch.empty((nz, oy, ox), dtype=torch.float32, device="cuda")
    for name, method in (("K4_bilinear_apply_dense", 1), ("K5_bicubic_apply", 2), ("K6_nearest_apply", 0)):
        c = fimex_b200.createCachedInterpolation("x", "y", method, px, py, S.MEPS_NX, S.MEPS_NY, ox, oy)
        sub = c.reduceInput(src, S.MEPS_NX, S.MEPS_NY)
        ms = timed(lambda: _lib.check(lib.fimexcu_cache_apply_dev_f(c._h, sub.data_ptr(), nz, dst.data_ptr(), nan, nan, stream), name))
        out[name] = entry(ms, alg, pts, "S->T1 slab, the reference's dense layout (row pitch 949: no tensor map possible)")
        if method in (0, 2):  # nearest and bicubic also through the staged kernels (padded layout)
            pin = c.toPadded(sub)
            ip, il, op, ol_ = c.paddedLayout()
            pout = torch.empty((nz, oy, op), dtype=torch.float32, device="cuda")
            ms = timed(lambda: _lib.check(lib.fimexcu_cache_apply_dev_strided_f(c._h, pin.data_ptr(), ip, il, nz, pout.data_ptr(), op, ol_, nan, nan, stream), name))
            out[name + "_padded_staged"] = entry(ms, alg, pts, "S->T1 slab, padded layout, TMA-staged" +
                                                 (" (bit-exact: the direct kernel; 40 separately rounded FP64 operations per value bound any bit-exact form at 0.41 ms = 40 %)" if method == 2 else ""))
            if method == 2:
                c.setFastBicubic(True)
                ms = timed(lambda: _lib.check(lib.fimexcu_cache_apply_dev_strided_f(c._h, pin.data_ptr(), ip, il, nz, pout.data_ptr(), op, ol_, nan, nan, stream), name))
                out[name + "_padded_staged_f32"] = entry(ms, alg, pts, "same, opt-in f32 FMA arithmetic (1e-5 contract, identical undefined mask)")
                c.setFastBicubic(False)
            del pin, pout
        c.close()
    # K7: fused u,v interpolation + rotation, and the stand-alone rotation
    ang = np.random.default_rng(1).uniform(-0.6, 0.6, ox * oy)
    cvr = I.CachedVectorReprojection(np.stack([np.cos(ang), np.sin(ang), np.degrees(ang)], axis=1).copy(), ox, oy)
    dst2 = torch.empty_like(dst)
    sub = ci.reduceInput(src, S.MEPS_NX, S.MEPS_NY)
    ms = timed(lambda: _lib.check(lib.fimexcu_cache_apply_uv_dev_f(ci._h, cvr._h, sub.data_ptr(), sub.data_ptr(), nz, dst.data_ptr(), dst2.data_ptr(),
                                                                   nan, nan, nan, nan, stream), "uv"))
    out["K7_fused_uv_bilinear_rotation"] = entry(ms, 2 * alg + 16.0 * ox * oy, 2 * pts, "u and v of one S->T1 slab in one pass")
    pin = ci.toPadded(sub)
    ip, il, op, ol_ = ci.paddedLayout()
    pu, pv = torch.empty((nz, oy, op), dtype=torch.float32, device="cuda"), torch.empty((nz, oy, op), dtype=torch.float32, device="cuda")
    ms = timed(lambda: _lib.check(lib.fimexcu_cache_apply_uv_dev_strided_f(ci._h, cvr._h, pin.data_ptr(), pin.data_ptr(), ip, il, nz, pu.data_ptr(),
                                                                           pv.data_ptr(), op, ol_, nan, nan, nan, nan, stream), "uv"))
    out["K7_fused_uv_bilinear_rotation_padded_staged"] = entry(ms, 2 * alg + 16.0 * ox * oy, 2 * pts, "same, padded layout, TMA-staged")
    pu.normal_()
    pv.normal_()
    ms = timed(lambda: _lib.check(lib.fimexcu_vecrot_apply_dev_strided_f(cvr._h, pu.data_ptr(), pv.data_ptr(), op, ol_, nz, stream), "rot"))
    out["K7_rotation_alone_padded"] = entry(ms, (16.0 + 16.0 / nz) * pts, 2 * pts, "in-place rotation of u,v in the padded layout (rows on 128 B: 16-byte accesses)")
    del pin, pu, pv
    ms = timed(lambda: _lib.check(lib.fimexcu_vecrot_apply_dev_f(cvr._h, dst.data_ptr(), dst2.data_ptr(), nz, stream), "rot"))
    out["K7_rotation_alone"] = entry(ms, (16.0 + 16.0 / nz) * pts, 2 * pts, "in-place rotation of u,v [65][1021][2891]")
    cvr.close()
    del dst2
    # K14+K15: hybrid-sigma -> 12 pressure levels, log, 30 members on the T2 grid (configs[4] shape, one time step)
    g = np.random.default_rng(2)
    ncol = 30 * 801 * 561
    eta = np.linspace(0.0, 1.0, 65)
    ap, b = 100.0 * (1 - eta) * 0.2 * eta * 4 + 0.1, eta ** 1.5  # hPa; smooth hybrid coefficients, model top ~0.1 hPa
    ps = torch.from_numpy((101325 - 12 * g.uniform(0, 2500, 801 * 561) + g.normal(0, 300, 801 * 561)).astype(np.float32)).cuda()
    ps = ps.repeat(30).reshape(1, ncol).contiguous()
    vin = torch.randn((1, 65, ncol), dtype=torch.float32, device="cuda")
    levels = np.array([1000, 925, 850, 700, 500, 400, 300, 250, 200, 150, 100, 50], dtype=np.float64)
    ms = timed(lambda: V.verticalHybridPressure(1, vin, ap, b, ps, levels, ps_scale=0.01), reps=3)
    out["K14_K15_vertical_hybrid_to_pressure_log"] = entry(ms, 312.0 * ncol, 12 * ncol, "65 -> 12 levels, 30 members x 801x561 columns, p formed on the fly")
    ms = timed(lambda: V.verticalHybridPressure(0, vin, ap, b, ps, levels, ps_scale=0.01), reps=3)
    out["K14_K15_vertical_hybrid_to_pressure_linear"] = entry(ms, 312.0 * ncol, 12 * ncol, "same, method linear")
    # explicit vertical coordinate (a materialised pressure field, as for non-hybrid sources): the one-pass sweep kernel; the
    # coordinate is read too: 4*65 more bytes per column
    ncol_e = 10 * 801 * 561
    p_e = (torch.from_numpy(ap).cuda().float()[None, :, None] + torch.from_numpy(b).cuda().float()[None, :, None] * (ps[:, None, :ncol_e] * 0.01)).contiguous()
    vin_e = vin[:, :, :ncol_e].contiguous()
    ms = timed(lambda: V.verticalInterpolate(1, vin_e, p_e, level1=levels), reps=3)
    out["K15_vertical_explicit_coordinate_sweep_log"] = entry(ms, (312.0 + 260.0) * ncol_e, 12 * ncol_e,
                                                               "65 -> 12 levels, 10 members x 801x561 columns, coordinate field read: 572 B per column")
    del vin, ps, p_e, vin_e
    # configs[4] second half, chained on the device: bilinear S->T2 of (65 levels x 6 members) and of ps, then hybrid -> 12 pressure levels
    lon2, lat2 = S.target_axes("T2")
    px2, py2, ox2, oy2 = S.meps_index_tables("T2")
    c2 = fimex_b200.CachedInterpolation("x", "y", 1, px2, py2, S.MEPS_NX, S.MEPS_NY, ox2, oy2)
    nmem = 6
    ip2, il2, op2, ol2 = c2.paddedLayout()
    f_in = torch.randn((65 * nmem, c2.getInY(), ip2), dtype=torch.float32, device="cuda")
    ps_in = 101325.0 + 300.0 * torch.randn((nmem, c2.getInY(), ip2), dtype=torch.float32, device="cuda")
    f_out = torch.empty((65 * nmem, oy2, op2), dtype=torch.float32, device="cuda")
    ps_out = torch.empty((nmem, oy2, op2), dtype=torch.float32, device="cuda")
    n_inner = nmem * oy2 * op2

    ms = timed(lambda: c2.interpolateValuesPadded(f_in, out=f_out), reps=5)
    out["K4_bilinear_T2_padded_staged"] = entry(ms, S.bilinear_algorithmic_bytes(c2.getInX() * c2.getInY(), ox2 * oy2, 65 * nmem), 65 * nmem * ox2 * oy2,
                                                "S->T2 (interior target, reduced source window 505x642, 7.13 B/pt), 390 levels per launch")

    def chain():
        c2.interpolateValuesPadded(f_in, out=f_out)
        c2.interpolateValuesPadded(ps_in, out=ps_out)
        return V.verticalHybridPressure(0, f_out.reshape(1, 65, n_inner), ap, b, ps_out.reshape(1, n_inner), levels, ps_scale=0.01, rows=(ox2, op2))
    ms = timed(chain, reps=3)
    alg_h = S.bilinear_algorithmic_bytes(c2.getInX() * c2.getInY(), ox2 * oy2, 65 * nmem) + S.bilinear_algorithmic_bytes(c2.getInX() * c2.getInY(), ox2 * oy2, nmem)
    alg_v = 312.0 * nmem * ox2 * oy2
    out["chained_bilinear_T2_then_vertical"] = entry(ms, alg_h + alg_v, 12 * nmem * ox2 * oy2,
                                                     "configs[4]: 6 members x 65 levels S->T2 (7.13 B/pt) + ps, then 65->12 pressure levels (linear); points = final output values")
    c2.close()
    del f_in, ps_in, f_out, ps_out
    # vertical coordinate converters (8(f) rank 2): hypsometric integration to altitude, 65 levels, with humidity
    ncol_v = 6 * 801 * 561
    p_f = (torch.linspace(0.998, 0.01, 65, device="cuda")[None, :, None] * (950.0 + 60.0 * torch.rand((1, 1, ncol_v), device="cuda"))).contiguous()
    sap_f = (p_f[:, 0, :] / 0.998).contiguous()
    sgp_f = 20000.0 * torch.rand((1, ncol_v), device="cuda")
    t_f = 220.0 + 80.0 * torch.rand((1, 65, ncol_v), device="cuda")
    q_f = 0.02 * torch.rand((1, 65, ncol_v), device="cuda")
    ms = timed(lambda: V.pressureIntegrationAltitude(p_f, sap_f, sgp_f, t_f, q_f, start_high_p=1), reps=5)
    out["vcoord_pressure_integration_altitude"] = entry(ms, (8.0 + 4.0 * 65 * 4) * ncol_v, 65 * ncol_v,
                                                        "hypsometric integration, 65 levels, p + T + q read, altitude written: 1048 B per column")
    lev_ms = timed(lambda: V.verticalLevelField(V.VCOORD_HYBRID_SIGMA_AP, ap, b, 0.0, sap_f, surface_scale=1.0, post=V.VPOST_STANDARD_ALTITUDE,
                                                topography=sgp_f[0], topography_factor=-1.0 / 9.80665), reps=5)
    out["vcoord_hybrid_pressure_to_height_field"] = entry(lev_ms, (8.0 + 4.0 * 65) * ncol_v, 65 * ncol_v,
                                                          "hybrid pressure -> standard altitude -> minus topography, one pass, 65 levels: 268 B per column")
    del p_f, sap_f, sgp_f, t_f, q_f
    # K12 / K13: forward aggregation of the 2030x1354 swath, 8 levels (level_compat off: every level aggregated), onto a 1000x1000 grid
    # (configs[3]-like, 2.7 sources per cell) and onto a coarse 100x68 grid (20:1 in each direction, ~400 sources per cell)
    gx, gy = np.meshgrid(np.arange(1354), np.arange(2030))
    fin = None
    for tag, (tox, toy) in (("", (1000, 1000)), ("_coarse_20to1", (68, 100))):
        for name, method in (("K12_forward_mean", 6), ("K12_forward_max", 8), ("K13_forward_median", 7)):
            fwd = I.CachedForwardInterpolation("x", "y", method, (gx * (tox / 1354.0)).ravel() - 0.3, (gy * (toy / 2030.0)).ravel() - 0.3, 1354, 2030, tox, toy,
                                               level_compat=False)
            if fin is None or fin.shape[1:] != (fwd.getInY(), fwd.getInX()):
                fin = torch.randn((8, fwd.getInY(), fwd.getInX()), dtype=torch.float32, device="cuda")
            fout = torch.empty((8, toy, tox), dtype=torch.float32, device="cuda")
            ms = timed(lambda: _lib.check(lib.fimexcu_forward_apply_dev_f(fwd._h, fin.data_ptr(), 8, fout.data_ptr(), nan, nan, 0, stream), "fwd"))
            n_in = fwd.getInX() * fwd.getInY()
            out[name + tag] = entry(ms, (4.0 + 4.0 * tox * toy / n_in + 8.0 / 8) * n_in * 8, n_in * 8,
                                    f"2030x1354 swath -> {tox}x{toy}, 8 levels; points = source points")
            if name == "K12_forward_mean" and not tag:  # the same aggregation over a full 65-level slab (one launch)
                fin65 = torch.randn((65, fwd.getInY(), fwd.getInX()), dtype=torch.float32, device="cuda")
                fout65 = torch.empty((65, toy, tox), dtype=torch.float32, device="cuda")
                ms = timed(lambda: _lib.check(lib.fimexcu_forward_apply_dev_f(fwd._h, fin65.data_ptr(), 65, fout65.data_ptr(), nan, nan, 0, stream), "fwd"))
                out["K12_forward_mean_65_levels"] = entry(ms, (4.0 + 4.0 * tox * toy / n_in + 8.0 / 65) * n_in * 65, n_in * 65,
                                                          f"2030x1354 swath -> {tox}x{toy}, 65 levels; points = source points")
                del fin65, fout65
            fwd.close()
    # packed slabs at the ends of a device-resident chain (8(f) rank 1): int16 -> f32 of one source slab, f32 -> int16 of one T1 slab
    try:
        from fimex_b200 import packed as PK
        n_src, n_dst = nz * inx * iny, nz * ox * oy
        raw16 = torch.randint(-30000, 30000, (n_src,), dtype=torch.int16, device="cuda")
        f32o = torch.empty((n_src,), dtype=torch.float32, device="cuda")
        ms = timed(lambda: _lib.check(lib.fimexcu_unpack_dev_f(1, raw16.data_ptr(), n_src, -32767.0, f32o.data_ptr(), stream), "unpack"))
        out["unpack_i16_source_slab"] = entry(ms, 6.0 * n_src, n_src, "int16 -> f32 (fill -> NaN) of one 949x1069x65 slab: 6 B per value")
        f32i = torch.randn((n_dst,), dtype=torch.float32, device="cuda") * 1000.0
        i16o = torch.empty((n_dst,), dtype=torch.int16, device="cuda")
        ms = timed(lambda: _lib.check(lib.fimexcu_pack_dev_f(1, f32i.data_ptr(), n_dst, -32767.0, i16o.data_ptr(), stream), "pack"))
        out["pack_i16_target_slab"] = entry(ms, 6.0 * n_dst, n_dst, "f32 -> int16 (NaN -> fill, lround) of one 2891x1021x65 slab: 6 B per value")
        # CDMTimeInterpolator between two resident T1 slabs (8(f) rank 4): A + f * (B - A), 12 B per value
        f32b = torch.randn((n_dst,), dtype=torch.float32, device="cuda")
        f32c = torch.empty((n_dst,), dtype=torch.float32, device="cuda")
        ms = timed(lambda: _lib.check(lib.fimexcu_get_values_two_level_dev_f(4, f32i.data_ptr(), f32b.data_ptr(), f32c.data_ptr(), n_dst, 0.0, 3.0, 1.0, stream), "blend"))
        out["time_blend_two_target_slabs"] = entry(ms, 12.0 * n_dst, n_dst, "linear_weak_extrapol between two 2891x1021x65 slabs: 12 B per value")
        del raw16, f32o, f32i, i16o, f32b, f32c
    except Exception as exc:
        out["pack_error"] = str(exc)[:200]
    # one-off builds and pre / post-processing of the path (wall clock of the call as a user makes it; no roofline: search- or
    # sweep-bound, not streams)
    def wall(fn, reps=3):
        fn()
        torch.cuda.synchronize()
        best = float("inf")
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            best = min(best, time.perf_counter() - t0)
        return best * 1e3

    try:
        from fimex_b200.coordinates import flannTranslatePointsToClosestInputCell, fastTranslatePointsToClosestInputCell
        slon, slat = S.swath_lonlat()
        tlon, tlat = np.meshgrid(np.arange(0, 20, 0.025), np.arange(58, 70, 0.025))
        maxdist = 0.025 * np.pi / 180 * 6371000.0
        what4 = f"configs[3]: 2030x1354 swath -> {tlon.shape[1]}x{tlon.shape[0]} targets, host arrays in and out, best of 3"
        out["K9_coord_kdtree_translate"] = {"ms": wall(lambda: flannTranslatePointsToClosestInputCell(maxdist, tlon, tlat, slon, slat)), "what": what4}
        out["K10_coord_nearestneighbor_translate"] = {"ms": wall(lambda: fastTranslatePointsToClosestInputCell(tlon, tlat, slon, slat)), "what": what4}
        fpx, fpy = (slon - 0.0) / 0.025, (slat - 58.0) / 0.025
        out["K11_forward_cache_build"] = {"ms": wall(lambda: I.CachedForwardInterpolation("x", "y", 6, fpx, fpy, 1354, 2030, tlon.shape[1], tlon.shape[0]).close()),
                                          "what": "CachedForwardInterpolation ctor, same swath and target (tables from the host), best of 3"}
        del slon, slat, tlon, tlat, fpx, fpy
        lonT, latT = S.target_axes("T1")
        out["K8_vector_rotation_matrix_build"] = {"ms": wall(lambda: RP.get_vector_reproject_matrix(S.MEPS_PROJ, S.LATLON_PROJ, lonT, latT)),
                                                  "what": "get_vector_reproject_matrix for the 2.95 M targets of T1 (host arrays out), best of 3"}
    except Exception as exc:  # keep the bench line even if one of the one-off builds fails
        out["one_off_builds_error"] = str(exc)[:200]
    try:
        yy, xx = np.meshgrid(np.arange(S.MEPS_NY), np.arange(S.MEPS_NX), indexing="ij")
        base = (280 + 15 * np.sin(xx / 90.0) * np.cos(yy / 70.0)).astype(np.float32)
        land = torch.from_numpy(np.sin(xx / 37.0) * np.cos(yy / 53.0) > 0.55).cuda()  # ~20 % undefined blobs
        fld = torch.from_numpy(base).cuda()[None].repeat(nz, 1, 1) + torch.randn((nz, S.MEPS_NY, S.MEPS_NX), device="cuda")
        fld[:, land] = nan
        out["K16_fill2d_65_levels"] = {"ms": wall(lambda: I.mifi_fill2d_f(fld.clone(), 0.01, 1.6, 100), reps=2),
                                       "what": "mifi_fill2d_f on 65 levels of 949x1069 with ~20 % undefined (device slab, incl. a clone of it)"}
        out["K16_creepfill2d_65_levels"] = {"ms": wall(lambda: I.mifi_creepfill2d_f(fld.clone(), 20, 2), reps=2),
                                            "what": "mifi_creepfill2d_f (repeat 20, weight 2), same slab"}
        del fld, land
    except Exception as exc:
        out["fill_error"] = str(exc)[:200]
    # K1+K2+K3: cache build from projection strings (one-off; points/s, no roofline)
    lon, lat = S.target_axes("T1")
    mx, my = S.meps_axes()
    dt = float("inf")
    for _ in range(4):  # the tables are plain cudaMalloc'd (long-lived): single builds vary between 2 and 17 ms with the allocator
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        c = RP.cachedInterpolationFromProjection(1, S.LATLON_PROJ, S.MEPS_PROJ, lon, lat, mx, my)
        torch.cuda.synchronize()
        dt = min(dt, time.perf_counter() - t0)
        c.close()
    out["K1_K2_K3_cache_build_from_projection"] = {"ms": dt * 1e3, "Mpoints_per_s": ox * oy / dt / 1e6,
                                                   "what": "reproject_axes + 2x points2position + reduced domain + tables + row bands, 2.95 M targets, wall clock incl. allocations, best of 4"}
    return out

# ------------------------------------------------------------------------------------------------ B200 arm
def main_b200(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import fimex_b200
    from fimex_b200 import _lib
    lib = fimex_b200.lib

    S, px, py, ox, oy = workload(args.target)
    nz = S.MEPS_NZ
    ci = fimex_b200.CachedInterpolation("x", "y", 1, px, py, S.MEPS_NX, S.MEPS_NY, ox, oy)
    inx, iny = ci.getInX(), ci.getInY()
    rd = ci.reducedDomain()
    n_slabs = 3  # distinct resident inputs, rotated: every step reads 264 MB and writes 767 MB, both far beyond the 126 MB L2
    # HBM layout of resident slabs: rows padded to 32 B (source) / 128 B (target), see DESIGN.md "data layout";
    # FIMEXCU_BENCH_DENSE=1 times the reference's dense [z][y][x] layout instead.
    dense = os.environ.get("FIMEXCU_BENCH_DENSE") == "1"
    slabs = []
    for k in range(n_slabs):
        full = S.meps_slab_torch(t=rank * 1000 + k, device="cuda")
        sub = ci.reduceInput(full, S.MEPS_NX, S.MEPS_NY) if rd else full
        slabs.append(sub if dense else ci.toPadded(sub))
        del full, sub
    ip, il, op, ol = ci.paddedLayout()
    if dense:
        ip, il, op, ol = inx, inx * iny, ox, ox * oy
    outs = [torch.empty((nz, oy, op), dtype=torch.float32, device="cuda") for _ in range(2)]
    stream = _lib.current_stream_ptr()
    points_per_step = nz * ox * oy
    launches = 0

    def step(i):
        _lib.check(lib.fimexcu_cache_apply_dev_strided_f(ci._h, slabs[i % n_slabs].data_ptr(), ip, il, nz, outs[i % 2].data_ptr(), op, ol,
                                                         float("nan"), float("nan"), stream), "apply")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(max(args.warmup, 3)):
        step(i)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for i in range(args.steps):
        step(i)
    ev1.record()
    barrier()
    launches += args.steps
    ms_total = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    value = world * points_per_step / (ms_step * 1e-3) / 1e9

    # roofline of the dominant (only) kernel of the step
    alg_bytes = S.bilinear_algorithmic_bytes(inx * iny, ox * oy, nz)
    peak, peak_src = peaks()
    my_ms = ms_total / args.steps  # this rank's mean launch duration (one launch per step)
    achieved = alg_bytes / (my_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "kernel": "fxcu::staged_apply_tma_kernel<bilinear>" if not os.environ.get("FIMEXCU_NO_TMA") else "fxcu::bilinear_apply_kernel", "algorithmic_bytes_per_launch": alg_bytes, "peak_source": peak_src,
                "bytes_per_point": alg_bytes / points_per_step}
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(prof):
        try:
            roofline["traffic"] = json.load(open(prof)).get("staged_apply_tma_kernel_dram_bytes_per_launch")
        except Exception:
            pass

    # end to end through the host-pointer entry point, exactly the call sequence of the reference-side plug-in
    # (fimex_b200/host/CudaCachedInterpolation.h::interpolateValues): the input is the upstream reader's ordinary pageable array,
    # the result array is allocated INSIDE the call (fimexcu_host_alloc, the pinned pool) and handed back, the caller reads it and
    # drops it (fimexcu_host_free) -- all inside the timed region.  The figure with both buffers pinned and allocated once up
    # front (what round 1 reported) is kept beside it as `pinned_preallocated`.
    e2e = None
    import ctypes as C
    n_in = nz * iny * inx
    in_bytes, out_bytes = n_in * 4, points_per_step * 4
    new_size = C.c_size_t()
    nanf = float("nan")
    if not args.no_e2e:
        dense_in = slabs[0] if dense else slabs[0][:, :, :inx].contiguous()
        pageable_in = dense_in.cpu().numpy()  # malloc'ed, not pinned: what `new float[]` gives the reference
        h_in = lib.fimexcu_host_alloc(in_bytes)
        if not h_in:
            raise RuntimeError("pinned allocation failed: " + _lib.last_error())
        C.memmove(h_in, pageable_in.ctypes.data, in_bytes)
        del dense_in
        e2e_steps = max(2, min(args.steps, 6))
        checksum = [0.0]
        frac_c = C.c_double(1.0)
        n_bands = int(lib.fimexcu_cache_row_bands(ci._h, None, 0, C.byref(frac_c)))
        band_frac = float(frac_c.value) if (n_bands and frac_c.value < 0.9 and os.environ.get("FIMEXCU_E2E_BANDS", "1") != "0") else 1.0

        def plugin_step():
            h_out = lib.fimexcu_host_alloc(out_bytes)
            if not h_out:
                raise RuntimeError("pinned allocation failed: " + _lib.last_error())
            _lib.check(lib.fimexcu_cache_apply_f(ci._h, pageable_in.ctypes.data, n_in, h_out, C.byref(new_size), nanf, nanf), "apply_f")
            res = (C.c_float * 2).from_address(h_out + (out_bytes // 2 // 4) * 4)  # the caller looks at the result it was handed
            checksum[0] += res[0]
            lib.fimexcu_host_free(h_out)

        h_fixed = lib.fimexcu_host_alloc(out_bytes)

        def pinned_step():
            _lib.check(lib.fimexcu_cache_apply_f(ci._h, h_in, n_in, h_fixed, C.byref(new_size), nanf, nanf), "apply_f")

        def timed_host(fn):
            fn()
            barrier()
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                fn()
            torch.cuda.synchronize()
            dt = torch.tensor([(time.perf_counter() - t0) / e2e_steps], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            return float(dt.item())

        dt_plugin = timed_host(plugin_step)
        dt_pinned = timed_host(pinned_step)
        chunks = 8
        launches += 2 * e2e_steps * chunks  # timed e2e steps only (the warm-up calls are not counted)
        e2e = {"value": world * points_per_step / dt_plugin / 1e9, "unit": UNIT, "h2d_bytes_per_step": in_bytes,
               "d2h_bytes_per_step": int(round(out_bytes * band_frac)), "result_bytes_per_step": out_bytes, "ms_per_step": dt_plugin * 1e3,
               "call": "plug-in sequence: fimexcu_host_alloc(result) + fimexcu_cache_apply_f(pageable input -> pinned bounce chunks -> H2D, kernel, "
                       "D2H pipelined over level chunks) + read + fimexcu_host_free, all timed",
               "row_bands": {"bands": n_bands, "fraction_copied": band_frac,
                             "what": "only the row bands of the target that can hold data are copied back (fimexcu_cache_row_bands); the "
                                     "undefined rest of the result array is written by host cores meanwhile"},
               "pinned_preallocated": {"value": world * points_per_step / dt_pinned / 1e9, "unit": UNIT, "ms_per_step": dt_pinned * 1e3,
                                       "call": "fimexcu_cache_apply_f with both host buffers pinned and allocated once, outside the timed region"}}
        lib.fimexcu_host_free(h_in)
        lib.fimexcu_host_free(h_fixed)
        # the same slab as a packed CDM_SHORT variable: raw int16 in, int16 out, unpack / pack fused in the kernel
        try:
            raw = torch.randint(-30000, 30000, (n_in,), dtype=torch.int16, device="cuda").cpu().numpy()  # pageable

            def packed_step():
                p_out = lib.fimexcu_host_alloc(points_per_step * 2)
                if not p_out:
                    raise RuntimeError("pinned allocation failed: " + _lib.last_error())
                _lib.check(lib.fimexcu_cache_apply_i16(ci._h, raw.ctypes.data, n_in, p_out, C.byref(new_size), -32767.0), "apply_i16")
                lib.fimexcu_host_free(p_out)

            dtp = timed_host(packed_step)
            launches += e2e_steps * chunks
            e2e["packed_i16"] = {"value": world * points_per_step / dtp / 1e9, "unit": UNIT, "h2d_bytes_per_step": n_in * 2,
                                 "d2h_bytes_per_step": int(round(points_per_step * 2 * band_frac)), "ms_per_step": dtp * 1e3,
                                 "call": "same plug-in sequence through fimexcu_cache_apply_i16 (CDM_SHORT slab: raw int16 in and out, fill <-> NaN and lround fused)"}
        except Exception as exc:
            e2e["packed_i16"] = {"error": str(exc)[:200]}

    # the whole job of BASELINE configs[1]: 67 time-step slabs dealt t % N over the ranks (the reference's MPI writer split,
    # NetCDF_CDMWriter.cc:607-618), wall of the slowest rank; 67 = 8*8 + 3, so at N = 8 three ranks carry 9 slabs (9/8 imbalance)
    job = None
    if not args.no_job:
        sharding = _load_by_path("sharding")
        mine = sharding.plan(S.MEPS_NT, nz, world)[rank]
        counts = [len(x) for x in sharding.plan(S.MEPS_NT, nz, world)]
        barrier()
        ev0.record()
        for k, (t, z0, z1) in enumerate(mine):
            step(k)
        ev1.record()
        barrier()
        launches += len(mine)
        tj = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tj, op=dist.ReduceOp.MAX)
        job = {"slabs": S.MEPS_NT, "slabs_per_rank": counts, "imbalance": max(counts) * world / S.MEPS_NT,
               "device_resident": {"value": S.MEPS_NT * points_per_step / (float(tj.item()) * 1e-3) / 1e9, "unit": UNIT, "ms": float(tj.item()),
                                   "what": "all 67 slabs, inputs resident in HBM (3 rotating slabs per rank), max over ranks"}}
        # the same with the slabs dealt over two streams, as two of the reference's writer threads would issue them
        # (NetCDF_CDMWriter.cc:599-606): the ramp of one launch fills the tail of the other
        two = [torch.cuda.Stream(), torch.cuda.Stream()]
        cur = torch.cuda.current_stream()
        barrier()
        ev0.record()
        for q in two:
            q.wait_event(ev0)
        for k, (t, z0, z1) in enumerate(mine):
            _lib.check(lib.fimexcu_cache_apply_dev_strided_f(ci._h, slabs[k % n_slabs].data_ptr(), ip, il, nz, outs[k % 2].data_ptr(), op, ol,
                                                             float("nan"), float("nan"), two[k % 2].cuda_stream), "apply")
        for q in two:
            cur.wait_stream(q)
        ev1.record()
        barrier()
        launches += len(mine)
        tj2 = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tj2, op=dist.ReduceOp.MAX)
        job["device_resident_two_streams"] = {"value": S.MEPS_NT * points_per_step / (float(tj2.item()) * 1e-3) / 1e9, "unit": UNIT, "ms": float(tj2.item()),
                                              "what": "the same 67 slabs issued alternately on two streams"}
        if not args.no_e2e:
            n_job = len(mine) if args.full_job else min(len(mine), 6)  # bounded sample of this rank's share unless --full-job
            barrier()
            t0 = time.perf_counter()
            for _ in range(n_job):
                plugin_step()
            torch.cuda.synchronize()
            per_slab = torch.tensor([(time.perf_counter() - t0) / max(n_job, 1)], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(per_slab, op=dist.ReduceOp.MAX)
            launches += n_job * 8
            wall = float(per_slab.item()) * max(counts)
            job["e2e"] = {"value": S.MEPS_NT * points_per_step / wall / 1e9, "unit": UNIT, "ms": wall * 1e3,
                          "what": f"plug-in call sequence per slab; slowest rank's mean slab time over {n_job} of its slabs x its {max(counts)} slabs"}

        # the reference's own process model on the same N GPUs: ONE process (rank 0; the other ranks wait), one cache handle whose
        # tables are replicated onto every GPU (fimexcu_cache_replicate), N host threads -- the writer's OpenMP threads -- each bound
        # to a GPU and pulling slabs t, t + N, ... through the plug-in call sequence
        if world > 1 and not args.no_e2e:
            barrier()
            if rank == 0:
                try:
                    ci.replicate(list(range(world)))
                    per_thread = 4
                    errs = []

                    def writer(d):
                        try:
                            _lib.check(lib.fimexcu_set_device(d), "set_device")
                            ns = C.c_size_t()
                            for _ in range(per_thread + 1 if d == 0 else per_thread):
                                h_out = lib.fimexcu_host_alloc(out_bytes)
                                _lib.check(lib.fimexcu_cache_apply_f(ci._h, pageable_in.ctypes.data, n_in, h_out, C.byref(ns), nanf, nanf), "apply_f")
                                lib.fimexcu_host_free(h_out)
                        except Exception as exc:  # noqa: BLE001
                            errs.append(repr(exc))

                    for timed_round in (False, True):  # first round warms the pools and streams of every (thread, device)
                        ths = [threading.Thread(target=writer, args=(d,)) for d in range(world)]
                        t0 = time.perf_counter()
                        for th in ths:
                            th.start()
                        for th in ths:
                            th.join()
                        wall_sp = time.perf_counter() - t0
                    _lib.check(lib.fimexcu_set_device(local), "set_device")
                    n_sp = world * per_thread + 1
                    job["single_process"] = ({"error": errs[0][:200]} if errs else
                                             {"value": n_sp * points_per_step / wall_sp / 1e9, "unit": UNIT, "ms": wall_sp * 1e3, "slabs": n_sp,
                                              "what": f"one process, one replicated cache handle, {world} host threads bound to {world} GPUs, plug-in call "
                                                      f"sequence; {n_sp} slabs dealt t % N (thread 0 carries one more)"})
                except Exception as exc:  # noqa: BLE001
                    job["single_process"] = {"error": str(exc)[:200]}
            barrier()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            ref = cpu_reference(px, py, S, ox, oy, budget_s=12.0)
            levels = ref["levels"]
            reps = 2 if ref["t_level"] * levels < 6 else 1
            dt = ref["run"](levels, reps)
            cpu = {"value": levels * ox * oy / dt / 1e9, "unit": UNIT, "cores": ref["threads"], "kind": ref["kind"],
                   "sample": f"{levels} of {nz} levels of one S->{args.target} slab x{reps} "
                             f"(CachedInterpolation::interpolateValues, {ref['threads']} OpenMP threads)"}
            try:
                cpu["vertical_hybrid_to_pressure_log"] = cpu_vertical_port()
            except Exception as exc:
                cpu["vertical_hybrid_to_pressure_log"] = {"error": str(exc)[:200]}
        except Exception as exc:  # the checker is optional on the GPU box; the GPU numbers stand on their own
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "unavailable", "sample": str(exc)[:200]}

    kernels = None
    if rank == 0 and world == 1 and not args.no_secondary:
        del slabs, outs
        torch.cuda.empty_cache()
        try:
            kernels = secondary_kernels(S, ci, px, py, ox, oy, peak)
        except Exception as exc:
            kernels = {"error": str(exc)[:300]}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "per_gpu": value / world,
            "config": {"workload": workload_name(S, args.target, ox, oy),
                       "reduced_source_window": [inx, iny], "sharding": f"independent time-step slabs, {world} rank(s), no collective",
                       "l2": "inputs larger than L2: 3 rotating 264 MB source slabs, 767 MB written per step (L2 126 MB)",
                       "hbm_layout": "dense [z][y][x]" if dense else f"rows padded: source pitch {ip} (32 B), target pitch {op} (128 B)"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "job": job, "gpu_launches": launches, "clocks": clocks, "kernels": kernels,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    a = parse_args()
    sys.exit(main_reference(a) if a.impl == "reference" else main_b200(a))
