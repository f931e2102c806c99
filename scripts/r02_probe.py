"""Round-2 development probe: device timings and accuracy (vs MPB_FLAG_STRICT = the reference's bits) of the fp64 modes
at one size.  usage: r02_probe.py [n] [m] [modes...]   modes: default carryoff old seqnoreseed f32"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))
import numpy as np
import mpb200 as mp

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
m = int(sys.argv[2]) if len(sys.argv) > 2 else 256
modes = sys.argv[3:] or ["default", "old", "f32"]
MODES = {"default": (mp.F64, 0), "carryoff": (mp.F64, mp.FLAG_NO_CHAIN_CARRY), "old": (mp.F64, mp.FLAG_NO_CHAIN_CARRY | mp.FLAG_WINDOW_MEAN),
         "seqnoreseed": (mp.F64, mp.FLAG_SEQ_MEAN | mp.FLAG_NO_RESEED), "f32": (mp.F32, 0), "strict": (mp.F64, mp.FLAG_STRICT)}
ctx = mp.Context(0)
peak, _ = ctx.measure_fma_peak(mp.F64, 1 << 15)
for kind in ("walk", "uniform"):
    rng = np.random.default_rng(1234)
    T = rng.random(n) if kind == "uniform" else np.cumsum(rng.standard_normal(n))
    ref = None
    if n <= (1 << 20) and os.environ.get("R02_NO_STRICT") is None:
        plan = mp.Plan(ctx, n, m, mp.F64, mp.FLAG_STRICT)
        plan.upload_series(T); plan.precompute(); plan.compute(); plan.finalize()
        ref = plan.download(); ts = plan.timings(); plan.close()
        print(json.dumps({"n": n, "m": m, "data": kind, "mode": "strict", "pre_ms": ts["precompute_ms"], "compute_ms": ts["compute_ms"]}), flush=True)
    for mode in modes:
        dt, flags = MODES[mode]
        plan = mp.Plan(ctx, n, m, dt, flags)
        if os.environ.get("R02_MARGIN"):
            plan.set_chain_margin(int(os.environ["R02_MARGIN"]))
        plan.upload_series(T)
        best = None
        for rep in range(3):
            plan.precompute(); plan.compute(); plan.finalize()
            MP, MPI, corr = plan.download(); t = plan.timings()
            if best is None or t["compute_ms"] < best["compute_ms"]:
                best = t
        t = best
        out = {"n": n, "m": m, "data": kind, "mode": mode, "pre_ms": round(t["precompute_ms"], 3), "compute_ms": round(t["compute_ms"], 3),
               "fin_ms": round(t["finalize_ms"], 3), "cells_per_s": t["cells"] / ((t["compute_ms"] + t["precompute_ms"]) * 1e-3),
               "frac4": 4 * t["cells"] / (t["compute_ms"] * 1e-3) / peak, "slow_updates": t["slow_updates"], "replays": t["replays"],
               "filter_hits": t["filter_hits"], "items": t["items"], "seg_rows": t["segment_rows"], "launches": t["launches"],
               "chain_candidates": t["chain_candidates"], "chain_max_correction": t["chain_max_correction"], "chain_bound": t["chain_bound"], "chain_margin": t["chain_margin"]}
        if ref is not None:
            rel = np.abs(corr - ref[2]) / ref[2]
            dist_ok = (1 - ref[2]) > 1e-6
            out["max_rel_corr_vs_strict"] = float(rel.max())
            out["max_rel_dist_vs_strict"] = float((np.abs(MP - ref[0])[dist_ok] / ref[0][dist_ok]).max())
            out["index_mismatches"] = int((MPI != ref[1]).sum())
        print(json.dumps(out), flush=True)
        plan.close()
