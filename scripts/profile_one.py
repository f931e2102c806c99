"""One whole-problem run (default n=2^20, m=256, fp64, uniform) for ncu captures: 2 steps, prints phase times."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))
import numpy as np
import mpb200 as mp

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
m = int(sys.argv[2]) if len(sys.argv) > 2 else 256
dt = mp.F32 if len(sys.argv) > 3 and sys.argv[3] == "f32" else mp.F64
kind = sys.argv[4] if len(sys.argv) > 4 else "uniform"
rng = np.random.default_rng(1234)
T = rng.random(n) if kind == "uniform" else np.cumsum(rng.standard_normal(n))
ctx = mp.Context(0)
plan = mp.Plan(ctx, n, m, dt, 0)
plan.upload_series(T)
for rep in range(2):
    plan.precompute(); plan.compute(); plan.finalize()
    MP, MPI, corr = plan.download()
    t = plan.timings()
    print(json.dumps({"n": n, "m": m, "pre_ms": t["precompute_ms"], "compute_ms": t["compute_ms"], "fin_ms": t["finalize_ms"],
                      "cells_per_s": t["cells"] / (t["compute_ms"] * 1e-3), "slow_updates": t["slow_updates"]}), flush=True)
plan.close()
