"""Quick device-side timing of the three kernels at a few sizes (development aid, not the bench)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import mpb200 as mp
from oracle_py import random_walk, uniform_series

ctx = mp.Context(0)
for dt, name in ((mp.F64, "f64"), (mp.F32, "f32")):
    r, ms = ctx.measure_fma_peak(dt, 1 << 15)
    print(f"fma peak {name}: {r:.4e} FMA/s ({ms:.2f} ms)", flush=True)
sizes = [int(a) for a in sys.argv[1:]] or [1 << 17, 1 << 18, 1 << 20]
for n in sizes:
    for gen, gname in ((uniform_series, "uniform"), (random_walk, "walk")):
        for dt, flags, label in ((mp.F64, 0, "f64"), (mp.F64, 1, "f64-strict"), (mp.F32, 0, "f32")):
            T = gen(n, 1234)
            plan = mp.Plan(ctx, n, 256, dt, flags)
            plan.upload_series(T)
            for rep in range(2):
                plan.precompute(); plan.compute(); plan.finalize()
                MP, MPI, corr = plan.download(); t = plan.timings()
            cps = t["cells"] / (t["compute_ms"] * 1e-3)
            print(json.dumps({"n": n, "data": gname, "mode": label, "pre_ms": round(t["precompute_ms"], 3), "compute_ms": round(t["compute_ms"], 3),
                              "fin_ms": round(t["finalize_ms"], 3), "cells_per_s": cps, "slow_updates": t["slow_updates"], "bands": t["bands"], "items": t["items"], "seg_rows": t["segment_rows"]}), flush=True)
            plan.close()
