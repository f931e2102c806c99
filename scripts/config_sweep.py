"""BASELINE.json configs 4 and 5 on ONE GPU (the multi-GPU forms shard these bands): the subsequence-length sweep
m = 32..4096 at n = 2^22 and, with --big, n = 2^24, m = 1024.  Random walk, fp64; one timed step per point.
Size-independent checks on every result: finite distances, indices outside the exclusion zone, symmetry of the best
pair (MPI[MPI[i]] reaches a correlation >= that of i)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))
import numpy as np
import mpb200 as mp

ctx = mp.Context(0)
points = [(1 << 22, m) for m in (32, 64, 128, 256, 512, 1024, 2048, 4096)]
if "--big" in sys.argv:
    points.append((1 << 24, 1024))
for n, m in points:
    T = np.cumsum(np.random.default_rng(1234).standard_normal(n))
    plan = mp.Plan(ctx, n, m, mp.F64, 0)
    plan.upload_series(T)
    plan.precompute(); plan.compute(); plan.finalize()
    MP, MPI, corr = plan.download(); t = plan.timings()
    idx = np.arange(MPI.size)
    ok = bool(np.isfinite(MP).all() and (MPI >= 0).all() and (np.abs(MPI - idx) >= m // 4).all() and (corr[MPI] >= corr - 1e-12).all())
    print(json.dumps({"n": n, "m": m, "data": "walk", "dtype": "f64", "compute_ms": round(t["compute_ms"], 2), "pre_ms": round(t["precompute_ms"], 3),
                      "cells_per_s": t["cells"] / (t["compute_ms"] * 1e-3), "seg_rows": t["segment_rows"], "items": t["items"], "replays": t["replays"],
                      "properties_ok": ok, "mp_min": float(MP.min()), "mp_max": float(MP.max())}), flush=True)
    plan.close()
ctx.close()
