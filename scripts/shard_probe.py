"""Runs every rank's band shard of a world-size-W job one after the other on ONE GPU and prints its tile-kernel time
and filter statistics (load-balance probe; no collectives)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))
import numpy as np
import mpb200 as mp
W = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
kind = sys.argv[3] if len(sys.argv) > 3 else "uniform"
rng = np.random.default_rng(1234)
T = rng.random(n) if kind == "uniform" else np.cumsum(rng.standard_normal(n))
ctx = mp.Context(0)
for r in range(W):
    plan = mp.Plan(ctx, n, 256, mp.F64, 0); plan.set_shard(r, W); plan.upload_series(T)
    for rep in range(2):
        plan.precompute(); plan.compute(); plan.finalize(); plan.download(); t = plan.timings()
    print(json.dumps({"rank": r, "world": W, "compute_ms": round(t["compute_ms"], 3), "items": t["items"], "bands": t["bands"], "hits": t["filter_hits"],
                      "wins": t["slow_updates"], "replays": t["replays"]}), flush=True)
    plan.close()
