"""Histogram of how many lanes of a warp pass the block filter when a block is replayed (library built with -DMPB_REPLAY_HIST).
usage: MPB200_LIB=.../libmpb200_hist.so python scripts/r02_replay_hist.py [n] [m]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))
import numpy as np
import mpb200 as mp
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
m = int(sys.argv[2]) if len(sys.argv) > 2 else 256
ctx = mp.Context(0)
for kind in ("walk", "uniform"):
    rng = np.random.default_rng(1234)
    T = rng.random(n) if kind == "uniform" else np.cumsum(rng.standard_normal(n))
    plan = mp.Plan(ctx, n, m, mp.F64, 0)
    plan.upload_series(T); plan.precompute(); plan.compute(); plan.finalize(); plan.download()
    h = plan.replay_histogram(); t = plan.timings()
    tot = sum(h)
    cum = np.cumsum(h) / max(tot, 1)
    print(json.dumps({"data": kind, "n": n, "m": m, "replays": t["replays"], "hist": h, "cum_le_1_2_4_8_16": [round(float(cum[k]), 3) for k in (1, 2, 4, 8, 16)],
                      "mean_lanes": sum(k * v for k, v in enumerate(h)) / max(tot, 1)}))
    plan.close()
