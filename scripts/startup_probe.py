"""Start-up cost of a step: device time of the first local bands one by one (sketch on/off), for a 1-GPU plan and for one
rank of an 8-way job.  usage: python scripts/startup_probe.py [n] [kind]"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))
import numpy as np
import mpb200 as mp

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
kind = sys.argv[2] if len(sys.argv) > 2 else "uniform"
rng = np.random.default_rng(1234)
T = rng.random(n) if kind == "uniform" else np.cumsum(rng.standard_normal(n))
ctx = mp.Context(0)
cuts = [0, 1, 2, 3, 4, 6, 8, 12, 16, 24, 32, 48, 64]
for world, rank in ((1, 0), (8, 3)):
    for sk in (0, 1):
        p = mp.Plan(ctx, n, 256, mp.F64, 0); p.set_shard(rank, world); p.set_sketch_bands(sk)
        p.upload_series(T)
        for rep in range(2):
            p.precompute(); ctx.finish()
            ms = []
            for a, b in zip(cuts[:-1], cuts[1:]):
                p.compute_bands(a, b); ms.append(round(ctx.finish() * 1e-6, 3))
            p.compute_bands(cuts[-1], -1); rest = round(ctx.finish() * 1e-6, 3)
        t = p.timings()
        print(json.dumps({"world": world, "rank": rank, "sketch": sk, "data": kind, "cuts": cuts, "ms": ms, "rest_ms": rest, "total_ms": round(sum(ms) + rest, 3),
                          "local_bands": p.local_bands(), "replays": t["replays"], "hits": t["filter_hits"], "wins": t["slow_updates"]}), flush=True)
        p.close()
ctx.close()
