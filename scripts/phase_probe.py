"""Emulates a world-size-W job on ONE GPU: W plans (one per virtual rank) walk their band shards phase by phase, the
profiles are max-merged between phases (mpb_plan_merge_peers on the same device), and the job time is taken as
sum over phases of the slowest rank's tile-kernel time (what a barrier per merge would cost).  Evaluates merge schedules
(how often the ranks must share their profiles for the filter thresholds to stay tight) without spending 8-GPU box time.
usage: python scripts/phase_probe.py W n kind sched [sched ...]   sched = comma list of LOCAL band breakpoints, e.g. 1,6,-1
       or "geo:<first>:<ratio>" (breakpoints first, first*ratio, ...), or "every:<k>"."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))
import numpy as np
import mpb200 as mp

W = int(sys.argv[1]); n = int(sys.argv[2]); kind = sys.argv[3]; scheds = sys.argv[4:]
m = 256
rng = np.random.default_rng(1234)
T = rng.random(n) if kind == "uniform" else np.cumsum(rng.standard_normal(n))
ctx = mp.Context(0)
plans = []
for r in range(W):
    p = mp.Plan(ctx, n, m, mp.F64, 0); p.set_shard(r, W); p.set_sketch_bands(int(os.environ.get("MPB_SKETCH", "-1")))
    p.upload_series(T); plans.append(p)
    if os.environ.get("MPB_SEG"): p.set_segment_rows(int(os.environ["MPB_SEG"]))
nlb = min(p.local_bands() for p in plans)


def breakpoints(s):
    if s == "auto":                       # the schedule mpb_plan_compute_merged uses without peer access (a function of the JOB's band count)
        from mpb200.dist import merge_points
        return merge_points(sum(p.local_bands() for p in plans), W) + [-1]
    if s.startswith("geo:"):
        _, first, ratio = s.split(":"); b, out = float(first), []
        while b < nlb: out.append(int(round(b))); b *= float(ratio)
        return sorted(set(out)) + [-1]
    if s.startswith("every:"):
        k = int(s.split(":")[1]); return list(range(k, nlb, k)) + [-1]
    return [int(x) for x in s.split(",")]


ref = None
for s in scheds:
    bps = breakpoints(s)
    for rep in range(2):
        for p in plans: p.precompute()
        ctx.finish()
        phase_ms, lo = [], 0
        for bp in bps:
            ts = []
            for p in plans:
                p.compute_bands(lo, bp); ts.append(ctx.finish() * 1e-6)
            phase_ms.append(ts)
            if W > 1:
                plans[0].merge_peers(plans[1:])
                for p in plans[1:]: p.merge_peers([plans[0]])
                ctx.finish()
            lo = bp
    plans[0].finalize(); MP, MPI, corr = plans[0].download()
    if ref is None: ref = (MP, MPI)
    stats = [p.timings() for p in plans]
    tot = sum(max(ts) for ts in phase_ms)
    print(json.dumps({"sketch": int(os.environ.get("MPB_SKETCH", "-1")), "breakpoints": bps, "world": W, "n": n, "data": kind, "sched": s, "phases": len(bps), "sum_of_phase_max_ms": round(tot, 3),
                      "sum_of_phase_mean_ms": round(sum(sum(ts) / len(ts) for ts in phase_ms), 3),
                      "phase_max_ms": [round(max(ts), 3) for ts in phase_ms][:12],
                      "replays_rank0": stats[0]["replays"], "replays_max": max(t["replays"] for t in stats),
                      "parked_max": max(t["filter_hits"] for t in stats), "same_result": bool(np.array_equal(MP, ref[0]) and np.array_equal(MPI, ref[1]))}), flush=True)
for p in plans: p.close()
ctx.close()
