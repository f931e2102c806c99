"""Tuning aid: times the tile kernel of several library builds (csrc/Makefile VARIANT=...) on the same box.
usage: python scripts/variant_perf.py [variant ...]   ("" or "default" = libmpb200.so); sizes via MPB_SIZES=n1,n2"""
import os, sys, json, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "matrix-profile-hls_b200")
CHILD = r'''
import os, sys, json
sys.path.insert(0, os.path.join(%r, "python"))
import numpy as np
import mpb200 as mp
ctx = mp.Context(0)
sizes = [int(x) for x in os.environ.get("MPB_SIZES", "1048576").split(",")]
for n in sizes:
    for kind in ("uniform", "walk"):
        for dt, name in ((mp.F64, "f64"), (mp.F32, "f32")):
            if name == "f32" and n != sizes[-1]: continue
            rng = np.random.default_rng(1234)
            T = rng.random(n) if kind == "uniform" else np.cumsum(rng.standard_normal(n))
            plan = mp.Plan(ctx, n, 256, dt, 0)
            if os.environ.get("MPB_SEG"): plan.set_segment_rows(int(os.environ["MPB_SEG"]))
            plan.upload_series(T)
            best = None
            for rep in range(3):
                plan.precompute(); plan.compute(); plan.finalize()
                MP, MPI, corr = plan.download(); t = plan.timings()
                best = t if best is None or t["compute_ms"] < best["compute_ms"] else best
            print(json.dumps({"variant": os.environ.get("MPB_VARIANT"), "n": n, "data": kind, "dtype": name, "compute_ms": round(best["compute_ms"], 3),
                              "pre_ms": round(best["precompute_ms"], 3), "cells_per_s": best["cells"] / (best["compute_ms"] * 1e-3),
                              "hits": best["filter_hits"], "wins": best["slow_updates"], "replays": best["replays"], "queued": best["replay_lane_blocks"], "items": best["items"], "seg_rows": best["segment_rows"],
                              "mp_sum": float(MP.sum()), "mpi_sum": int(MPI.astype(np.int64).sum())}), flush=True)
            plan.close()
''' % PKG
for v in (sys.argv[1:] or ["default"]):
    env = dict(os.environ)
    env["MPB_VARIANT"] = v
    if v not in ("", "default"):
        env["MPB200_LIB"] = os.path.join(PKG, f"libmpb200_{v}.so")
    r = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True, timeout=1200)
    sys.stdout.write(r.stdout)
    if r.returncode != 0:
        sys.stdout.write(json.dumps({"variant": v, "error": r.stderr[-2000:]}) + "\n")
    sys.stdout.flush()
