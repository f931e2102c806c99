"""Fast path (window means + re-seeded segments + nested FMA) against the chain-faithful mode (MPB_FLAG_STRICT,
bit-identical to the reference kernels -- pinned by tests/golden) at sizes the CPU oracle cannot reach.
Prints max relative deviation of the correlation and of the distance, and index mismatches."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))
import numpy as np
import mpb200 as mp

ctx = mp.Context(0)
m = 256
sizes = [int(a) for a in sys.argv[1:]] or [1 << 16, 1 << 18, 1 << 20]


def run(T, dtype, flags, seg=None):
    plan = mp.Plan(ctx, T.size, m, dtype, flags)
    if seg is not None:
        plan.set_segment_rows(seg)
    plan.upload_series(T)
    plan.precompute(); plan.compute(); plan.finalize()
    MP, MPI, corr = plan.download()
    t = plan.timings()
    plan.close()
    return MP, MPI, corr, t


for n in sizes:
    for kind in ("uniform", "walk"):
        rng = np.random.default_rng(1234)
        T = rng.random(n) if kind == "uniform" else np.cumsum(rng.standard_normal(n))
        MP0, MPI0, c0, t0 = run(T, mp.F64, mp.FLAG_STRICT)
        for label, dtype, flags, seg in (("fast", mp.F64, 0, None), ("fast-seg1024", mp.F64, 0, 1024), ("noreseed", mp.F64, 4, None),
                                         ("seqmean-noreseed", mp.F64, 6, None), ("f32-fast", mp.F32, 0, None), ("f32-noreseed", mp.F32, 4, None)):
            MP, MPI, c, t = run(T, dtype, flags, seg)
            ok = (1 - c0) > 1e-6
            print(json.dumps({"n": n, "data": kind, "mode": label, "seg_rows": t["segment_rows"], "items": t["items"],
                              "max_abs_dcorr": float(np.abs(c - c0).max()), "max_rel_dcorr": float((np.abs(c - c0) / c0).max()),
                              "max_rel_ddist": float((np.abs(MP - MP0)[ok] / MP0[ok]).max()), "idx_mismatch": int((MPI != MPI0).sum()),
                              "min_corr": float(c0.min()), "compute_ms": round(t["compute_ms"], 3), "strict_ms": round(t0["compute_ms"], 3),
                              "cells_per_s": t["cells"] / (t["compute_ms"] * 1e-3), "slow": t["slow_updates"]}), flush=True)
