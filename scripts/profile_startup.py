"""ncu target: the first 8 local bands of rank 3 of an 8-way job (the threshold-discovery phase), twice."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))
import numpy as np
import mpb200 as mp
n = 1 << 20
T = np.random.default_rng(1234).random(n)
ctx = mp.Context(0)
p = mp.Plan(ctx, n, 256, mp.F64, 0); p.set_shard(3, 8); p.upload_series(T)
for rep in range(2):
    p.precompute(); ctx.finish(); p.compute_bands(0, 8); print(ctx.finish() * 1e-6)
p.close(); ctx.close()
