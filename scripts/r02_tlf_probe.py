"""One MatrixProfileKernelTLF chunk at the reference's README size (n = 2^20, m = 256, w = 4096) on device buffers: device time
and processed cells per second of the chunk-compatible kernel (the drop-in path; the fast path is mpb_plan_compute)."""
import ctypes as C, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))
import numpy as np
import mpb200 as mp
n, m, w = 1 << 20, 256, 4096
T = np.cumsum(np.random.default_rng(1234).standard_normal(n))
ctx = mp.Context(0)
plan = mp.Plan(ctx, n, m, mp.F64, mp.FLAG_STRICT)
plan.upload_series(T); plan.precompute()
inp = plan.download_input(); plan.close()
sublen = n - m + 1
bin_, bout = ctx.make_buffer(inp.nbytes), ctx.make_buffer(sublen * 32)
bin_.copy_from_host(inp); ctx.finish()
L = mp.lib()
for it in (0, 128, 255):
    for rep in range(2):
        ctx.finish()
        rc = L.mpb_kernel_tlf_enqueue(ctx._h, C.c_int(mp.F64), C.c_int(n), C.c_int(m), C.c_int(w), C.c_int(it), bin_._h, bout._h)
        assert rc == 0, L.mpb_last_error()
        ns = ctx.finish()
    rows = sublen - it * w
    print(json.dumps({"n": n, "m": m, "w": w, "iteration": it, "rows": rows, "ms": ns * 1e-6, "processed_cells_per_s": rows * w / (ns * 1e-9)}), flush=True)
