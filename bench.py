#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 matrix-profile engine (BASELINE.json metric).

Metric: useful matrix-profile cell updates per second, cells = L(L+1)/2, L = (n-m+1) - m/4.
Workload (config.workload): BASELINE.json configs[1] -- double, n = 1048576, m = 256 on a
data/binary/1048576.tsb-shaped synthetic series (i.i.d. U[0,1) float64), fixed total work whatever the
GPU count ("strong" scaling: the diagonal bands are dealt round-robin over the ranks and the per-rank
profiles are max-merged over NCCL).

A step = one whole matrix profile: precompute kernel + diagonal-band kernel [+ NCCL merge] + finalize.
  value  : series already resident in HBM when the step starts (device time, CUDA events, max over ranks)
  e2e    : the same through the C ABI's one-call host-buffer entry (mpb_compute / plan upload+download):
           pinned host series -> H2D -> kernels -> D2H of profile + index, every step.
  roofline: the diagonal-band kernel against the FP64 pipe issue peak measured in this run (DFMA
           microbenchmark), 4 FP64-pipe ops per useful cell; HBM traffic is
           reported beside it as evidence that HBM is not the limiter.
  cpu_baseline / --impl reference: the reference's own Vanilla kernel sources (oracle/_ref, compiled
           from /root/reference at build time) on the host cores, on a bounded sample of diagonal chunks.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))

METRIC = "matrix-profile cell updates/sec"
UNIT = "cells/s"
_CAL = {}
REF_W = 4096          # MP_SIZE_W of the reference's README configuration (README.md:35)


def synth_series(n, kind, seed=1234):
    import numpy as np
    rng = np.random.default_rng(seed)
    return rng.random(n) if kind == "uniform" else np.cumsum(rng.standard_normal(n))


def useful_cells(n, m):
    L = (n - m + 1) - m // 4
    return L * (L + 1) // 2


def chunk_useful_cells(n, m, w, it):
    """Useful cells of diagonal chunk `it` (diagonals [it*w, (it+1)*w), exclusion zone and the
    out-of-range triangle removed)."""
    sublen, excl = n - m + 1, m // 4
    lo, hi = max(it * w, excl), min((it + 1) * w, sublen)
    return sum(sublen - d for d in range(lo, hi)) if hi > lo else 0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.proc, self.lines = device, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(2)
        sm, smax, power, reasons = [], None, [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax = float(f[2]); power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": smax, "power_w_max": max(power) if power else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------------------------------
# CPU baseline: the reference's own kernel sources on the host cores (bounded sample of chunks)
# --------------------------------------------------------------------------------------------------
def cpu_reference_sample(n, m, kind, target_s=15.0, threads=None, est_rate=None):
    """Times the reference's Vanilla kernel (kernel/MatrixProfileKernelVanilla.cpp, compiled unmodified into
    oracle/_ref/libmpref_fast.so with -O3 -march=native; falls back to the C restatement when the reference
    was not compiled here) on one diagonal chunk of width w = 4096 per host thread, chunks chosen so that
    each costs about target_s seconds.  Returns the cpu_baseline object."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "oracle"))      # the oracle: bench's cpu_baseline leg may execute it
    from concurrent.futures import ThreadPoolExecutor
    from oracle_py import Oracle, RefShim
    threads = threads or os.cpu_count() or 1
    sublen = n - m + 1
    w = REF_W if n >= 4 * REF_W else 256
    nit = (n - m + w) // w
    if est_rate is None:
        est_rate = _CAL.get("rate", 2.0e8)             # cells/s per thread; refined after every sample
    want_rows = max(w, int(target_s * est_rate / w))
    it0 = max(0, min(nit - threads, (sublen - want_rows) // w))
    its = [it for it in range(it0, min(nit, it0 + threads))]
    T = synth_series(n, kind)
    orc = Oracle(fast=True)
    inp = orc.precompute(T, m)
    use_ref = RefShim.available(fast=True)
    eng = RefShim(fast=True) if use_ref else orc

    def one(it):
        return eng.kernel("vanilla", n, m, w, it, inp).size     # ctypes releases the GIL during the call

    t0 = time.perf_counter()
    with ThreadPoolExecutor(len(its)) as ex:
        list(ex.map(one, its))
    dt = time.perf_counter() - t0
    cells = sum(chunk_useful_cells(n, m, w, it) for it in its)
    _CAL["rate"] = cells / dt / len(its)
    return {"value": cells / dt, "unit": UNIT, "cores": len(its), "kind": "reference" if use_ref else "port",
            "sample": f"Vanilla kernel, double, w={w}: diagonal chunks {its[0]}..{its[-1]} of {nit} of the same n={n}, m={m} series, "
                      f"one chunk per host thread, {cells:.4g} useful cells in {dt:.1f} s (-O3 -march=native)",
            "seconds": dt}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    vals = []
    base = None
    for i in range(args.warmup + args.steps):
        # warm-up steps are short samples (page the library and the series in); timed ones are the bounded sample
        base = cpu_reference_sample(args.n, args.m, args.data, target_s=2.0 if i < args.warmup else args.ref_seconds)
        if i >= args.warmup:
            vals.append(base)
    cells_s = sum(b["value"] * b["seconds"] for b in vals) / sum(b["seconds"] for b in vals)
    ms = 1e3 * sum(b["seconds"] for b in vals) / len(vals)
    base = dict(vals[-1]); base["value"] = cells_s
    line = {"impl": "reference", "metric": METRIC, "value": cells_s, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args), "cpu_baseline": base,
            "e2e": {"value": cells_s, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def workload_config(args):
    return {"workload": f"Tiled-kernel config, {'double' if args.dtype == 'f64' else 'float'}, n={args.n}, m={args.m}, "
                        f"{'tsb-shaped U[0,1)' if args.data == 'uniform' else 'random-walk'} synthetic series (BASELINE.json configs[1])",
            "n": args.n, "m": args.m, "series": args.data, "partition": f"diagonal bands round-robin over {args.gpus} GPU(s)",
            "l2": "flushed between timed steps (256 MiB write)"}


# --------------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------------
def run_gpu_arm(args):
    import numpy as np
    import torch
    import mpb200 as mp

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, m = args.n, args.m
    dtype = mp.F64 if args.dtype == "f64" else mp.F32
    sublen = n - m + 1
    cells = useful_cells(n, m)

    ctx = mp.Context(local)
    stream = torch.cuda.ExternalStream(ctx.stream, device=local)
    plan = mp.Plan(ctx, n, m, dtype, 0)
    plan.set_shard(rank, world)
    T_pinned = torch.from_numpy(synth_series(n, args.data)).pin_memory()
    MP_pinned = torch.empty(sublen, dtype=torch.float64).pin_memory()
    MPI_pinned = torch.empty(sublen, dtype=torch.int32).pin_memory()
    plan.upload_series_ptr(T_pinned.data_ptr())
    ctx.finish()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")     # > 126 MB of L2
    corr_buf = torch.empty(sublen, dtype=torch.float64, device="cuda")
    index_buf = torch.empty(sublen, dtype=torch.int32, device="cuda")
    if world > 1:
        from mpb200.dist import compute_merged
        with torch.cuda.stream(stream):
            dist.all_reduce(index_buf)          # communicator creation stays outside the timed region
        ctx.finish()

    def step(e2e):
        """One whole matrix profile on this rank's band shard; returns (t0, t1, launches) events on the engine stream."""
        with torch.cuda.stream(stream):
            flush.zero_()
            t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
            t0.record(stream)
            if e2e:
                plan.upload_series_ptr(T_pinned.data_ptr())
            plan.precompute()
            extra = 0
            if world > 1:
                extra = compute_merged(plan, corr_buf, index_buf, world)   # first bands, merge, remaining bands, merge
            else:
                plan.compute()
            plan.finalize()
            if e2e:
                plan.download_into(MP_pinned.data_ptr(), MPI_pinned.data_ptr())
            t1.record(stream)
        return t0, t1, extra

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(e2e, steps, warmup):
        for _ in range(warmup):
            step(e2e)
        sync_all()
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        evs, kern_ms, launches = [], [], 0
        for _ in range(steps):
            t0, t1, extra = step(e2e)
            torch.cuda.synchronize()
            tm = plan.timings()
            kern_ms.append(tm["compute_ms"])
            launches += tm["launches"]
            evs.append(t0.elapsed_time(t1))
            if world > 1:
                dist.barrier()
        sync_all()
        clocks = sampler.stop() if rank == 0 else None
        total = torch.tensor([sum(evs)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(total, op=dist.ReduceOp.MAX)
        return total.item() / steps, sum(kern_ms) / len(kern_ms), launches, clocks, tm

    ms_dev, band_ms, launches, clocks, tm = timed(False, args.steps, args.warmup)
    ms_e2e, _, _, _, _ = timed(True, args.steps, max(1, args.warmup // 2))
    # Secondary workload, reported beside the headline (not the headline): the same n and m on a random walk
    # (BASELINE.json north_star / configs[2..3]); smooth series keep far more cells near the running best.
    other = "walk" if args.data == "uniform" else "uniform"
    T_other = torch.from_numpy(synth_series(n, other)).pin_memory()
    plan.upload_series_ptr(T_other.data_ptr())
    ctx.finish()
    ms_other, band_other, _, _, tm_other = timed(False, max(2, args.steps // 2), 1)
    plan.upload_series_ptr(T_pinned.data_ptr())
    ctx.finish()

    fma_peak, _ = ctx.measure_fma_peak(dtype, 1 << 15)
    fma_peak, _ = ctx.measure_fma_peak(dtype, 1 << 16)
    out = None
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        my_cells = cells / world                                   # round-robin bands: equal shares to within one band
        ops = 4.0 * my_cells / (band_ms * 1e-3)
        prof_path = os.path.join(ROOT, "profiles", "tile_kernel_traffic.json")
        traffic = None
        if os.path.exists(prof_path):
            traffic = json.load(open(prof_path)).get(f"{args.dtype}_n{n}_m{m}")
        hbm_alg = 8.0 * n + 12.0 * sublen                          # read T once, write MP + MPI once
        out = {"metric": METRIC, "value": cells / (ms_dev * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": args.dtype,
               "data": "synthetic", "config": workload_config(args),
               "e2e": {"value": cells / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e, "h2d_bytes_per_step": 8 * n, "d2h_bytes_per_step": 12 * sublen},
               "gpu_launches": launches,
               "roofline": {"bound": "fp64_fma_pipe" if args.dtype == "f64" else "fp32_fma_pipe", "kernel": "tile_kernel",
                            "achieved": ops / 1e12, "peak": fma_peak / 1e12, "unit": "T pipe-ops/s (4 per cell: 2 FMA + 2 MUL)",
                            "frac": ops / fma_peak, "peak_source": "DFMA/FFMA issue-rate microbenchmark measured in this run (mpb_measure_fma_peak)",
                            "kernel_ms": band_ms, "kernel_share_of_step": band_ms / ms_dev, "traffic": traffic,
                            "hbm": {"algorithmic_bytes_per_step": hbm_alg, "achieved_gbs": hbm_alg / (band_ms * 1e-3) / 1e9,
                                    "peak_gbs": peaks.get("hbm_gbs", 6650.0), "peak_source": "measured" if peaks else "fallback"}},
               "clocks": clocks, "phases_ms": {k: tm[k] for k in ("precompute_ms", "compute_ms", "finalize_ms")},
               "slow_updates": tm["slow_updates"], "filter_hits": tm["filter_hits"], "replays": tm["replays"], "bands_rank0": tm["bands"],
               "secondary": {"series": other, "value": cells / (ms_other * 1e-3), "unit": UNIT, "ms_per_step": ms_other,
                             "roofline_frac": 4.0 * my_cells / (band_other * 1e-3) / fma_peak, "slow_updates": tm_other["slow_updates"],
                             "filter_hits": tm_other["filter_hits"], "replays": tm_other["replays"]}}
        out["roofline"]["executed_pipe_ops_per_cell"] = 2.03     # 2 DFMA per cell + 2 DMUL per 64-cell block (the filter never forms P)
    plan.close()
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu_reference_sample(n, m, args.data, target_s=1.5)          # calibrates the per-thread rate (untimed warm-up)
        cb = cpu_reference_sample(n, m, args.data, target_s=args.ref_seconds)
        cb.pop("seconds", None)
        out["cpu_baseline"] = cb
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n", type=int, default=1 << 20)
    ap.add_argument("--m", type=int, default=256)
    ap.add_argument("--dtype", default="f64", choices=["f64", "f32"])
    ap.add_argument("--data", default="uniform", choices=["uniform", "walk"])
    ap.add_argument("--ref-seconds", type=float, default=12.0, help="CPU seconds per reference sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
