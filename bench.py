#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 matrix-profile engine (BASELINE.json metric).

Metric: useful matrix-profile cell updates per second, cells = L(L+1)/2, L = (n-m+1) - m/4.
Workload (config.workload): double, n = 1048576, m = 256 on a synthetic RANDOM WALK (BASELINE.json north_star: "throughput on
synthetic random-walk series of the named n and m"), fixed total work whatever the GPU count ("strong" scaling: the diagonal
bands are dealt round-robin over the ranks and the per-rank profiles are max-merged over NCCL inside libmpb200.so).
The tsb-shaped series of BASELINE.json configs[1] (i.i.d. U[0,1), what the reference's data/binary/*.tsb hold) is measured in
the same run with the same steps and reported as a full second block, `tsb_shaped`.

A step = one whole matrix profile in the default (chain-faithful) fp64 mode: precompute kernels + diagonal-band kernel
+ chain-carry resolve [+ NCCL merges] + finalize.
  value  : series already resident in HBM when the step starts (device time, CUDA events, max over ranks)
  e2e    : the same through the C ABI with HOST buffers: pinned host series -> H2D -> kernels -> D2H of profile + index, every
           step; `one_call_ms` is mpb_compute() itself (pageable buffers, plan create/destroy inside) timed once on the host clock
  roofline: the diagonal-band kernel against the FP64 pipe issue peak measured in this run (DFMA microbenchmark).
           `frac` follows SURVEY.md section 8-d (4 pipe ops per useful cell: the reference formulation's 2 DFMA + 2 DMUL);
           `frac_physical` counts what the kernel EXECUTES, from its own counters (cells walked, blocks replayed, cells parked,
           diagonals re-seeded, sketch) -- the filter never forms P in its hot loop, so this is about half of `frac`.
  cpu_baseline / --impl reference: the reference's own Vanilla kernel sources (oracle/_ref, compiled from /root/reference at build
           time) on the host cores, on a bounded sample of diagonal chunks -- the SAME chunks in both legs for a given thread count.
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
REF_W = 4096              # MP_SIZE_W of the reference's README configuration (README.md:35)
REF_NOMINAL_RATE = 4.5e8  # useful cells/s of one host thread of the GPU boxes running the reference's Vanilla kernel (-O3 -march=native; measured 4.6-4.8e8 in round 2): sizes the sample


def synth_series(n, kind, seed=1234):
    import numpy as np
    rng = np.random.default_rng(seed)
    if kind == "uniform":
        return rng.random(n)
    if kind == "periodic":      # worst case of the data-dependent filter: every subsequence has thousands of near-best neighbours
        i = np.arange(n)
        return np.sin(2 * np.pi * i / 97.0) + 0.5 * np.sin(2 * np.pi * i / 1013.0) + 0.05 * rng.standard_normal(n)
    return np.cumsum(rng.standard_normal(n))


def useful_cells(n, m):
    L = (n - m + 1) - m // 4
    return L * (L + 1) // 2


def chunk_useful_cells(n, m, w, it):
    """Useful cells of diagonal chunk `it` (diagonals [it*w, (it+1)*w), exclusion zone and the out-of-range triangle removed)."""
    sublen, excl = n - m + 1, m // 4
    lo, hi = max(it * w, excl), min((it + 1) * w, sublen)
    return (hi - lo) * (2 * sublen - lo - hi + 1) // 2 if hi > lo else 0


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
def reference_sample_chunks(n, m, threads, target_s):
    """The diagonal chunks (width w) the CPU legs time: one per host thread, consecutive, placed where a chunk costs about
    target_s seconds at the nominal per-thread rate -- a pure function of (n, m, threads, target_s), so the cpu_baseline leg and
    the --impl reference arm time the same sample."""
    sublen = n - m + 1
    w = REF_W if n >= 4 * REF_W else 256
    nit = (n - m + w) // w
    want_rows = max(w, int(target_s * REF_NOMINAL_RATE / w))
    it0 = max(0, min(nit - threads, (sublen - want_rows) // w))
    return w, nit, list(range(it0, min(nit, it0 + threads)))


def cpu_reference_sample(n, m, kind, target_s=8.0, threads=None):
    """Times the reference's Vanilla kernel (kernel/MatrixProfileKernelVanilla.cpp, compiled unmodified into
    oracle/_ref/libmpref_fast.so with -O3 -march=native; the C restatement when the reference was not compiled here) on one
    diagonal chunk per host thread.  Returns the cpu_baseline object."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))      # the oracle: bench's CPU legs may execute it
    from concurrent.futures import ThreadPoolExecutor
    from oracle_py import Oracle, RefShim
    threads = threads or os.cpu_count() or 1
    w, nit, its = reference_sample_chunks(n, m, threads, target_s)
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
    return {"value": cells / dt, "unit": UNIT, "cores": len(its), "kind": "reference" if use_ref else "port",
            "sample": f"Vanilla kernel, double, w={w}: diagonal chunks {its[0]}..{its[-1]} of {nit} of the same n={n}, m={m} {kind} series, "
                      f"one chunk per host thread, {cells:.4g} useful cells in {dt:.1f} s (-O3 -march=native)",
            "seconds": dt}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    vals = []
    for i in range(args.warmup + args.steps):
        # warm-up steps are short samples (page the library and the series in); timed ones are the bounded sample
        base = cpu_reference_sample(args.n, args.m, args.data, target_s=1.0 if i < args.warmup else args.ref_seconds)
        if i >= args.warmup:
            vals.append(base)
    cells_s = sum(b["value"] * b["seconds"] for b in vals) / sum(b["seconds"] for b in vals)
    ms = 1e3 * sum(b["seconds"] for b in vals) / len(vals)
    base = dict(vals[-1]); base["value"] = cells_s
    line = {"impl": "reference", "metric": METRIC, "value": cells_s, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, args.data), "cpu_baseline": base,
            "e2e": {"value": cells_s, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def workload_config(args, kind):
    series = {"walk": "random-walk (north_star)", "uniform": "tsb-shaped U[0,1) (BASELINE.json configs[1])",
              "periodic": "two sines + 5 % noise (worst case of the filter: near-best neighbours everywhere)"}[kind]
    return {"workload": f"matrix profile, {'double' if args.dtype == 'f64' else 'float'}, n={args.n}, m={args.m}, synthetic {series} series, "
                        f"default mode ({'chain-faithful: running mean + telescoped chain carry' if args.dtype == 'f64' else 'periodic exact QT re-seed'})",
            "n": args.n, "m": args.m, "series": kind, "partition": f"diagonal bands round-robin over {args.gpus} GPU(s), NCCL max-merge in libmpb200.so",
            "l2": "flushed between timed steps (256 MiB write)"}


def executed_pipe_ops(tm, m, sketch_cells):
    """FP64/FP32-pipe operations the step EXECUTES, from the kernel's own counters:
    hot loop 128 FMA + 2 MUL per 64-cell block of every walked cell; a replay pass (32 lanes, one queued 8 x 8-cell block each) walks
    2048 cells again with 2 FMA + 2 MUL each; a parked cell costs 2 MUL in the drain; a re-seeded diagonal m FMA; a sketched cell
    2 FMA + 2 MUL."""
    return (tm["cells_walked"] * (130.0 / 64.0) + tm["replays"] * 2048 * 4.0 + tm["filter_hits"] * 8 * 2.0
            + tm["reseeded_diagonals"] * float(m) + sketch_cells * 4.0)


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
    # stdout carries the JSON line(s) and nothing else: native libraries write there too (NCCL prints its version banner to stdout
    # when NCCL_DEBUG is set), so file descriptor 1 is pointed at stderr for the run and the JSON goes to the original
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = args.n
    dtype = mp.F64 if args.dtype == "f64" else mp.F32
    ctx = mp.Context(local)
    stream = torch.cuda.ExternalStream(ctx.stream, device=local)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")     # > 126 MB of L2
    comm = None
    if world > 1:
        from mpb200.dist import make_comm
        comm = make_comm(ctx)               # communicator creation stays outside the timed region
    fma_peak, _ = ctx.measure_fma_peak(dtype, 1 << 15)
    fma_peak, _ = ctx.measure_fma_peak(dtype, 1 << 16)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    for m in ([int(x) for x in args.m_list.split(",")] if args.m_list else [args.m]):
        args.m = m
        out = bench_config(args, mp, torch, dist, ctx, stream, flush, comm, fma_peak, peaks, rank, world, local, dtype)
        if rank == 0:
            os.write(json_fd, (json.dumps(out) + "\n").encode())
    if comm is not None:
        comm.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def bench_config(args, mp, torch, dist, ctx, stream, flush, comm, fma_peak, peaks, rank, world, local, dtype):
    """One (n, m) configuration: both series, device-resident and end-to-end; returns rank 0's JSON object."""
    n, m = args.n, args.m
    sublen = n - m + 1
    cells = useful_cells(n, m)
    plan = mp.Plan(ctx, n, m, dtype, 0)
    plan.set_shard(rank, world)
    if args.margin:
        plan.set_chain_margin(args.margin)
    MP_pinned = torch.empty(sublen, dtype=torch.float64).pin_memory()
    MPI_pinned = torch.empty(sublen, dtype=torch.int32).pin_memory()
    if world > 1:
        plan.connect_peers(comm)            # every rank maps the others' threshold arrays (CUDA IPC over NVLink), once
        ctx.finish()

    def step(e2e, T_pinned):
        """One whole matrix profile on this rank's band shard; returns (t0, t1) events on the engine stream."""
        with torch.cuda.stream(stream):
            flush.zero_()
            t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
            t0.record(stream)
            if e2e:
                plan.upload_series_ptr(T_pinned.data_ptr())
            plan.precompute()
            if world > 1:
                plan.compute_merged(comm)   # shared sketch, first bands, threshold merge, remaining bands, result merge (csrc/multi_gpu.cu)
            else:
                plan.compute()
            plan.finalize()
            if e2e:
                try:
                    plan.download_into(MP_pinned.data_ptr(), MPI_pinned.data_ptr())
                except mp.MpbError as e:                     # MPB_E_MARGIN: keep in step with the other ranks; plan.check() below reports it
                    if e.code != -5:
                        raise
            t1.record(stream)
        return t0, t1

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(e2e, steps, warmup, T_pinned):
        for _ in range(warmup):
            step(e2e, T_pinned)
        sync_all()
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        evs, kern_ms, launches, tm = [], [], 0, None
        for _ in range(steps):
            t0, t1 = step(e2e, T_pinned)
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

    def measure(kind):
        T_pinned = torch.from_numpy(synth_series(n, kind)).pin_memory()
        plan.upload_series_ptr(T_pinned.data_ptr())
        ctx.finish()
        ms_dev, band_ms, launches, clocks, tm = timed(False, args.steps, args.warmup, T_pinned)
        ms_e2e, _, _, clocks_e2e, _ = timed(True, args.steps, max(min(3, args.warmup), args.warmup // 2), T_pinned)
        bad = 0                                             # every rank: the chain-carry margin covered what it had to
        try:
            plan.check()
        except mp.MpbError as e:
            if e.code != -5:                                # MPB_E_MARGIN: measure again with a wider margin (all ranks together)
                raise
            bad = 1
        if world > 1:
            flag = torch.tensor([bad], dtype=torch.int32, device="cuda")
            dist.all_reduce(flag, op=dist.ReduceOp.MAX)
            bad = int(flag.item())
        if bad:
            import math
            return -int(round(math.log2(plan.timings()["chain_margin"])))
        one_call_ms = None
        if world == 1 and rank == 0:                        # mpb_compute itself: pageable host buffers, plan create / destroy inside
            Tn = T_pinned.numpy().copy()
            mp.matrix_profile(Tn, m, dtype, 0, ctx=ctx)     # first call pays the allocator's page faults
            t0 = time.perf_counter()
            mp.matrix_profile(Tn, m, dtype, 0, ctx=ctx)
            one_call_ms = (time.perf_counter() - t0) * 1e3
        if rank != 0:
            return None
        my_cells = cells / world                                   # round-robin bands: equal shares to within one band
        ops = 4.0 * my_cells / (band_ms * 1e-3)
        sketch_cells = (sublen - m // 4) * 512.0 / world if dtype == mp.F64 or world == 1 else (sublen - m // 4) * 512.0
        executed = executed_pipe_ops(tm, m, sketch_cells)
        hbm_alg = 8.0 * n + 12.0 * sublen                          # read T once, write MP + MPI once
        traffic = None
        prof_path = os.path.join(ROOT, "profiles", "tile_kernel_traffic.json")
        if world == 1 and os.path.exists(prof_path):              # a 1-GPU ncu figure: not repeated for sharded runs
            traffic = json.load(open(prof_path)).get(f"{args.dtype}_n{n}_m{m}_{kind}")
        return {"value": cells / (ms_dev * 1e-3), "ms_per_step": ms_dev,
                "e2e": {"value": cells / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e, "h2d_bytes_per_step": 8 * n, "d2h_bytes_per_step": 12 * sublen,
                        "one_call_ms": one_call_ms, "clocks": clocks_e2e},
                "gpu_launches": launches,
                "roofline": {"bound": "fp64_fma_pipe" if args.dtype == "f64" else "fp32_fma_pipe", "kernel": "tile_kernel (+ chain-carry scan / resolve, sketch)",
                             "achieved": ops / 1e12, "peak": fma_peak / 1e12, "unit": "T pipe-ops/s (4 per cell: 2 FMA + 2 MUL)",
                             "frac": ops / fma_peak, "peak_source": "DFMA/FFMA issue-rate microbenchmark measured in this run (mpb_measure_fma_peak)",
                             "executed_pipe_ops_per_cell": executed / my_cells, "frac_physical": executed / (band_ms * 1e-3) / fma_peak,
                             "executed_source": "kernel counters of rank 0: cells walked, blocks replayed, cells parked, diagonals re-seeded, sketch",
                             "kernel_ms": band_ms, "kernel_share_of_step": band_ms / ms_dev, "traffic": traffic,
                             "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of one 1-GPU ncu --set full capture (profiles/)" if traffic else None,
                             "hbm": {"algorithmic_bytes_per_step": hbm_alg, "achieved_gbs": hbm_alg / (band_ms * 1e-3) / 1e9,
                                     "peak_gbs": peaks.get("hbm_gbs", 6650.0), "peak_source": "measured" if peaks else "fallback"}},
                "clocks": clocks, "phases_ms": {k: tm[k] for k in ("precompute_ms", "compute_ms", "finalize_ms")},
                "counters": {k: tm[k] for k in ("slow_updates", "filter_hits", "replays", "replay_lane_blocks", "cells_walked", "reseeded_diagonals", "chain_candidates",
                                                "chain_max_correction", "chain_bound", "chain_margin", "items", "segment_rows", "bands")}}

    def measure_with_margin(kind):
        while True:
            res = measure(kind)
            if not isinstance(res, int):
                return res
            plan.set_chain_margin(max(8, res - 2))          # res = the margin's log2 that proved too narrow

    head = measure_with_margin(args.data)
    other = "uniform" if args.data == "walk" else "walk"
    second = measure_with_margin(other) if not args.no_second_series else None
    out = None
    if rank == 0:
        out = {"metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": args.dtype,
               "data": "synthetic", "config": workload_config(args, args.data)}
        out.update({k: head[k] for k in ("e2e", "gpu_launches", "roofline", "clocks", "phases_ms", "counters")})
        if second:
            second["config"] = workload_config(args, other)
            out["tsb_shaped" if other == "uniform" else "random_walk"] = second
    plan.close()
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu_reference_sample(n, m, args.data, target_s=1.0)          # pages the library and the series in (untimed)
        cb = cpu_reference_sample(n, m, args.data, target_s=args.ref_seconds)
        cb.pop("seconds", None)
        out["cpu_baseline"] = cb
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n", type=int, default=1 << 20)
    ap.add_argument("--m", type=int, default=256)
    ap.add_argument("--dtype", default="f64", choices=["f64", "f32"])
    ap.add_argument("--data", default="walk", choices=["uniform", "walk", "periodic"])
    ap.add_argument("--ref-seconds", type=float, default=8.0, help="CPU seconds per reference sample (one diagonal chunk per host thread; the same sample in the cpu_baseline leg and the --impl reference arm)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-second-series", action="store_true", help="skip the second series block (long configs)")
    ap.add_argument("--m-list", default="", help="comma list of subsequence lengths: one JSON line per m (BASELINE.json config 5)")
    ap.add_argument("--margin", type=int, default=0, help="chain-carry near-tie margin 2^-k (0 = the library's default for n)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
