"""Multi-GPU GPU tests.  The schedule's pieces are emulated on ONE device (any GPU box); the tests that need two or more GPUs
(NCCL inside libmpb200.so: torchrun ranks, the in-process mpb_compute_multi, the host binary with --gpus) skip on a 1-GPU box and
run under `gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`."""
import os
import subprocess
import sys

import numpy as np
import pytest
from conftest import ROOT, load_golden

pytestmark = pytest.mark.gpu
PKG = os.path.join(ROOT, "matrix-profile-hls_b200")


@pytest.fixture(scope="module")
def mp():
    import mpb200
    assert mpb200.device_count() >= 1
    return mpb200


def ngpus():
    import mpb200
    return mpb200.device_count()


def single(mp, T, m, dtype=0):
    ctx = mp.Context(0)
    plan = mp.Plan(ctx, T.size, m, dtype, 0)
    plan.upload_series(T); plan.precompute(); plan.compute(); plan.finalize()
    res = plan.download()
    plan.close(); ctx.close()
    return res


def test_merged_schedule_emulated_on_one_device(mp):
    """mpb_plan_compute_merged's steps with torch doing the reductions: every shard sketches its share of band 0, the hints are
    max-reduced, the first bands are walked, the results so far merged (which lifts every shard's thresholds), the rest walked,
    the results merged -- the profile equals the single-plan one bit for bit (fp64 chain-carry mode)."""
    import torch
    g = load_golden("walk65536_m256_w4096_t128")
    T, m = g["T"], g["m"]
    MP0, MPI0, corr0 = single(mp, T, m)
    ctx = mp.Context(0)
    stream = torch.cuda.ExternalStream(ctx.stream)
    world = 3
    plans = []
    for r in range(world):
        p = mp.Plan(ctx, T.size, m, mp.F64, 0)
        p.set_shard(r, world); p.upload_series(T); p.precompute()
        plans.append(p)
    total = plans[0].total_bands()
    pt = mp.merge_point(total, world)
    assert pt > 0
    np_count = plans[0].hint_count
    with torch.cuda.stream(stream):
        for r, p in enumerate(plans):
            p.sketch_share(0, r, world)
        bufs = [torch.empty(np_count, dtype=torch.int32, device="cuda") for _ in plans]
        for p, b in zip(plans, bufs):
            _copy_d2d(b.data_ptr(), p.hint_ptr, 4 * np_count)
        hmax = bufs[0]
        for b in bufs[1:]:
            hmax = torch.maximum(hmax, b)
        hmax = hmax.contiguous()
        for p in plans:
            _copy_d2d(p.hint_ptr, hmax.data_ptr(), 4 * np_count)
            p.apply_hint()
        sub = T.size - m + 1

        def merge():
            corr = [torch.empty(sub, dtype=torch.float64, device="cuda") for _ in plans]
            idx = [torch.empty(sub, dtype=torch.int32, device="cuda") for _ in plans]
            for p, c, i in zip(plans, corr, idx):
                p.export(c.data_ptr(), i.data_ptr())
            cmax = corr[0].view(torch.int64)
            for c in corr[1:]:
                cmax = torch.maximum(cmax, c.view(torch.int64))
            cmax = cmax.view(torch.float64).contiguous()
            for p, i in zip(plans, idx):
                p.mask_index(cmax.data_ptr(), i.data_ptr())
            imin = idx[0]
            for i in idx[1:]:
                imin = torch.minimum(imin, i)
            imin = imin.contiguous()
            for p in plans:
                p.import_(cmax.data_ptr(), imin.data_ptr())
            stream.synchronize()

        for p in plans:
            p.compute_bands(0, pt)
        merge()
        for p in plans:
            p.compute_bands(pt, -1)
        merge()
        plans[1].finalize()
        stream.synchronize()
    MP, MPI, corr = plans[1].download()
    for p in plans:
        p.check(); p.close()
    ctx.close()
    assert np.array_equal(corr, corr0) and np.array_equal(MPI, MPI0) and np.array_equal(MP, MP0)


@pytest.mark.parametrize("dtype", [0, 1])
def test_pushed_thresholds_do_not_change_the_result(mp, dtype):
    """Plans that push their improvements into one another's threshold arrays (what connected ranks do over NVLink; here four band
    shards on one device, run one after the other so the later ones start from the earlier ones' thresholds) merge to the
    single-plan profile bit for bit, with far fewer exact updates than four isolated shards."""
    g = load_golden("walk65536_m256_w4096_t128")
    T, m = g["T"], g["m"]
    MP0, MPI0, corr0 = single(mp, T, m, dtype)
    ctx = mp.Context(0)
    world = 4
    updates = {}
    for shared in (False, True):
        plans = []
        for r in range(world):
            p = mp.Plan(ctx, T.size, m, dtype, 0)
            p.set_shard(r, world); p.upload_series(T); p.precompute()
            plans.append(p)
        if shared:
            for p in plans:
                p.share_thresholds([q for q in plans if q is not p])
                assert p.shared_with == world - 1
        for p in plans:
            p.compute()
        plans[0].merge_peers(plans[1:])
        plans[0].finalize()
        MP, MPI, corr = plans[0].download()
        updates[shared] = sum(p.timings()["slow_updates"] for p in plans)
        for p in plans:
            p.check(); p.close()
        assert np.array_equal(corr, corr0) and np.array_equal(MPI, MPI0) and np.array_equal(MP, MP0), shared
    ctx.close()
    assert updates[True] < 0.8 * updates[False]


def _copy_d2d(dst, src, nbytes):
    """cudaMemcpy device -> device between a torch tensor and a raw library pointer (the runtime torch already loaded)."""
    import ctypes
    import torch
    torch.cuda.synchronize()
    rt = None
    for name in ("libcudart.so.12", "libcudart.so.13", "libcudart.so"):
        try:
            rt = ctypes.CDLL(name)
            break
        except OSError:
            continue
    assert rt is not None, "libcudart not loadable"
    rt.cudaMemcpy.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
    assert rt.cudaMemcpy(ctypes.c_void_p(dst), ctypes.c_void_p(src), ctypes.c_size_t(nbytes), 3) == 0      # cudaMemcpyDeviceToDevice
    torch.cuda.synchronize()


@pytest.mark.parametrize("kind,n,m,dtype", [("walk", 1 << 17, 256, 0), ("uniform", 1 << 17, 256, 0), ("walk", 1 << 17, 256, 1), ("walk", 50000, 64, 0)])
def test_compute_multi_matches_single_gpu(mp, kind, n, m, dtype):
    """mpb_compute_multi (one host thread per GPU, NCCL merge inside the library) == the single-GPU profile, bit for bit."""
    if ngpus() < 2:
        pytest.skip("needs two GPUs")
    rng = np.random.default_rng(1234)
    T = rng.random(n) if kind == "uniform" else np.cumsum(rng.standard_normal(n))
    MP0, MPI0, _ = single(mp, T, m, dtype)
    for G in sorted({2, ngpus()}):
        MP, MPI, tms = mp.matrix_profile_multi(T, m, G, dtype, return_timings=True)
        assert np.array_equal(MP, MP0) and np.array_equal(MPI, MPI0), G
        assert len(tms) == G and all(t["bands"] > 0 for t in tms)


@pytest.mark.parametrize("kind,dtype", [("walk", 0), ("uniform", 0), ("walk", 1)])
def test_torchrun_nccl_merge_bit_equal_to_single_gpu(mp, tmp_path, kind, dtype):
    """One process per GPU (torchrun), the product's own multi-GPU step: the merged profile on hardware equals the 1-GPU bits,
    and every rank ends up holding it."""
    if ngpus() < 2:
        pytest.skip("needs two GPUs")
    n, m = 1 << 18, 256
    rng = np.random.default_rng(1234)
    T = rng.random(n) if kind == "uniform" else np.cumsum(rng.standard_normal(n))
    MP0, MPI0, corr0 = single(mp, T, m, dtype)
    for G in sorted({2, ngpus()}):
        out = str(tmp_path / f"res_{kind}_{dtype}_{G}.npz")
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={G}", "--master-addr", "127.0.0.1", "--master-port", str(29500 + G),
               os.path.join(ROOT, "tests", "_dist_worker.py"), out, kind, str(n), str(m), str(dtype)]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-3000:]
        d = np.load(out)
        assert int(d["all_ranks_equal"]) == 1
        assert np.array_equal(d["corr"], corr0) and np.array_equal(d["MPI"], MPI0) and np.array_equal(d["MP"], MP0), G


def test_host_binary_gpus2_writes_the_single_gpu_files(mp, tmp_path):
    """MatrixProfileHost --gpus 2 (host threads + NCCL) writes .mpb / .mpib byte-identical to --gpus 1."""
    if ngpus() < 2:
        pytest.skip("needs two GPUs")
    exe = os.path.join(PKG, "MatrixProfileHost")
    rng = np.random.default_rng(5)
    T = np.cumsum(rng.standard_normal(100000))
    inp = str(tmp_path / "series.tsb")
    T.astype("<f8").tofile(inp)
    outs = []
    for G in (1, 2):
        o = str(tmp_path / f"out{G}")
        r = subprocess.run([exe, "-i", inp, "-o", o, "-m", "128", "--gpus", str(G), "--verbose"], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        outs.append((open(o + ".mpb", "rb").read(), open(o + ".mpib", "rb").read()))
    assert outs[0] == outs[1]
