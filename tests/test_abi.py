"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/mp_b200.h declares, fails loudly without a device (no CPU fallback), and the multi-GPU merge
logic is right (world_size 2, gloo)."""
import ctypes
import os
import re

import numpy as np
import pytest
from conftest import ROOT


def header_symbols():
    src = open(os.path.join(ROOT, "include", "mp_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mpb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import mpb200
    if not os.path.exists(mpb200.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    L = ctypes.CDLL(mpb200.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 35
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/mp_b200.h but not exported"
    assert sorted(mpb200.EXPORTS) == syms      # the binding lists exactly the header's surface


def test_no_cpu_fallback_without_device():
    import torch
    import mpb200
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(mpb200.MpbError) as e:
        mpb200.Context(0)
    assert e.value.code == -2          # MPB_E_NO_DEVICE
    inp = np.zeros(93, mpb200.IN_F64)
    out = np.zeros(93, mpb200.OUT_F64)
    with pytest.raises(mpb200.MpbError):
        mpb200.MatrixProfileKernelTLF(100, 8, 0, inp, out, w=8)


def test_useful_cells_and_record_layouts():
    import mpb200
    assert mpb200.useful_cells(65536, 256) == (65281 - 64) * (65281 - 64 + 1) // 2
    assert mpb200.useful_cells(1 << 20, 256) == (1048321 - 64) * (1048321 - 64 + 1) // 2
    assert mpb200.IN_F64.itemsize == 32 and mpb200.OUT_F64.itemsize == 32
    assert mpb200.IN_F32.itemsize == 16 and mpb200.OUT_F32.itemsize == 16


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from mpb200.dist import allreduce_profile
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(42)
    L = 5000
    # every rank derives ALL per-rank partial profiles, keeps its own, and knows the expected merge
    corr_all = rng.random((world, L))
    idx_all = rng.integers(0, L, (world, L)).astype(np.int32)
    corr_all[:, :50] = 0.0; idx_all[:, :50] = -1                     # slots nobody updated
    corr_all[1:, 50:100] = corr_all[0, 50:100]                       # exact ties -> smallest index wins
    corr_all[0, 100:150] = 0.0; idx_all[0, 100:150] = -1             # unset on rank 0 only
    corr = torch.from_numpy(corr_all[rank].copy())
    index = torch.from_numpy(idx_all[rank].copy())
    allreduce_profile(corr, index)
    want_corr = corr_all.max(0)
    cand = np.where((corr_all == want_corr) & (idx_all >= 0), idx_all, 2**31 - 1).min(0)
    want_idx = np.where(cand == 2**31 - 1, -1, cand).astype(np.int32)
    q.put((rank, bool(np.array_equal(corr.numpy(), want_corr)), bool(np.array_equal(index.numpy(), want_idx))))
    dist.destroy_process_group()


def test_allreduce_profile_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(60)
    assert all(ok1 and ok2 for _, ok1, ok2 in res), res


def test_band_round_robin_partition_is_exact_cover():
    from mpb200.dist import band_owner
    for nb in (1, 7, 128, 1024):
        for world in (1, 2, 4, 8):
            owners = [band_owner(b, world) for b in range(nb)]
            assert all(0 <= o < world for o in owners)
            counts = [owners.count(r) for r in range(world)]
            assert max(counts) - min(counts) <= 1
