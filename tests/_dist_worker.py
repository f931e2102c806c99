"""torchrun worker of tests/test_gpu_multi.py: one rank per GPU computes the matrix profile of a seeded series through the
product's multi-GPU step (mpb_plan_compute_merged: NCCL merges inside libmpb200.so); rank 0 saves MP / MPI / corr."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))
import mpb200 as mp                      # noqa: E402
from mpb200.dist import make_comm        # noqa: E402

out, kind, n, m, dtype = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rng = np.random.default_rng(1234)
T = rng.random(n) if kind == "uniform" else np.cumsum(rng.standard_normal(n))
ctx = mp.Context(local)
comm = make_comm(ctx)
plan = mp.Plan(ctx, n, m, dtype, 0)
plan.set_shard(rank, world)
if os.environ.get("MPB_TEST_NO_PEERS") is None:
    plan.connect_peers(comm)
    assert plan.shared_with == world - 1
plan.upload_series(T)
for _ in range(2):                       # twice: a second step on the same plan must give the same bits
    plan.precompute()
    plan.compute_merged(comm)
    plan.finalize()
    MP, MPI, corr = plan.download()
plan.check()
res = torch.from_numpy(corr).cuda()
ref = res.clone()
dist.broadcast(ref, 0)
same = torch.tensor([int(torch.equal(ref, res))], device="cuda")
dist.all_reduce(same, op=dist.ReduceOp.MIN)           # every rank holds the job's profile after the result merge
if rank == 0:
    np.savez(out, MP=MP, MPI=MPI, corr=corr, all_ranks_equal=int(same.item()), bands=plan.timings()["bands"])
plan.close(); comm.close(); ctx.close()
dist.barrier()
dist.destroy_process_group()
