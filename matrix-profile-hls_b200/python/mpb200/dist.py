"""Multi-GPU plumbing for one-process-per-GPU jobs (torchrun): the compute step, the NCCL max-reduction and the merge
schedule live in libmpb200.so (csrc/multi_gpu.cu, C ABI mpb_comm_* / mpb_plan_compute_merged); torch.distributed is only
used to hand rank 0's NCCL unique id to the other ranks.

Diagonal bands are dealt round-robin over the ranks (the reference's chunk -> compute-unit round-robin,
host/MatrixProfileHost.cpp:101-128) and there is ONE exchange step: an element-wise max of the per-rank profiles (the
reference's UpdateAggregates, include/host/HostSideComputation.hpp:72-90, which its host does serially).  A (float64 corr,
int32 index) pair does not fit one 64-bit word exactly, so the reduction runs as two all-reduces: MAX over the correlations
viewed as int64 (corr >= +0.0, so the IEEE bit pattern is order-preserving), then MIN over the indices with every rank that
does not hold the maximum contributing INT32_MAX.  allreduce_profile / mask_index_torch restate that rule on torch tensors
for the CPU (gloo) tests of the host logic.
"""
import torch
import torch.distributed as dist

from . import COMM_ID_BYTES, Comm, comm_unique_id, merge_point

INT32_MAX = 2**31 - 1
WARM_BANDS_TOTAL = 48     # mirrors MPB_WARM_BANDS_TOTAL (csrc/multi_gpu.cu)


def mask_index_torch(corr, index, corr_max):
    """Reference implementation of mpb_plan_mask_index for CPU tensors (gloo tests)."""
    return torch.where((corr == corr_max) & (index >= 0), index, torch.full_like(index, INT32_MAX))


def allreduce_profile(corr, index, mask_fn=mask_index_torch, group=None):
    """In-place exact (corr, index) max across ranks. corr float64[sublen], index int32[sublen]."""
    assert corr.dtype == torch.float64 and index.dtype == torch.int32
    mine = corr.clone()
    dist.all_reduce(corr.view(torch.int64), op=dist.ReduceOp.MAX, group=group)
    masked = mask_fn(mine, index, corr)
    dist.all_reduce(masked, op=dist.ReduceOp.MIN, group=group)
    index.copy_(torch.where(masked == INT32_MAX, torch.full_like(masked, -1), masked))
    return corr, index


def band_owner(band, world):
    return band % world


def merge_points(total_bands, world):
    """Local-band indices after which the ranks max-merge their thresholds before the final merge (mpb_merge_point): once,
    when about WARM_BANDS_TOTAL bands have been walked in total.  A function of the JOB's band count, so that every rank takes
    the same decision even when the bands do not divide evenly over the ranks."""
    pt = merge_point(total_bands, world)
    return [pt] if pt > 0 else []


def make_comm(ctx, group=None):
    """An mpb200.Comm for this rank of the default (or given) process group: rank 0 draws the NCCL unique id, the group's
    broadcast hands it out (CPU tensor on gloo, CUDA tensor on nccl), every rank then joins ncclCommInitRank."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    dev = torch.device("cuda", ctx.device) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    buf = torch.zeros(COMM_ID_BYTES, dtype=torch.uint8, device=dev)
    if rank == 0:
        buf.copy_(torch.frombuffer(bytearray(comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(buf, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    return Comm(ctx, bytes(buf.cpu().numpy().tobytes()), rank, world)
