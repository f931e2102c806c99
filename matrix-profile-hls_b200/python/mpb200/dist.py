"""Multi-GPU plumbing: one process per GPU, diagonal bands dealt round-robin over the ranks
(the reference's chunk -> compute-unit round-robin, host/MatrixProfileHost.cpp:101-128), and ONE exchange
step: an element-wise max of the per-rank profiles (the reference's UpdateAggregates,
include/host/HostSideComputation.hpp:72-90, which its host does serially).

A (float64 corr, int32 index) pair does not fit one 64-bit word exactly, so the max-reduction runs as
two NCCL all-reduces over NVLink: MAX over the correlations viewed as int64 (corr >= +0.0, so the IEEE
bit pattern is order-preserving), then MIN over the indices with every rank that does not hold the
maximum contributing INT32_MAX.  The result is exact and independent of rank count and arrival order
(ties -> smallest index, the same rule the band kernel's 128-bit CAS applies).
"""
import torch
import torch.distributed as dist

INT32_MAX = 2**31 - 1


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


def merge_plan_across_ranks(plan, corr_buf, index_buf, group=None):
    """GPU path: export the plan's packed profile, all-reduce it over NCCL, import the merged profile.
    corr_buf / index_buf are CUDA tensors (float64 / int32, sublen) owned by the caller and reused."""
    plan.export(corr_buf.data_ptr(), index_buf.data_ptr())

    def mask(mine, index, corr_max):
        plan.mask_index(corr_max.data_ptr(), index.data_ptr())   # kernel: index <- mine==max ? index : INT32_MAX
        return index

    allreduce_profile(corr_buf, index_buf, mask_fn=mask, group=group)
    plan.import_(corr_buf.data_ptr(), index_buf.data_ptr())


WARM_BANDS_TOTAL = 48     # diagonal bands (512 diagonals each) after which the filter thresholds are tight


def merge_points(nlb, world):
    """Local-band indices after which the ranks max-merge their profiles before the final merge: once, when about
    WARM_BANDS_TOTAL bands have been walked in total."""
    warm = max(1, -(-WARM_BANDS_TOTAL // world))
    return [warm] if world > 1 and warm < nlb else []


def compute_merged(plan, corr_buf, index_buf, world, group=None):
    """The multi-GPU compute step: every rank walks its first few bands, the profiles are max-merged (each rank's
    filter thresholds then come from about WARM_BANDS_TOTAL bands of ALL ranks instead of its own few), then every
    rank walks the rest of its bands and the profiles are merged again.  Returns the number of kernels launched
    besides the tile kernel."""
    nlb = plan.local_bands()
    extra = 0
    lo = 0
    for pt in merge_points(nlb, world):
        plan.compute_bands(lo, pt)
        merge_plan_across_ranks(plan, corr_buf, index_buf, group=group)
        extra += 3
        lo = pt
    plan.compute_bands(lo, -1)
    if world > 1:
        merge_plan_across_ranks(plan, corr_buf, index_buf, group=group)
        extra += 3
    return extra
