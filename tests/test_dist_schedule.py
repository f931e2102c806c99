"""Host logic of the multi-GPU merge schedule (csrc/multi_gpu.cu mpb_merge_point, mirrored by python/mpb200/dist.py): where the
ranks max-merge their thresholds, and that every rank of a job takes the same decision."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))
from mpb200.dist import merge_points, WARM_BANDS_TOTAL   # noqa: E402


def local_bands(total, rank, world):
    """Bands rank `rank` owns when `total` bands are dealt round-robin (csrc/scamp_tile.cu build_item_table)."""
    return (total - rank + world - 1) // world if total > rank else 0


def test_single_gpu_never_merges():
    assert merge_points(2048, 1) == []


def test_band_sharded_job_merges_once_when_the_thresholds_are_warm():
    for world in (2, 4, 8):
        pts = merge_points(2048, world)
        assert pts == [-(-WARM_BANDS_TOTAL // world)]          # WARM_BANDS_TOTAL bands walked in total
        assert 0 < pts[0] < 2048 // world


def test_tiny_problems_skip_the_early_merge():
    assert merge_points(8, 8) == [] and merge_points(48, 8) == []


def test_every_rank_takes_the_same_decision_and_owns_the_bands_it_walks_first():
    """Round 1 derived the merge point from the rank-local band count: with 8 GPUs and 49..55 bands some ranks issued four
    all-reduces and the others two (a hang).  The schedule is now a function of the job's band count only, and the bands a
    rank is told to walk before the early merge must exist on EVERY rank."""
    for world in (2, 3, 4, 8):
        for total in range(1, 400):
            pts = merge_points(total, world)
            assert len(pts) <= 1
            for pt in pts:
                assert all(0 < pt < local_bands(total, r, world) for r in range(world)), (total, world, pt)
