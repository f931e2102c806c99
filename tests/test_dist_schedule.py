"""Host logic of the multi-GPU merge schedule (python/mpb200/dist.py): where the ranks max-merge their profiles."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))
from mpb200.dist import merge_points, WARM_BANDS_TOTAL   # noqa: E402


def test_single_gpu_never_merges():
    assert merge_points(2048, 1) == []


def test_band_sharded_job_merges_once_when_the_thresholds_are_warm():
    for world in (2, 4, 8):
        nlb = 2048 // world
        pts = merge_points(nlb, world)
        assert pts == [-(-WARM_BANDS_TOTAL // world)]          # WARM_BANDS_TOTAL bands walked in total
        assert 0 < pts[0] < nlb


def test_tiny_problems_skip_the_early_merge():
    assert merge_points(1, 8) == [] and merge_points(6, 8) == []
