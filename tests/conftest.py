import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def load_golden(name):
    """Load tests/golden/<name>.npz; synthetic inputs are regenerated from their seed and checked."""
    import numpy as np
    from oracle_py import random_walk, uniform_series

    d = dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    if "T" not in d:
        gen = {"walk": random_walk, "uniform": uniform_series}[str(d["gen"])]
        T = gen(int(d["gen_n"]), int(d["gen_seed"]))
        assert T.sum() == d["T_sum"] and (T[:4] == d["T_head"]).all() and (T[-4:] == d["T_tail"]).all(), \
            "synthetic generator no longer reproduces the golden input"
        d["T"] = T
    for k in ("m", "w", "t"):
        d[k] = int(d[k])
    return d


@pytest.fixture(scope="session")
def oracle():
    from oracle_py import Oracle
    return Oracle()
