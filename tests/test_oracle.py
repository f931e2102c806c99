"""CPU tests: the C oracle (oracle/mp_oracle.c) against golden vectors produced by the reference's
own sources (oracle/gen_golden.py), and -- when oracle/_ref was built in this container -- against
the reference code directly, plus the reference's own GoogleTest suites."""
import os
import subprocess

import numpy as np
import pytest
from conftest import ROOT, load_golden
from oracle_py import IN_F32, RefShim

KAT = ["small128_m16_w32_t16", "b1024_m16_w32_t16", "b1024_m16_w256_t128", "b1024_m32_w512_t128",
       "b1024_m30_w64_t16", "b16384_m128_w1024_t128"]
SYN = ["walk4096_m64_w256_t128", "walk8192_m100_w1024_t256"]
BIG = ["walk65536_m256_w4096_t128", "unif65536_m256_w4096_t128",
       "walk32768_m1024_w4096_t128", "walk20480_m4096_w4096_t128", "unif20480_m32_w4096_t128"]   # m of BASELINE.json configs 4 and 5


@pytest.mark.parametrize("name", KAT + SYN)
def test_precompute_bit_exact(oracle, name):
    g = load_golden(name)
    inp = oracle.precompute(g["T"], g["m"])
    assert inp.tobytes() == g["input"].tobytes()


@pytest.mark.parametrize("kind", ["vanilla", "tiled"])
@pytest.mark.parametrize("name", [n for n in KAT if n != "b16384_m128_w1024_t128"])
def test_chunk_outputs_bit_exact(oracle, name, kind):
    """Per-chunk OutputDataPacks (values AND indices, incl. the P=0 clamp and tie order) equal the
    reference kernel's, chunk by chunk."""
    g = load_golden(name)
    T, m, w, t = g["T"], g["m"], g["w"], g["t"]
    n = T.size
    inp = oracle.precompute(T, m)
    for it in range((n - m + w) // w):
        out = oracle.kernel(kind, n, m, w, it, inp, t)
        assert np.array_equal(out["rowAggregate"]["value"], g[f"{kind}_chunk{it}_rv"])
        assert np.array_equal(out["rowAggregate"]["index"], g[f"{kind}_chunk{it}_ri"])
        assert np.array_equal(out["columnAggregate"]["value"], g[f"{kind}_chunk{it}_cv"])
        assert np.array_equal(out["columnAggregate"]["index"], g[f"{kind}_chunk{it}_ci"])


@pytest.mark.parametrize("kind", ["vanilla", "tiled"])
@pytest.mark.parametrize("name", KAT + SYN)
def test_full_run_matches_reference_kernels(oracle, name, kind):
    g = load_golden(name)
    MP, MPI, corr, cells = oracle.run(g["T"], g["m"], g["w"], kind, g["t"], nthreads=4)
    assert np.array_equal(corr, g[f"{kind}_corr"]) and np.array_equal(MPI, g[f"{kind}_MPI"])
    assert np.array_equal(MP, g[f"{kind}_MP"])
    assert corr.min() > 0          # the P=0 clamp is never what is compared


@pytest.mark.parametrize("name", KAT + SYN)
def test_reference_scamp_known_answers(oracle, name):
    """The reference's pass criterion (test/include/MatrixProfileTests.hpp:204-212): |dMP| <= 1e-8 on the
    distance and identical indices, against its independent SCAMP."""
    g = load_golden(name)
    MP, MPI, corr = oracle.reference_scamp(g["T"], g["m"])
    assert np.array_equal(MP, g["scamp_MP"]) and np.array_equal(MPI, g["scamp_MPI"])
    kMP, kMPI, _, _ = oracle.run(g["T"], g["m"], g["w"], "vanilla")
    assert np.abs(kMP - g["scamp_MP"]).max() <= 1e-8
    assert np.array_equal(kMPI, g["scamp_MPI"])


@pytest.mark.parametrize("name", BIG)
def test_config1_full(oracle, name):
    """BASELINE.json config 1 (n=65536, m=256) and the m = 32 / 1024 / 4096 cases in full, multi-threaded chunks."""
    g = load_golden(name)
    assert bool(g["tiled_equals_vanilla"])
    MP, MPI, corr, cells = oracle.run(g["T"], g["m"], g["w"], "vanilla", nthreads=os.cpu_count() or 4)
    assert np.array_equal(corr, g["vanilla_corr"]) and np.array_equal(MPI, g["vanilla_MPI"])
    sublen, w = g["T"].size - g["m"] + 1, g["w"]
    assert cells == sum((sublen - k * w) * w for k in range((sublen + w - 1) // w))   # processed cells, SURVEY section 8-d


@pytest.mark.parametrize("name", ["f32_b1024_m16_w256_t128", "f32_walk4096_m64_w256_t128"])
@pytest.mark.parametrize("kind", ["vanilla", "tiled"])
def test_float_chunks_bit_exact(oracle, name, kind):
    g = load_golden(name)
    T, m, w, t = g["T"], g["m"], g["w"], g["t"]
    n = T.size
    inp = oracle.precompute(T, m, "f32")
    assert inp.dtype == IN_F32 and inp.tobytes() == g["input"].tobytes()
    for it in range((n - m + w) // w):
        out = oracle.kernel(kind, n, m, w, it, inp, t)
        assert np.array_equal(out["rowAggregate"]["value"], g[f"{kind}_chunk{it}_rv"])
        assert np.array_equal(out["rowAggregate"]["index"], g[f"{kind}_chunk{it}_ri"])
        assert np.array_equal(out["columnAggregate"]["value"], g[f"{kind}_chunk{it}_cv"])
        assert np.array_equal(out["columnAggregate"]["index"], g[f"{kind}_chunk{it}_ci"])


def test_exact_corr_sampled_rows(oracle):
    """The oracle chain stays within ~1e-10 of exact long-double correlations on sampled rows."""
    g = load_golden("walk65536_m256_w4096_t128")
    T, m = g["T"], g["m"]
    rows = np.linspace(0, T.size - m, 24).astype(int)
    err = max(abs(oracle.exact_corr(T, m, i, g["vanilla_MPI"][i]) - g["vanilla_corr"][i]) for i in rows)
    assert err < 5e-10


# ---- directly against the reference's code, when it was compiled here (never on the GPU box) ----
needs_ref = pytest.mark.skipif(not RefShim.available(), reason="oracle/_ref not built (no /root/reference here)")


@needs_ref
def test_oracle_vs_ref_shim_random_shapes(oracle):
    ref = RefShim()
    rng = np.random.default_rng(5)
    for n, m, w, t in [(1024, 16, 64, 16), (1024, 30, 128, 32), (2048, 64, 256, 128), (4096, 64, 512, 128)]:
        T = np.cumsum(rng.standard_normal(n))
        inp = ref.precompute(T, m)
        assert oracle.precompute(T, m).tobytes() == inp.tobytes()
        for it in range((n - m + w) // w):
            for kind in ("vanilla", "tiled"):
                a, b = oracle.kernel(kind, n, m, w, it, inp, t), ref.kernel(kind, n, m, w, it, inp, t)
                for f in ("rowAggregate", "columnAggregate"):
                    assert np.array_equal(a[f]["value"], b[f]["value"]) and np.array_equal(a[f]["index"], b[f]["index"])


@needs_ref
def test_p_zero_clamp_names_an_excluded_column(oracle):
    """The edge the GPU engine documents a divergence for (tests/test_gpu_parity.py::
    test_slots_without_a_positive_correlation): when every valid correlation of a slot is negative, the reference's P = 0 substitution
    for excluded / out-of-range cells (Vanilla.cpp:72-84) wins -- value 0, index of a column INSIDE the exclusion zone.  Pinned here on
    the reference's own kernel and on the restatement."""
    T, m = np.random.default_rng(3).standard_normal(24), 16
    for w in (8, 32):
        MP, MPI, corr, _ = RefShim().run(T, m, w)
        oMP, oMPI, ocorr, _ = oracle.run(T, m, w)
        assert np.array_equal(corr, ocorr) and np.array_equal(MPI, oMPI) and np.array_equal(MP, oMP)
        zero = np.nonzero(corr == 0)[0]
        assert zero.size >= 1 and (corr >= 0).all()
        assert all(abs(int(MPI[i]) - int(i)) < m // 4 for i in zero)          # the "neighbour" is an excluded column
        assert np.array_equal(MP[zero], np.full(zero.size, np.sqrt(2.0 * m)))


@needs_ref
@pytest.mark.parametrize("exe", ["testVanillaKernel", "testTiledKernel"])
def test_reference_own_gtest_suites_pass(exe):
    """The reference's eight known-answer tests, unmodified, compiled by oracle/Makefile."""
    path = os.path.join(ROOT, "oracle", "_ref", exe)
    if not os.path.exists(path) or not os.path.isdir("/root/reference/test"):
        pytest.skip("gtest binaries / fixtures not present")
    r = subprocess.run([path], cwd="/root/reference/test", capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:]
    assert "PASSED" in r.stdout and "FAILED" not in r.stdout
