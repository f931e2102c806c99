"""The C++ host driver MatrixProfileHost: the reference's CLI surface (host/MatrixProfileHost.cpp:162-217),
file formats (include/host/FileIO.hpp) and loud failure without a device.  The GPU test runs the binary end to end
and compares its .mpb/.mpib with the golden vectors."""
import os
import subprocess

import numpy as np
import pytest
from conftest import ROOT, load_golden

EXE = os.path.join(ROOT, "matrix-profile-hls_b200", "MatrixProfileHost")


@pytest.fixture(scope="module")
def exe():
    if not os.path.exists(EXE):
        import __graft_entry__
        __graft_entry__.build()
    assert os.path.exists(EXE)
    return EXE


def run(exe, *args):
    return subprocess.run([exe, *args], capture_output=True, text=True, timeout=300)


def test_cli_help_version_and_errors(exe):
    r = run(exe, "--help")
    assert r.returncode == 0 and "-i, --input" in r.stdout and "-b, --xclbin" in r.stdout and "--verbose" in r.stdout
    r = run(exe, "-v")
    assert r.returncode == 0 and "MP_SIZE_M" in r.stdout and "MP_KERNEL" in r.stdout
    r = run(exe)                                   # --input required -> [ERROR] + help + failure (reference :196-200)
    assert r.returncode != 0 and "[ERROR] --input required" in r.stderr and "Usage" in r.stdout
    r = run(exe, "--no-such-flag")                 # unknown argument (reference :212-216)
    assert r.returncode != 0 and "[ERROR] Unknown argument" in r.stderr
    r = run(exe, "-i", "/nonexistent/file.tsb")
    assert r.returncode != 0 and "[ERROR]" in r.stderr


def test_rejects_ragged_file(exe, tmp_path):
    p = tmp_path / "bad.tsb"
    p.write_bytes(b"\0" * 1001)                    # not a multiple of 8 bytes
    r = run(exe, "-i", str(p))
    assert r.returncode != 0 and "[ERROR]" in r.stderr


def test_fails_loudly_without_device(exe, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = tmp_path / "s.tsb"
    np.random.default_rng(0).random(2048).tofile(p)
    r = run(exe, "-i", str(p), "-o", str(tmp_path / "out"))
    assert r.returncode != 0 and "[ERROR]" in r.stderr and "no CUDA device" in r.stderr
    assert not (tmp_path / "out.mpb").exists()     # no CPU fallback wrote anything


@pytest.mark.gpu
@pytest.mark.parametrize("extra", [[], ["--strict"], ["-b", "ignored.xclbin", "--verbose"]])
def test_end_to_end_files_match_golden(exe, tmp_path, extra):
    g = load_golden("walk8192_m100_w1024_t256")
    p = tmp_path / "walk.tsb"
    g["T"].astype("<f8").tofile(p)
    out = tmp_path / "res"
    r = run(exe, "-i", str(p), "-o", str(out), "-m", str(g["m"]), *extra)
    assert r.returncode == 0, r.stderr
    assert "[INFO] Kernel Execution Completed Successfully." in r.stdout and "Benchmark Profile" in r.stdout
    MP = np.fromfile(str(out) + ".mpb", "<f8")
    MPI = np.fromfile(str(out) + ".mpib", "<i4")
    want = np.sqrt(2 * g["m"] * (1 - g["vanilla_corr"]))
    assert MP.size == want.size and MPI.size == want.size
    if "--strict" in extra:
        assert np.array_equal(MP, want)
    assert np.allclose(MP, want, rtol=1e-9, atol=0)
    assert np.array_equal(MPI, g["vanilla_MPI"])


@pytest.mark.gpu
def test_no_output_warns_and_discards(exe, tmp_path):
    p = tmp_path / "u.tsb"
    np.random.default_rng(1).random(4096).tofile(p)
    r = run(exe, "-i", str(p), "-m", "64")
    assert r.returncode == 0 and "[WARN] No output" in r.stdout
