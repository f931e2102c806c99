"""tsbin codec: mirrors the reference tool util/tsbin.py (flags :113-190, formats :18-112)."""
import gzip
import os
import subprocess
import sys

import numpy as np
import pytest

from mpb200 import tsbin

REF = "/root/reference"
TOOL = os.path.join(os.path.dirname(tsbin.__file__), "tsbin.py")


def run(*args):
    return subprocess.run([sys.executable, TOOL, *args], capture_output=True, text=True)


def test_encode_decode_round_trip_double(tmp_path):
    rng = np.random.default_rng(5)
    x = np.concatenate([rng.standard_normal(1000), [0.0, -0.0, 1e-310, 1.7976931348623157e308, 0.1]])
    txt = tmp_path / "s.txt"
    txt.write_text("\n".join(repr(float(v)) for v in x) + "\n")
    assert run("-e", "-t", "double", str(txt), "-o", str(tmp_path / "s.tsb")).returncode == 0
    raw = np.fromfile(tmp_path / "s.tsb", "<f8")
    assert raw.tobytes() == x.astype("<f8").tobytes()                     # bit-exact, headerless, little-endian
    assert run("-d", "-t", "double", str(tmp_path / "s.tsb"), "-o", str(tmp_path / "back.txt")).returncode == 0
    back = np.array((tmp_path / "back.txt").read_text().split(), dtype=np.float64)
    assert back.tobytes() == x.tobytes()


def test_int_files_and_oneline_and_gz(tmp_path):
    idx = np.array([-1, 0, 7, 2**31 - 1, -2**31], np.int32)
    with gzip.open(tmp_path / "i.txt.gz", "wb") as f:
        f.write((" ".join(str(int(v)) for v in idx) + "\n").encode())      # all numbers on one line
    assert run("-e", "-t", "int", str(tmp_path / "i.txt.gz"), "-o", str(tmp_path / "i.mpib")).returncode == 0
    assert np.array_equal(np.fromfile(tmp_path / "i.mpib", "<i4"), idx)
    r = run("-d", "-t", "int", "--oneline", str(tmp_path / "i.mpib"))
    assert r.returncode == 0 and r.stdout.strip() == ", ".join(str(int(v)) for v in idx)
    r = run("-d", "-t", "int", str(tmp_path / "i.mpib"), "--offset", "1", "-l", "2")
    assert r.stdout.split() == ["0", "7"]


def test_errors(tmp_path):
    (tmp_path / "odd.bin").write_bytes(b"\0" * 12)
    assert run("-d", "-t", "double", str(tmp_path / "odd.bin")).returncode != 0          # 12 bytes is not a multiple of 8
    assert run("-d", "-t", "int", "-n", "4", str(tmp_path / "odd.bin")).returncode != 0   # 3 ints, 4 expected
    assert run("-d", "-t", "int", "-n", "3", str(tmp_path / "odd.bin")).returncode == 0
    (tmp_path / "bad.txt").write_text("1.0 two 3.0\n")
    assert run("-e", "-t", "double", str(tmp_path / "bad.txt"), "-o", str(tmp_path / "x.tsb")).returncode != 0
    (tmp_path / "f.txt").write_text("1.5\n")
    assert run("-e", "-t", "int", str(tmp_path / "f.txt"), "-o", str(tmp_path / "x.mpib")).returncode != 0   # int() rejects "1.5"
    assert run("-e", "-t", "double", str(tmp_path / "f.txt")).returncode != 0               # -o is mandatory for -e
    assert run(str(tmp_path / "f.txt")).returncode != 0                                      # neither -e nor -d
    assert run("-e", "-d", str(tmp_path / "f.txt")).returncode != 0
    assert run("-d", str(tmp_path / "missing.tsb")).returncode != 0


def test_api_feeds_the_host_driver_format(tmp_path):
    x = np.arange(10, dtype=np.float64) / 3
    tsbin.write_bin(str(tmp_path / "a.tsb"), x, "double")
    assert os.path.getsize(tmp_path / "a.tsb") == 80
    assert np.array_equal(tsbin.read_bin(str(tmp_path / "a.tsb")), x)
    with pytest.raises(tsbin.TsbinError):
        tsbin.read_bin(str(tmp_path / "a.tsb"), "double", 11)


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not mounted (GPU box)")
def test_against_reference_tool_and_shipped_files(tmp_path):
    # the shipped pair: data/benchmark/1024.txt.gz encodes to data/binary/1024.tsb byte for byte
    assert run("-e", "-t", "double", f"{REF}/data/benchmark/1024.txt.gz", "-o", str(tmp_path / "1024.tsb")).returncode == 0
    assert (tmp_path / "1024.tsb").read_bytes() == open(f"{REF}/data/binary/1024.tsb", "rb").read()
    # the reference tool itself on a real dataset, both directions
    src = f"{REF}/data/real/italianpowerdemand.txt"
    ref_tool = f"{REF}/util/tsbin.py"
    assert subprocess.run([sys.executable, ref_tool, "-e", "-t", "double", src, "-o", str(tmp_path / "ref.tsb"), "--offset", "5", "-l", "2000"], capture_output=True).returncode == 0
    assert run("-e", "-t", "double", src, "-o", str(tmp_path / "mine.tsb"), "--offset", "5", "-l", "2000").returncode == 0
    assert (tmp_path / "ref.tsb").read_bytes() == (tmp_path / "mine.tsb").read_bytes()
    assert subprocess.run([sys.executable, ref_tool, "-d", "-t", "double", str(tmp_path / "ref.tsb"), "-o", str(tmp_path / "ref.txt")], capture_output=True).returncode == 0
    assert run("-d", "-t", "double", str(tmp_path / "ref.tsb"), "-o", str(tmp_path / "mine.txt")).returncode == 0
    assert (tmp_path / "ref.txt").read_text() == (tmp_path / "mine.txt").read_text()
