#!/usr/bin/env python3
"""Generate tests/golden/*.npz from the REFERENCE's own sources (TEST INFRASTRUCTURE ONLY).

Run in the build container, where /root/reference exists:
    make -C oracle ref && python oracle/gen_golden.py
Every value below is produced by code compiled from /root/reference (oracle/_ref/libmpref.so, built
by oracle/Makefile with -O2 -ffp-contract=off): the Vanilla and Tiled HLS kernels as plain C++
(kernel/MatrixProfileKernel{Vanilla,Tiled}.cpp), the test-side precompute
(test/include/MatrixProfileTests.hpp:117-139) and the test-side SCAMP
(test/include/MatrixProfileReference.hpp:90-136).  The cases are the reference's eight known-answer
tests (test/TestVanillaKernel.cpp:29-42, test/TestTiledKernel.cpp:29-52) plus BASELINE.json config 1
(n=65536, m=256 random walk) and a few edge shapes.  The inputs of the known-answer cases are the
reference's fixtures data/test/*.txt, stored here as float64.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from oracle_py import IN_F32, Oracle, RefShim, random_walk, uniform_series  # noqa: E402

REF = os.environ.get("MP_REFERENCE_DIR", "/root/reference")
OUT = os.path.join(HERE, "..", "tests", "golden")


def fields(out):
    return (out["rowAggregate"]["value"], out["rowAggregate"]["index"], out["columnAggregate"]["value"], out["columnAggregate"]["index"])


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = RefShim()
    load = lambda f: np.loadtxt(os.path.join(REF, "data", "test", f))
    T128, T1024, T16k = load("small128_syn.txt"), load("1024.txt"), load("16384.txt")
    # name, T, m, w, t, keep per-chunk outputs?
    cases = [
        ("small128_m16_w32_t16", T128, 16, 32, 16, True, None),       # TestSmall128SynM16(W32)
        ("b1024_m16_w32_t16", T1024, 16, 32, 16, True, None),         # TestBenchmark1024SynM16W32(T16)
        ("b1024_m16_w256_t128", T1024, 16, 256, 128, True, None),     # TestBenchmark1024SynM16W256T128
        ("b1024_m32_w512_t128", T1024, 32, 512, 128, True, None),     # TestBenchmark1024SynM32W256T128 (w is 512)
        ("b16384_m128_w1024_t128", T16k, 128, 1024, 128, False, None),  # TestBenchmark16384SynM128W1024(T128)
        ("b1024_m30_w64_t16", T1024, 30, 64, 16, True, None),         # edge: m not a multiple of 4
        ("walk4096_m64_w256_t128", random_walk(4096, 7), 64, 256, 128, False, ("walk", 7)),
        ("walk8192_m100_w1024_t256", random_walk(8192, 11), 100, 1024, 256, False, ("walk", 11)),
        ("walk65536_m256_w4096_t128", random_walk(65536, 1234), 256, 4096, 128, False, ("walk", 1234)),  # BASELINE.json config 1
        ("unif65536_m256_w4096_t128", uniform_series(65536, 1234), 256, 4096, 128, False, ("uniform", 1234)),
        ("walk32768_m1024_w4096_t128", random_walk(32768, 21), 1024, 4096, 128, False, ("walk", 21)),     # BASELINE.json config 4's m
        ("walk20480_m4096_w4096_t128", random_walk(20480, 22), 4096, 4096, 128, False, ("walk", 22)),     # config 5's largest m
        ("unif20480_m32_w4096_t128", uniform_series(20480, 23), 32, 4096, 128, False, ("uniform", 23)),   # config 5's smallest m
    ]
    only = set(sys.argv[1:])          # optional: regenerate just the named cases
    if only:
        cases = [c for c in cases if c[0] in only]
    for name, T, m, w, t, keep, gen in cases:
        n = T.size
        big = n > 16384
        inp = ref.precompute(T, m)
        nit = (n - m + w) // w
        d = {"m": m, "w": w, "t": t}
        if gen is None:
            d["T"] = T
        else:   # synthetic series are regenerated from the seed by the tests; a checksum guards the generator
            d["gen"] = np.array(gen[0]); d["gen_n"] = n; d["gen_seed"] = gen[1]
            d["T_sum"] = T.sum(); d["T_head"] = T[:4].copy(); d["T_tail"] = T[-4:].copy()
        res = {}
        for kind in ("vanilla", "tiled"):
            MP, MPI, corr, cells = ref.run(T, m, w, kind, t)
            res[kind] = (MP, MPI, corr)
            if big and kind == "tiled":   # Vanilla and Tiled agree bit for bit on these inputs: store one copy
                assert all(np.array_equal(a, b) for a, b in zip(res["vanilla"], res["tiled"]))
                d["tiled_equals_vanilla"] = True
                continue
            d[f"{kind}_MPI"], d[f"{kind}_corr"] = MPI, corr
            if not big:
                d[f"{kind}_MP"] = MP
            if keep:
                for it in range(nit):
                    rv, ri, cv, ci = fields(ref.kernel(kind, n, m, w, it, inp, t))
                    d[f"{kind}_chunk{it}_rv"], d[f"{kind}_chunk{it}_ri"] = rv.copy(), ri.copy()
                    d[f"{kind}_chunk{it}_cv"], d[f"{kind}_chunk{it}_ci"] = cv.copy(), ci.copy()
        sMP, sMPI = ref.scamp(T, m)
        if big:
            d["scamp_max_abs_dMP"] = np.abs(sMP - res["vanilla"][0]).max()
            d["scamp_idx_mismatches"] = int((sMPI != res["vanilla"][1]).sum())
        else:
            d["scamp_MP"], d["scamp_MPI"] = sMP, sMPI
        if n <= 16384:
            d["input"] = inp
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
        print(name, "n", n, "chunks", nit, "min corr %.4f" % d["vanilla_corr"].min(), flush=True)

    # float: reference float kernels on double precompute cast to float (what the host does,
    # include/host/HostSideComputation.hpp:30-41).  Unpinned upstream (no float tests); recorded
    # so the fp32 engine can be compared with the reference's own float path.
    if only:
        return
    orc = Oracle()
    for name, T, m, w, t in [("f32_b1024_m16_w256_t128", T1024, 16, 256, 128), ("f32_walk4096_m64_w256_t128", random_walk(4096, 7), 64, 256, 128)]:
        n = T.size
        inp = orc.precompute(T, m, "f32")
        assert inp.dtype == IN_F32
        d = {"T": T, "m": m, "w": w, "t": t, "input": inp}
        nit = (n - m + w) // w
        for kind in ("vanilla", "tiled"):
            for it in range(nit):
                rv, ri, cv, ci = fields(ref.kernel(kind, n, m, w, it, inp, t))
                d[f"{kind}_chunk{it}_rv"], d[f"{kind}_chunk{it}_ri"] = rv.copy(), ri.copy()
                d[f"{kind}_chunk{it}_cv"], d[f"{kind}_chunk{it}_ci"] = cv.copy(), ci.copy()
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
        print(name, flush=True)


if __name__ == "__main__":
    main()
