"""ctypes bindings for the CPU oracle (oracle/libmporacle.so) and, when built, the reference shim
(oracle/_ref/libmpref.so).  TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs -- never from the product package.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

# record layouts: reference include/Config.hpp.in:43-73
AGG_F64 = np.dtype([("value", "<f8"), ("index", "<i4"), ("_pad", "<i4")])
IN_F64 = np.dtype([("QT", "<f8"), ("df", "<f8"), ("dg", "<f8"), ("inv", "<f8")])
OUT_F64 = np.dtype([("rowAggregate", AGG_F64), ("columnAggregate", AGG_F64)])
AGG_F32 = np.dtype([("value", "<f4"), ("index", "<i4")])
IN_F32 = np.dtype([("QT", "<f4"), ("df", "<f4"), ("dg", "<f4"), ("inv", "<f4")])
OUT_F32 = np.dtype([("rowAggregate", AGG_F32), ("columnAggregate", AGG_F32)])
assert AGG_F64.itemsize == 16 and IN_F64.itemsize == 32 and OUT_F64.itemsize == 32
assert AGG_F32.itemsize == 8 and IN_F32.itemsize == 16 and OUT_F32.itemsize == 16


def build_oracle():
    subprocess.run(["make", "-s", "-C", HERE, "oracle"], check=True)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    """The plain-C restatement (mp_oracle.c). fast=True loads the -O3 -march=native build (timing only)."""

    def __init__(self, fast=False):
        name = "libmporacle_fast.so" if fast else "libmporacle.so"
        path = os.path.join(HERE, name)
        if not os.path.exists(path):
            build_oracle()
        self.lib = C.CDLL(path)
        L = self.lib
        for suf in ("f64", "f32"):
            getattr(L, f"mpo_run_{suf}").restype = C.c_int64
            getattr(L, f"mpo_run_{suf}").argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                                     C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.mpo_exact_corr.restype = C.c_double
        L.mpo_exact_corr.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]

    def precompute(self, T, m, dtype="f64"):
        T = np.ascontiguousarray(T, dtype=np.float64)
        n = T.size
        out = np.zeros(n - m + 1, dtype=IN_F64 if dtype == "f64" else IN_F32)
        getattr(self.lib, f"mpo_precompute_{dtype}")(_ptr(T), C.c_int(n), C.c_int(m), _ptr(out))
        return out

    def kernel(self, kind, n, m, w, iteration, inp, t=0):
        """One diagonal chunk. kind: 'vanilla' | 'tiled'. Returns OutputDataPack[nRows]."""
        f64 = inp.dtype == IN_F64
        sublen = n - m + 1
        out = np.zeros(sublen, dtype=OUT_F64 if f64 else OUT_F32)
        suf = "f64" if f64 else "f32"
        if kind == "vanilla":
            getattr(self.lib, f"mpo_kernel_vanilla_{suf}")(C.c_int(n), C.c_int(m), C.c_int(w), C.c_int(iteration), _ptr(inp), _ptr(out))
        else:
            getattr(self.lib, f"mpo_kernel_tiled_{suf}")(C.c_int(n), C.c_int(m), C.c_int(w), C.c_int(t), C.c_int(iteration), _ptr(inp), _ptr(out))
        return out[: sublen - iteration * w]

    def run(self, T, m, w, kind="vanilla", t=0, dtype="f64", nthreads=1, max_chunks=-1):
        """Whole driver flow (host/MatrixProfileHost.cpp:54-160). Returns (MP, MPI, corr, cells)."""
        T = np.ascontiguousarray(T, dtype=np.float64)
        n = T.size
        sublen = n - m + 1
        MP = np.zeros(sublen, np.float64)
        MPI = np.zeros(sublen, np.int32)
        corr = np.zeros(sublen, np.float64)
        cells = getattr(self.lib, f"mpo_run_{dtype}")(_ptr(T), n, m, w, t, 0 if kind == "vanilla" else 1, nthreads, max_chunks,
                                                      _ptr(MP), _ptr(MPI), _ptr(corr))
        return MP, MPI, corr, int(cells)

    def reference_scamp(self, T, m):
        T = np.ascontiguousarray(T, dtype=np.float64)
        sublen = T.size - m + 1
        MP = np.zeros(sublen, np.float64)
        MPI = np.zeros(sublen, np.int32)
        corr = np.zeros(sublen, np.float64)
        self.lib.mpo_reference_scamp_f64(_ptr(T), C.c_int(T.size), C.c_int(m), _ptr(MP), _ptr(MPI), _ptr(corr))
        return MP, MPI, corr

    def exact_corr(self, T, m, i, j):
        T = np.ascontiguousarray(T, dtype=np.float64)
        return float(self.lib.mpo_exact_corr(_ptr(T), T.size, m, int(i), int(j)))


class RefShim:
    """The reference's own sources compiled by oracle/Makefile `ref` into oracle/_ref (needs
    /root/reference at build time only)."""

    @staticmethod
    def available(fast=False):
        return os.path.exists(os.path.join(HERE, "_ref", "libmpref_fast.so" if fast else "libmpref.so"))

    def __init__(self, fast=False):
        self.lib = C.CDLL(os.path.join(HERE, "_ref", "libmpref_fast.so" if fast else "libmpref.so"))

    def precompute(self, T, m):
        T = np.ascontiguousarray(T, dtype=np.float64)
        out = np.zeros(T.size - m + 1, dtype=IN_F64)
        rc = self.lib.ref_precompute_f64(C.c_int(T.size), C.c_int(m), _ptr(T), _ptr(out))
        if rc != 0:
            raise KeyError(f"(n={T.size}, m={m}) not instantiated in ref_shim_host.cpp")
        return out

    def scamp(self, T, m):
        T = np.ascontiguousarray(T, dtype=np.float64)
        sublen = T.size - m + 1
        MP = np.zeros(sublen, np.float64)
        MPI = np.zeros(sublen, np.int32)
        rc = self.lib.ref_scamp_f64(C.c_int(T.size), C.c_int(m), _ptr(T), _ptr(MP), _ptr(MPI))
        if rc != 0:
            raise KeyError(f"(n={T.size}, m={m}) not instantiated in ref_shim_host.cpp")
        return MP, MPI

    def kernel(self, kind, n, m, w, iteration, inp, t=0):
        f64 = inp.dtype == IN_F64
        sublen = n - m + 1
        out = np.zeros(sublen, dtype=OUT_F64 if f64 else OUT_F32)
        suf = "f64" if f64 else "f32"
        if kind == "vanilla":
            rc = getattr(self.lib, f"ref_vanilla_{suf}")(C.c_int(w), C.c_int(n), C.c_int(m), C.c_int(iteration), _ptr(inp), _ptr(out))
        else:
            rc = getattr(self.lib, f"ref_tiled_{suf}")(C.c_int(w), C.c_int(t), C.c_int(n), C.c_int(m), C.c_int(iteration), _ptr(inp), _ptr(out))
        if rc < 0:
            raise KeyError(f"(w={w}, t={t}) not instantiated in the ref shim")
        if rc != 0:
            raise RuntimeError(f"reference kernel stream check failed (rc={rc})")
        return out[: sublen - iteration * w]

    def run(self, T, m, w, kind="vanilla", t=0, max_chunks=-1):
        """Reference kernels driven by the reference's chunk loop + merges (host/MatrixProfileHost.cpp:101-142),
        merge arithmetic in numpy (strict >, earlier chunk wins; column wins row on ties)."""
        T = np.ascontiguousarray(T, dtype=np.float64)
        n = T.size
        sublen = n - m + 1
        try:
            inp = self.precompute(T, m)
        except KeyError:
            inp = Oracle().precompute(T, m)   # bit-identical restatement (pinned by tests/test_oracle.py)
        rv = np.full(sublen, -1e12); ri = np.full(sublen, -1, np.int32)
        cv = np.full(sublen, -1e12); ci = np.full(sublen, -1, np.int32)
        nit = (n - m + w) // w
        if max_chunks >= 0:
            nit = min(nit, max_chunks)
        cells = 0
        for it in range(nit):
            out = self.kernel(kind, n, m, w, it, inp, t)
            off = it * w
            nr = sublen - off
            a = out["rowAggregate"]; b = out["columnAggregate"]
            k = a["value"] > rv[:nr]
            rv[:nr][k] = a["value"][k]; ri[:nr][k] = a["index"][k]
            k = b["value"] > cv[off:]
            cv[off:][k] = b["value"][k]; ci[off:][k] = b["index"][k]
            cells += nr * w
        pick = rv > cv
        corr = np.where(pick, rv, cv)
        MPI = np.where(pick, ri, ci).astype(np.int32)
        MP = np.sqrt(2 * m * (1 - corr))
        return MP, MPI, corr, cells


def random_walk(n, seed=1234):
    """Synthetic random walk: T[0] = N(0,1), T[i] = T[i-1] + N(0,1)."""
    return np.cumsum(np.random.default_rng(seed).standard_normal(n))


def uniform_series(n, seed=1234):
    """`.tsb`-shaped synthetic input: i.i.d. U[0,1) float64 (what the reference's data/binary files hold)."""
    return np.random.default_rng(seed).random(n)
