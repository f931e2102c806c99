"""Host-side logic of bench.py and of the boundary that needs no GPU: the metric's cell accounting, the bounded CPU sample, the
C-ness of the public header, and the Timings mirror of mpb_timings."""
import ctypes
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "matrix-profile-hls_b200", "python"))
import bench  # noqa: E402


def test_chunk_cells_add_up_to_the_useful_cells():
    """cells/s is quoted on L(L+1)/2 useful cells; the per-chunk figure the CPU legs use must partition exactly that."""
    for n, m, w in [(1 << 20, 256, 4096), (65536, 256, 4096), (16384, 128, 1024), (1000, 30, 64), (4096, 32, 256)]:
        nit = (n - m + w) // w
        assert sum(bench.chunk_useful_cells(n, m, w, it) for it in range(nit)) == bench.useful_cells(n, m), (n, m, w)
    assert bench.chunk_useful_cells(1024, 16, 32, 0) == sum((1024 - 16 + 1) - d for d in range(4, 32))   # diagonals 4..31 of chunk 0 (0..3 excluded)


def test_cpu_sample_is_a_pure_function_of_its_arguments():
    """The cpu_baseline leg and the --impl reference arm must time the same chunks for a given thread count."""
    a = bench.reference_sample_chunks(1 << 20, 256, 16, 8.0)
    b = bench.reference_sample_chunks(1 << 20, 256, 16, 8.0)
    assert a == b
    w, nit, its = a
    assert w == 4096 and nit == 256 and len(its) == 16 and its == list(range(its[0], its[0] + 16)) and 0 <= its[0] and its[-1] < nit
    w, nit, its = bench.reference_sample_chunks(4096, 32, 64, 8.0)       # small problems: fewer chunks than threads, narrower chunks
    assert w == 256 and 1 <= len(its) <= nit and its[-1] < nit


def test_public_header_is_plain_c(tmp_path):
    src = tmp_path / "hdr.c"
    src.write_text('#include "mp_b200.h"\nint main(void) { mpb_timings t; (void)t; return MPB_TLF_NOALLOC_SCRATCH_BYTES > 0 ? 0 : 1; }\n')
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), str(src)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_timings_mirror_matches_the_header():
    """python/mpb200 Timings must list the fields of mpb_timings in the header's order (the struct crosses the ABI by pointer)."""
    import mpb200
    hdr = open(os.path.join(ROOT, "include", "mp_b200.h")).read()
    body = hdr[hdr.index("typedef struct {"):hdr.index("} mpb_timings;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = []
    for decl in body.split(";"):
        decl = decl.replace("typedef struct {", "").strip()
        if not decl:
            continue
        ctype, names = decl.split(None, 1)
        for nm in names.split(","):
            fields.append((nm.strip(), ctype))
    mirror = [(n, {ctypes.c_double: "double", ctypes.c_int64: "int64_t", ctypes.c_int32: "int32_t"}[t]) for n, t in mpb200.Timings._fields_]
    assert mirror == fields
    assert ctypes.sizeof(mpb200.Timings) == sum(8 if t != "int32_t" else 4 for _, t in fields)     # no padding: the int32 fields come in pairs
