"""mpb200 -- thin Python binding (ctypes) over the C ABI of the B200 matrix-profile engine
(include/mp_b200.h, libmpb200.so).  It mirrors the reference's interface for the hot path:

* ``MatrixProfileKernelTLF(n, m, iteration, input, output)`` -- the kernel-level contract
  (reference kernel/MatrixProfileKernelVanilla.cpp:16-17; plugin interface
  test/include/MatrixProfileTests.hpp:67), with the configure-time ``w`` / ``data_t`` as keywords;
* ``Context`` / ``Buffer`` -- what the host driver uses of include/host/OpenCL.hpp
  (host/MatrixProfileHost.cpp:69-97,110-122);
* ``matrix_profile(T, m)`` / ``Plan`` -- the device section of RunMatrixProfileKernel
  (host/MatrixProfileHost.cpp:66-142) as one call.

There is no CPU fallback: a missing library or device raises.  PyTorch is not imported here; the
multi-GPU plumbing that needs torch.distributed lives in mpb200.dist.
"""
import atexit
import ctypes as C
import os
import weakref

import numpy as np

F64, F32 = 0, 1
FLAG_STRICT, FLAG_SEQ_MEAN, FLAG_NO_RESEED, FLAG_NO_SEED, FLAG_NO_CHAIN_CARRY, FLAG_WINDOW_MEAN, FLAG_MU_WALK = 1, 2, 4, 8, 16, 32, 64
_LIB = None
TLF_NOALLOC_SCRATCH_BYTES = 16 << 20     # MPB_TLF_NOALLOC_SCRATCH_BYTES (include/mp_b200.h)
_PKG_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))   # matrix-profile-hls_b200/
LIB_PATH = os.environ.get("MPB200_LIB") or os.path.join(_PKG_ROOT, "libmpb200.so")   # MPB200_LIB: tuning builds (csrc/Makefile VARIANT=)

# record layouts: reference include/Config.hpp.in:43-73
AGG_F64 = np.dtype([("value", "<f8"), ("index", "<i4"), ("_pad", "<i4")])
IN_F64 = np.dtype([("QT", "<f8"), ("df", "<f8"), ("dg", "<f8"), ("inv", "<f8")])
OUT_F64 = np.dtype([("rowAggregate", AGG_F64), ("columnAggregate", AGG_F64)])
AGG_F32 = np.dtype([("value", "<f4"), ("index", "<i4")])
IN_F32 = np.dtype([("QT", "<f4"), ("df", "<f4"), ("dg", "<f4"), ("inv", "<f4")])
OUT_F32 = np.dtype([("rowAggregate", AGG_F32), ("columnAggregate", AGG_F32)])


class MpbError(RuntimeError):
    """Non-zero status from the C ABI (the reference throws ConfigurationError / RuntimeError, OpenCL.hpp:48-60)."""

    def __init__(self, code, msg):
        super().__init__(f"mpb200 error {code}: {msg}")
        self.code = code


class Timings(C.Structure):
    _fields_ = [("h2d_ms", C.c_double), ("precompute_ms", C.c_double), ("compute_ms", C.c_double), ("finalize_ms", C.c_double),
                ("d2h_ms", C.c_double), ("total_ms", C.c_double), ("cells", C.c_int64), ("slow_updates", C.c_int64),
                ("launches", C.c_int32), ("bands", C.c_int32), ("items", C.c_int32), ("segment_rows", C.c_int32), ("filter_hits", C.c_int64), ("replays", C.c_int64),
                ("chain_candidates", C.c_int64), ("chain_max_correction", C.c_double), ("chain_bound", C.c_double), ("chain_margin", C.c_double),
                ("cells_walked", C.c_int64), ("reseeded_diagonals", C.c_int64), ("replay_lane_blocks", C.c_int64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


EXPORTS = [
    "mpb_last_error", "mpb_version", "mpb_device_count", "mpb_context_create", "mpb_context_destroy", "mpb_context_finish",
    "mpb_context_stream", "mpb_context_device", "mpb_buffer_create", "mpb_buffer_destroy", "mpb_buffer_copy_from_host",
    "mpb_buffer_copy_to_host", "mpb_buffer_device_ptr", "mpb_buffer_size", "mpb_kernel_tlf_enqueue", "mpb_kernel_tlf_enqueue_noalloc", "mpb_kernel_tlf_f64",
    "mpb_kernel_tlf_f32", "mpb_compute", "mpb_plan_create", "mpb_plan_destroy", "mpb_plan_set_shard", "mpb_plan_set_segment_rows", "mpb_plan_segment_rows", "mpb_plan_set_sketch_bands", "mpb_plan_upload_series",
    "mpb_plan_series_device_ptr", "mpb_plan_precompute", "mpb_plan_compute", "mpb_plan_compute_bands", "mpb_plan_local_bands", "mpb_plan_export", "mpb_plan_mask_index",
    "mpb_plan_import", "mpb_plan_share_thresholds", "mpb_plan_shared_with", "mpb_plan_connect_peers", "mpb_plan_set_chain_margin", "mpb_plan_set_log_entries", "mpb_plan_merge_peers", "mpb_plan_finalize", "mpb_plan_download", "mpb_plan_mp_device_ptr",
    "mpb_plan_mpi_device_ptr", "mpb_plan_download_input", "mpb_plan_replay_histogram", "mpb_plan_check", "mpb_plan_timings", "mpb_plan_mu_walked", "mpb_comm_unique_id", "mpb_comm_create", "mpb_comm_destroy",
    "mpb_comm_rank", "mpb_comm_world", "mpb_plan_allreduce", "mpb_plan_compute_merged", "mpb_merge_point", "mpb_plan_total_bands",
    "mpb_plan_sketch_share", "mpb_plan_hint_device_ptr", "mpb_plan_hint_count", "mpb_plan_apply_hint", "mpb_compute_multi", "mpb_useful_cells", "mpb_measure_fma_peak",
]


def lib():
    """Load libmpb200.so (built in-tree by __graft_entry__.build()); fail loudly if it is missing."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise MpbError(-2, f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` (there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.mpb_last_error.restype = C.c_char_p
        L.mpb_version.restype = C.c_char_p
        for f in ("mpb_context_stream", "mpb_buffer_device_ptr", "mpb_plan_series_device_ptr", "mpb_plan_mp_device_ptr", "mpb_plan_mpi_device_ptr",
                  "mpb_plan_hint_device_ptr"):
            getattr(L, f).restype = C.c_void_p
            getattr(L, f).argtypes = [C.c_void_p]
        L.mpb_buffer_size.restype = C.c_size_t
        L.mpb_useful_cells.restype = C.c_int64
        L.mpb_useful_cells.argtypes = [C.c_int, C.c_int]
        _LIB = L
    return _LIB


def _chk(rc):
    if rc != 0:
        raise MpbError(rc, lib().mpb_last_error().decode())


def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


def device_count():
    n = C.c_int(0)
    _chk(lib().mpb_device_count(C.byref(n)))
    return n.value


def useful_cells(n, m):
    return int(lib().mpb_useful_cells(n, m))


_LIVE_CONTEXTS = weakref.WeakSet()


@atexit.register
def _close_all():
    """Release device state while the interpreter and the CUDA runtime are still intact (object finalisers run in
    arbitrary order at shutdown; the library also ignores handles that are already gone)."""
    for ctx in list(_LIVE_CONTEXTS):
        ctx.close()


class Context:
    """Replaces OpenCL::Context (OpenCL.hpp:73-107): one device + one command stream."""

    def __init__(self, ordinal=0):
        self._h = C.c_void_p()
        self._children = weakref.WeakSet()    # plans / buffers must be destroyed before their context
        _chk(lib().mpb_context_create(C.c_int(ordinal), C.byref(self._h)))
        _LIVE_CONTEXTS.add(self)

    def finish(self):
        """Context::Finish(): returns device nanoseconds since the previous finish()."""
        ns = C.c_int64(0)
        _chk(lib().mpb_context_finish(self._h, C.byref(ns)))
        return ns.value

    @property
    def stream(self):
        return lib().mpb_context_stream(self._h)

    @property
    def device(self):
        return lib().mpb_context_device(self._h)

    def make_buffer(self, nbytes):
        return Buffer(self, nbytes)

    def kernel_tlf_enqueue(self, dtype, n, m, w, iteration, input_buf, output_buf, *, noalloc=False):
        """Kernel::EnqueueTask() of the chunk kernel on device buffers (asynchronous).  noalloc=True takes the entry point whose
        work space is the library's static device array (no cudaMalloc inside; size limit TLF_NOALLOC_SCRATCH_BYTES)."""
        fn = lib().mpb_kernel_tlf_enqueue_noalloc if noalloc else lib().mpb_kernel_tlf_enqueue
        _chk(fn(self._h, C.c_int(dtype), C.c_int(n), C.c_int(m), C.c_int(w), C.c_int(iteration), input_buf._h, output_buf._h))

    def measure_fma_peak(self, dtype=F64, iters=1 << 15):
        r, ms = C.c_double(0), C.c_double(0)
        _chk(lib().mpb_measure_fma_peak(self._h, C.c_int(dtype), C.c_int(iters), C.byref(r), C.byref(ms)))
        return r.value, ms.value

    def close(self):
        if self._h:
            for child in list(self._children):
                child.close()
            lib().mpb_context_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Buffer:
    """Replaces OpenCL::Buffer<T,Access> (OpenCL.hpp:109-211)."""

    def __init__(self, ctx, nbytes):
        self.ctx = ctx
        self._h = C.c_void_p()
        _chk(lib().mpb_buffer_create(ctx._h, C.c_size_t(nbytes), C.byref(self._h)))
        ctx._children.add(self)
        self.nbytes = nbytes

    def copy_from_host(self, arr):
        arr = np.ascontiguousarray(arr)
        _chk(lib().mpb_buffer_copy_from_host(self._h, _vp(arr), C.c_size_t(arr.nbytes)))
        self._keep = arr   # asynchronous: keep the source alive until finish()

    def copy_to_host(self, arr):
        assert arr.flags["C_CONTIGUOUS"]
        _chk(lib().mpb_buffer_copy_to_host(self._h, _vp(arr), C.c_size_t(arr.nbytes)))
        return arr

    @property
    def ptr(self):
        return lib().mpb_buffer_device_ptr(self._h)

    def close(self):
        if self._h:
            lib().mpb_buffer_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def MatrixProfileKernelTLF(n, m, iteration, input, output, *, w):
    """The reference's kernel contract with host arrays: ``input`` is InputDataPack[n-m+1]
    (IN_F64 / IN_F32), ``output`` OutputDataPack[>= nRows]; output[0:nRows] is overwritten."""
    if input.dtype == IN_F64 and output.dtype == OUT_F64:
        fn = lib().mpb_kernel_tlf_f64
    elif input.dtype == IN_F32 and output.dtype == OUT_F32:
        fn = lib().mpb_kernel_tlf_f32
    else:
        raise MpbError(-1, "input/output must be IN_F64/OUT_F64 or IN_F32/OUT_F32 record arrays")
    if input.size != n - m + 1 or output.size < n - m + 1 - iteration * w:
        raise MpbError(-1, "input must hold n-m+1 packs and output at least nRows packs")
    _chk(fn(C.c_int(n), C.c_int(m), C.c_int(w), C.c_int(iteration), _vp(input), _vp(output)))
    return output


class Plan:
    """Whole-problem engine state resident in HBM (C ABI mpb_plan_*)."""

    def __init__(self, ctx, n, m, dtype=F64, flags=0):
        self.ctx, self.n, self.m, self.dtype = ctx, n, m, dtype
        self.sublen = n - m + 1
        self._h = C.c_void_p()
        _chk(lib().mpb_plan_create(ctx._h, C.c_int(n), C.c_int(m), C.c_int(dtype), C.c_uint(flags), C.byref(self._h)))
        ctx._children.add(self)

    def set_shard(self, rank, world):
        _chk(lib().mpb_plan_set_shard(self._h, C.c_int(rank), C.c_int(world)))

    def set_segment_rows(self, rows):
        """Rows per re-seeded segment (multiple of the tile height, 192; 0 = never; -1 = automatic)."""
        _chk(lib().mpb_plan_set_segment_rows(self._h, C.c_int(rows)))

    def set_sketch_bands(self, bands):
        """Local bands covered by the sketch pass before the first tile launch of a step (0 = off; -1 = default)."""
        _chk(lib().mpb_plan_set_sketch_bands(self._h, C.c_int(bands)))

    @property
    def segment_rows(self):
        return lib().mpb_plan_segment_rows(self._h)

    def upload_series(self, T):
        T = np.ascontiguousarray(T, dtype=np.float64)
        assert T.size == self.n
        self._keep = T
        _chk(lib().mpb_plan_upload_series(self._h, _vp(T)))

    def upload_series_ptr(self, host_ptr):
        _chk(lib().mpb_plan_upload_series(self._h, C.c_void_p(host_ptr)))

    @property
    def series_ptr(self):
        return lib().mpb_plan_series_device_ptr(self._h)

    @property
    def mp_ptr(self):
        return lib().mpb_plan_mp_device_ptr(self._h)

    @property
    def mpi_ptr(self):
        return lib().mpb_plan_mpi_device_ptr(self._h)

    def precompute(self):
        _chk(lib().mpb_plan_precompute(self._h))

    def compute(self):
        _chk(lib().mpb_plan_compute(self._h))

    def compute_bands(self, lb_begin, lb_end=-1):
        """Kernel (2) over local bands [lb_begin, lb_end) only (lb_end < 0 = to the end)."""
        _chk(lib().mpb_plan_compute_bands(self._h, C.c_int(lb_begin), C.c_int(lb_end)))

    def local_bands(self):
        c = C.c_int(0)
        _chk(lib().mpb_plan_local_bands(self._h, C.byref(c)))
        return c.value

    def export(self, corr_ptr, index_ptr):
        _chk(lib().mpb_plan_export(self._h, C.c_void_p(corr_ptr), C.c_void_p(index_ptr)))

    def mask_index(self, corr_max_ptr, index_ptr):
        _chk(lib().mpb_plan_mask_index(self._h, C.c_void_p(corr_max_ptr), C.c_void_p(index_ptr)))

    def import_(self, corr_ptr, index_ptr):
        _chk(lib().mpb_plan_import(self._h, C.c_void_p(corr_ptr), C.c_void_p(index_ptr)))

    def share_thresholds(self, others):
        """Push this plan's profile improvements to the threshold arrays of `others` too (same process; one direction)."""
        arr = (C.c_void_p * len(others))(*[o._h for o in others])
        _chk(lib().mpb_plan_share_thresholds(self._h, arr, C.c_int(len(others))))

    def connect_peers(self, comm):
        """Collective: map the threshold arrays of all ranks of `comm` (P2P / CUDA IPC) and push improvements to them."""
        _chk(lib().mpb_plan_connect_peers(self._h, comm._h))

    @property
    def shared_with(self):
        return int(lib().mpb_plan_shared_with(self._h))

    def set_chain_margin(self, log2_inverse):
        _chk(lib().mpb_plan_set_chain_margin(self._h, C.c_int(log2_inverse)))

    def set_log_entries(self, entries):
        _chk(lib().mpb_plan_set_log_entries(self._h, C.c_int64(entries)))

    def sketch_share(self, band, share_rank, share_world):
        """This rank's share (row segments share_rank, share_rank + share_world, ...) of the sketch of global band `band` into hint[]."""
        _chk(lib().mpb_plan_sketch_share(self._h, C.c_int(band), C.c_int(share_rank), C.c_int(share_world)))

    @property
    def hint_ptr(self):
        return lib().mpb_plan_hint_device_ptr(self._h)

    @property
    def hint_count(self):
        return int(lib().mpb_plan_hint_count(self._h))

    def apply_hint(self):
        _chk(lib().mpb_plan_apply_hint(self._h))

    def total_bands(self):
        c = C.c_int(0)
        _chk(lib().mpb_plan_total_bands(self._h, C.byref(c)))
        return c.value

    def allreduce(self, comm):
        """Exact (corr, index) max of the ranks' result profiles over NCCL, in place (thresholds rise to the merged values)."""
        _chk(lib().mpb_plan_allreduce(self._h, comm._h))

    def compute_merged(self, comm):
        """One rank's compute step of a multi-GPU job: shared sketch, first bands, threshold merge, rest, result merge."""
        _chk(lib().mpb_plan_compute_merged(self._h, comm._h))

    def merge_peers(self, others):
        arr = (C.c_void_p * len(others))(*[o._h for o in others])
        _chk(lib().mpb_plan_merge_peers(self._h, arr, C.c_int(len(others))))

    def finalize(self):
        _chk(lib().mpb_plan_finalize(self._h))

    def download(self, want_corr=True):
        MP = np.empty(self.sublen, np.float64)
        MPI = np.empty(self.sublen, np.int32)
        corr = np.empty(self.sublen, np.float64) if want_corr else None
        _chk(lib().mpb_plan_download(self._h, _vp(MP), _vp(MPI), _vp(corr) if want_corr else C.c_void_p()))
        return MP, MPI, corr

    def download_into(self, mp_host_ptr, mpi_host_ptr):
        _chk(lib().mpb_plan_download(self._h, C.c_void_p(mp_host_ptr), C.c_void_p(mpi_host_ptr), C.c_void_p()))

    def download_input(self):
        out = np.empty(self.sublen, IN_F64 if self.dtype == F64 else IN_F32)
        _chk(lib().mpb_plan_download_input(self._h, _vp(out)))
        return out

    def replay_histogram(self):
        h = (C.c_int64 * 33)()
        _chk(lib().mpb_plan_replay_histogram(self._h, h))
        return list(h)

    def check(self):
        """Waits for the plan's work; raises if the near-tie log overflowed or the corrections exceed the margin's reach."""
        _chk(lib().mpb_plan_check(self._h))

    def mu_walked(self):
        w = C.c_int(0)
        _chk(lib().mpb_plan_mu_walked(self._h, C.byref(w)))
        return bool(w.value)

    def timings(self):
        t = Timings()
        _chk(lib().mpb_plan_timings(self._h, C.byref(t)))
        return t.as_dict()

    def close(self):
        if self._h:
            lib().mpb_plan_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


COMM_ID_BYTES = 128


def comm_unique_id():
    """ncclGetUniqueId: 128 bytes to be handed to every rank of the job (e.g. torch.distributed.broadcast)."""
    buf = (C.c_ubyte * COMM_ID_BYTES)()
    _chk(lib().mpb_comm_unique_id(buf, C.c_size_t(COMM_ID_BYTES)))
    return bytes(buf)


def merge_point(total_bands, world):
    return int(lib().mpb_merge_point(C.c_int(total_bands), C.c_int(world)))


class Comm:
    """One rank of a multi-GPU job: an NCCL communicator on the context's device (C ABI mpb_comm_*)."""

    def __init__(self, ctx, unique_id, rank, world):
        assert len(unique_id) == COMM_ID_BYTES
        self.ctx, self.rank, self.world = ctx, rank, world
        self._h = C.c_void_p()
        _chk(lib().mpb_comm_create(ctx._h, C.c_char_p(bytes(unique_id)), C.c_int(rank), C.c_int(world), C.byref(self._h)))
        ctx._children.add(self)

    def close(self):
        if self._h:
            lib().mpb_comm_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def matrix_profile_multi(T, m, ngpus, dtype=F64, flags=0, devices=None, return_timings=False):
    """mpb_compute_multi: HOST buffers, ngpus GPUs of this process (one host thread each), NCCL merge."""
    T = np.ascontiguousarray(T, dtype=np.float64)
    sublen = T.size - m + 1
    MP = np.empty(max(sublen, 0), np.float64)
    MPI = np.empty(max(sublen, 0), np.int32)
    tms = (Timings * ngpus)()
    dev = (C.c_int * ngpus)(*devices) if devices is not None else None
    _chk(lib().mpb_compute_multi(C.c_int(ngpus), dev, _vp(T), C.c_int(T.size), C.c_int(m), C.c_int(dtype), C.c_uint(flags), _vp(MP), _vp(MPI), tms))
    return (MP, MPI, [t.as_dict() for t in tms]) if return_timings else (MP, MPI)


def matrix_profile(T, m, dtype=F64, flags=0, ctx=None, return_timings=False):
    """Host buffers in, host buffers out: T[n] float64 -> (MP[n-m+1] float64 distance, MPI[n-m+1] int32).
    One call into mpb_compute (H2D, kernels (1)-(3), D2H)."""
    T = np.ascontiguousarray(T, dtype=np.float64)
    own = ctx is None
    ctx = ctx or Context(0)
    sublen = T.size - m + 1
    MP = np.empty(max(sublen, 0), np.float64)
    MPI = np.empty(max(sublen, 0), np.int32)
    t = Timings()
    _chk(lib().mpb_compute(ctx._h, _vp(T), C.c_int(T.size), C.c_int(m), C.c_int(dtype), C.c_uint(flags), _vp(MP), _vp(MPI), C.byref(t)))
    if own:
        ctx.close()
    return (MP, MPI, t.as_dict()) if return_timings else (MP, MPI)
