"""mpb_kernel_tlf_enqueue_noalloc (include/mp_b200.h): the allocation-free variant of the chunk kernel entry, whose work space is
a static device array inside libmpb200.so.  Collected last on purpose.  Checked through the
C ABI against the existing entry point (mpb_kernel_tlf_enqueue) on the same device buffers: the two must write the same bytes
-- the chunk contract is bit-exact (tests/test_gpu_parity.py::test_tlf_chunks_bit_exact), so the tolerance is zero -- on two
calls in a row (the second starts from the array the first one left behind)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _input_packs(mp, T, m, dtype):
    """InputDataPack[n-m+1] of T from the engine's own precompute (kernel 1), through the C ABI."""
    ctx = mp.Context(0)
    plan = mp.Plan(ctx, T.size, m, mp.F64, mp.FLAG_STRICT)
    plan.upload_series(T)
    plan.precompute()
    packs = plan.download_input()
    plan.close()
    ctx.close()
    if dtype == mp.F32:
        f = np.zeros(packs.size, mp.IN_F32)
        for k in ("QT", "df", "dg", "inv"):
            f[k] = packs[k].astype(np.float32)
        packs = f
    return packs


@pytest.mark.parametrize("dtype_name", ["f64", "f32"])
def test_noalloc_entry_equals_the_allocating_entry_twice(dtype_name):
    import mpb200 as mp
    dtype = mp.F64 if dtype_name == "f64" else mp.F32
    out_t = mp.OUT_F64 if dtype == mp.F64 else mp.OUT_F32
    rng = np.random.default_rng(77)
    ctx = mp.Context(0)
    try:
        # two different problems back to back: the second call meets the first call's leftovers in the shared array
        for n, m, w, it in [(20000, 64, 1024, 3), (12000, 100, 512, 0)]:
            T = np.cumsum(rng.standard_normal(n))
            packs = _input_packs(mp, T, m, dtype)
            sublen = n - m + 1
            nr = sublen - it * w
            bin_ = ctx.make_buffer(packs.nbytes)
            bout_a = ctx.make_buffer(sublen * out_t.itemsize)
            bout_b = ctx.make_buffer(sublen * out_t.itemsize)
            bin_.copy_from_host(packs)
            want = np.zeros(sublen, out_t)
            got = np.zeros(sublen, out_t)
            ctx.kernel_tlf_enqueue(dtype, n, m, w, it, bin_, bout_a)                       # existing entry (allocates its work space)
            bout_a.copy_to_host(want)
            for call in range(2):
                ctx.kernel_tlf_enqueue(dtype, n, m, w, it, bin_, bout_b, noalloc=True)     # static work space
                bout_b.copy_to_host(got)
                for agg in ("rowAggregate", "columnAggregate"):
                    assert np.array_equal(got[agg]["value"][:nr], want[agg]["value"][:nr]), (n, call, agg)
                    assert np.array_equal(got[agg]["index"][:nr], want[agg]["index"][:nr]), (n, call, agg)
                got[:] = 0
            assert np.isfinite(want["rowAggregate"]["value"][:nr]).all() and (want["rowAggregate"]["value"][:nr] > 0).any()
            for b in (bin_, bout_a, bout_b):
                b.close()
    finally:
        ctx.close()


def test_noalloc_entry_rejects_sizes_beyond_the_static_array():
    import mpb200 as mp
    ctx = mp.Context(0)
    try:
        n, m, w = 1 << 20, 256, 4096                  # needs ~200 MB of work space: far beyond the static array
        small = ctx.make_buffer(64)
        with pytest.raises(mp.MpbError) as e:
            ctx.kernel_tlf_enqueue(mp.F64, n, m, w, 0, small, small, noalloc=True)
        assert e.value.code == -1 and "MPB_TLF_NOALLOC_SCRATCH_BYTES" in str(e.value)
        small.close()
    finally:
        ctx.close()
