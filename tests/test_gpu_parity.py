"""GPU parity tests (run with -m gpu on the B200 box): the CUDA engine, called through the C ABI
(include/mp_b200.h via the ctypes binding), against the CPU oracle and the golden vectors that the
reference's own sources produced.  Tolerances (BASELINE.json north_star): fp64 1e-9 relative on the
profile, index exact except ties within that tolerance; fp32 1e-4."""
import numpy as np
import pytest
from conftest import load_golden

pytestmark = pytest.mark.gpu

KAT = ["small128_m16_w32_t16", "b1024_m16_w32_t16", "b1024_m16_w256_t128", "b1024_m32_w512_t128",
       "b1024_m30_w64_t16", "b16384_m128_w1024_t128"]
SYN = ["walk4096_m64_w256_t128", "walk8192_m100_w1024_t256"]
BIG = ["walk65536_m256_w4096_t128", "unif65536_m256_w4096_t128",
       "walk32768_m1024_w4096_t128", "walk20480_m4096_w4096_t128", "unif20480_m32_w4096_t128"]   # m of BASELINE.json configs 4 and 5
RTOL = 1e-9


@pytest.fixture(scope="module")
def mp():
    import mpb200
    assert mpb200.device_count() >= 1
    return mpb200


@pytest.fixture(scope="module")
def ctx(mp):
    c = mp.Context(0)
    yield c
    c.close()


def run_plan(mp, ctx, T, m, dtype=0, flags=0, shard=None):
    plan = mp.Plan(ctx, T.size, m, dtype, flags)
    if shard:
        plan.set_shard(*shard)
    plan.upload_series(T)
    plan.precompute()
    plan.compute()
    return plan


def check_profile(oracle, T, m, corr, MPI, MP, g_corr, g_MPI, rtol=RTOL):
    """value within rtol (relative) on the correlation, and on the distance after the 1/(2(1-P)) amplification
    that the Pearson -> distance map applies; index equal, or the two candidates' exact (long double)
    correlations within the tolerance of each other."""
    assert g_corr.min() > 0                      # the P=0 clamp is never what is being compared
    assert (np.abs(corr - g_corr) / np.abs(g_corr)).max() <= rtol
    g_MP = np.sqrt(2 * m * (1 - g_corr))
    ok = (1 - g_corr) > 1e-6
    assert (np.abs(MP - g_MP)[ok] / g_MP[ok]).max() <= rtol
    bad = np.nonzero(MPI != g_MPI)[0]
    for i in bad:
        a, b = oracle.exact_corr(T, m, i, MPI[i]), oracle.exact_corr(T, m, i, g_MPI[i])
        assert abs(a - b) <= max(rtol, RTOL) * max(abs(a), abs(b)), (i, MPI[i], g_MPI[i], a, b)   # the tie rule keeps the north_star tolerance
    return len(bad)


@pytest.mark.parametrize("flags", [0, 2])
@pytest.mark.parametrize("name", KAT + SYN + BIG)
def test_precompute_bit_exact(mp, ctx, oracle, name, flags):
    """kernel (1) == PreComputeStatistics, bit for bit (incl. the sequential running mean): the fp64 default and
    MPB_FLAG_SEQ_MEAN."""
    g = load_golden(name)
    T, m = g["T"], g["m"]
    plan = mp.Plan(ctx, T.size, m, mp.F64, flags)
    plan.upload_series(T)
    plan.precompute()
    got = plan.download_input()
    want = g["input"] if "input" in g else oracle.precompute(T, m)
    assert got.tobytes() == want.tobytes()
    plan.close()


@pytest.mark.parametrize("name", KAT + SYN + BIG)
def test_precompute_window_mean_close(mp, ctx, oracle, name):
    """MPB_FLAG_WINDOW_MEAN (the fp32 default): exact window means instead of the running recurrence.  df is independent
    of mu (bit-equal); dg, QT, inv agree with PreComputeStatistics to the running mean's own drift."""
    g = load_golden(name)
    T, m = g["T"], g["m"]
    plan = mp.Plan(ctx, T.size, m, mp.F64, mp.FLAG_WINDOW_MEAN)
    plan.upload_series(T)
    plan.precompute()
    got = plan.download_input()
    plan.close()
    want = g["input"] if "input" in g else oracle.precompute(T, m)
    assert np.array_equal(got["df"], want["df"])
    scale = np.abs(T).max()
    assert np.abs(got["dg"] - want["dg"]).max() <= 1e-11 * scale
    assert (np.abs(got["inv"] - want["inv"]) / want["inv"]).max() <= 1e-9
    assert np.abs(got["QT"] - want["QT"]).max() <= 1e-9 * np.abs(want["QT"]).max()


def _check_chunks(mp, g, inp, out, kind_prefix="vanilla"):
    T, m, w = g["T"], g["m"], g["w"]
    n = T.size
    for it in range((n - m + w) // w):
        mp.MatrixProfileKernelTLF(n, m, it, inp, out, w=w)
        nr = n - m + 1 - it * w
        assert np.array_equal(out["rowAggregate"]["value"][:nr], g[f"{kind_prefix}_chunk{it}_rv"]), it
        assert np.array_equal(out["rowAggregate"]["index"][:nr], g[f"{kind_prefix}_chunk{it}_ri"]), it
        assert np.array_equal(out["columnAggregate"]["value"][:nr], g[f"{kind_prefix}_chunk{it}_cv"]), it
        assert np.array_equal(out["columnAggregate"]["index"][:nr], g[f"{kind_prefix}_chunk{it}_ci"]), it


@pytest.mark.parametrize("name", [n for n in KAT if n != "b16384_m128_w1024_t128"])
def test_tlf_chunks_bit_exact(mp, oracle, name):
    """MatrixProfileKernelTLF contract: per-chunk OutputDataPacks equal the reference Vanilla kernel's,
    values and indices (P=0 clamp and tie order included)."""
    g = load_golden(name)
    inp = oracle.precompute(g["T"], g["m"])
    _check_chunks(mp, g, inp, np.zeros(inp.size, mp.OUT_F64))


@pytest.mark.parametrize("name", ["f32_b1024_m16_w256_t128", "f32_walk4096_m64_w256_t128"])
def test_tlf_chunks_bit_exact_float(mp, name):
    g = load_golden(name)
    inp = np.ascontiguousarray(g["input"])
    _check_chunks(mp, g, inp, np.zeros(inp.size, mp.OUT_F32))


@pytest.mark.parametrize("name", KAT)
def test_reference_harness_flow_through_tlf(mp, oracle, name):
    """The reference's TestMatrixProfileKernel flow (test/include/MatrixProfileTests.hpp:141-213) with the
    GPU kernel plugged in: chunk loop, UpdateAggregates, PostCompute, then its own pass criterion
    (|dMP| <= 1e-8 on the distance, identical index) against its independent SCAMP."""
    g = load_golden(name)
    T, m, w = g["T"], g["m"], g["w"]
    n = T.size
    sublen = n - m + 1
    inp = oracle.precompute(T, m)
    out = np.zeros(sublen, mp.OUT_F64)
    rv = np.full(sublen, -1e12); ri = np.full(sublen, -1, np.int32)
    cv = np.full(sublen, -1e12); ci = np.full(sublen, -1, np.int32)
    for it in range((n - m + w) // w):
        mp.MatrixProfileKernelTLF(n, m, it, inp, out, w=w)
        off, nr = it * w, sublen - it * w
        a, b = out["rowAggregate"][:nr], out["columnAggregate"][:nr]
        k = a["value"] > rv[:nr]; rv[:nr][k] = a["value"][k]; ri[:nr][k] = a["index"][k]
        k = b["value"] > cv[off:]; cv[off:][k] = b["value"][k]; ci[off:][k] = b["index"][k]
    pick = rv > cv
    MP = np.sqrt(2 * m * (1 - np.where(pick, rv, cv)))
    MPI = np.where(pick, ri, ci)
    assert np.abs(MP - g["scamp_MP"]).max() <= 1e-8
    assert np.array_equal(MPI, g["scamp_MPI"])
    assert np.array_equal(MP, g["vanilla_MP"])     # and bit-identical to the reference kernel's own result


@pytest.mark.parametrize("flags", [0, 1, 4, 6, 16, 48])
@pytest.mark.parametrize("name", KAT + SYN + BIG)
def test_fast_path_vs_reference_kernels(mp, ctx, oracle, name, flags):
    """Kernels (1)-(3) vs the reference kernels' merged result. flags=1 (MPB_FLAG_STRICT: sequential mean, single
    seed, un-fused chain) must reproduce the reference's bits; the default (0: sequential mean, nested FMA, re-seeded
    row segments + telescoped chain carry) and 6 (sequential mean, one full-length chain per diagonal, nested FMA)
    follow the reference's chain to rounding (1e-11); 4 (window means, no re-seed), 16 (re-seeded segments without the
    carry) and 48 (window means + plain re-seeding: round 1's fast mode) must stay within 1e-9 relative."""
    g = load_golden(name)
    T, m = g["T"], g["m"]
    plan = run_plan(mp, ctx, T, m, mp.F64, flags)
    plan.finalize()
    MP, MPI, corr = plan.download()
    plan.close()
    if flags == mp.FLAG_STRICT:
        assert np.array_equal(corr, g["vanilla_corr"])
        assert np.array_equal(MP, np.sqrt(2 * m * (1 - g["vanilla_corr"])))
    nbad = check_profile(oracle, T, m, corr, MPI, MP, g["vanilla_corr"], g["vanilla_MPI"], rtol=1e-11 if flags in (0, 6) else RTOL)
    if flags == mp.FLAG_STRICT and name != "small128_m16_w32_t16":
        assert nbad == 0


def test_one_call_host_api(mp, oracle):
    g = load_golden("walk8192_m100_w1024_t256")
    MP, MPI, t = mp.matrix_profile(g["T"], g["m"], return_timings=True)
    assert np.allclose(MP, np.sqrt(2 * g["m"] * (1 - g["vanilla_corr"])), rtol=1e-9, atol=0)
    assert np.array_equal(MPI, g["vanilla_MPI"])
    assert t["cells"] == mp.useful_cells(8192, 100) and t["launches"] >= 5 and t["bands"] >= 1


@pytest.mark.parametrize("world", [2, 3, 8])
def test_band_shards_merge_to_unsharded_result(mp, ctx, world):
    """Multi-GPU path on one device: each rank's plan walks bands b % world == rank; the peer merge
    (exact (corr, index) max) must reproduce the single-plan profile bit for bit."""
    g = load_golden("walk65536_m256_w4096_t128")
    T, m = g["T"], g["m"]
    full = run_plan(mp, ctx, T, m)
    assert full.segment_rows > 0 and full.timings()["items"] > full.timings()["bands"]    # re-seeded row segments in effect
    full.finalize()
    MP0, MPI0, corr0 = full.download()
    plans = [run_plan(mp, ctx, T, m, shard=(r, world)) for r in range(world)]
    plans[0].merge_peers(plans[1:])
    plans[0].finalize()
    MP, MPI, corr = plans[0].download()
    assert np.array_equal(corr, corr0) and np.array_equal(MPI, MPI0) and np.array_equal(MP, MP0)
    for p in plans + [full]:
        p.close()


def test_export_mask_import_roundtrip(mp, ctx):
    """The NCCL-merge building blocks (export -> [all-reduce] -> mask -> [all-reduce] -> import) on one
    device with torch doing the reductions of two shards."""
    import torch
    g = load_golden("walk8192_m100_w1024_t256")
    T, m = g["T"], g["m"]
    sublen = T.size - m + 1
    stream = torch.cuda.ExternalStream(ctx.stream)
    plans = [run_plan(mp, ctx, T, m, shard=(r, 2)) for r in range(2)]
    with torch.cuda.stream(stream):
        corr = [torch.empty(sublen, dtype=torch.float64, device="cuda") for _ in plans]
        idx = [torch.empty(sublen, dtype=torch.int32, device="cuda") for _ in plans]
        for p, c, i in zip(plans, corr, idx):
            p.export(c.data_ptr(), i.data_ptr())
        cmax = torch.maximum(corr[0].view(torch.int64), corr[1].view(torch.int64)).view(torch.float64)
        for p, i in zip(plans, idx):
            p.mask_index(cmax.data_ptr(), i.data_ptr())
        imin = torch.minimum(idx[0], idx[1])
        plans[0].import_(cmax.data_ptr(), imin.data_ptr())
        plans[0].finalize()
        stream.synchronize()
    MP, MPI, c = plans[0].download()
    assert (np.abs(c - g["vanilla_corr"]) / g["vanilla_corr"]).max() <= RTOL
    assert np.array_equal(MPI, g["vanilla_MPI"])
    for p in plans:
        p.close()


def test_full_size_properties(mp, ctx, oracle):
    """BASELINE.json headline size (n=2^20, m=256, fp64): size-independent properties.
    (a) the neighbour's own best is at least as good: corr[MPI[i]] >= corr[i] (the same cell feeds both);
    (b) no neighbour inside the exclusion zone; (c) sampled rows agree with exact long-double
    correlations; (d) a two-shard run merges to the same bits."""
    from oracle_py import uniform_series
    n, m = 1 << 20, 256
    T = uniform_series(n, 99)
    plan = run_plan(mp, ctx, T, m)
    plan.finalize()
    MP, MPI, corr = plan.download()
    assert (MPI >= 0).all() and (MPI < n - m + 1).all()
    assert (corr[MPI] >= corr).all()
    assert (np.abs(MPI - np.arange(n - m + 1)) >= m // 4).all()
    assert np.array_equal(MP, np.sqrt(2 * m * (1 - corr)))
    rows = np.linspace(0, n - m, 48).astype(int)
    err = max(abs(oracle.exact_corr(T, m, i, MPI[i]) - corr[i]) / corr[i] for i in rows)
    assert err < 1e-9
    a, b = (run_plan(mp, ctx, T, m, shard=(r, 2)) for r in range(2))
    a.merge_peers([b]); a.finalize()
    MP2, MPI2, corr2 = a.download()
    assert np.array_equal(corr2, corr) and np.array_equal(MPI2, MPI)
    for p in (plan, a, b):
        p.close()


def test_fast_path_float_vs_double_oracle(mp, ctx, oracle):
    """fp32 engine vs the DOUBLE reference kernels at 1e-4 absolute on correlation;
    the reference's own float path is untested upstream."""
    g = load_golden("walk8192_m100_w1024_t256")
    T, m = g["T"], g["m"]
    plan = run_plan(mp, ctx, T, m, mp.F32)
    plan.finalize()
    MP, MPI, corr = plan.download()
    plan.close()
    assert np.abs(corr - g["vanilla_corr"]).max() <= 1e-4
    for i in np.nonzero(MPI != g["vanilla_MPI"])[0]:
        a, b = oracle.exact_corr(T, m, i, MPI[i]), oracle.exact_corr(T, m, i, g["vanilla_MPI"][i])
        assert abs(a - b) <= 1e-4


def test_edge_shapes(mp, ctx, oracle):
    """Smallest legal problems, a single partial band, m not a multiple of 4."""
    rng = np.random.default_rng(3)
    for n, m in [(24, 8), (64, 8), (200, 30), (1100, 16), (1500, 5), (3000, 64)]:
        T = np.cumsum(rng.standard_normal(n))
        w = 8
        while w * 2 <= min(n, 4096):
            w *= 2
        eMP, eMPI, ecorr, _ = oracle.run(T, m, w)
        plan = run_plan(mp, ctx, T, m, mp.F64, mp.FLAG_STRICT)
        plan.finalize()
        MP, MPI, corr = plan.download()
        plan.close()
        pos = ecorr > 0      # rows with no positive neighbour keep the kernels' P=0 floor (index then arbitrary upstream, -1 here)
        assert np.array_equal(corr[pos], ecorr[pos]) and np.array_equal(MPI[pos], eMPI[pos]), (n, m)
        assert (corr[~pos] == 0).all() and (MPI[~pos] == -1).all()


def test_error_behaviour(mp, ctx):
    with pytest.raises(mp.MpbError):
        mp.Plan(ctx, 100, 200)               # m > n (reference: CMakeLists.txt:86-89 rejects it at configure time)
    with pytest.raises(mp.MpbError):
        mp.Plan(ctx, 16, 16)                 # n-m+1 inside the exclusion zone
    with pytest.raises(mp.MpbError):
        mp.Context(10_000)
    inp = np.zeros(1024 - 16 + 1, mp.IN_F64)
    out = np.zeros(1024 - 16 + 1, mp.OUT_F64)
    with pytest.raises(mp.MpbError):
        mp.MatrixProfileKernelTLF(1024, 16, 99, inp, out, w=32)    # iteration out of range
    with pytest.raises(mp.MpbError):
        mp.MatrixProfileKernelTLF(1024, 16, 0, inp, out, w=4)      # w < 8 (row-reduce lanes, Vanilla.cpp:29-36)


def test_compute_in_band_pieces_equals_one_launch(mp, ctx):
    """mpb_plan_compute_bands: walking the bands in three launches gives the profile of the single launch, bit for bit
    (the phased multi-GPU step relies on it)."""
    g = load_golden("walk65536_m256_w4096_t128")
    T, m = g["T"], g["m"]
    whole = run_plan(mp, ctx, T, m)
    whole.finalize()
    MP0, MPI0, corr0 = whole.download()
    plan = mp.Plan(ctx, T.size, m, mp.F64, 0)
    plan.upload_series(T)
    plan.precompute()
    nlb = plan.local_bands()
    assert nlb == whole.timings()["bands"] and nlb > 4
    plan.compute_bands(0, 1)
    plan.compute_bands(1, 4)
    plan.compute_bands(4, -1)
    plan.finalize()
    MP, MPI, corr = plan.download()
    assert np.array_equal(corr, corr0) and np.array_equal(MPI, MPI0) and np.array_equal(MP, MP0)
    with pytest.raises(mp.MpbError):
        plan.compute_bands(5, 2)
    plan.close(); whole.close()


@pytest.mark.parametrize("dtype", [0, 1])
def test_seed_profile_does_not_change_the_result(mp, ctx, oracle, dtype):
    """The lag-m/4 seed diagonal is an ordinary diagonal computed twice: with and without it the profile agrees to
    rounding (the seed's dot product and the chain round differently) and the indices are the same."""
    g = load_golden("walk65536_m256_w4096_t128")
    T, m = g["T"], g["m"]
    out = []
    for flags in (0, mp.FLAG_NO_SEED):
        plan = run_plan(mp, ctx, T, m, dtype, flags)
        plan.finalize()
        out.append(plan.download())
        plan.close()
    tol = 0.0 if dtype == 0 else 2e-6      # fp64 chain-carry mode: the seed only ever is a filter threshold
    assert np.abs(out[0][2] - out[1][2]).max() <= tol
    differ = out[0][1] != out[1][1]
    assert differ.mean() < 1e-3          # only exact-tie / last-bit cases may pick the other neighbour
    if dtype == 0:
        check_profile(oracle, T, m, out[0][2], out[0][1], out[0][0], g["vanilla_corr"], g["vanilla_MPI"])


def test_filter_statistics_are_sane(mp, ctx):
    """The QT-domain block filter must reject almost everything on a converged profile: replayed blocks and parked
    thread-rows are a small fraction of the work, and every profile improvement went through the exact path."""
    rng = np.random.default_rng(5)
    n, m = 1 << 17, 256
    T = rng.random(n)
    plan = run_plan(mp, ctx, T, m)
    plan.finalize()
    MP, MPI, corr = plan.download()
    t = plan.timings()
    plan.close()
    blocks = t["cells"] / (64 * 32)                     # 8 rows x 8 diagonals x 32 lanes per block
    assert 0 < t["replays"] < 0.25 * blocks
    assert t["slow_updates"] >= (MPI >= 0).sum() * 0.5  # at least about one exact update per subsequence
    assert np.isfinite(MP).all() and corr.min() > 0 and corr.max() <= 1 + 1e-12


def test_fast_path_matches_strict_on_tsb_shaped_series(mp, ctx):
    """n = 2^18 uniform series (the shape of the reference's data/binary/*.tsb): fast path vs MPB_FLAG_STRICT
    (= the reference kernels' bits) within the 1e-9 budget, indices identical where the values are not tied."""
    rng = np.random.default_rng(1234)
    n, m = 1 << 18, 256
    T = rng.random(n)
    res = []
    for flags in (0, mp.FLAG_STRICT):
        plan = run_plan(mp, ctx, T, m, mp.F64, flags)
        plan.finalize()
        res.append(plan.download())
        plan.close()
    rel = np.abs(res[0][2] - res[1][2]) / res[1][2]
    assert rel.max() <= RTOL
    bad = res[0][1] != res[1][1]
    assert bad.mean() < 1e-4
    assert (rel[bad] <= RTOL).all()


@pytest.mark.parametrize("dtype", [0, 1])
@pytest.mark.parametrize("series", ["walk65536_m256_w4096_t128", "unif65536_m256_w4096_t128", "walk8192_m100_w1024_t256"])
def test_sketch_pass_does_not_change_the_result(mp, ctx, series, dtype):
    """The sketch pass leaves borrowed lower bounds {bound, index -1} in the profile; the tile kernel walks the same
    bands, records the cell behind every bound exactly and replaces it.  Whatever number of bands is sketched, the
    profile is bit-identical to the one computed without the sketch, and no borrowed entry survives."""
    g = load_golden(series)
    T, m = g["T"], g["m"]
    res = []
    for sk in (0, 1, 3, 1000):
        plan = mp.Plan(ctx, T.size, m, dtype, 0)
        plan.set_sketch_bands(sk)
        plan.upload_series(T)
        plan.precompute(); plan.compute(); plan.finalize()
        res.append(plan.download())
        plan.close()
    for MP, MPI, corr in res[1:]:
        assert (MPI >= 0).all()
        assert np.array_equal(corr, res[0][2]) and np.array_equal(MPI, res[0][1]) and np.array_equal(MP, res[0][0])


def test_sketch_pass_in_sharded_phased_compute(mp, ctx):
    """Band shards with the sketch on, walked in pieces with a merge in between (the multi-GPU step): same profile as
    one unsharded plan without the sketch."""
    g = load_golden("unif65536_m256_w4096_t128")
    T, m = g["T"], g["m"]
    ref = mp.Plan(ctx, T.size, m, mp.F64, 0)
    ref.set_sketch_bands(0); ref.upload_series(T); ref.precompute(); ref.compute(); ref.finalize()
    MP0, MPI0, corr0 = ref.download()
    ref.close()
    world = 3
    plans = []
    for r in range(world):
        p = mp.Plan(ctx, T.size, m, mp.F64, 0)
        p.set_shard(r, world); p.set_sketch_bands(2); p.upload_series(T); p.precompute()
        p.compute_bands(0, 3)
        plans.append(p)
    plans[0].merge_peers(plans[1:])
    for p in plans[1:]:
        p.merge_peers([plans[0]])
    for p in plans:
        p.compute_bands(3, -1)
    plans[0].merge_peers(plans[1:])
    plans[0].finalize()
    MP, MPI, corr = plans[0].download()
    for p in plans:
        p.close()
    assert np.array_equal(corr, corr0) and np.array_equal(MPI, MPI0) and np.array_equal(MP, MP0)


# ---------------------------------------------------------------------------------------------------------------------
# Parity at the benchmarked size (n = 2^20, m = 256).  The CPU oracle needs an hour per thread there, so the reference's
# bits come from MPB_FLAG_STRICT, which the tests above pin bit-for-bit to the reference kernels for n <= 65536.
# ---------------------------------------------------------------------------------------------------------------------
def _series(kind, n, seed=1234):
    rng = np.random.default_rng(seed)
    return rng.random(n) if kind == "uniform" else np.cumsum(rng.standard_normal(n))


def _profile(mp, ctx, T, m, dtype, flags):
    plan = run_plan(mp, ctx, T, m, dtype, flags)
    plan.finalize()
    MP, MPI, corr = plan.download()
    t = plan.timings()
    plan.close()
    return MP, MPI, corr, t


@pytest.fixture(scope="module")
def strict_2p20(mp, ctx):
    """MPB_FLAG_STRICT profiles (the reference's bits) of the two benchmark series, computed once."""
    out = {}
    for kind in ("walk", "uniform"):
        T = _series(kind, 1 << 20)
        out[kind] = (T,) + _profile(mp, ctx, T, 256, mp.F64, mp.FLAG_STRICT)[:3]
    return out


@pytest.mark.parametrize("kind", ["walk", "uniform"])
def test_default_fp64_follows_the_reference_chain_at_2p20(mp, ctx, oracle, strict_2p20, kind):
    """BASELINE.json north_star: fp64 within 1e-9 relative of the reference on a random walk at n = 2^20, m = 256 -- the
    default (benchmarked) mode against MPB_FLAG_STRICT, ALL rows: correlation and distance, and every index mismatch must
    be a tie within the tolerance (exact long-double correlations of the two candidates)."""
    T, MP0, MPI0, corr0 = strict_2p20[kind]
    MP, MPI, corr, t = _profile(mp, ctx, T, 256, mp.F64, 0)
    assert t["segment_rows"] > 0 and t["items"] > 10 * t["bands"]            # the parallel (band x segment) decomposition is in effect
    assert t["chain_candidates"] > 0
    nbad = check_profile(oracle, T, 256, corr, MPI, MP, corr0, MPI0, rtol=RTOL)
    rel = (np.abs(corr - corr0) / corr0).max()
    print(f"{kind}: max rel corr err {rel:.3e}, index mismatches (all ties) {nbad}, max chain correction {t['chain_max_correction']:.3e}")
    assert rel <= 1e-10                                                      # rounding (nested FMA vs mul/mul/add/add), not drift
    assert t["chain_max_correction"] < 0.5 * t["chain_margin"]                # the near-tie margin covered every correction applied


def test_plain_reseeding_drifts_at_2p20(mp, ctx, strict_2p20):
    """What the chain carry is for: the same decomposition WITHOUT it (MPB_FLAG_NO_CHAIN_CARRY | MPB_FLAG_WINDOW_MEAN,
    round 1's default) is more than 1e-9 away from the reference on the random walk."""
    T, MP0, MPI0, corr0 = strict_2p20["walk"]
    MP, MPI, corr, t = _profile(mp, ctx, T, 256, mp.F64, mp.FLAG_NO_CHAIN_CARRY | mp.FLAG_WINDOW_MEAN)
    rel = (np.abs(corr - corr0) / corr0).max()
    assert 1e-9 < rel < 1e-7


@pytest.mark.parametrize("kind", ["walk", "uniform"])
def test_fp32_at_2p20_against_the_double_chain(mp, ctx, oracle, strict_2p20, kind):
    """BASELINE.json config 3: fp32 with periodic exact re-seeding vs the DOUBLE reference chain at 1e-4 absolute on the
    correlation, every index mismatch passed through the tie rule (exact correlations of the two candidates within 1e-4)."""
    T, MP0, MPI0, corr0 = strict_2p20[kind]
    MP, MPI, corr, t = _profile(mp, ctx, T, 256, mp.F32, 0)
    err = np.abs(corr - corr0).max()
    bad = np.nonzero(MPI != MPI0)[0]
    print(f"{kind}: fp32 max abs corr err {err:.3e}, index mismatches {bad.size}")
    assert err <= 1e-4
    for i in bad:
        a, b = oracle.exact_corr(T, 256, i, MPI[i]), oracle.exact_corr(T, 256, i, MPI0[i])
        assert abs(a - b) <= 1e-4, (i, MPI[i], MPI0[i], a, b)


def _mu_series(kind, n, rng):
    if kind == "uniform": return rng.random(n)
    if kind == "walk": return np.cumsum(rng.standard_normal(n))
    if kind == "noise0": return rng.standard_normal(n)                       # mean hovers around zero: mostly hard steps
    if kind == "near1": return np.cumsum(rng.standard_normal(n)) * 1e-3 + 1.0  # mean keeps crossing a binade boundary
    if kind == "dyadic": return np.round(rng.random(n) * 8) / 8                # exact ties of d / ulp
    if kind == "tiny": return (rng.random(n) - 0.5) * 1e-300
    if kind == "const": return np.full(n, 0.3)
    if kind == "big": return np.cumsum(rng.standard_normal(n)) * 1e6
    raise KeyError(kind)


@pytest.mark.parametrize("kind,n,m", [("uniform", 1 << 20, 256), ("walk", 1 << 20, 256), ("noise0", 1 << 18, 64), ("near1", 1 << 18, 128),
                                      ("dyadic", 1 << 18, 32), ("tiny", 1 << 17, 32), ("const", 1 << 16, 32), ("big", 1 << 18, 1024),
                                      ("walk", 70001, 4096), ("uniform", 5000, 9)])
def test_running_mean_scan_equals_the_walk(mp, ctx, kind, n, m):
    """The reference's running mean (HostSideComputation.hpp:23,29) as a verified segmented scan == the same recurrence
    walked element by element (MPB_FLAG_MU_WALK, which the golden tests pin to the reference), byte for byte over the
    whole InputDataPack array, and the scan must have passed its own verification (no fallback to the walk)."""
    T = _mu_series(kind, n, np.random.default_rng(7))
    packs = []
    for flags in (mp.FLAG_SEQ_MEAN, mp.FLAG_SEQ_MEAN | mp.FLAG_MU_WALK):
        plan = mp.Plan(ctx, n, m, mp.F64, flags)
        plan.upload_series(T)
        plan.precompute()
        packs.append(plan.download_input())
        walked = plan.mu_walked()
        assert walked == bool(flags & mp.FLAG_MU_WALK)
        plan.close()
    assert packs[0].tobytes() == packs[1].tobytes()


def test_tlf_chunks_at_w4096_n65536(mp, oracle):
    """MatrixProfileKernelTLF at the reference README's chunk width (w = 4096) on the n = 65536 random-walk golden: three chunks
    byte-for-byte against the CPU restatement of the Vanilla kernel, and the reference's host flow over ALL 16 chunks
    (UpdateAggregates + PostComputeAggregates, host/MatrixProfileHost.cpp:101-142) bit-equal to the golden merged profile."""
    g = load_golden("walk65536_m256_w4096_t128")
    T, m, w = g["T"], g["m"], 4096
    n = T.size
    sublen = n - m + 1
    inp = oracle.precompute(T, m)
    out = np.zeros(sublen, mp.OUT_F64)
    nit = (n - m + w) // w
    rv = np.full(sublen, -1e12); ri = np.full(sublen, -1, np.int32)
    cv = np.full(sublen, -1e12); ci = np.full(sublen, -1, np.int32)
    for it in range(nit):
        mp.MatrixProfileKernelTLF(n, m, it, inp, out, w=w)
        off, nr = it * w, sublen - it * w
        if it in (0, 7, nit - 1):
            want = oracle.kernel("vanilla", n, m, w, it, inp)
            for agg in ("rowAggregate", "columnAggregate"):          # field bytes (the struct padding is not part of the contract)
                assert np.array_equal(out[agg]["value"][:nr], want[agg]["value"][:nr]), (it, agg)
                assert np.array_equal(out[agg]["index"][:nr], want[agg]["index"][:nr]), (it, agg)
        a, b = out["rowAggregate"][:nr], out["columnAggregate"][:nr]
        k = a["value"] > rv[:nr]; rv[:nr][k] = a["value"][k]; ri[:nr][k] = a["index"][k]
        k = b["value"] > cv[off:]; cv[off:][k] = b["value"][k]; ci[off:][k] = b["index"][k]
    pick = rv > cv
    assert np.array_equal(np.where(pick, rv, cv), g["vanilla_corr"])
    assert np.array_equal(np.where(pick, ri, ci), g["vanilla_MPI"])


def test_slots_without_a_positive_correlation(mp, ctx, oracle):
    """Documented divergence: a slot whose valid correlations are all <= 0.  The reference's P = 0
    substitution for excluded / out-of-range cells (Vanilla.cpp:72-84) then wins the max, so its value is 0 (distance
    sqrt(2m)) and its index names an excluded column; the engine never walks excluded cells: same value, index -1."""
    T, m = np.random.default_rng(3).standard_normal(24), 16
    eMP, eMPI, ecorr, _ = oracle.run(T, m, 8)
    zero = ecorr <= 0
    assert zero.any() and not zero.all()
    for flags in (0, mp.FLAG_STRICT):
        plan = run_plan(mp, ctx, T, m, mp.F64, flags)
        plan.finalize()
        MP, MPI, corr = plan.download()
        plan.close()
        assert np.array_equal(corr[zero], ecorr[zero]) and np.array_equal(MP[zero], eMP[zero]), flags
        assert (MPI[zero] == -1).all(), flags
        assert np.abs(corr[~zero] - ecorr[~zero]).max() <= 1e-9 and np.array_equal(MPI[~zero], eMPI[~zero]), flags


def test_window_longer_than_the_precompute_staging_is_rejected(mp, ctx):
    """m beyond what the precompute kernels can stage in shared memory fails with MPB_E_INVALID and a message, not with an
    opaque launch error."""
    with pytest.raises(mp.MpbError) as e:
        mp.Plan(ctx, 60000, 15000)
    assert e.value.code == -1 and "14336" in str(e.value)
