// precompute.cu -- kernel (1): the reference's PreComputeStatistics
// (include/host/HostSideComputation.hpp:20-45) on the GPU, bit-faithful in fp64:
//   mu_i  : the reference's *running* mean recurrence (:29), evaluated strictly in order (mu_scan_kernel);
//   df_i, dg_i (:30-33), QT_i = sum_k (T[i+k]-mu_i)(T[k]-mu_0) (:37,:40), inv_i = 1/sqrt(sum_k (T[i+k]-mu_i)^2)
//   (:38,:41) with the reference's summation order and un-fused multiplies/adds (stats_kernel).
// IEEE division and sqrt are correctly rounded on both sides, so the output equals the host's bits.
#include "mp_common.cuh"
#include "mp_kernels.h"

namespace mpb {

// ---------------------------------------------------------------------------------------------
// Running mean.  The recurrence mean += (T[i+m-1]-T[i-1])/m is a chain of dependent roundings, so one
// thread walks it; the rest of the CTA computes the increments (independent) one chunk ahead into
// shared memory and drains finished chunks to HBM, so the scanning thread only sees LDS/STS latency.
// ---------------------------------------------------------------------------------------------
constexpr int SCAN_CHUNK = 1024;
constexpr int SCAN_THREADS = 256;

__global__ void __launch_bounds__(SCAN_THREADS, 1)
mu_scan_kernel(const double *__restrict__ T, int n, int m, double *__restrict__ mu, const int *__restrict__ only_if) {
    if (only_if && *only_if == 0) return;        // the parallel path verified: nothing to do
    __shared__ __align__(16) double delta[2][SCAN_CHUNK];
    __shared__ __align__(16) double out[2][SCAN_CHUNK];
    __shared__ double carry;
    const int sublen = n - m + 1;
    const int tid = threadIdx.x;
    const double dm = (double)m;
    if (tid == 0) {
        double mean = 0.0;                       // std::accumulate(T.begin(), T.begin()+m, 0.0), :23
        for (int k = 0; k < m; ++k) mean = __dadd_rn(mean, T[k]);
        carry = __ddiv_rn(mean, dm);             // mean /= m
    }
    const int nchunks = (sublen + SCAN_CHUNK - 1) / SCAN_CHUNK;
    // element e of chunk c is subsequence i = c*SCAN_CHUNK + e; increment for i = 0 is 0 (:29)
    auto fill = [&](int c, int buf) {
        for (int e = tid; e < SCAN_CHUNK; e += SCAN_THREADS) {
            const int i = c * SCAN_CHUNK + e;
            double d = 0.0;
            if (i > 0 && i < sublen) d = __ddiv_rn(__dsub_rn(T[i + m - 1], T[i - 1]), dm);
            delta[buf][e] = d;
        }
    };
    fill(0, 0);
    __syncthreads();
    for (int c = 0; c < nchunks; ++c) {
        const int buf = c & 1;
        if (tid == 0) {
            // the chain itself: 16 increments in registers (LDS.128 one group ahead), 16 dependent DADDs, results back
            // through STS.128 -- the chain only ever waits for the previous DADD
            double mean = carry;
            const double2 *din = reinterpret_cast<const double2 *>(delta[buf]);
            double2 *dout = reinterpret_cast<double2 *>(out[buf]);
            double2 d[8], nd[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) d[k] = din[k];
#pragma unroll 1
            for (int e = 0; e < SCAN_CHUNK / 2; e += 8) {
                const int en = e + 8 < SCAN_CHUNK / 2 ? e + 8 : e;
#pragma unroll
                for (int k = 0; k < 8; ++k) nd[k] = din[en + k];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    double2 o;
                    mean = __dadd_rn(mean, d[k].x); o.x = mean;
                    mean = __dadd_rn(mean, d[k].y); o.y = mean;
                    dout[e + k] = o;
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) d[k] = nd[k];
            }
            carry = mean;
        } else if (tid >= 32) {                  // warps 1.. prefetch the next chunk and drain the previous one
            if (c + 1 < nchunks) {
                for (int e = tid - 32; e < SCAN_CHUNK; e += SCAN_THREADS - 32) {
                    const int i = (c + 1) * SCAN_CHUNK + e;
                    double d = 0.0;
                    if (i < sublen) d = __ddiv_rn(__dsub_rn(T[i + m - 1], T[i - 1]), dm);
                    delta[buf ^ 1][e] = d;
                }
            }
            if (c > 0) {
                for (int e = tid - 32; e < SCAN_CHUNK; e += SCAN_THREADS - 32) {
                    const int i = (c - 1) * SCAN_CHUNK + e;
                    if (i < sublen) mu[i] = out[buf ^ 1][e];
                }
            }
        }
        __syncthreads();
    }
    for (int e = tid; e < SCAN_CHUNK; e += SCAN_THREADS) {
        const int i = (nchunks - 1) * SCAN_CHUNK + e;
        if (i < sublen) mu[i] = out[(nchunks - 1) & 1][e];
    }
}

// ---------------------------------------------------------------------------------------------
// The same running mean, bit for bit, WITHOUT walking the chain element by element (20 ms -> well under 1 ms at n = 2^20).
// mean_i = RN(mean_{i-1} + d_i) is a chain of dependent roundings, but a rounding only looks at the bits below the
// result's ulp u: while the running mean stays inside one binade (same exponent, same sign) every mean is a multiple of
// u and RN(mean + d) = mean + t with t = d rounded to a multiple of u -- a TRANSLATION that does not depend on the mean
// (exact ties of d / u aside).  Such "easy" steps form segments inside which the means are exact prefix sums of the t's
// (every partial sum is a difference of two doubles of one binade, hence exact in fp64), a segmented scan.  Steps that
// change binade or sign, tie, or whose binade cannot be told from an approximate mean (the exact window mean, which
// differs from the running mean by its drift, ~1e-10 relative, against a 6e-8 guard band) are "hard": one thread walks
// only those, in order, with the real fp64 operation (mu_hard_kernel: 35 k of 1 M steps on the U[0,1) series, 6 k on
// the random walk).  Nothing here is trusted: mu_fill_kernel re-checks EVERY step, mean_i == RN(mean_{i-1} + d_i), in
// parallel -- by induction from mean_0 the result is then the reference's chain -- and mu_scan_kernel (the plain
// sequential walk) runs if any check fails.  Series whose mean hovers around zero make most steps hard and fall
// back towards the sequential speed; the result is the same bits either way.
// ---------------------------------------------------------------------------------------------
constexpr int MUP_CHUNK = 1024;          // elements per CTA of the classify / fill kernels
constexpr int MUP_THREADS = 256;
constexpr int MUP_PER = MUP_CHUNK / MUP_THREADS;

struct MuScratch {
    double *plocal;          // [i] sum of t over the easy steps since the last hard step of i's chunk (or the chunk start)
    double *hd, *hg, *hm;    // dense, in order, per hard step k: its increment d, the exact sum of the easy steps since hard step k-1, its mean
    unsigned short *hrank;   // [i] hard steps of i's chunk at positions <= i
    unsigned short *hpos;    // [chunk * MUP_CHUNK + r] position in the chunk of its r-th hard step
    int *hcnt;               // [chunk] hard steps in the chunk
    int *hbase;              // [chunk] hard steps before the chunk
    double *tail;            // [chunk] plocal of the chunk's last element
    double *carry;           // [chunk] exact sum of the easy steps between the last hard step before the chunk and the chunk's first element
    unsigned long long *absmax;  // bits of max |T|
    int *bad;                // != 0: a step failed verification (or the parallel path was skipped): mu_scan_kernel takes over
    int *total;              // hard steps in all
};

__global__ void mu_absmax_kernel(const double *__restrict__ T, int n, unsigned long long *out) {
    double v = 0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) v = fmax(v, fabs(T[i]));
    for (int sft = 16; sft; sft >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, sft));
    if ((threadIdx.x & 31) == 0) atomicMax(out, (unsigned long long)__double_as_longlong(v));   // non-negative doubles order like integers
}

// exponent field and sign of a double, and "too close to a binade boundary (or to zero) to trust"
__device__ __forceinline__ int mup_binade(double x) { return (__double2hiint(x) >> 20) & 0xfff; }   // sign bit + 11 exponent bits
__device__ __forceinline__ bool mup_uncertain(double x, double theta) {
    const unsigned frac = ((unsigned)__double2hiint(x) & 0xfffffu);           // top 20 mantissa bits
    return !(fabs(x) >= theta) || frac < 16u || frac > 0xfffffu - 16u;        // within 2^-16 of a power of two (guard band >> drift)
}

__device__ __forceinline__ double mup_delta(const double *__restrict__ T, int i, int m, double dm) {
    return i > 0 ? __ddiv_rn(__dsub_rn(T[i + m - 1], T[i - 1]), dm) : 0.0;    // HostSideComputation.hpp:29
}

__global__ void __launch_bounds__(MUP_THREADS)
mu_classify_kernel(const double *__restrict__ T, int n, int m, MuScratch sc) {
    extern __shared__ double sm[];                       // T[lo - 1 .. lo - 1 + MUP_CHUNK + m]   (window means of lo-1 .. hi-1)
    __shared__ double wsum[MUP_THREADS / 32];
    __shared__ int wflag[MUP_THREADS / 32], wcnt[MUP_THREADS / 32];
    const int sublen = n - m + 1, lo = blockIdx.x * MUP_CHUNK, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const double dm = (double)m;
    for (int k = tid; k < MUP_CHUNK + 1 + m; k += MUP_THREADS) { const int g = lo - 1 + k; sm[k] = (g >= 0 && g < n) ? T[g] : 0.0; }
    __syncthreads();
    const double theta = ldexp(__longlong_as_double((long long)*sc.absmax), -16);
    // approximate means (exact window means) of elements e0-1 .. e0+MUP_PER-1, e0 = first element of this thread
    const int e0 = tid * MUP_PER;
    double approx[MUP_PER];
#pragma unroll
    for (int q = 0; q < MUP_PER; ++q) {
        const double *w = sm + e0 + q;                   // window of element lo + e0 + q - 1
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
        int k = 0;
        for (; k + 3 < m; k += 4) { s0 += w[k]; s1 += w[k + 1]; s2 += w[k + 2]; s3 += w[k + 3]; }
        for (; k < m; ++k) s0 += w[k];
        approx[q] = ((s0 + s1) + (s2 + s3)) / dm;
    }
    double t[MUP_PER];
    bool hard[MUP_PER];
#pragma unroll
    for (int q = 0; q < MUP_PER; ++q) {
        const int i = lo + e0 + q;
        const double d = (i > 0 && i < sublen) ? __ddiv_rn(__dsub_rn(sm[e0 + q + m], sm[e0 + q]), dm) : 0.0;   // T[i+m-1] - T[i-1]
        const double prev = approx[q], s = prev + d;
        const int bp = mup_binade(prev), bs = mup_binade(s);
        bool h = i == 0 || bp != bs || mup_uncertain(prev, theta) || mup_uncertain(s, theta);
        const int e = (bs & 0x7ff) - 1023 - 52;          // ulp of the binade = 2^e
        const double qd = ldexp(d, -e), r = rint(qd);
        h = h || fabs(qd - r) == 0.5 || !(fabs(qd) < 4.0e18);
        hard[q] = h || i >= sublen;
        t[q] = hard[q] ? 0.0 : ldexp(r, e);
    }
    // segmented inclusive scan of (t, hard) over the chunk: thread-local, then across the warp, then across the warps
    double run = 0.0; bool any = false; int cnt = 0;
    double loc[MUP_PER]; int lcnt[MUP_PER]; bool lany[MUP_PER];
#pragma unroll
    for (int q = 0; q < MUP_PER; ++q) {
        if (hard[q]) { run = 0.0; any = true; ++cnt; } else run += t[q];
        loc[q] = run; lcnt[q] = cnt; lany[q] = any;
    }
    double wsumv = run; int wflagv = any, wcntv = cnt;   // inclusive over lanes
#pragma unroll
    for (int sft = 1; sft < 32; sft <<= 1) {
        const double os = __shfl_up_sync(0xffffffffu, wsumv, sft);
        const int of = __shfl_up_sync(0xffffffffu, wflagv, sft), oc = __shfl_up_sync(0xffffffffu, wcntv, sft);
        if (lane >= sft) { if (!wflagv) wsumv += os; wflagv |= of; wcntv += oc; }
    }
    if (lane == 31) { wsum[warp] = wsumv; wflag[warp] = wflagv; wcnt[warp] = wcntv; }
    // exclusive prefix for this thread within its warp
    double ex = __shfl_up_sync(0xffffffffu, wsumv, 1); int exf = __shfl_up_sync(0xffffffffu, wflagv, 1), exc = __shfl_up_sync(0xffffffffu, wcntv, 1);
    if (lane == 0) { ex = 0.0; exf = 0; exc = 0; }
    __syncthreads();
    double bsum = 0.0; int bflag = 0, bcnt = 0;          // carry from the warps before this one
    for (int w = 0; w < warp; ++w) { if (wflag[w]) { bsum = wsum[w]; bflag = 1; } else bsum += wsum[w]; bcnt += wcnt[w]; }
    if (!exf) { ex += bsum; exf = bflag; }
    exc += bcnt;
#pragma unroll
    for (int q = 0; q < MUP_PER; ++q) {
        const int i = lo + e0 + q;
        if (i >= sublen) break;
        const double pl = lany[q] ? loc[q] : loc[q] + ex;
        const int rk = lcnt[q] + exc;
        sc.plocal[i] = pl;
        sc.hrank[i] = (unsigned short)rk;
        if (hard[q]) sc.hpos[(size_t)lo + rk - 1] = (unsigned short)(e0 + q);
        if (i == min(lo + MUP_CHUNK, sublen) - 1) { sc.tail[blockIdx.x] = pl; sc.hcnt[blockIdx.x] = rk; }
    }
}

// Chunk-level scan (one CTA; the chunk count is n / 1024): hard steps before each chunk, and the easy-step sum carried into it.
__global__ void __launch_bounds__(1024, 1)
mu_chunk_scan_kernel(int nch, MuScratch sc) {
    __shared__ int wc[32]; __shared__ double ws[32]; __shared__ int wf[32];
    __shared__ int basec; __shared__ double bases; __shared__ int basef;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) { basec = 0; bases = 0.0; basef = 1; }     // "a hard step" precedes chunk 0 (element 0 is hard anyway)
    __syncthreads();
    for (int c0 = 0; c0 < nch; c0 += 1024) {
        const int c = c0 + tid;
        int cnt = c < nch ? sc.hcnt[c] : 0, flag = cnt > 0;
        double sum = c < nch ? sc.tail[c] : 0.0;             // the chunk's contribution to the running easy-step sum (restarts at its last hard step)
        int icnt = cnt;
#pragma unroll
        for (int sft = 1; sft < 32; sft <<= 1) {
            const int oc = __shfl_up_sync(0xffffffffu, icnt, sft), of = __shfl_up_sync(0xffffffffu, flag, sft);
            const double os = __shfl_up_sync(0xffffffffu, sum, sft);
            if (lane >= sft) { icnt += oc; if (!flag) sum += os; flag |= of; }
        }
        if (lane == 31) { wc[warp] = icnt; ws[warp] = sum; wf[warp] = flag; }
        int ec = __shfl_up_sync(0xffffffffu, icnt, 1), ef = __shfl_up_sync(0xffffffffu, flag, 1);
        double es = __shfl_up_sync(0xffffffffu, sum, 1);
        if (lane == 0) { ec = 0; ef = 0; es = 0.0; }
        __syncthreads();
        int pc = basec, pf = basef; double ps = bases;       // everything before this warp: previous rounds, then the warps before
        for (int w = 0; w < warp; ++w) { pc += wc[w]; if (wf[w]) { ps = ws[w]; pf = 1; } else ps += ws[w]; }
        if (!ef) es += ps;
        ec += pc;
        if (c < nch) { sc.hbase[c] = ec; sc.carry[c] = es; }
        __syncthreads();
        if (tid == 1023) { basec = ec + cnt; if (cnt > 0) bases = sc.tail[c < nch ? c : nch - 1]; else bases = es + (c < nch ? sc.tail[c] : 0.0); basef = 1; }
        __syncthreads();
    }
    if (tid == 0) *sc.total = basec;
}

// dense operand arrays of the hard steps
__global__ void mu_compact_kernel(const double *__restrict__ T, int n, int m, MuScratch sc) {
    const int c = blockIdx.x, lo = c * MUP_CHUNK, cnt = sc.hcnt[c], base = sc.hbase[c];
    const double carry = sc.carry[c], dm = (double)m;
    for (int r = threadIdx.x; r < cnt; r += blockDim.x) {
        const int pos = sc.hpos[(size_t)lo + r], i = lo + pos;
        sc.hd[base + r] = mup_delta(T, i, m, dm);
        const double pl = pos > 0 ? sc.plocal[i - 1] : 0.0;   // easy steps between the previous hard step (or the chunk start) and this one
        sc.hg[base + r] = r == 0 ? __dadd_rn(carry, pl) : pl;
    }
}

// The hard steps, in order, with the real operation: M_k = RN((M_{k-1} + G_k) + d_k), the first sum being exact.
// Thread 0 walks; the whole CTA streams the operands through shared memory one stage ahead.
constexpr int MUH_STAGE = 1024;
__global__ void __launch_bounds__(MUP_THREADS, 1)
mu_hard_kernel(const double *__restrict__ T, int m, MuScratch sc) {
    __shared__ __align__(16) double sd[2][MUH_STAGE], sg[2][MUH_STAGE], so[MUH_STAGE];
    const int tid = threadIdx.x, total = *sc.total;
    auto fetch = [&](int st, int buf) {
        for (int r = tid; r < MUH_STAGE; r += MUP_THREADS) {
            const int k = st * MUH_STAGE + r;
            sd[buf][r] = k < total ? sc.hd[k] : 0.0;
            sg[buf][r] = k < total ? sc.hg[k] : 0.0;
        }
    };
    double M = 0.0;
    if (tid == 0) {
        double mean = 0.0;                   // std::accumulate(T.begin(), T.begin()+m, 0.0), HostSideComputation.hpp:23
        for (int k = 0; k < m; ++k) mean = __dadd_rn(mean, T[k]);
        M = __ddiv_rn(mean, (double)m);
    }
    fetch(0, 0);
    __syncthreads();
    const int nst = (total + MUH_STAGE - 1) / MUH_STAGE;
    for (int st = 0; st < nst; ++st) {
        const int buf = st & 1;
        if (tid == 0) {
            const double2 *pd = reinterpret_cast<const double2 *>(sd[buf]), *pg = reinterpret_cast<const double2 *>(sg[buf]);
            double2 *po = reinterpret_cast<double2 *>(so);
#pragma unroll 4
            for (int r = 0; r < MUH_STAGE / 2; ++r) {
                const double2 d = pd[r], g = pg[r];
                double2 o;
                M = __dadd_rn(__dadd_rn(M, g.x), d.x); o.x = M;
                M = __dadd_rn(__dadd_rn(M, g.y), d.y); o.y = M;
                po[r] = o;
            }
        } else if (tid >= 32 && st + 1 < nst) {
            for (int r = tid - 32; r < MUH_STAGE; r += MUP_THREADS - 32) {
                const int k = (st + 1) * MUH_STAGE + r;
                sd[buf ^ 1][r] = k < total ? sc.hd[k] : 0.0;
                sg[buf ^ 1][r] = k < total ? sc.hg[k] : 0.0;
            }
        }
        __syncthreads();
        for (int r = tid; r < MUH_STAGE; r += MUP_THREADS) { const int k = st * MUH_STAGE + r; if (k < total) sc.hm[k] = so[r]; }
        __syncthreads();
    }
}

__device__ __forceinline__ double mup_value(const MuScratch &sc, int i) {
    const int c = i / MUP_CHUNK, rk = sc.hrank[i], base = sc.hbase[c];
    // after the chunk's rk-th hard step: its mean + the easy steps since; before the chunk's first: the last hard step before
    // the chunk + (the easy steps carried into the chunk + those inside it), every sum exact (module header)
    if (rk > 0) return __dadd_rn(sc.hm[base + rk - 1], sc.plocal[i]);
    return __dadd_rn(sc.hm[base - 1], __dadd_rn(sc.carry[c], sc.plocal[i]));
}

// mu[i] for every i, and the proof: every step must satisfy the recurrence exactly
__global__ void mu_fill_kernel(const double *__restrict__ T, int n, int m, MuScratch sc, double *__restrict__ mu) {
    const int sublen = n - m + 1, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sublen) return;
    const double v = mup_value(sc, i);
    mu[i] = v;
    if (i > 0) {
        const double want = __dadd_rn(mup_value(sc, i - 1), mup_delta(T, i, m, (double)m));
        if (__double_as_longlong(want) != __double_as_longlong(v)) atomicOr(sc.bad, 1);
    }
}

// ---------------------------------------------------------------------------------------------
// Exact window means (default mode): mu_i = (sum_k T[i+k]) / m, one thread per subsequence, k in order.
// Unlike the running recurrence above this has no sequential chain (0.1 ms instead of 20 ms at n = 2^20)
// and no drift: |mu_i - exact| <= m * eps * max|T| instead of growing with i.  The fast path re-seeds QT
// from exact dot products anyway, so its parity budget against the reference (1e-9) already contains the
// reference's own running-mean drift (SURVEY.md section 7 "hard parts", App. B-9 (iii)).
// ---------------------------------------------------------------------------------------------
constexpr int MUW_THREADS = 256;
__global__ void __launch_bounds__(MUW_THREADS)
mu_window_kernel(const double *__restrict__ T, int n, int m, double *__restrict__ mu) {
    extern __shared__ double sm[];
    const int sublen = n - m + 1;
    const int i0 = blockIdx.x * MUW_THREADS;
    for (int k = threadIdx.x; k < MUW_THREADS + m; k += MUW_THREADS) sm[k] = (i0 + k < n) ? T[i0 + k] : 0.0;
    __syncthreads();
    const int i = i0 + threadIdx.x;
    if (i >= sublen) return;
    const double *w = sm + threadIdx.x;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;     // 4 interleaved partial sums: shorter chains, smaller error
    int k = 0;
    for (; k + 3 < m; k += 4) { s0 += w[k]; s1 += w[k + 1]; s2 += w[k + 2]; s3 += w[k + 3]; }
    for (; k < m; ++k) s0 += w[k];
    mu[i] = ((s0 + s1) + (s2 + s3)) / (double)m;
}

// ---------------------------------------------------------------------------------------------
// df, dg, QT seeds, inverse norms.  One thread per subsequence, k-loop in the reference's order.
// ---------------------------------------------------------------------------------------------
constexpr int STATS_THREADS = 256;

template <typename OutT>
__global__ void __launch_bounds__(STATS_THREADS)
stats_kernel(const double *__restrict__ T, const double *__restrict__ mu, int n, int m,
             OutT *__restrict__ df_dg /* interleaved {df,dg} */, OutT *__restrict__ inv, OutT *__restrict__ qt0) {
    extern __shared__ double sm[];
    double *win = sm;                        // T[i0 .. i0 + STATS_THREADS + m)
    double *a0 = sm + STATS_THREADS + m;     // T[k] - mu_0
    const int sublen = n - m + 1;
    const int i0 = blockIdx.x * STATS_THREADS;
    const double mu0 = mu[0];
    for (int k = threadIdx.x; k < STATS_THREADS + m; k += STATS_THREADS) win[k] = (i0 + k < n) ? T[i0 + k] : 0.0;
    for (int k = threadIdx.x; k < m; k += STATS_THREADS) a0[k] = __dsub_rn(T[k], mu0);
    __syncthreads();
    const int i = i0 + threadIdx.x;
    if (i >= sublen) return;
    const double mean = mu[i];
    double df = 0.0, dg = 0.0;
    if (i > 0) {
        const double prev_mu = mu[i - 1];
        const double hi = win[threadIdx.x + m - 1], lo = T[i - 1];
        df = __ddiv_rn(__dsub_rn(hi, lo), 2.0);                                   // :30
        dg = __dadd_rn(__dsub_rn(hi, mean), __dsub_rn(lo, prev_mu));              // :32
    }
    double qt = 0.0, ss = 0.0;
    const double *w = win + threadIdx.x;
#pragma unroll 4
    for (int k = 0; k < m; ++k) {
        const double x = __dsub_rn(w[k], mean);
        qt = __dadd_rn(qt, __dmul_rn(x, a0[k]));                                  // :37
        ss = __dadd_rn(ss, __dmul_rn(x, x));                                      // :38
    }
    const double iv = __ddiv_rn(1.0, __dsqrt_rn(ss));                            // :41
    df_dg[2 * (size_t)i] = (OutT)df;                                              // cast to data_t on store (:30-41)
    df_dg[2 * (size_t)i + 1] = (OutT)dg;
    inv[i] = (OutT)iv;
    qt0[i] = (OutT)qt;
}

// Threshold operands of the tile kernel: {norm_i = 1 / inv_i, inv_i} pairs, and the minimum of norm over the MNC_WIN
// columns that start at i (the columns one lane covers in one block of the hot loop).  Beyond sublen inv is the zero padding, so norm = +inf there ("never a candidate").
template <typename OutT>
__global__ void norm_kernel(const OutT *__restrict__ inv, int sublen, int np, OutT *__restrict__ nm, OutT *__restrict__ mnc) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np) return;
    const OutT inf = (OutT)__longlong_as_double(0x7ff0000000000000LL);
    OutT first = inf, mn = inf;
#pragma unroll
    for (int k = 0; k < MNC_WIN; ++k) {
        const int c = i + k;
        const OutT v = c < sublen ? (OutT)1 / inv[c] : inf;
        if (k == 0) first = v;
        mn = v < mn ? v : mn;
    }
    nm[2 * (size_t)i] = first;
    nm[2 * (size_t)i + 1] = i < sublen ? inv[i] : (OutT)0;
    mnc[i] = mn;
}

// ---------------------------------------------------------------------------------------------
// Seed profile: the first diagonal outside the exclusion zone (lag m/4), every cell from its own centred dot
// product, P = (QT * inv_i) * inv_j (kernel/MatrixProfileKernelVanilla.cpp:84), applied to both the row's and the
// column's slot.  These are ordinary cells of the matrix (the tile kernel computes them again), so the result is
// unchanged; but subsequences at lag m/4 overlap in 3m/4 samples, so every slot starts from a high correlation
// and the tile kernel's filter rejects almost everything from its first row on -- on every GPU alike, whichever
// bands it owns.  Cost: m FMAs per subsequence (about 0.1 ms at n = 2^20).  Skipped in MPB_FLAG_STRICT, where the
// value must carry the reference chain's exact rounding.
// ---------------------------------------------------------------------------------------------
template <typename OutT>
__global__ void __launch_bounds__(STATS_THREADS)
seed_thresholds_kernel(const double *__restrict__ T, const double *__restrict__ mu, const OutT *__restrict__ inv, int n, int m, int lag,
                       unsigned long long *thr, double rel, double mu_abs) {
    extern __shared__ double sm[];                   // T[i0 .. i0 + STATS_THREADS + m + lag)
    const int sublen = n - m + 1;
    const int i0 = blockIdx.x * STATS_THREADS;
    for (int k = threadIdx.x; k < STATS_THREADS + m + lag; k += STATS_THREADS) sm[k] = (i0 + k < n) ? T[i0 + k] : 0.0;
    __syncthreads();
    const int i = i0 + threadIdx.x, j = i + lag;
    if (j >= sublen) return;
    const double mi = mu[i], mj = mu[j];
    const double *a = sm + threadIdx.x, *b = a + lag;
    double q0 = 0.0, q1 = 0.0;
    int k = 0;
    for (; k + 1 < m; k += 2) {
        q0 = fma(a[k] - mi, b[k] - mj, q0);
        q1 = fma(a[k + 1] - mi, b[k + 1] - mj, q1);
    }
    if (k < m) q0 = fma(a[k] - mi, b[k] - mj, q0);
    const OutT qt = (OutT)(q0 + q1);
    const double p = (double)((qt * inv[i]) * inv[j]);
    if (p > 0.0) {
        const double lb = p - fmax(rel * p, mu_abs * fmax(p, MPB_CHAIN_FLOOR));
        if (lb > 0.0) {
            const unsigned long long bits = (unsigned long long)__double_as_longlong(lb);
            atomicMax(thr + i, bits);
            atomicMax(thr + j, bits);
        }
    }
}

__global__ void profile_init_kernel(ProfileEntry *prof, int count) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) { ProfileEntry e; e.corr = 0.0; e.index = -1; e.pad = 0; prof[i] = e; }
}

// Gathers the engine's SoA arrays into the reference's InputDataPack[sublen] (parity checks only).
template <typename OutT>
__global__ void pack_input_kernel(const OutT *__restrict__ df_dg, const OutT *__restrict__ inv, const OutT *__restrict__ qt0,
                                  int sublen, InputDataPack<OutT> *__restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < sublen) { InputDataPack<OutT> p; p.QT = qt0[i]; p.df = df_dg[2 * (size_t)i]; p.dg = df_dg[2 * (size_t)i + 1]; p.inv = inv[i]; out[i] = p; }
}

struct MuLayout { size_t plocal, hd, hg, hm, tail, carry, absmax, bad, total, hcnt, hbase, hrank, hpos, bytes; };
static MuLayout mu_layout(int n, int m) {
    const size_t sub = (size_t)(n - m + 1), nch = (sub + MUP_CHUNK - 1) / MUP_CHUNK, padded = nch * MUP_CHUNK, nch2 = (nch + 1) & ~(size_t)1;
    MuLayout L; size_t o = 0;
    L.plocal = o; o += 8 * padded;  L.hd = o; o += 8 * padded;  L.hg = o; o += 8 * padded;  L.hm = o; o += 8 * padded;
    L.tail = o; o += 8 * nch;  L.carry = o; o += 8 * nch;  L.absmax = o; o += 8;  L.bad = o; o += 4;  L.total = o; o += 4;
    L.hcnt = o; o += 4 * nch2;  L.hbase = o; o += 4 * nch2;  L.hrank = o; o += 2 * padded;  L.hpos = o; o += 2 * padded;
    L.bytes = o + 256;
    return L;
}
size_t mu_scan_scratch_bytes(int n, int m) { return mu_layout(n, m).bytes; }
size_t mu_scan_bad_offset(int n, int m) { return mu_layout(n, m).bad; }   // where the "verification failed" flag of mode 1 lives inside the scratch

// mode 0: the plain sequential walk; 1 (default): parallel + verification, sequential walk only if a check fails
cudaError_t launch_mu_scan(const double *T, int n, int m, double *mu, void *scratch, int mode, cudaStream_t s) {
    const size_t smem = sizeof(double) * (size_t)(MUP_CHUNK + 1 + m);
    if (mode == 0 || scratch == nullptr || smem > 200 * 1024) {
        mu_scan_kernel<<<1, SCAN_THREADS, 0, s>>>(T, n, m, mu, nullptr);
        return cudaGetLastError();
    }
    const int sublen = n - m + 1, nch = (sublen + MUP_CHUNK - 1) / MUP_CHUNK;
    const MuLayout L = mu_layout(n, m);
    unsigned char *b = (unsigned char *)scratch;
    MuScratch sc;
    sc.plocal = (double *)(b + L.plocal); sc.hd = (double *)(b + L.hd); sc.hg = (double *)(b + L.hg); sc.hm = (double *)(b + L.hm);
    sc.tail = (double *)(b + L.tail); sc.carry = (double *)(b + L.carry); sc.absmax = (unsigned long long *)(b + L.absmax);
    sc.bad = (int *)(b + L.bad); sc.total = (int *)(b + L.total); sc.hcnt = (int *)(b + L.hcnt); sc.hbase = (int *)(b + L.hbase);
    sc.hrank = (unsigned short *)(b + L.hrank); sc.hpos = (unsigned short *)(b + L.hpos);
    cudaError_t e = cudaMemsetAsync(sc.absmax, 0, 16, s);          // absmax, bad, total
    if (e != cudaSuccess) return e;
    mu_absmax_kernel<<<296, 256, 0, s>>>(T, n, sc.absmax);
    if (smem > 48 * 1024) {
        e = cudaFuncSetAttribute(mu_classify_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    mu_classify_kernel<<<nch, MUP_THREADS, smem, s>>>(T, n, m, sc);
    mu_chunk_scan_kernel<<<1, 1024, 0, s>>>(nch, sc);
    mu_compact_kernel<<<nch, 128, 0, s>>>(T, n, m, sc);
    mu_hard_kernel<<<1, MUP_THREADS, 0, s>>>(T, m, sc);
    mu_fill_kernel<<<(sublen + 255) / 256, 256, 0, s>>>(T, n, m, sc, mu);
    mu_scan_kernel<<<1, SCAN_THREADS, 0, s>>>(T, n, m, mu, sc.bad);
    return cudaGetLastError();
}

cudaError_t launch_mu_window(const double *T, int n, int m, double *mu, cudaStream_t s) {
    const int sublen = n - m + 1;
    const size_t smem = sizeof(double) * (size_t)(MUW_THREADS + m);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(mu_window_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    mu_window_kernel<<<(sublen + MUW_THREADS - 1) / MUW_THREADS, MUW_THREADS, smem, s>>>(T, n, m, mu);
    return cudaGetLastError();
}

template <typename OutT>
static cudaError_t launch_stats_t(const double *T, const double *mu, int n, int m, OutT *df_dg, OutT *inv, OutT *qt0, cudaStream_t s) {
    const int sublen = n - m + 1;
    const size_t smem = sizeof(double) * (size_t)(STATS_THREADS + 2 * m);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(stats_kernel<OutT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    stats_kernel<OutT><<<(sublen + STATS_THREADS - 1) / STATS_THREADS, STATS_THREADS, smem, s>>>(T, mu, n, m, df_dg, inv, qt0);
    return cudaGetLastError();
}
cudaError_t launch_stats_f64(const double *T, const double *mu, int n, int m, double *df_dg, double *inv, double *qt0, cudaStream_t s) {
    return launch_stats_t<double>(T, mu, n, m, df_dg, inv, qt0, s);
}
cudaError_t launch_stats_f32(const double *T, const double *mu, int n, int m, float *df_dg, float *inv, float *qt0, cudaStream_t s) {
    return launch_stats_t<float>(T, mu, n, m, df_dg, inv, qt0, s);
}
cudaError_t launch_norm_f64(const double *inv, int sublen, int np, double *nm, double *mnc, cudaStream_t s) {
    norm_kernel<double><<<(np + 255) / 256, 256, 0, s>>>(inv, sublen, np, nm, mnc);
    return cudaGetLastError();
}
cudaError_t launch_norm_f32(const float *inv, int sublen, int np, float *nm, float *mnc, cudaStream_t s) {
    norm_kernel<float><<<(np + 255) / 256, 256, 0, s>>>(inv, sublen, np, nm, mnc);
    return cudaGetLastError();
}
template <typename OutT>
static cudaError_t launch_seed_t(const double *T, const double *mu, const OutT *inv, int n, int m, int lag, unsigned long long *thr, double rel, double mu_abs, cudaStream_t s) {
    const int sublen = n - m + 1;
    if (sublen - lag <= 0) return cudaSuccess;
    const size_t smem = sizeof(double) * (size_t)(STATS_THREADS + m + lag);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(seed_thresholds_kernel<OutT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    seed_thresholds_kernel<OutT><<<(sublen - lag + STATS_THREADS - 1) / STATS_THREADS, STATS_THREADS, smem, s>>>(T, mu, inv, n, m, lag, thr, rel, mu_abs);
    return cudaGetLastError();
}
cudaError_t launch_seed_thresholds_f64(const double *T, const double *mu, const double *inv, int n, int m, int lag, unsigned long long *thr, double rel, double mu_abs, cudaStream_t s) {
    return launch_seed_t<double>(T, mu, inv, n, m, lag, thr, rel, mu_abs, s);
}
cudaError_t launch_seed_thresholds_f32(const double *T, const double *mu, const float *inv, int n, int m, int lag, unsigned long long *thr, double rel, double mu_abs, cudaStream_t s) {
    return launch_seed_t<float>(T, mu, inv, n, m, lag, thr, rel, mu_abs, s);
}
cudaError_t launch_profile_init(ProfileEntry *prof, int count, cudaStream_t s) {
    profile_init_kernel<<<(count + 255) / 256, 256, 0, s>>>(prof, count);
    return cudaGetLastError();
}
cudaError_t launch_pack_input_f64(const double *df_dg, const double *inv, const double *qt0, int sublen, void *out, cudaStream_t s) {
    pack_input_kernel<double><<<(sublen + 255) / 256, 256, 0, s>>>(df_dg, inv, qt0, sublen, (InputDataPack<double> *)out);
    return cudaGetLastError();
}
cudaError_t launch_pack_input_f32(const float *df_dg, const float *inv, const float *qt0, int sublen, void *out, cudaStream_t s) {
    pack_input_kernel<float><<<(sublen + 255) / 256, 256, 0, s>>>(df_dg, inv, qt0, sublen, (InputDataPack<float> *)out);
    return cudaGetLastError();
}

}  // namespace mpb
