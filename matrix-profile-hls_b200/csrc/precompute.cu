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
mu_scan_kernel(const double *__restrict__ T, int n, int m, double *__restrict__ mu) {
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
seed_profile_kernel(const double *__restrict__ T, const double *__restrict__ mu, const OutT *__restrict__ inv, int n, int m, int lag,
                    ProfileEntry *prof, double borrow) {
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
        if (borrow > 0.0) {     // chain-carry mode (borrow = the near-tie margin): a lower bound, four margins below the cell; the
            const double lb = chain_lower(p, 4.0 * borrow);                        // tile kernel meets the cell itself and replaces it
            if (lb > 0.0) { profile_update(prof + i, lb, -1); profile_update(prof + j, lb, -1); }
        } else {
            profile_update(prof + i, p, j);
            profile_update(prof + j, p, i);
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

cudaError_t launch_mu_scan(const double *T, int n, int m, double *mu, cudaStream_t s) {
    mu_scan_kernel<<<1, SCAN_THREADS, 0, s>>>(T, n, m, mu);
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
static cudaError_t launch_seed_t(const double *T, const double *mu, const OutT *inv, int n, int m, int lag, ProfileEntry *prof, double borrow, cudaStream_t s) {
    const int sublen = n - m + 1;
    if (sublen - lag <= 0) return cudaSuccess;
    const size_t smem = sizeof(double) * (size_t)(STATS_THREADS + m + lag);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(seed_profile_kernel<OutT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    seed_profile_kernel<OutT><<<(sublen - lag + STATS_THREADS - 1) / STATS_THREADS, STATS_THREADS, smem, s>>>(T, mu, inv, n, m, lag, prof, borrow);
    return cudaGetLastError();
}
cudaError_t launch_seed_profile_f64(const double *T, const double *mu, const double *inv, int n, int m, int lag, ProfileEntry *prof, double borrow, cudaStream_t s) {
    return launch_seed_t<double>(T, mu, inv, n, m, lag, prof, borrow, s);
}
cudaError_t launch_seed_profile_f32(const double *T, const double *mu, const float *inv, int n, int m, int lag, ProfileEntry *prof, double borrow, cudaStream_t s) {
    return launch_seed_t<float>(T, mu, inv, n, m, lag, prof, borrow, s);
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
