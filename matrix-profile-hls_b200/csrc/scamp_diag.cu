// scamp_diag.cu -- kernel (2): the SCAMP diagonal-band kernel.
//
// What it replaces: both HLS kernels (kernel/MatrixProfileKernelVanilla.cpp:57-106,
// kernel/MatrixProfileKernelTiled.cpp:149-223) and the host-side chunk merge UpdateAggregates
// (include/host/HostSideComputation.hpp:72-90), fused.
//
// Mapping.  A CTA owns a band of BAND_W = 1024 consecutive diagonals and walks it from row 0 to its
// last row; each thread keeps the QT values of DPT = 8 adjacent diagonals in registers for the whole
// walk, so every diagonal is ONE unbroken chain seeded once from QT_0 (precompute.cu) -- the same
// sequence of operations as the reference (kernel/MatrixProfileKernelVanilla.cpp:43-55,81), no re-seed.
// Per cell: QT <- fma(df_r, dg_c, fma(df_c, dg_r, QT)); P = QT * inv_r * inv_c  (2 DFMA + 2 DMUL).
// Rows are staged TILE_H at a time in shared memory together with the TILE_H + BAND_W columns the
// tile touches; column data is stored de-interleaved by (column mod 8) so that the 32 lanes of a warp,
// whose columns are 8 apart, read consecutive shared-memory words (conflict-free LDS.128/LDS.64).
//
// Max/argmax.  The profile (one 16-byte {corr, index} slot per subsequence, row and column roles
// merged) lives in HBM/L2.  Comparing every P against two running maxima in fp64 would cost two more
// FP64-pipe instructions per cell (DSETP), so instead each tile snapshots the HIGH 32 BITS of the
// current best correlation of its rows and columns into shared memory, and the inner loop tests
// hi32(P) >= threshold with integer min/compare instructions on the INT pipe (for P > 0 the IEEE bit
// pattern is monotone; negative P has the sign bit set and never passes; the profile starts at +0.0,
// which is the reference kernels' own P = 0 floor).  Only the rare candidates take
// the exact path: a 128-bit compare-and-swap on the profile slot (mp_common.cuh) plus an atomicMax of
// the shared threshold so the rest of the CTA sees the tighter bound at once.
#include "mp_common.cuh"
#include "mp_kernels.h"

namespace mpb {

template <typename T> struct Vec2;
template <> struct Vec2<double> { using type = double2; };
template <> struct Vec2<float> { using type = float2; };

template <typename T> struct __align__(16) RowPack;
template <> struct __align__(16) RowPack<double> { double df, dg, inv; int thr; int pad; };
template <> struct __align__(16) RowPack<float> { float df, dg, inv; int thr; };

// Filter key: monotone in P for P > 0; never passes for P <= 0 (thresholds are >= 1).
__device__ __forceinline__ int filter_key(double p) { return __double2hiint(p); }
__device__ __forceinline__ int filter_key(float p) { return __float_as_int(p); }
__device__ __forceinline__ int threshold_of(double best, double) { const int k = __double2hiint(best); return k < 1 ? 1 : k; }
__device__ __forceinline__ int threshold_of(double best, float) { const int k = __float_as_int((float)best); return k < 1 ? 1 : k; }

template <bool STRICT> __device__ __forceinline__ double qt_step(double qt, double dfr, double dgr, double dfc, double dgc) {
    if (STRICT) return __dadd_rn(qt, __dadd_rn(__dmul_rn(dfr, dgc), __dmul_rn(dfc, dgr)));   // the reference's rounding sequence (:81)
    return fma(dfr, dgc, fma(dfc, dgr, qt));
}
template <bool STRICT> __device__ __forceinline__ float qt_step(float qt, float dfr, float dgr, float dfc, float dgc) {
    if (STRICT) return __fadd_rn(qt, __fadd_rn(__fmul_rn(dfr, dgc), __fmul_rn(dfc, dgr)));
    return fmaf(dfr, dgc, fmaf(dfc, dgr, qt));
}
__device__ __forceinline__ double corr_of(double qt, double ir, double ic) { return __dmul_rn(__dmul_rn(qt, ir), ic); }   // :84
__device__ __forceinline__ float corr_of(float qt, float ir, float ic) { return __fmul_rn(__fmul_rn(qt, ir), ic); }

// Exact path (rare): update the HBM profile slot, then tighten the shared threshold.
template <typename T>
__device__ __noinline__ void exact_update(ProfileEntry *slot, T p, int idx, int *shared_thr, unsigned long long *counter) {
    const double best = profile_update(slot, (double)p, idx);
    atomicMax(shared_thr, threshold_of(best, T(0)));
    if (counter) atomicAdd(counter, 1ull);
}

template <typename T, bool STRICT>
__global__ void __launch_bounds__(BAND_THREADS, 3)
band_kernel(const typename Vec2<T>::type *__restrict__ df_dg, const T *__restrict__ inv, const T *__restrict__ qt0,
            ProfileEntry *__restrict__ prof, int sublen, int excl, int rank, int world, unsigned long long *counter) {
    using T2 = typename Vec2<T>::type;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    RowPack<T> *rows = reinterpret_cast<RowPack<T> *>(smem_raw);
    T2 *c_dfdg = reinterpret_cast<T2 *>(rows + TILE_H);
    T *c_inv = reinterpret_cast<T *>(c_dfdg + DPT * COL_STRIDE);
    int *c_thr = reinterpret_cast<int *>(c_inv + DPT * COL_STRIDE);

    const int t = threadIdx.x;
    const int band = rank + blockIdx.x * world;
    const int d0 = excl + band * BAND_W;          // first (longest) diagonal of the band
    const int nrows = sublen - d0;                // rows of the longest diagonal
    if (nrows <= 0) return;

    T qt[DPT];
#pragma unroll
    for (int j = 0; j < DPT; ++j) qt[j] = qt0[d0 + t * DPT + j];   // zero padding beyond sublen

    for (int r0 = 0; r0 < nrows; r0 += TILE_H) {
        const int c0 = r0 + d0;
        __syncthreads();   // previous tile fully consumed
        for (int i = t; i < TILE_H; i += BAND_THREADS) {
            const T2 a = df_dg[r0 + i];
            RowPack<T> rp;
            rp.df = a.x; rp.dg = a.y; rp.inv = inv[r0 + i];
            rp.thr = threshold_of(prof[r0 + i].corr, T(0));
            rows[i] = rp;
        }
        for (int x = t; x < COL_WIN; x += BAND_THREADS) {
            const int p = (x & 7) * COL_STRIDE + (x >> 3);
            c_dfdg[p] = df_dg[c0 + x];
            c_inv[p] = inv[c0 + x];
            c_thr[p] = threshold_of(prof[c0 + x].corr, T(0));
        }
        __syncthreads();

        // column registers: slot k holds window column x = i + 8t + ((k - i) mod 8)
        T cdf[DPT], cdg[DPT], civ[DPT];
        int cth[DPT];
#pragma unroll
        for (int k = 0; k < DPT; ++k) {
            const int p = k * COL_STRIDE + t;
            const T2 a = c_dfdg[p];
            cdf[k] = a.x; cdg[k] = a.y; civ[k] = c_inv[p]; cth[k] = c_thr[p];
        }
        for (int I = 0; I < TILE_H / 8; ++I) {
#pragma unroll
            for (int ii = 0; ii < 8; ++ii) {
                const int i = I * 8 + ii;
                const RowPack<T> rp = rows[i];
                // next column entering this thread's window: x = i + 8t + 8 -> class ii, slot t + I + 1
                const int pn = ii * COL_STRIDE + t + I + 1;
                const T2 nd = c_dfdg[pn];
                const T niv = c_inv[pn];
                const int nth = c_thr[pn];
                T P[DPT];
                bool hit = false;
#pragma unroll
                for (int j = 0; j < DPT; ++j) {
                    const int k = (ii + j) & 7;
                    qt[j] = qt_step<STRICT>(qt[j], rp.df, rp.dg, cdf[k], cdg[k]);
                    P[j] = corr_of(qt[j], rp.inv, civ[k]);
                    hit |= filter_key(P[j]) >= min(rp.thr, cth[k]);
                }
                if (hit) {
                    const int row = r0 + i;
#pragma unroll
                    for (int j = 0; j < DPT; ++j) {
                        const int k = (ii + j) & 7;
                        const int key = filter_key(P[j]);
                        const int col = row + d0 + t * DPT + j;
                        if (key >= rp.thr) exact_update<T>(prof + row, P[j], col, &rows[i].thr, counter);
                        if (key >= cth[k]) {
                            const int x = i + t * DPT + j;
                            exact_update<T>(prof + col, P[j], row, &c_thr[(x & 7) * COL_STRIDE + (x >> 3)], counter);
                        }
                    }
                }
                cdf[ii] = nd.x; cdg[ii] = nd.y; civ[ii] = niv; cth[ii] = nth;
            }
        }
    }
}

int band_count(int sublen, int excl) {
    const int ndiag = sublen - excl;
    return ndiag <= 0 ? 0 : (ndiag + BAND_W - 1) / BAND_W;
}

template <typename T>
static size_t band_smem_bytes() {
    return sizeof(RowPack<T>) * TILE_H + (size_t)DPT * COL_STRIDE * (2 * sizeof(T) + sizeof(T) + sizeof(int));
}

template <typename T>
static cudaError_t launch_bands_t(const T *df_dg, const T *inv, const T *qt0, ProfileEntry *prof, const BandLaunch &p,
                                  int *launched_ctas, cudaStream_t s) {
    using T2 = typename Vec2<T>::type;
    const int nb = band_count(p.sublen, p.excl);
    const int mine = nb > p.rank ? (nb - p.rank + p.world - 1) / p.world : 0;
    if (launched_ctas) *launched_ctas = mine;
    if (mine == 0) return cudaSuccess;
    const size_t smem = band_smem_bytes<T>();
    auto kern = p.strict ? band_kernel<T, true> : band_kernel<T, false>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    kern<<<mine, BAND_THREADS, smem, s>>>(reinterpret_cast<const T2 *>(df_dg), inv, qt0, prof, p.sublen, p.excl, p.rank, p.world, p.slow_counter);
    return cudaGetLastError();
}

cudaError_t launch_bands_f64(const double *df_dg, const double *inv, const double *qt0, ProfileEntry *prof,
                             const BandLaunch &p, int *launched_ctas, cudaStream_t s) {
    return launch_bands_t<double>(df_dg, inv, qt0, prof, p, launched_ctas, s);
}
cudaError_t launch_bands_f32(const float *df_dg, const float *inv, const float *qt0, ProfileEntry *prof,
                             const BandLaunch &p, int *launched_ctas, cudaStream_t s) {
    return launch_bands_t<float>(df_dg, inv, qt0, prof, p, launched_ctas, s);
}

}  // namespace mpb
