// scamp_tile.cu -- kernel (2): the SCAMP diagonal-tile kernel.
//
// What it replaces: both HLS kernels (kernel/MatrixProfileKernelVanilla.cpp:57-106,
// kernel/MatrixProfileKernelTiled.cpp:149-223) and the host-side chunk merge UpdateAggregates
// (include/host/HostSideComputation.hpp:72-90), fused.
//
// Decomposition.  The upper triangle is cut into parallelogram work items: BAND_W = 1024 consecutive
// diagonals x seg_rows rows.  One CTA walks one item; each of its 4 warps owns 256 of the diagonals and
// lane l keeps the QT values of DPT = 8 adjacent diagonals in registers, so inside an item every diagonal
// is one unbroken chain
//     QT <- fma(df_r, dg_c, fma(df_c, dg_r, QT));   P = (QT * inv_r) * inv_c        (2 DFMA + 2 DMUL)
// (kernel/MatrixProfileKernelVanilla.cpp:81,84).  The first item of a band starts from the precomputed
// row-0 seeds QT_0 (:43-55); later items re-seed every diagonal with the exact centred dot product
//     QT(r, c) = sum_k (T[r+k] - mu_r)(T[c+k] - mu_c)
// computed in fp64 from the series (BASELINE.json north_star: "periodic QT re-seeding to bound drift").
// With seg_rows = 0 (MPB_FLAG_NO_RESEED / MPB_FLAG_STRICT) a band is a single item and every diagonal is
// the reference's own chain from row 0 to the end.
// Items are ordered band-major, so the CTAs that run at the same time work on the same band and
// different row segments: disjoint rows AND disjoint columns, which keeps the thresholds below fresh.
//
// Staging.  Every WARP stages its own tile: TILE_H = 128 rows and the TILE_H + 256 columns they touch, in
// its private slice of shared memory, synchronised with __syncwarp only -- there is no CTA-wide barrier in
// the main loop (ncu of the CTA-shared version showed 48 % of warp time parked at __syncthreads).  Column
// data is stored de-interleaved by (column mod 8) so that the 32 lanes, whose columns are 8 apart, read
// consecutive shared-memory words (conflict-free LDS.128).
//
// Max/argmax.  The profile (one 16-byte {corr, index} slot per subsequence, row and column roles merged,
// include/host/HostSideComputation.hpp:61) lives in HBM/L2.  An fp64 compare is a DSETP on the FP64 pipe,
// so the hot loop never compares doubles.  Each tile snapshots the HIGH 32 BITS of the current best
// correlation of its rows and columns (for P > 0 the IEEE pattern is monotone; P <= 0 has the sign bit set
// or is zero and never passes; the profile starts at +0.0, the reference kernels' own P = 0 floor) and, per
// column, the minimum of that snapshot over the 8-column window that ends there (cm8).  Per thread-row
// (8 cells) the loop takes max3-trees of hi32(P) (VIMNMX3) and tests it once against
// min(row threshold, cm8 of the lane's current 8-column window).  Only thread-rows that pass take the
// cold path: exact per-cell tests, a 128-bit compare-and-swap on the profile slot (mp_common.cuh), and a
// shared-memory atomicMax that tightens the thresholds for the rest of the warp at once; the profile updates
// themselves are queued per warp and applied in parallel at tile boundaries (flush_queue).
#include "mp_common.cuh"
#include "mp_kernels.h"

namespace mpb {

// Shared-memory records: rowA[i] = {df, dg}, rowB[i] = {inv, threshold}; colA[p] = {df, dg},
// colB[p] = {inv, cm8, exact column threshold}.  Each is fetched with one vector load.
template <typename T> struct Tr;
template <> struct Tr<double> {
    using T2 = double2;
    struct __align__(16) RowB { double inv; int thr; int pad; };
    struct __align__(16) ColB { double inv; int cm8; int cthr; };
    struct RowBV { double inv; int thr; };
    struct ColV { double inv; int cm8; };
    static __device__ __forceinline__ RowBV load_rowb(const RowB *r) {
        const double2 b = *reinterpret_cast<const double2 *>(r);
        return {b.x, __double2loint(b.y)};
    }
    static __device__ __forceinline__ ColV load_colb(const ColB *c) {
        const double2 b = *reinterpret_cast<const double2 *>(c);
        return {b.x, __double2loint(b.y)};
    }
};
template <> struct Tr<float> {
    using T2 = float2;
    struct __align__(8) RowB { float inv; int thr; };
    struct __align__(16) ColB { float inv; int cm8; int cthr; int pad; };
    struct RowBV { float inv; int thr; };
    struct ColV { float inv; int cm8; };
    static __device__ __forceinline__ RowBV load_rowb(const RowB *r) {
        const float2 b = *reinterpret_cast<const float2 *>(r);
        return {b.x, __float_as_int(b.y)};
    }
    static __device__ __forceinline__ ColV load_colb(const ColB *c) {
        const float2 b = *reinterpret_cast<const float2 *>(c);
        return {b.x, __float_as_int(b.y)};
    }
};

struct WarpCtx {          // per-warp context the cold path reads back from shared memory
    ProfileEntry *prof;
    int r0, c0;           // global row of tile row 0, global column of window column 0
    unsigned slow;        // profile updates attempted by this warp
    unsigned qn;          // candidates pushed since the last flush (may exceed QCAP: the excess went straight to HBM)
};
static_assert(sizeof(WarpCtx) <= 32, "ctx slot");

// Candidate queue.  A candidate that passed the exact per-cell test is parked here (16 bytes) and the warp
// goes straight back to the FP64 loop; the 128-bit compare-and-swaps on the profile are issued later, by all
// lanes in parallel, so their L2 round trips overlap instead of stalling the warp once per candidate
// (ncu: 28 % of warp time sat in the blocking version of this path).
struct __align__(16) QEntry { double val; int slot; int idx; };
constexpr int QCAP = 128;
constexpr int QFLUSH = 64;    // flush at a tile boundary once this many candidates wait

constexpr int SEED_KC = TILE_H;   // k-chunk of the seed dot products (series samples staged per pass)

constexpr int ROWS_STAGED = TILE_H + 8;   // TILE_H rows + the first row of the next tile (pipelined QT step) + prefetch slack

template <typename T>
struct WarpSmem {
    using T2 = typename Tr<T>::T2;
    using RowB = typename Tr<T>::RowB;
    using ColB = typename Tr<T>::ColB;
    static constexpr int BYTES = 32 + ROWS_STAGED * (int)(sizeof(T2) + sizeof(RowB)) + DPT * COL_STRIDE * (int)(sizeof(T2) + sizeof(ColB)) + QCAP * (int)sizeof(QEntry);
    WarpCtx *ctx; T2 *rowA; RowB *rowB; T2 *colA; ColB *colB; QEntry *queue;
    __device__ __forceinline__ explicit WarpSmem(unsigned char *base) {
        ctx = reinterpret_cast<WarpCtx *>(base);
        rowA = reinterpret_cast<T2 *>(base + 32);
        rowB = reinterpret_cast<RowB *>(rowA + ROWS_STAGED);
        colA = reinterpret_cast<T2 *>(rowB + ROWS_STAGED);
        colB = reinterpret_cast<ColB *>(colA + DPT * COL_STRIDE);
        queue = reinterpret_cast<QEntry *>(colB + DPT * COL_STRIDE);
    }
};
static_assert(WarpSmem<double>::BYTES % 16 == 0 && WarpSmem<float>::BYTES % 16 == 0, "warp slices stay 16-byte aligned");
static_assert(ROWS_STAGED * sizeof(float2) >= SEED_KC * sizeof(double), "seed operand A aliases the rowA buffer");
static_assert(DPT * COL_STRIDE * sizeof(float2) >= (SEED_KC + WARP_DIAGS + DPT) * sizeof(double), "seed window U aliases colA");

__device__ __forceinline__ int col_slot(int x) { return (x & 7) * COL_STRIDE + (x >> 3); }

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

extern __shared__ __align__(16) unsigned char smem_raw[];

// ---------------------------------------------------------------------------------------------
// Cold path: thread-row (tile row i, this lane's 8 diagonals) passed the hi32 filter.  Exact per-cell tests
// against the shared thresholds; survivors are queued and the thresholds raised to the candidate's own key
// at once (the candidate WILL be applied, so anything below it can be dropped from now on).
// ---------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void push_candidate(WarpSmem<T> &S, double val, int slot, int idx) {
    const unsigned pos = atomicAdd(&S.ctx->qn, 1u);
    if (pos < (unsigned)QCAP) { QEntry e; e.val = val; e.slot = slot; e.idx = idx; S.queue[pos] = e; }
    else profile_update(S.ctx->prof + slot, val, idx);     // queue full (start-up burst): blocking update
}

template <typename T>
__device__ __noinline__ void cold_row(T p0, T p1, T p2, T p3, T p4, T p5, T p6, T p7, int i) {
    WarpSmem<T> S(smem_raw + (threadIdx.x >> 5) * WarpSmem<T>::BYTES);
    const T P[DPT] = {p0, p1, p2, p3, p4, p5, p6, p7};
    const int x0 = i + DPT * (threadIdx.x & 31);    // window column of diagonal j = 0
    const int row = S.ctx->r0 + i, c0 = S.ctx->c0;
    volatile int *rthr_p = &S.rowB[i].thr;
    const int rthr = *rthr_p;
    int bestj = -1;
    T bestp = 0;
#pragma unroll
    for (int j = 0; j < DPT; ++j) {
        const int key = filter_key(P[j]);
        if (key >= rthr && (bestj < 0 || P[j] > bestp)) { bestj = j; bestp = P[j]; }   // best of the row; ties keep the smaller column
        volatile int *cp = &S.colB[col_slot(x0 + j)].cthr;
        if (key >= *cp) {
            push_candidate<T>(S, (double)P[j], c0 + x0 + j, row);
            atomicMax((int *)cp, key);
        }
    }
    if (bestj >= 0) {
        push_candidate<T>(S, (double)bestp, row, c0 + x0 + bestj);
        atomicMax((int *)rthr_p, filter_key(bestp));
    }
}

// Applies the queued candidates to the profile: one candidate per lane, all lanes at once. Warp-collective.
template <typename T>
__device__ __noinline__ void flush_queue() {
    WarpSmem<T> S(smem_raw + (threadIdx.x >> 5) * WarpSmem<T>::BYTES);
    __syncwarp();
    const unsigned pushed = S.ctx->qn;
    const int nq = pushed < (unsigned)QCAP ? (int)pushed : QCAP;
    ProfileEntry *prof = S.ctx->prof;
    for (int e = threadIdx.x & 31; e < nq; e += 32) {
        const QEntry q = S.queue[e];
        profile_update(prof + q.slot, q.val, q.idx);
    }
    __syncwarp();
    if ((threadIdx.x & 31) == 0) { S.ctx->slow += pushed; S.ctx->qn = 0; }
    __syncwarp();
}

// ---------------------------------------------------------------------------------------------
// Exact re-seed of this lane's DPT diagonals at row rs (segment start - 1), in fp64 from the series:
//   QT(rs, c) = sum_k A[k] * (T[c+k] - mu_ref) - (mu_c - mu_ref) * sum_k A[k],   A[k] = T[rs+k] - mu_rs,
// algebraically sum_k (T[rs+k]-mu_rs)(T[c+k]-mu_c) (include/host/HostSideComputation.hpp:37 generalised from
// row 0 to row rs); mu_ref is one mean shared by the warp so the shifted window is staged once for all lanes.
// ---------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void reseed(T (&qt)[DPT], WarpSmem<T> &S, const double *__restrict__ Tser, const double *__restrict__ mu,
                                       int n, int m, int sublen, int rs, int d0w, int lane) {
    double *A = reinterpret_cast<double *>(S.rowA);      // aliases the rowA buffer
    double *U = reinterpret_cast<double *>(S.colA);      // aliases colA (de-interleaved like the columns)
    const double mur = mu[rs];
    const double muref = mu[min(rs + d0w, sublen - 1)];
    double acc[DPT], sumA = 0.0;
#pragma unroll
    for (int j = 0; j < DPT; ++j) acc[j] = 0.0;
    for (int kc = 0; kc < m; kc += SEED_KC) {
        __syncwarp();
        for (int k = lane; k < SEED_KC; k += 32) A[k] = (kc + k < m) ? __dsub_rn(Tser[rs + kc + k], mur) : 0.0;
        for (int y = lane; y < SEED_KC + WARP_DIAGS + DPT; y += 32) {
            const long long idx = (long long)rs + d0w + kc + y;
            U[col_slot(y)] = idx < n ? __dsub_rn(Tser[idx], muref) : 0.0;
        }
        __syncwarp();
        double w[DPT];
#pragma unroll
        for (int k = 0; k < DPT; ++k) w[k] = U[k * COL_STRIDE + lane];
        for (int K = 0; K < SEED_KC / DPT; ++K) {
#pragma unroll
            for (int kk = 0; kk < DPT; ++kk) {
                const double a = A[K * DPT + kk];
#pragma unroll
                for (int j = 0; j < DPT; ++j) acc[j] = fma(a, w[(kk + j) & 7], acc[j]);
                sumA += a;
                w[kk] = U[kk * COL_STRIDE + lane + K + 1];
            }
        }
    }
#pragma unroll
    for (int j = 0; j < DPT; ++j) {
        const int c = min(rs + d0w + lane * DPT + j, sublen - 1);
        qt[j] = (T)(acc[j] - (mu[c] - muref) * sumA);
    }
    __syncwarp();   // A/U are about to be overwritten by the first tile
}

// item_start[lb] = first work item of this rank's local band lb (nlb + 1 entries)
template <typename T, bool STRICT>
__global__ void __launch_bounds__(BAND_THREADS, 3)
tile_kernel(const typename Tr<T>::T2 *__restrict__ df_dg, const T *__restrict__ inv, const T *__restrict__ qt0,
            const double *__restrict__ Tser, const double *__restrict__ mu, ProfileEntry *prof,
            const int *__restrict__ item_start, int nlb, int n, int m, int sublen, int excl, int rank, int world, int seg_rows,
            unsigned long long *counter) {
    using T2 = typename Tr<T>::T2;
    using RowB = typename Tr<T>::RowB;
    using ColB = typename Tr<T>::ColB;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WarpSmem<T> S(smem_raw + warp * WarpSmem<T>::BYTES);

    // work item -> (local band, segment): binary search in the prefix table
    int lo = 0, hi = nlb;
    const int item = blockIdx.x;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (item_start[mid] <= item) lo = mid; else hi = mid; }
    const int seg = item - item_start[lo];
    const int band = rank + lo * world;
    const int d0w = excl + band * BAND_W + warp * WARP_DIAGS;   // first (longest) diagonal of this warp
    const int nrows = sublen - d0w;                             // rows of that diagonal
    const int rbeg = seg_rows > 0 ? seg * seg_rows : 0;
    const int rend = seg_rows > 0 ? min(rbeg + seg_rows, nrows) : nrows;
    if (rbeg >= rend) return;                                   // warp-uniform; no CTA-wide barrier follows

    T qt[DPT];
    if (seg == 0) {
#pragma unroll
        for (int j = 0; j < DPT; ++j) qt[j] = qt0[d0w + lane * DPT + j];   // zero padding beyond sublen
    } else {
        reseed<T>(qt, S, Tser, mu, n, m, sublen, rbeg - 1, d0w, lane);
    }
    if (lane == 0) { S.ctx->prof = prof; S.ctx->slow = 0; S.ctx->qn = 0; }

    // Software pipeline: on entry to row i the registers hold QT(row i) already; the loop body turns it into
    // P(row i), then issues the QT step of row i+1 BEFORE testing P(row i), so the max-tree / compare / branch
    // latency of one row hides behind the FMAs of the next.  Each tile therefore stages TILE_H + 1 rows.
    bool first_tile = true;
    for (int r0 = rbeg; r0 < rend; r0 += TILE_H) {
        const int c0 = r0 + d0w;
        __syncwarp();   // previous tile fully consumed
        if (S.ctx->qn >= (unsigned)QFLUSH) flush_queue<T>();
        if (lane == 0) { S.ctx->r0 = r0; S.ctx->c0 = c0; }
#pragma unroll
        for (int i = lane; i < ROWS_STAGED; i += 32) {
            S.rowA[i] = df_dg[r0 + i];
            RowB rb;
            rb.inv = inv[r0 + i];
            rb.thr = threshold_of(__ldcg(&prof[r0 + i].corr), T(0));
            S.rowB[i] = rb;
        }
#pragma unroll 4
        for (int x = lane; x < COL_WIN; x += 32) {
            const int p = col_slot(x);
            S.colA[p] = df_dg[c0 + x];
            S.colB[p].inv = inv[c0 + x];
            S.colB[p].cthr = threshold_of(__ldcg(&prof[c0 + x].corr), T(0));
        }
        __syncwarp();
        for (int x = DPT + lane; x < COL_WIN; x += 32) {   // cm8[x] = min of the exact thresholds of columns [x-8, x)
            int mn = S.colB[col_slot(x - 1)].cthr;
#pragma unroll
            for (int e = 2; e <= DPT; ++e) mn = min(mn, S.colB[col_slot(x - e)].cthr);
            S.colB[col_slot(x)].cm8 = mn;
        }
        __syncwarp();

        // column registers: slot k holds window column x = i + 8*lane + ((k - i) mod 8)
        T cdf[DPT], cdg[DPT], civ[DPT];
#pragma unroll
        for (int k = 0; k < DPT; ++k) {
            const int p = k * COL_STRIDE + lane;
            const T2 a = S.colA[p];
            cdf[k] = a.x; cdg[k] = a.y; civ[k] = S.colB[p].inv;
        }
        if (first_tile) {                                  // pipeline prologue: QT(row rbeg)
            const T2 ra = S.rowA[0];
#pragma unroll
            for (int j = 0; j < DPT; ++j) qt[j] = qt_step<STRICT>(qt[j], ra.x, ra.y, cdf[j], cdg[j]);
            first_tile = false;
        }
        const T2 *rowA = S.rowA + 1;
        const RowB *rowB = S.rowB;
        const T2 *colA = S.colA + lane + 1;
        const ColB *colB = S.colB + lane + 1;
        typename Tr<T>::RowBV rb = Tr<T>::load_rowb(rowB);
#pragma unroll 1
        for (int I = 0; I < TILE_H / DPT; ++I) {
#pragma unroll
            for (int ii = 0; ii < DPT; ++ii) {
                const T2 ra = rowA[ii];                                             // {df, dg} of row i + 1
                const typename Tr<T>::RowBV rbn = Tr<T>::load_rowb(rowB + ii + 1);  // {inv, thr} of row i + 1
                // next column entering this lane's window: x = i + 8*lane + 8 -> class ii, slot lane + I + 1;
                // its cm8 is the threshold minimum over the CURRENT window [i + 8*lane, i + 8*lane + 8)
                const T2 nd = colA[ii * COL_STRIDE];
                const typename Tr<T>::ColV nb = Tr<T>::load_colb(colB + ii * COL_STRIDE);
                T P[DPT];
#pragma unroll
                for (int j = 0; j < DPT; ++j) P[j] = corr_of(qt[j], rb.inv, civ[(ii + j) & 7]);
                const int thr = min(rb.thr, nb.cm8);
                cdf[ii] = nd.x; cdg[ii] = nd.y; civ[ii] = nb.inv;
#pragma unroll
                for (int j = 0; j < DPT; ++j) {
                    const int k = (ii + 1 + j) & 7;
                    qt[j] = qt_step<STRICT>(qt[j], ra.x, ra.y, cdf[k], cdg[k]);     // QT(row i + 1)
                }
                const int m0 = __vimax3_s32(filter_key(P[0]), filter_key(P[1]), filter_key(P[2]));
                const int m1 = __vimax3_s32(filter_key(P[3]), filter_key(P[4]), filter_key(P[5]));
                const int m2 = __vimax3_s32(filter_key(P[6]), filter_key(P[7]), m0);
                if (__builtin_expect(max(m1, m2) >= thr, 0))
                    cold_row<T>(P[0], P[1], P[2], P[3], P[4], P[5], P[6], P[7], I * DPT + ii);
                rb = rbn;
            }
            rowA += DPT; rowB += DPT; colA += 1; colB += 1;
        }
    }
    flush_queue<T>();
    if (lane == 0 && counter && S.ctx->slow) atomicAdd(counter, (unsigned long long)S.ctx->slow);
}

int band_count(int sublen, int excl) {
    const int ndiag = sublen - excl;
    return ndiag <= 0 ? 0 : (ndiag + BAND_W - 1) / BAND_W;
}

// Prefix table of work items for the bands of `rank`: band b = rank + lb * world owns
// ceil(rows(b) / seg_rows) items (one when seg_rows == 0). Returns the total item count.
long long build_item_table(int sublen, int excl, int rank, int world, int seg_rows, std::vector<int> &item_start) {
    const int nb = band_count(sublen, excl);
    const int nlb = nb > rank ? (nb - rank + world - 1) / world : 0;
    item_start.assign((size_t)nlb + 1, 0);
    long long total = 0;
    for (int lb = 0; lb < nlb; ++lb) {
        const int band = rank + lb * world;
        const int nrows = sublen - (excl + band * BAND_W);
        item_start[lb] = (int)total;
        total += seg_rows > 0 ? (nrows + seg_rows - 1) / seg_rows : 1;
    }
    item_start[nlb] = (int)total;
    return total;
}

// Default segment length: long enough that the re-seed (m FMAs per diagonal per segment against
// 4 * seg_rows pipe operations) stays under ~2 %, short enough that a rank has several waves of items.
int default_seg_rows(int sublen, int excl, int m, int world, int sm_count) {
    int floor_rows = ((16 * m + 255) / 256) * 256;
    if (floor_rows < 1024) floor_rows = 1024;
    int seg = floor_rows < 4096 ? 4096 : floor_rows;
    std::vector<int> tab;
    const long long want = 6LL * 3 * sm_count;            // >= 6 waves of 3 CTAs per SM
    while (seg / 2 >= floor_rows && build_item_table(sublen, excl, 0, world, seg, tab) < want) seg /= 2;
    return seg;
}

template <typename T>
static cudaError_t launch_tiles_t(const T *df_dg, const T *inv, const T *qt0, ProfileEntry *prof, const TileLaunch &p, cudaStream_t s) {
    using T2 = typename Tr<T>::T2;
    if (p.items <= 0) return cudaSuccess;
    if (p.seg_rows % TILE_H != 0 || p.seg_rows < 0) return cudaErrorInvalidValue;
    const size_t smem = (size_t)(BAND_THREADS / 32) * WarpSmem<T>::BYTES;
    auto kern = p.strict ? tile_kernel<T, true> : tile_kernel<T, false>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    kern<<<(unsigned)p.items, BAND_THREADS, smem, s>>>(reinterpret_cast<const T2 *>(df_dg), inv, qt0, p.series, p.mu, prof, p.item_start, p.nlb,
                                                       p.n, p.m, p.sublen, p.excl, p.rank, p.world, p.seg_rows, p.slow_counter);
    return cudaGetLastError();
}

cudaError_t launch_tiles_f64(const double *df_dg, const double *inv, const double *qt0, ProfileEntry *prof, const TileLaunch &p, cudaStream_t s) {
    return launch_tiles_t<double>(df_dg, inv, qt0, prof, p, s);
}
cudaError_t launch_tiles_f32(const float *df_dg, const float *inv, const float *qt0, ProfileEntry *prof, const TileLaunch &p, cudaStream_t s) {
    return launch_tiles_t<float>(df_dg, inv, qt0, prof, p, s);
}

}  // namespace mpb
