// scamp_tile.cu -- kernel (2): the SCAMP diagonal-tile kernel.
//
// What it replaces: both HLS kernels (kernel/MatrixProfileKernelVanilla.cpp:57-106,
// kernel/MatrixProfileKernelTiled.cpp:149-223) and the host-side chunk merge UpdateAggregates
// (include/host/HostSideComputation.hpp:72-90), fused.
//
// Decomposition.  The upper triangle is cut into parallelogram work items: BAND_W consecutive diagonals x
// seg_rows rows.  One CTA walks one item; each of its warps owns WARP_DIAGS of the diagonals and lane l
// keeps the QT values of DPT adjacent diagonals in registers, so inside an item every diagonal is one
// unbroken chain
//     QT <- fma(df_r, dg_c, fma(df_c, dg_r, QT))                                      (2 DFMA per cell)
// (kernel/MatrixProfileKernelVanilla.cpp:81).  The first item of a band starts from the precomputed row-0
// seeds QT_0 (:43-55); later items re-seed every diagonal with the exact centred dot product
//     QT(r, c) = sum_k (T[r+k] - mu_r)(T[c+k] - mu_c)
// computed in fp64 from the series (BASELINE.json north_star: "periodic QT re-seeding to bound drift").
// With seg_rows = 0 (MPB_FLAG_NO_RESEED / MPB_FLAG_STRICT) a band is a single item and every diagonal is
// the reference's own chain from row 0 to the end.
//
// Max/argmax without touching P.  The reference turns every QT into P = QT * inv_r * inv_c (:84, 2 DMUL per
// cell) and compares it with the row and column aggregates (:87-96).  Here the comparison is moved to the QT
// domain: a cell can only improve row r or column c if
//     QT >= norm_r * norm_c * min(best_r, best_c),        norm = 1 / inv,
// and for a BLOCK of cells -- the DPT = 8 diagonals of a lane over RB = 8 consecutive rows, i.e. rows
// [i0, i0 + 8) x columns [y, y + 15) -- that is implied by
//     QT >= min( min_r(best_r * norm_r) * min_c norm_c ,  min_r norm_r * min_c (best_c * norm_c) )   =: theta
// with the minima taken over the block's rows and its 15-column window.  The hot loop therefore only keeps a
// running maximum of the HIGH 32 BITS of its QT values (VIMNMX3 on the ALU pipe: 4 per 8 cells; an fp64
// compare would be a DSETP on the FP64 pipe) and tests it ONCE per block against theta (2 DMUL per 64 cells).
// Row-block minima and column-window minima are staged per tile; the row terms are scaled by (1 - 2^-18),
// which more than covers every rounding on the way, so the filter has no false negatives: every cell that
// beats the current profile makes its block pass.  The FP64 pipe carries 2 DFMA per cell + 2 DMUL per 64
// cells (2.03 pipe ops per cell against the reference formulation's 4).
//
// Hits.  A block that passes is replayed from its starting QT values (same operations, same bits), this time
// with the per-row form of the test; the thread-rows that pass it park their 8 QT values in a small per-warp
// hit buffer.  At the end of the tile (or when the buffer fills) the warp drains the buffer with all lanes in
// parallel: each cell repeats the test on its own against the tile's per-column keys; survivors compute exact
// P = (QT * inv_r) * inv_c (:84, the reference's rounding order) and apply it to the column's and the row's
// profile slot with a 128-bit compare-and-swap (mp_common.cuh: larger corr wins, equal corr -> smaller index;
// deterministic for any arrival order or GPU count).  The profile (one 16-byte {corr, index} slot per
// subsequence, row and column roles merged, HostSideComputation.hpp:61) lives in HBM/L2 and is only ever
// touched by drains and by the per-tile threshold snapshots.
//
// Staging.  Every WARP stages its own tile (TILE_H rows and the TILE_H + WARP_DIAGS columns they touch) in
// its private slice of shared memory, synchronised with __syncwarp only -- there is no CTA-wide barrier in
// the main loop.  Column data is stored de-interleaved by (column mod DPT) so that the 32 lanes, whose
// columns are DPT apart, read consecutive shared-memory words (conflict-free LDS.128).
#include "mp_common.cuh"
#include "mp_kernels.h"

namespace mpb {

template <typename T> struct Tr;
template <> struct Tr<double> { using T2 = double2; };
template <> struct Tr<float> { using T2 = float2; };

static_assert(DPT == 8 && DPT % RB == 0 && NWIN == 1, "this kernel version: 8 diagonals per lane, blocks of 8 or 4 rows");
constexpr int BPG = DPT / RB;             // filter blocks per group of DPT rows (one rotation of the column registers)
constexpr int HCAP = 48;                  // hit buffer capacity (thread-rows of DPT cells)
constexpr int HDRAIN = HCAP - 32;         // drain once more than this many thread-rows wait (a row can add up to 32)
constexpr int SEED_KC = TILE_H;           // k-chunk of the seed dot products (series samples staged per pass)
constexpr int ROWS_STAGED = TILE_H + 8;   // TILE_H rows + the first row of the next tile (pipelined QT step) + prefetch slack
constexpr int ROW_BLOCKS = ROWS_STAGED / RB;
constexpr int COL_SLOTS = DPT * COL_STRIDE;
constexpr unsigned FULL = 0xffffffffu;

// A cell that improved (or, in the chain-carry mode, came within the margin of) the warp's view of a slot:
// code = index of the cell in the hit buffer (hit * DPT + j) | role << 16 (0: the cell's column slot, 1: its row slot).
struct __align__(16) Winner { double v; int code; int pad; };
struct __align__(16) WarpLog { ChainLogEntry *log; unsigned long long *count; unsigned cap; int km; double mu; };   // chain-carry mode: this warp's sub-log, the near-tie margin, and the margin in high-word ulps (lower_key)
#ifndef MPB_WCAP
#define MPB_WCAP 96
#endif
#ifndef MPB_STAGE_BATCH
#define MPB_STAGE_BATCH 1
#endif
// Blocks of the hot loop per loop iteration.  With 2 the checkpoint / QT register sets alternate and ptxas drops ~27 of
// the ~38 loop-carried register copies per block; the fp64 loop is issue-bound (a DFMA costs two issue slots, every
// other instruction one: scripts/dfma_occupancy_probe.cu), so that is 4 % of the step.  The fp32 loop measured slower
// unrolled (profiles/r01_v8_variants_unroll_occupancy.jsonl).
#ifndef MPB_HOT_UNROLL_F64
#define MPB_HOT_UNROLL_F64 2
#endif
#ifndef MPB_HOT_UNROLL_F32
#define MPB_HOT_UNROLL_F32 1
#endif
template <typename T> __host__ __device__ constexpr int hot_unroll() { return sizeof(T) == 8 ? MPB_HOT_UNROLL_F64 : MPB_HOT_UNROLL_F32; }
constexpr int WCAP = MPB_WCAP;                  // winner list capacity (a drain pass adds up to 32 column + 4 row winners; 32 + 32 in the chain-carry mode)

template <typename T>
struct WarpSmem {
    using T2 = typename Tr<T>::T2;
    static constexpr int NK = ((COL_WIN + RB - 1) / RB + 3) / 4 * 4;   // block windows start every RB columns
    static constexpr int HITS_BYTES = HCAP * DPT * (int)sizeof(T) + HCAP * 8;
    static constexpr int BYTES = ROWS_STAGED * (int)sizeof(T2) + ROWS_STAGED * 16 + COL_SLOTS * (int)sizeof(T2) + COL_SLOTS * (int)sizeof(T) + COL_SLOTS * 8 +
                                 HITS_BYTES + (ROW_BLOCKS + 1) / 2 * 2 * (int)sizeof(T2) + 2 * NK * (int)sizeof(T) + 16 + WCAP * 16 + (int)sizeof(WarpLog) + (int)((sizeof(ThrShare) + 15) / 16 * 16);
    T2 *rowA;         // [i] = {df_r, dg_r}
    double2 *rowC;    // [i] = {best correlation of the row as this warp knows it, inv_r}
    T2 *colA;         // [slot(x)] = {df_c, dg_c}
    T *colInv;        // [slot(x)] = inv_c
    double *colBest;  // [slot(x)] = best correlation of the column as this warp knows it
    T *hitQ;          // [h][DPT] QT values of a parked thread-row
    int2 *hitM;       // [h] = {tile row i, window column x of its first cell}
    int *ck;          // staging scratch (aliases the empty hit buffer): [slot(x)] = filter key of best_c * norm_c
    T2 *rowK;         // [b] = minima of {best_r * norm_r * s, norm_r * s} over the rows [RB*b, RB*b + RB),  s = 1 - 2^-18
    T *mncK;          // [q] = min of norm over the MNC_WIN columns starting at window column RB*q
    T *wK;            // [q] = min of best_c * norm_c over the same columns (key precision)
    unsigned *stat;   // [0] thread-rows parked, [1] profile slots improved, [2] replay passes (32 lane-blocks wide), [3] lane-blocks queued for the replay
    Winner *wlist;    // cells that improved the warp's view of a slot, waiting for the profile update
    WarpLog *wlog;    // chain-carry mode: where this warp appends its candidates
    ThrShare *wsh;    // the job's threshold arrays (own + peers): where improvements are pushed
    __device__ __forceinline__ explicit WarpSmem(unsigned char *base) {
        rowA = reinterpret_cast<T2 *>(base);
        rowC = reinterpret_cast<double2 *>(rowA + ROWS_STAGED);
        colA = reinterpret_cast<T2 *>(rowC + ROWS_STAGED);
        colInv = reinterpret_cast<T *>(colA + COL_SLOTS);
        colBest = reinterpret_cast<double *>(colInv + COL_SLOTS);
        hitQ = reinterpret_cast<T *>(colBest + COL_SLOTS);
        hitM = reinterpret_cast<int2 *>(hitQ + HCAP * DPT);
        ck = reinterpret_cast<int *>(hitQ);
        rowK = reinterpret_cast<T2 *>(hitM + HCAP);
        mncK = reinterpret_cast<T *>(rowK + (ROW_BLOCKS + 1) / 2 * 2);
        wK = mncK + NK;
        stat = reinterpret_cast<unsigned *>(wK + NK);
        wlist = reinterpret_cast<Winner *>(stat + 4);
        wlog = reinterpret_cast<WarpLog *>(wlist + WCAP);
        wsh = reinterpret_cast<ThrShare *>(wlog + 1);
    }
};
static_assert(WarpSmem<double>::BYTES % 16 == 0 && WarpSmem<float>::BYTES % 16 == 0, "warp slices stay 16-byte aligned");
static_assert((ROWS_STAGED * sizeof(float2)) % 16 == 0 && (COL_SLOTS * sizeof(float)) % 16 == 0 && (HCAP * DPT * sizeof(float)) % 16 == 0, "sub-arrays stay 16-byte aligned");
static_assert(WarpSmem<float>::HITS_BYTES >= COL_SLOTS * 4, "the column-key scratch fits the hit buffer");
static_assert(ROWS_STAGED * sizeof(float2) >= SEED_KC * sizeof(double), "seed operand A aliases rowA");
static_assert(COL_SLOTS * sizeof(float2) >= COL_SLOTS * sizeof(double), "seed window U aliases colA");
static_assert(SEED_KC + WARP_DIAGS + DPT <= COL_WIN && SEED_KC % DPT == 0, "seed window fits the column slots");

template <typename T> constexpr int cta_smem_bytes() { return BAND_WARPS * WarpSmem<T>::BYTES; }
static_assert(4 * (BAND_WARPS * WarpSmem<double>::BYTES + 1024) <= 227 * 1024 || BAND_WARPS != 2, "four 2-warp fp64 CTAs per SM must fit the 227 KB of shared memory");
template <typename T> constexpr int ctas_per_sm() {
    return (227 * 1024) / (cta_smem_bytes<T>() + 1024) < 1 ? 1 : ((227 * 1024) / (cta_smem_bytes<T>() + 1024) > 8 ? 8 : (227 * 1024) / (cta_smem_bytes<T>() + 1024));
}

__device__ __forceinline__ int col_slot(int x) { return (x % DPT) * COL_STRIDE + (x / DPT); }

// Filter key: the high word of the IEEE pattern, monotone for positive values; negative values have the sign
// bit set (negative as an int) and never pass.
__device__ __forceinline__ int key_of(double v) { return __double2hiint(v); }
__device__ __forceinline__ int key_of(float v) { return __float_as_int(v); }
__device__ __forceinline__ double slack(double) { return 1.0 - 1.0 / 262144.0; }
__device__ __forceinline__ float slack(float) { return 1.0f - 1.0f / 262144.0f; }
// +inf as a threshold means "never passes" and survives every product with a positive finite factor; the staged
// terms are arranged so that inf never meets 0 (0 * inf is a NaN with the sign bit set, i.e. a key that always passes).
__device__ __forceinline__ double pos_inf(double) { return __longlong_as_double(0x7ff0000000000000LL); }
__device__ __forceinline__ float pos_inf(float) { return __int_as_float(0x7f800000); }
__device__ __forceinline__ double tiny_best(double) { return 1e-300; }
__device__ __forceinline__ float tiny_best(float) { return 1e-30f; }
__device__ __forceinline__ void set_key(double &dst, int key) { dst = __hiloint2double(key, 0); }
__device__ __forceinline__ void set_key(float &dst, int key) { dst = __int_as_float(key); }

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
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }

// Chain-carry mode: a threshold lowered by (at least) the near-tie margin, on the ALU pipe.  hw = high word of a
// non-negative double; km = margin in high-word ulps (one ulp is 2^-21 .. 2^-20 of the value): the result is >= km ulps
// lower for values above CHAIN_FLOOR = 2^-6 (exponent field 1017) and a constant km * 2^-26 lower below it (the shift
// grows by one per binade), which is chain_lower()'s rule rounded towards "lower".
__device__ __forceinline__ int lower_key(int hw, int km) {
    const int sh = min(max(1017 - (hw >> 20), 0), 10);
    return max(hw - (km << sh), 0);
}
__device__ __forceinline__ double lower_thr(double v, int km) { return __hiloint2double(lower_key(__double2hiint(v), km), __double2loint(v)); }

extern __shared__ __align__(16) unsigned char smem_raw[];

// 64-bit max on a shared-memory double that is known to be >= +0.0 (positive IEEE patterns order like integers)
__device__ __forceinline__ double smem_max_pos(double *addr, double v) {
    return __longlong_as_double((long long)atomicMax(reinterpret_cast<unsigned long long *>(addr), (unsigned long long)__double_as_longlong(v)));
}

// ---------------------------------------------------------------------------------------------
// Drain: resolves the parked thread-rows, one CELL per lane (hit e / DPT, diagonal e % DPT), 32 / DPT thread-rows
// per pass.  Everything a cell needs is in the tile's shared memory: exact P = (QT * inv_r) * inv_c (:84, the
// reference's rounding order) is compared with the warp's own view of the column's best (atomic max, so of the
// cells that compete for a column in one pass only improvements survive) and -- after a shuffle max over the
// thread-row -- of the row's best.  Only cells that improve that view touch the profile in HBM/L2, with a 128-bit
// compare-and-swap (mp_common.cuh: larger corr wins, equal corr -> smaller index).  Warp-collective.
// ---------------------------------------------------------------------------------------------
// Chain-carry mode: (corr, index) max into a threshold slot with the three-way rule of mp_common.cuh.  A candidate
// more than the margin above the slot's best replaces it, one more than the margin below is dropped, and within the
// margin the loser of the comparison -- the candidate, or the entry it replaces -- goes to the near-tie log.  Entries
// with index -1 (borrowed lower bounds) are thresholds, not cells: they are never logged.
__device__ __forceinline__ void log_near_tie(const WarpLog *wl, double corr, int slot, int idx) {
    const unsigned long long pos = atomicAdd(wl->count, 1ull);
    if (pos < wl->cap) { ChainLogEntry le; le.corr = corr; le.slot = slot; le.index = idx; wl->log[pos] = le; }
}
__device__ __forceinline__ double profile_update_log(ProfileEntry *prof, int slot_id, double v, int idx, const WarpLog *wl) {
    ProfileEntry *slot = prof + slot_id;
    const double mu = wl->mu;
    unsigned long long clo, chi, olo, ohi;
    asm volatile("ld.global.relaxed.gpu.v2.u64 {%0,%1}, [%2];" : "=l"(clo), "=l"(chi) : "l"(slot) : "memory");
    const unsigned long long nlo = (unsigned long long)__double_as_longlong(v);
    const unsigned long long nhi = (unsigned long long)(unsigned)idx;
    while (true) {
        const double cv = __longlong_as_double((long long)clo);
        const unsigned ci = (unsigned)chi;
        const bool near = v >= chain_lower(cv, mu) && cv >= chain_lower(v, mu);
        if (!(v > cv || (v == cv && (unsigned)idx < ci))) {
            if (near && !(v == cv && (unsigned)idx == ci)) log_near_tie(wl, v, slot_id, idx);
            break;
        }
        asm volatile(
            "{\n .reg .b128 c, d, o;\n mov.b128 c, {%2, %3};\n mov.b128 d, {%4, %5};\n"
            " atom.global.relaxed.gpu.cas.b128 o, [%6], c, d;\n mov.b128 {%0, %1}, o;\n}"
            : "=l"(olo), "=l"(ohi) : "l"(clo), "l"(chi), "l"(nlo), "l"(nhi), "l"(slot) : "memory");
        if (olo == clo && ohi == chi) {
            if (near && ci != 0xffffffffu) log_near_tie(wl, cv, slot_id, (int)ci);
            clo = nlo;
            break;
        }
        clo = olo; chi = ohi;
    }
    return __longlong_as_double((long long)clo);
}

// applies the winner list to the profile: one entry per lane, so the memory round trips of a whole batch overlap.
template <typename T, bool LOG>
__device__ __forceinline__ unsigned apply_winners(WarpSmem<T> &S, int cnt, int r0, int c0, ProfileEntry *prof, int lane) {
    unsigned wins = 0;
    __syncwarp();
    for (int e = lane; e < cnt; e += 32) {
        const Winner w = S.wlist[e];
        const int cell = w.code & 0xffff, h = cell / DPT, j = cell % DPT;
        const int2 mt = S.hitM[h];
        const int row = r0 + mt.x, col = c0 + mt.y + j;
        const int slot = (w.code >> 16) ? row : col, idx = (w.code >> 16) ? col : row;
        const double after = LOG ? profile_update_log(prof, slot, w.v, idx, S.wlog) : profile_update(prof + slot, w.v, idx);
        if (after == w.v) {                      // the rank's best for this slot: raise the job's thresholds (own copy and every peer's)
            ++wins;
            const unsigned long long bits = (unsigned long long)__double_as_longlong(w.v);
            const int np = S.wsh->n;
            for (int k = 0; k < np; ++k) atomicMax(S.wsh->thr[k] + slot, bits);
        }
    }
    __syncwarp();
    return wins;
}

template <typename T, bool LOG>
__device__ __noinline__ void drain_hits(int nhits, int r0, int c0, ProfileEntry *prof) {
    const int lane = threadIdx.x & 31;
    WarpSmem<T> S(smem_raw + (threadIdx.x >> 5) * WarpSmem<T>::BYTES);
    __syncwarp();
    unsigned wins = 0;
    int wcnt = 0;                                            // warp-uniform
    const int ncell = nhits * DPT;
    const unsigned lt = (1u << lane) - 1u;
#pragma unroll 1
    for (int base = 0; base < ncell; base += 32) {
        const int e = base + lane;
        const bool act = e < ncell;
        const int h = act ? e / DPT : 0, j = e % DPT;
        const int2 mt = S.hitM[h];
        const int i = mt.x, x = mt.y + j;
        const int p = col_slot(x);
        const double2 rc = S.rowC[i];
        double pv = (double)corr_of(S.hitQ[h * DPT + j], (T)rc.y, S.colInv[p]);
        if (!act || !(pv > 0.0)) pv = -1.0;                  // P <= 0 and NaN never enter the profile (it starts at +0.0)
        // chain-carry mode: the warp's views hold the bests LOWERED by the near-tie margin (lower_thr is monotone, so the
        // max commutes with it): near misses are candidates too, and the replay's key tests need no extra work
        const double pl = LOG ? lower_thr(fmax(pv, 0.0), S.wlog->km) : pv;
        bool cwin = pv > 0.0 && pv >= S.colBest[p];
        if (cwin) cwin = pv >= smem_max_pos(&S.colBest[p], pl);
        double bv = pv;                                      // best of the thread-row; ties keep the smaller column
        int bx = x;
#pragma unroll
        for (int sft = 1; sft < DPT; sft <<= 1) {
            const double ov = __shfl_xor_sync(FULL, bv, sft);
            const int ox = __shfl_xor_sync(FULL, bx, sft);
            if (ov > bv || (ov == bv && ox < bx)) { bv = ov; bx = ox; }
        }
        bool rwin;
        int rcode = (h * DPT + (bx - mt.y)) | 0x10000;
        if (!LOG) {
            rwin = j == 0 && bv > 0.0 && bv >= rc.x;
            if (rwin) rwin = bv >= smem_max_pos(&S.rowC[i].x, bv);
        } else {
            // the thread-row's best goes into the warp's view of the row; every cell of the thread-row within the
            // margin of both that best and the row's previous best is a candidate
            double old = rc.x;
            const double bl = lower_thr(fmax(bv, 0.0), S.wlog->km);
            if (j == 0 && bv > 0.0 && bv >= rc.x) old = smem_max_pos(&S.rowC[i].x, bl);
            old = __shfl_sync(FULL, old, lane & ~(DPT - 1));
            rwin = pv > 0.0 && pv >= fmax(old, bl);
            rcode = e | 0x10000;
        }
        const unsigned cm = __ballot_sync(FULL, cwin), rm = __ballot_sync(FULL, rwin);
        if (cm | rm) {
            if (wcnt + __popc(cm) + __popc(rm) > WCAP) { wins += apply_winners<T, LOG>(S, wcnt, r0, c0, prof, lane); wcnt = 0; }
            if (cwin) { Winner w; w.v = pv; w.code = e; w.pad = 0; S.wlist[wcnt + __popc(cm & lt)] = w; }
            wcnt += __popc(cm);
            if (rwin) { Winner w; w.v = LOG ? pv : bv; w.code = rcode; w.pad = 0; S.wlist[wcnt + __popc(rm & lt)] = w; }
            wcnt += __popc(rm);
        }
    }
    if (wcnt) wins += apply_winners<T, LOG>(S, wcnt, r0, c0, prof, lane);
    wins = __reduce_add_sync(FULL, wins);
    if (lane == 0) { S.stat[0] += nhits; S.stat[1] += wins; }
    __syncwarp();
}

// ---------------------------------------------------------------------------------------------
// max of the filter keys of 8 values, optionally folded with a running maximum
template <typename T>
__device__ __forceinline__ int max_key8(const T *q) {
    const int a = __vimax3_s32(key_of(q[0]), key_of(q[1]), key_of(q[2]));
    const int b = __vimax3_s32(key_of(q[3]), key_of(q[4]), key_of(q[5]));
    const int c = __vimax3_s32(key_of(q[6]), key_of(q[7]), a);
    return max(b, c);
}
template <typename T>
__device__ __forceinline__ int max_key8(const T *q, int run) {
    const int a = __vimax3_s32(key_of(q[0]), key_of(q[1]), key_of(q[2]));
    const int b = __vimax3_s32(key_of(q[3]), key_of(q[4]), key_of(q[5]));
    const int c = __vimax3_s32(key_of(q[6]), key_of(q[7]), run);
    return __vimax3_s32(a, b, c);
}

template <typename T> struct QVec { T v[DPT]; };

// ---------------------------------------------------------------------------------------------
// Replay of one block.  The hot loop only knows that SOME cell of SOME lane's block passed the block-wide test.
// The block is walked again from the QT values it started with -- same operands, same order, bit-identical QT --
// and this time every cell's P = (QT * inv_r) * inv_c is formed (all lanes, 2 DMUL per cell) and its high word
// compared, as an integer, with the high word of the warp's view of the column's best and of the row's best.
// On smooth series (random walks) the QT-domain window tests are too loose to be worth an intermediate step:
// P lives within a percent of the best for whole neighbourhoods, and only a per-cell test in the P domain tells
// improvements from near misses.  Thread-rows in which a lane passes are parked for the drain.
// Column operands slide through registers exactly as in the hot loop.  Warp-collective; returns the new hit count.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int hi_word(const double *p) { return reinterpret_cast<const int *>(p)[1]; }

// PH = first row of the block inside its group of DPT rows (0, or RB when a group holds two blocks): the column registers are
// indexed by residue class, and the class of cell j in row PH + t is (PH + t + j) mod DPT.
template <typename T, bool STRICT, bool LOG, int PH>
__device__ __noinline__ int replay_block(QVec<T> qc, int g0, int r0, int c0, int hitcnt, ProfileEntry *prof) {
    using T2 = typename Tr<T>::T2;
    const int lane = threadIdx.x & 31;
    WarpSmem<T> S(smem_raw + (threadIdx.x >> 5) * WarpSmem<T>::BYTES);
    const int e0 = g0 / DPT + lane;                          // g0 (the group's first row) is a multiple of DPT: class k of entry e0 is column g0 + DPT*lane + k
    if (lane == 0) S.stat[2] += 1;
    T cdf[DPT], cdg[DPT], civ[DPT];
    int ckb[DPT];                                            // high word of the column's best (live view; >= 1 so that P <= 0 never passes)
#pragma unroll
    for (int k = 0; k < DPT; ++k) {                          // classes below PH already hold the next entry's column (they left the window earlier in the group)
        const int p = k * COL_STRIDE + e0 + (k < PH ? 1 : 0);
        const T2 a = S.colA[p];
        cdf[k] = a.x; cdg[k] = a.y;
        civ[k] = S.colInv[p];
        ckb[k] = max(hi_word(&S.colBest[p]), 1);
    }
#pragma unroll
    for (int t = 0; t < RB; ++t) {
        const int i = g0 + PH + t;
        const double2 rc = S.rowC[i];                        // {row's best, inv_r}
        const T ir = (T)rc.y;
        int kp[DPT];
        bool cand = false;
#pragma unroll
        for (int j = 0; j < DPT; ++j) {
            kp[j] = __double2hiint((double)corr_of(qc.v[j], ir, civ[(PH + t + j) % DPT]));
            cand |= kp[j] >= ckb[(PH + t + j) % DPT];
        }
        const int a = __vimax3_s32(kp[0], kp[1], kp[2]), b = __vimax3_s32(kp[3], kp[4], kp[5]);
        cand |= __vimax3_s32(kp[6], kp[7], max(a, b)) >= max(__double2hiint(rc.x), 1);
        // Thread-rows that pass the high-word test are parked as they are: the drain repeats the comparison exactly, cell
        // by cell, and a false candidate needs a P within 2^-20 of the current best (rare).
        const unsigned mk = __ballot_sync(FULL, cand);
        if (mk) {
            if (cand) {
                const int pos = hitcnt + __popc(mk & ((1u << lane) - 1u));
#pragma unroll
                for (int j = 0; j < DPT; ++j) S.hitQ[pos * DPT + j] = qc.v[j];
                S.hitM[pos] = make_int2(i, i + DPT * lane);
            }
            hitcnt += __popc(mk);
            if (hitcnt > HDRAIN) { drain_hits<T, LOG>(hitcnt, r0, c0, prof); hitcnt = 0; }
        }
        if (t + 1 < RB) {
            const T2 ra = S.rowA[i + 1];
            const int p = (PH + t) * COL_STRIDE + e0 + 1;    // column entering the window: g0 + PH + t + DPT*lane + DPT
            const T2 nd = S.colA[p];
            cdf[PH + t] = nd.x; cdg[PH + t] = nd.y;
            civ[PH + t] = S.colInv[p];
            ckb[PH + t] = max(hi_word(&S.colBest[p]), 1);
#pragma unroll
            for (int j = 0; j < DPT; ++j) {
                const int k = (PH + t + 1 + j) % DPT;
                qc.v[j] = qt_step<STRICT>(qc.v[j], ra.x, ra.y, cdf[k], cdg[k]);
            }
        }
    }
    return hitcnt;
}

// ---------------------------------------------------------------------------------------------
// Deferred, lane-compacted replay (RB == DPT builds).  On smooth series the block filter is loose: a fifth of the
// (warp, block) pairs have SOME lane passing, but only ~1 lane in 32 does (tests/tools/filter_sim.py; 70 % of the warp-wide
// replays parked nothing), so a warp-wide replay spends 31 lanes on blocks that did not pass.  Instead, a lane whose block
// passes only QUEUES it -- the block's starting QT values and its position, in the (empty) hit buffer -- and carries on.
// When QCAP lane-blocks wait, or at the end of the tile, the warp replays the queue with ONE QUEUED BLOCK PER LANE: the
// same walk as replay_block (same operands, same order, bit-identical QT; every cell's P = (QT * inv_r) * inv_c compared
// by its high word with the warp's view of the column's and the row's best), only that row and column positions now
// differ from lane to lane.  Blocks start at multiples of DPT rows, so the residue class of cell j in row t is (t + j)
// mod DPT in every lane and the column registers rotate statically, as in the hot loop.  The entries are in registers
// before the first thread-row is parked, so the queue and the hit buffer share their storage.
// Warp-collective; leaves the hit buffer drained.
// ---------------------------------------------------------------------------------------------
constexpr int QCAP = 32;                  // queued lane-blocks replayed in one pass
template <typename T, bool STRICT, bool LOG>
__device__ __noinline__ void flush_queue(int qcount, int r0, int c0, ProfileEntry *prof) {
    using T2 = typename Tr<T>::T2;
    const int lane = threadIdx.x & 31;
    WarpSmem<T> S(smem_raw + (threadIdx.x >> 5) * WarpSmem<T>::BYTES);
    __syncwarp();
    const bool act = lane < qcount;
    const int2 q = act ? S.hitM[lane] : make_int2(0, 0);     // {first row of the block (tile row, multiple of DPT), window column of its first cell}
    const int i0 = q.x, e0 = q.y / DPT;
    T qc[DPT];
#pragma unroll
    for (int j = 0; j < DPT; ++j) qc[j] = act ? S.hitQ[lane * DPT + j] : T(0);
    if (lane == 0) S.stat[2] += 1;
    __syncwarp();                                            // every entry is in registers: the buffer now takes parked thread-rows
    int hitcnt = 0;
    T cdf[DPT], cdg[DPT], civ[DPT];
    int ckb[DPT];                                            // high word of the column's best (live view; >= 1 so that P <= 0 never passes)
#pragma unroll
    for (int k = 0; k < DPT; ++k) {
        const int p = k * COL_STRIDE + e0;
        const T2 a = S.colA[p];
        cdf[k] = a.x; cdg[k] = a.y;
        civ[k] = S.colInv[p];
        ckb[k] = max(hi_word(&S.colBest[p]), 1);
    }
#pragma unroll
    for (int t = 0; t < RB; ++t) {
        const int i = i0 + t;
        const double2 rc = S.rowC[i];                        // {row's best, inv_r}
        const T ir = (T)rc.y;
        int kp[DPT];
        bool cand = false;
#pragma unroll
        for (int j = 0; j < DPT; ++j) {
            kp[j] = __double2hiint((double)corr_of(qc[j], ir, civ[(t + j) % DPT]));
            cand |= kp[j] >= ckb[(t + j) % DPT];
        }
        const int a = __vimax3_s32(kp[0], kp[1], kp[2]), b = __vimax3_s32(kp[3], kp[4], kp[5]);
        cand |= __vimax3_s32(kp[6], kp[7], max(a, b)) >= max(__double2hiint(rc.x), 1);
        cand &= act;
        const unsigned mk = __ballot_sync(FULL, cand);
        if (mk) {
            if (cand) {
                const int pos = hitcnt + __popc(mk & ((1u << lane) - 1u));
#pragma unroll
                for (int j = 0; j < DPT; ++j) S.hitQ[pos * DPT + j] = qc[j];
                S.hitM[pos] = make_int2(i, q.y + t);
            }
            hitcnt += __popc(mk);
            if (hitcnt > HDRAIN) { drain_hits<T, LOG>(hitcnt, r0, c0, prof); hitcnt = 0; }
        }
        if (t + 1 < RB) {
            const T2 ra = S.rowA[i + 1];
            const int p = t * COL_STRIDE + e0 + 1;           // column entering the lane's window
            const T2 nd = S.colA[p];
            cdf[t] = nd.x; cdg[t] = nd.y;
            civ[t] = S.colInv[p];
            ckb[t] = max(hi_word(&S.colBest[p]), 1);
#pragma unroll
            for (int j = 0; j < DPT; ++j) {
                const int k = (t + 1 + j) % DPT;
                qc[j] = qt_step<STRICT>(qc[j], ra.x, ra.y, cdf[k], cdg[k]);
            }
        }
    }
    if (hitcnt > 0) drain_hits<T, LOG>(hitcnt, r0, c0, prof);
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
    double *A = reinterpret_cast<double *>(S.rowA);      // aliases rowA
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
                for (int j = 0; j < DPT; ++j) acc[j] = fma(a, w[(kk + j) % DPT], acc[j]);
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
// counters[0] += thread-rows parked for the exact test, counters[1] += profile slots improved, counters[2] += blocks replayed
// LOG = chain-carry mode (mp_common.cuh): the item stores the QT values it started from and ended with, and the drain
// logs near-best candidates for the resolve kernel; the profile `prof` then only carries the filter thresholds.
template <typename T, bool STRICT, bool LOG>
__global__ void __launch_bounds__(BAND_THREADS, ctas_per_sm<T>())
tile_kernel(const typename Tr<T>::T2 *__restrict__ df_dg, const typename Tr<T>::T2 *__restrict__ nm, const T *__restrict__ mnc,
            const T *__restrict__ qt0, const double *__restrict__ Tser, const double *__restrict__ mu, ProfileEntry *prof,
            const int *__restrict__ item_start, int item_begin, int nlb, int n, int m, int sublen, int excl, int rank, int world, int seg_rows,
            unsigned long long *counters, const ChainFix cf, const ThrShare sh) {
    using T2 = typename Tr<T>::T2;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WarpSmem<T> S(smem_raw + warp * WarpSmem<T>::BYTES);

    // work item -> (local band, segment): 32-way search in the prefix table (three dependent loads for up to 32768 bands
    // instead of a binary search's fifteen: the item prologue was 2.4 % of the warp samples)
    int lo = 0, hi = nlb;
    const int item = blockIdx.x + item_begin;
    while (hi - lo > 1) {
        const int span = hi - lo, stepw = (span + 31) >> 5;
        const int probe = lo + min(lane * stepw, span - 1);      // probes lo, lo + stepw, ...: item_start[lo] <= item always holds
        const unsigned le = __ballot_sync(FULL, item_start[probe] <= item);
        const int k = 31 - __clz((int)le);                      // last probe that is <= item (bit 0 is always set)
        lo = lo + min(k * stepw, span - 1);
        hi = min(hi, lo + stepw);
    }
    const int seg = item - item_start[lo];
    const int band = item_start[nlb + 1 + lo];                  // band ids follow the prefix table
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

    int hitcnt = 0;                 // warp-uniform: lane-blocks queued for the replay (BPG == 1) / thread-rows parked (BPG > 1)
    if (lane < 4) S.stat[lane] = 0;
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < MAX_PEERS; ++k) S.wsh->thr[k] = sh.thr[k];
        S.wsh->n = sh.n; S.wsh->pad = 0;
    }
    const double *__restrict__ thr = reinterpret_cast<const double *>(sh.thr[0]);
    if (LOG && lane == 0) {
        const unsigned shard = (blockIdx.x * BAND_WARPS + warp) % LOG_SHARDS;
        WarpLog wl; wl.log = cf.log + (size_t)shard * cf.shard_cap; wl.count = cf.log_count + shard * 16; wl.cap = cf.shard_cap;
        wl.km = cf.key_margin; wl.mu = cf.margin;
        *S.wlog = wl;
    }
    const T sl = slack(T(0));
    const int km = LOG ? cf.key_margin : 0;               // chain-carry mode: every staged best is lowered by the near-tie margin

    // Software pipeline: on entry to row i the registers hold QT(row i) already; the loop body takes the filter
    // keys of QT(row i), then issues the QT step to row i+1 BEFORE branching on the test, so the max-tree /
    // compare / vote latency of one row hides behind the FMAs of the next.  Each tile stages TILE_H + 1 rows.
    {   // pipeline prologue: QT(row rbeg), operands straight from HBM/L2 (once per item)
        const T2 ra = df_dg[rbeg];
        const T2 *cp = df_dg + rbeg + d0w + lane * DPT;
#pragma unroll
        for (int j = 0; j < DPT; ++j) { const T2 c = cp[j]; qt[j] = qt_step<STRICT>(qt[j], ra.x, ra.y, c.x, c.y); }
        if (LOG && seg > 0) {                              // S_k: what this segment's chain holds at its first row
            double2 *dst = reinterpret_cast<double2 *>(cf.seg_seed + ((size_t)(item - cf.item_base) * BAND_W + warp * WARP_DIAGS + lane * DPT));
#pragma unroll
            for (int j = 0; j < DPT; j += 2) dst[j / 2] = make_double2((double)qt[j], (double)qt[j + 1]);
        }
    }
    for (int r0 = rbeg; r0 < rend; r0 += TILE_H) {
        const int c0 = r0 + d0w;
        __syncwarp();   // previous tile fully consumed
#pragma unroll
        for (int i0s = 0; i0s < ROWS_STAGED; i0s += 32) {  // rows: operands, the warp's view of their best, block minima of the threshold terms
            const int i = i0s + lane;
            T2 rb;
            rb.x = rb.y = pos_inf(T(0));
            if (i < ROWS_STAGED) {
                S.rowA[i] = df_dg[r0 + i];
                const T2 nv = nm[r0 + i];                       // {norm, inv}; norm = +inf on the padding
                double rbest = __ldcg(thr + r0 + i);
                if (LOG) rbest = lower_thr(rbest, km);
                S.rowC[i] = make_double2(rbest, (double)nv.y);
                const T best = max((T)rbest, tiny_best(T(0)));  // never 0: 0 * inf would be a NaN
                rb.y = mul_rn(nv.x, sl);
                rb.x = mul_rn(best, rb.y);
            }
            // minima over each block of RB = 8 consecutive rows (8 adjacent lanes), taken on the filter keys: the terms are
            // positive (or +inf), so their high words order like the values, and a key with the low word dropped is a
            // slightly smaller threshold term -- conservative.  (A double min is DSETP + SEL + FSEL + NaN fix-up and a
            // 64-bit shuffle two SHFLs: 75 instructions per 32 rows against 12 here.)
            int kx = key_of(rb.x), ky = key_of(rb.y);
#pragma unroll
            for (int sft = 1; sft < RB; sft <<= 1) {
                kx = min(kx, __shfl_xor_sync(FULL, kx, sft));
                ky = min(ky, __shfl_xor_sync(FULL, ky, sft));
            }
            if (lane % RB == 0 && i < ROWS_STAGED) { set_key(rb.x, kx); set_key(rb.y, ky); S.rowK[i / RB] = rb; }
        }
#if MPB_STAGE_BATCH
        {   // every global load of the column window is issued before the first one is consumed: one L2 round trip per
            // tile instead of one per unrolled group (the ncu source view of v6 had 20 % of the samples in this staging)
            constexpr int NCI = (COL_WIN + 31) / 32, NQI = (WarpSmem<T>::NK + 31) / 32;
            T2 ca[NCI], nv[NCI];
            double cb[NCI];
            T mq[NQI];
#pragma unroll
            for (int k = 0; k < NCI; ++k) {
                const int x = min(lane + 32 * k, COL_WIN - 1);
                ca[k] = df_dg[c0 + x];
                nv[k] = nm[c0 + x];                         // {norm_c, inv_c}
                cb[k] = __ldcg(thr + c0 + x);
                if (LOG) cb[k] = lower_thr(cb[k], km);
            }
#pragma unroll
            for (int k = 0; k < NQI; ++k) { const int q = lane + 32 * k; mq[k] = RB * q < COL_WIN ? mnc[c0 + RB * q] : pos_inf(T(0)); }
#pragma unroll
            for (int k = 0; k < NCI; ++k) {
                const int x = lane + 32 * k;
                if (x < COL_WIN) {
                    const int p = col_slot(x);
                    S.colA[p] = ca[k];
                    S.colBest[p] = cb[k];
                    S.colInv[p] = nv[k].y;
                    S.ck[p] = key_of(nv[k].x) < key_of(pos_inf(T(0))) ? key_of(mul_rn((T)cb[k], nv[k].x)) : key_of(pos_inf(T(0)));   // padding (norm = +inf): never a candidate
                }
            }
#pragma unroll
            for (int k = 0; k < NQI; ++k) { const int q = lane + 32 * k; if (q < WarpSmem<T>::NK) S.mncK[q] = mq[k]; }
        }
#else
#pragma unroll 4
        for (int x = lane; x < COL_WIN; x += 32) {
            const int p = col_slot(x);
            S.colA[p] = df_dg[c0 + x];
            const T2 nv = nm[c0 + x];                       // {norm_c, inv_c}
            double cbest = __ldcg(thr + c0 + x);
            if (LOG) cbest = lower_thr(cbest, km);
            S.colBest[p] = cbest;
            S.colInv[p] = nv.y;
            S.ck[p] = nv.x < pos_inf(T(0)) ? key_of(mul_rn((T)cbest, nv.x)) : key_of(pos_inf(T(0)));   // padding: never a candidate
        }
        for (int q = lane; q < WarpSmem<T>::NK; q += 32) S.mncK[q] = RB * q < COL_WIN ? mnc[c0 + RB * q] : pos_inf(T(0));
#endif
        __syncwarp();
        // minima of the column keys over the MNC_WIN columns of each block window: lane owns window q = columns
        // [RB*q, RB*q + MNC_WIN) of the staged window (what one lane covers in one block of RB rows); beyond it: skipped
        for (int q = lane; q < WarpSmem<T>::NK; q += 32) {
            int mn = key_of(pos_inf(T(0)));
            const int ph = (q % BPG) * RB, e = q / BPG;       // the window starts in class ph of entry e (ph = 0 when a group is one block)
#pragma unroll
            for (int k = 0; k < MNC_WIN; ++k) {
                const int cls = (ph + k) % DPT, ent = e + (ph + k) / DPT;
                const int v = S.ck[cls * COL_STRIDE + min(ent, COL_STRIDE - 1)];      // branch-free: a clamped load, discarded beyond the window
                mn = ent < COL_STRIDE ? min(mn, v) : mn;
            }
            set_key(S.wK[q], mn);
        }
        __syncwarp();

        // column registers: slot k holds window column x = i + DPT*lane + ((k - i) mod DPT)
        T cdf[DPT], cdg[DPT];
#pragma unroll
        for (int k = 0; k < DPT; ++k) {
            const T2 a = S.colA[k * COL_STRIDE + lane];
            cdf[k] = a.x; cdg[k] = a.y;
        }
        const T2 *rowA = S.rowA + 2;                       // [ii] = {df, dg} of row i + 2 (consumed one sub-step later)
        const T2 *colA = S.colA + lane + 1;
        const T *mncK = S.mncK + BPG * lane;               // [b] = window terms of block b: columns [RB*b + DPT*lane, + MNC_WIN)
        const T *wK = S.wK + BPG * lane;
        // operands of sub-step 0, loaded one sub-step ahead like all the others
        T2 ra = S.rowA[1];                                 // {df, dg} of row i + 1
        T2 nd = colA[0];                                   // column entering the lane's window: x = i + DPT*lane + DPT
#pragma unroll hot_unroll<T>()
        for (int I = 0; I < TILE_H / DPT; ++I) {           // DPT rows: one rotation of the column registers
#pragma unroll
            for (int h = 0; h < BPG; ++h) {                // RB rows: one filter test
                QVec<T> qc;                                // QT at the start of the block (replay checkpoint)
#pragma unroll
                for (int j = 0; j < DPT; ++j) qc.v[j] = qt[j];
                const T2 rk = S.rowK[I * BPG + h];
                const T cwn = mncK[I * BPG + h];
                const T cww = wK[I * BPG + h];
                int run = 0;
#pragma unroll
                for (int u = 0; u < RB; ++u) {
                    const int ii = h * RB + u;
                    const T2 ra_n = rowA[ii];              // operands of sub-step ii + 1
                    const T2 nd_n = colA[((ii + 1) % DPT) * COL_STRIDE + (ii + 1) / DPT];
                    run = u == 0 ? max_key8(qt) : max_key8(qt, run);      // keys of QT(row i)
                    cdf[ii] = nd.x; cdg[ii] = nd.y;
#pragma unroll
                    for (int j = 0; j < DPT; ++j) {
                        const int k = (ii + 1 + j) % DPT;
                        qt[j] = qt_step<STRICT>(qt[j], ra.x, ra.y, cdf[k], cdg[k]);     // QT(row i + 1)
                    }
                    ra = ra_n; nd = nd_n;
                }
                const int ka = key_of(mul_rn(rk.x, cwn));
                const int kb = key_of(mul_rn(rk.y, cww));
                const bool mine = run > __vimin_s32_relu(ka, kb);
#ifdef MPB_REPLAY_HIST
                { const unsigned bm = __ballot_sync(FULL, mine); if (bm && lane == 0 && counters) atomicAdd(counters + 4 + __popc(bm), 1ull); }
#endif
                if (__builtin_expect(__any_sync(FULL, mine), 0)) {
                    if constexpr (BPG == 1) {                  // queue the lane-blocks that passed (flush_queue replays them one per lane)
                        const unsigned bm = __ballot_sync(FULL, mine);
                        const int nq = __popc(bm);
                        if (hitcnt + nq > QCAP) { flush_queue<T, STRICT, LOG>(hitcnt, r0, c0, prof); hitcnt = 0; }
                        if (mine) {
                            const int pos = hitcnt + __popc(bm & ((1u << lane) - 1u));
#pragma unroll
                            for (int j = 0; j < DPT; ++j) S.hitQ[pos * DPT + j] = qc.v[j];
                            S.hitM[pos] = make_int2(I * DPT, I * DPT + DPT * lane);
                        }
                        if (lane == 0) S.stat[3] += nq;
                        hitcnt += nq;
                    } else {
                        hitcnt = h == 0 ? replay_block<T, STRICT, LOG, 0>(qc, I * DPT, r0, c0, hitcnt, prof)
                                        : replay_block<T, STRICT, LOG, (BPG > 1 ? RB : 0)>(qc, I * DPT, r0, c0, hitcnt, prof);
                    }
                }
            }
            rowA += DPT; colA += 1;
        }
        if (hitcnt > 0) {                                  // candidates never outlive their tile
            if constexpr (BPG == 1) flush_queue<T, STRICT, LOG>(hitcnt, r0, c0, prof);
            else drain_hits<T, LOG>(hitcnt, r0, c0, prof);
            hitcnt = 0;
        }
    }
    if (LOG) {                                             // E_k: the chain's value at the first row of the next segment
        const int item_e = blockIdx.x + item_begin;
        double2 *dst = reinterpret_cast<double2 *>(cf.seg_end + ((size_t)(item_e - cf.item_base) * BAND_W + (threadIdx.x >> 5) * WARP_DIAGS + (threadIdx.x & 31) * DPT));
#pragma unroll
        for (int j = 0; j < DPT; j += 2) dst[j / 2] = make_double2((double)qt[j], (double)qt[j + 1]);
    }
    __syncwarp();
    if (counters && lane == 0 && S.stat[2]) {
        atomicAdd(counters, (unsigned long long)S.stat[0]); atomicAdd(counters + 1, (unsigned long long)S.stat[1]);
        atomicAdd(counters + 2, (unsigned long long)S.stat[2]); atomicAdd(counters + 3, (unsigned long long)S.stat[3]);
    }
}

// ---------------------------------------------------------------------------------------------
// Sketch pass.  The filter is only as tight as the profile, and in the first bands of a step the profile is empty:
// nearly every cell of band 0 improves its row or its column, every block is replayed, every row parked and drained
// (measured: band 0 = 1/2048 of the cells but 6-7 ms of a 156 ms step, and the same 6-7 ms on EVERY rank of a
// multi-GPU job, where it is a quarter of the rank's time).  The sketch walks the first bands once WITHOUT the
// hit machinery: the same work items, the same chains (bit-identical QT), P = (QT * inv_r) * inv_c for every cell
// (the reference's per-cell form, kernel/MatrixProfileKernelVanilla.cpp:84), but only the HIGH WORD of P is kept --
// a running integer max per column (registers while the column is in the lane's window, then shared memory) and
// per row (one REDUX.MAX per row) -- and pushed into hint[] with a 32-bit atomic max per row / column and tile.
// hint[s] is then a lower bound of slot s' best correlation over the sketched bands; apply_hint_kernel stores it,
// lowered by two high-word ulps, as a "borrowed" entry {bound, index -1} wherever it beats the slot.  The tile
// kernel then walks the SAME bands with a tight filter: the cell behind each bound passes (P >= bound), is
// recorded exactly and replaces the borrowed entry (equal corr -> any real index beats -1), so the result is
// unchanged.  Cost: 4 FP64-pipe ops per sketched cell, no CAS, no replay.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(BAND_THREADS, ctas_per_sm<T>())
sketch_kernel(const typename Tr<T>::T2 *__restrict__ df_dg, const typename Tr<T>::T2 *__restrict__ nm, const T *__restrict__ qt0,
              const double *__restrict__ Tser, const double *__restrict__ mu, int *__restrict__ hint,
              const int *__restrict__ item_start, int item_begin, int nlb, int n, int m, int sublen, int excl, int seg_rows, int direct_band,
              int seg_first, int seg_stride) {
    using T2 = typename Tr<T>::T2;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WarpSmem<T> S(smem_raw + warp * WarpSmem<T>::BYTES);
    int *rowK = reinterpret_cast<int *>(S.wlist);               // [i] high word of the row's best P in this tile
    static_assert(WCAP * 16 >= ROWS_STAGED * 4, "row maxima alias the winner list");

    // direct_band >= 0: one band, cut into this launch's own (shorter) segments, block b = segment b -- the bounds only
    // have to be true lower bounds, so the sketch need not walk the tile kernel's chains (a re-seeded QT differs from
    // the running one by rounding, ~1e-13, and the bounds are lowered by 2e-6), and one band then fills the GPU.
    int lo = 0, hi = nlb;
    const int item = blockIdx.x + item_begin;
    if (direct_band < 0) while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (item_start[mid] <= item) lo = mid; else hi = mid; }
    const int seg = direct_band >= 0 ? seg_first + (int)blockIdx.x * seg_stride : item - item_start[lo];   // direct: this rank's share of the band's segments
    const int band = direct_band >= 0 ? direct_band : item_start[nlb + 1 + lo];
    const int d0w = excl + band * BAND_W + warp * WARP_DIAGS;
    const int nrows = sublen - d0w;
    const int rbeg = seg_rows > 0 ? seg * seg_rows : 0;
    const int rend = seg_rows > 0 ? min(rbeg + seg_rows, nrows) : nrows;
    if (rbeg >= rend) return;

    T qt[DPT];
    if (seg == 0) {
#pragma unroll
        for (int j = 0; j < DPT; ++j) qt[j] = qt0[d0w + lane * DPT + j];
    } else {
        reseed<T>(qt, S, Tser, mu, n, m, sublen, rbeg - 1, d0w, lane);
    }
    bool first_tile = true;
    for (int r0 = rbeg; r0 < rend; r0 += TILE_H) {
        const int c0 = r0 + d0w;
        __syncwarp();
        for (int i = lane; i < ROWS_STAGED; i += 32) {
            S.rowA[i] = df_dg[r0 + i];
            S.rowC[i] = make_double2(0.0, (double)nm[r0 + i].y);
        }
#pragma unroll 4
        for (int x = lane; x < COL_WIN; x += 32) {
            const int p = col_slot(x);
            S.colA[p] = df_dg[c0 + x];
            S.colInv[p] = nm[c0 + x].y;
            S.ck[p] = 0;
        }
        __syncwarp();
        T cdf[DPT], cdg[DPT], civ[DPT];
        int cmax[DPT];
#pragma unroll
        for (int k = 0; k < DPT; ++k) {
            const T2 a = S.colA[k * COL_STRIDE + lane];
            cdf[k] = a.x; cdg[k] = a.y;
            civ[k] = S.colInv[k * COL_STRIDE + lane];
            cmax[k] = 0;
        }
        if (first_tile) {
            const T2 ra = S.rowA[0];
#pragma unroll
            for (int j = 0; j < DPT; ++j) qt[j] = qt_step<false>(qt[j], ra.x, ra.y, cdf[j], cdg[j]);
            first_tile = false;
        }
#pragma unroll 1
        for (int I = 0; I < TILE_H / DPT; ++I) {
            const int e0 = I + lane;
#pragma unroll
            for (int ii = 0; ii < DPT; ++ii) {
                const int i = I * DPT + ii;
                const T ir = (T)S.rowC[i].y;
                const T2 ra = S.rowA[i + 1];
                int kp[DPT];
#pragma unroll
                for (int j = 0; j < DPT; ++j) {
                    const int k = (ii + j) % DPT;
                    kp[j] = __double2hiint((double)corr_of(qt[j], ir, civ[k]));
                    cmax[k] = max(cmax[k], kp[j]);
                }
                const int a = __vimax3_s32(kp[0], kp[1], kp[2]), b = __vimax3_s32(kp[3], kp[4], kp[5]);
                const int rm = __reduce_max_sync(FULL, __vimax3_s32(kp[6], kp[7], max(a, b)));
                if (lane == ii) rowK[i] = rm;
                // the column leaving the lane's window (tile column i + DPT*lane) hands its maximum to shared memory,
                // the one entering (i + DPT*lane + DPT) takes its register slot
                const int p = ii * COL_STRIDE + e0;
                S.ck[p] = max(S.ck[p], cmax[ii]);
                cmax[ii] = 0;
                const T2 nd = S.colA[p + 1];
                cdf[ii] = nd.x; cdg[ii] = nd.y;
                civ[ii] = S.colInv[p + 1];
#pragma unroll
                for (int j = 0; j < DPT; ++j) {
                    const int k = (ii + 1 + j) % DPT;
                    qt[j] = qt_step<false>(qt[j], ra.x, ra.y, cdf[k], cdg[k]);
                }
            }
            __syncwarp();      // the next block's lanes update the shared-memory maxima this block's neighbours wrote
        }
#pragma unroll
        for (int k = 0; k < DPT; ++k) {                        // columns still in the windows: TILE_H + DPT*lane + k
            const int p = k * COL_STRIDE + TILE_H / DPT + lane;
            S.ck[p] = max(S.ck[p], cmax[k]);
        }
        __syncwarp();
        for (int i = lane; i < TILE_H; i += 32) { const int v = rowK[i]; if (v > 0) atomicMax(hint + r0 + i, v); }
        for (int x = lane; x < COL_WIN; x += 32) { const int v = S.ck[col_slot(x)]; if (v > 0) atomicMax(hint + c0 + x, v); }
    }
}

// hint[i] (high word of a lower bound of slot i's best correlation) -> borrowed profile entry {bound, -1}
__global__ void apply_hint_kernel(const int *__restrict__ hint, unsigned long long *__restrict__ thr, int sublen) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sublen) return;
    const int h = hint[i] - 2;                                  // two high-word ulps (2e-6 relative) below any rounding concern
    if (h <= 0) return;
    atomicMax(thr + i, (unsigned long long)(unsigned)h << 32);
}

template <typename T>
static cudaError_t launch_sketch_t(const T *df_dg, const T *nm, const T *qt0, int *hint, const TileLaunch &p, bool apply, cudaStream_t s) {
    using T2 = typename Tr<T>::T2;
    if (p.items <= 0) return cudaSuccess;
    const size_t smem = (size_t)cta_smem_bytes<T>();
    cudaError_t e = cudaFuncSetAttribute(sketch_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(hint, 0, sizeof(int) * (size_t)padded_len(p.sublen), s);
    if (e != cudaSuccess) return e;
    long long grid = p.items;
    int seg_rows = p.seg_rows;
    const bool direct = p.sketch_band >= 0 && p.sketch_seg_rows > 0;
    const int first = direct ? p.sketch_share_rank : 0, stride = direct ? p.sketch_share_world : 1;
    if (direct) {
        seg_rows = p.sketch_seg_rows;
        const long long nseg = (p.sublen - (p.excl + p.sketch_band * BAND_W) + seg_rows - 1) / seg_rows;
        grid = nseg > first ? (nseg - first + stride - 1) / stride : 0;
    }
    if (grid > 0) {
        sketch_kernel<T><<<(unsigned)grid, BAND_THREADS, smem, s>>>(reinterpret_cast<const T2 *>(df_dg), reinterpret_cast<const T2 *>(nm), qt0, p.series, p.mu, hint,
                                                                p.item_start, p.item_begin, p.nlb, p.n, p.m, p.sublen, p.excl, seg_rows,
                                                                direct ? p.sketch_band : -1, first, stride);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    if (!apply) return cudaSuccess;
    return launch_apply_hint(hint, p.share.thr[0], p.sublen, s);
}
cudaError_t launch_apply_hint(const int *hint, unsigned long long *thr, int sublen, cudaStream_t s) {
    apply_hint_kernel<<<(sublen + 255) / 256, 256, 0, s>>>(hint, thr, sublen);
    return cudaGetLastError();
}
cudaError_t launch_sketch_f64(const double *df_dg, const double *nm, const double *qt0, int *hint, const TileLaunch &p, bool apply, cudaStream_t s) {
    return launch_sketch_t<double>(df_dg, nm, qt0, hint, p, apply, s);
}
cudaError_t launch_sketch_f32(const float *df_dg, const float *nm, const float *qt0, int *hint, const TileLaunch &p, bool apply, cudaStream_t s) {
    return launch_sketch_t<float>(df_dg, nm, qt0, hint, p, apply, s);
}

int band_count(int sublen, int excl) {
    const int ndiag = sublen - excl;
    return ndiag <= 0 ? 0 : (ndiag + BAND_W - 1) / BAND_W;
}

// Work-item table of `rank`, whose bands are b = rank, rank + world, ...: entries [0, nlb] are the prefix sums of
// the items per band (ceil(rows(b) / seg_rows), one when seg_rows == 0) and entries [nlb + 1, 2 nlb] the band ids
// in visiting order (increasing lag: measured better than a bit-reversed order, which delays the first visit of
// many rows and triples the filter hits at n = 2^18).  Returns the total item count.
// (Up to v7 every rank walked band 0 first, ranks > 0 redundantly, to warm its thresholds; the sketch pass does that
// better: 8-way emulation 23.4 -> 18.9 ms without the redundant band.  Dealing the first bands over the ranks by row
// segments, or cutting them into shorter segments, measured no gain: profiles/r01_v8_phase_probe_*.jsonl.)
long long build_item_table(int sublen, int excl, int rank, int world, int seg_rows, std::vector<int> &item_start) {
    const int nb = band_count(sublen, excl);
    const int nlb = nb > rank ? (nb - rank + world - 1) / world : 0;
    item_start.assign(2 * (size_t)nlb + 1, 0);
    long long total = 0;
    for (int lb = 0; lb < nlb; ++lb) {
        const int band = rank + lb * world;
        const int nrows = sublen - (excl + band * BAND_W);
        item_start[lb] = (int)total;
        item_start[(size_t)nlb + 1 + lb] = band;
        total += seg_rows > 0 ? (nrows + seg_rows - 1) / seg_rows : 1;
    }
    item_start[nlb] = (int)total;
    return total;
}

// Default segment length: long enough that the re-seed (m FMAs per diagonal per segment against
// ~2.25 * seg_rows pipe operations) stays around 3 %, short enough that a rank has several waves of items.
int default_seg_rows(int sublen, int excl, int m, int world, int sm_count, int base_rows) {
    int floor_rows = ((16 * m + 255) / 256) * 256;
    if (floor_rows < 1024) floor_rows = 1024;
    int seg = floor_rows < base_rows ? base_rows : floor_rows;
    std::vector<int> tab;
    const long long want = 6LL * 3 * sm_count;            // >= 4.5 waves of the 4 CTAs an SM holds
    while (seg / 2 >= floor_rows && build_item_table(sublen, excl, 0, world, seg, tab) < want) seg /= 2;
    return (seg + TILE_H - 1) / TILE_H * TILE_H;          // whole tiles of 192 rows
}

template <typename T>
static cudaError_t launch_tiles_t(const T *df_dg, const T *nm, const T *mnc, const T *qt0, ProfileEntry *prof, const TileLaunch &p, cudaStream_t s) {
    using T2 = typename Tr<T>::T2;
    if (p.items <= 0) return cudaSuccess;
    if (p.seg_rows % TILE_H != 0 || p.seg_rows < 0 || p.share.n < 1 || p.share.n > MAX_PEERS) return cudaErrorInvalidValue;
    const bool carry = p.chain.log != nullptr;
    if (carry && (sizeof(T) != 8 || p.strict || p.seg_rows == 0)) return cudaErrorInvalidValue;
    const size_t smem = (size_t)cta_smem_bytes<T>();
    auto kern = p.strict ? tile_kernel<T, true, false> : tile_kernel<T, false, false>;
    if (carry) kern = tile_kernel<T, false, sizeof(T) == 8>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    kern<<<(unsigned)p.items, BAND_THREADS, smem, s>>>(reinterpret_cast<const T2 *>(df_dg), reinterpret_cast<const T2 *>(nm), mnc, qt0, p.series, p.mu,
                                                       prof, p.item_start, p.item_begin, p.nlb, p.n, p.m, p.sublen, p.excl, p.rank, p.world, p.seg_rows, p.counters, p.chain, p.share);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Chain carry, step 2: D_k(d) for every (band, warp, diagonal) of the chunk, in place of S_k.
//   D_0 = 0,  D_k = D_{k-1} + (E_{k-1} - S_k)          (E - S first: the two are equal to ~1e-9 relative, so the
// difference is exact and the carry accumulates without losing what it is made of)
// One thread per diagonal of the chunk's bands; consecutive threads read consecutive doubles.
// stats[3] = max over (k, d) of |D_k(d)| * (largest inv among the rows of segment k) * (largest inv among the columns
// those rows meet on diagonal d), as double bits: an upper bound of the correction any P of the chunk can receive
// (imax[j] = largest inv over rows [j * seg_rows, (j + 1) * seg_rows), from inv_block_max_kernel).
// ---------------------------------------------------------------------------------------------
__global__ void inv_block_max_kernel(const double *__restrict__ inv, int sublen, int seg_rows, double *__restrict__ imax) {
    const int j = blockIdx.x;
    double v = 0.0;
    for (int i = j * seg_rows + threadIdx.x; i < min((j + 1) * seg_rows, sublen); i += blockDim.x) v = fmax(v, inv[i]);
    __shared__ double red[32];
    for (int sft = 16; sft; sft >>= 1) v = fmax(v, __shfl_xor_sync(FULL, v, sft));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) { for (int w = 1; w < (int)blockDim.x / 32; ++w) v = fmax(v, red[w]); imax[j] = v; }
}

__global__ void chain_scan_kernel(ChainFix cf, const int *__restrict__ item_start, int lb_begin, int lb_end, int nlb, int sublen, int excl, int seg_rows,
                                  const double *__restrict__ imax, unsigned long long *stats) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int lb = lb_begin + (int)(t / BAND_W), off = (int)(t % BAND_W);
    double bound = 0.0;
    if (lb < lb_end) {
        const int band = item_start[nlb + 1 + lb];
        const int d0w = excl + band * BAND_W + (off / WARP_DIAGS) * WARP_DIAGS;
        const int d = excl + band * BAND_W + off;
        const int nrows = sublen - d0w;                          // rows of the warp's first diagonal: its segment count
        const int nseg = nrows > 0 ? (nrows + seg_rows - 1) / seg_rows : 0;
        const int nblk = (sublen + seg_rows - 1) / seg_rows;
        const size_t base = (size_t)(item_start[lb] - cf.item_base) * BAND_W + off;
        double carry = 0.0;
        if (nseg > 0) cf.seg_seed[base] = 0.0;
        for (int k = 1; k < nseg; ++k) {
            const double ek = cf.seg_end[base + (size_t)(k - 1) * BAND_W];
            const double sk = cf.seg_seed[base + (size_t)k * BAND_W];
            carry = __dadd_rn(carry, __dsub_rn(ek, sk));
            cf.seg_seed[base + (size_t)k * BAND_W] = carry;
            if ((long long)k * seg_rows + d < sublen) {          // the diagonal has cells in this segment
                const int jc = (k * seg_rows + d) / seg_rows;
                const double ic = fmax(imax[min(jc, nblk - 1)], imax[min(jc + 1, nblk - 1)]);
                bound = fmax(bound, fabs(carry) * imax[k] * ic);
            }
        }
    }
    const unsigned hi = __reduce_max_sync(FULL, (unsigned)(__double_as_longlong(bound) >> 32) + (bound > 0.0 ? 1u : 0u));   // high word, rounded up
    if ((threadIdx.x & 31) == 0 && hi) atomicMax(stats + 3, (unsigned long long)hi << 32);
}

// Chain carry, step 3: the carry of cell (r, c): P = P' + D_k(d) * inv_r * inv_c, the value the reference's unbroken chain
// gives the cell (kernel/MatrixProfileKernelVanilla.cpp:81,84) to rounding.  Returns false when the cell's band is not one
// of this launch chunk's (another rank's entry after a threshold merge, or an earlier chunk's: already resolved).
struct ChainGeom { const int *item_start; int nlb, lb_begin, lb_end, excl, rank, world, seg_rows; };
__device__ __forceinline__ bool chain_value(const ChainFix &cf, const ChainGeom &g, const double *__restrict__ inv, double corr, int slot, int index,
                                            double *out, double *rel) {
    if (index < 0) return false;
    const int r = min(slot, index), c = max(slot, index);
    const int gd = c - r - g.excl;
    if (gd < 0) return false;
    const int band = gd / BAND_W;
    if (band % g.world != g.rank) return false;
    const int lb = band / g.world;
    if (lb < g.lb_begin || lb >= g.lb_end) return false;
    const int k = r / g.seg_rows;
    const double d = cf.seg_seed[(size_t)(g.item_start[lb] + k - cf.item_base) * BAND_W + gd % BAND_W];
    const double corr_fix = d * inv[r] * inv[c];
    *out = corr + corr_fix;
    *rel = fabs(corr_fix) / corr;
    return true;
}

// (a) every slot's surviving threshold entry, (b) every logged near tie -> the result profile, exact (corr, index) max.
// stats[0] = largest relative correction applied (double bits), stats[1] += values resolved, stats[2] |= log overflow
__global__ void chain_resolve_kernel(ChainFix cf, ChainGeom g, const double *__restrict__ inv, const ProfileEntry *__restrict__ prof, ProfileEntry *fin,
                                     int sublen, unsigned long long *stats) {
    double worst = 0.0;
    unsigned mine = 0;
    const long long stride = (long long)gridDim.x * blockDim.x, t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (long long i = t0; i < sublen; i += stride) {
        const ProfileEntry e = prof[i];
        double pv, rel;
        if (e.corr > 0.0 && chain_value(cf, g, inv, e.corr, (int)i, e.index, &pv, &rel) && pv > 0.0) {
            profile_update(fin + i, pv, e.index); worst = fmax(worst, rel); ++mine;
        }
    }
    for (int shard = 0; shard < LOG_SHARDS; ++shard) {
        unsigned long long cnt = cf.log_count[shard * 16];
        if (cnt > cf.shard_cap) {                                // near ties were dropped: the host turns this into an error
            if (t0 == 0) atomicOr(stats + 2, 1ull);
            cnt = cf.shard_cap;
        }
        const ChainLogEntry *log = cf.log + (size_t)shard * cf.shard_cap;
        for (long long e = t0; e < (long long)cnt; e += stride) {
            const ChainLogEntry le = log[e];
            double pv, rel;
            if (chain_value(cf, g, inv, le.corr, le.slot, le.index, &pv, &rel) && pv > 0.0) {
                profile_update(fin + le.slot, pv, le.index); worst = fmax(worst, rel); ++mine;
            }
        }
    }
    const unsigned hi = __reduce_max_sync(FULL, (unsigned)(__double_as_longlong(worst) >> 32));   // the high word is enough for a statistic
    mine = __reduce_add_sync(FULL, mine);
    if ((threadIdx.x & 31) == 0) {
        if (hi) atomicMax(stats, (unsigned long long)hi << 32);
        if (mine) atomicAdd(stats + 1, (unsigned long long)mine);
    }
}

cudaError_t launch_inv_block_max(const double *inv, int sublen, int seg_rows, double *imax, cudaStream_t s) {
    if (seg_rows <= 0) return cudaSuccess;
    inv_block_max_kernel<<<(sublen + seg_rows - 1) / seg_rows, 256, 0, s>>>(inv, sublen, seg_rows, imax);
    return cudaGetLastError();
}

cudaError_t launch_chain_resolve(const TileLaunch &p, int lb_begin, int lb_end, const double *inv, const double *imax, const ProfileEntry *prof, ProfileEntry *fin,
                                 unsigned long long *stats, cudaStream_t s) {
    if (p.chain.log == nullptr || lb_end <= lb_begin) return cudaSuccess;
    const long long threads = (long long)(lb_end - lb_begin) * BAND_W;
    chain_scan_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(p.chain, p.item_start, lb_begin, lb_end, p.nlb, p.sublen, p.excl, p.seg_rows, imax, stats);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    ChainGeom g; g.item_start = p.item_start; g.nlb = p.nlb; g.lb_begin = lb_begin; g.lb_end = lb_end; g.excl = p.excl; g.rank = p.rank; g.world = p.world; g.seg_rows = p.seg_rows;
    chain_resolve_kernel<<<1184, 256, 0, s>>>(p.chain, g, inv, prof, fin, p.sublen, stats);
    return cudaGetLastError();
}

cudaError_t launch_tiles_f64(const double *df_dg, const double *nm, const double *mnc, const double *qt0, ProfileEntry *prof, const TileLaunch &p, cudaStream_t s) {
    return launch_tiles_t<double>(df_dg, nm, mnc, qt0, prof, p, s);
}
cudaError_t launch_tiles_f32(const float *df_dg, const float *nm, const float *mnc, const float *qt0, ProfileEntry *prof, const TileLaunch &p, cudaStream_t s) {
    return launch_tiles_t<float>(df_dg, nm, mnc, qt0, prof, p, s);
}

}  // namespace mpb
