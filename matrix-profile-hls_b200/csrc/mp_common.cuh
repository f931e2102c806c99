// mp_common.cuh -- shared device types and helpers of the B200 matrix-profile engine.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mpb {

// One profile slot: best Pearson correlation seen so far for subsequence i (as a row OR as a column,
// i.e. the reference's rowAggregates[i] and columnAggregates[i] already merged,
// include/host/HostSideComputation.hpp:61) and the neighbour that achieved it.  16 bytes so that a
// single 128-bit compare-and-swap updates value and index together, exactly.
struct __align__(16) ProfileEntry {
    double corr;
    int32_t index;
    int32_t pad;
};

// Byte-identical to include/Config.hpp.in:43-73.
template <typename T> struct Aggregate { T value; int32_t index; };
template <typename T> struct InputDataPack { T QT, df, dg, inv; };
template <typename T> struct OutputDataPack { Aggregate<T> rowAggregate, columnAggregate; };
static_assert(sizeof(Aggregate<double>) == 16 && sizeof(InputDataPack<double>) == 32 && sizeof(OutputDataPack<double>) == 32, "layout");
static_assert(sizeof(Aggregate<float>) == 8 && sizeof(InputDataPack<float>) == 16 && sizeof(OutputDataPack<float>) == 16, "layout");

// Geometry of the diagonal-tile kernel (scamp_tile.cu). A work item is BAND_W consecutive diagonals x one row
// segment, walked by one CTA; each of its warps owns WARP_DIAGS diagonals (DPT per lane) and stages its OWN
// tile of TILE_H rows + the columns they touch, so the warps of a CTA never wait for one another.
// The three knobs can be overridden at build time (make EXTRA="-DMPB_DPT=16 ...") for tuning runs.
#ifndef MPB_DPT
#define MPB_DPT 8
#endif
#ifndef MPB_TILE_H
#define MPB_TILE_H 192
#endif
#ifndef MPB_WARPS
#define MPB_WARPS 2
#endif
constexpr int DPT = MPB_DPT;              // diagonals per thread (register micro-tile width): 8 or 16
constexpr int FWIN = 8;                   // columns a lane covers in one row
constexpr int NWIN = DPT / FWIN;
#ifndef MPB_RB
#define MPB_RB 8
#endif
constexpr int RB = MPB_RB;                // rows per branch-free block of the hot loop (one filter test per block and lane): 8 or 4
constexpr int MNC_WIN = FWIN + RB - 1;    // columns a lane covers in one block = width of the staged column-window minima
constexpr int BAND_WARPS = MPB_WARPS;     // warps per CTA
constexpr int BAND_THREADS = 32 * BAND_WARPS;
constexpr int WARP_DIAGS = DPT * 32;      // diagonals per warp
constexpr int BAND_W = WARP_DIAGS * BAND_WARPS;   // diagonals per band (= per CTA)
constexpr int TILE_H = MPB_TILE_H;        // rows staged per warp tile
constexpr int COL_WIN = TILE_H + WARP_DIAGS + DPT;   // columns (and filter-window starts) a warp tile can touch
constexpr int COL_STRIDE = ((COL_WIN + DPT - 1) / DPT) | 1;  // entries per residue class, odd => conflict-free staging stores
static_assert(DPT == 8 || DPT == 16, "DPT");
static_assert(RB == 8 || RB == 4, "RB");
static_assert(TILE_H % DPT == 0 && TILE_H % 32 == 0, "TILE_H");
// Arrays indexed by subsequence are padded so that tiles may run past n-m+1 without bounds checks.
constexpr int PAD = BAND_W + 4 * TILE_H + 64;

__host__ __device__ inline int padded_len(int sublen) { return ((sublen + PAD + 63) / 64) * 64; }

// ---------------------------------------------------------------------------------------------
// Telescoped chain carry (fp64 default mode).  The reference seeds every diagonal ONCE at row 0 and walks it to the
// end (kernel/MatrixProfileKernelVanilla.cpp:43-55,81); its QT at row r is QT_0 + the sum of all updates up to r,
// drift of the running mean included.  The tile kernel cuts each band into independent row segments that start from
// an exact dot product S_k (so that thousands of work items run at once).  Because the update is ADDITIVE, the value
// the reference's chain holds anywhere inside segment k is
//     C(r, d) = Q'_k(r, d) + D_k(d),     D_0 = 0,   D_{k+1}(d) = D_k(d) + E_k(d) - S_{k+1}(d),
// where Q'_k is the segment's own chain, E_k its value at the first row of segment k+1 and S_{k+1} the value segment
// k+1 started from at that same row: the carry of the unbroken chain, recovered by telescoping AFTER the fact.
// Every work item stores S_k and E_k for its diagonals and a scan over k gives D_k.  While the tile kernel runs, the
// profile holds the segments' own values P' = Q' * inv_r * inv_c; they differ from the chain's P by a relative
// delta = D / Q' (1e-8 typically, 1e-6 where a window of the series is nearly flat).  Decisions that a correction of
// that size could reverse are deferred: a candidate that beats a slot's best by more than the MARGIN replaces it for
// good, one that loses by more than the margin is dropped for good, and a candidate within the margin of the best goes
// (or sends the entry it replaces) to a log of near ties.  After the scan the resolve kernel gives every surviving slot
// entry and every logged near tie its carry, P = P' + D * inv_r * inv_c, and takes the exact (corr, index) max.
// The filter thresholds are lowered by the margin, so every cell within the margin of a slot's best gets that far.
// ---------------------------------------------------------------------------------------------
struct __align__(16) ChainLogEntry { double corr; int32_t slot; int32_t index; };   // P' of cell (slot, index) for slot `slot`
constexpr int LOG_SHARDS = 64;                       // sub-logs, one append counter each (no single hot address)
struct ChainFix {
    double *seg_seed;                // [item - item_base][BAND_W]  S_k (k >= 1), overwritten by D_k in the scan
    double *seg_end;                 // [item - item_base][BAND_W]  E_k
    ChainLogEntry *log;              // LOG_SHARDS x shard_cap entries
    unsigned long long *log_count;   // [LOG_SHARDS * 16] (one counter per 128-byte line); may exceed shard_cap: overflow
    unsigned int shard_cap;
    int item_base;                   // first work item of the launch chunk the arrays are laid out for
    double margin;                   // near-tie margin mu: two values are a near tie when they differ by <= mu * max(value, CHAIN_FLOOR)
    int key_margin;                  // mu in ulps of a double's high word (one ulp is >= 2^-21 relative), rounded up
};
// The margin is relative for correlations above CHAIN_FLOOR and absolute (mu * CHAIN_FLOOR) below it, so that ONE bound
// on the absolute correction any cell can receive (chain_scan_kernel, mpb_timings.chain_bound) proves it sufficient:
// 2 * bound <= mu * CHAIN_FLOOR  ==>  no pair of candidates further apart than the margin can change order.
#define MPB_CHAIN_FLOOR 0.015625
__host__ __device__ inline double chain_lower(double v, double mu) { return v - mu * (v > MPB_CHAIN_FLOOR ? v : MPB_CHAIN_FLOOR); }   // v minus the margin


// Filter thresholds.  thr[i] = a lower bound of the best correlation of subsequence i known to the JOB, as the bit pattern of a
// non-negative double (orders like an integer, so one 64-bit atomic max raises it): the seed diagonal and the sketch pass put
// borrowed bounds there, and every improvement the tile kernel's drain applies to its rank's profile is pushed to the thr[]
// array of EVERY rank of the job -- its own in HBM, the peers' over NVLink (P2P / CUDA-IPC mapped memory, fire-and-forget
// red.max.u64) -- so a slot's threshold is discovered once per job, not once per GPU.  The tile kernel's staging reads only
// its own copy.  Thresholds are hints: stale or lost ones cost time, never correctness.
constexpr int MAX_PEERS = 8;
struct ThrShare {
    unsigned long long *thr[MAX_PEERS];   // [0] this rank's array (np entries), [1 .. n) the peers'
    int n;
    int pad;
};

// Exact (corr, index) max into a global profile slot: larger corr wins, equal corr -> smaller index
// (deterministic whatever the arrival order / GPU count).  Returns the slot's corr after the update
// (the band kernel derives its filter threshold from it).
__device__ __forceinline__ double profile_update(ProfileEntry *slot, double v, int idx) {
    unsigned long long clo, chi, olo, ohi;
    asm volatile("ld.global.relaxed.gpu.v2.u64 {%0,%1}, [%2];" : "=l"(clo), "=l"(chi) : "l"(slot) : "memory");
    const unsigned long long nlo = (unsigned long long)__double_as_longlong(v);
    const unsigned long long nhi = (unsigned long long)(unsigned)idx;
    while (true) {
        const double cv = __longlong_as_double((long long)clo);
        const unsigned ci = (unsigned)chi;   // index -1 (unset) compares as the largest unsigned
        if (!(v > cv || (v == cv && (unsigned)idx < ci))) break;
        asm volatile(
            "{\n .reg .b128 c, d, o;\n mov.b128 c, {%2, %3};\n mov.b128 d, {%4, %5};\n"
            " atom.global.relaxed.gpu.cas.b128 o, [%6], c, d;\n mov.b128 {%0, %1}, o;\n}"
            : "=l"(olo), "=l"(ohi) : "l"(clo), "l"(chi), "l"(nlo), "l"(nhi), "l"(slot) : "memory");
        if (olo == clo && ohi == chi) { clo = nlo; break; }
        clo = olo; chi = ohi;
    }
    return __longlong_as_double((long long)clo);
}


}  // namespace mpb
