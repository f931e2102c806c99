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
constexpr int RB = 8;                     // rows per branch-free block of the hot loop (one filter test per block and lane)
constexpr int MNC_WIN = FWIN + RB - 1;    // columns a lane covers in one block = width of the staged column-window minima
constexpr int BAND_WARPS = MPB_WARPS;     // warps per CTA
constexpr int BAND_THREADS = 32 * BAND_WARPS;
constexpr int WARP_DIAGS = DPT * 32;      // diagonals per warp
constexpr int BAND_W = WARP_DIAGS * BAND_WARPS;   // diagonals per band (= per CTA)
constexpr int TILE_H = MPB_TILE_H;        // rows staged per warp tile
constexpr int COL_WIN = TILE_H + WARP_DIAGS + DPT;   // columns (and filter-window starts) a warp tile can touch
constexpr int COL_STRIDE = ((COL_WIN + DPT - 1) / DPT) | 1;  // entries per residue class, odd => conflict-free staging stores
static_assert(DPT == 8 || DPT == 16, "DPT");
static_assert(TILE_H % DPT == 0 && TILE_H % 32 == 0, "TILE_H");
// Arrays indexed by subsequence are padded so that tiles may run past n-m+1 without bounds checks.
constexpr int PAD = BAND_W + 4 * TILE_H + 64;

__host__ __device__ inline int padded_len(int sublen) { return ((sublen + PAD + 63) / 64) * 64; }

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
