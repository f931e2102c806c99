// mp_kernels.h -- host-side launchers of the engine's kernels (internal to libmpb200.so).
#pragma once
#include "mp_common.cuh"
#include <cuda_runtime.h>
#include <stdint.h>
#include <vector>

namespace mpb {

// kernel (1): precompute.cu
// the reference's sequential running mean, bit for bit: mode 0 = one thread walks the chain; mode 1 = segmented-scan formulation with
// every step verified (precompute.cu), the walk only as its fallback.  scratch: mu_scan_scratch_bytes(n, m) device bytes (mode 1)
cudaError_t launch_mu_scan(const double *T, int n, int m, double *mu, void *scratch, int mode, cudaStream_t s);
size_t mu_scan_scratch_bytes(int n, int m);
size_t mu_scan_bad_offset(int n, int m);
cudaError_t launch_mu_window(const double *T, int n, int m, double *mu, cudaStream_t s);    // exact window means (parallel)
cudaError_t launch_stats_f64(const double *T, const double *mu, int n, int m, double *df_dg, double *inv, double *qt0, cudaStream_t s);
cudaError_t launch_stats_f32(const double *T, const double *mu, int n, int m, float *df_dg, float *inv, float *qt0, cudaStream_t s);
// nm[i] = {norm_i = 1 / inv_i, inv_i}, mnc[i] = min of norm over [i, i + MNC_WIN), for i < np; norm = +inf on the padding (inv = 0)
cudaError_t launch_norm_f64(const double *inv, int sublen, int np, double *nm, double *mnc, cudaStream_t s);
cudaError_t launch_norm_f32(const float *inv, int sublen, int np, float *nm, float *mnc, cudaStream_t s);
cudaError_t launch_profile_init(ProfileEntry *prof, int count, cudaStream_t s);
// seeds every THRESHOLD with the cell at lag `lag` (the first diagonal outside the exclusion zone), computed directly and lowered by
// max(rel * P, mu * max(P, CHAIN_FLOOR)): a borrowed lower bound -- the tile kernel meets the cell itself on its way
cudaError_t launch_seed_thresholds_f64(const double *T, const double *mu, const double *inv, int n, int m, int lag, unsigned long long *thr, double rel, double mu_abs, cudaStream_t s);
cudaError_t launch_seed_thresholds_f32(const double *T, const double *mu, const float *inv, int n, int m, int lag, unsigned long long *thr, double rel, double mu_abs, cudaStream_t s);
cudaError_t launch_pack_input_f64(const double *df_dg, const double *inv, const double *qt0, int sublen, void *out, cudaStream_t s);
cudaError_t launch_pack_input_f32(const float *df_dg, const float *inv, const float *qt0, int sublen, void *out, cudaStream_t s);

// kernel (2): scamp_tile.cu -- work items = (diagonal band of BAND_W, row segment of seg_rows); bands b = rank, rank+world, ...
struct TileLaunch {
    int n, m, sublen, excl, rank, world;
    int seg_rows;                       // rows per re-seeded segment (multiple of TILE_H); 0 = one full-length chain per diagonal
    bool strict;                        // un-fused QT update (the reference's rounding sequence)
    const double *series, *mu;          // T[n], mu[sublen] (re-seed operands)
    const int *item_start;              // device: prefix table (nlb + 1 entries) followed by the band ids in visiting order (nlb entries)
    int nlb;                            // this rank's band count
    long long items;                    // work items of this launch (= grid size)
    int item_begin;                     // first work item of this launch (index into the prefix table)
    unsigned long long *counters;       // device: [0] cells parked, [1] profile slots improved, [2] blocks replayed (may be null)
    int sketch_band = -1, sketch_seg_rows = 0;   // sketch launches only: walk this one band in segments of sketch_seg_rows rows (own cut; -1 / 0 = the items of the table)
    int sketch_share_rank = 0, sketch_share_world = 1;   // one-band sketch shared by a multi-GPU job: this launch walks segments rank, rank + world, ...
    ChainFix chain = {};                // chain-carry mode (fp64, re-seeded segments): log != nullptr switches it on
    ThrShare share = {};                // threshold arrays of the job: [0] this rank's (np entries), then the peers'; n >= 1
};
int band_count(int sublen, int excl);                    // total number of bands of the problem
long long build_item_table(int sublen, int excl, int rank, int world, int seg_rows, std::vector<int> &item_start);
// base_rows: the segment length aimed at when the problem has items to spare (shorter for small problems, never below 16 m)
int default_seg_rows(int sublen, int excl, int m, int world, int sm_count, int base_rows);
cudaError_t launch_tiles_f64(const double *df_dg, const double *nm, const double *mnc, const double *qt0, ProfileEntry *prof, const TileLaunch &p, cudaStream_t s);
cudaError_t launch_tiles_f32(const float *df_dg, const float *nm, const float *mnc, const float *qt0, ProfileEntry *prof, const TileLaunch &p, cudaStream_t s);

// chain-carry steps 2 and 3 for the local bands [lb_begin, lb_end) of the launch chunk p.chain was laid out for: carry scan, then the
// surviving threshold entries of `prof` and the logged near ties into `fin`.  imax = launch_inv_block_max's output.
// stats: [0] largest relative correction applied (double bits) [1] += values resolved [2] |= log overflow [3] bound of any correction
cudaError_t launch_inv_block_max(const double *inv, int sublen, int seg_rows, double *imax, cudaStream_t s);
cudaError_t launch_chain_resolve(const TileLaunch &p, int lb_begin, int lb_end, const double *inv, const double *imax, const ProfileEntry *prof, ProfileEntry *fin,
                                 unsigned long long *stats, cudaStream_t s);

// sketch pass over the items of `p` (first bands of a step): hint[np] <- high words of per-slot lower bounds, then into the thresholds p.share.thr[0]
// apply = false: leave hint[] for a max-reduction across the ranks that share the sketch, then launch_apply_hint
cudaError_t launch_sketch_f64(const double *df_dg, const double *nm, const double *qt0, int *hint, const TileLaunch &p, bool apply, cudaStream_t s);
cudaError_t launch_sketch_f32(const float *df_dg, const float *nm, const float *qt0, int *hint, const TileLaunch &p, bool apply, cudaStream_t s);
cudaError_t launch_apply_hint(const int *hint, unsigned long long *thr, int sublen, cudaStream_t s);

// kernel (3): merge.cu
cudaError_t launch_finalize(const ProfileEntry *prof, int sublen, int m, double *MP, int32_t *MPI, double *corr, cudaStream_t s);
cudaError_t launch_export(const ProfileEntry *prof, int sublen, double *corr, int32_t *index, cudaStream_t s);
cudaError_t launch_mask_index(const ProfileEntry *prof, int sublen, const double *corr_max, int32_t *index, cudaStream_t s);
// import / merge_peer also raise the thresholds thr[] to the merged correlations
cudaError_t launch_import(ProfileEntry *prof, int sublen, const double *corr, const int32_t *index, unsigned long long *thr, cudaStream_t s);
cudaError_t launch_merge_peer(ProfileEntry *prof, const ProfileEntry *peer, int sublen, unsigned long long *thr, cudaStream_t s);

// chunk-compatible MatrixProfileKernelTLF: chunk_tlf.cu
cudaError_t launch_chunk_tlf_f64(int n, int m, int w, int iteration, const void *input, void *output, void *scratch, size_t scratch_bytes, cudaStream_t s);
cudaError_t launch_chunk_tlf_f32(int n, int m, int w, int iteration, const void *input, void *output, void *scratch, size_t scratch_bytes, cudaStream_t s);
size_t chunk_tlf_scratch_bytes(int n, int m, int w);

// measurement: peak_fma.cu
cudaError_t launch_fma_peak(int dtype, int iters, int blocks, double *sink, cudaStream_t s);
}  // namespace mpb
