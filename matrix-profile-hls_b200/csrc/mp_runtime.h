// mp_runtime.h -- internal: handle layouts and error plumbing shared by runtime.cu and multi_gpu.cu (not part of the C ABI).
#pragma once
#include "../../include/mp_b200.h"
#include "mp_common.cuh"
#include "mp_kernels.h"

#include <cstdarg>
#include <cstdio>
#include <vector>

int mpb_fail(int code, const char *fmt, ...);      // records the message for mpb_last_error() (thread-local) and returns code
#define fail mpb_fail
#define CU(expr)                                                                                     \
    do {                                                                                             \
        cudaError_t e__ = (expr);                                                                    \
        if (e__ != cudaSuccess)                                                                      \
            return fail(e__ == cudaErrorMemoryAllocation ? MPB_E_NOMEM : MPB_E_CUDA, "%s: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

struct mpb_context {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t mark = nullptr, now = nullptr;   // Finish() reports device time since the last Finish()
    void *tlf_scratch = nullptr;
    size_t tlf_scratch_bytes = 0;
    int sm_count = 0;
};

struct mpb_buffer {
    mpb_context *ctx = nullptr;
    void *ptr = nullptr;
    size_t bytes = 0;
};

struct mpb_plan {
    mpb_context *ctx = nullptr;
    int n = 0, m = 0, sublen = 0, excl = 0, np = 0, dtype = MPB_F64;
    unsigned flags = 0;
    int rank = 0, world = 1;
    double *T = nullptr, *mu = nullptr;
    void *df_dg = nullptr, *inv = nullptr, *qt0 = nullptr;    // data_t arrays, zero-padded to np
    void *nm = nullptr, *mnc = nullptr;                       // data_t {norm, inv} pairs and window-min norms, np entries
    mpb::ProfileEntry *prof = nullptr;                        // np slots: running bests = the tile kernel's filter thresholds
    mpb::ProfileEntry *fin = nullptr;                         // np slots: the result profile of the chain-carry mode (else unused; the result is prof)
    bool seq_mean = false, chain_fix = false;                 // mu by the reference's running recurrence; telescoped chain carry (mp_common.cuh)
    unsigned long long *thr = nullptr;                        // np filter thresholds (mp_common.cuh ThrShare): own allocation, so that peers can map it
    mpb::ThrShare share = {};                                 // [0] = thr, then the connected peers' arrays
    void *ipc_opened[mpb::MAX_PEERS] = {};                    // peers' arrays opened through CUDA IPC (closed with the plan)
    double *seg_seed = nullptr, *seg_end = nullptr;           // chain carry: S_k / E_k of the work items of one launch chunk, [chunk_items][BAND_W]
    long long chunk_items = 0;                                // capacity of those arrays in work items
    mpb::ChainLogEntry *chain_log = nullptr;                  // LOG_SHARDS sub-logs of log_shard_cap candidates
    unsigned long long *log_count = nullptr;                  // LOG_SHARDS counters, one per 128-byte line
    unsigned long long *chain_stats = nullptr;                // [0] largest relative correction (double bits) [1] candidates resolved [2] log overflow seen
    unsigned log_shard_cap = 0;
    long long log_entries_req = -1;                           // requested total log capacity (-1 = automatic)
    double *inv_blockmax = nullptr;                           // largest inv per block of seg_rows rows (bound of the chain-carry corrections)
    int inv_blocks = 0;
    int margin_log2 = 17;                                     // near-tie margin = 2^-margin_log2
    double *MP = nullptr, *corr = nullptr;
    int32_t *MPI = nullptr;
    unsigned long long *slow_counter = nullptr;
    void *mu_scratch = nullptr;           // work space of the parallel running-mean kernels (seq_mean only)
    int *hint = nullptr;                  // np high words: lower bounds of the slots' best correlation from the sketch pass
    int sketch_bands = -1;                // local bands the sketch pass covers (-1 = default, 0 = off)
    int seg_rows = -1;                    // requested (-1 = auto)
    int seg_rows_eff = 0;                 // in effect for the current table
    int *item_start = nullptr;            // device prefix table of work items
    int nlb = 0;
    long long items = 0;
    bool table_dirty = true;
    std::vector<int> table_host;          // host copy of the work-item table (prefix sums, then band ids)
    cudaEvent_t ev[8] = {};
    mpb_timings tm = {};
    bool ran_h2d = false;
    bool step_open = false;               // the compute timer / statistics of the current step have been started
    bool sketched = false;                // a sketch pass has run since the last precompute (mpb_plan_compute_bands then skips its own)
};

