/*
 * mp_b200.h -- C ABI of the B200-native matrix-profile engine (libmpb200.so).
 *
 * This is the drop-in boundary for the hot path of jlscheerer/matrix-profile-hls: it replaces the
 * OpenCL/XRT wrapper include/host/OpenCL.hpp and the FPGA kernel behind it.  Plain pointers and
 * sizes only; no exceptions cross it; every function returns MPB_OK (0) or a negative MPB_E_* code
 * and mpb_last_error() describes the failure.  Citations are file:line relative to the reference.
 *
 * There is NO CPU fallback: every compute entry point fails with MPB_E_NO_DEVICE when no CUDA
 * device is usable.
 *
 * Record layouts are byte-identical to include/Config.hpp.in:43-73:
 *   aggregate_t    { data_t value; int32 index; }            16 B (double, 4 B padding) / 8 B (float)
 *   InputDataPack  { data_t QT, df, dg, inv; }               32 B / 16 B
 *   OutputDataPack { aggregate_t rowAggregate, columnAggregate; }   32 B / 16 B
 */
#ifndef MP_B200_H
#define MP_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define MPB_OK 0
#define MPB_E_INVALID (-1)    /* bad argument (the reference's ConfigurationError, OpenCL.hpp:48-52) */
#define MPB_E_NO_DEVICE (-2)  /* no usable CUDA device: there is no CPU fallback */
#define MPB_E_CUDA (-3)       /* CUDA runtime failure (the reference's RuntimeError, OpenCL.hpp:54-60) */
#define MPB_E_NOMEM (-4)
#define MPB_E_MARGIN (-5)     /* chain-carry mode: the corrections found exceed what the near-tie margin covers; the result was not
                                 proven -- run again with a wider margin (mpb_plan_set_chain_margin; mpb_compute does so itself) */

/* MP_DATA_TYPE (CMakeLists.txt:35-37) */
#define MPB_F64 0
#define MPB_F32 1

/* mpb_plan flags.  Default (0):
 *   fp64 -- chain-faithful AND parallel: mu by the reference's running-mean recurrence (HostSideComputation.hpp:23,29),
 *           nested-FMA QT update, the matrix cut into independent (band x row segment) work items that start from
 *           exact dot products, and the reference's unbroken single-seed chain (Vanilla.cpp:43-55,81) recovered
 *           from them by telescoping ("chain carry", csrc/mp_common.cuh): ~1e-12 from the reference's bits.
 *   fp32 -- exact window means, nested FMA, exact fp64 re-seed of every diagonal each `segment rows` rows, which is
 *           what bounds the float chain's drift (BASELINE.json north_star). */
#define MPB_FLAG_STRICT 1u     /* chain-faithful: implies SEQ_MEAN | NO_RESEED and the un-fused QT update (mul,mul,add,add):
                                  bit-identical to the reference kernels' chain, 6 instead of 4 FP64 ops per cell */
#define MPB_FLAG_SEQ_MEAN 2u   /* mu by the reference's sequential running-mean recurrence (HostSideComputation.hpp:29) */
#define MPB_FLAG_NO_RESEED 4u  /* one full-length chain per diagonal, seeded at row 0 only (Vanilla.cpp:43-55) */
#define MPB_FLAG_NO_SEED 8u    /* do not pre-load the profile with the lag-m/4 diagonal (measurement aid; results are the same) */
#define MPB_FLAG_NO_CHAIN_CARRY 16u /* fp64: keep the re-seeded segments' own values (round 1's fast mode: drifts from the
                                  reference by up to 6e-9 on a 2^20-sample random walk) */
#define MPB_FLAG_WINDOW_MEAN 32u /* fp64: exact window means instead of the reference's running mean (no sequential scan) */
#define MPB_FLAG_MU_WALK 64u   /* running mean by one thread walking the chain (measurement aid: the default evaluates the same
                                  recurrence as a verified segmented scan, csrc/precompute.cu, and gives the same bits) */

typedef struct mpb_context mpb_context; /* replaces OpenCL::Context   (OpenCL.hpp:73-107, 294-317) */
typedef struct mpb_buffer mpb_buffer;   /* replaces OpenCL::Buffer<T> (OpenCL.hpp:109-211) */
typedef struct mpb_plan mpb_plan;       /* whole-problem engine state (no reference equivalent: the
                                           reference sizes everything at configure time) */
typedef struct mpb_comm mpb_comm;       /* one rank of a multi-GPU job: an NCCL communicator bound to a context (replaces the
                                           replication over MP_KERNEL_REPL compute units, host/MatrixProfileHost.cpp:78-97) */

const char *mpb_last_error(void);
const char *mpb_version(void);

/* ---- device runtime: what host/MatrixProfileHost.cpp:69-97,110-122 uses of OpenCL.hpp ---- */
int mpb_device_count(int *count);
/* OpenCL::Context() (OpenCL.hpp:294-298): opens device `ordinal`, creates its command stream. */
int mpb_context_create(int ordinal, mpb_context **ctx);
int mpb_context_destroy(mpb_context *ctx);
/* Context::Finish() (OpenCL.hpp:85-89): waits for all enqueued work; *elapsed_ns (may be NULL) gets the
 * device time spent since the previous Finish (CUDA events), as the reference returns nanoseconds. */
int mpb_context_finish(mpb_context *ctx, int64_t *elapsed_ns);
/* The CUDA stream (cudaStream_t) work is enqueued on, for callers that time with their own events. */
void *mpb_context_stream(mpb_context *ctx);
int mpb_context_device(mpb_context *ctx);
/* Context::MakeBuffer<T,Access>(bank, count) (OpenCL.hpp:305-308; ctor :113-127). No memory banks on a GPU. */
int mpb_buffer_create(mpb_context *ctx, size_t bytes, mpb_buffer **buf);
int mpb_buffer_destroy(mpb_buffer *buf);
/* Buffer::CopyFromHost(begin,end) (OpenCL.hpp:131-141): asynchronous enqueue; completes by Finish(). */
int mpb_buffer_copy_from_host(mpb_buffer *buf, const void *src, size_t bytes);
/* Buffer::CopyToHost(ptr,count) (OpenCL.hpp:155-161): blocking read. */
int mpb_buffer_copy_to_host(mpb_buffer *buf, void *dst, size_t bytes);
void *mpb_buffer_device_ptr(mpb_buffer *buf);
size_t mpb_buffer_size(mpb_buffer *buf);

/* ---- kernel-level contract: MatrixProfileKernelTLF(n, m, iteration, input, output)
 *      (kernel/MatrixProfileKernelVanilla.cpp:16-17, kernel/MatrixProfileKernelTiled.cpp:250-251;
 *      plugin interface test/include/MatrixProfileTests.hpp:67).  w (MP_SIZE_W) and data_t are
 *      configure-time constants upstream and explicit arguments here.
 * Computes diagonal chunk `iteration` ([iteration*w, (iteration+1)*w)) over rows [0, n-m+1-iteration*w):
 * output[i] = { row aggregate of row i, column aggregate of column iteration*w + i }, including the
 * kernels' P=0 substitution for excluded / out-of-range cells and the Vanilla kernel's tie order, with
 * un-fused arithmetic -- byte-for-byte what the reference kernel compiled as plain C++ writes
 * (field bytes; struct padding is zeroed).
 * Kernel::SetKernelArguments(0, n, m, iteration, bufIn, bufOut) + Kernel::EnqueueTask()
 * (OpenCL.hpp:246-267, 379-381; call sites host/MatrixProfileHost.cpp:110-113). Asynchronous. */
int mpb_kernel_tlf_enqueue(mpb_context *ctx, int dtype, int n, int m, int w, int iteration,
                           const mpb_buffer *input, mpb_buffer *output);
/* Allocation-free variant of mpb_kernel_tlf_enqueue.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, same bytes written.  The chunk's work space -- 32 bytes per subsequence + 8*w bytes per 192 rows -- is a
 * file-scope __device__ array of MPB_TLF_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit; a problem
 * that needs more is rejected with MPB_E_INVALID (n = 65536, m = 256 at w = 4096 needs 12.6 MiB; n = 2^20 does not fit).
 * The call itself performs no cudaMalloc / cudaFree and calls no allocating library routine (the first call on a device looks
 * the array's address up with cudaGetSymbolAddress: make that one outside a graph capture).  All calls share the one array
 * and it keeps the previous call's contents: what must start at zero is cleared by the call itself, and two calls that
 * overlap on the GPU (different streams / contexts of one device) race -- serialise them. */
#define MPB_TLF_NOALLOC_SCRATCH_BYTES (16u << 20)
int mpb_kernel_tlf_enqueue_noalloc(mpb_context *ctx, int dtype, int n, int m, int w, int iteration,
                                   const mpb_buffer *input, mpb_buffer *output);
/* Host-pointer form of the same contract (what a MatrixProfileKernel<data_t,int,w> subclass calls):
 * copies input (n-m+1 InputDataPacks) to the device, runs the chunk, copies back nRows OutputDataPacks. */
int mpb_kernel_tlf_f64(int n, int m, int w, int iteration, const void *input, void *output);
int mpb_kernel_tlf_f32(int n, int m, int w, int iteration, const void *input, void *output);

/* ---- whole-problem engine: the device section of RunMatrixProfileKernel
 *      (host/MatrixProfileHost.cpp:66-142) in three kernels: precompute (PreComputeStatistics,
 *      include/host/HostSideComputation.hpp:20-45), SCAMP diagonal bands (both kernels + the
 *      UpdateAggregates merge, :72-90, fused), finalize (PostComputeAggregates, :47-69). ---- */
typedef struct {
    double h2d_ms, precompute_ms, compute_ms, finalize_ms, d2h_ms, total_ms; /* device times (CUDA events) */
    int64_t cells;        /* useful cells L(L+1)/2, L = n-m+1 - m/4 */
    int64_t slow_updates; /* profile slots improved through the exact 128-bit CAS path */
    int32_t launches;     /* kernels launched by the last run */
    int32_t bands;        /* diagonal bands this rank walked */
    int32_t items;        /* (band, row segment) work items = CTAs of the tile kernel */
    int32_t segment_rows; /* rows per re-seeded segment (0 = full-length chains) */
    int64_t filter_hits;  /* cells that passed the QT-domain filter and were checked exactly against the profile */
    int64_t replays;      /* replay passes: 32 lanes x (8 rows x 8 diagonals) walked a second time in the per-cell form (one queued lane-block per lane) */
    int64_t chain_candidates;    /* chain-carry mode: values (slot entries + logged near ties) resolved with their carry */
    double chain_max_correction; /* chain-carry mode: largest relative correction |carry * inv_r * inv_c| / P applied (how far the
                                    independent segments were from the unbroken chain); 0 in the other modes */
    double chain_bound;          /* chain-carry mode: upper bound of the ABSOLUTE correction any cell's P could receive */
    double chain_margin;         /* chain-carry mode: the near-tie margin in effect (relative) */
    int64_t cells_walked;        /* cells this rank's tile kernel walks (useful cells + exclusion-free padding of the last tiles / bands) */
    int64_t reseeded_diagonals;  /* (diagonal, segment) pairs that start from an exact dot product of m terms */
    int64_t replay_lane_blocks;  /* (lane, 8 x 8-cell block) pairs that passed the block filter and were queued for a replay pass */
} mpb_timings;

/* One-call form with HOST buffers (copies inside): T[n] float64 in; MP[n-m+1] float64 (z-normalised
 * Euclidean distance) and MPI[n-m+1] int32 out -- the .tsb -> .mpb/.mpib contract
 * (include/host/FileIO.hpp:24-48,105-121; host/MatrixProfileHost.cpp:62,144-153). */
int mpb_compute(mpb_context *ctx, const double *T, int n, int m, int dtype, unsigned flags,
                double *MP, int32_t *MPI, mpb_timings *timings);

/* Plan form: buffers stay resident in HBM between steps. */
int mpb_plan_create(mpb_context *ctx, int n, int m, int dtype, unsigned flags, mpb_plan **plan);
int mpb_plan_destroy(mpb_plan *plan);
/* Diagonal-band shard: this plan walks bands b with b % world == rank (the reference's round-robin of
 * chunks over MP_KERNEL_REPL compute units, host/MatrixProfileHost.cpp:101-128). Default (0, 1). */
int mpb_plan_set_shard(mpb_plan *plan, int rank, int world);
/* Rows per re-seeded segment (multiple of the tile height, 192; 0 = never re-seed; negative = choose automatically, the default). */
int mpb_plan_set_segment_rows(mpb_plan *plan, int rows);
int mpb_plan_segment_rows(mpb_plan *plan);
/* Local bands covered by the sketch pass that precedes the first tile-kernel launch of a step (0 = off, negative =
 * default): it walks those bands once without the hit machinery and leaves a lower bound of every slot's best
 * correlation in the profile, so the filter is tight from the first tile on.  Result-neutral. */
int mpb_plan_set_sketch_bands(mpb_plan *plan, int bands);
int mpb_plan_upload_series(mpb_plan *plan, const double *T_host);   /* async H2D of T[n] */
void *mpb_plan_series_device_ptr(mpb_plan *plan);                   /* double[n] in HBM, caller may fill it */
int mpb_plan_precompute(mpb_plan *plan);                            /* kernel (1) */
int mpb_plan_compute(mpb_plan *plan);                               /* kernel (2) over this rank's bands */
/* The same in pieces: local bands [lb_begin, lb_end) of this rank (lb_end < 0 = all). A multi-GPU job walks a few
 * bands, max-merges the profiles across ranks (export / mask_index / import below), then walks the rest: the filter
 * thresholds of every rank then come from all ranks' first bands instead of its own only. lb_begin == 0 starts the
 * step's compute timer and statistics. */
int mpb_plan_compute_bands(mpb_plan *plan, int lb_begin, int lb_end);
int mpb_plan_local_bands(mpb_plan *plan, int *count);                /* number of diagonal bands this rank walks */
/* Split the packed (corr, index) profile into two dense device arrays for an external max-reduction
 * (NCCL through torch.distributed): corr_out float64[n-m+1] (non-negative, so its int64 bit pattern is
 * order-preserving), index_out int32[n-m+1]. */
int mpb_plan_export(mpb_plan *plan, void *corr_out_dev, void *index_out_dev);
/* After corr was max-reduced across ranks: index_dev[i] = (my corr == corr_max_dev[i]) ? my index : INT32_MAX,
 * so that a min-reduction of index_dev picks the smallest index among the ranks holding the maximum. */
int mpb_plan_mask_index(mpb_plan *plan, const void *corr_max_dev, void *index_dev);
int mpb_plan_import(mpb_plan *plan, const void *corr_dev, const void *index_dev);
/* Threshold sharing inside one process: every improvement this plan's tile kernel makes to its profile is also pushed to the
 * threshold arrays of `others` (same problem; this device or peer-accessible ones), so a slot's threshold is discovered once per
 * job.  One direction: call it on every plan of the job.  Thresholds are hints -- results do not depend on them. */
int mpb_plan_share_thresholds(mpb_plan *plan, mpb_plan *const *others, int count);
int mpb_plan_shared_with(mpb_plan *plan);           /* number of other plans / ranks this plan pushes to */
/* Near-tie margin of the chain-carry mode: 2^-log2_inverse relative (8..21, default 18).  Candidates within the margin of a
 * slot's best are deferred to the resolve step; the margin must exceed the corrections (mpb_timings.chain_max_correction). */
int mpb_plan_set_chain_margin(mpb_plan *plan, int log2_inverse);
/* Total capacity of the chain-carry near-tie log in entries of 16 bytes (<= 0: automatic, 16 per subsequence). */
int mpb_plan_set_log_entries(mpb_plan *plan, int64_t entries);
/* In-library merge over NVLink peer memory (single process, several devices): folds the profiles of
 * `others` into `plan` with an exact (corr, index) max; peers are read through P2P loads. */
int mpb_plan_merge_peers(mpb_plan *plan, mpb_plan *const *others, int count);
int mpb_plan_finalize(mpb_plan *plan);                              /* kernel (3): Pearson -> distance */
int mpb_plan_download(mpb_plan *plan, double *MP_host, int32_t *MPI_host, double *corr_host /* may be NULL */);
void *mpb_plan_mp_device_ptr(mpb_plan *plan);                       /* double[n-m+1] */
void *mpb_plan_mpi_device_ptr(mpb_plan *plan);                      /* int32[n-m+1] */
/* The precomputed InputDataPack[n-m+1] (reference layout), for parity checks of kernel (1). */
int mpb_plan_download_input(mpb_plan *plan, void *input_host);
/* Waits for the plan's work and reports what only the device knows: MPB_E_NOMEM if the near-tie log overflowed,
 * MPB_E_MARGIN if the chain-carry corrections exceed what the margin covers (mpb_plan_download runs the same check). */
int mpb_plan_check(mpb_plan *plan);
int mpb_plan_timings(mpb_plan *plan, mpb_timings *out);
/* After mpb_plan_precompute with the reference's running mean: *walked = 1 when the chain was walked element by element
 * (MPB_FLAG_MU_WALK, or the verified segmented-scan formulation failed a check and fell back), 0 when the scan stood. */
int mpb_plan_mu_walked(mpb_plan *plan, int *walked);
int64_t mpb_useful_cells(int n, int m);

/* ---- multi-GPU: diagonal bands round-robin over the ranks, profiles max-merged over NVLink with NCCL
 *      (the reference: chunk k on compute unit k % MP_KERNEL_REPL, host-side UpdateAggregates;
 *      host/MatrixProfileHost.cpp:101-137, include/host/HostSideComputation.hpp:72-90).
 *      NCCL is bound at run time (libnccl.so.2); single-GPU use needs none. ---- */
#define MPB_COMM_ID_BYTES 128
/* ncclGetUniqueId: call on one rank, hand the 128 bytes to all ranks (MPI, torch.distributed, a file ...). */
int mpb_comm_unique_id(void *id, size_t bytes);
/* ncclCommInitRank on ctx's device.  Collective: every rank of the job must call it. */
int mpb_comm_create(mpb_context *ctx, const void *id, int rank, int world, mpb_comm **comm);
int mpb_comm_destroy(mpb_comm *comm);
int mpb_comm_rank(mpb_comm *comm);
int mpb_comm_world(mpb_comm *comm);
/* Threshold sharing across the ranks (collective, one node): every rank maps the others' threshold arrays -- peer pointers inside
 * one process, CUDA-IPC handles across processes, exchanged with one ncclAllGather -- and its tile kernel pushes every profile
 * improvement to all of them over NVLink (red.max.u64 into peer memory) while it computes. */
int mpb_plan_connect_peers(mpb_plan *plan, mpb_comm *comm);
/* Exact (corr, index) max of the ranks' result profiles, in place, stream-ordered: ncclAllReduce MAX on the correlations' int64
 * bit patterns, then MIN on the indices of the ranks that hold the maximum (ties -> smallest index). */
int mpb_plan_allreduce(mpb_plan *plan, mpb_comm *comm);
/* One rank's compute step of a multi-GPU job (after mpb_plan_set_shard(rank, world) and mpb_plan_precompute): shared sketch
 * of band 0 (each rank walks 1/world of its rows; one max-reduction), this rank's bands (thresholds flowing between the ranks
 * all along when they are connected; else one max-merge after the first bands), max-merge of the result.  Every rank then
 * holds the job's profile; call mpb_plan_finalize on any. */
int mpb_plan_compute_merged(mpb_plan *plan, mpb_comm *comm);
int mpb_merge_point(int total_bands, int world);   /* local bands walked before the threshold merge (0 = no early merge) */
int mpb_plan_total_bands(mpb_plan *plan, int *count);   /* diagonal bands of the whole problem (all ranks) */
/* Pieces of the shared sketch, for callers that bring their own collective: this rank's share of band `band` into hint[]
 * (int32[mpb_plan_hint_count], max-reduce it across the sharing ranks), then mpb_plan_apply_hint. */
int mpb_plan_sketch_share(mpb_plan *plan, int band, int share_rank, int share_world);
void *mpb_plan_hint_device_ptr(mpb_plan *plan);
int mpb_plan_hint_count(mpb_plan *plan);
int mpb_plan_apply_hint(mpb_plan *plan);
/* One call, ngpus GPUs of this process (devices[] or 0..ngpus-1), HOST buffers: one host thread per GPU, NCCL merge, rank 0
 * writes MP / MPI.  timings (may be NULL) receives ngpus entries.  What MatrixProfileHost --gpus G runs. */
int mpb_compute_multi(int ngpus, const int *devices, const double *T, int n, int m, int dtype, unsigned flags,
                      double *MP, int32_t *MPI, mpb_timings *timings);

/* ---- measurement support ---- */
/* Development aid (libraries built with -DMPB_REPLAY_HIST; all zeros otherwise): hist33[k] = replayed blocks of the last step
 * in which k of the warp's 32 lanes passed the block filter, k = 0..32. */
int mpb_plan_replay_histogram(mpb_plan *plan, int64_t *hist33);
/* Register-resident DFMA (dtype MPB_F64) / FFMA (MPB_F32) issue-rate microbenchmark on ctx's device:
 * *fma_per_s = fused multiply-adds retired per second over `iters` dependent-chain iterations per thread.
 * This is the denominator of the FMA roofline. */
int mpb_measure_fma_peak(mpb_context *ctx, int dtype, int iters, double *fma_per_s, double *ms);

#ifdef __cplusplus
}
#endif
#endif
