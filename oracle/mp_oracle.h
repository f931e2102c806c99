/*
 * mp_oracle.h -- CPU oracle for the matrix-profile hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This is a plain-C restatement of the reference algorithm (jlscheerer/matrix-profile-hls).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * call it; the product path (matrix-profile-hls_b200/csrc) never links or loads it.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function here against
 *   (1) the reference's own kernel sources compiled from /root/reference into oracle/_ref
 *       (oracle/ref_shim.cpp; byte-for-byte on per-chunk OutputDataPacks), and
 *   (2) the reference's test-side SCAMP (test/include/MatrixProfileReference.hpp:90-136) on the
 *       reference's eight known-answer cases (test/TestVanillaKernel.cpp:29-42,
 *       test/TestTiledKernel.cpp:29-52), via golden vectors in tests/golden/ produced by
 *       oracle/gen_golden.py from those reference sources.
 *
 * All record layouts are byte-identical to the reference's (include/Config.hpp.in:43-73).
 */
#ifndef MP_ORACLE_H
#define MP_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* include/Config.hpp.in:43-57 (aggregate_t), :61-66 (InputDataPack), :68-73 (OutputDataPack) */
typedef struct { double value; int32_t index; } mpo_aggregate_f64;          /* 16 B (4 B tail padding) */
typedef struct { double QT, df, dg, inv; } mpo_input_f64;                   /* 32 B */
typedef struct { mpo_aggregate_f64 rowAggregate, columnAggregate; } mpo_output_f64; /* 32 B */
typedef struct { float value; int32_t index; } mpo_aggregate_f32;           /*  8 B */
typedef struct { float QT, df, dg, inv; } mpo_input_f32;                    /* 16 B */
typedef struct { mpo_aggregate_f32 rowAggregate, columnAggregate; } mpo_output_f32; /* 16 B */

/* include/AggregateTypeTraits.hpp:14-22,35-38 */
#define MPO_AGGREGATE_INIT (-1e12)
#define MPO_INDEX_INIT (-1)

/* include/host/HostSideComputation.hpp:20-45 -- PreComputeStatistics: all arithmetic in double,
 * cast to data_t on store.  T has n elements, out has n-m+1. */
void mpo_precompute_f64(const double *T, int n, int m, mpo_input_f64 *out);
void mpo_precompute_f32(const double *T, int n, int m, mpo_input_f32 *out);
/* test/include/MatrixProfileTests.hpp:117-139 -- test-harness twin, all arithmetic in data_t
 * (identical to the above for double). */
void mpo_precompute_datat_f32(const float *T, int n, int m, mpo_input_f32 *out);

/* kernel/MatrixProfileKernelVanilla.cpp:16-107 -- one diagonal chunk `iteration` of width w.
 * output[0 .. n-m+1-iteration*w) is overwritten. Requires w >= 8. */
void mpo_kernel_vanilla_f64(int n, int m, int w, int iteration, const mpo_input_f64 *in, mpo_output_f64 *out);
void mpo_kernel_vanilla_f32(int n, int m, int w, int iteration, const mpo_input_f32 *in, mpo_output_f32 *out);
/* kernel/MatrixProfileKernelTiled.cpp:38-277 -- same chunk through the systolic chain of
 * ceil(w/t) DiagonalComputeElements. Requires w % t == 0 and t >= 8 (the reference's own tests
 * never leave that regime; ragged w reads uninitialised PE state upstream). */
void mpo_kernel_tiled_f64(int n, int m, int w, int t, int iteration, const mpo_input_f64 *in, mpo_output_f64 *out);
void mpo_kernel_tiled_f32(int n, int m, int w, int t, int iteration, const mpo_input_f32 *in, mpo_output_f32 *out);

/* include/host/HostSideComputation.hpp:72-90 -- UpdateAggregates for chunk `iteration`. */
void mpo_update_aggregates_f64(int n, int m, int w, int iteration, const mpo_output_f64 *chunk,
                               mpo_aggregate_f64 *rowAgg, mpo_aggregate_f64 *colAgg);
void mpo_update_aggregates_f32(int n, int m, int w, int iteration, const mpo_output_f32 *chunk,
                               mpo_aggregate_f32 *rowAgg, mpo_aggregate_f32 *colAgg);
/* include/host/HostSideComputation.hpp:47-69 -- PostComputeAggregates (+ Pearson -> distance).
 * corr (optional, may be NULL) receives the merged Pearson correlation before conversion. */
void mpo_post_compute_f64(int n, int m, const mpo_aggregate_f64 *rowAgg, const mpo_aggregate_f64 *colAgg,
                          double *MP, int32_t *MPI, double *corr);
void mpo_post_compute_f32(int n, int m, const mpo_aggregate_f32 *rowAgg, const mpo_aggregate_f32 *colAgg,
                          double *MP, int32_t *MPI, double *corr);

/* host/MatrixProfileHost.cpp:54-160 -- the whole driver flow (precompute, chunk loop, merges,
 * post-compute). kernel: 0 = Vanilla, 1 = Tiled (t used only then). nthreads > 1 computes chunks
 * concurrently (one private OutputDataPack buffer per thread, merged in chunk order, which is
 * what the reference's MP_KERNEL_REPL replication does, host/MatrixProfileHost.cpp:101-137).
 * max_chunks < 0 runs every chunk; otherwise only the first max_chunks (bounded bench sample).
 * Returns the number of (row, diagonal) cells the kernels processed. */
int64_t mpo_run_f64(const double *T, int n, int m, int w, int t, int kernel, int nthreads, int max_chunks,
                    double *MP, int32_t *MPI, double *corr);
int64_t mpo_run_f32(const double *T, int n, int m, int w, int t, int kernel, int nthreads, int max_chunks,
                    double *MP, int32_t *MPI, double *corr);

/* test/include/MatrixProfileReference.hpp:90-136 -- Reference::ComputeMatrixProfile, the
 * reference's independent textbook SCAMP (true-max semantics, first-seen wins ties). */
void mpo_reference_scamp_f64(const double *T, int n, int m, double *MP, int32_t *MPI, double *corr);

/* Exact (long double, direct dot product) correlation of subsequences i and j: sampled-row check. */
double mpo_exact_corr(const double *T, int n, int m, int i, int j);

#ifdef __cplusplus
}
#endif
#endif
