/*
 * mp_oracle.c -- CPU oracle for the matrix-profile hot path.  TEST INFRASTRUCTURE ONLY
 * (see mp_oracle.h for who may call it and how parity is pinned).
 * Build: gcc -O2 -ffp-contract=off -pthread -shared -fPIC (oracle/Makefile).
 * Every function cites the reference file:line (relative to /root/reference) it restates.
 */
#include "mp_oracle.h"
#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* include/host/HostSideComputation.hpp:20-45.  mean is the reference's *running* mean (:29):
 * a strictly sequential recurrence whose rounding the engine must reproduce bit for bit. */
#define DEFINE_PRECOMPUTE(NAME, OUT_T, CAST_T)                                                   \
    void NAME(const double *T, int n, int m, OUT_T *data) {                                      \
        double mean = 0;                                                                         \
        for (int k = 0; k < m; ++k) mean += T[k]; /* std::accumulate, :23 */                     \
        mean /= m;                                                                               \
        double prev_mu; const double mu0 = mean;                                                 \
        for (int i = 0; i < n - m + 1; ++i) {                                                    \
            prev_mu = mean;                                                                      \
            mean += (i > 0) ? ((T[i + m - 1] - T[i - 1]) / m) : 0;                    /* :29 */  \
            data[i].df = (i > 0) ? (CAST_T)((T[i + m - 1] - T[i - 1]) / 2) : (CAST_T)0; /* :30 */\
            data[i].dg = (i > 0) ? (CAST_T)((T[i + m - 1] - mean) + (T[i - 1] - prev_mu)) : (CAST_T)0; /* :32 */ \
            double qt = 0, inv = 0;                                                              \
            for (int k = 0; k < m; ++k) {                                                        \
                qt += (T[i + k] - mean) * (T[k] - mu0);                               /* :37 */  \
                inv += (T[i + k] - mean) * (T[i + k] - mean);                         /* :38 */  \
            }                                                                                    \
            data[i].QT = (CAST_T)qt;                                                  /* :40 */  \
            data[i].inv = (CAST_T)(1 / sqrt(inv));                                    /* :41 */  \
        }                                                                                        \
    }
DEFINE_PRECOMPUTE(mpo_precompute_f64, mpo_input_f64, double)
DEFINE_PRECOMPUTE(mpo_precompute_f32, mpo_input_f32, float)

/* test/include/MatrixProfileTests.hpp:117-139, data_t = float */
void mpo_precompute_datat_f32(const float *T, int n, int m, mpo_input_f32 *data) {
    float mean = 0;
    for (int k = 0; k < m; ++k) mean += T[k];
    mean /= m;
    float prev_mu; const float mu0 = mean;
    for (int i = 0; i < n - m + 1; ++i) {
        prev_mu = mean;
        mean += (i > 0) ? ((T[i + m - 1] - T[i - 1]) / m) : 0;
        data[i].df = (i > 0) ? (T[i + m - 1] - T[i - 1]) / 2 : 0.0f;
        data[i].dg = (i > 0) ? ((T[i + m - 1] - mean) + (T[i - 1] - prev_mu)) : 0.0f;
        data[i].QT = 0; data[i].inv = 0;
        for (int k = 0; k < m; ++k) {
            data[i].QT += (T[i + k] - mean) * (T[k] - mu0);
            data[i].inv += (T[i + k] - mean) * (T[i + k] - mean);
        }
        data[i].inv = 1.0f / sqrtf(data[i].inv);
    }
}

#define DATA_T double
#define SUF f64
#include "mp_oracle_impl.h"
#undef DATA_T
#undef SUF
#define DATA_T float
#define SUF f32
#include "mp_oracle_impl.h"
#undef DATA_T
#undef SUF

/* test/include/MatrixProfileReference.hpp:90-136 (helpers :30-86).  Row-major walk over the upper
 * triangle; excluded cells are skipped (:112,:124), init -1e12/-1 (:99-100), strict > (:83). */
void mpo_reference_scamp_f64(const double *T, int n, int m, double *MP, int32_t *MPI, double *corr) {
    const int sublen = n - m + 1;
    double *mu = (double *)calloc((size_t)sublen, sizeof(double));
    double *df = (double *)calloc((size_t)sublen, sizeof(double));
    double *dg = (double *)calloc((size_t)sublen, sizeof(double));
    double *inv = (double *)calloc((size_t)sublen, sizeof(double));
    double *QT = (double *)calloc((size_t)sublen, sizeof(double));
    for (int i = 0; i < m; ++i) mu[0] += T[i];                                    /* MovMean :30-41 */
    mu[0] /= m;
    for (int i = 1; i < sublen; ++i) mu[i] = mu[i - 1] + (T[i + m - 1] - T[i - 1]) / m;
    for (int i = 1; i < sublen; ++i) df[i] = (T[i + m - 1] - T[i - 1]) / 2;      /* ComputeDf :43-49 */
    for (int i = 1; i < sublen; ++i) dg[i] = (T[i + m - 1] - mu[i]) + (T[i - 1] - mu[i - 1]); /* ComputeDg :51-57 */
    for (int i = 0; i < sublen; ++i) {                                            /* ComputeInv :59-69 */
        double sum = 0;
        for (int k = 0; k < m; ++k) sum += (T[i + k] - mu[i]) * (T[i + k] - mu[i]);
        inv[i] = 1 / sqrt(sum);
    }
    for (int i = 0; i < sublen; ++i) { MP[i] = MPO_AGGREGATE_INIT; MPI[i] = MPO_INDEX_INIT; } /* :99-100 */
#define EXCL(row, column) (((column) - (row)) < m / 4)                            /* :73-79 */
#define UPD(i, j, P) do { if ((P) > MP[i]) { MP[i] = (P); MPI[i] = (j); } } while (0) /* :82-86 */
    for (int i = 0; i < sublen; ++i) {                                            /* first row :105-118 */
        double sum = 0;
        for (int k = 0; k < m; ++k) sum += (T[i + k] - mu[i]) * (T[k] - mu[0]);
        QT[i] = sum;
        const double P = QT[i] * inv[0] * inv[i];
        if (!EXCL(0, i)) { UPD(0, i, P); UPD(i, 0, P); }
    }
    for (int row = 1; row < sublen; ++row)                                        /* :120-131 */
        for (int k = 0; k < sublen - row; ++k) {
            QT[k] += df[row] * dg[k + row] + df[k + row] * dg[row];
            const double P = QT[k] * inv[row] * inv[k + row];
            if (!EXCL(row, k + row)) { UPD(row, k + row, P); UPD(k + row, row, P); }
        }
#undef EXCL
#undef UPD
    for (int i = 0; i < sublen; ++i) {                                            /* :134-135 */
        if (corr) corr[i] = MP[i];
        MP[i] = sqrt(2 * m * (1 - MP[i]));
    }
    free(mu); free(df); free(dg); free(inv); free(QT);
}

/* Not from the reference: exact Pearson correlation in long double, for sampled-row checks at
 * sizes where no full CPU run is feasible. */
double mpo_exact_corr(const double *T, int n, int m, int i, int j) {
    (void)n;
    long double si = 0, sj = 0;
    for (int k = 0; k < m; ++k) { si += T[i + k]; sj += T[j + k]; }
    const long double mi = si / m, mj = sj / m;
    long double qq = 0, ii = 0, jj = 0;
    for (int k = 0; k < m; ++k) {
        const long double a = T[i + k] - mi, b = T[j + k] - mj;
        qq += a * b; ii += a * a; jj += b * b;
    }
    return (double)(qq / (sqrtl(ii) * sqrtl(jj)));
}
