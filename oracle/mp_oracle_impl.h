/*
 * mp_oracle_impl.h -- type-generic body of the CPU oracle (TEST INFRASTRUCTURE ONLY).
 * Included twice by mp_oracle.c with DATA_T = double / float and SUF = f64 / f32.
 * Compile with -ffp-contract=off: the reference's shipped build has no FMA contraction.
 */

#define CAT_(a, b) a##_##b
#define CAT(a, b) CAT_(a, b)
#define AGG CAT(mpo_aggregate, SUF)
#define INP CAT(mpo_input, SUF)
#define OUTP CAT(mpo_output, SUF)

static inline AGG CAT(agg_make, SUF)(DATA_T v, int32_t idx) {
    AGG a; memset(&a, 0, sizeof a); a.value = v; a.index = idx; return a;
}
#define AGG_INIT CAT(agg_make, SUF)((DATA_T)MPO_AGGREGATE_INIT, MPO_INDEX_INIT)

/* include/kernel/TreeReduce.hpp:16-75 -- pairwise max tree over 8 lanes; on equal values the
 * HIGHER lane survives (`arr[2i] > arr[2i+1] ? arr[2i] : arr[2i+1]`, :28,:42). */
static AGG CAT(tree_reduce8, SUF)(const AGG *arr) {
    AGG r4[4], r2[2];
    for (int i = 0; i < 4; ++i) r4[i] = arr[2 * i].value > arr[2 * i + 1].value ? arr[2 * i] : arr[2 * i + 1];
    for (int i = 0; i < 2; ++i) r2[i] = r4[2 * i].value > r4[2 * i + 1].value ? r4[2 * i] : r4[2 * i + 1];
    return r2[0].value > r2[1].value ? r2[0] : r2[1];
}

/* kernel/MatrixProfileKernelVanilla.cpp:16-107 */
void CAT(mpo_kernel_vanilla, SUF)(int n, int m, int w, int iteration, const INP *input, OUTP *output) {
    const int sublen = n - m + 1;
    const int nOffset = iteration * w;              /* :23 */
    const int nRows = sublen - nOffset;             /* :24 */
    DATA_T *QT = (DATA_T *)malloc(sizeof(DATA_T) * (size_t)w);
    AGG rowReduce[8][8];                            /* :29-36 */
    AGG *colAgg = (AGG *)malloc(sizeof(AGG) * (size_t)w * 2); /* :40, [w][2] */
    for (int a = 0; a < 8; ++a) for (int b = 0; b < 8; ++b) rowReduce[a][b] = AGG_INIT;

    for (int i = 0; i < w; ++i) {                   /* :43-55 MatrixProfileInitQT */
        const int columnIndex = nOffset + i;
        QT[i] = columnIndex < sublen ? input[columnIndex].QT : (DATA_T)0;
    }
    for (int i = 0; i < nRows; ++i) {               /* :57-106 */
        const int rowIndex = i;
        const INP row = input[rowIndex];
        for (int j = 0; j < w; ++j) {
            const int columnIndex = nOffset + i + j;                     /* :68 */
            const int columnInBounds = columnIndex < sublen;             /* :72 */
            const int exclusionZone = rowIndex > columnIndex - m / 4;    /* :73 */
            const int inBounds = columnInBounds && !exclusionZone;       /* :74 */
            INP column;
            if (columnInBounds) column = input[columnIndex];
            else { column.QT = 0; column.df = 0; column.dg = 0; column.inv = 0; }   /* :77-78 */
            QT[j] += row.df * column.dg + column.df * row.dg;            /* :81 */
            const DATA_T P = inBounds ? (DATA_T)(QT[j] * row.inv * column.inv) : (DATA_T)0;   /* :84 */
            const AGG prevRow = (j < 8) ? AGG_INIT : rowReduce[i % 8][j % 8];               /* :87-88 */
            rowReduce[i % 8][j % 8] = prevRow.value > P ? prevRow : CAT(agg_make, SUF)(P, columnIndex); /* :89-90 */
            const AGG prevColumn = (j < w - 1 && i > 0) ? colAgg[(j + 1) * 2 + (i % 2)] : AGG_INIT; /* :93-94 */
            colAgg[j * 2 + ((i + 1) % 2)] = P > prevColumn.value ? CAT(agg_make, SUF)(P, rowIndex) : prevColumn; /* :95-96 */
        }
        output[i].rowAggregate = CAT(tree_reduce8, SUF)(rowReduce[i % 8]);  /* :100 */
        output[i].columnAggregate = colAgg[0 * 2 + ((i + 1) % 2)];          /* :102,:105 */
    }
    free(QT); free(colAgg);
}

/* kernel/MatrixProfileKernelTiled.cpp:38-277.  The hls::stream FIFOs (include/kernel/Stream.hpp:14-17)
 * carry one ComputePack per row between neighbouring DiagonalComputeElements; here each element is
 * run to completion over all rows and its per-row output packs are kept in an array that the next
 * element consumes -- the same values in the same order, without the queue objects. */
typedef struct { DATA_T df, dg, inv; } CAT(dpack, SUF);                 /* include/kernel/DataPacks.hpp:11-18 */
typedef struct { CAT(dpack, SUF) row; AGG rowAggregate; CAT(dpack, SUF) column; AGG columnAggregate; } CAT(cpack, SUF); /* :20-28 */

void CAT(mpo_kernel_tiled, SUF)(int n, int m, int w, int t, int iteration, const INP *input, OUTP *output) {
    const int sublen = n - m + 1;
    const int nOffset = iteration * w;          /* :44 */
    const int nRows = sublen - nOffset;         /* :45 */
    const int nPE = (w + t - 1) / t;            /* :258 */
    if (w % t != 0 || t < 8) { fprintf(stderr, "mpo_kernel_tiled: need w %% t == 0 and t >= 8\n"); abort(); }
    typedef CAT(cpack, SUF) CP;
    typedef CAT(dpack, SUF) DP;
    CP *stream = (CP *)malloc(sizeof(CP) * (size_t)(nRows > 0 ? nRows : 1));
    /* MemoryToStream compute phase, :61-88 */
    for (int i = 0; i < nRows; ++i) {
        const INP r = input[i];
        CP p; memset(&p, 0, sizeof p);
        p.row.df = r.df; p.row.dg = r.dg; p.row.inv = r.inv;
        p.rowAggregate = AGG_INIT;
        const int columnIndex = nOffset + w + i;                 /* :76 */
        if (columnIndex < sublen) { const INP c = input[columnIndex]; p.column.df = c.df; p.column.dg = c.dg; p.column.inv = c.inv; }
        p.columnAggregate = AGG_INIT;
        stream[i] = p;
    }
    DP *columns = (DP *)malloc(sizeof(DP) * (size_t)t);
    AGG *columnAggregates = (AGG *)malloc(sizeof(AGG) * (size_t)t);
    DATA_T *QT = (DATA_T *)malloc(sizeof(DATA_T) * (size_t)t);
    for (int stage = 0; stage < nPE; ++stage) {                  /* :267-273, DiagonalComputeElement :93-224 */
        const int revStage = (w - 1) / t - stage;                /* :106 */
        const int afterMe = t * revStage;                        /* :116 */
        /* scatter phase :124-137 (scatter stream carries input[nOffset + i], zeros out of bounds :47-58) */
        for (int j = 0; j < t; ++j) {
            const int columnIndex = nOffset + afterMe + j;
            if (columnIndex < sublen) {
                const INP c = input[columnIndex];
                QT[j] = c.QT; columns[j].df = c.df; columns[j].dg = c.dg; columns[j].inv = c.inv;
            } else { QT[j] = 0; columns[j].df = 0; columns[j].dg = 0; columns[j].inv = 0; }
            columnAggregates[j] = AGG_INIT;
        }
        AGG rowReduce[8][8];
        for (int a = 0; a < 8; ++a) for (int b = 0; b < 8; ++b) rowReduce[a][b] = AGG_INIT;
        for (int i = 0; i < nRows; ++i) {                        /* :149-223 */
            const CP read = stream[i];
            const DP row = read.row;
            const AGG rowAggregateBackward = read.rowAggregate;
            for (int j = 0; j < t; ++j) {
                const DP column = columns[j];
                QT[j] += row.df * column.dg + column.df * row.dg;               /* :172 */
                const int rowIndex = i;
                const int columnIndex = nOffset + afterMe + i + j;              /* :176 */
                const int columnInBounds = columnIndex < sublen;
                const int exclusionZone = rowIndex > columnIndex - m / 4;       /* :181 */
                const int inBounds = columnInBounds && !exclusionZone;
                const DATA_T P = inBounds ? (DATA_T)(QT[j] * row.inv * column.inv) : (DATA_T)0;   /* :185 */
                const AGG prevRow = (j < 8) ? rowAggregateBackward : rowReduce[i % 8][j % 8];     /* :188-189 */
                rowReduce[i % 8][j % 8] = prevRow.value > P ? prevRow : CAT(agg_make, SUF)(P, columnIndex); /* :190-191 */
                const AGG prevColumn = (i > 0) ? columnAggregates[j] : AGG_INIT;                  /* :194-195 */
                columnAggregates[j] = (prevColumn.value > P) ? prevColumn : CAT(agg_make, SUF)(P, rowIndex); /* :196-197 */
            }
            const AGG rowAggregate = CAT(tree_reduce8, SUF)(rowReduce[i % 8]);  /* :201 */
            const DP columnForward = columns[0];                                  /* :204 */
            const AGG columnAggregateForward = columnAggregates[0];               /* :205 */
            for (int j = 0; j < t - 1; ++j) { columns[j] = columns[j + 1]; columnAggregates[j] = columnAggregates[j + 1]; } /* :210-216 */
            columns[t - 1] = read.column;                                         /* :218 */
            columnAggregates[t - 1] = read.columnAggregate;                       /* :219 */
            CP o; memset(&o, 0, sizeof o);
            o.row = row; o.rowAggregate = rowAggregate; o.column = columnForward; o.columnAggregate = columnAggregateForward;
            stream[i] = o;                                                        /* :222 */
        }
    }
    for (int i = 0; i < nRows; ++i) {                            /* StreamToMemory :234-247 */
        output[i].rowAggregate = stream[i].rowAggregate;
        output[i].columnAggregate = stream[i].columnAggregate;
    }
    free(stream); free(columns); free(columnAggregates); free(QT);
}

/* include/host/HostSideComputation.hpp:72-90 */
void CAT(mpo_update_aggregates, SUF)(int n, int m, int w, int iteration, const OUTP *chunk, AGG *rowAgg, AGG *colAgg) {
    const int nOffset = iteration * w;
    const int nRows = n - m + 1 - nOffset;
    for (int i = 0; i < nRows; ++i) {
        const AGG prevRow = iteration > 0 ? rowAgg[i] : AGG_INIT;               /* :79 */
        const AGG prevCol = iteration > 0 ? colAgg[i + nOffset] : AGG_INIT;     /* :80 */
        const AGG currRow = chunk[i].rowAggregate, currCol = chunk[i].columnAggregate;
        rowAgg[i] = currRow.value > prevRow.value ? currRow : prevRow;          /* :85 */
        colAgg[i + nOffset] = currCol.value > prevCol.value ? currCol : prevCol;/* :86 */
    }
}

/* include/host/HostSideComputation.hpp:47-69 */
void CAT(mpo_post_compute, SUF)(int n, int m, const AGG *rowAgg, const AGG *colAgg, double *MP, int32_t *MPI, double *corr) {
    for (int i = 0; i < n - m + 1; ++i) {
        const AGG a = rowAgg[i].value > colAgg[i].value ? rowAgg[i] : colAgg[i];  /* :61 */
        if (corr) corr[i] = (double)a.value;
        MP[i] = sqrt(2 * m * (1 - (double)a.value));                               /* :48,:64 -- P promoted to double */
        MPI[i] = a.index;                                                          /* :65 */
    }
}

typedef struct { int n, m, w, t, kernel, iteration; const INP *input; OUTP *output; } CAT(chunk_job, SUF);
static void *CAT(chunk_worker, SUF)(void *arg) {
    CAT(chunk_job, SUF) *j = (CAT(chunk_job, SUF) *)arg;
    if (j->kernel == 0) CAT(mpo_kernel_vanilla, SUF)(j->n, j->m, j->w, j->iteration, j->input, j->output);
    else CAT(mpo_kernel_tiled, SUF)(j->n, j->m, j->w, j->t, j->iteration, j->input, j->output);
    return NULL;
}

/* host/MatrixProfileHost.cpp:54-160 */
int64_t CAT(mpo_run, SUF)(const double *T, int n, int m, int w, int t, int kernel, int nthreads, int max_chunks,
                          double *MP, int32_t *MPI, double *corr) {
    const int sublen = n - m + 1;
    INP *input = (INP *)malloc(sizeof(INP) * (size_t)sublen);
    CAT(mpo_precompute, SUF)(T, n, m, input);                                    /* :66 */
    AGG *rowAgg = (AGG *)malloc(sizeof(AGG) * (size_t)sublen), *colAgg = (AGG *)malloc(sizeof(AGG) * (size_t)sublen);
    for (int i = 0; i < sublen; ++i) { rowAgg[i] = AGG_INIT; colAgg[i] = AGG_INIT; }
    int nIterations = (n - m + w) / w;                                           /* :101 */
    if (max_chunks >= 0 && max_chunks < nIterations) nIterations = max_chunks;
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 64) nthreads = 64;
    OUTP **bufs = (OUTP **)malloc(sizeof(OUTP *) * (size_t)nthreads);
    for (int k = 0; k < nthreads; ++k) bufs[k] = (OUTP *)malloc(sizeof(OUTP) * (size_t)sublen);
    int64_t cells = 0;
    for (int base = 0; base < nIterations; base += nthreads) {                   /* batches of kNumKernels, :101-128 */
        const int cnt = nIterations - base < nthreads ? nIterations - base : nthreads;
        CAT(chunk_job, SUF) jobs[64]; pthread_t th[64];
        for (int k = 0; k < cnt; ++k) {
            jobs[k].n = n; jobs[k].m = m; jobs[k].w = w; jobs[k].t = t; jobs[k].kernel = kernel;
            jobs[k].iteration = base + k; jobs[k].input = input; jobs[k].output = bufs[k];
            if (k > 0) pthread_create(&th[k], NULL, CAT(chunk_worker, SUF), &jobs[k]);
        }
        CAT(chunk_worker, SUF)(&jobs[0]);
        for (int k = 1; k < cnt; ++k) pthread_join(th[k], NULL);
        for (int k = 0; k < cnt; ++k) {                                          /* :125-126, in chunk order */
            CAT(mpo_update_aggregates, SUF)(n, m, w, base + k, bufs[k], rowAgg, colAgg);
            cells += (int64_t)(sublen - (base + k) * w) * w;
        }
    }
    CAT(mpo_post_compute, SUF)(n, m, rowAgg, colAgg, MP, MPI, corr);             /* :142 */
    for (int k = 0; k < nthreads; ++k) free(bufs[k]);
    free(bufs); free(input); free(rowAgg); free(colAgg);
    return cells;
}

#undef AGG_INIT
#undef AGG
#undef INP
#undef OUTP
#undef CAT
#undef CAT_
