// runtime.cu -- implementation of the C ABI in include/mp_b200.h over the CUDA runtime.
// Replaces the reference's OpenCL/XRT wrapper include/host/OpenCL.hpp (Context / Buffer / Program /
// Kernel) and hosts the whole-problem engine (plan) that drives kernels (1)-(3).
#include "mp_runtime.h"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <set>
#include <vector>

#ifndef MPB_SKETCH_SEG_ROWS
#define MPB_SKETCH_SEG_ROWS (4 * mpb::TILE_H)   // rows per work item of a one-band sketch pass
#endif
#ifndef MPB_DEFAULT_SKETCH_BANDS
#define MPB_DEFAULT_SKETCH_BANDS 1
#endif
#ifndef MPB_CHAIN_MARGIN_LOG2
#define MPB_CHAIN_MARGIN_LOG2 0                  // near-tie margin of the chain-carry mode = 2^-k; 0 = by series length (default_margin_log2)
#endif
// Rows per re-seeded segment when the problem is large enough.  The re-seed costs m FMAs per diagonal and segment (6 % of a step at
// 4224 rows, m = 256).  fp32 and the plain re-seeding modes rely on it to bound the chain's drift (fp32: 4.6e-5 at 4224 rows, n = 2^20,
// against a 1e-4 budget), so they keep the shorter length; the chain-carry mode reproduces the reference's unbroken chain whatever
// the cut, so it takes the longer one (-3 % on both series).
#ifndef MPB_SEG_BASE
#define MPB_SEG_BASE 4096
#endif
#ifndef MPB_SEG_BASE_CARRY
#define MPB_SEG_BASE_CARRY 8192
#endif
#ifndef MPB_CHAIN_CHUNK_ITEMS
#define MPB_CHAIN_CHUNK_ITEMS (1 << 19)          // work items per launch chunk of the chain-carry mode (4 KB of S_k + 4 KB of E_k each)
#endif

namespace mpb {
int fma_peak_chains();
int fma_peak_threads();
}

namespace {
thread_local char g_err[512] = "";
}  // namespace

int mpb_fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

// Live-handle registry: destroying a context destroys the plans and buffers created on it first, and destroying a
// handle twice (or after its context went away) is a no-op -- language bindings finalise objects in any order.
namespace {
std::mutex g_reg_mu;
std::set<mpb_context *> g_ctxs;
std::set<mpb_plan *> g_plans;
std::set<mpb_buffer *> g_bufs;
template <typename H> bool reg_take(std::set<H *> &set, H *h) { std::lock_guard<std::mutex> l(g_reg_mu); return set.erase(h) > 0; }
template <typename H> void reg_add(std::set<H *> &set, H *h) { std::lock_guard<std::mutex> l(g_reg_mu); set.insert(h); }
}  // namespace

extern "C" {

const char *mpb_last_error(void) { return g_err; }
const char *mpb_version(void) { return "mpb200 0.5 (sm_100a; tile kernel: 8 diagonals per lane, 512-diagonal bands, per-warp tiles of 192 rows, block-wide QT-domain filter, (band x row segment) work items, fp64 chain carry, NCCL merge)"; }

int mpb_device_count(int *count) {
    if (!count) return fail(MPB_E_INVALID, "count is null");
    int c = 0;
    cudaError_t e = cudaGetDeviceCount(&c);
    if (e != cudaSuccess || c == 0) {
        *count = 0;
        cudaGetLastError();
        return fail(MPB_E_NO_DEVICE, "no CUDA device: %s", e == cudaSuccess ? "count is 0" : cudaGetErrorString(e));
    }
    *count = c;
    return MPB_OK;
}

int mpb_context_create(int ordinal, mpb_context **out) {
    if (!out) return fail(MPB_E_INVALID, "ctx out-pointer is null");
    *out = nullptr;
    int count = 0;
    int rc = mpb_device_count(&count);
    if (rc != MPB_OK) return rc;
    if (ordinal < 0 || ordinal >= count) return fail(MPB_E_INVALID, "device ordinal %d out of range [0,%d)", ordinal, count);
    CU(cudaSetDevice(ordinal));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, ordinal));
    if (prop.major < 10) return fail(MPB_E_NO_DEVICE, "device %d is sm_%d%d; this library carries sm_100a code only", ordinal, prop.major, prop.minor);
    mpb_context *c = new (std::nothrow) mpb_context();
    if (!c) return fail(MPB_E_NOMEM, "out of host memory");
    c->device = ordinal;
    c->sm_count = prop.multiProcessorCount;
    reg_add(g_ctxs, c);               // from here on mpb_context_destroy(c) releases whatever was created
    cudaError_t ce = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (ce == cudaSuccess) ce = cudaEventCreate(&c->mark);
    if (ce == cudaSuccess) ce = cudaEventCreate(&c->now);
    if (ce == cudaSuccess) ce = cudaEventRecord(c->mark, c->stream);
    if (ce != cudaSuccess) { mpb_context_destroy(c); return fail(MPB_E_CUDA, "creating the context's stream / events: %s", cudaGetErrorString(ce)); }
    *out = c;
    return MPB_OK;
}

int mpb_context_destroy(mpb_context *c) {
    if (!c || !reg_take(g_ctxs, c)) return MPB_OK;
    std::vector<mpb_plan *> plans;
    std::vector<mpb_buffer *> bufs;
    {
        std::lock_guard<std::mutex> l(g_reg_mu);
        for (mpb_plan *p : g_plans) if (p->ctx == c) plans.push_back(p);
        for (mpb_buffer *b : g_bufs) if (b->ctx == c) bufs.push_back(b);
    }
    for (mpb_plan *p : plans) mpb_plan_destroy(p);
    for (mpb_buffer *b : bufs) mpb_buffer_destroy(b);
    cudaSetDevice(c->device);
    if (c->tlf_scratch) cudaFree(c->tlf_scratch);
    if (c->mark) cudaEventDestroy(c->mark);
    if (c->now) cudaEventDestroy(c->now);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return MPB_OK;
}

int mpb_context_finish(mpb_context *c, int64_t *elapsed_ns) {
    if (!c) return fail(MPB_E_INVALID, "ctx is null");
    CU(cudaSetDevice(c->device));
    CU(cudaEventRecord(c->now, c->stream));
    CU(cudaEventSynchronize(c->now));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, c->mark, c->now));
    if (elapsed_ns) *elapsed_ns = (int64_t)((double)ms * 1e6);
    CU(cudaEventRecord(c->mark, c->stream));
    return MPB_OK;
}

void *mpb_context_stream(mpb_context *c) { return c ? (void *)c->stream : nullptr; }
int mpb_context_device(mpb_context *c) { return c ? c->device : -1; }

int mpb_buffer_create(mpb_context *c, size_t bytes, mpb_buffer **out) {
    if (!c || !out) return fail(MPB_E_INVALID, "null argument");
    *out = nullptr;
    CU(cudaSetDevice(c->device));
    mpb_buffer *b = new (std::nothrow) mpb_buffer();
    if (!b) return fail(MPB_E_NOMEM, "out of host memory");
    b->ctx = c;
    b->bytes = bytes;
    cudaError_t e = cudaMalloc(&b->ptr, bytes ? bytes : 1);
    if (e != cudaSuccess) { delete b; return fail(MPB_E_NOMEM, "cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e)); }
    reg_add(g_bufs, b);
    *out = b;
    return MPB_OK;
}

int mpb_buffer_destroy(mpb_buffer *b) {
    if (!b || !reg_take(g_bufs, b)) return MPB_OK;
    cudaSetDevice(b->ctx->device);
    cudaFree(b->ptr);
    delete b;
    return MPB_OK;
}

int mpb_buffer_copy_from_host(mpb_buffer *b, const void *src, size_t bytes) {
    if (!b || !src) return fail(MPB_E_INVALID, "null argument");
    if (bytes > b->bytes) return fail(MPB_E_INVALID, "copy of %zu bytes into a %zu-byte buffer", bytes, b->bytes);
    CU(cudaSetDevice(b->ctx->device));
    CU(cudaMemcpyAsync(b->ptr, src, bytes, cudaMemcpyHostToDevice, b->ctx->stream));
    return MPB_OK;
}

int mpb_buffer_copy_to_host(mpb_buffer *b, void *dst, size_t bytes) {
    if (!b || !dst) return fail(MPB_E_INVALID, "null argument");
    if (bytes > b->bytes) return fail(MPB_E_INVALID, "copy of %zu bytes out of a %zu-byte buffer", bytes, b->bytes);
    CU(cudaSetDevice(b->ctx->device));
    CU(cudaMemcpyAsync(dst, b->ptr, bytes, cudaMemcpyDeviceToHost, b->ctx->stream));
    CU(cudaStreamSynchronize(b->ctx->stream));
    return MPB_OK;
}

void *mpb_buffer_device_ptr(mpb_buffer *b) { return b ? b->ptr : nullptr; }
size_t mpb_buffer_size(mpb_buffer *b) { return b ? b->bytes : 0; }

// ------------------------------------------------------------------ MatrixProfileKernelTLF contract
static int tlf_check(int dtype, int n, int m, int w, int iteration) {
    if (dtype != MPB_F64 && dtype != MPB_F32) return fail(MPB_E_INVALID, "dtype must be MPB_F64 or MPB_F32");
    if (m < 4 || n < m) return fail(MPB_E_INVALID, "need n >= m >= 4 (n=%d, m=%d)", n, m);   // CMakeLists.txt:83-89
    if (w < 8 || w > n || w >= (1 << 24)) return fail(MPB_E_INVALID, "need 8 <= w <= n (w=%d)", w);   // :83-85
    if (iteration < 0 || (long long)iteration * w >= (long long)(n - m + 1)) return fail(MPB_E_INVALID, "iteration %d outside [0, ceil((n-m+1)/w))", iteration);
    return MPB_OK;
}

int mpb_kernel_tlf_enqueue(mpb_context *c, int dtype, int n, int m, int w, int iteration, const mpb_buffer *input, mpb_buffer *output) {
    if (!c || !input || !output) return fail(MPB_E_INVALID, "null argument");
    int rc = tlf_check(dtype, n, m, w, iteration);
    if (rc != MPB_OK) return rc;
    const size_t sublen = (size_t)(n - m + 1), esz = dtype == MPB_F64 ? 32 : 16;
    if (input->bytes < sublen * esz) return fail(MPB_E_INVALID, "input buffer holds %zu bytes, need %zu", input->bytes, sublen * esz);
    const size_t nRows = sublen - (size_t)iteration * w;
    if (output->bytes < nRows * esz) return fail(MPB_E_INVALID, "output buffer holds %zu bytes, need %zu", output->bytes, nRows * esz);
    CU(cudaSetDevice(c->device));
    const size_t need = mpb::chunk_tlf_scratch_bytes(n, m, w);
    if (c->tlf_scratch_bytes < need) {
        CU(cudaStreamSynchronize(c->stream));
        if (c->tlf_scratch) cudaFree(c->tlf_scratch);
        c->tlf_scratch = nullptr; c->tlf_scratch_bytes = 0;
        CU(cudaMalloc(&c->tlf_scratch, need));
        c->tlf_scratch_bytes = need;
    }
    if (dtype == MPB_F64) CU(mpb::launch_chunk_tlf_f64(n, m, w, iteration, input->ptr, output->ptr, c->tlf_scratch, c->tlf_scratch_bytes, c->stream));
    else CU(mpb::launch_chunk_tlf_f32(n, m, w, iteration, input->ptr, output->ptr, c->tlf_scratch, c->tlf_scratch_bytes, c->stream));
    return MPB_OK;
}

static int tlf_host(int dtype, int n, int m, int w, int iteration, const void *input, void *output) {
    if (!input || !output) return fail(MPB_E_INVALID, "null argument");
    int rc = tlf_check(dtype, n, m, w, iteration);
    if (rc != MPB_OK) return rc;
    static thread_local mpb_context *ctx = nullptr;   // one implicit context per calling thread, device 0
    if (!ctx) { rc = mpb_context_create(0, &ctx); if (rc != MPB_OK) return rc; }
    const size_t sublen = (size_t)(n - m + 1), esz = dtype == MPB_F64 ? 32 : 16;
    const size_t nRows = sublen - (size_t)iteration * w;
    mpb_buffer *bin = nullptr, *bout = nullptr;
    rc = mpb_buffer_create(ctx, sublen * esz, &bin);
    if (rc == MPB_OK) rc = mpb_buffer_create(ctx, sublen * esz, &bout);
    if (rc == MPB_OK) rc = mpb_buffer_copy_from_host(bin, input, sublen * esz);
    if (rc == MPB_OK) rc = mpb_kernel_tlf_enqueue(ctx, dtype, n, m, w, iteration, bin, bout);
    if (rc == MPB_OK) rc = mpb_buffer_copy_to_host(bout, output, nRows * esz);
    mpb_buffer_destroy(bin);
    mpb_buffer_destroy(bout);
    return rc;
}
int mpb_kernel_tlf_f64(int n, int m, int w, int iteration, const void *input, void *output) { return tlf_host(MPB_F64, n, m, w, iteration, input, output); }
int mpb_kernel_tlf_f32(int n, int m, int w, int iteration, const void *input, void *output) { return tlf_host(MPB_F32, n, m, w, iteration, input, output); }

// ------------------------------------------------------------------ whole-problem engine
int64_t mpb_useful_cells(int n, int m) {
    const int64_t L = (int64_t)(n - m + 1) - m / 4;
    return L > 0 ? L * (L + 1) / 2 : 0;
}

int mpb_plan_destroy(mpb_plan *p) {
    if (!p || !reg_take(g_plans, p)) return MPB_OK;
    cudaSetDevice(p->ctx->device);
    cudaStreamSynchronize(p->ctx->stream);
    for (void *q : p->ipc_opened) if (q) cudaIpcCloseMemHandle(q);
    void *ptrs[] = {p->thr, p->T, p->mu, p->df_dg, p->inv, p->qt0, p->nm, p->mnc, p->prof, p->MP, p->corr, p->MPI, p->slow_counter, p->item_start, p->hint,
                    p->fin, p->seg_seed, p->seg_end, p->chain_log, p->log_count, p->chain_stats, p->inv_blockmax, p->mu_scratch};
    for (void *q : ptrs) if (q) cudaFree(q);
    for (cudaEvent_t e : p->ev) if (e) cudaEventDestroy(e);
    delete p;
    return MPB_OK;
}

int mpb_plan_create(mpb_context *c, int n, int m, int dtype, unsigned flags, mpb_plan **out) {
    if (!c || !out) return fail(MPB_E_INVALID, "null argument");
    *out = nullptr;
    if (dtype != MPB_F64 && dtype != MPB_F32) return fail(MPB_E_INVALID, "dtype must be MPB_F64 or MPB_F32");
    if (m < 4 || n < m) return fail(MPB_E_INVALID, "need n >= m >= 4 (n=%d, m=%d)", n, m);
    if (n - m + 1 <= m / 4) return fail(MPB_E_INVALID, "n-m+1 = %d does not exceed the exclusion zone m/4 = %d: no valid pair", n - m + 1, m / 4);
    if (m > 14336) return fail(MPB_E_INVALID, "m = %d exceeds 14336: the precompute kernels stage 2 m doubles per CTA in shared memory (227 KB)", m);
    CU(cudaSetDevice(c->device));
    mpb_plan *p = new (std::nothrow) mpb_plan();
    if (!p) return fail(MPB_E_NOMEM, "out of host memory");
    reg_add(g_plans, p);              // from here on mpb_plan_destroy(p) releases whatever was allocated
    if (flags & MPB_FLAG_STRICT) flags |= MPB_FLAG_SEQ_MEAN | MPB_FLAG_NO_RESEED;
    p->ctx = c; p->n = n; p->m = m; p->dtype = dtype; p->flags = flags;
    // fp64 default: the reference's running mean and its unbroken single-seed chain (recovered by telescoping over the
    // re-seeded segments); fp32 default: exact window means and plain re-seeding, which is what bounds ITS drift
    p->seq_mean = (flags & MPB_FLAG_SEQ_MEAN) || (dtype == MPB_F64 && !(flags & MPB_FLAG_WINDOW_MEAN));
    p->chain_fix = dtype == MPB_F64 && !(flags & (MPB_FLAG_STRICT | MPB_FLAG_NO_RESEED | MPB_FLAG_NO_CHAIN_CARRY));
    // The corrections grow with the series length (measured bound, random walk, m = 256: 4e-11 / 7e-10 / 9e-9 / 6e-8 at
    // n = 2^16 / 2^18 / 2^20 / 2^22) and the margin must be 128 x the bound (chain_check); a wide margin only costs log
    // entries on very smooth series (m in the thousands), so it is not simply set to the largest value.
    p->margin_log2 = MPB_CHAIN_MARGIN_LOG2 ? MPB_CHAIN_MARGIN_LOG2 : (n <= (1 << 20) ? 17 : n <= (1 << 22) ? 15 : n <= (1 << 24) ? 13 : 11);
    // short windows: a nearly flat window has a tiny norm, and the bound pairs the largest inverse norms of whole row / column blocks
    // (measured bound at m = 32, n = 2^22, random walk: 2.4e-6, which needs 2^-11)
    if (!MPB_CHAIN_MARGIN_LOG2) p->margin_log2 = std::max(9, p->margin_log2 - (m < 64 ? 4 : m < 128 ? 2 : 0));
    p->sublen = n - m + 1; p->excl = m / 4; p->np = mpb::padded_len(p->sublen);
    const size_t es = dtype == MPB_F64 ? 8 : 4, np = (size_t)p->np;
#define ALLOC(ptr, bytes)                                                                            \
    do {                                                                                             \
        cudaError_t e__ = cudaMalloc((void **)&(ptr), (bytes));                                      \
        if (e__ != cudaSuccess) { mpb_plan_destroy(p); return fail(MPB_E_NOMEM, "cudaMalloc(%zu) for " #ptr ": %s", (size_t)(bytes), cudaGetErrorString(e__)); } \
    } while (0)
    ALLOC(p->T, sizeof(double) * (size_t)n);
    ALLOC(p->mu, sizeof(double) * (size_t)p->sublen);
    ALLOC(p->df_dg, 2 * es * np);
    ALLOC(p->inv, es * np);
    ALLOC(p->qt0, es * np);
    ALLOC(p->nm, 2 * es * np);
    ALLOC(p->mnc, es * np);
    ALLOC(p->prof, sizeof(mpb::ProfileEntry) * np);
    ALLOC(p->MP, sizeof(double) * (size_t)p->sublen);
    ALLOC(p->corr, sizeof(double) * (size_t)p->sublen);
    ALLOC(p->MPI, sizeof(int32_t) * (size_t)p->sublen);
    ALLOC(p->slow_counter, 40 * sizeof(unsigned long long));
    ALLOC(p->hint, sizeof(int) * np);
    ALLOC(p->thr, sizeof(unsigned long long) * np);
    p->share.thr[0] = p->thr; p->share.n = 1;
    if (p->seq_mean && !(flags & MPB_FLAG_MU_WALK)) ALLOC(p->mu_scratch, mpb::mu_scan_scratch_bytes(n, m));
    if (p->chain_fix) {
        ALLOC(p->fin, sizeof(mpb::ProfileEntry) * np);
        ALLOC(p->log_count, sizeof(unsigned long long) * 16 * mpb::LOG_SHARDS);
        ALLOC(p->chain_stats, 4 * sizeof(unsigned long long));
    }
#undef ALLOC
    // zero padding beyond sublen is what lets tiles run off the matrix edge without bounds checks
    cudaError_t ie = cudaMemsetAsync(p->df_dg, 0, 2 * es * np, c->stream);
    if (ie == cudaSuccess) ie = cudaMemsetAsync(p->inv, 0, es * np, c->stream);
    if (ie == cudaSuccess) ie = cudaMemsetAsync(p->qt0, 0, es * np, c->stream);
    if (ie == cudaSuccess) ie = cudaMemsetAsync(p->slow_counter, 0, 40 * sizeof(unsigned long long), c->stream);
    if (ie == cudaSuccess && p->chain_fix) ie = cudaMemsetAsync(p->chain_stats, 0, 4 * sizeof(unsigned long long), c->stream);
    for (cudaEvent_t &e : p->ev) if (ie == cudaSuccess) ie = cudaEventCreate(&e);
    if (ie != cudaSuccess) { mpb_plan_destroy(p); return fail(MPB_E_CUDA, "initialising the plan's buffers / events: %s", cudaGetErrorString(ie)); }
    p->tm.cells = mpb_useful_cells(n, m);
    *out = p;
    return MPB_OK;
}

int mpb_plan_set_shard(mpb_plan *p, int rank, int world) {
    if (!p) return fail(MPB_E_INVALID, "plan is null");
    if (world < 1 || rank < 0 || rank >= world) return fail(MPB_E_INVALID, "bad shard %d/%d", rank, world);
    p->rank = rank; p->world = world; p->table_dirty = true;
    return MPB_OK;
}

int mpb_plan_set_segment_rows(mpb_plan *p, int rows) {
    if (!p) return fail(MPB_E_INVALID, "plan is null");
    if (rows > 0 && rows % mpb::TILE_H != 0) return fail(MPB_E_INVALID, "segment rows must be a multiple of %d (got %d)", mpb::TILE_H, rows);
    p->seg_rows = rows < 0 ? -1 : rows; p->table_dirty = true;
    return MPB_OK;
}

int mpb_plan_segment_rows(mpb_plan *p) { return p ? p->seg_rows_eff : -1; }

int mpb_plan_set_sketch_bands(mpb_plan *p, int bands) {
    if (!p) return fail(MPB_E_INVALID, "plan is null");
    p->sketch_bands = bands < 0 ? -1 : bands;
    return MPB_OK;
}

// (band, segment) work-item table of this rank: rebuilt when the shard or the segment length changes.
static int plan_build_table(mpb_plan *p) {
    if (!p->table_dirty) return MPB_OK;
    int seg = 0;
    if (!(p->flags & MPB_FLAG_NO_RESEED)) seg = p->seg_rows >= 0 ? p->seg_rows : mpb::default_seg_rows(p->sublen, p->excl, p->m, 8 /* nominal rank count: keeps results independent of the GPU count */, p->ctx->sm_count, p->chain_fix ? MPB_SEG_BASE_CARRY : MPB_SEG_BASE);
    std::vector<int> tab;
    const long long items = mpb::build_item_table(p->sublen, p->excl, p->rank, p->world, seg, tab);
    if (items > 0x7fffffffLL) return fail(MPB_E_INVALID, "%lld work items exceed the grid limit; raise the segment length", items);
    CU(cudaStreamSynchronize(p->ctx->stream));
    if (p->item_start) { cudaFree(p->item_start); p->item_start = nullptr; }
    CU(cudaMalloc((void **)&p->item_start, sizeof(int) * tab.size()));
    CU(cudaMemcpy(p->item_start, tab.data(), sizeof(int) * tab.size(), cudaMemcpyHostToDevice));
    p->table_host = tab;
    p->nlb = ((int)tab.size() - 1) / 2; p->items = items; p->seg_rows_eff = seg; p->table_dirty = false;
    {   // cells the tile kernel walks for this rank: every warp (WARP_DIAGS diagonals) runs whole tiles of TILE_H rows down its longest diagonal
        long long walked = 0, seeded = 0;
        for (int lb = 0; lb < p->nlb; ++lb)
            for (int w = 0; w < mpb::BAND_WARPS; ++w) {
                const long long rows = (long long)p->sublen - (p->excl + (long long)tab[(size_t)p->nlb + 1 + lb] * mpb::BAND_W + w * mpb::WARP_DIAGS);
                if (rows <= 0) continue;
                walked += (rows + mpb::TILE_H - 1) / mpb::TILE_H * mpb::TILE_H * mpb::WARP_DIAGS;
                if (seg > 0) seeded += ((rows + seg - 1) / seg - 1) * mpb::WARP_DIAGS;
            }
        p->tm.cells_walked = walked; p->tm.reseeded_diagonals = seeded;
    }
    if (p->chain_fix && seg == 0) p->chain_fix = false;          // segment length forced to 0: full-length chains, nothing to carry
    if (p->chain_fix) {
        // S_k / E_k arrays: one launch chunk at a time (whole bands; at least the longest band), at most MPB_CHAIN_CHUNK_ITEMS items
        long long longest = 0;
        for (int lb = 0; lb < p->nlb; ++lb) longest = std::max(longest, (long long)tab[lb + 1] - tab[lb]);
        const long long want = std::max(longest, std::min(items, (long long)MPB_CHAIN_CHUNK_ITEMS));
        if (want > p->chunk_items) {
            if (p->seg_seed) cudaFree(p->seg_seed);
            if (p->seg_end) cudaFree(p->seg_end);
            p->seg_seed = p->seg_end = nullptr; p->chunk_items = 0;
            const size_t bytes = sizeof(double) * (size_t)want * mpb::BAND_W;
            if (cudaMalloc((void **)&p->seg_seed, bytes) != cudaSuccess || cudaMalloc((void **)&p->seg_end, bytes) != cudaSuccess) {
                cudaGetLastError();
                return fail(MPB_E_NOMEM, "cudaMalloc(2 x %zu) for the chain-carry segment arrays", bytes);
            }
            p->chunk_items = want;
        }
        if (!p->chain_log) {
            long long ent = p->log_entries_req > 0 ? p->log_entries_req : std::max(4LL << 20, 64LL * p->sublen);
            ent = (ent + mpb::LOG_SHARDS - 1) / mpb::LOG_SHARDS;
            if (ent > 0x7fffffffLL) ent = 0x7fffffffLL;
            const size_t bytes = sizeof(mpb::ChainLogEntry) * (size_t)ent * mpb::LOG_SHARDS;
            if (cudaMalloc((void **)&p->chain_log, bytes) != cudaSuccess) { cudaGetLastError(); return fail(MPB_E_NOMEM, "cudaMalloc(%zu) for the chain-carry candidate log", bytes); }
            p->log_shard_cap = (unsigned)ent;
        }
        const int blocks = (p->sublen + seg - 1) / seg;
        if (blocks > p->inv_blocks) {
            if (p->inv_blockmax) cudaFree(p->inv_blockmax);
            p->inv_blockmax = nullptr; p->inv_blocks = 0;
            if (cudaMalloc((void **)&p->inv_blockmax, sizeof(double) * (size_t)blocks) != cudaSuccess) { cudaGetLastError(); return fail(MPB_E_NOMEM, "cudaMalloc for the inverse-norm block maxima"); }
            p->inv_blocks = blocks;
        }
    }
    return MPB_OK;
}

int mpb_plan_upload_series(mpb_plan *p, const double *T_host) {
    if (!p || !T_host) return fail(MPB_E_INVALID, "null argument");
    CU(cudaSetDevice(p->ctx->device));
    CU(cudaEventRecord(p->ev[0], p->ctx->stream));
    CU(cudaMemcpyAsync(p->T, T_host, sizeof(double) * (size_t)p->n, cudaMemcpyHostToDevice, p->ctx->stream));
    CU(cudaEventRecord(p->ev[1], p->ctx->stream));
    p->ran_h2d = true;
    return MPB_OK;
}

void *mpb_plan_series_device_ptr(mpb_plan *p) { return p ? p->T : nullptr; }

int mpb_plan_precompute(mpb_plan *p) {
    if (!p) return fail(MPB_E_INVALID, "plan is null");
    cudaStream_t s = p->ctx->stream;
    CU(cudaSetDevice(p->ctx->device));
    int rc = plan_build_table(p);
    if (rc != MPB_OK) return rc;
    CU(cudaEventRecord(p->ev[2], s));
    if (p->seq_mean) CU(mpb::launch_mu_scan(p->T, p->n, p->m, p->mu, p->mu_scratch, (p->flags & MPB_FLAG_MU_WALK) ? 0 : 1, s));
    else CU(mpb::launch_mu_window(p->T, p->n, p->m, p->mu, s));
    if (p->dtype == MPB_F64) CU(mpb::launch_stats_f64(p->T, p->mu, p->n, p->m, (double *)p->df_dg, (double *)p->inv, (double *)p->qt0, s));
    else CU(mpb::launch_stats_f32(p->T, p->mu, p->n, p->m, (float *)p->df_dg, (float *)p->inv, (float *)p->qt0, s));
    if (p->dtype == MPB_F64) CU(mpb::launch_norm_f64((const double *)p->inv, p->sublen, p->np, (double *)p->nm, (double *)p->mnc, s));
    else CU(mpb::launch_norm_f32((const float *)p->inv, p->sublen, p->np, (float *)p->nm, (float *)p->mnc, s));
    CU(mpb::launch_profile_init(p->prof, p->np, s));
    CU(cudaMemsetAsync(p->thr, 0, sizeof(unsigned long long) * (size_t)p->np, s));
    p->tm.launches = 4;
    if (p->chain_fix) {
        CU(mpb::launch_profile_init(p->fin, p->np, s));
        CU(cudaMemsetAsync(p->chain_stats, 0, 4 * sizeof(unsigned long long), s));
        CU(mpb::launch_inv_block_max((const double *)p->inv, p->sublen, p->seg_rows_eff, p->inv_blockmax, s));
        p->tm.launches += 2;
    }
    if (!(p->flags & (MPB_FLAG_STRICT | MPB_FLAG_NO_SEED))) {
        // the seed is a borrowed lower bound in the thresholds, never a result: lowered by what separates its direct dot product from
        // the chain's value of the same cell (rounding; in the chain-carry mode four near-tie margins)
        const double rel = p->dtype == MPB_F64 ? std::ldexp(1.0, -20) : std::ldexp(1.0, -12);
        const double mu_abs = p->chain_fix ? 4.0 * std::ldexp(1.0, -p->margin_log2) : 0.0;
        if (p->dtype == MPB_F64) CU(mpb::launch_seed_thresholds_f64(p->T, p->mu, (const double *)p->inv, p->n, p->m, p->excl, p->thr, rel, mu_abs, s));
        else CU(mpb::launch_seed_thresholds_f32(p->T, p->mu, (const float *)p->inv, p->n, p->m, p->excl, p->thr, rel, mu_abs, s));
        p->tm.launches += 1;
    }
    CU(cudaEventRecord(p->ev[3], s));
    p->step_open = false; p->sketched = false;
    return MPB_OK;
}

int mpb_plan_local_bands(mpb_plan *p, int *count) {
    if (!p || !count) return fail(MPB_E_INVALID, "null argument");
    CU(cudaSetDevice(p->ctx->device));
    int rc = plan_build_table(p);
    if (rc != MPB_OK) return rc;
    *count = p->nlb;
    return MPB_OK;
}

// first launch of a step: statistics and the compute timer start here
static int plan_step_begin(mpb_plan *p) {
    if (p->step_open) return MPB_OK;
    cudaStream_t s = p->ctx->stream;
    CU(cudaMemsetAsync(p->slow_counter, 0, 40 * sizeof(unsigned long long), s));
    CU(cudaEventRecord(p->ev[4], s));
    p->tm.bands = p->nlb; p->tm.items = (int32_t)p->items; p->tm.segment_rows = p->seg_rows_eff;
    p->step_open = true;
    return MPB_OK;
}
static bool plan_sketch_allowed(const mpb_plan *p) { return !(p->flags & (MPB_FLAG_STRICT | MPB_FLAG_NO_SEED | MPB_FLAG_NO_RESEED)); }

int mpb_plan_total_bands(mpb_plan *p, int *count) {
    if (!p || !count) return fail(MPB_E_INVALID, "null argument");
    *count = mpb::band_count(p->sublen, p->excl);
    return MPB_OK;
}

// One-band sketch shared by the ranks of a multi-GPU job: this rank walks segments share_rank, share_rank + share_world, ... of
// global band `band` and leaves the high words in hint[]; the caller max-reduces hint[] across the ranks, then mpb_plan_apply_hint.
int mpb_plan_sketch_share(mpb_plan *p, int band, int share_rank, int share_world) {
    if (!p) return fail(MPB_E_INVALID, "plan is null");
    if (share_world < 1 || share_rank < 0 || share_rank >= share_world || band < 0 || band >= mpb::band_count(p->sublen, p->excl)) return fail(MPB_E_INVALID, "bad sketch share");
    if (p->dtype != MPB_F64 || !plan_sketch_allowed(p)) return fail(MPB_E_INVALID, "the shared sketch is for the fp64 re-seeded modes");
    cudaStream_t s = p->ctx->stream;
    CU(cudaSetDevice(p->ctx->device));
    int rc = plan_build_table(p);
    if (rc == MPB_OK) rc = plan_step_begin(p);
    if (rc != MPB_OK) return rc;
    mpb::TileLaunch sl;
    sl.n = p->n; sl.m = p->m; sl.sublen = p->sublen; sl.excl = p->excl; sl.rank = p->rank; sl.world = p->world;
    sl.seg_rows = p->seg_rows_eff; sl.strict = false; sl.series = p->T; sl.mu = p->mu; sl.item_start = p->item_start; sl.nlb = p->nlb;
    sl.item_begin = 0; sl.items = 1; sl.counters = nullptr;
    sl.sketch_band = band; sl.sketch_seg_rows = MPB_SKETCH_SEG_ROWS; sl.sketch_share_rank = share_rank; sl.sketch_share_world = share_world;
    sl.share = p->share;
    CU(mpb::launch_sketch_f64((const double *)p->df_dg, (const double *)p->nm, (const double *)p->qt0, p->hint, sl, false, s));
    p->tm.launches += 1;
    p->sketched = true;
    return MPB_OK;
}
void *mpb_plan_hint_device_ptr(mpb_plan *p) { return p ? p->hint : nullptr; }
int mpb_plan_hint_count(mpb_plan *p) { return p ? p->np : 0; }
int mpb_plan_apply_hint(mpb_plan *p) {
    if (!p) return fail(MPB_E_INVALID, "plan is null");
    CU(cudaSetDevice(p->ctx->device));
    CU(mpb::launch_apply_hint(p->hint, p->thr, p->sublen, p->ctx->stream));
    p->tm.launches += 1;
    return MPB_OK;
}

int mpb_plan_compute_bands(mpb_plan *p, int lb_begin, int lb_end) {
    if (!p) return fail(MPB_E_INVALID, "plan is null");
    cudaStream_t s = p->ctx->stream;
    CU(cudaSetDevice(p->ctx->device));
    int rc = plan_build_table(p);
    if (rc != MPB_OK) return rc;
    if (lb_end < 0 || lb_end > p->nlb) lb_end = p->nlb;
    if (lb_begin < 0 || lb_begin > lb_end) return fail(MPB_E_INVALID, "bad local band range [%d, %d) of %d", lb_begin, lb_end, p->nlb);
    mpb::TileLaunch tl;
    tl.n = p->n; tl.m = p->m; tl.sublen = p->sublen; tl.excl = p->excl; tl.rank = p->rank; tl.world = p->world;
    tl.seg_rows = p->seg_rows_eff;
    tl.strict = (p->flags & MPB_FLAG_STRICT) != 0;
    tl.series = p->T; tl.mu = p->mu;
    tl.item_start = p->item_start; tl.nlb = p->nlb;
    tl.item_begin = p->nlb > 0 ? p->table_host[lb_begin] : 0;
    tl.items = p->nlb > 0 ? (long long)p->table_host[lb_end] - tl.item_begin : 0;
    tl.counters = p->slow_counter;
    tl.share = p->share;
    if (lb_begin == 0) {                      // first launch of a step: statistics and the compute timer start here
        rc = plan_step_begin(p);
        if (rc != MPB_OK) return rc;
        // Sketch pass over the first bands of this call (scamp_tile.cu): lower bounds for every slot, so that the tile
        // kernel meets a tight filter in the bands where the profile would still be empty.  Result-neutral; never in
        // the chain-faithful modes (their seg_rows = 0 makes a band a single item anyway).
        int sk = p->sketch_bands < 0 ? MPB_DEFAULT_SKETCH_BANDS : p->sketch_bands;
        if (p->sketched || !plan_sketch_allowed(p)) sk = 0;
        if (sk > lb_end) sk = lb_end;
        if (sk > 0 && p->nlb > 0) {
            mpb::TileLaunch sl = tl;
            sl.item_begin = p->table_host[0];
            sl.items = (long long)p->table_host[sk] - sl.item_begin;
            // one band: cut it finer than the tile kernel does, so that it fills the GPU.  fp64 only: a bound from a chain
            // with other re-seed points is safe when the chains agree far better than the 2e-6 by which the bounds are
            // lowered (fp64: ~1e-12); fp32 chains differ by up to 1e-5 and keep the tile kernel's own cut (bit-equal QT).
            if (sk == 1 && p->dtype == MPB_F64 && p->seg_rows_eff > MPB_SKETCH_SEG_ROWS) {
                sl.sketch_band = p->table_host[(size_t)p->nlb + 1];
                sl.sketch_seg_rows = MPB_SKETCH_SEG_ROWS;
            }
            if (p->dtype == MPB_F64) CU(mpb::launch_sketch_f64((const double *)p->df_dg, (const double *)p->nm, (const double *)p->qt0, p->hint, sl, true, s));
            else CU(mpb::launch_sketch_f32((const float *)p->df_dg, (const float *)p->nm, (const float *)p->qt0, p->hint, sl, true, s));
            p->tm.launches += sl.items > 0 ? 2 : 0;
            p->sketched = true;
        }
    }
    if (p->chain_fix && tl.items > 0) {
        // chain-carry mode: launch chunks of whole bands that fit the S_k / E_k arrays; each chunk = tile kernel (candidates into
        // the log) + carry scan + resolve into the result profile
        int lb = lb_begin;
        while (lb < lb_end) {
            int hi = lb + 1;
            while (hi < lb_end && (long long)p->table_host[hi + 1] - p->table_host[lb] <= p->chunk_items) ++hi;
            mpb::TileLaunch cl = tl;
            cl.item_begin = p->table_host[lb];
            cl.items = (long long)p->table_host[hi] - cl.item_begin;
            cl.chain.seg_seed = p->seg_seed; cl.chain.seg_end = p->seg_end; cl.chain.log = p->chain_log; cl.chain.log_count = p->log_count;
            cl.chain.shard_cap = p->log_shard_cap; cl.chain.item_base = cl.item_begin;
            const double margin = std::ldexp(1.0, -p->margin_log2);
            cl.chain.margin = margin;
            cl.chain.key_margin = (1 << std::max(0, 21 - p->margin_log2)) + 1;
            CU(cudaMemsetAsync(p->log_count, 0, sizeof(unsigned long long) * 16 * mpb::LOG_SHARDS, s));
            CU(mpb::launch_tiles_f64((const double *)p->df_dg, (const double *)p->nm, (const double *)p->mnc, (const double *)p->qt0, p->prof, cl, s));
            CU(mpb::launch_chain_resolve(cl, lb, hi, (const double *)p->inv, p->inv_blockmax, p->prof, p->fin, p->chain_stats, s));
            p->tm.launches += cl.items > 0 ? 3 : 0;
            lb = hi;
        }
    } else {
        if (p->dtype == MPB_F64) CU(mpb::launch_tiles_f64((const double *)p->df_dg, (const double *)p->nm, (const double *)p->mnc, (const double *)p->qt0, p->prof, tl, s));
        else CU(mpb::launch_tiles_f32((const float *)p->df_dg, (const float *)p->nm, (const float *)p->mnc, (const float *)p->qt0, p->prof, tl, s));
        p->tm.launches += tl.items > 0 ? 1 : 0;
    }
    CU(cudaEventRecord(p->ev[5], s));
    return MPB_OK;
}

int mpb_plan_compute(mpb_plan *p) { return mpb_plan_compute_bands(p, 0, -1); }

// the profile export / mask_index / import / merge_peers address: the result
static mpb::ProfileEntry *merge_profile(mpb_plan *p) { return p->chain_fix ? p->fin : p->prof; }

// Threshold sharing inside one process: improvements this plan's tile kernel makes are also pushed to the threshold arrays of
// `others` (plans of the same problem on this or on peer-accessible devices).  One direction: call it on every plan of the job.
int mpb_plan_share_thresholds(mpb_plan *p, mpb_plan *const *others, int count) {
    if (!p || (count > 0 && !others)) return fail(MPB_E_INVALID, "null argument");
    if (count + 1 > mpb::MAX_PEERS) return fail(MPB_E_INVALID, "at most %d plans can share thresholds", mpb::MAX_PEERS);
    CU(cudaSetDevice(p->ctx->device));
    p->share.n = 1;
    for (int k = 0; k < count; ++k) {
        mpb_plan *o = others[k];
        if (!o || o == p || o->sublen != p->sublen) return fail(MPB_E_INVALID, "peer plan %d does not match", k);
        if (o->ctx->device != p->ctx->device) {
            int can = 0;
            CU(cudaDeviceCanAccessPeer(&can, p->ctx->device, o->ctx->device));
            if (!can) return fail(MPB_E_CUDA, "device %d cannot access device %d over P2P", p->ctx->device, o->ctx->device);
            cudaError_t e = cudaDeviceEnablePeerAccess(o->ctx->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(MPB_E_CUDA, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
            cudaGetLastError();
        }
        p->share.thr[p->share.n++] = o->thr;
    }
    return MPB_OK;
}
int mpb_plan_shared_with(mpb_plan *p) { return p ? p->share.n - 1 : 0; }

int mpb_plan_set_chain_margin(mpb_plan *p, int log2_inverse) {
    if (!p) return fail(MPB_E_INVALID, "plan is null");
    if (log2_inverse < 8 || log2_inverse > 21) return fail(MPB_E_INVALID, "the chain-carry margin is 2^-k with 8 <= k <= 21 (got %d)", log2_inverse);
    p->margin_log2 = log2_inverse;
    return MPB_OK;
}

int mpb_plan_set_log_entries(mpb_plan *p, int64_t entries) {
    if (!p) return fail(MPB_E_INVALID, "plan is null");
    if (p->chain_log) { CU(cudaSetDevice(p->ctx->device)); CU(cudaStreamSynchronize(p->ctx->stream)); cudaFree(p->chain_log); p->chain_log = nullptr; }
    p->log_entries_req = entries > 0 ? entries : -1; p->table_dirty = true;
    return MPB_OK;
}

int mpb_plan_export(mpb_plan *p, void *corr_out, void *index_out) {
    if (!p || !corr_out || !index_out) return fail(MPB_E_INVALID, "null argument");
    CU(cudaSetDevice(p->ctx->device));
    CU(mpb::launch_export(merge_profile(p), p->sublen, (double *)corr_out, (int32_t *)index_out, p->ctx->stream));
    p->tm.launches += 1;
    return MPB_OK;
}

int mpb_plan_mask_index(mpb_plan *p, const void *corr_max, void *index) {
    if (!p || !corr_max || !index) return fail(MPB_E_INVALID, "null argument");
    CU(cudaSetDevice(p->ctx->device));
    CU(mpb::launch_mask_index(merge_profile(p), p->sublen, (const double *)corr_max, (int32_t *)index, p->ctx->stream));
    p->tm.launches += 1;
    return MPB_OK;
}

int mpb_plan_import(mpb_plan *p, const void *corr, const void *index) {
    if (!p || !corr || !index) return fail(MPB_E_INVALID, "null argument");
    CU(cudaSetDevice(p->ctx->device));
    CU(mpb::launch_import(merge_profile(p), p->sublen, (const double *)corr, (const int32_t *)index, p->thr, p->ctx->stream));
    p->tm.launches += 1;
    return MPB_OK;
}

int mpb_plan_merge_peers(mpb_plan *p, mpb_plan *const *others, int count) {
    if (!p || (count > 0 && !others)) return fail(MPB_E_INVALID, "null argument");
    CU(cudaSetDevice(p->ctx->device));
    for (int k = 0; k < count; ++k) {
        mpb_plan *o = others[k];
        if (!o || o->sublen != p->sublen) return fail(MPB_E_INVALID, "peer plan %d does not match", k);
        if (o->ctx->device != p->ctx->device) {
            int can = 0;
            CU(cudaDeviceCanAccessPeer(&can, p->ctx->device, o->ctx->device));
            if (!can) return fail(MPB_E_CUDA, "device %d cannot access device %d over P2P", p->ctx->device, o->ctx->device);
            cudaError_t e = cudaDeviceEnablePeerAccess(o->ctx->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(MPB_E_CUDA, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
            cudaGetLastError();
        }
        // the peer's band kernel must be complete before its slots are read
        CU(cudaSetDevice(o->ctx->device));
        CU(cudaEventRecord(o->ev[7], o->ctx->stream));
        CU(cudaSetDevice(p->ctx->device));
        CU(cudaStreamWaitEvent(p->ctx->stream, o->ev[7], 0));
        if (p->chain_fix != o->chain_fix) return fail(MPB_E_INVALID, "peer plan %d runs in a different mode", k);
        CU(mpb::launch_merge_peer(merge_profile(p), merge_profile(o), p->sublen, p->thr, p->ctx->stream));
        p->tm.launches += 1;
    }
    return MPB_OK;
}

int mpb_plan_finalize(mpb_plan *p) {
    if (!p) return fail(MPB_E_INVALID, "plan is null");
    CU(cudaSetDevice(p->ctx->device));
    CU(mpb::launch_finalize(p->chain_fix ? p->fin : p->prof, p->sublen, p->m, p->MP, p->MPI, p->corr, p->ctx->stream));
    CU(cudaEventRecord(p->ev[6], p->ctx->stream));
    p->tm.launches += 1;
    return MPB_OK;
}

// chain-carry mode: a candidate log that overflowed dropped candidates -- the result may miss a maximum, so that is an error
static int chain_check(mpb_plan *p) {
    if (!p->chain_fix) return MPB_OK;
    unsigned long long st[4] = {0, 0, 0, 0};
    CU(cudaMemcpy(st, p->chain_stats, sizeof st, cudaMemcpyDeviceToHost));
    if (st[2]) return fail(MPB_E_NOMEM, "the chain-carry near-tie log overflowed (%u entries per sub-log): raise it with mpb_plan_set_log_entries", p->log_shard_cap);
    double bound; memcpy(&bound, &st[3], sizeof bound);
    const double margin = std::ldexp(1.0, -p->margin_log2);
    if (2.0 * bound > margin * MPB_CHAIN_FLOOR)      // two candidates further apart than the margin could have changed order
        return fail(MPB_E_MARGIN, "chain-carry corrections are bounded by %.3g, too large for the near-tie margin 2^-%d (needs 2 * bound <= margin / 64): "
                                  "widen it with mpb_plan_set_chain_margin, or use MPB_FLAG_NO_RESEED", bound, p->margin_log2);
    return MPB_OK;
}

int mpb_plan_check(mpb_plan *p) {
    if (!p) return fail(MPB_E_INVALID, "plan is null");
    CU(cudaSetDevice(p->ctx->device));
    CU(cudaStreamSynchronize(p->ctx->stream));
    return chain_check(p);
}

int mpb_plan_download(mpb_plan *p, double *MP_host, int32_t *MPI_host, double *corr_host) {
    if (!p || !MP_host || !MPI_host) return fail(MPB_E_INVALID, "null argument");
    cudaStream_t s = p->ctx->stream;
    CU(cudaSetDevice(p->ctx->device));
    CU(cudaMemcpyAsync(MP_host, p->MP, sizeof(double) * (size_t)p->sublen, cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(MPI_host, p->MPI, sizeof(int32_t) * (size_t)p->sublen, cudaMemcpyDeviceToHost, s));
    if (corr_host) CU(cudaMemcpyAsync(corr_host, p->corr, sizeof(double) * (size_t)p->sublen, cudaMemcpyDeviceToHost, s));
    CU(cudaEventRecord(p->ev[7], s));
    CU(cudaStreamSynchronize(s));
    return chain_check(p);
}

void *mpb_plan_mp_device_ptr(mpb_plan *p) { return p ? p->MP : nullptr; }
void *mpb_plan_mpi_device_ptr(mpb_plan *p) { return p ? p->MPI : nullptr; }

int mpb_plan_download_input(mpb_plan *p, void *input_host) {
    if (!p || !input_host) return fail(MPB_E_INVALID, "null argument");
    cudaStream_t s = p->ctx->stream;
    CU(cudaSetDevice(p->ctx->device));
    const size_t bytes = (size_t)p->sublen * (p->dtype == MPB_F64 ? 32 : 16);
    void *tmp = nullptr;
    CU(cudaMalloc(&tmp, bytes));
    cudaError_t e = p->dtype == MPB_F64
        ? mpb::launch_pack_input_f64((const double *)p->df_dg, (const double *)p->inv, (const double *)p->qt0, p->sublen, tmp, s)
        : mpb::launch_pack_input_f32((const float *)p->df_dg, (const float *)p->inv, (const float *)p->qt0, p->sublen, tmp, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(input_host, tmp, bytes, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFree(tmp);
    CU(e);
    return MPB_OK;
}

int mpb_plan_mu_walked(mpb_plan *p, int *walked) {
    if (!p || !walked) return fail(MPB_E_INVALID, "null argument");
    *walked = 0;
    if (!p->seq_mean) return MPB_OK;
    if (!p->mu_scratch) { *walked = 1; return MPB_OK; }
    CU(cudaSetDevice(p->ctx->device));
    CU(cudaStreamSynchronize(p->ctx->stream));
    int bad = 0;
    CU(cudaMemcpy(&bad, (const char *)p->mu_scratch + mpb::mu_scan_bad_offset(p->n, p->m), sizeof bad, cudaMemcpyDeviceToHost));
    *walked = bad != 0;
    return MPB_OK;
}

// development aid (builds with -DMPB_REPLAY_HIST): hist[k] = replayed blocks in which k lanes passed the block filter, k = 0..32
int mpb_plan_replay_histogram(mpb_plan *p, int64_t *hist33) {
    if (!p || !hist33) return fail(MPB_E_INVALID, "null argument");
    CU(cudaSetDevice(p->ctx->device));
    CU(cudaStreamSynchronize(p->ctx->stream));
    unsigned long long h[40];
    CU(cudaMemcpy(h, p->slow_counter, sizeof h, cudaMemcpyDeviceToHost));
    for (int k = 0; k <= 32; ++k) hist33[k] = (int64_t)h[4 + k];
    return MPB_OK;
}

int mpb_plan_timings(mpb_plan *p, mpb_timings *out) {
    if (!p || !out) return fail(MPB_E_INVALID, "null argument");
    CU(cudaSetDevice(p->ctx->device));
    CU(cudaStreamSynchronize(p->ctx->stream));
    auto span = [&](int a, int b) -> double {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, p->ev[a], p->ev[b]) != cudaSuccess) { cudaGetLastError(); return 0.0; }
        return (double)ms;
    };
    mpb_timings t = p->tm;
    t.h2d_ms = p->ran_h2d ? span(0, 1) : 0.0;
    t.precompute_ms = span(2, 3);
    t.compute_ms = span(4, 5);
    t.finalize_ms = span(5, 6);
    t.d2h_ms = span(6, 7);
    t.total_ms = span(p->ran_h2d ? 0 : 2, 7);
    unsigned long long cnt[4] = {0, 0, 0, 0};
    CU(cudaMemcpy(cnt, p->slow_counter, sizeof cnt, cudaMemcpyDeviceToHost));
    t.filter_hits = (int64_t)cnt[0];
    t.slow_updates = (int64_t)cnt[1];
    t.replays = (int64_t)cnt[2];
    t.replay_lane_blocks = (int64_t)cnt[3];
    t.chain_candidates = 0; t.chain_max_correction = 0.0; t.chain_bound = 0.0; t.chain_margin = 0.0;
    if (p->chain_fix) {
        unsigned long long st[4] = {0, 0, 0, 0};
        CU(cudaMemcpy(st, p->chain_stats, sizeof st, cudaMemcpyDeviceToHost));
        memcpy(&t.chain_max_correction, &st[0], sizeof(double));
        memcpy(&t.chain_bound, &st[3], sizeof(double));
        t.chain_candidates = (int64_t)st[1];
        t.chain_margin = std::ldexp(1.0, -p->margin_log2);
    }
    *out = t;
    return MPB_OK;
}

int mpb_compute(mpb_context *c, const double *T, int n, int m, int dtype, unsigned flags, double *MP, int32_t *MPI, mpb_timings *timings) {
    if (!c || !T || !MP || !MPI) return fail(MPB_E_INVALID, "null argument");
    int rc = MPB_OK, margin = 0;
    for (int attempt = 0; attempt < 4; ++attempt) {       // MPB_E_MARGIN: the same problem again with a 4x wider near-tie margin
        mpb_plan *p = nullptr;
        rc = mpb_plan_create(c, n, m, dtype, flags, &p);
        if (rc == MPB_OK && margin) rc = mpb_plan_set_chain_margin(p, margin);
        if (rc == MPB_OK) rc = mpb_plan_upload_series(p, T);
        if (rc == MPB_OK) rc = mpb_plan_precompute(p);
        if (rc == MPB_OK) rc = mpb_plan_compute(p);
        if (rc == MPB_OK) rc = mpb_plan_finalize(p);
        if (rc == MPB_OK) rc = mpb_plan_download(p, MP, MPI, nullptr);
        if (rc == MPB_OK && timings) rc = mpb_plan_timings(p, timings);
        if (rc == MPB_E_MARGIN) margin = std::max(8, p->margin_log2 - 2);
        mpb_plan_destroy(p);
        if (rc != MPB_E_MARGIN) break;
    }
    return rc;
}

int mpb_measure_fma_peak(mpb_context *c, int dtype, int iters, double *fma_per_s, double *ms_out) {
    if (!c || !fma_per_s) return fail(MPB_E_INVALID, "null argument");
    if (iters < 1) return fail(MPB_E_INVALID, "iters must be positive");
    CU(cudaSetDevice(c->device));
    double *sink = nullptr;
    CU(cudaMalloc(&sink, sizeof(double)));
    cudaEvent_t a, b;
    CU(cudaEventCreate(&a));
    CU(cudaEventCreate(&b));
    const int blocks = c->sm_count * 8;
    CU(mpb::launch_fma_peak(dtype, iters / 8 + 1, blocks, sink, c->stream));   // warm-up
    CU(cudaEventRecord(a, c->stream));
    CU(mpb::launch_fma_peak(dtype, iters, blocks, sink, c->stream));
    CU(cudaEventRecord(b, c->stream));
    CU(cudaEventSynchronize(b));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, a, b));
    cudaEventDestroy(a); cudaEventDestroy(b); cudaFree(sink);
    const double fmas = (double)blocks * mpb::fma_peak_threads() * mpb::fma_peak_chains() * (double)iters;
    *fma_per_s = fmas / ((double)ms * 1e-3);
    if (ms_out) *ms_out = ms;
    return MPB_OK;
}

}  // extern "C"
