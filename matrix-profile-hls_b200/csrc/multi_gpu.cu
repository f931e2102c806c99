// multi_gpu.cu -- the multi-GPU step inside the product: NCCL max-reduction of the per-rank profiles over NVLink and the
// phased compute schedule, behind the C ABI (include/mp_b200.h, mpb_comm_* / mpb_plan_compute_merged / mpb_compute_multi).
//
// What it replaces: the reference's replication over MP_KERNEL_REPL compute units -- chunk k on unit k % kNumKernels, then
// the host-side element-wise max merge UpdateAggregates (host/MatrixProfileHost.cpp:101-137,
// include/host/HostSideComputation.hpp:72-90).  Here rank r walks the diagonal bands b = r (mod G) on its own GPU and the
// merge is two ncclAllReduce calls on the device: MAX over the correlations viewed as int64 (non-negative doubles order
// like integers) and MIN over the indices with every rank that does not hold the maximum contributing INT32_MAX -- an
// exact (corr, index) max, independent of rank count and arrival order (ties -> smallest index, the tile kernel's rule).
//
// NCCL is bound at run time (dlopen of libnccl.so.2): a process that already carries NCCL (torch) shares that copy, a
// single-GPU user needs none, and the library itself links only the CUDA runtime.
#include "mp_runtime.h"

#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <condition_variable>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <unistd.h>

namespace {

struct NcclApi {
    void *handle = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    decltype(&ncclGetVersion) GetVersion = nullptr;
    std::string error;
};

NcclApi &nccl() {
    static NcclApi api = [] {
        NcclApi a;
        for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
            a.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (a.handle) break;
        }
        if (!a.handle) { a.error = std::string("libnccl.so.2 not found: ") + (dlerror() ? dlerror() : "?"); return a; }
#define MPB_SYM(field, sym)                                                                        \
    a.field = reinterpret_cast<decltype(a.field)>(dlsym(a.handle, #sym));                          \
    if (!a.field) a.error = "symbol " #sym " missing from libnccl"
        MPB_SYM(GetUniqueId, ncclGetUniqueId);
        MPB_SYM(CommInitRank, ncclCommInitRank);
        MPB_SYM(CommDestroy, ncclCommDestroy);
        MPB_SYM(AllReduce, ncclAllReduce);
        MPB_SYM(AllGather, ncclAllGather);
        MPB_SYM(GetErrorString, ncclGetErrorString);
        MPB_SYM(GetVersion, ncclGetVersion);
#undef MPB_SYM
        return a;
    }();
    return api;
}

#define NC(expr)                                                                                                   \
    do {                                                                                                           \
        ncclResult_t r__ = (expr);                                                                                 \
        if (r__ != ncclSuccess) return fail(MPB_E_CUDA, "%s: %s (%s:%d)", #expr, nccl().GetErrorString(r__), __FILE__, __LINE__); \
    } while (0)

}  // namespace

struct mpb_comm {
    mpb_context *ctx = nullptr;
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    double *corr = nullptr;      // staging of the reduction: sublen doubles + sublen int32
    int32_t *index = nullptr;
    size_t cap = 0;              // subsequences the staging holds
};

extern "C" {

int mpb_comm_unique_id(void *id, size_t bytes) {
    if (!id || bytes < NCCL_UNIQUE_ID_BYTES) return fail(MPB_E_INVALID, "the id buffer must hold %d bytes", NCCL_UNIQUE_ID_BYTES);
    if (!nccl().error.empty()) return fail(MPB_E_CUDA, "NCCL unavailable: %s", nccl().error.c_str());
    ncclUniqueId u;
    NC(nccl().GetUniqueId(&u));
    memcpy(id, &u, NCCL_UNIQUE_ID_BYTES);
    return MPB_OK;
}

int mpb_comm_create(mpb_context *ctx, const void *id, int rank, int world, mpb_comm **out) {
    if (!ctx || !id || !out) return fail(MPB_E_INVALID, "null argument");
    *out = nullptr;
    if (world < 1 || rank < 0 || rank >= world) return fail(MPB_E_INVALID, "bad rank %d of %d", rank, world);
    if (!nccl().error.empty()) return fail(MPB_E_CUDA, "NCCL unavailable: %s", nccl().error.c_str());
    CU(cudaSetDevice(ctx->device));
    mpb_comm *c = new (std::nothrow) mpb_comm();
    if (!c) return fail(MPB_E_NOMEM, "out of host memory");
    c->ctx = ctx; c->rank = rank; c->world = world;
    ncclUniqueId u;
    memcpy(&u, id, NCCL_UNIQUE_ID_BYTES);
    ncclResult_t r = nccl().CommInitRank(&c->comm, world, u, rank);
    if (r != ncclSuccess) { delete c; return fail(MPB_E_CUDA, "ncclCommInitRank(%d of %d): %s", rank, world, nccl().GetErrorString(r)); }
    *out = c;
    return MPB_OK;
}

int mpb_comm_destroy(mpb_comm *c) {
    if (!c) return MPB_OK;
    cudaSetDevice(c->ctx->device);
    cudaStreamSynchronize(c->ctx->stream);
    if (c->corr) cudaFree(c->corr);
    if (c->index) cudaFree(c->index);
    if (c->comm) nccl().CommDestroy(c->comm);
    delete c;
    return MPB_OK;
}

int mpb_comm_rank(mpb_comm *c) { return c ? c->rank : -1; }
int mpb_comm_world(mpb_comm *c) { return c ? c->world : 0; }

// Threshold sharing across the ranks of the communicator (mp_common.cuh ThrShare): every rank maps the threshold arrays of all
// the others -- raw peer pointers inside one process, CUDA-IPC handles across processes, exchanged with one ncclAllGather --
// and from then on pushes every profile improvement to all of them over NVLink.  Collective; one node.
int mpb_plan_connect_peers(mpb_plan *p, mpb_comm *c) {
    if (!p || !c) return fail(MPB_E_INVALID, "null argument");
    if (p->ctx != c->ctx) return fail(MPB_E_INVALID, "plan and communicator belong to different contexts");
    if (c->world > mpb::MAX_PEERS) return fail(MPB_E_INVALID, "threshold sharing supports up to %d ranks", mpb::MAX_PEERS);
    if (c->world == 1) return MPB_OK;
    struct Rec { long long pid; int device; int sublen; unsigned long long ptr; cudaIpcMemHandle_t handle; char pad[96 - 24 - sizeof(cudaIpcMemHandle_t)]; };
    static_assert(sizeof(Rec) == 96, "record size");
    cudaStream_t s = p->ctx->stream;
    CU(cudaSetDevice(p->ctx->device));
    std::vector<Rec> recs((size_t)c->world);
    Rec &me = recs[(size_t)c->rank];
    memset(&me, 0, sizeof me);
    me.pid = (long long)getpid(); me.device = p->ctx->device; me.sublen = p->sublen; me.ptr = (unsigned long long)(uintptr_t)p->thr;
    CU(cudaIpcGetMemHandle(&me.handle, p->thr));
    Rec *dev = nullptr;
    CU(cudaMalloc((void **)&dev, sizeof(Rec) * recs.size()));
    cudaError_t e = cudaMemcpyAsync(dev + c->rank, &me, sizeof me, cudaMemcpyHostToDevice, s);
    ncclResult_t nr = ncclSuccess;
    if (e == cudaSuccess) nr = nccl().AllGather(dev + c->rank, dev, sizeof(Rec), ncclInt8, c->comm, s);
    if (e == cudaSuccess && nr == ncclSuccess) e = cudaMemcpyAsync(recs.data(), dev, sizeof(Rec) * recs.size(), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess && nr == ncclSuccess) e = cudaStreamSynchronize(s);
    cudaFree(dev);
    if (nr != ncclSuccess) return fail(MPB_E_CUDA, "ncclAllGather: %s", nccl().GetErrorString(nr));
    CU(e);
    for (void *&q : p->ipc_opened) if (q) { cudaIpcCloseMemHandle(q); q = nullptr; }
    p->share.n = 1;
    for (int r = 0; r < c->world; ++r) {
        if (r == c->rank) continue;
        const Rec &o = recs[(size_t)r];
        if (o.sublen != p->sublen) return fail(MPB_E_INVALID, "rank %d runs a different problem", r);
        void *ptr = nullptr;
        if (o.pid == me.pid) {                       // same process (host thread per GPU): the pointer is valid here once peer access is on
            if (o.device != p->ctx->device) {
                cudaError_t pe = cudaDeviceEnablePeerAccess(o.device, 0);
                if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) return fail(MPB_E_CUDA, "cudaDeviceEnablePeerAccess(%d): %s", o.device, cudaGetErrorString(pe));
                cudaGetLastError();
            }
            ptr = (void *)(uintptr_t)o.ptr;
        } else {
            cudaError_t ie = cudaIpcOpenMemHandle(&ptr, o.handle, cudaIpcMemLazyEnablePeerAccess);
            if (ie != cudaSuccess) { cudaGetLastError(); return fail(MPB_E_CUDA, "cudaIpcOpenMemHandle(rank %d): %s", r, cudaGetErrorString(ie)); }
            p->ipc_opened[p->share.n] = ptr;
        }
        p->share.thr[p->share.n++] = (unsigned long long *)ptr;
    }
    return MPB_OK;
}

// Exact (corr, index) max of the plans' result profiles across the communicator, in place (every rank's thresholds rise to the
// merged correlations as well).  Stream-ordered on the context's stream; no host synchronisation.
int mpb_plan_allreduce(mpb_plan *p, mpb_comm *c) {
    if (!p || !c) return fail(MPB_E_INVALID, "null argument");
    if (p->ctx != c->ctx) return fail(MPB_E_INVALID, "plan and communicator belong to different contexts");
    if (c->world == 1) return MPB_OK;
    cudaStream_t s = p->ctx->stream;
    CU(cudaSetDevice(p->ctx->device));
    const size_t sub = (size_t)p->sublen;
    if (c->cap < sub) {
        CU(cudaStreamSynchronize(s));
        if (c->corr) cudaFree(c->corr);
        if (c->index) cudaFree(c->index);
        c->corr = nullptr; c->index = nullptr; c->cap = 0;
        CU(cudaMalloc((void **)&c->corr, sizeof(double) * sub));
        CU(cudaMalloc((void **)&c->index, sizeof(int32_t) * sub));
        c->cap = sub;
    }
    int rc = mpb_plan_export(p, c->corr, c->index);
    if (rc != MPB_OK) return rc;
    NC(nccl().AllReduce(c->corr, c->corr, sub, ncclInt64, ncclMax, c->comm, s));       // corr >= +0.0: the bit pattern orders like the value
    rc = mpb_plan_mask_index(p, c->corr, c->index);                                    // index <- (my corr == max) ? my index : INT32_MAX
    if (rc != MPB_OK) return rc;
    NC(nccl().AllReduce(c->index, c->index, sub, ncclInt32, ncclMin, c->comm, s));
    return mpb_plan_import(p, c->corr, c->index);
}

#ifndef MPB_WARM_BANDS_TOTAL
#define MPB_WARM_BANDS_TOTAL 48   // diagonal bands (512 diagonals each) after which the filter thresholds are tight
#endif

// Local bands after which the ranks max-merge their THRESHOLDS before walking the rest: derived from the job's total band
// count, so every rank takes the same decision (a rank-local count would let ranks disagree when the bands do not divide evenly).
int mpb_merge_point(int total_bands, int world) {
    if (world <= 1) return 0;
    const int warm = std::max(1, (MPB_WARM_BANDS_TOTAL + world - 1) / world);
    return warm < total_bands / world ? warm : 0;
}

// The multi-GPU compute step of one rank (after mpb_plan_precompute): job-wide sketch of band 0 (each rank 1/G of its rows, one
// max-reduction of the bounds), the first bands, a max-merge of the thresholds, the remaining bands, the max-merge of the result.
int mpb_plan_compute_merged(mpb_plan *p, mpb_comm *c) {
    if (!p || !c) return fail(MPB_E_INVALID, "null argument");
    if (p->ctx != c->ctx) return fail(MPB_E_INVALID, "plan and communicator belong to different contexts");
    if (p->rank != c->rank || p->world != c->world) return fail(MPB_E_INVALID, "plan shard %d/%d does not match communicator rank %d/%d", p->rank, p->world, c->rank, c->world);
    cudaStream_t s = p->ctx->stream;
    CU(cudaSetDevice(p->ctx->device));
    int total = 0, rc = mpb_plan_total_bands(p, &total);
    if (rc != MPB_OK) return rc;
    const bool reseeded = !(p->flags & (MPB_FLAG_STRICT | MPB_FLAG_NO_SEED | MPB_FLAG_NO_RESEED));
    if (c->world > 1 && p->dtype == MPB_F64 && reseeded && p->sketch_bands != 0 && total > 0) {
        rc = mpb_plan_sketch_share(p, 0, c->rank, c->world);
        if (rc != MPB_OK) return rc;
        NC(nccl().AllReduce(p->hint, p->hint, (size_t)p->np, ncclInt32, ncclMax, c->comm, s));
        rc = mpb_plan_apply_hint(p);
        if (rc != MPB_OK) return rc;
    }
    // ranks that push their improvements to one another (mpb_plan_connect_peers) need no merge on the way; otherwise the
    // results so far are max-merged once the first bands are walked, which also lifts every rank's thresholds
    const int pt = p->share.n > 1 ? 0 : mpb_merge_point(total, c->world);
    if (pt > 0) {
        rc = mpb_plan_compute_bands(p, 0, pt);
        if (rc == MPB_OK) rc = mpb_plan_allreduce(p, c);
        if (rc == MPB_OK) rc = mpb_plan_compute_bands(p, pt, -1);
    } else {
        rc = mpb_plan_compute_bands(p, 0, -1);
    }
    if (rc == MPB_OK) rc = mpb_plan_allreduce(p, c);
    if (rc == MPB_OK) CU(cudaEventRecord(p->ev[5], s));      // the merges belong to the compute phase
    return rc;
}

// ---------------------------------------------------------------------------------------------
// One call, G GPUs of this process, HOST buffers: the device section of RunMatrixProfileKernel with MP_KERNEL_REPL = G
// (host/MatrixProfileHost.cpp:66-153).  One host thread per GPU (the reference drives its replicas from one thread and
// merges on the host); rank 0 finalises and downloads.  timings (may be NULL) receives G entries.
// ---------------------------------------------------------------------------------------------
namespace {
struct Rendezvous {
    std::mutex mu;
    std::condition_variable cv;
    int waiting = 0, generation = 0, parties = 0;
    int failed = MPB_OK;
    std::string message;
    void fail_with(int rc, const char *msg) { std::lock_guard<std::mutex> l(mu); if (failed == MPB_OK) { failed = rc; message = msg; } }
    // returns the job's status: every thread learns of a failure at the same barrier and leaves together
    int arrive() {
        std::unique_lock<std::mutex> l(mu);
        const int gen = generation;
        if (++waiting == parties) { waiting = 0; ++generation; cv.notify_all(); }
        else cv.wait(l, [&] { return gen != generation; });
        return failed;
    }
};
}  // namespace

int mpb_compute_multi(int ngpus, const int *devices, const double *T, int n, int m, int dtype, unsigned flags, double *MP, int32_t *MPI, mpb_timings *timings) {
    if (ngpus < 1 || !T || !MP || !MPI) return fail(MPB_E_INVALID, "bad argument");
    int count = 0, rc = mpb_device_count(&count);
    if (rc != MPB_OK) return rc;
    for (int g = 0; g < ngpus; ++g) {
        const int dev = devices ? devices[g] : g;
        if (dev < 0 || dev >= count) return fail(MPB_E_INVALID, "device %d out of range [0,%d)", dev, count);
    }
    if (ngpus == 1) {
        mpb_context *ctx = nullptr;
        rc = mpb_context_create(devices ? devices[0] : 0, &ctx);
        if (rc == MPB_OK) rc = mpb_compute(ctx, T, n, m, dtype, flags, MP, MPI, timings);
        mpb_context_destroy(ctx);
        return rc;
    }
    int margin = 0;
    for (int attempt = 0; attempt < 4; ++attempt) {        // MPB_E_MARGIN on any rank: the whole job again with a 4x wider near-tie margin
    char id[NCCL_UNIQUE_ID_BYTES];
    rc = mpb_comm_unique_id(id, sizeof id);
    if (rc != MPB_OK) return rc;
    Rendezvous rv;
    rv.parties = ngpus;
    int margin_next = 0;
    auto worker = [&](int g) {
        mpb_context *ctx = nullptr; mpb_plan *plan = nullptr; mpb_comm *comm = nullptr;
        auto step = [&](int r) { if (r != MPB_OK) rv.fail_with(r, mpb_last_error()); };
        int r = mpb_context_create(devices ? devices[g] : g, &ctx);
        if (r == MPB_OK) r = mpb_plan_create(ctx, n, m, dtype, flags, &plan);
        if (r == MPB_OK) r = mpb_plan_set_shard(plan, g, ngpus);
        if (r == MPB_OK && margin) r = mpb_plan_set_chain_margin(plan, margin);
        if (r == MPB_OK && g == 0) margin_next = std::max(8, plan->margin_log2 - 2);
        step(r);
        if (rv.arrive() == MPB_OK) {                       // nobody enters the NCCL rendezvous unless everybody can
            step(mpb_comm_create(ctx, id, g, ngpus, &comm));
            if (rv.arrive() == MPB_OK && ngpus <= mpb::MAX_PEERS) {
                step(mpb_plan_connect_peers(plan, comm));
                rv.arrive();
            }
            if (rv.failed == MPB_OK) {
                r = mpb_plan_upload_series(plan, T);
                if (r == MPB_OK) r = mpb_plan_precompute(plan);
                if (r == MPB_OK) r = mpb_plan_compute_merged(plan, comm);
                if (r == MPB_OK && g == 0) r = mpb_plan_finalize(plan);
                if (r == MPB_OK && g == 0) r = mpb_plan_download(plan, MP, MPI, nullptr);
                if (r == MPB_OK && g != 0) r = mpb_plan_check(plan);
                if (r == MPB_OK && timings) r = mpb_plan_timings(plan, &timings[g]);
                step(r);
            }
        }
        mpb_comm_destroy(comm);
        mpb_plan_destroy(plan);
        mpb_context_destroy(ctx);
    };
    std::vector<std::thread> th;
    for (int g = 1; g < ngpus; ++g) th.emplace_back(worker, g);
    worker(0);
    for (auto &t : th) t.join();
    if (rv.failed == MPB_E_MARGIN && attempt < 3) { margin = margin_next; continue; }
    if (rv.failed != MPB_OK) return fail(rv.failed, "%s", rv.message.c_str());
    return MPB_OK;
    }
    return rc;
}

}  // extern "C"
