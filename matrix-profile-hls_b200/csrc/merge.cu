// merge.cu -- kernel (3): final merge / conversion, and the pieces of the multi-GPU max-reduction.
//
// finalize_kernel is the reference's PostComputeAggregates + PearsonCorrelationToEuclideanDistance
// (include/host/HostSideComputation.hpp:47-69): MP[i] = sqrt(2*m*(1-P)) in double, MPI[i] = index.  The
// row-vs-column merge of :61 and the chunk merge of :72-90 already happened inside the band kernel
// (one profile slot per subsequence).  The exclusion zone needs no handling here because excluded
// diagonals (d < m/4) are never walked.
#include "mp_common.cuh"
#include "mp_kernels.h"
#include <limits.h>

namespace mpb {

__global__ void finalize_kernel(const ProfileEntry *__restrict__ prof, int sublen, int m, double *__restrict__ MP,
                                int32_t *__restrict__ MPI, double *__restrict__ corr) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sublen) return;
    const ProfileEntry e = prof[i];
    if (corr) corr[i] = e.corr;
    MP[i] = __dsqrt_rn(__dmul_rn((double)(2 * m), __dsub_rn(1.0, e.corr)));   // :48
    MPI[i] = e.index;                                                          // :65
}

__global__ void export_kernel(const ProfileEntry *__restrict__ prof, int sublen, double *__restrict__ corr, int32_t *__restrict__ index) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < sublen) { const ProfileEntry e = prof[i]; corr[i] = e.corr; index[i] = e.index; }
}

// index[i] = INT32_MAX unless this rank holds the global maximum, so min-reduce(index) = smallest index
// among the holders (the same tie rule as profile_update). An unset slot (index -1) never holds a
// maximum that another rank set, and if nobody set it the merged index must stay -1.
__global__ void mask_index_kernel(const ProfileEntry *__restrict__ prof, int sublen, const double *__restrict__ corr_max, int32_t *__restrict__ index) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sublen) return;
    const ProfileEntry e = prof[i];
    index[i] = (e.corr == corr_max[i] && e.index >= 0) ? e.index : INT_MAX;
}

__global__ void import_kernel(ProfileEntry *__restrict__ prof, int sublen, const double *__restrict__ corr, const int32_t *__restrict__ index,
                              unsigned long long *__restrict__ thr) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sublen) return;
    ProfileEntry e; e.corr = corr[i]; e.index = index[i] == INT_MAX ? -1 : index[i]; e.pad = 0;
    prof[i] = e;
    if (thr && e.corr > 0.0) atomicMax(thr + i, (unsigned long long)__double_as_longlong(e.corr));   // what the job knows is a threshold for this rank too
}

// Exact (corr, index) max of a peer GPU's profile into ours; `peer` is a P2P-mapped pointer, so the
// loads cross NVLink directly (no staging copy).
__global__ void merge_peer_kernel(ProfileEntry *__restrict__ prof, const ProfileEntry *__restrict__ peer, int sublen, unsigned long long *__restrict__ thr) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= sublen) return;
    const ProfileEntry a = prof[i], b = peer[i];
    if (b.corr > a.corr || (b.corr == a.corr && (unsigned)b.index < (unsigned)a.index)) {
        prof[i] = b;
        if (thr && b.corr > 0.0) atomicMax(thr + i, (unsigned long long)__double_as_longlong(b.corr));
    }
}

#define GRID(n) ((n) + 255) / 256, 256
cudaError_t launch_finalize(const ProfileEntry *prof, int sublen, int m, double *MP, int32_t *MPI, double *corr, cudaStream_t s) {
    finalize_kernel<<<GRID(sublen), 0, s>>>(prof, sublen, m, MP, MPI, corr);
    return cudaGetLastError();
}
cudaError_t launch_export(const ProfileEntry *prof, int sublen, double *corr, int32_t *index, cudaStream_t s) {
    export_kernel<<<GRID(sublen), 0, s>>>(prof, sublen, corr, index);
    return cudaGetLastError();
}
cudaError_t launch_mask_index(const ProfileEntry *prof, int sublen, const double *corr_max, int32_t *index, cudaStream_t s) {
    mask_index_kernel<<<GRID(sublen), 0, s>>>(prof, sublen, corr_max, index);
    return cudaGetLastError();
}
cudaError_t launch_import(ProfileEntry *prof, int sublen, const double *corr, const int32_t *index, unsigned long long *thr, cudaStream_t s) {
    import_kernel<<<GRID(sublen), 0, s>>>(prof, sublen, corr, index, thr);
    return cudaGetLastError();
}
cudaError_t launch_merge_peer(ProfileEntry *prof, const ProfileEntry *peer, int sublen, unsigned long long *thr, cudaStream_t s) {
    merge_peer_kernel<<<GRID(sublen), 0, s>>>(prof, peer, sublen, thr);
    return cudaGetLastError();
}

}  // namespace mpb
