// chunk_tlf.cu -- the chunk-compatible kernel entry: MatrixProfileKernelTLF(n, m, iteration, input, output)
// (kernel/MatrixProfileKernelVanilla.cpp:16-107) with the reference's exact per-chunk semantics, so that
// the reference's own harness (test/include/MatrixProfileTests.hpp:141-213) can drive the GPU:
//   * un-fused arithmetic in data_t (QT += row.df*column.dg + column.df*row.dg, :81; P = QT*row.inv*column.inv, :84);
//   * P = 0 substituted for excluded / out-of-range cells, which still compete (:72-84);
//   * row aggregate over all w cells of the row with the Vanilla kernel's tie order: the 8-lane partial
//     maxima keep the LATER column on ties (:87-90) and the tree keeps the HIGHER lane
//     (include/kernel/TreeReduce.hpp:28,42) => among equal maxima the winner maximises (j mod 8, j);
//   * column aggregate keeps the EARLIER row on ties (:93-96) => minimises i.
// This is the drop-in/compat path (bit-for-bit the reference kernel's OutputDataPacks); the fast path is scamp_tile.cu.
#include "mp_common.cuh"
#include "mp_kernels.h"
#include "mp_runtime.h"

#include <mutex>

namespace mpb {

// order-preserving map fp -> unsigned 64 (total order incl. negatives; 0 is below every value)
__device__ __forceinline__ unsigned long long ord_key(double v) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double ord_unkey(unsigned long long k, double) {
    const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
    return __longlong_as_double((long long)b);
}
__device__ __forceinline__ unsigned long long ord_key(float v) {
    const unsigned b = (unsigned)__float_as_int(v);
    return (unsigned long long)((b >> 31) ? ~b : (b | 0x80000000u));
}
__device__ __forceinline__ float ord_unkey(unsigned long long k, float) {
    const unsigned kk = (unsigned)k;
    const unsigned b = (kk >> 31) ? (kk & 0x7fffffffu) : ~kk;
    return __int_as_float((int)b);
}

__device__ __forceinline__ double strict_step(double qt, double rdf, double rdg, double cdf, double cdg) {
    return __dadd_rn(qt, __dadd_rn(__dmul_rn(rdf, cdg), __dmul_rn(cdf, rdg)));
}
__device__ __forceinline__ float strict_step(float qt, float rdf, float rdg, float cdf, float cdg) {
    return __fadd_rn(qt, __fadd_rn(__fmul_rn(rdf, cdg), __fmul_rn(cdf, rdg)));
}
__device__ __forceinline__ double strict_corr(double qt, double ri, double ci) { return __dmul_rn(__dmul_rn(qt, ri), ci); }
__device__ __forceinline__ float strict_corr(float qt, float ri, float ci) { return __fmul_rn(__fmul_rn(qt, ri), ci); }

// ---------------------------------------------------------------------------------------------
// Decomposition.  The chain QT_i = QT_{i-1} + (df_i*dg_c + df_c*dg_i) must be walked in order with the reference's
// roundings, but only ONE addition per row sits on the chain; everything else (the two products, P, the maxima) does not.
//   pass 0  one thread per diagonal walks the bare chain (2 DMUL + 2 DADD per row) and drops a checkpoint every TLF_R rows;
//   pass 1  a CTA per (TLF_D diagonals x TLF_R rows) tile restarts every diagonal from its checkpoint (same operations, same
//           order: the same bits), forms P with the reference's clamp, and reduces (value, tie-order) PAIRS: per row across
//           the warp with REDUX, per column by handing the running pair of a column from lane j+1 to lane j as the rows advance
//           (the Tiled kernel's systolic shift, kernel/MatrixProfileKernelTiled.cpp:204-219); one 128-bit compare-and-swap
//           per row / column and tile merges the tile's pairs into the chunk's;
//   pass 2  writes the OutputDataPacks.
// A pair is (order-preserving integer image of P, tie code): the tie code of a row cell is ((j mod 8) << 24 | j) + 1 (largest
// wins), that of a column cell ~i (earliest row wins), so one lexicographic max reproduces value AND index.
// ---------------------------------------------------------------------------------------------
constexpr int TLF_R = 192;             // rows per tile = checkpoint distance
constexpr int TLF_D = 256;             // diagonals per tile (one per thread)
struct __align__(16) TlfPair { unsigned long long key; unsigned tie; unsigned pad; };
struct TlfScratch {
    TlfPair *row, *col;                // [nRows] running (value, tie) maxima of the chunk
    void *ckpt;                        // [ceil(nRows / TLF_R)][w] QT of every diagonal just before the tile's first row
};

__device__ __forceinline__ bool pair_less(unsigned long long ka, unsigned ta, unsigned long long kb, unsigned tb) { return ka < kb || (ka == kb && ta < tb); }

// lexicographic max into a global pair, exact whatever the arrival order
__device__ __forceinline__ void pair_max_global(TlfPair *slot, unsigned long long key, unsigned tie) {
    unsigned long long clo, chi, olo, ohi;
    asm volatile("ld.global.relaxed.gpu.v2.u64 {%0,%1}, [%2];" : "=l"(clo), "=l"(chi) : "l"(slot) : "memory");
    const unsigned long long nhi = (unsigned long long)tie;
    while (pair_less(clo, (unsigned)chi, key, tie)) {
        asm volatile(
            "{\n .reg .b128 c, d, o;\n mov.b128 c, {%2, %3};\n mov.b128 d, {%4, %5};\n"
            " atom.global.relaxed.gpu.cas.b128 o, [%6], c, d;\n mov.b128 {%0, %1}, o;\n}"
            : "=l"(olo), "=l"(ohi) : "l"(clo), "l"(chi), "l"(key), "l"(nhi), "l"(slot) : "memory");
        if (olo == clo && ohi == chi) break;
        clo = olo; chi = ohi;
    }
}

__global__ void tlf_init_kernel(TlfScratch s, int nRows) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nRows) { TlfPair z; z.key = 0; z.tie = 0; z.pad = 0; s.row[i] = z; s.col[i] = z; }
}

// pass 0: the bare chain of every diagonal of the chunk, checkpointed every TLF_R rows
template <typename T>
__global__ void __launch_bounds__(128) tlf_chain_kernel(int n, int m, int w, int iteration, const InputDataPack<T> *__restrict__ input, TlfScratch s) {
    const int sublen = n - m + 1, nOffset = iteration * w, nRows = sublen - nOffset;
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= w) return;
    T *ck = reinterpret_cast<T *>(s.ckpt);
    T qt = nOffset + j < sublen ? input[nOffset + j].QT : (T)0;                    // :43-55
    for (int i0 = 0; i0 < nRows; i0 += TLF_R) {
        ck[(size_t)(i0 / TLF_R) * w + j] = qt;
        const int cnt = min(TLF_R, nRows - i0);
#pragma unroll 4
        for (int r = 0; r < cnt; ++r) {
            const int i = i0 + r, col = nOffset + i + j;
            const T rdf = input[i].df, rdg = input[i].dg;
            T cdf = 0, cdg = 0;
            if (col < sublen) { cdf = input[col].df; cdg = input[col].dg; }        // columns beyond the series read as zero packs (:52-53,77-78)
            qt = strict_step(qt, rdf, rdg, cdf, cdg);
        }
    }
}

// pass 1: one (TLF_D diagonals x TLF_R rows) tile per CTA
template <typename T>
__global__ void __launch_bounds__(TLF_D) tlf_tile_kernel(int n, int m, int w, int iteration, const InputDataPack<T> *__restrict__ input, TlfScratch s) {
    __shared__ unsigned long long rowK[TLF_D / 32][TLF_R], colK[TLF_D / 32][TLF_R + 32];
    __shared__ unsigned rowT[TLF_D / 32][TLF_R], colT[TLF_D / 32][TLF_R + 32];
    const unsigned FULL = 0xffffffffu;
    const int sublen = n - m + 1, nOffset = iteration * w, nRows = sublen - nOffset, excl = m / 4;
    const int j0 = blockIdx.x * TLF_D, i0 = blockIdx.y * TLF_R;
    const int t = threadIdx.x, lane = t & 31, wq = t >> 5, j = j0 + t;
    const bool active = j < w;
    const int cnt = min(TLF_R, nRows - i0);
    T qt = active ? reinterpret_cast<const T *>(s.ckpt)[(size_t)blockIdx.y * w + j] : (T)0;
    unsigned long long pk = 0; unsigned pt = 0;                                    // running pair of the column this lane is on
    for (int e = lane; e < TLF_R + 32; e += 32) { colK[wq][e] = 0; colT[wq][e] = 0; }
    __syncwarp();
#pragma unroll 2
    for (int r = 0; r < cnt; ++r) {
        const int i = i0 + r, col = nOffset + i + j;
        const InputDataPack<T> row = input[i];
        const bool colIn = active && col < sublen;
        InputDataPack<T> c; c.QT = 0; c.df = 0; c.dg = 0; c.inv = 0;
        if (colIn) c = input[col];
        qt = strict_step(qt, row.df, row.dg, c.df, c.dg);                          // :81
        const bool inb = colIn && !(i > col - excl);                               // :72-74
        const T P = inb ? strict_corr(qt, row.inv, c.inv) : (T)0;                   // :84
        const unsigned long long key = active ? ord_key(P) : 0ull;
        // column: take over the pair lane + 1 carried for this column, add this row's cell
        unsigned long long ik = __shfl_down_sync(FULL, pk, 1); unsigned it = __shfl_down_sync(FULL, pt, 1);
        if (lane == 31) { ik = 0; it = 0; }
        pk = ik; pt = it;
        if (active && i + j < nRows && pair_less(pk, pt, key, ~(unsigned)i)) { pk = key; pt = ~(unsigned)i; }   // :93-96: the earlier row wins ties
        if (lane == 0) { colK[wq][r] = pk; colT[wq][r] = pt; }                      // lane 0's column leaves the warp's window: complete for these 32 diagonals
        // row: (value, tie code) max over the warp's 32 cells (:87-90 + TreeReduce: maximise (j mod 8, j) among equal values)
        const unsigned kh = __reduce_max_sync(FULL, (unsigned)(key >> 32));
        const unsigned kl = __reduce_max_sync(FULL, (unsigned)(key >> 32) == kh ? (unsigned)key : 0u);
        const unsigned long long kmax = (unsigned long long)kh << 32 | kl;
        const unsigned tie = (active && key == kmax) ? ((((unsigned)j & 7u) << 24) | (unsigned)j) + 1u : 0u;
        const unsigned tmax = __reduce_max_sync(FULL, tie);
        if (lane == 0) { rowK[wq][r] = kmax; rowT[wq][r] = tmax; }
    }
    if (cnt > 0 && lane > 0) { colK[wq][cnt - 1 + lane] = pk; colT[wq][cnt - 1 + lane] = pt; }   // the columns still inside the window
    __syncthreads();
    for (int r = t; r < cnt; r += TLF_D) {                                         // rows of the tile: max over the warps, then into the chunk's
        unsigned long long k = rowK[0][r]; unsigned tt = rowT[0][r];
#pragma unroll
        for (int q = 1; q < TLF_D / 32; ++q) if (pair_less(k, tt, rowK[q][r], rowT[q][r])) { k = rowK[q][r]; tt = rowT[q][r]; }
        pair_max_global(s.row + i0 + r, k, tt);
    }
    for (int cc = t; cc < cnt + TLF_D - 1; cc += TLF_D) {                          // columns i0 + j0 + cc (relative to the chunk) the tile touched
        if (i0 + j0 + cc >= nRows) break;
        unsigned long long k = 0; unsigned tt = 0;
#pragma unroll
        for (int q = 0; q < TLF_D / 32; ++q) {
            const int e = cc - 32 * q;                                             // warp q wrote the column at entry cc - 32 q
            if (e >= 0 && e < cnt + 31 && pair_less(k, tt, colK[q][e], colT[q][e])) { k = colK[q][e]; tt = colT[q][e]; }
        }
        if (k) pair_max_global(s.col + i0 + j0 + cc, k, tt);
    }
}

template <typename T>
__global__ void tlf_write_kernel(int nRows, int nOffset, TlfScratch s, OutputDataPack<T> *__restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nRows) return;
    OutputDataPack<T> o;
    memset(&o, 0, sizeof o);
    const TlfPair r = s.row[i], c = s.col[i];
    o.rowAggregate.value = ord_unkey(r.key, T(0));
    o.rowAggregate.index = nOffset + i + (int)((r.tie - 1u) & 0xffffffu);
    o.columnAggregate.value = ord_unkey(c.key, T(0));
    o.columnAggregate.index = (int)~c.tie;
    out[i] = o;
}

size_t chunk_tlf_scratch_bytes(int n, int m, int w) {
    const size_t sublen = (size_t)(n - m + 1), tiles = (sublen + TLF_R - 1) / TLF_R;
    return sublen * 2 * sizeof(TlfPair) + tiles * (size_t)w * 8 + 256;
}

template <typename T>
static cudaError_t launch_chunk_tlf_t(int n, int m, int w, int iteration, const void *input, void *output, void *scratch, size_t scratch_bytes, cudaStream_t st) {
    const int sublen = n - m + 1;
    const int nOffset = iteration * w;
    const int nRows = sublen - nOffset;
    if (nRows <= 0) return cudaSuccess;
    if (scratch_bytes < chunk_tlf_scratch_bytes(n, m, w) || w < 8 || w >= (1 << 24)) return cudaErrorInvalidValue;
    TlfScratch s;
    unsigned char *p = (unsigned char *)scratch;
    s.row = (TlfPair *)p; p += sizeof(TlfPair) * (size_t)sublen;
    s.col = (TlfPair *)p; p += sizeof(TlfPair) * (size_t)sublen;
    s.ckpt = p;
    const int gr = (nRows + 255) / 256;
    tlf_init_kernel<<<gr, 256, 0, st>>>(s, nRows);
    tlf_chain_kernel<T><<<(w + 127) / 128, 128, 0, st>>>(n, m, w, iteration, (const InputDataPack<T> *)input, s);
    tlf_tile_kernel<T><<<dim3((w + TLF_D - 1) / TLF_D, (nRows + TLF_R - 1) / TLF_R), TLF_D, 0, st>>>(n, m, w, iteration, (const InputDataPack<T> *)input, s);
    tlf_write_kernel<T><<<gr, 256, 0, st>>>(nRows, nOffset, s, (OutputDataPack<T> *)output);
    return cudaGetLastError();
}
cudaError_t launch_chunk_tlf_f64(int n, int m, int w, int iteration, const void *input, void *output, void *scratch, size_t scratch_bytes, cudaStream_t s) {
    return launch_chunk_tlf_t<double>(n, m, w, iteration, input, output, scratch, scratch_bytes, s);
}
cudaError_t launch_chunk_tlf_f32(int n, int m, int w, int iteration, const void *input, void *output, void *scratch, size_t scratch_bytes, cudaStream_t s) {
    return launch_chunk_tlf_t<float>(n, m, w, iteration, input, output, scratch, scratch_bytes, s);
}

}  // namespace mpb

// ---------------------------------------------------------------------------------------------
// Allocation-free variant of mpb_kernel_tlf_enqueue (include/mp_b200.h: no counterpart in the reference; added
// for callers that cannot call cudaMalloc -- CUDA graph capture, tight serving loops, harnesses that
// forbid it).  The chunk's work space (running (value, tie) pairs + chain checkpoints) lives in ONE file-scope device array
// of MPB_TLF_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit.  In host code the array's name is
// only a host-side stand-in: its device address is taken once per device with cudaGetSymbolAddress (that device current)
// and that pointer is what the kernels receive.  All calls share the array: tlf_init_kernel clears the pairs of the chunk's
// rows at the start of every call (the checkpoints are written before they are read), and calls that overlap on the GPU
// (two streams, two contexts of one device) race on it.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_tlf_static_scratch[MPB_TLF_NOALLOC_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
void *g_static_addr[kMaxDevices] = {};      // device address of g_tlf_static_scratch, per device ordinal

// device must be current
cudaError_t tlf_static_scratch(int device, void **ptr) {
    std::lock_guard<std::mutex> l(g_static_mu);
    if (!g_static_addr[device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_tlf_static_scratch);
        if (e != cudaSuccess) return e;
        g_static_addr[device] = p;
    }
    *ptr = g_static_addr[device];
    return cudaSuccess;
}
}  // namespace

extern "C" int mpb_kernel_tlf_enqueue_noalloc(mpb_context *c, int dtype, int n, int m, int w, int iteration, const mpb_buffer *input, mpb_buffer *output) {
    if (!c || !input || !output) return fail(MPB_E_INVALID, "null argument");
    if (dtype != MPB_F64 && dtype != MPB_F32) return fail(MPB_E_INVALID, "dtype must be MPB_F64 or MPB_F32");
    if (m < 4 || n < m) return fail(MPB_E_INVALID, "need n >= m >= 4 (n=%d, m=%d)", n, m);
    if (w < 8 || w > n || w >= (1 << 24)) return fail(MPB_E_INVALID, "need 8 <= w <= n (w=%d)", w);
    if (iteration < 0 || (long long)iteration * w >= (long long)(n - m + 1)) return fail(MPB_E_INVALID, "iteration %d outside [0, ceil((n-m+1)/w))", iteration);
    const size_t need = mpb::chunk_tlf_scratch_bytes(n, m, w);
    if (need > (size_t)MPB_TLF_NOALLOC_SCRATCH_BYTES)
        return fail(MPB_E_INVALID, "mpb_kernel_tlf_enqueue_noalloc: n=%d, m=%d, w=%d needs %zu bytes of work space, the static array holds %zu "
                                   "(MPB_TLF_NOALLOC_SCRATCH_BYTES); use mpb_kernel_tlf_enqueue for larger problems", n, m, w, need, (size_t)MPB_TLF_NOALLOC_SCRATCH_BYTES);
    const size_t sublen = (size_t)(n - m + 1), esz = dtype == MPB_F64 ? 32 : 16;
    if (input->bytes < sublen * esz) return fail(MPB_E_INVALID, "input buffer holds %zu bytes, need %zu", input->bytes, sublen * esz);
    const size_t nRows = sublen - (size_t)iteration * w;
    if (output->bytes < nRows * esz) return fail(MPB_E_INVALID, "output buffer holds %zu bytes, need %zu", output->bytes, nRows * esz);
    if (c->device < 0 || c->device >= kMaxDevices) return fail(MPB_E_INVALID, "device ordinal %d beyond the %d this entry point tracks", c->device, kMaxDevices);
    CU(cudaSetDevice(c->device));
    void *scratch = nullptr;
    CU(tlf_static_scratch(c->device, &scratch));
    if (dtype == MPB_F64) CU(mpb::launch_chunk_tlf_f64(n, m, w, iteration, input->ptr, output->ptr, scratch, MPB_TLF_NOALLOC_SCRATCH_BYTES, c->stream));
    else CU(mpb::launch_chunk_tlf_f32(n, m, w, iteration, input->ptr, output->ptr, scratch, MPB_TLF_NOALLOC_SCRATCH_BYTES, c->stream));
    return MPB_OK;
}
