// chunk_tlf.cu -- the chunk-compatible kernel entry: MatrixProfileKernelTLF(n, m, iteration, input, output)
// (kernel/MatrixProfileKernelVanilla.cpp:16-107) with the reference's exact per-chunk semantics, so that
// the reference's own harness (test/include/MatrixProfileTests.hpp:141-213) can drive the GPU:
//   * un-fused arithmetic in data_t (QT += row.df*column.dg + column.df*row.dg, :81; P = QT*row.inv*column.inv, :84);
//   * P = 0 substituted for excluded / out-of-range cells, which still compete (:72-84);
//   * row aggregate over all w cells of the row with the Vanilla kernel's tie order: the 8-lane partial
//     maxima keep the LATER column on ties (:87-90) and the tree keeps the HIGHER lane
//     (include/kernel/TreeReduce.hpp:28,42) => among equal maxima the winner maximises (j mod 8, j);
//   * column aggregate keeps the EARLIER row on ties (:93-96) => minimises i.
// One thread per diagonal of the chunk walks all rows (same chain as the reference). Exactness without
// 128-bit atomics: pass 1 takes the maxima of an order-preserving integer image of P, pass 2 re-walks
// the chain and resolves ties among the cells that equal the maximum, pass 3 writes OutputDataPacks.
// This is the drop-in/compat path; the fast path is scamp_diag.cu.
#include "mp_common.cuh"
#include "mp_kernels.h"

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

struct TlfScratch {
    unsigned long long *rowKey, *colKey;   // [nRows] maxima of ord_key(P)
    unsigned *rowTie, *colTie;             // [nRows] tie winners: max ((j&7)<<24 | j) / min i
};

__global__ void tlf_init_kernel(TlfScratch s, int nRows) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nRows) { s.rowKey[i] = 0; s.colKey[i] = 0; s.rowTie[i] = 0; s.colTie[i] = 0xffffffffu; }
}

template <typename T, int PASS>
__global__ void __launch_bounds__(256) tlf_walk_kernel(int n, int m, int w, int iteration, const InputDataPack<T> *__restrict__ input, TlfScratch s) {
    const int sublen = n - m + 1;
    const int nOffset = iteration * w;
    const int nRows = sublen - nOffset;
    const int j = blockIdx.x * blockDim.x + threadIdx.x;     // diagonal within the chunk
    const bool active = j < w;
    const int excl = m / 4;
    T qt = 0;
    if (active && nOffset + j < sublen) qt = input[nOffset + j].QT;      // :43-55
    for (int i = 0; i < nRows; ++i) {
        const InputDataPack<T> row = input[i];
        const int col = nOffset + i + j;
        const bool colIn = active && col < sublen;
        InputDataPack<T> c; c.QT = 0; c.df = 0; c.dg = 0; c.inv = 0;
        if (colIn) c = input[col];
        qt = strict_step(qt, row.df, row.dg, c.df, c.dg);
        const bool inb = colIn && !(i > col - excl);
        const T P = inb ? strict_corr(qt, row.inv, c.inv) : (T)0;
        const unsigned long long key = active ? ord_key(P) : 0ull;
        if (PASS == 1) {
            unsigned long long k = key;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) { const unsigned long long other = __shfl_xor_sync(0xffffffffu, k, o); k = other > k ? other : k; }
            if ((threadIdx.x & 31) == 0) atomicMax(&s.rowKey[i], k);
            if (active && i + j < nRows) atomicMax(&s.colKey[i + j], key);
        } else {
            const unsigned long long rk = s.rowKey[i];
            unsigned tie = (active && key == rk) ? ((((unsigned)j & 7u) << 24) | (unsigned)j) + 1u : 0u;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) tie = max(tie, __shfl_xor_sync(0xffffffffu, tie, o));
            if ((threadIdx.x & 31) == 0 && tie) atomicMax(&s.rowTie[i], tie);
            if (active && i + j < nRows && key == s.colKey[i + j]) atomicMin(&s.colTie[i + j], (unsigned)i);
        }
    }
}

template <typename T>
__global__ void tlf_write_kernel(int nRows, int nOffset, TlfScratch s, OutputDataPack<T> *__restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nRows) return;
    OutputDataPack<T> o;
    memset(&o, 0, sizeof o);
    o.rowAggregate.value = ord_unkey(s.rowKey[i], T(0));
    o.rowAggregate.index = nOffset + i + (int)((s.rowTie[i] - 1u) & 0xffffffu);
    o.columnAggregate.value = ord_unkey(s.colKey[i], T(0));
    o.columnAggregate.index = (int)s.colTie[i];
    out[i] = o;
}

size_t chunk_tlf_scratch_bytes(int n, int m) {
    const size_t sublen = (size_t)(n - m + 1);
    return sublen * (8 + 8 + 4 + 4) + 256;
}

template <typename T>
static cudaError_t launch_chunk_tlf_t(int n, int m, int w, int iteration, const void *input, void *output, void *scratch, size_t scratch_bytes, cudaStream_t st) {
    const int sublen = n - m + 1;
    const int nOffset = iteration * w;
    const int nRows = sublen - nOffset;
    if (nRows <= 0) return cudaSuccess;
    if (scratch_bytes < chunk_tlf_scratch_bytes(n, m) || w < 8 || w >= (1 << 24)) return cudaErrorInvalidValue;
    TlfScratch s;
    unsigned char *p = (unsigned char *)scratch;
    s.rowKey = (unsigned long long *)p; p += 8 * (size_t)sublen;
    s.colKey = (unsigned long long *)p; p += 8 * (size_t)sublen;
    s.rowTie = (unsigned *)p; p += 4 * (size_t)sublen;
    s.colTie = (unsigned *)p;
    const int gr = (nRows + 255) / 256, gw = (w + 255) / 256;
    tlf_init_kernel<<<gr, 256, 0, st>>>(s, nRows);
    tlf_walk_kernel<T, 1><<<gw, 256, 0, st>>>(n, m, w, iteration, (const InputDataPack<T> *)input, s);
    tlf_walk_kernel<T, 2><<<gw, 256, 0, st>>>(n, m, w, iteration, (const InputDataPack<T> *)input, s);
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
