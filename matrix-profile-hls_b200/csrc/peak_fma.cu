// peak_fma.cu -- register-resident FMA issue-rate microbenchmark: the denominator of the FMA roofline
// (MEASURED_PEAKS.json carries no FP64/FP32 CUDA-core peak).
#include "mp_kernels.h"

namespace mpb {

constexpr int PEAK_CHAINS = 16;   // independent accumulators per thread: enough ILP to cover FMA latency
constexpr int PEAK_THREADS = 256;

template <typename T>
__global__ void __launch_bounds__(PEAK_THREADS) fma_peak_kernel(int iters, double *sink) {
    T acc[PEAK_CHAINS];
    const T a = (T)1.0000001 + (T)threadIdx.x * (T)1e-9, b = (T)1e-7;
#pragma unroll
    for (int c = 0; c < PEAK_CHAINS; ++c) acc[c] = (T)c;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int c = 0; c < PEAK_CHAINS; ++c) acc[c] = acc[c] * a + b;   // contracts to one FMA
    }
    T sum = 0;
#pragma unroll
    for (int c = 0; c < PEAK_CHAINS; ++c) sum += acc[c];
    if (sum == (T)123.456) sink[0] = (double)sum;   // keeps the chains alive, practically never taken
}

cudaError_t launch_fma_peak(int dtype, int iters, int blocks, double *sink, cudaStream_t s) {
    if (dtype == 0) fma_peak_kernel<double><<<blocks, PEAK_THREADS, 0, s>>>(iters, sink);
    else fma_peak_kernel<float><<<blocks, PEAK_THREADS, 0, s>>>(iters, sink);
    return cudaGetLastError();
}
int fma_peak_chains() { return PEAK_CHAINS; }
int fma_peak_threads() { return PEAK_THREADS; }

}  // namespace mpb
