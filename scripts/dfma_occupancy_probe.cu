// dfma_occupancy_probe.cu -- how many warps per SM sub-partition does the FP64 pipe of a B200 need?
// The tile kernel (scamp_tile.cu) runs 2 warps per sub-partition, each with 8 independent QT chains of 2 dependent
// DFMAs per row plus the filter's integer max tree and the operand LDS.128s.  This probe reproduces that instruction
// pattern register-resident (no global traffic) and sweeps warps per sub-partition x extra instruction mix, printing
// the achieved fraction of 64 DFMA lanes per SM per clock.  Tuning evidence only (profiles/), not part of the library.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/dfma_occupancy_probe scripts/dfma_occupancy_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

constexpr int CH = 8;      // chains per thread (DPT)
constexpr int ROWS = 8;    // rows per block of the hot loop

template <int MIX>   // 0: DFMA only, 1: + max3 key tree (4 VIMNMX3 per row), 2: + 2 LDS.128 per row and operand rotation, 3: + 16 register moves per block
__global__ void __launch_bounds__(1024, 1) probe(int iters, double *sink, int *isink) {
    extern __shared__ double2 sm[];
    double acc[CH], ck[CH];
    const int lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) sm[i] = make_double2(1.0 + 1e-9 * i, 1e-7 * i);
    __syncthreads();
    double2 ra = sm[lane], ca = sm[lane + 32];
#pragma unroll
    for (int c = 0; c < CH; ++c) acc[c] = c + threadIdx.x * 1e-3;
    int run = 0, hits = 0;
    const double2 *rp = sm + (threadIdx.x >> 5) * 8, *cp = sm + 64 + lane;
    for (int it = 0; it < iters; ++it) {
        if (MIX >= 3) {
#pragma unroll
            for (int c = 0; c < CH; ++c) ck[c] = acc[c];
        }
#pragma unroll
        for (int r = 0; r < ROWS; ++r) {
            double2 ra_n = ra, ca_n = ca;
            if (MIX >= 2) { ra_n = rp[(it * ROWS + r) & 63]; ca_n = cp[((it * ROWS + r) & 63) * 9]; }
            if (MIX >= 1) {
                const int a = __vimax3_s32(__double2hiint(acc[0]), __double2hiint(acc[1]), __double2hiint(acc[2]));
                const int b = __vimax3_s32(__double2hiint(acc[3]), __double2hiint(acc[4]), __double2hiint(acc[5]));
                const int c = __vimax3_s32(__double2hiint(acc[6]), __double2hiint(acc[7]), run);
                run = __vimax3_s32(a, b, c);
            }
#pragma unroll
            for (int c = 0; c < CH; ++c) acc[c] = fma(ra.x, ca.y, fma(ca.x, ra.y, acc[c]));
            ra = ra_n; ca = ca_n;
        }
        if (MIX >= 1) {
            if (__any_sync(0xffffffffu, run > 0x7fe00000)) {
                hits++;
                if (MIX >= 3) {
#pragma unroll
                    for (int c = 0; c < CH; ++c) acc[c] = ck[c];
                }
            }
            run = 0;
        }
    }
    double s = 0;
#pragma unroll
    for (int c = 0; c < CH; ++c) s += acc[c];
    if (s == 123.456) { sink[0] = s; isink[0] = hits + run; }
}

template <int MIX>
static double run_one(int warps_per_sm, int iters, int sms, double clk_hz) {
    double *sink; int *isink;
    cudaMalloc(&sink, 8); cudaMalloc(&isink, 4);
    const size_t smem = 120 * 1024;   // one CTA per SM
    cudaFuncSetAttribute(probe<MIX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    probe<MIX><<<sms, 32 * warps_per_sm, smem>>>(iters / 8 + 1, sink, isink);
    cudaEventRecord(a);
    probe<MIX><<<sms, 32 * warps_per_sm, smem>>>(iters, sink, isink);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    const double dfma = (double)sms * 32 * warps_per_sm * (double)iters * ROWS * CH * 2;
    const double frac = dfma / (ms * 1e-3) / (sms * 64.0 * clk_hz);
    cudaFree(sink); cudaFree(isink); cudaEventDestroy(a); cudaEventDestroy(b);
    if (cudaGetLastError() != cudaSuccess) return -1;
    return frac;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double clk = khz * 1e3;
    const int iters = 20000;
    const int wps[] = {4, 8, 12, 16, 24, 32};
    for (int w : wps) {
        printf("{\"probe\": \"dfma_occupancy\", \"warps_per_sm\": %d, \"warps_per_subpartition\": %.1f, \"clock_mhz\": %.0f, "
               "\"frac_dfma_only\": %.4f, \"frac_with_key_tree\": %.4f, \"frac_with_lds\": %.4f, \"frac_with_checkpoint\": %.4f}\n",
               w, w / 4.0, clk / 1e6, run_one<0>(w, iters, p.multiProcessorCount, clk), run_one<1>(w, iters, p.multiProcessorCount, clk),
               run_one<2>(w, iters, p.multiProcessorCount, clk), run_one<3>(w, iters, p.multiProcessorCount, clk));
        fflush(stdout);
    }
    return 0;
}
