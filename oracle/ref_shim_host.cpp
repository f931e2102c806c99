// ref_shim_host.cpp -- TEST INFRASTRUCTURE ONLY: the reference's test-side precompute
// (test/include/MatrixProfileTests.hpp:117-139, the verbatim twin of
// include/host/HostSideComputation.hpp:20-45 for data_t = double) and its independent SCAMP
// (test/include/MatrixProfileReference.hpp:90-136) behind a C ABI.  Both are templated on (n, m),
// so a fixed list of sizes is instantiated; unknown sizes return -1.
#include "ref_shim.hpp"
#include <memory>

namespace {
using Pack = MatrixProfileTests::MatrixProfileKernel<double, int, 8>::InputDataPack;

template <int N, int M>
int precompute(const double *T, void *out) {
    auto t = std::make_unique<std::array<double, N>>();
    auto d = std::make_unique<std::array<Pack, N - M + 1>>();
    std::memcpy(t->data(), T, sizeof(double) * N);
    MatrixProfileTests::PrecomputeStatistics<double, int, N, M, Pack>(*t, *d);
    std::memcpy(out, d->data(), sizeof(Pack) * (N - M + 1));
    return 0;
}
template <int N, int M>
int scamp(const double *T, double *MP, int *MPI) {
    auto t = std::make_unique<std::array<double, N>>();
    auto mp = std::make_unique<std::array<double, N - M + 1>>();
    auto mpi = std::make_unique<std::array<int, N - M + 1>>();
    std::memcpy(t->data(), T, sizeof(double) * N);
    Reference::ComputeMatrixProfile<double, int, N, M>(*t, *mp, *mpi);
    std::memcpy(MP, mp->data(), sizeof(double) * (N - M + 1));
    std::memcpy(MPI, mpi->data(), sizeof(int) * (N - M + 1));
    return 0;
}
}  // namespace

#define NM_LIST(X) X(128, 16) X(1024, 16) X(1024, 30) X(1024, 32) X(2048, 64) X(4096, 64) X(8192, 100) X(16384, 128) X(65536, 256) X(32768, 1024) X(20480, 4096) X(20480, 32)

extern "C" int ref_precompute_f64(int n, int m, const double *T, void *out) {
#define X(N, M) if (n == N && m == M) return precompute<N, M>(T, out);
    NM_LIST(X)
#undef X
    return -1;
}
extern "C" int ref_scamp_f64(int n, int m, const double *T, double *MP, int *MPI) {
#define X(N, M) if (n == N && m == M) return scamp<N, M>(T, MP, MPI);
    NM_LIST(X)
#undef X
    return -1;
}
