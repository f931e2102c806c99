// ref_shim_tiled.cpp -- TEST INFRASTRUCTURE ONLY: the reference's Tiled kernel behind a C ABI.
#include "ref_shim.hpp"

namespace MatrixProfileTests {
// Wrapper idiom of test/TestTiledKernel.cpp:14-27 (t is a free name the kernel source expects).
template <typename data_t, typename index_t, int w, int t>
struct RefTiled : public MatrixProfileKernel<data_t, index_t, w> {
    using Base = MatrixProfileKernel<data_t, index_t, w>;
    using aggregate_t = typename Base::aggregate_t;
    using InputDataPack = typename Base::InputDataPack;
    using OutputDataPack = typename Base::OutputDataPack;
    static constexpr data_t aggregate_init = Base::aggregate_init;
    static constexpr index_t index_init = Base::index_init;
    static constexpr index_t nColumns = Base::nColumns;
#include "MatrixProfileKernelTiled.cpp"     // reference: kernel/
};
}  // namespace MatrixProfileTests

using MatrixProfileTests::RefTiled;
#define WT_LIST(X) X(32, 16) X(64, 16) X(128, 32) X(256, 128) X(512, 128) X(1024, 128) X(1024, 256) X(4096, 128) X(4096, 256)

extern "C" int ref_tiled_f64(int w, int t, int n, int m, int iteration, const void *in, void *out) {
#define X(W, T) if (w == W && t == T) { int rc = refshim::run_chunk<RefTiled<double, int, W, T>>(n, m, iteration, in, out); return rc | (Mock::all_streams_empty ? 0 : 2); }
    WT_LIST(X)
#undef X
    return -1;
}
extern "C" int ref_tiled_f32(int w, int t, int n, int m, int iteration, const void *in, void *out) {
#define X(W, T) if (w == W && t == T) { int rc = refshim::run_chunk<RefTiled<float, int, W, T>>(n, m, iteration, in, out); return rc | (Mock::all_streams_empty ? 0 : 2); }
    WT_LIST(X)
#undef X
    return -1;
}
