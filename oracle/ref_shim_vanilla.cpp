// ref_shim_vanilla.cpp -- TEST INFRASTRUCTURE ONLY: the reference's Vanilla kernel behind a C ABI.
#include "ref_shim.hpp"

namespace MatrixProfileTests {
// Same wrapper idiom as test/TestVanillaKernel.cpp:14-27: the kernel source is textually included
// inside a struct that supplies data_t / index_t / nColumns and the record types.
template <typename data_t, typename index_t, int w>
struct RefVanilla : public MatrixProfileKernel<data_t, index_t, w> {
    using Base = MatrixProfileKernel<data_t, index_t, w>;
    using aggregate_t = typename Base::aggregate_t;
    using InputDataPack = typename Base::InputDataPack;
    using OutputDataPack = typename Base::OutputDataPack;
    static constexpr data_t aggregate_init = Base::aggregate_init;
    static constexpr index_t index_init = Base::index_init;
    static constexpr index_t nColumns = Base::nColumns;
#include "MatrixProfileKernelVanilla.cpp"   // reference: kernel/
};
}  // namespace MatrixProfileTests

using MatrixProfileTests::RefVanilla;
#define W_LIST(X) X(8) X(16) X(32) X(64) X(128) X(256) X(512) X(1024) X(2048) X(4096)

extern "C" int ref_vanilla_f64(int w, int n, int m, int iteration, const void *in, void *out) {
    switch (w) {
#define X(W) case W: return refshim::run_chunk<RefVanilla<double, int, W>>(n, m, iteration, in, out);
        W_LIST(X)
#undef X
    }
    return -1;
}
extern "C" int ref_vanilla_f32(int w, int n, int m, int iteration, const void *in, void *out) {
    switch (w) {
#define X(W) case W: return refshim::run_chunk<RefVanilla<float, int, W>>(n, m, iteration, in, out);
        W_LIST(X)
#undef X
    }
    return -1;
}
