// ref_shim.hpp -- TEST INFRASTRUCTURE ONLY.
// Glue that lets the *unmodified* reference sources under $(MP_REFERENCE_DIR) (default
// /root/reference) be compiled as plain C++ and called through a C ABI, so that the C oracle
// (mp_oracle.c) and the CUDA engine can be checked against the reference's own code.
// No reference source is copied: the kernel .cpp files are #included from where they lie, one
// kernel per translation unit (their support headers use #pragma once), exactly the way the
// reference's own tests consume them (test/TestVanillaKernel.cpp:14-27, test/TestTiledKernel.cpp:14-27).
#pragma once
#include "MatrixProfileTests.hpp"   // reference: test/include (defines TEST_MOCK_SW, Mock::stream, the
                                    // abstract MatrixProfileKernel<data_t,index_t,w>, PrecomputeStatistics)
#include <cstdint>
#include <cstring>

namespace refshim {
// Calls kernel K on raw buffers whose layout is include/Config.hpp.in:43-73 ( == the nested types
// of MatrixProfileTests::MatrixProfileKernel, test/include/MatrixProfileTests.hpp:42-63 ).
template <typename K>
int run_chunk(int n, int m, int iteration, const void *in, void *out) {
    K kernel;
    Mock::Reset();
    kernel.MatrixProfileKernelTLF(n, m, iteration, static_cast<const typename K::InputDataPack *>(in),
                                  static_cast<typename K::OutputDataPack *>(out));
    return Mock::read_from_empty_stream ? 1 : 0;
}
}  // namespace refshim
