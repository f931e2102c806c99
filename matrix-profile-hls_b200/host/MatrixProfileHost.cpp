// MatrixProfileHost -- C++ host driver of the B200 matrix-profile engine.
//
// Same command line, same files and same report as the reference driver host/MatrixProfileHost.cpp:
//   -b,--xclbin  -i,--input  -o,--output  --verbose  -v,--version  -h,--help          (:166-175)
//   input  : raw little-endian float64 series (.tsb, include/host/FileIO.hpp:24-48)
//   output : <out>.mpb (float64 distances) and <out>.mpib (int32 indices)  (:144-153, FileIO.hpp:105-121)
// The OpenCL/XRT wrapper (include/host/OpenCL.hpp) is replaced by the C ABI in include/mp_b200.h; the
// chunk loop + UpdateAggregates + PostComputeAggregates (:101-142) are three kernels behind mpb_plan_*.
// Configure-time constants keep their reference names (CMakeLists.txt:28-46): MP_SIZE_N / MP_SIZE_M /
// MP_DATA_TYPE / MP_KERNEL / MP_SIZE_W / MP_SIZE_T / MP_KERNEL_REPL arrive as -D macros (Config.hpp.in or the
// Makefile).  MP_SIZE_N = 0 means "take n from the file size" (the reference needs one binary per size).
// There is no CPU fallback: without a usable CUDA device the driver prints [ERROR] and exits with failure.
#include "../../include/mp_b200.h"
#ifdef MP_HAVE_CONFIG_HPP
#include "Config.hpp"
#endif

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <optional>
#include <string>
#include <vector>

#ifndef MP_SIZE_N
#define MP_SIZE_N 0
#endif
#ifndef MP_SIZE_M
#define MP_SIZE_M 256
#endif
#ifndef MP_SIZE_W
#define MP_SIZE_W 4096
#endif
#ifndef MP_SIZE_T
#define MP_SIZE_T 128
#endif
#ifndef MP_KERNEL_REPL
#define MP_KERNEL_REPL 1
#endif
#ifndef MP_DATA_TYPE_IS_FLOAT
#define MP_DATA_TYPE_IS_FLOAT 0
#endif
#ifndef MP_KERNEL_NAME
#define MP_KERNEL_NAME "Tiled"
#endif

namespace {

// ---- Logger: the reference's prefixes (include/host/Logger.hpp:30-58) ----
bool g_verbose = false;
enum class Level { Info, Verbose, Warning, Error, Debug };
void Log(Level lvl, const std::string &msg) {
    switch (lvl) {
        case Level::Info: std::cout << "[INFO] " << msg << std::endl; break;
        case Level::Verbose: if (g_verbose) std::cout << "[INFO/V] " << msg << std::endl; break;
        case Level::Warning: std::cout << "[WARN] " << msg << std::endl; break;
        case Level::Error: std::cerr << "[ERROR] " << msg << std::endl; break;
        case Level::Debug: std::cout << "[DEBUG] " << msg << std::endl; break;
    }
}

// ---- BenchmarkProfile: named categories with percentages (include/host/BenchmarkProfile.hpp:18-57) ----
struct BenchmarkProfile {
    std::vector<std::string> order;
    std::map<std::string, std::vector<std::pair<std::string, double>>> entries;   // category -> (name, ms)
    void Push(const std::string &category, const std::string &name, double ms) {
        if (!entries.count(category)) order.push_back(category);
        entries[category].emplace_back(name, ms);
    }
    void Report() const {
        double total = 0;
        for (const auto &c : entries) for (const auto &e : c.second) total += e.second;
        std::cout << std::endl << "Benchmark Profile (total " << total << " ms)" << std::endl;
        for (const auto &cat : order) {
            double sum = 0;
            for (const auto &e : entries.at(cat)) sum += e.second;
            std::printf("  %-58s %12.3f ms  %6.2f %%\n", cat.c_str(), sum, total > 0 ? 100.0 * sum / total : 0.0);
            if (g_verbose)
                for (const auto &e : entries.at(cat)) std::printf("      %-54s %12.3f ms\n", e.first.c_str(), e.second);
        }
    }
};

struct Timer {
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    double ElapsedMs() const { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); }
};

// ---- FileIO: raw arrays, no header (include/host/FileIO.hpp:24-48, 105-121) ----
bool ReadBinaryFile(const std::string &path, std::vector<double> &out, long long expect_n) {
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    if (!f) { Log(Level::Error, "Unable to open input file \"" + path + "\""); return false; }
    const long long bytes = (long long)f.tellg();
    if (bytes % (long long)sizeof(double) != 0) { Log(Level::Error, "Input file size is not a multiple of 8 bytes (float64 series expected)"); return false; }
    const long long n = bytes / (long long)sizeof(double);
    if (expect_n > 0 && n != expect_n) {   // FileIO.hpp:38-42: the file must hold exactly n values
        Log(Level::Error, "Input file holds " + std::to_string(n) + " values, but this binary was configured for MP_SIZE_N=" + std::to_string(expect_n));
        return false;
    }
    out.resize((size_t)n);
    f.seekg(0);
    f.read(reinterpret_cast<char *>(out.data()), bytes);
    return (bool)f;
}
template <typename T>
bool WriteBinaryFile(const std::string &path, const std::vector<T> &v) {
    std::ofstream f(path, std::ios::binary);
    if (!f) { Log(Level::Error, "Unable to open output file \"" + path + "\""); return false; }
    f.write(reinterpret_cast<const char *>(v.data()), (std::streamsize)(v.size() * sizeof(T)));
    return (bool)f;
}

const char *kHelp =
    "Matrix Profile Host - (C++/CUDA, B200)\n"
    "Usage:\n"
    "  MatrixProfileHost [OPTION...]\n\n"
    "  -b, --xclbin arg   accepted for compatibility with the FPGA driver (no bitstream is loaded) [optional]\n"
    "  -i, --input arg    input file (time series) [required]\n"
    "  -o, --output arg   output file (matrix profile / matrix profile index)\n"
    "      --verbose      increase output verbosity\n"
    "  -v, --version      prints version information and exits\n"
    "  -h, --help         shows help message and exits\n"
    "  -m, --sublen arg   subsequence length (default: MP_SIZE_M)\n"
    "      --gpus arg     number of GPUs to shard the diagonal bands over (default: MP_KERNEL_REPL)\n"
    "      --float        compute in fp32 (MP_DATA_TYPE=float) with periodic exact QT re-seed\n"
    "      --strict       chain-faithful mode: bit-identical to the reference kernels (slower)\n";

#define MPB_CHECK(call)                                                                   \
    do {                                                                                  \
        const int rc__ = (call);                                                          \
        if (rc__ != MPB_OK) {                                                             \
            Log(Level::Error, std::string(#call) + " failed: " + mpb_last_error());        \
            return EXIT_FAILURE;                                                          \
        }                                                                                 \
    } while (0)

struct Options {
    std::string input;
    std::optional<std::string> output, xclbin;
    int m = MP_SIZE_M;
    int gpus = MP_KERNEL_REPL;
    bool use_float = MP_DATA_TYPE_IS_FLOAT != 0;
    bool strict = false;
};

// The device section of the reference's RunMatrixProfileKernel (host/MatrixProfileHost.cpp:54-160).
int RunMatrixProfileKernel(const Options &opt) {
    BenchmarkProfile profile;
    if (!opt.output) Log(Level::Warning, "No output (-o, --output) parameter provided. Results will be discarded!");   // :57-58

    Log(Level::Verbose, "Loading input time series...");
    std::vector<double> T;
    if (!ReadBinaryFile(opt.input, T, MP_SIZE_N)) return EXIT_FAILURE;                                                  // :62-63
    if (T.size() >= (size_t)1 << 31) { Log(Level::Error, "series too long for index_t = int (CMakeLists.txt:39-41)"); return EXIT_FAILURE; }
    const int n = (int)T.size(), m = opt.m;
    const int sublen = n - m + 1;

    Log(Level::Verbose, "Initializing CUDA context(s)...");
    int ndev = 0;
    MPB_CHECK(mpb_device_count(&ndev));
    const int G = opt.gpus;
    if (G < 1 || G > ndev) { Log(Level::Error, "requested " + std::to_string(G) + " GPU(s), " + std::to_string(ndev) + " available"); return EXIT_FAILURE; }
    const unsigned flags = opt.strict ? MPB_FLAG_STRICT : 0u;
    const int dtype = opt.use_float ? MPB_F32 : MPB_F64;
    std::vector<double> MP((size_t)(sublen > 0 ? sublen : 0));
    std::vector<int32_t> MPI(MP.size());
    std::vector<mpb_timings> tms((size_t)G);

    // The reference copies the input to every compute unit, enqueues the chunks round-robin, and merges on the host
    // (:78-137).  Here: one host thread per GPU (mpb_compute_multi), each uploads the series, pre-computes, walks its
    // diagonal bands, and the profiles are max-merged over NVLink with NCCL; rank 0 converts and downloads.
    Log(Level::Verbose, "Copying memory to device...");                                                                // :89-92
    Log(Level::Info, "Pre-Computing Statistics on Device");                                                           // was: on Host (:65-66)
    Log(Level::Verbose, "Starting Kernel Execution(s)...");
    Timer wall;
    const int rc = mpb_compute_multi(G, nullptr, T.data(), n, m, dtype, flags, MP.data(), MPI.data(), tms.data());    // chunk loop :101-137 + :142
    Log(Level::Info, "Performing Post-Computation on Device");
    if (rc != MPB_OK) {
        Log(Level::Error, std::string("matrix profile computation failed: ") + mpb_last_error());
        return EXIT_FAILURE;
    }
    const double wall_ms = wall.ElapsedMs();
    Log(Level::Info, "Kernel Execution Completed Successfully.");

    double pre = 0, comp = 0, fin = 0, h2d = 0, d2h = 0;
    long long items = 0;
    for (int g = 0; g < G; ++g) {
        const mpb_timings &tm = tms[(size_t)g];
        if (tm.precompute_ms > pre) pre = tm.precompute_ms;
        if (tm.compute_ms > comp) comp = tm.compute_ms;
        if (tm.h2d_ms > h2d) h2d = tm.h2d_ms;
        items += tm.items;
        if (g == 0) { fin = tm.finalize_ms; d2h = tm.d2h_ms; }
    }
    const std::string kname = std::string("v") + MP_KERNEL_NAME + ", w=" + std::to_string(MP_SIZE_W);
    profile.Push("0. Host <-> Device Copies", "series H2D (" + std::to_string(G) + " GPU)", h2d);
    profile.Push("0. Host <-> Device Copies", "profile + index D2H", d2h);
    profile.Push("1. Device-Side [Pre-Computation]", "mean / df / dg / inv / QT seeds", pre);                        // HostSideComputation.hpp:44
    profile.Push("2. GPU Computation [" + kname + "]", "MatrixProfileKernelTLF [" + std::to_string(items) + " tiles on " + std::to_string(G) + " GPU, NCCL merge included]", comp);   // :121-122
    profile.Push("3/4. Device-Side [Update-Aggregates + Post-Computation]", "merge + Pearson -> distance", fin);      // HostSideComputation.hpp:68,89
    if (g_verbose) {
        const double cells = (double)mpb_useful_cells(n, m);
        Log(Level::Verbose, "n=" + std::to_string(n) + " m=" + std::to_string(m) + " useful cells=" + std::to_string((long long)cells) +
                                " -> " + std::to_string(cells / (comp * 1e-3)) + " cells/s (kernel), wall " + std::to_string(wall_ms) + " ms");
    }

    if (opt.output) {
        Log(Level::Verbose, "Saving results (MP/MPI) to file...");
        if (!WriteBinaryFile(*opt.output + ".mpb", MP) || !WriteBinaryFile(*opt.output + ".mpib", MPI)) return EXIT_FAILURE;   // :147-152
    }
    profile.Report();
    Log(Level::Verbose, "Terminating Host.");
    return EXIT_SUCCESS;
}

}  // namespace

int main(int argc, char *argv[]) {
    Options opt;
    bool want_help = false, want_version = false;
    auto value = [&](int &i, const std::string &a, std::string &out) -> bool {   // "--opt value" and "--opt=value"
        const size_t eq = a.find('=');
        if (eq != std::string::npos && a.rfind("--", 0) == 0) { out = a.substr(eq + 1); return true; }
        if (i + 1 >= argc) return false;
        out = argv[++i];
        return true;
    };
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        const std::string key = a.substr(0, a.find('='));
        std::string v;
        bool ok = true;
        if (key == "-h" || key == "--help") want_help = true;
        else if (key == "-v" || key == "--version") want_version = true;
        else if (key == "--verbose") g_verbose = true;
        else if (key == "--float") opt.use_float = true;
        else if (key == "--strict") opt.strict = true;
        else if (key == "-b" || key == "--xclbin") { ok = value(i, a, v); opt.xclbin = v; }
        else if (key == "-i" || key == "--input") { ok = value(i, a, v); opt.input = v; }
        else if (key == "-o" || key == "--output") { ok = value(i, a, v); opt.output = v; }
        else if (key == "-m" || key == "--sublen") { ok = value(i, a, v); opt.m = std::atoi(v.c_str()); }
        else if (key == "--gpus") { ok = value(i, a, v); opt.gpus = std::atoi(v.c_str()); }
        else ok = false;
        if (!ok) {                                                   // reference: :212-216
            Log(Level::Error, "Unknown argument\n");
            std::cout << kHelp << std::endl;
            return EXIT_FAILURE;
        }
    }
    if (want_help) { std::cout << kHelp << std::endl; return EXIT_SUCCESS; }
    if (want_version) {
        std::cout << "Matrix Profile Host (B200) - " << mpb_version() << " [MP_KERNEL=" << MP_KERNEL_NAME << ", MP_DATA_TYPE="
                  << (MP_DATA_TYPE_IS_FLOAT ? "float" : "double") << ", MP_SIZE_N=" << MP_SIZE_N << ", MP_SIZE_M=" << MP_SIZE_M
                  << ", MP_SIZE_W=" << MP_SIZE_W << ", MP_SIZE_T=" << MP_SIZE_T << ", MP_KERNEL_REPL=" << MP_KERNEL_REPL << "]" << std::endl;
        return EXIT_SUCCESS;
    }
    if (opt.input.empty()) {                                         // reference: :196-200
        Log(Level::Error, "--input required\n");
        std::cout << kHelp << std::endl;
        return EXIT_FAILURE;
    }
    return RunMatrixProfileKernel(opt);
}
