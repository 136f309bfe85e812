// lz77_sss_bench_tau <file> <result_file> [min_tau] [max_tau] — the harness shape of the reference's
// bench/lz77_sss_bench_tau.cpp:36-101 on the B200 build: factorize_approximate<greedy, lpf_opt, tau> for tau = 4, 8, ...,
// 4096 at num_threads = 1, one `RESULT key=value ...` line per tau appended to <result_file> (keys of the reference's
// own lines, lz77_sss.hpp:298-311,357-365, so the authors' tooling reads both).
#include <fstream>
#include <string>
std::string text_name;
std::string result_file_path;
std::ofstream result_file;
#define LZ77_SSS_BENCH 1
#include <lz77_sss/lz77_sss.hpp>

#include <filesystem>
#include <iostream>
#include <utility>

template <uint64_t tau>
void run_tau(std::string& T, uint64_t n, uint64_t min_tau, uint64_t max_tau) {
    if (tau < min_tau || tau > max_tau) return;
    std::cout << "running LZ77 SSS 3-approximation with tau = " << tau << std::endl;
    uint64_t z = 0;
    if (n <= 0xFFFFFFFFull)
        lz77_sss<uint32_t>::factorize_approximate<greedy, lpf_opt, tau>(T.data(), (uint32_t)n, [&](auto) { z++; }, {.num_threads = 1, .log = true});
    else
        lz77_sss<uint64_t>::factorize_approximate<greedy, lpf_opt, tau>(T.data(), n, [&](auto) { z++; }, {.num_threads = 1, .log = true});
}

template <size_t... K>
void run_all(std::string& T, uint64_t n, uint64_t min_tau, uint64_t max_tau, std::index_sequence<K...>) {
    (run_tau<(uint64_t(4) << K)>(T, n, min_tau, max_tau), ...);
}

int main(int argc, char** argv) {
    if (argc < 2 || argc > 5) {
        std::cout << "usage: lz77_sss_bench_tau <file> <result_file> <min_tau> <max_tau>\n       the last three parameters are optional" << std::endl;
        return 2;
    }
    const std::string path = argv[1];
    std::ifstream in(path, std::ios::binary);
    if (!in.good()) { std::cout << "error: could not read <file>" << std::endl; return 1; }
    if (argc >= 3) {
        text_name = path.substr(path.find_last_of("/\\") + 1);
        result_file_path = argv[2];
        result_file.open(result_file_path, std::ofstream::app);
    }
    const uint64_t min_tau = argc >= 4 ? std::atol(argv[3]) : 4, max_tau = argc >= 5 ? std::atol(argv[4]) : 4096;
    const uint64_t n = std::filesystem::file_size(path);
    std::string T(n, '\0');
    in.read(T.data(), (std::streamsize)n);
    run_all(T, n, min_tau, max_tau, std::make_index_sequence<11>{});     // 4 << 0 ... 4 << 10
    return 0;
}
