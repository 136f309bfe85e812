// lz77_sss_3_aprx <input_file> <output_file> [threads] — file-level drop-in for the reference's cli/lz77_sss_3_aprx.cpp:28-90:
// reads the text, runs factorize_approximate<greedy, lpf_opt> on the GPU through the header mirror and writes the
// factors in the reference's Huffman file format (misc/huffman.hpp), readable by the reference's lz77-sss-decode.
#include <lz77_sss/lz77_sss.hpp>
#include <lz77_sss/misc/huffman.hpp>

#include <chrono>
#include <filesystem>
#include <fstream>
#include <iostream>

int main(int argc, char** argv) {
    if (argc < 3 || argc > 4) {
        std::cout << "usage: lz77_sss_3_aprx <input_file> <output_file> <threads>\n       the last parameter is optional" << std::endl;
        return 2;
    }
    std::ifstream in(argv[1], std::ios::binary);
    std::ofstream out(argv[2], std::ios::binary);
    if (!in.good()) { std::cout << "error: could not read <input_file>" << std::endl; return 1; }
    if (!out.good()) { std::cout << "error: could not write to <output_file>" << std::endl; return 1; }
    const uint16_t p = argc == 4 ? (uint16_t)std::atoi(argv[3]) : 0;
    const uint64_t n = std::filesystem::file_size(argv[1]);
    std::string T(n, '\0');
    in.read(T.data(), (std::streamsize)n);
    const auto t0 = std::chrono::steady_clock::now();
    huff_writer writer(out, n);
    uint64_t z = 0;
    if (n <= 0xFFFFFFFFull)
        lz77_sss<uint32_t>::factorize_approximate<greedy, lpf_opt>(T.data(), (uint32_t)n, [&](auto f) { writer.add(f); z++; },
                                                                    {.num_threads = p, .log = true});
    else
        lz77_sss<uint64_t>::factorize_approximate<greedy, lpf_opt>(T.data(), n, [&](auto f) { writer.add(f); z++; },
                                                                    {.num_threads = p, .log = true});
    writer.finish();
    out.close();
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    const uint64_t bytes = std::filesystem::file_size(argv[2]);
    std::cout << "num. of factors: " << z << "\noutput file size: " << bytes << " bytes\ncompression ratio: " << n / (double)bytes
              << "\ntime: " << sec << " s" << std::endl;
    return 0;
}
