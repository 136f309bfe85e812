// lz77_sss_decode <input_file> <output_file> — file-level drop-in for the reference's cli/lz77_sss_decode.cpp: reads a
// factor file in the reference's Huffman format and decodes it.  The reference streams through a ring buffer and file
// seeks because a sequential decoder needs its own output as input; here the factors go to the GPU and the whole text is
// resolved in HBM (lz77_sss<>::decode, csrc/decode.cu), then written once.
#include <lz77_sss/lz77_sss.hpp>
#include <lz77_sss/misc/huffman.hpp>

#include <fstream>
#include <iostream>

int main(int argc, char** argv) {
    if (argc != 3) { std::cout << "usage: lz77_sss_decode <input_file> <output_file>" << std::endl; return 2; }
    std::ifstream in(argv[1], std::ios::binary);
    if (!in.good()) { std::cout << "error: could not read <input_file>" << std::endl; return 1; }
    uint64_t n = 0;
    in.read(reinterpret_cast<char*>(&n), 5);
    bit_reader reader(in);
    huffman len_huff, dist_huff;
    using factor = lz77_sss<uint64_t>::factor;
    huff_factor_iterator<factor> it(reader, len_huff, dist_huff, n);
    std::string T(n, '\0');
    lz77_sss<uint64_t>::decode(it, T.data(), n);
    std::ofstream out(argv[2], std::ios::binary);
    if (!out.good()) { std::cout << "error: could not write to <output_file>" << std::endl; return 1; }
    out.write(T.data(), (std::streamsize)n);
    std::cout << "decoded " << n << " bytes" << std::endl;
    return 0;
}
