// factor_file — the reference's file formats through the B200 build's host headers (no GPU needed).
//   factor_file huff-write  <factors.bin> <n> <out.lz77>      factors.bin: z pairs of little-endian u64 (src, len)
//   factor_file huff-read   <in.lz77> <factors.bin>           prints n and z
//   factor_file vbyte-write <values.bin> <32|64> <out>        values.bin: u64 values
//   factor_file vbyte-read  <in> <32|64> <values.bin>
// Used by tests/test_formats_cpu.py against files written by the reference's own headers (oracle/_ref) and against the
// committed fixtures under tests/golden/formats/.
#include <lz77_sss/misc/huffman.hpp>
#include <lz77_sss/misc/vbyte.hpp>

#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

struct factor64 {
    uint64_t src, len;
    uint64_t length() const { return len ? len : 1; }
};

static std::vector<uint64_t> read_u64(const char* path) {
    std::ifstream in(path, std::ios::binary | std::ios::ate);
    if (!in.good()) { std::fprintf(stderr, "cannot read %s\n", path); std::exit(2); }
    const size_t bytes = (size_t)in.tellg();
    std::vector<uint64_t> v(bytes / 8);
    in.seekg(0);
    in.read(reinterpret_cast<char*>(v.data()), (std::streamsize)(v.size() * 8));
    return v;
}

int main(int argc, char** argv) {
    if (argc < 4) { std::fprintf(stderr, "usage: see the head of tools/cpp/factor_file.cpp\n"); return 2; }
    const std::string cmd = argv[1];
    if (cmd == "huff-write" && argc == 5) {
        const auto f = read_u64(argv[2]);
        const uint64_t n = std::strtoull(argv[3], nullptr, 10);
        std::ofstream out(argv[4], std::ios::binary);
        huff_writer w(out, n);
        for (size_t k = 0; k + 1 < f.size(); k += 2) w.add(factor64{f[k], f[k + 1]});
        w.finish();
        return out.good() ? 0 : 1;
    }
    if (cmd == "huff-read" && argc == 4) {
        std::ifstream in(argv[2], std::ios::binary);
        if (!in.good()) return 2;
        uint64_t n = 0;
        in.read(reinterpret_cast<char*>(&n), 5);
        bit_reader r(in);
        huffman lh, dh;
        huff_factor_iterator<factor64> it(r, lh, dh, n);
        std::ofstream out(argv[3], std::ios::binary);
        uint64_t z = 0;
        for (uint64_t pos = 0; pos < n; z++) {
            const factor64 f = *it++;
            out.write(reinterpret_cast<const char*>(&f), 16);
            pos += f.length();
        }
        std::printf("%llu %llu\n", (unsigned long long)n, (unsigned long long)z);
        return 0;
    }
    if (cmd == "vbyte-write" && argc == 5) {
        const auto v = read_u64(argv[2]);
        std::ofstream out(argv[4], std::ios::binary);
        for (uint64_t x : v) {
            if (std::atoi(argv[3]) == 32) encode_vbyte<uint32_t>(out, (uint32_t)x);
            else encode_vbyte<uint64_t>(out, x);
        }
        return 0;
    }
    if (cmd == "vbyte-read" && argc == 5) {
        std::ifstream in(argv[2], std::ios::binary);
        std::ofstream out(argv[4], std::ios::binary);
        while (in.peek() != EOF) {
            const uint64_t x = std::atoi(argv[3]) == 32 ? decode_vbyte<uint32_t>(in) : decode_vbyte<uint64_t>(in);
            out.write(reinterpret_cast<const char*>(&x), 8);
        }
        return 0;
    }
    std::fprintf(stderr, "bad arguments\n");
    return 2;
}
