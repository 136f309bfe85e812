// ssszip_gapped encode <input_file> <output_file> | decode <input_file> <output_file> — the gapped container of the
// reference's ssszip (cli/ssszip.cpp:120-180, encode_gapped): factorize_approximate<skip_phrases, lpf_opt> on the GPU, then
//   1 byte: positions are 64 bit;  8 bytes: text length;  then alternating
//   gap   : vbyte(gap length) vbyte(0) + the gap's raw bytes      (phrases shorter than min_lpf_len are folded into gaps)
//   phrase: vbyte(distance = position - source) vbyte(length)
// with the variable-byte integers of misc/vbyte.hpp.  (ssszip pipes this container through an external compressor; that
// part is outside the hot path.)  decode restores the text from the container on the host: it is the format check.
#include <lz77_sss/lz77_sss.hpp>
#include <lz77_sss/misc/vbyte.hpp>

#include <filesystem>
#include <fstream>
#include <iostream>

static constexpr uint64_t min_lpf_len = 64;       // cli/ssszip.cpp:37: shorter phrases inside a gap stay part of the gap

template <typename pos_t>
void encode(const std::string& input, std::ofstream& out) {
    using factor = typename lz77_sss<pos_t>::factor;
    const bool wide = sizeof(pos_t) == 8;
    const uint64_t n = input.size();
    out.write(reinterpret_cast<const char*>(&wide), 1);
    out.write(reinterpret_cast<const char*>(&n), 8);
    pos_t i = 0, gap_beg = 0, gap_len = 0;
    auto close_gap = [&]() {
        if (!gap_len) return;
        encode_vbyte<pos_t>(out, gap_len);
        encode_vbyte<pos_t>(out, 0);
        out.write(&input[gap_beg], gap_len);
        gap_len = 0;
    };
    // skip_phrases emits {gap length, 0} for gaps and {source, length} for phrases (factorize/skip_gaps.cpp:34-61)
    auto output = [&](factor f) {
        const pos_t adv = f.len == 0 ? f.src : f.len;
        if (f.len == 0 || (gap_len && f.len < min_lpf_len)) {
            if (!gap_len) gap_beg = i;
            gap_len += adv;
        } else {
            close_gap();
            encode_vbyte<pos_t>(out, i - f.src);
            encode_vbyte<pos_t>(out, f.len);
        }
        i += adv;
    };
    lz77_sss<pos_t>::template factorize_approximate<skip_phrases, lpf_opt>(input.data(), (pos_t)n, output, {.num_threads = 1});
    close_gap();
}

template <typename pos_t>
std::string decode(std::ifstream& in, uint64_t n) {
    std::string T;
    T.reserve(n);
    while (T.size() < n) {
        const pos_t a = decode_vbyte<pos_t>(in), b = decode_vbyte<pos_t>(in);
        if (b == 0) {
            const size_t at = T.size();
            T.resize(at + a);
            in.read(&T[at], a);
        } else {
            const size_t from = T.size() - a;
            for (pos_t k = 0; k < b; k++) T.push_back(T[from + k]);
        }
    }
    return T;
}

int main(int argc, char** argv) {
    if (argc != 4) { std::cout << "usage: ssszip_gapped encode|decode <input_file> <output_file>" << std::endl; return 2; }
    const std::string cmd = argv[1];
    std::ifstream in(argv[2], std::ios::binary);
    std::ofstream out(argv[3], std::ios::binary);
    if (!in.good() || !out.good()) { std::cout << "error: cannot open the files" << std::endl; return 1; }
    if (cmd == "encode") {
        const uint64_t n = std::filesystem::file_size(argv[2]);
        std::string T(n, '\0');
        in.read(T.data(), (std::streamsize)n);
        if (n <= 0xFFFFFFFFull) encode<uint32_t>(T, out);
        else encode<uint64_t>(T, out);
    } else {
        bool wide = false;
        uint64_t n = 0;
        in.read(reinterpret_cast<char*>(&wide), 1);
        in.read(reinterpret_cast<char*>(&n), 8);
        const std::string T = wide ? decode<uint64_t>(in, n) : decode<uint32_t>(in, n);
        out.write(T.data(), (std::streamsize)T.size());
    }
    return 0;
}
