// huffman.hpp — the reference's on-disk factor format (row f1 of SURVEY.md §8), host side of the B200 build.
//
// Same class names and public members as /root/reference/include/lz77_sss/misc/huffman.hpp:42-436 (bit_writer,
// bit_reader, put/get_elias_delta, huffman, huff_writer, huff_factor_iterator), so the reference's CLIs
// (cli/lz77_sss_3_aprx.cpp:73-84, cli/lz77_sss_decode.cpp:61-66) compile against it, and byte-identical files:
//   file   = n as 5 little-endian bytes, then blocks until the factors end, zero bits up to the last byte
//   block  = Elias-delta(count <= 2^14), 66 x 4-bit code lengths of the length classes, 66 x 4-bit code lengths of the
//            distance classes, then per factor:  class(len) [0 = literal: 8 raw bits of the byte;
//            k > 0: the k-1 low bits of len], and for a copy class(dist) + the low bits of dist = pos - src
//   class  = bit_width(value); codes = canonical Huffman, code lengths limited to 15
// Bits are packed most-significant first.  The code LENGTHS depend on tie-breaking inside the reference's construction
// (huffman.hpp:184-247: merge order by (weight, node number), depth cap by moving the deepest codes down, lengths
// handed out along std::sort by frequency); this file reproduces those choices so that encoders agree bit for bit.
#pragma once
#include <algorithm>
#include <array>
#include <bit>
#include <cstdint>
#include <cstring>
#include <istream>
#include <iterator>
#include <ostream>
#include <set>
#include <span>
#include <string>
#include <utility>
#include <vector>

// MSB-first bit sink over an ostream
class bit_writer {
    std::ostream& out;
    std::string buf;            // whole bytes not yet handed to the stream
    uint64_t acc = 0;           // pending bits, left-aligned
    unsigned fill = 0;          // number of pending bits (< 64)

    void drain(bool all) {
        if (all || buf.size() >= (1u << 16)) {
            out.write(buf.data(), (std::streamsize)buf.size());
            buf.clear();
        }
    }
    void push_word(uint64_t w, unsigned bytes) {   // the `bytes` most significant bytes of w
        for (unsigned b = 0; b < bytes; b++) buf.push_back((char)(w >> (56 - 8 * b)));
    }

public:
    explicit bit_writer(std::ostream& o) : out(o) { buf.reserve((1u << 16) + 16); }

    // the n low bits of x (n <= 64), most significant first
    void put(uint64_t x, uint64_t n) {
        if (n == 0) return;
        if (n < 64) x &= (uint64_t(1) << n) - 1;
        const unsigned room = 64 - fill;
        if (n < room) {
            acc |= x << (room - n);
            fill += (unsigned)n;
            return;
        }
        // the accumulator fills up: emit 8 bytes, keep the rest
        const unsigned rest = (unsigned)n - room;
        push_word(acc | (rest ? x >> rest : x), 8);
        acc = rest ? x << (64 - rest) : 0;
        fill = rest;
        drain(false);
    }

    // zero-pad to a byte boundary and hand everything to the stream
    void flush() {
        push_word(acc, (fill + 7) / 8);
        acc = 0;
        fill = 0;
        drain(true);
    }
};

// MSB-first bit source over an istream; reading past the end yields zero bits
class bit_reader {
    std::istream& in;
    std::vector<unsigned char> buf;
    size_t at = 0, have = 0;
    uint64_t acc = 0;           // unread bits, left-aligned
    unsigned fill = 0;

    bool next_byte(unsigned& b) {
        if (at == have) {
            in.read(reinterpret_cast<char*>(buf.data()), (std::streamsize)buf.size());
            have = (size_t)in.gcount();
            at = 0;
            if (!have) return false;
        }
        b = buf[at++];
        return true;
    }
    void top_up() {             // at least 57 bits unless the stream ends
        unsigned b;
        while (fill <= 56 && next_byte(b)) {
            acc |= uint64_t(b) << (56 - fill);
            fill += 8;
        }
    }

public:
    explicit bit_reader(std::istream& i) : in(i), buf(1u << 16) {}

    template <typename T>
    T get_bit() {
        if (fill == 0) top_up();
        const T bit = (T)(acc >> 63);
        acc <<= 1;
        if (fill) fill--;
        return bit;
    }

    uint64_t get(uint64_t n) {
        if (n == 0) return 0;
        if (n > 56) {           // two pieces: the accumulator is topped up to at least 57 bits at a time
            const uint64_t hi = get(n - 32);
            return (hi << 32) | get(32);
        }
        if (fill < n) top_up();
        const uint64_t v = acc >> (64 - n);
        acc <<= n;
        fill = fill >= n ? fill - (unsigned)n : 0;
        return v;
    }
};

// Elias delta code of x >= 1 (x == 0 writes nothing, like the reference)
inline void put_elias_delta(bit_writer& out, uint64_t x) {
    if (x == 0) return;
    const uint64_t w = std::bit_width(x);          // x = 1 followed by w-1 bits
    const uint64_t ww = std::bit_width(w);
    out.put(0, ww - 1);                            // unary part
    out.put(w, ww);                                // w, starting with its leading one
    out.put(x, w - 1);                             // x without its leading one
}

inline uint64_t get_elias_delta(bit_reader& in) {
    uint64_t zeros = 0;
    while (zeros < 64 && in.get_bit<uint8_t>() == 0) zeros++;   // (a stream that ends early reads as zero bits)
    const uint64_t w = (uint64_t(1) << zeros) | in.get(zeros);
    if (w <= 1) return w;
    return (uint64_t(1) << (w - 1)) | in.get(w - 1);
}

// canonical, length-limited (15) Huffman code over a small alphabet
class huffman {
public:
    static constexpr uint64_t max_len = 15;

private:
    std::vector<uint8_t> len_of;                            // code length per symbol, 0 = unused
    std::vector<uint16_t> code_of;
    std::array<uint32_t, max_len + 2> count{};              // codes per length
    std::array<uint32_t, max_len + 2> first_code{};         // smallest code of a length
    std::array<uint32_t, max_len + 2> first_rank{};         // its index in `by_code`
    std::vector<uint16_t> by_code;                          // symbols by (length, symbol)

    void assign_codes() {
        const size_t sigma = len_of.size();
        count.fill(0);
        for (uint8_t l : len_of)
            if (l) count[l]++;
        uint32_t code = 0, rank = 0;
        for (uint64_t l = 1; l <= max_len; l++) {
            code = (code + count[l - 1]) << 1;
            first_code[l] = code;
            first_rank[l] = rank;
            rank += count[l];
        }
        code_of.assign(sigma, 0);
        by_code.assign(rank, 0);
        std::array<uint32_t, max_len + 2> used{};
        for (size_t s = 0; s < sigma; s++) {
            const uint8_t l = len_of[s];
            if (!l) continue;
            code_of[s] = (uint16_t)(first_code[l] + used[l]);
            by_code[first_rank[l] + used[l]] = (uint16_t)s;
            used[l]++;
        }
    }

public:
    void build_from_freq(std::span<const uint64_t> freq) {
        const size_t sigma = freq.size();
        len_of.assign(sigma, 0);
        std::vector<uint64_t> live;
        for (size_t s = 0; s < sigma; s++)
            if (freq[s]) live.push_back(s);
        if (live.size() == 1) len_of[live[0]] = 1;
        if (live.size() < 2) { assign_codes(); return; }

        // Huffman tree: always merge the two smallest (weight, node number); a merged node gets the next number
        const size_t m = live.size();
        std::vector<uint64_t> weight(m);
        std::vector<uint32_t> parent(2 * m, 0);
        std::set<std::pair<uint64_t, uint32_t>> open;
        for (uint32_t k = 0; k < m; k++) {
            weight[k] = freq[live[k]];
            open.insert({weight[k], k});
        }
        uint32_t next = (uint32_t)m;
        while (open.size() > 1) {
            const auto a = *open.begin();
            open.erase(open.begin());
            const auto b = *open.begin();
            open.erase(open.begin());
            parent[a.second] = parent[b.second] = next;
            open.insert({a.first + b.first, next});
            next++;
        }
        const uint32_t root = next - 1;
        // natural depths, capped; the cap may break the Kraft sum, which is repaired by pushing codes of the largest
        // length below the cap one level down until the code is feasible again
        std::array<uint64_t, max_len + 2> per_len{};
        for (uint32_t k = 0; k < m; k++) {
            uint64_t d = 0;
            for (uint32_t v = k; v != root; v = parent[v]) d++;
            per_len[std::min<uint64_t>(std::max<uint64_t>(d, 1), max_len)]++;
        }
        uint64_t kraft = 0;
        for (uint64_t l = 1; l <= max_len; l++) kraft += per_len[l] << (max_len - l);
        while (kraft > (uint64_t(1) << max_len)) {
            uint64_t l = max_len - 1;
            while (l >= 1 && per_len[l] == 0) l--;
            per_len[l]--;
            per_len[l + 1]++;
            kraft -= uint64_t(1) << (max_len - l - 1);
        }
        // lengths are handed out along the order of std::sort by frequency: the rarest symbols get the longest codes
        std::sort(live.begin(), live.end(), [&](uint64_t a, uint64_t b) { return freq[a] < freq[b]; });
        size_t k = 0;
        for (uint64_t l = max_len; l >= 1; l--)
            for (uint64_t c = 0; c < per_len[l]; c++) len_of[live[k++]] = (uint8_t)l;
        assign_codes();
    }

    void write_table(bit_writer& out) const {
        for (uint8_t l : len_of) out.put(l, 4);
    }

    void read_table(bit_reader& in, uint64_t alphabet) {
        len_of.assign(alphabet, 0);
        for (auto& l : len_of) l = (uint8_t)in.get(4);
        assign_codes();
    }

    void encode(bit_writer& out, uint64_t sym) const { out.put(code_of[sym], len_of[sym]); }

    uint64_t decode(bit_reader& in) const {
        uint32_t c = 0;
        for (uint64_t l = 1; l <= max_len; l++) {
            c = (c << 1) | in.get_bit<uint32_t>();
            if (count[l] && c - first_code[l] < count[l]) return by_code[first_rank[l] + (c - first_code[l])];
        }
        return 0;
    }
};

static constexpr uint32_t huff_sigma = 66;                 // classes 0..65: bit widths of values below 2^65
static constexpr uint64_t huff_block_size = 1 << 14;

// factor stream -> file (cli/lz77_sss_3_aprx.cpp:73-84: construct with the text length, add() every factor, finish())
class huff_writer {
    bit_writer out;
    std::vector<std::pair<uint64_t, uint64_t>> pending;    // (distance or literal byte, length)
    uint64_t pos = 0;

    void emit_block() {
        if (pending.empty()) return;
        std::array<uint64_t, huff_sigma> len_freq{}, dist_freq{};
        for (const auto& [val, len] : pending) {
            len_freq[len ? std::bit_width(len) : 0]++;
            if (len) dist_freq[std::bit_width(val)]++;
        }
        huffman len_code, dist_code;
        len_code.build_from_freq(len_freq);
        dist_code.build_from_freq(dist_freq);
        put_elias_delta(out, pending.size());
        len_code.write_table(out);
        dist_code.write_table(out);
        for (const auto& [val, len] : pending) {
            const uint64_t lc = len ? std::bit_width(len) : 0;
            len_code.encode(out, lc);
            if (!lc) {
                out.put(val, 8);
                continue;
            }
            out.put(len, lc - 1);
            const uint64_t dc = std::bit_width(val);
            dist_code.encode(out, dc);
            out.put(val, dc - 1);
        }
        pending.clear();
    }

public:
    huff_writer(std::ostream& o, uint64_t n) : out(o) {
        for (int b = 0; b < 5; b++) o.put((char)(n >> (8 * b)));
        pending.reserve(huff_block_size);
    }

    template <typename factor_t>
    void add(const factor_t& f) {
        if (f.len == 0) {
            pending.emplace_back((uint64_t)f.src & 0xFFu, 0);
            pos += 1;
        } else {
            pending.emplace_back(pos - (uint64_t)f.src, (uint64_t)f.len);
            pos += f.len;
        }
        if (pending.size() == huff_block_size) emit_block();
    }

    void finish() {
        emit_block();
        out.flush();
    }
};

// file -> factor stream (cli/lz77_sss_decode.cpp:61-66): input iterator over the factors of a text of length n; the
// caller reads the 5 length bytes, constructs the reader and two code objects, and iterates until n characters are out
template <typename factor_t>
class huff_factor_iterator {
    bit_reader* r = nullptr;
    huffman* len_code = nullptr;
    huffman* dist_code = nullptr;
    uint64_t left_in_block = 0, pos = 0, n = 0;
    factor_t cur{};

    void advance() {
        if (pos >= n) return;
        if (!left_in_block) {
            left_in_block = get_elias_delta(*r);
            len_code->read_table(*r, huff_sigma);
            dist_code->read_table(*r, huff_sigma);
        }
        const uint64_t lc = len_code->decode(*r);
        if (!lc) {
            cur.src = (decltype(cur.src))r->get(8);
            cur.len = 0;
            pos += 1;
        } else {
            const uint64_t len = (uint64_t(1) << (lc - 1)) | r->get(lc - 1);
            const uint64_t dc = dist_code->decode(*r);
            const uint64_t dist = (uint64_t(1) << (dc - 1)) | r->get(dc - 1);
            cur.len = (decltype(cur.len))len;
            cur.src = (decltype(cur.src))(pos - dist);
            pos += len;
        }
        left_in_block--;
    }

public:
    using iterator_category = std::input_iterator_tag;
    using value_type = factor_t;
    using difference_type = std::ptrdiff_t;
    using pointer = const factor_t*;
    using reference = const factor_t&;

    huff_factor_iterator() = default;
    huff_factor_iterator(bit_reader& reader, huffman& lh, huffman& dh, uint64_t text_len)
        : r(&reader), len_code(&lh), dist_code(&dh), n(text_len) { advance(); }

    reference operator*() const { return cur; }
    huff_factor_iterator& operator++() { advance(); return *this; }
    huff_factor_iterator operator++(int) { auto old = *this; advance(); return old; }
    bool operator==(const huff_factor_iterator& o) const { return pos == o.pos && n == o.n; }
};
