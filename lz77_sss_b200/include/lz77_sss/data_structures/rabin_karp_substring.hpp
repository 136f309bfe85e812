// rabin_karp_substring.hpp — B200 host-side mirror of the reference's fingerprint structure.
//
// Same class template, constructor arguments and member functions as
// /root/reference/include/lz77_sss/data_structures/rabin_karp_substring.hpp:35-240 (used by the exact path's sample
// index with mers_exp = 31, sample_index.hpp:139,373).  The pass over the text — the fingerprints of all prefixes
// T[0..16k) — runs on the GPU (liblz77sss_b200.so, lz77sss_rks_create / lz77sss_rks_table); the scalar queries
// (substring, substring_up_to, push, roll, concat) are answered on the host from that table exactly like the
// reference answers them from its `fps` vector, and substring_batch() sends many queries to the device at once.
//
// Differences a caller can see:
//  * only mers_exp = 31 (the instantiation the reference uses) is built;
//  * the table always has one entry per 16 characters; the sampling rate argument is kept and reported by
//    sampling_rate(), it does not change any result;
//  * the base is drawn like the reference's (std::random_device, uniform in [257, 2^31 - 2], :89-96) unless
//    lz77_sss_b200::rks_base() pins it.
#pragma once
#include <cmath>
#include <cstdint>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>

#include <lz77sss_b200.h>

enum direction { LEFT, RIGHT };     // misc/utils.hpp

namespace lz77_sss_b200 {
inline uint64_t& rks_base() {       // 0 = draw from std::random_device
    static thread_local uint64_t b = 0;
    return b;
}
inline int& rks_device() {
    static thread_local int d = 0;
    return d;
}
}  // namespace lz77_sss_b200

template <uint8_t mers_exp = 31, typename pos_t = uint32_t>
class rabin_karp_substring {
    static_assert(mers_exp == 31, "lz77sss_b200 builds rabin_karp_substring<31, pos_t> (the reference's instantiation)");
    using fp_t = uint32_t;
    using fp_conc_t = uint64_t;
    static constexpr fp_t mers_prim = (fp_t(1) << mers_exp) - 1;

    const uint8_t* T = nullptr;
    pos_t n = 0, s = 16, w = 0;
    fp_conc_t b = 0;
    fp_conc_t pop_prec[256] = {};
    std::vector<fp_t> table;        // table[g] = fingerprint of T[0 .. 16 (g + 1)), from the device
    lz77sss_handle* h = nullptr;
    lz77sss_rks* r = nullptr;

    static fp_t mod(fp_conc_t v) {
        fp_conc_t x = (v >> mers_exp) + (v & mers_prim);
        x = (x >> mers_exp) + (x & mers_prim);
        x = (x >> mers_exp) + (x & mers_prim);
        return x >= mers_prim ? (fp_t)(x - mers_prim) : (fp_t)x;
    }
    fp_t b_pow(pos_t e) const {
        fp_conc_t res = 1, x = b;
        for (uint64_t k = e; k; k >>= 1) {
            if (k & 1) res = mod(res * x);
            x = mod(x * x);
        }
        return (fp_t)res;
    }
    static void check(int rc) {
        if (rc != LZ77SSS_OK) throw std::runtime_error(std::string("lz77sss_b200: ") + lz77sss_last_error());
    }
    void release() {
        if (r) lz77sss_rks_destroy(r);
        if (h) lz77sss_destroy(h);
        r = nullptr;
        h = nullptr;
    }

public:
    rabin_karp_substring() = default;
    rabin_karp_substring(const rabin_karp_substring&) = delete;
    rabin_karp_substring& operator=(const rabin_karp_substring&) = delete;
    rabin_karp_substring(rabin_karp_substring&& o) noexcept { *this = std::move(o); }
    rabin_karp_substring& operator=(rabin_karp_substring&& o) noexcept {
        if (this != &o) {
            release();
            T = o.T; n = o.n; s = o.s; w = o.w; b = o.b;
            for (int c = 0; c < 256; c++) pop_prec[c] = o.pop_prec[c];
            table = std::move(o.table);
            h = o.h; r = o.r;
            o.h = nullptr; o.r = nullptr;
        }
        return *this;
    }
    ~rabin_karp_substring() { release(); }

    // rabin_karp_substring.hpp:79-172; p (the reference's thread count) has no meaning on the device
    template <typename char_t>
    rabin_karp_substring(const char_t* text, const pos_t size, const pos_t sampling = 16, const pos_t window = 0,
                         const uint16_t /*p*/ = 1)
        : T(reinterpret_cast<const uint8_t*>(text)), n(size), s(sampling), w(window) {
        static_assert(sizeof(char_t) == 1);
        b = lz77_sss_b200::rks_base();
        if (b == 0) {
            std::random_device rd;
            std::mt19937_64 mt(rd());
            b = std::uniform_int_distribution<fp_t>(257, mers_prim - 1)(mt);
        }
        if (w != 0) {
            const fp_conc_t top = b_pow(w);
            for (fp_conc_t c = 0; c < 256; c++) pop_prec[c] = mod(fp_conc_t(mers_prim) * mers_prim - top * c);
        }
        lz77sss_config cfg{};
        cfg.struct_size = sizeof cfg;
        cfg.tau = 512;
        cfg.num_threads = 1;
        cfg.phr_mode = LZ77SSS_PHR_LPF_OPT;
        cfg.fact_mode = LZ77SSS_FACT_GREEDY;
        cfg.sss_qmax = 0xFFFFFFFFu;
        for (int i = 0; i < 5; i++) cfg.gap_bases[i] = 257 + i;
        cfg.device = lz77_sss_b200::rks_device();
        check(lz77sss_create(&cfg, &h));
        check(lz77sss_rks_create(h, T, n, b, &r));
        uint64_t cnt = 0;
        check(lz77sss_rks_table(r, nullptr, 0, &cnt));
        table.resize(cnt);
        check(lz77sss_rks_table(r, table.data(), cnt, &cnt));
    }

    pos_t sampling_rate() const { return s; }
    pos_t window_size() const { return w; }
    uint64_t size_in_bytes() const { return sizeof(*this) + table.size() * sizeof(fp_t); }

    template <typename char_t>
    fp_t push(const fp_t fp, const char_t chr) const { return mod(b * fp_conc_t(fp) + fp_conc_t((uint8_t)chr)); }      // :187-191
    template <typename char_t>
    fp_t roll(const fp_t fp, const char_t pop, const char_t chr) const {                                                 // :193-197
        return mod(b * fp_conc_t(fp) + pop_prec[(uint8_t)pop] + fp_conc_t((uint8_t)chr));
    }
    fp_t concat(const fp_t fp_lft, const fp_t fp_rght, const pos_t len_rght) const {                                     // :199-202
        return mod(fp_conc_t(b_pow(len_rght)) * fp_conc_t(fp_lft) + fp_conc_t(fp_rght));
    }
    template <direction dir = RIGHT>
    fp_t substring_naive(pos_t pos, const pos_t len) const {                                                             // :204-214
        if constexpr (dir == LEFT) pos -= len - 1;
        fp_t fp = 0;
        for (pos_t i = pos; i < pos + len; i++) fp = push(fp, T[i]);
        return fp;
    }
    fp_t substring_up_to(const pos_t pos) const {                                                                        // :216-223
        const pos_t g = pos / 16;
        fp_t fp = g ? table[g - 1] : 0;
        for (pos_t i = g * 16; i < pos; i++) fp = push(fp, T[i]);
        return fp;
    }
    template <direction dir = RIGHT>
    fp_t substring(pos_t pos, const pos_t len) const {                                                                   // :225-239
        if constexpr (dir == LEFT) pos -= len - 1;
        const fp_t beg = mod(fp_conc_t(substring_up_to(pos)) * fp_conc_t(b_pow(len)));
        const fp_t end = substring_up_to(pos + len);
        return end >= beg ? end - beg : mers_prim - (beg - end);
    }
    // many substring() queries at once, on the device
    std::vector<fp_t> substring_batch(const std::vector<uint64_t>& pos, const std::vector<uint64_t>& len) const {
        std::vector<fp_t> out(pos.size());
        check(lz77sss_rks_substring(r, pos.data(), len.data(), pos.size(), out.data()));
        return out;
    }
    fp_conc_t base() const { return b; }
};
