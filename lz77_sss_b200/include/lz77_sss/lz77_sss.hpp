// lz77_sss.hpp — B200 host-side mirror of the reference's public API.
//
// Same names, template parameters, argument meaning and callback contract as
// /root/reference/include/lz77_sss/lz77_sss.hpp:47-75,129-203 (enums, parameters, lz77_sss<pos_t>::factor,
// factorize_approximate / factorize_exact / decode), so code written against the reference compiles against this
// header unchanged; the work is forwarded to liblz77sss_b200.so through the C ABI of include/lz77sss_b200.h.
//
// Differences a caller can see:
//  * errors: the reference has no error path; here a failed device call throws std::runtime_error.
//  * hash parameters: the reference draws the five gap-index bases from std::random_device
//    (rolling_hash.hpp:30-39,127-130).  lz77_sss_b200::options().gap_bases does the same when left at 0 and can be
//    pinned for reproducible factor streams (the parity tests do).
//  * the input needs no zero padding; it is only read in [0, n).
//  * implemented: factorize_approximate<greedy, lpf_opt, tau> and decode.  Other modes throw (scope: DESIGN.md).
#pragma once
#include <algorithm>
#include <array>
#include <cstdint>
#include <iostream>
#include <iterator>
#include <random>
#include <stdexcept>
#include <string>
#include <thread>
#include <type_traits>
#include <vector>

#include <lz77sss_b200.h>

enum phrase_mode { lpf_naive, lpf_lnf_naive, lpf_opt, lpf_lnf_opt };      // lz77_sss.hpp:47-52
enum factorize_mode { greedy_naive, greedy, skip_phrases };               // lz77_sss.hpp:54-58
enum transform_mode { naive, with_samples, without_samples };             // lz77_sss.hpp:60-64

struct parameters {                                                       // lz77_sss.hpp:67-70
    uint16_t num_threads = 0;
    bool log = false;
};

namespace lz77_sss_b200 {
struct options_t {
    std::array<uint64_t, 5> gap_bases{};   // 0 = draw from std::random_device in [257, 2^20-1], like the reference
    uint32_t sss_base = 0;                 // 0 = library default
    uint32_t sss_qmax = 0xFFFFFFFFu;       // library default
    int device = 0;
};
inline options_t& options() {
    static thread_local options_t o;
    return o;
}
inline void check(int rc) {
    if (rc != LZ77SSS_OK) throw std::runtime_error(std::string("lz77sss_b200: ") + lz77sss_last_error());
}
}  // namespace lz77_sss_b200

template <typename pos_t = uint32_t>
class lz77_sss {
public:
    static_assert(std::is_same_v<pos_t, uint32_t> || std::is_same_v<pos_t, uint64_t>);

    static constexpr phrase_mode default_phr_mode = lpf_opt;
    static constexpr factorize_mode default_fact_mode = greedy;
    static constexpr transform_mode default_transf_mode = without_samples;
    static constexpr uint64_t default_tau = 512;

    struct factor {                                                       // lz77_sss.hpp:129-174
        pos_t src;
        pos_t len;

        pos_t length() const { return std::max<pos_t>(1, len); }

        static constexpr pos_t size_of() { return std::is_same_v<pos_t, uint32_t> ? 8 : 10; }

        friend std::istream& operator>>(std::istream& in, factor& f) {
            if constexpr (std::is_same_v<pos_t, uint32_t>) {
                in.read((char*)&f, 8);
            } else {
                f.src = 0;
                f.len = 0;
                in.read((char*)&f.src, 5);
                in.read((char*)&f.len, 5);
            }
            return in;
        }
        friend std::ostream& operator<<(std::ostream& out, const factor& f) {
            if constexpr (std::is_same_v<pos_t, uint32_t>) {
                out.write((char*)&f, 8);
            } else {
                out.write((char*)&f.src, 5);
                out.write((char*)&f.len, 5);
            }
            return out;
        }
    };

    template <factorize_mode fact_mode = default_fact_mode, phrase_mode phr_mode = default_phr_mode,
              uint64_t tau = default_tau, typename char_t, typename output_fnc_t>
    static void factorize_approximate(char_t* input, pos_t input_size, output_fnc_t output, parameters params = {}) {
        static_assert(sizeof(char_t) == 1);
        handle h(tau, params, (uint32_t)phr_mode, (uint32_t)fact_mode);
        auto tramp = [](void* ctx, uint64_t src, uint64_t len) {
            (*static_cast<output_fnc_t*>(ctx))(factor{(pos_t)src, (pos_t)len});
        };
        lz77_sss_b200::check(lz77sss_factorize_approximate(h.h, reinterpret_cast<const uint8_t*>(input), input_size,
                                                          +tramp, &output));
    }

    template <factorize_mode fact_mode = default_fact_mode, phrase_mode phr_mode = default_phr_mode,
              transform_mode transf_mode = default_transf_mode, uint64_t tau = default_tau, typename char_t,
              typename output_fnc_t>
    static void factorize_exact(char_t*, pos_t, output_fnc_t, parameters = {}) {
        static_assert(fact_mode != skip_phrases);                         // lz77_sss.hpp:319
        throw std::runtime_error("lz77sss_b200: factorize_exact is not built yet (DESIGN.md scope table)");
    }

    template <std::input_iterator fact_it_t, typename out_it_t>
    static void decode(fact_it_t fact_it, out_it_t out_it, pos_t output_size) {     // algorithms/common.cpp:33-54
        using char_t = std::iter_value_t<out_it_t>;
        static_assert(sizeof(char_t) == 1);
        std::vector<uint64_t> src, len;
        for (uint64_t pos = 0; pos < output_size;) {
            factor f = *fact_it++;
            src.push_back(f.src);
            len.push_back(f.len);
            pos += f.length();
        }
        std::vector<uint8_t> out(output_size);
        handle h(default_tau, {}, lpf_opt, greedy);
        lz77_sss_b200::check(lz77sss_decode(h.h, src.data(), len.data(), src.size(), out.data(), output_size));
        for (uint64_t i = 0; i < output_size; i++) out_it[i] = (char_t)out[i];
    }

private:
    struct handle {
        lz77sss_handle* h = nullptr;
        handle(uint64_t tau, parameters params, uint32_t phr_mode, uint32_t fact_mode) {
            auto& o = lz77_sss_b200::options();
            lz77sss_config c{};
            c.struct_size = sizeof c;
            c.tau = (uint32_t)tau;
            // num_threads == 0: the reference takes omp_get_max_threads() (lz77_sss.hpp:289-291)
            c.num_threads = params.num_threads ? params.num_threads : std::max(1u, std::thread::hardware_concurrency());
            c.pos_bits = sizeof(pos_t) * 8;
            c.phr_mode = phr_mode;
            c.fact_mode = fact_mode;
            c.sss_base = o.sss_base;
            c.sss_qmax = o.sss_qmax;
            std::random_device rd;
            std::mt19937_64 gen(rd());
            std::uniform_int_distribution<uint64_t> dist(257, (uint64_t(1) << 20) - 1);   // rolling_hash.hpp:31-36
            for (int i = 0; i < 5; i++) c.gap_bases[i] = o.gap_bases[i] ? o.gap_bases[i] : dist(gen);
            c.device = o.device;
            c.log = params.log;
            lz77_sss_b200::check(lz77sss_create(&c, &h));
        }
        ~handle() { lz77sss_destroy(h); }
        handle(const handle&) = delete;
        handle& operator=(const handle&) = delete;
    };
};
