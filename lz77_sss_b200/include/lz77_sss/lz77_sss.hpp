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
//  * implemented: factorize_approximate for all three factorize modes x four phrase modes, decode (any n), and
//    factorize_exact for texts below 4 GiB (same factor lengths as the reference for every transform_mode / range_ds_t;
//    sources are valid earlier occurrences but not necessarily the reference's, whose choice depends on its range structure).
//  * LZ77_SSS_BENCH: like the reference (lz77_sss.hpp:298-311,357-365), a translation unit that defines LZ77_SSS_BENCH
//    and the globals text_name / result_file_path / result_file before including this header gets one
//    `RESULT key=value ...` line per logged factorization, with the reference's keys; the phase keys carry the
//    device pipeline's times under the reference's names (mapping at write_result below).
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

#ifdef LZ77_SSS_BENCH
#include <fstream>
extern std::string text_name;             // defined by the bench program, like bench/lz77_sss_bench.cpp:29-31
extern std::string result_file_path;
extern std::ofstream result_file;
#endif

// the range structures of the exact path are template template parameters of factorize_exact (lz77_sss.hpp:188-196);
// the names exist so that the reference's call sites compile
template <typename sidx_t> struct decomposed_static_weighted_square_grid {};
template <typename sidx_t> struct decomposed_static_weighted_kd_tree {};
template <typename sidx_t> struct decomposed_semi_dynamic_square_grid {};
template <typename sidx_t> struct static_weighted_square_grid {};
template <typename sidx_t> struct static_weighted_kd_tree {};
template <typename sidx_t> struct semi_dynamic_square_grid {};

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
    template <typename sidx_t> using default_range_ds_t = decomposed_static_weighted_square_grid<sidx_t>;
    static constexpr uint64_t default_tau = 512;                          // lz77_sss.hpp:82-94
    static constexpr uint64_t max_delta = 256;
    static constexpr uint64_t rks_sample_rate = 16;
    static constexpr uint64_t num_patt_lens = 5;
    static constexpr uint64_t min_rh_index_size = 1 << 20;
    static constexpr uint64_t max_rh_index_size = 1 << 30;
    static constexpr double min_rel_rh_index_size = 0.1;
    static constexpr uint64_t num_par_sect_per_thr = 16;

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
#ifdef LZ77_SSS_BENCH
        if (params.log) write_result(h, input_size, tau, (uint32_t)phr_mode, (uint32_t)fact_mode);
#endif
    }

    template <factorize_mode fact_mode = default_fact_mode, phrase_mode phr_mode = default_phr_mode,
              transform_mode transf_mode = default_transf_mode,
              template <typename> typename range_ds_t = default_range_ds_t, uint64_t tau = default_tau, typename char_t,
              typename output_fnc_t>
    static void factorize_exact(char_t* input, pos_t input_size, output_fnc_t output, parameters params = {}) {   // lz77_sss.hpp:188-200
        static_assert(sizeof(char_t) == 1);
        static_assert(fact_mode != skip_phrases);                         // lz77_sss.hpp:319
        // transf_mode and range_ds_t choose HOW the reference searches; the factor lengths do not depend on them, and the
        // device computes them from the suffix array of the text (csrc/exact.cu)
        handle h(tau, params, (uint32_t)phr_mode, (uint32_t)fact_mode);
        auto tramp = [](void* ctx, uint64_t src, uint64_t len) {
            (*static_cast<output_fnc_t*>(ctx))(factor{(pos_t)src, (pos_t)len});
        };
        lz77_sss_b200::check(lz77sss_factorize_exact(h.h, reinterpret_cast<const uint8_t*>(input), input_size, +tramp, &output));
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
    struct handle;
#ifdef LZ77_SSS_BENCH
    // the reference's RESULT line (lz77_sss.hpp:298-311,357-365; phase keys of SURVEY.md §5).  Phase mapping of the device
    // pipeline: lce_t = synchronizing set + suffix order + LCP/range-minimum structures (build_lce), nsv_psv = 0 (the
    // nearest-smaller searches run inside the candidate kernel), lpf = candidates + their copy + the coverage chain,
    // phrase_info = 0 (inside lpf), init_gap_idx = 0 (the index is built slice by slice inside factorize), factorize =
    // gap index + parse + callbacks; h2d = upload of the text.  Times in nanoseconds, throughput in the reference's unit.
    static void write_result(handle& h, uint64_t n, uint64_t tau, uint32_t phr_mode, uint32_t fact_mode) {
        if (result_file_path == "") return;
        lz77sss_stats st;
        lz77_sss_b200::check(lz77sss_get_stats(h.h, &st));
        const double ratio = n / (double)st.num_factors;
        result_file << "RESULT text_name=" << text_name << " n=" << n << " alg=lz77_sss_b200 num_threads=" << h.num_threads
                    << " tau=" << tau << " phr_mode=" << phr_mode << " fact_mode=" << fact_mode << " transf_mode=0"
                    << " h2d=" << st.t_h2d_ns << " lce_t=" << st.t_sss_ns + st.t_sort_ns + st.t_lce_ns << " nsv_psv=0"
                    << " lpf=" << st.t_lpf_ns + st.t_d2h_ns + st.t_chain_ns << " phrase_info=0 init_gap_idx=0"
                    << " factorize=" << st.t_gap_hash_ns + st.t_emit_ns << " num_factors=" << st.num_factors
                    << " comp_ratio=" << ratio << " time=" << st.t_total_ns
                    << " throughput=" << 1000.0 * n / (double)st.t_total_ns << " mem_peak=0" << std::endl;
    }
#endif
    struct handle {
        uint32_t num_threads = 0;
        lz77sss_handle* h = nullptr;
        handle(uint64_t tau, parameters params, uint32_t phr_mode, uint32_t fact_mode) {
            auto& o = lz77_sss_b200::options();
            lz77sss_config c{};
            c.struct_size = sizeof c;
            c.tau = (uint32_t)tau;
            // num_threads == 0: the reference takes omp_get_max_threads() (lz77_sss.hpp:289-291)
            c.num_threads = params.num_threads ? params.num_threads : std::max(1u, std::thread::hardware_concurrency());
            num_threads = c.num_threads;
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
