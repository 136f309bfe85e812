// ORACLE (test infrastructure).  Thin C-ABI driver around the UNMODIFIED reference headers
// (/root/reference/include + patched-files, compiled where they lie) against oracle/shim.
// Built once per tau (-DREF_TAU=...) into oracle/_ref/libref_tau<tau>.so, because the reference makes
// tau a template parameter (lz77_sss.hpp:176-186) and its misc/utils.hpp defines non-inline
// functions (one translation unit per library).
// LZ77_SSS_BENCH makes the reference append its own `RESULT key=value` line (phase times in ns: lce_t, nsv_psv,
// lpf, phrase_info, init_gap_idx, factorize, ...; lz77_sss.hpp:298-311,357-365, misc/utils.hpp:642-648) to
// `result_file` when parameters::log is set.  The three globals are the ones bench/lz77_sss_bench.cpp:29-31 defines.
#define LZ77_SSS_BENCH
#include <fstream>
#include <string>
std::string text_name;
std::string result_file_path;
std::ofstream result_file;
#include <lz77_sss/lz77_sss.hpp>
#include <lz77_sss/misc/huffman.hpp>
#include <lz77_sss/misc/vbyte.hpp>
#include <omp.h>
#include <unistd.h>
#include <iostream>
#include <sstream>
#include <cstdint>
#include <cstring>
#include <random>
#include <vector>

#ifndef REF_TAU
#error "compile with -DREF_TAU=<tau>"
#endif

namespace {

struct sink_t {
    uint64_t* src;
    uint64_t* len;
    uint64_t cap;
    uint64_t z;
};

template <typename pos_t, uint64_t tau>
int64_t run_approximate(const uint8_t* text, uint64_t n, int phr_mode, int fact_mode,
                        int num_threads, sink_t& sink, double* seconds, bool log = false) {
    // every reference driver pads the input with zero bytes (cli/lz77_sss_3_aprx.cpp:66,
    // bench/lz77_sss_bench_tau.cpp:72)
    std::vector<char> buf(n + 4 * 4096, 0);
    std::memcpy(buf.data(), text, n);
    auto out = [&](typename lz77_sss<pos_t>::factor f) {
        if (sink.z < sink.cap) {
            sink.src[sink.z] = f.src;
            sink.len[sink.z] = f.len;
        }
        sink.z++;
    };
    parameters prm { .num_threads = (uint16_t)num_threads, .log = log };
    char* T = buf.data();
    pos_t N = (pos_t)n;
    using L = lz77_sss<pos_t>;
    auto t0 = std::chrono::steady_clock::now();   // timed like lz77_sss.hpp:314-353: around the call only
#define RUN(FM, PM) L::template factorize_approximate<FM, PM, tau>(T, N, out, prm)
#define RUN_PM(FM)                                                   \
    switch (phr_mode) {                                              \
        case lpf_naive: RUN(FM, lpf_naive); break;                   \
        case lpf_lnf_naive: RUN(FM, lpf_lnf_naive); break;           \
        case lpf_opt: RUN(FM, lpf_opt); break;                       \
        case lpf_lnf_opt: RUN(FM, lpf_lnf_opt); break;               \
        default: return -1;                                          \
    }
    if (fact_mode == greedy) { RUN_PM(greedy) }
    else if (fact_mode == greedy_naive) { RUN_PM(greedy_naive) }
    else if (fact_mode == skip_phrases) { RUN_PM(skip_phrases) }
    else return -1;
#undef RUN_PM
#undef RUN
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return (int64_t)sink.z;
}

// Mirror of the generator inside rk_prime<107>::random64 (rolling_hash.hpp:127-130): a
// function-local static mt19937_64 seeded once from (the pinned) std::random_device, one draw
// from uniform_int_distribution<uint64_t>(257, 2^20-1) per rk_prime constructed with base 0.
// rolling_hash_index_107 constructs num_patt_lens = 5 of them per factorization
// (rolling_hash_index_107.hpp:72-75), in order of pattern-length index.
std::mt19937_64& mirror() {
    static std::mt19937_64 g(ORACLE_PINNED_SEED);
    return g;
}
uint64_t draw(std::mt19937_64& g) {
    return std::uniform_int_distribution<uint64_t>(257, (uint64_t(1) << 20) - 1)(g);
}
}  // namespace

extern "C" {

uint64_t refdrv_tau() { return REF_TAU; }
int refdrv_max_threads() { return omp_get_max_threads(); }
uint64_t refdrv_sizeof_rk_prime() { return sizeof(lce::rolling_hash::rk_prime<107>); }

void refdrv_set_sss_params(uint32_t base, uint32_t qmax) {
    oracle::g_sss_base = base;
    oracle::g_sss_qmax = qmax;
}

// The 5 gap-index bases the NEXT refdrv_factorize_approximate call of this library will use.
void refdrv_peek_next_bases(uint64_t out[5]) {
    std::mt19937_64 g = mirror();
    for (int i = 0; i < 5; i++) out[i] = draw(g);
}

// phr_mode / fact_mode: the reference's enum values (lz77_sss.hpp:47-58).  pos_t = uint32_t iff
// n <= UINT32_MAX, as the reference CLIs choose (cli/lz77_sss_3_aprx.cpp:73).
// Returns z (number of factors; may exceed cap, then only the first cap are stored) or <0.
int64_t refdrv_factorize_approximate(const uint8_t* text, uint64_t n, int phr_mode, int fact_mode,
                                     int num_threads, uint64_t* src, uint64_t* len, uint64_t cap,
                                     double* seconds) {
    if (n == 0) return -2;
    sink_t s { src, len, cap, 0 };
    int64_t z;
    if (n <= 0xFFFFFFFFull) z = run_approximate<uint32_t, REF_TAU>(text, n, phr_mode, fact_mode, num_threads, s, seconds);
    else z = run_approximate<uint64_t, REF_TAU>(text, n, phr_mode, fact_mode, num_threads, s, seconds);
    if (z >= 0 && fact_mode != skip_phrases) {   // skip_phrases builds no gap index (lz77_sss.hpp:417-418)
        int builds = (phr_mode == lpf_lnf_naive || phr_mode == lpf_lnf_opt) ? 1 : 1;
        for (int i = 0; i < 5 * builds; i++) draw(mirror());   // advance past the bases that run consumed
    }
    return z;
}

// Same run with parameters::log = true: the reference's console log is discarded, its RESULT line (phase times in
// nanoseconds) is returned in result_out.  Timed region as above.
int64_t refdrv_factorize_approximate_logged(const uint8_t* text, uint64_t n, int phr_mode, int fact_mode, int num_threads,
                                            uint64_t* src, uint64_t* len, uint64_t cap, double* seconds,
                                            char* result_out, uint64_t result_cap) {
    if (n == 0) return -2;
    char path[] = "/tmp/refdrv_result_XXXXXX";
    int fd = mkstemp(path);
    if (fd < 0) return -3;
    close(fd);
    text_name = "synthetic";
    result_file_path = path;
    result_file.open(path, std::ios_base::trunc);
    std::ostringstream devnull;
    std::streambuf* old = std::cout.rdbuf(devnull.rdbuf());
    sink_t s { src, len, cap, 0 };
    int64_t z;
    if (n <= 0xFFFFFFFFull) z = run_approximate<uint32_t, REF_TAU>(text, n, phr_mode, fact_mode, num_threads, s, seconds, true);
    else z = run_approximate<uint64_t, REF_TAU>(text, n, phr_mode, fact_mode, num_threads, s, seconds, true);
    std::cout.rdbuf(old);
    result_file.close();
    result_file_path = "";
    if (z >= 0 && fact_mode != skip_phrases)
        for (int i = 0; i < 5; i++) draw(mirror());
    if (result_out && result_cap) {
        std::ifstream in(path);
        std::string line((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
        std::strncpy(result_out, line.c_str(), result_cap - 1);
        result_out[result_cap - 1] = 0;
    }
    unlink(path);
    return z;
}

// rabin_karp_substring<31, uint32_t> (data_structures/rabin_karp_substring.hpp:35-240), built by the reference's own
// constructor with p threads; the base it drew (from the pinned std::random_device) is returned so that the device
// build can use the same one.  fps_out receives the reference's table for sampling rate s, qout[k] = substring(qpos[k], qlen[k]).
int64_t refdrv_rks(const uint8_t* text, uint64_t n, uint64_t s, int p, uint64_t* base_out, uint32_t* fps_out,
                   uint64_t fps_cap, const uint64_t* qpos, const uint64_t* qlen, uint64_t nq, uint32_t* qout) {
    struct probe : rabin_karp_substring<31, uint32_t> {
        using rabin_karp_substring<31, uint32_t>::rabin_karp_substring;
        uint64_t base() const { return (uint64_t)this->b; }
        const std::vector<uint32_t>& table() const { return this->fps; }
    };
    std::vector<char> buf(n + 64, 0);
    std::memcpy(buf.data(), text, n);
    probe R(buf.data(), (uint32_t)n, (uint32_t)s, (uint32_t)0, (uint16_t)p);
    if (base_out) *base_out = R.base();
    const auto& f = R.table();
    for (uint64_t k = 0; k < f.size() && k < fps_cap; k++) fps_out[k] = f[k];
    for (uint64_t k = 0; k < nq; k++) qout[k] = R.substring((uint32_t)qpos[k], (uint32_t)qlen[k]);
    return (int64_t)f.size();
}

// lz77_sss<pos_t>::factorize_exact<greedy, lpf_opt, transf_mode> (lz77_sss.hpp:188-200) with the default range structure, as
// cli/lz77_sss_exact_smpl.cpp:46-56 and cli/lz77_sss_exact.cpp call it.  transf_mode: 0 naive, 1 with_samples,
// 2 without_samples.  Only built for tau = 64 and 512 (compile time).
int64_t refdrv_factorize_exact(const uint8_t* text, uint64_t n, int transf_mode, int num_threads, uint64_t* src,
                               uint64_t* len, uint64_t cap, double* seconds) {
#if REF_TAU == 64 || REF_TAU == 512
    if (n == 0 || n > 0xFFFFFFFFull) return -2;
    std::vector<char> buf(n + 4 * 4096, 0);
    std::memcpy(buf.data(), text, n);
    sink_t sink { src, len, cap, 0 };
    auto out = [&](lz77_sss<uint32_t>::factor f) {
        if (sink.z < sink.cap) { sink.src[sink.z] = f.src; sink.len[sink.z] = f.len; }
        sink.z++;
    };
    parameters prm { .num_threads = (uint16_t)num_threads, .log = false };
    using L = lz77_sss<uint32_t>;
    auto t0 = std::chrono::steady_clock::now();
    if (transf_mode == 0) L::template factorize_exact<greedy, lpf_opt, naive, L::template default_range_ds_t, REF_TAU>(buf.data(), (uint32_t)n, out, prm);
    else if (transf_mode == 1) L::template factorize_exact<greedy, lpf_opt, with_samples, L::template default_range_ds_t, REF_TAU>(buf.data(), (uint32_t)n, out, prm);
    else if (transf_mode == 2) L::template factorize_exact<greedy, lpf_opt, without_samples, L::template default_range_ds_t, REF_TAU>(buf.data(), (uint32_t)n, out, prm);
    else return -1;
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    for (int i = 0; i < 5; i++) draw(mirror());        // the approximation inside drew the five gap-index bases
    return (int64_t)sink.z;
#else
    (void)text; (void)n; (void)transf_mode; (void)num_threads; (void)src; (void)len; (void)cap; (void)seconds;
    return -3;
#endif
}

// The reference's own file formats (misc/huffman.hpp:321-436, misc/vbyte.hpp:64-101), used as cli/lz77_sss_3_aprx.cpp:73-84
// and cli/lz77_sss_decode.cpp:61-66 use them.
int refdrv_huff_write(const uint64_t* src, const uint64_t* len, uint64_t z, uint64_t n, const char* path) {
    std::ofstream out(path, std::ios::binary);
    if (!out.good()) return -1;
    huff_writer writer(out, n);
    for (uint64_t i = 0; i < z; i++) writer.add(lz77_sss<uint64_t>::factor{src[i], len[i]});
    writer.finish();
    return 0;
}
int64_t refdrv_huff_read(const char* path, uint64_t* src, uint64_t* len, uint64_t cap, uint64_t* n_out) {
    std::ifstream in(path, std::ios::binary);
    if (!in.good()) return -1;
    uint64_t n = 0;
    in.read((char*)&n, 5);
    bit_reader reader(in);
    huffman len_huff, dist_huff;
    using factor = lz77_sss<uint64_t>::factor;
    huff_factor_iterator<factor> it(reader, len_huff, dist_huff, n);
    uint64_t z = 0;
    for (uint64_t pos = 0; pos < n;) {
        const factor f = *it++;
        if (z < cap) { src[z] = f.src; len[z] = f.len; }
        z++;
        pos += f.length();
    }
    if (n_out) *n_out = n;
    return (int64_t)z;
}
int refdrv_vbyte_write(const uint64_t* vals, uint64_t cnt, int bits, const char* path) {
    std::ofstream out(path, std::ios::binary);
    if (!out.good()) return -1;
    for (uint64_t i = 0; i < cnt; i++) {
        if (bits == 32) encode_vbyte<uint32_t>(out, (uint32_t)vals[i]);
        else encode_vbyte<uint64_t>(out, vals[i]);
    }
    return 0;
}

// lz77_sss<>::decode (algorithms/common.cpp:33-54)
int refdrv_decode(const uint64_t* src, const uint64_t* len, uint64_t z, uint8_t* out, uint64_t n) {
    std::vector<lz77_sss<uint64_t>::factor> f(z);
    for (uint64_t i = 0; i < z; i++) f[i] = { src[i], len[i] };
    lz77_sss<uint64_t>::decode(f.begin(), reinterpret_cast<char*>(out), n);
    return 0;
}
}
