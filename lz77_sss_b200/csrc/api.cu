// C-ABI (include/lz77sss_b200.h) and host orchestration of the device pipeline.
//
// Mirrors lz77_sss<pos_t>::factorizer::compute_approximation<greedy, lpf_opt>
// (/root/reference/include/lz77_sss/lz77_sss.hpp:374-491):
//   build_lce           -> K1+K2 (sss.cu), K3 (suffix.cu), range-min/successor structures (lpf.cu)
//   build_LPF_opt       -> lpf_candidates kernel (all LCE probes, data-parallel) + the coverage chain on the
//                          host (lpf_opt.cpp:58-63,87-89: each step depends on the previous phrase end)
//   get_phrase_info / pattern lengths / table size -> host scalars, same double arithmetic (lz77_sss.hpp:112-122,425-461)
//   gap index           -> K5 (gap_hash.cu) computes every bucket id; the host keeps only the table of last writers
//   factorize_greedy    -> thin sequential pass on the host (greedy.cpp:34-140), calls the factor callback
// The host never computes a fingerprint, never sorts and never builds an index: it consumes device results.
#include <algorithm>
#include <array>
#include <atomic>
#include <exception>
#include <mutex>
#include <thread>
#include <chrono>
#include <cmath>
#include <cstring>
#include <limits>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/lz77sss_b200.h"
#include "common.cuh"
#include "stages.cuh"

namespace lz77sss {
thread_local unsigned long long g_kernel_launches = 0;
bool g_sync_every_kernel = getenv("LZ77SSS_SYNC_KERNELS") != nullptr;
bool g_prof_on = false;
namespace {
struct prof_row { std::string name; double ms = 0; uint64_t launches = 0, bytes = 0; };
std::vector<prof_row> g_prof;
cudaEvent_t g_prof_e0 = nullptr, g_prof_e1 = nullptr;
}
void prof_begin(cudaStream_t st) {
    if (!g_prof_e0) {
        CUDA_CHECK(cudaEventCreate(&g_prof_e0));
        CUDA_CHECK(cudaEventCreate(&g_prof_e1));
    }
    CUDA_CHECK(cudaEventRecord(g_prof_e0, st));
}
void prof_end(cudaStream_t st, const char* name, uint64_t bytes) {
    CUDA_CHECK(cudaEventRecord(g_prof_e1, st));
    CUDA_CHECK(cudaEventSynchronize(g_prof_e1));
    float ms = 0;
    CUDA_CHECK(cudaEventElapsedTime(&ms, g_prof_e0, g_prof_e1));
    for (auto& r : g_prof)
        if (r.name == name) { r.ms += ms; r.launches++; r.bytes += bytes; return; }
    prof_row r;
    r.name = name; r.ms = ms; r.launches = 1; r.bytes = bytes;
    g_prof.push_back(r);
}
}

using namespace lz77sss;

namespace {

thread_local std::string g_last_error;

// cfg.pos_bits == 0 means "choose per input" (cli/lz77_sss_3_aprx.cpp:73); the choice must not outlive the call, also
// when the call throws
struct pos_bits_guard {
    lz77sss_config& cfg;
    const uint32_t saved;
    pos_bits_guard(lz77sss_config& c, uint64_t n) : cfg(c), saved(c.pos_bits) {
        if (cfg.pos_bits == 0) cfg.pos_bits = n <= 0xFFFFFFFFull ? 32 : 64;
    }
    ~pos_bits_guard() { cfg.pos_bits = saved; }
};

struct lpf_t { uint64_t beg, end, src; };

using clk = std::chrono::steady_clock;
inline uint64_t ns_since(clk::time_point t0) {
    return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(clk::now() - t0).count();
}

// pinned host buffer, grow-only
template <typename T>
struct hbuf {
    T* p = nullptr;
    size_t cap = 0;
    ~hbuf() { if (p) cudaFreeHost(p); }
    void reserve(size_t n) {
        if (n <= cap) return;
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;                    // (a failed allocation below must not leave a capacity without a buffer)
        size_t want = n + n / 8 + 64;
        CUDA_CHECK(cudaMallocHost(&p, want * sizeof(T)));
        cap = want;
    }
    T& operator[](size_t i) { return p[i]; }
    const T& operator[](size_t i) const { return p[i]; }
};

constexpr uint64_t FRONT_PAD = 256;   // readable zero bytes before T[0] (leftward LCE loads)

template <typename pos_t>
struct pipeline {
    dbuf<pos_t> S;
    sss_scratch sss;
    suffix_state<pos_t> suf;
    lce_state<pos_t> lce;
    hbuf<pos_t> hS, h_plce, h_nlce, h_plcel, h_nlcel;
    hbuf<pos_t> h_psrc, h_nsrc;        // text position of the candidate's source, ~0 = none
    uint64_t s = 0;
};

}  // namespace

struct lz77sss_handle {
    lz77sss_config cfg;
    cudaStream_t st = nullptr;
    dbuf<uint8_t> text_buf;           // FRONT_PAD + n + TEXT_PAD
    dbuf<uint8_t> rev_buf;            // the reversed text of the lpf_lnf phrase modes, same padding
    dbuf<uint64_t> d_lq;              // leftward-LCE queries (qi, qj, out)
    const uint8_t* d_text = nullptr;  // text of the last factorization (device)
    uint64_t n = 0;
    int pos_bytes = 0;                // internal position width of the last factorization
    pipeline<uint32_t> p32;
    pipeline<uint64_t> p64;
    lz77sss_stats stats;
    bool have_result = false;
    std::vector<std::vector<lpf_t>> LPF;   // per logical block, without the sentinel
    // gap bucket ids
    dbuf<uint64_t> d_rng;
    dbuf<uint32_t> d_ids;
    hbuf<uint32_t> h_ids;
    hbuf<uint64_t> h_rng;
    std::vector<uint8_t*> user_texts;
    // device emission
    gap_parse_scratch gp;
    decode_scratch dec;
    exact_scratch ex;
    dbuf<uint64_t> dec_src, dec_len;
    dbuf<uint8_t> dec_out;
    pinned_vec<uint8_t> d2h_stage[2];      // decode output on its way to pageable caller memory
    pinned_vec<uint64_t> f_src, f_len;  // factors of the device emission (pinned: they arrive by D2H copies)
    uint64_t slice_len = 1ull << 28;
    uint32_t block_len = 1024;        // positions per parse block (measured at 1 GiB: 2048 / 1024 / 512 -> 65.9 / 50.5 / 50.6 ms of parse)
    bool host_emit = false;
    // sharded run (lz77sss_shard_*): what lz77sss_shard_prepare leaves for the range parses
    bool prepare_only = false, prepared = false;
    const uint8_t* host_text = nullptr;
    hbuf<uint64_t> ml_buf;            // merged phrase list incl. the sentinel: three planes (beg, end, src) of ml_count entries
    size_t ml_count = 0;
    uint64_t table_entries = 0;
};

struct lz77sss_rks {
    lz77sss_handle* h = nullptr;
    dbuf<uint8_t> text;
    uint8_t* d_text = nullptr;
    rks_state st;
    unsigned long long launches = 0;
};

namespace {

int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

template <typename F>
int guarded(F f) {
    try {
        return f();
    } catch (const cuda_error& e) {
        return fail(LZ77SSS_E_CUDA, e.what());
    } catch (const std::invalid_argument& e) {
        return fail(LZ77SSS_E_ARG, e.what());
    } catch (const std::exception& e) {
        return fail(LZ77SSS_E_INTERNAL, e.what());
    }
}

bool valid_tau(uint32_t tau) { return tau >= 4 && tau <= 4096 && (tau & (tau - 1)) == 0; }

uint8_t* alloc_text(dbuf<uint8_t>& buf, uint64_t n, cudaStream_t st) {
    uint64_t total = FRONT_PAD + n + TEXT_PAD;
    buf.reserve(total);
    CUDA_CHECK(cudaMemsetAsync(buf.get(), 0, FRONT_PAD, st));
    CUDA_CHECK(cudaMemsetAsync(buf.get() + FRONT_PAD + n, 0, TEXT_PAD, st));
    return buf.get() + FRONT_PAD;
}

template <typename pos_t>
pipeline<pos_t>& pipe(lz77sss_handle* h);
template <> pipeline<uint32_t>& pipe<uint32_t>(lz77sss_handle* h) { return h->p32; }
template <> pipeline<uint64_t>& pipe<uint64_t>(lz77sss_handle* h) { return h->p64; }

// runs f(0) .. f(p-1) on up to hardware_concurrency host threads; the first exception is rethrown
template <typename F>
void run_blocks(uint32_t p, F f) {
    unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    if (const char* e = getenv("LZ77SSS_HOST_THREADS"))      // several ranks on one box share its cores (bench.py sets cores / ranks)
        hw = std::max(1u, (unsigned)atoi(e));
    const unsigned nt = std::min<unsigned>(p, hw);
    if (nt <= 1) {
        for (uint32_t i = 0; i < p; i++) f(i);
        return;
    }
    std::atomic<uint32_t> next{0};
    std::exception_ptr err;
    std::mutex mu;
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; t++)
        th.emplace_back([&]() {
            try {
                for (uint32_t i = next++; i < p; i = next++) f(i);
            } catch (...) {
                std::lock_guard<std::mutex> g(mu);
                if (!err) err = std::current_exception();
            }
        });
    for (auto& t : th) t.join();
    if (err) std::rethrow_exception(err);
}

// ---- host side of build_LPF_opt: the coverage chain (lpf_opt.cpp:46-145) over the device-computed candidates ----
template <typename pos_t>
void lpf_chain(lz77sss_handle* h, pipeline<pos_t>& P) {
    constexpr pos_t NONEP = ~(pos_t)0;
    const uint64_t n = h->n, s = P.s;
    const uint32_t p = h->cfg.num_threads;
    h->LPF.resize(p);
    for (auto& v : h->LPF) v.clear();          // keeps the capacity of the last run
    const pos_t* S = P.hS.p;
    // the logical blocks are independent (that is how the reference's threads run them, lpf_opt.cpp:46-56): one host
    // thread per block, at most the hardware's
    auto run_block = [&](uint32_t i_p) {
        const uint64_t b = (uint64_t)i_p * (n / p);
        const uint64_t e = i_p == p - 1 ? n : (uint64_t)(i_p + 1) * (n / p);
        const uint64_t i_min = std::lower_bound(S, S + s, (pos_t)b) - S;
        const uint64_t i_max = std::lower_bound(S, S + s, (pos_t)e) - S;
        uint64_t max_end = b;
        auto& out = h->LPF[i_p];
        for (uint64_t i = i_min; i < i_max; i++) {
            while (i + 1 < i_max && (uint64_t)S[i + 1] <= max_end) i++;
            const uint64_t lst_end = max_end, pos = S[i];
            lpf_t phr{0, 0, 0};
            auto candidate = [&](uint64_t src, uint64_t lce, uint64_t lcel_max, bool is_psv) {
                const uint64_t end = pos + lce;
                if (end <= lst_end) return;
                uint64_t beg = pos;
                if (pos > lst_end && src != 0 && pos != 0) {
                    const uint64_t cap = pos - lst_end;
                    // the kernel extended left up to pos - max(block start, S[i-1]); a visit can never ask for more
                    uint64_t kcap;
                    {
                        uint64_t prevpos = b;
                        if (i >= 2 && (uint64_t)S[i - 1] >= b) prevpos = S[i - 1];
                        kcap = std::min<uint64_t>(pos - prevpos, src);
                    }
                    if (std::min<uint64_t>(cap, src) > kcap) throw std::logic_error("lpf_chain: left-extension cap exceeded");
                    const uint64_t l = std::min<uint64_t>(lcel_max, cap);
                    beg -= l;
                    src -= l;
                }
                if (beg < lst_end) {
                    const uint64_t exc = lst_end - beg;
                    beg += exc;
                    src += exc;
                }
                if (end > max_end) max_end = end;
                if (is_psv) {
                    if (end - beg > 1) phr = {beg, end, src};
                } else if (end - beg > phr.end - phr.beg) {
                    phr = {beg, end, src};
                }
            };
            if (P.h_psrc[i] != NONEP) candidate(P.h_psrc[i], P.h_plce[i], P.h_plcel[i], true);
            if (P.h_nsrc[i] != NONEP) {
                candidate(P.h_nsrc[i], P.h_nlce[i], P.h_nlcel[i], false);
                if (phr.end - phr.beg > 1) out.push_back(phr);     // pushed only inside the NSV branch (lpf_opt.cpp:141-143)
            }
        }
    };
    run_blocks(p, run_block);
}

// host side of build_LPF_naive (lpf_naive.cpp:33-122): the longer of the two candidates without left extension,
// then skip the synchronizing positions the phrase covers
template <typename pos_t>
void lpf_chain_naive(lz77sss_handle* h, pipeline<pos_t>& P) {
    constexpr pos_t NONEP = ~(pos_t)0;
    const uint64_t n = h->n, s = P.s;
    const uint32_t p = h->cfg.num_threads;
    h->LPF.resize(p);
    for (auto& v : h->LPF) v.clear();          // keeps the capacity of the last run
    const pos_t* S = P.hS.p;
    for (uint32_t i_p = 0; i_p < p; i_p++) {
        const uint64_t b = (uint64_t)i_p * (n / p);
        const uint64_t e = i_p == p - 1 ? n : (uint64_t)(i_p + 1) * (n / p);
        const uint64_t i_min = std::lower_bound(S, S + s, (pos_t)b) - S;
        const uint64_t i_max = std::lower_bound(S, S + s, (pos_t)e) - S;
        auto& out = h->LPF[i_p];
        // the reference's loop header increments i once more after the skip loop below (lpf_naive.cpp:58,106-110):
        // the synchronizing position right behind every visited one is never visited
        for (uint64_t i = i_min; i < i_max; i++) {
            uint64_t src = 0, len = 0;
            if (P.h_psrc[i] != NONEP && (uint64_t)P.h_plce[i] > len) { src = P.h_psrc[i]; len = P.h_plce[i]; }
            if (P.h_nsrc[i] != NONEP && (uint64_t)P.h_nlce[i] > len) { src = P.h_nsrc[i]; len = P.h_nlce[i]; }
            const uint64_t pos = S[i];
            if (len > 0) out.push_back(lpf_t{pos, pos + len, src});
            do { i++; } while (i < i_max && (uint64_t)S[i] < pos + len);
        }
    }
}

// host side of build_LPF_all / build_LNF_all (lpf_lnf.cpp:34-134,139-249) over the device candidates of one text
// direction.  A candidate is dropped when it starts inside the last kept phrase of its kind and has the same
// offset; that rule needs only LCE_R and the offset, so it runs first, and the uncapped left extensions
// (lpf_lnf_opt) are then fetched for the survivors in one batch.
constexpr uint64_t LNF_LCAP = 64;     // left extension computed with the candidates; longer ones go to the batch
template <typename pos_t>
void all_chain(lz77sss_handle* h, pipeline<pos_t>& P, const uint8_t* d_text, bool lnf, bool opt) {
    constexpr pos_t NONEP = ~(pos_t)0;
    const uint64_t n = h->n, s = P.s;
    const uint32_t p = h->cfg.num_threads;
    const pos_t* S = P.hS.p;
    struct pending { uint32_t i_p; size_t slot; };
    std::vector<pending> pend;
    std::vector<uint64_t> q;          // qi plane, then qj plane
    std::vector<size_t> first_new(p);
    // the logical blocks are independent: one host thread each; their pending left-extension queries are collected per block
    std::vector<std::vector<pending>> pend_b(p);
    std::vector<std::vector<uint64_t>> q_b(p);
    run_blocks(p, [&](uint32_t i_p) {
        uint64_t b, e;
        if (lnf) {
            b = i_p == p - 1 ? 0 : n - (uint64_t)(i_p + 1) * (n / p);
            e = n - (uint64_t)i_p * (n / p);
        } else {
            b = (uint64_t)i_p * (n / p);
            e = i_p == p - 1 ? n : (uint64_t)(i_p + 1) * (n / p);
        }
        const uint64_t i_min = std::lower_bound(S, S + s, (pos_t)b) - S;
        const uint64_t i_max = e > (uint64_t)std::numeric_limits<pos_t>::max() ? s : std::lower_bound(S, S + s, (pos_t)e) - S;
        auto& out = h->LPF[i_p];
        auto& pend = pend_b[i_p];
        auto& q = q_b[i_p];
        first_new[i_p] = out.size();
        uint64_t lst_end[2] = {n, n}, lst_off[2] = {0, 0};
        for (uint64_t i = i_min; i < i_max; i++) {
            for (int side = 0; side < 2; side++) {
                const pos_t srcp = side == 0 ? P.h_psrc[i] : P.h_nsrc[i];
                if (srcp == NONEP) continue;
                const uint64_t pos = S[i], src = srcp;
                const uint64_t off = lnf ? src - pos : pos - src;
                if (pos < lst_end[side] && off == lst_off[side]) continue;
                const uint64_t lce = side == 0 ? P.h_plce[i] : P.h_nlce[i];
                const uint64_t end = pos + lce;
                uint64_t l = 0;
                bool full = false;
                if (opt && src != 0 && pos != 0) {
                    l = side == 0 ? P.h_plcel[i] : P.h_nlcel[i];
                    full = l == LNF_LCAP && std::min(src, pos) > LNF_LCAP;
                }
                if (end - (pos - l) > 1) {
                    lst_end[side] = end;
                    lst_off[side] = off;
                    if (full) {
                        pend.push_back({i_p, out.size()});
                        q.push_back(std::min(src, pos) - 1);
                        q.push_back(std::max(src, pos) - 1);
                    }
                    out.push_back(lpf_t{pos - l, end, src - l});
                }
            }
        }
    });
    for (uint32_t i_p = 0; i_p < p; i_p++) {
        pend.insert(pend.end(), pend_b[i_p].begin(), pend_b[i_p].end());
        q.insert(q.end(), q_b[i_p].begin(), q_b[i_p].end());
    }
    if (!pend.empty()) {
        const uint64_t nq = pend.size();
        std::vector<uint64_t> planes(2 * nq), res(nq);
        for (uint64_t k = 0; k < nq; k++) { planes[k] = q[2 * k]; planes[nq + k] = q[2 * k + 1]; }
        h->d_lq.reserve(3 * nq);
        CUDA_CHECK(cudaMemcpyAsync(h->d_lq.get(), planes.data(), 2 * nq * 8, cudaMemcpyHostToDevice, h->st));
        left_lce_batch(d_text, h->d_lq.get(), h->d_lq.get() + nq, nq, h->d_lq.get() + 2 * nq, h->st);
        CUDA_CHECK(cudaMemcpyAsync(res.data(), h->d_lq.get() + 2 * nq, nq * 8, cudaMemcpyDeviceToHost, h->st));
        CUDA_CHECK(cudaStreamSynchronize(h->st));
        h->stats.h2d_bytes += 16 * nq;
        h->stats.d2h_bytes += 8 * nq;
        for (uint64_t k = 0; k < nq; k++) {
            lpf_t& ph = h->LPF[pend[k].i_p][pend[k].slot];
            const uint64_t more = res[k] - LNF_LCAP;       // the first LNF_LCAP characters are already in
            ph.beg -= more;
            ph.src -= more;
        }
    }
    if (lnf)                                               // back to positions of the forward text
        for (uint32_t i_p = 0; i_p < p; i_p++) {
            auto& out = h->LPF[i_p];
            for (size_t k = first_new[i_p]; k < out.size(); k++) {
                const lpf_t r = out[k];
                out[k] = lpf_t{n - r.end, n - r.beg, n - (r.src + (r.end - r.beg))};
            }
        }
}

// greedy_phrase_selection (approximate/common.cpp:33-96): phrases ordered by (beg ascending, end descending; ties in
// insertion order, see oracle/shim/ips4o.hpp), then a chain of phrases that each reach furthest among those
// starting inside (or right behind) the current one; an overlapped phrase is cut at the start of its successor.
void select_phrases(std::vector<lpf_t>& P) {
    if (P.empty()) return;
    std::stable_sort(P.begin(), P.end(), [](const lpf_t& a, const lpf_t& b) {
        return a.beg < b.beg || (a.beg == b.beg && a.end > b.end);
    });
    const size_t cnt = P.size();
    size_t k = 0, i = 1;
    while (i < cnt && P[i].end < P[k].end) i++;
    while (i < cnt) {
        size_t x = cnt;
        if (i + 1 < cnt) {
            for (x = i + 1; x < cnt && P[x].beg <= P[k].end; x++)
                if (P[x].end > P[i].end) i = x;
            if (P[i].end <= P[k].end) i = x;
        }
        if (i == cnt) break;
        if (P[i].beg < P[k].end) P[k].end = P[i].beg;
        if (P[k].end > P[k].beg) k++;
        P[k] = P[i];
        i = x;
    }
    P.resize(k + 1);
}

// get_phrase_info (approximate/common.cpp:100-157)
void phrase_info(lz77sss_handle* h) {
    auto& st = h->stats;
    const uint64_t n = h->n;
    const uint32_t p = h->cfg.num_threads;
    st.num_gaps = st.num_lpf = st.len_lpf_phr = 0;
    for (uint32_t i_p = 0; i_p < p; i_p++) {
        uint64_t gaps = 0, len = 0, cnt = 0;
        uint64_t b = (uint64_t)i_p * (n / p);
        uint64_t e = i_p == p - 1 ? n : (uint64_t)(i_p + 1) * (n / p);
        auto& L = h->LPF[i_p];
        if (!L.empty()) e = std::max(e, L.back().end);
        for (uint32_t j = 0; j < i_p; j++)
            if (!h->LPF[j].empty()) b = std::max(b, h->LPF[j].back().end);
        size_t i = 0;
        while (i < L.size() && L[i].end <= b) i++;
        cnt = L.size() - i;
        if (cnt > 0) {
            len += L[i].end - std::max(L[i].beg, b);
            if (L[i].beg > b) gaps = 1;
            for (i++; i < L.size(); i++) {
                len += L[i].end - L[i].beg;
                if (L[i].beg > L[i - 1].end) gaps++;
            }
            if (L.back().end < e) gaps++;
        } else {
            gaps = 1;
        }
        st.num_gaps += gaps;
        st.num_lpf += cnt;
        st.len_lpf_phr += len;
    }
}

struct gap_params {
    std::array<uint64_t, 5> patt_lens{};
    uint64_t roll_threshold = 0, table_entries = 0;
};

// pattern lengths and table size (lz77_sss.hpp:99-122,425-461; rolling_hash_index_107.hpp:59-67), malloc_count == 0
gap_params choose_gap_params(lz77sss_handle* h) {
    using entry_t = std::pair<double, std::array<uint64_t, 5>>;
    static const double infty = std::numeric_limits<double>::max();
    static const std::array<entry_t, 10> table{{{6, {2, 3, 4, 5, 6}}, {8, {2, 3, 4, 6, 8}}, {12, {2, 3, 4, 8, 12}},
        {16, {2, 4, 6, 9, 16}}, {32, {2, 4, 6, 10, 20}}, {64, {2, 4, 7, 12, 28}}, {128, {2, 4, 8, 16, 36}},
        {256, {2, 5, 10, 20, 42}}, {1024, {2, 6, 12, 24, 48}}, {infty, {2, 8, 16, 32, 64}}}};
    auto& st = h->stats;
    const uint64_t n = h->n;
    const unsigned sizeof_pos = h->cfg.pos_bits / 8;
    gap_params g;
    uint64_t len_gaps = n - st.len_lpf_phr;
    if (sizeof_pos == 4) len_gaps = (uint32_t)len_gaps;
    const double rel_len_gaps = len_gaps / (double)n;
    const double avg_gap_len = len_gaps / (double)st.num_gaps;
    const double avg_lpf_phr_len = st.len_lpf_phr / (double)st.num_lpf;
    const uint64_t target = std::min<uint64_t>(uint64_t(1) << 30,
        std::max<uint64_t>({uint64_t(1) << 20, uint64_t(0), (uint64_t)((n / 3.0) * rel_len_gaps)}));
    const double guess = std::min<double>({avg_gap_len, avg_lpf_phr_len, 8.0 * std::pow(128, 1.0 - rel_len_gaps)});
    for (auto& [threshold, lens] : table)
        if (guess <= threshold) { g.patt_lens = lens; break; }
    for (int j = 0; j < 5; j++) g.roll_threshold += g.patt_lens[j];
    g.roll_threshold /= 5;
    // sizeof(rk_prime<107>) = 80 on x86-64 libstdc++ (checked against the reference build in tests)
    const int64_t rolling_hash_size = (80 + 16 * 256 * 256) * 5;
    uint64_t a = 1 << 20, bb;
    if (sizeof_pos == 4) bb = (uint32_t)(n * 0.1);
    else bb = (uint64_t)(n * 0.1);
    const int64_t min_index = std::max<uint64_t>(a, bb) / sizeof_pos;
    const int64_t max_index = (int64_t)(1 << 30) / sizeof_pos;
    const int64_t tgt_index = std::max<int64_t>(0, (int64_t)target - rolling_hash_size) / sizeof_pos;
    const uint64_t size_h = std::min<int64_t>(max_index, std::max<int64_t>(min_index, tgt_index));
    const uint8_t lg = std::round(std::log2(size_h));
    g.table_entries = uint64_t(1) << lg;
    return g;
}

// Bucket ids of the gap index for the text ranges the emission can touch, fetched from the device.
struct bucket_provider {
    lz77sss_handle* h;
    gap_hash_params gp;
    std::vector<uint64_t> beg, end, off;   // main ranges
    size_t cur = 0;
    // on-demand range
    uint64_t t_beg = 0, t_end = 0;
    std::vector<uint32_t> t_ids;
    uint64_t hashed = 0;

    void fetch_main() {
        const uint64_t nr = beg.size();
        uint64_t total = 0;
        off.resize(nr);
        for (uint64_t k = 0; k < nr; k++) { off[k] = total; total += end[k] - beg[k]; }
        hashed += total;
        if (!total) return;
        h->h_rng.reserve(3 * nr);
        std::memcpy(h->h_rng.p, beg.data(), nr * 8);
        std::memcpy(h->h_rng.p + nr, end.data(), nr * 8);
        std::memcpy(h->h_rng.p + 2 * nr, off.data(), nr * 8);
        h->d_rng.reserve(3 * nr);
        h->d_ids.reserve(5 * total);
        h->h_ids.reserve(5 * total);
        CUDA_CHECK(cudaMemcpyAsync(h->d_rng.get(), h->h_rng.p, 3 * nr * 8, cudaMemcpyHostToDevice, h->st));
        gap_bucket_ids(h->d_text, h->n, gp, h->d_rng.get(), h->d_rng.get() + nr, h->d_rng.get() + 2 * nr, nr, total,
                       h->d_ids.get(), h->st);
        CUDA_CHECK(cudaMemcpyAsync(h->h_ids.p, h->d_ids.get(), 5 * total * 4, cudaMemcpyDeviceToHost, h->st));
        CUDA_CHECK(cudaStreamSynchronize(h->st));
    }
    void fetch_temp(uint64_t pos) {
        // position outside the precomputed ranges (a gap factor ran over whole LPF phrases, greedy.cpp:68-77,
        // or the stale-fingerprint window at n - patt_len): hash a fresh stretch on the device
        t_beg = pos;
        t_end = std::min<uint64_t>(h->n, pos + 8192);
        const uint64_t total = t_end - t_beg;
        hashed += total;
        dbuf<uint64_t> rng;
        dbuf<uint32_t> ids;
        rng.reserve(3);
        ids.reserve(5 * total);
        uint64_t hr[3] = {t_beg, t_end, 0};
        CUDA_CHECK(cudaMemcpyAsync(rng.get(), hr, 24, cudaMemcpyHostToDevice, h->st));
        gap_bucket_ids(h->d_text, h->n, gp, rng.get(), rng.get() + 1, rng.get() + 2, 1, total, ids.get(), h->st);
        t_ids.resize(5 * total);
        CUDA_CHECK(cudaMemcpyAsync(t_ids.data(), ids.get(), 5 * total * 4, cudaMemcpyDeviceToHost, h->st));
        CUDA_CHECK(cudaStreamSynchronize(h->st));
    }
    // bucket of the window of pattern x starting at pos (pos + patt_len[x] <= n)
    uint32_t get(uint64_t pos, int x) {
        if (pos >= t_beg && pos < t_end) return t_ids[5 * (pos - t_beg) + x];
        if (!beg.empty()) {
            if (cur >= beg.size() || pos < beg[cur] || pos >= end[cur]) {
                // ranges are sorted and disjoint
                size_t k = std::upper_bound(beg.begin(), beg.end(), pos) - beg.begin();
                if (k > 0 && pos < end[k - 1]) cur = k - 1;
                else { fetch_temp(pos); return t_ids[x]; }
            }
            return h->h_ids[5 * (off[cur] + (pos - beg[cur])) + x];
        }
        fetch_temp(pos);
        return t_ids[x];
    }
};

// rolling_hash_index_107 without the fingerprints: cur_pos, the table of last writers, and the rule for stale
// fingerprints at the end of the text (rolling_hash_index_107.hpp:80-150).  table_t: dense vector (whole text on the
// host) or the sparse view of the device table that the tail of the text needs.
template <typename ent_t>
struct dense_table {
    std::vector<ent_t> H;
    explicit dense_table(uint64_t entries) { H.assign(entries, (ent_t)~(ent_t)0); }
    uint64_t get(uint32_t b) const { return H[b] == (ent_t)~(ent_t)0 ? ~0ull : (uint64_t)H[b]; }
    void set(uint32_t b, uint64_t pos) { H[b] = (ent_t)pos; }
};
struct sparse_table {
    std::unordered_map<uint32_t, uint64_t> H;   // every bucket the tail can touch is pre-loaded from the device
    uint64_t get(uint32_t b) const {
        auto it = H.find(b);
        if (it == H.end()) throw std::logic_error("tail emission touched a bucket that was not fetched");
        return it->second;
    }
    void set(uint32_t b, uint64_t pos) {
        auto it = H.find(b);
        if (it == H.end()) throw std::logic_error("tail emission touched a bucket that was not fetched");
        it->second = pos;
    }
};
template <typename table_t>
struct gap_index {
    bucket_provider& bp;
    table_t& H;
    uint64_t n;
    std::array<uint64_t, 5> L;
    uint64_t cur_pos = 0;
    bool zero[5];   // fingerprint x was reset at a position where the window no longer fits: it stays 0
    gap_index(bucket_provider& bp_, table_t& H_, uint64_t n_, const gap_params& g, uint64_t start_pos)
        : bp(bp_), H(H_), n(n_), L(g.patt_lens) {
        reinit(start_pos);
    }
    void reinit(uint64_t pos) {
        cur_pos = pos;
        for (int x = 0; x < 5; x++) zero[x] = !(pos + L[x] < n);
    }
    void roll() { cur_pos++; }
    uint32_t bucket(int x) {
        if (cur_pos + L[x] < n) return bp.get(cur_pos, x);
        if (zero[x]) return 0;
        return bp.get(n - L[x], x);     // stale: the last fingerprint that was rolled in
    }
    void advance() {
        for (int x = 0; x < 5; x++)
            if (cur_pos + L[x] < n) H.set(bp.get(cur_pos, x), cur_pos);
        cur_pos++;
    }
    void advance_i(int x) {
        if (cur_pos + L[x] < n) H.set(bp.get(cur_pos, x), cur_pos);
    }
    uint64_t advance_and_get_occ(int x) {
        const uint32_t b = bucket(x);
        const uint64_t occ = H.get(b);
        H.set(b, cur_pos);
        return occ;
    }
};

inline uint64_t host_lce(const uint8_t* T, uint64_t n, uint64_t src, uint64_t pos) {   // src < pos
    const uint64_t lim = n - pos;
    uint64_t k = 0;
    while (k + 8 <= lim) {
        uint64_t a, b;
        std::memcpy(&a, T + src + k, 8);
        std::memcpy(&b, T + pos + k, 8);
        if (a != b) return k + (__builtin_ctzll(a ^ b) >> 3);
        k += 8;
    }
    while (k < lim && T[src + k] == T[pos + k]) k++;
    return k;
}

// The merged phrase list as three planes (the layout the device kernels read); element access for the host loops.
struct phrase_list {
    const uint64_t* p = nullptr;
    size_t m = 0;
    size_t size() const { return m; }
    lpf_t operator[](size_t k) const { return lpf_t{p[k], p[m + k], p[2 * m + k]}; }
};

// next_lpf (approximate/factorize/common.cpp:74-104) takes no argument, so the sequence of phrases it returns is
// fixed before the emission starts: the per-block lists concatenated, where a block continues behind the last phrase of
// the one before it (phrases that end at or before it are dropped, the first survivor is trimmed to start there), and
// the sentinel {n, n+1, 0} (lz77_sss.hpp:418,423) at the end.  Written as planes into the handle's pinned buffer (reused
// between calls: a fresh 100 MB allocation per call costs more in page faults than the whole device pipeline).
phrase_list merged_phrase_list(lz77sss_handle* h) {
    const uint64_t n = h->n;
    const uint32_t p = h->cfg.num_threads;
    const auto& LPF = h->LPF;
    struct part { size_t first, out; uint64_t cut; };      // first phrase taken, its slot in the output, where it must start
    std::vector<part> parts(p);
    size_t m = 0;
    uint64_t last_end = 0;
    bool any = false;
    for (uint32_t i_p = 0; i_p < p; i_p++) {
        const auto& V = LPF[i_p];
        size_t li = 0;
        if (any) while (li < V.size() && V[li].end <= last_end) li++;
        parts[i_p] = part{li, m, any ? last_end : 0};
        if (li < V.size()) {
            m += V.size() - li;
            last_end = V.back().end;                        // ends ascend inside a block's list
            any = true;
        }
    }
    m += 1;                                                 // the sentinel
    h->ml_buf.reserve(3 * m);
    uint64_t* B = h->ml_buf.p;
    run_blocks(p, [&](uint32_t i_p) {
        const auto& V = LPF[i_p];
        size_t o = parts[i_p].out;
        for (size_t li = parts[i_p].first; li < V.size(); li++, o++) {
            lpf_t ph = V[li];
            if (li == parts[i_p].first && ph.beg < parts[i_p].cut) {
                const uint64_t offs = parts[i_p].cut - ph.beg;
                ph.beg += offs;
                ph.src += offs;
            }
            B[o] = ph.beg;
            B[m + o] = ph.end;
            B[2 * m + o] = ph.src;
        }
    });
    B[m - 1] = n;
    B[2 * m - 1] = n + 1;
    B[3 * m - 1] = 0;
    h->ml_count = m;
    return phrase_list{B, m};
}

// factorize (approximate/factorize/common.cpp:66-111) + factorize_greedy (greedy.cpp:34-140) from the state
// (i, idx.cur_pos, k): ML[k] is the first phrase that ends behind i.
template <typename table_t>
uint64_t greedy_emit_from(const uint8_t* T, uint64_t n, const phrase_list& ML, const gap_params& g,
                          gap_index<table_t>& idx, uint64_t i, size_t k, lz77sss_emit_fn emit, void* ctx) {
    uint64_t num_fact = 0;
    const size_t last = ML.size() - 1;
    auto next_lpf = [&]() {
        if (k < last) k++;
        return ML[k];
    };
    auto longest_prev_occ = [&](uint64_t pos, uint64_t& f_src, uint64_t& f_len) {   // factorize/common.cpp:33-61
        f_src = T[pos];
        f_len = 0;
        for (int x = 4; x >= 0; x--) {
            if (f_len == 0) {
                const uint64_t src = idx.advance_and_get_occ(x);
                if (src < pos && T[src] == T[pos]) {
                    f_len = host_lce(T, n, src, pos);
                    f_src = src;
                }
            } else {
                idx.advance_i(x);
            }
        }
        idx.cur_pos++;
    };
    lpf_t ph = ML[k];
    while (true) {
        uint64_t gap_end = ph.beg;
        if (i < gap_end) {
            if (idx.cur_pos < i) {
                if (i - idx.cur_pos <= g.roll_threshold) {
                    do { idx.roll(); } while (idx.cur_pos < i);
                } else {
                    idx.reinit(i);
                }
            }
            do {
                uint64_t fs, fl;
                longest_prev_occ(i, fs, fl);
                i += std::max<uint64_t>(1, fl);
                if (i > gap_end) {
                    if (i <= ph.end) {
                        fl -= i - gap_end;
                        i = gap_end;
                    } else {
                        do { ph = next_lpf(); } while (ph.end <= i);
                        while (idx.cur_pos < gap_end) idx.advance();
                        gap_end = ph.beg;
                    }
                }
                emit(ctx, fs, fl);
                num_fact++;
                while (idx.cur_pos < i) idx.advance();
            } while (i < gap_end);
        }
        if (i == n) break;
        const uint64_t exc = i - gap_end;
        uint64_t ls = ph.src + exc, ll = (ph.end - ph.beg) - exc;
        if (idx.cur_pos == i) {
            uint64_t fs, fl;
            longest_prev_occ(i, fs, fl);
            if (fl > ll) { ls = fs; ll = fl; }
        }
        emit(ctx, ls, ll);
        num_fact++;
        i += ll;
        while (ph.end <= i) ph = next_lpf();
    }
    return num_fact;
}

// factorize_greedy_naive (greedy_naive.cpp:34-113) from the state (i, idx.cur_pos, k): ML[k] is the next phrase
// (ML[k].beg >= i).  Every position is inserted before the table is probed behind it.
template <typename table_t>
uint64_t greedy_naive_emit_from(const uint8_t* T, uint64_t n, const phrase_list& ML, gap_index<table_t>& idx,
                                uint64_t i, size_t k, lz77sss_emit_fn emit, void* ctx) {
    uint64_t num_fact = 0;
    const size_t last = ML.size() - 1;
    auto longest_prev_occ = [&](uint64_t pos, uint64_t& f_src, uint64_t& f_len) {
        f_src = T[pos];
        f_len = 0;
        for (int x = 4; x >= 0; x--) {
            if (f_len == 0) {
                const uint64_t src = idx.advance_and_get_occ(x);
                if (src < pos && T[src] == T[pos]) {
                    f_len = host_lce(T, n, src, pos);
                    f_src = src;
                }
            } else {
                idx.advance_i(x);
            }
        }
        idx.cur_pos++;
    };
    lpf_t ph = ML[k];
    while (true) {
        const uint64_t gap_end = ph.beg;
        if (i < gap_end) {
            while (idx.cur_pos < i) idx.advance();
            do {
                uint64_t fs, fl;
                longest_prev_occ(i, fs, fl);
                fl = std::min<uint64_t>(fl, gap_end - i);
                emit(ctx, fs, fl);
                num_fact++;
                i += std::max<uint64_t>(1, fl);
                while (idx.cur_pos < i) idx.advance();
            } while (i < gap_end);
        }
        if (gap_end == n) break;
        uint64_t ls = ph.src, ll = ph.end - ph.beg;
        if (k < last) k++;
        const lpf_t nxt = ML[k];
        if (idx.cur_pos == gap_end) {
            uint64_t fs, fl;
            longest_prev_occ(gap_end, fs, fl);
            fl = std::min<uint64_t>(fl, nxt.beg - gap_end);
            if (fl > ll) { ls = fs; ll = fl; }
        }
        emit(ctx, ls, ll);
        num_fact++;
        i += ll;
        ph = nxt;
    }
    return num_fact;
}

// factorize_skip_gaps (skip_gaps.cpp:34-61): the phrases and the lengths of the gaps between them
uint64_t skip_gaps_emit(uint64_t n, const phrase_list& ML, lz77sss_emit_fn emit, void* ctx) {
    uint64_t num = 0;
    size_t k = 0;
    const size_t last = ML.size() - 1;
    lpf_t nxt = ML[0];
    emit(ctx, nxt.beg, 0);
    num++;
    while (nxt.beg < n) {
        const lpf_t phr = nxt;
        if (k < last) k++;
        nxt = ML[k];
        emit(ctx, phr.src, phr.end - phr.beg);
        num++;
        if (nxt.beg > phr.end) {
            emit(ctx, nxt.beg - phr.end, 0);
            num++;
        }
    }
    return num;
}

// The last 65 positions (end-of-text rules of the gap index, rolling_hash_index_107.hpp:93,112,123,145) on the host,
// from the state the device parse handed over, with the table entries of every bucket they can touch.
uint64_t emit_tail(lz77sss_handle* h, const uint8_t* host_text, uint64_t n, const phrase_list& ML, const gap_params& g,
                   bucket_provider& bp, const uint64_t ho[4], bool naive, lz77sss_emit_fn emit, void* ctx) {
    auto& st = h->stats;
    cudaStream_t s = h->st;
    const uint64_t n_cut = n > 65 ? n - 65 : 0;
    uint64_t z = 0;
    {
        // tail: table entries of every bucket the remaining positions can touch
        // (greedy_naive: the host re-inserts position n_cut - 1, so that "table cursor == i" keeps its meaning)
        const uint64_t tb = n_cut - (naive && n_cut > 0 ? 1 : 0);
        bp.beg.push_back(tb);
        bp.end.push_back(n);
        bp.fetch_main();
        std::vector<uint32_t> buckets;
        buckets.push_back(0);
        for (uint64_t pos = tb; pos < n; pos++)
            for (int x = 0; x < 5; x++)
                if (pos + g.patt_lens[x] <= n) buckets.push_back(bp.get(pos, x));
        std::sort(buckets.begin(), buckets.end());
        buckets.erase(std::unique(buckets.begin(), buckets.end()), buckets.end());
        std::vector<uint64_t> vals(buckets.size());
        gap_parse_query(h->gp, buckets.data(), (uint32_t)buckets.size(), vals.data(), s);
        sparse_table H;
        for (size_t q = 0; q < buckets.size(); q++) H.H[buckets[q]] = vals[q] ? vals[q] - 1 : ~0ull;
        gap_index<sparse_table> idx(bp, H, n, g, 0);
        if (n_cut > 0) {                           // the device parsed up to n - 65: every window fitted at its last
            idx.cur_pos = ho[1];                   // reinit, so no fingerprint is stuck at zero
            for (int x = 0; x < 5; x++) idx.zero[x] = false;
            // greedy_naive: the device table already holds every position below n_cut
            if (naive && ho[1] != ho[0]) idx.cur_pos = std::max<uint64_t>(ho[1], n_cut - 1);
        }                                          // else (n <= 65): reinit(0) of the constructor set the flags
        if (naive) z += greedy_naive_emit_from(host_text, n, ML, idx, ho[0], (size_t)ho[2], emit, ctx);
        else z += greedy_emit_from(host_text, n, ML, g, idx, ho[0], (size_t)ho[2], emit, ctx);
        st.d2h_bytes += 20 * bp.hashed + 8 * buckets.size();
    }
    return z;
}

// hands z factors to the callback in order; the library's own array sink (lz77sss_sink_emit) is filled with two block
// copies instead of z indirect calls
void emit_all(const uint64_t* src, const uint64_t* len, uint64_t z, lz77sss_emit_fn emit, void* ctx) {
    if (emit == &lz77sss_sink_emit) {
        auto* s = static_cast<lz77sss_factor_sink*>(ctx);
        const uint64_t room = s->count < s->cap ? s->cap - s->count : 0, m = std::min(z, room);
        if (m < (1u << 20)) {
            std::memcpy(s->src + s->count, src, m * 8);
            std::memcpy(s->len + s->count, len, m * 8);
        } else {                  // tens of MB into (usually untouched) caller memory: page faults and copy on several threads
            const uint32_t parts = 8;
            const uint64_t at = s->count;
            run_blocks(parts, [&](uint32_t t) {
                const uint64_t b = m * t / parts, e = m * (t + 1) / parts;
                std::memcpy(s->src + at + b, src + b, (e - b) * 8);
                std::memcpy(s->len + at + b, len + b, (e - b) * 8);
            });
        }
        s->count += z;
        return;
    }
    for (uint64_t q = 0; q < z; q++) emit(ctx, src[q], len[q]);
}

template <typename pos_t>
int factorize_impl(lz77sss_handle* h, const uint8_t* d_text, const uint8_t* host_text, uint64_t n,
                   lz77sss_emit_fn emit, void* ctx) {
    auto& P = pipe<pos_t>(h);
    auto& st = h->stats;
    const uint32_t tau = h->cfg.tau;
    cudaStream_t s = h->st;
    h->d_text = d_text;
    h->n = n;
    h->pos_bytes = sizeof(pos_t);
    st.n = n;

    // build_lce (algorithms/common.cpp:58-80) on one direction of the text, the candidates of the phrase mode, and
    // their copy to the host
    uint64_t size_s = 0;
    auto build_and_fetch = [&](const uint8_t* text_x, int cand_mode /* 0 lpf_opt/naive, 1 LPF_all, 2 LNF_all */) {
        // K1+K2
        auto t0 = clk::now();
        uint64_t s_cap = n / 2 + 1024;                 // |S| <= number of candidate positions; typical 2n/tau
        if (tau >= 16) s_cap = std::min<uint64_t>(n, 8 * (n / tau) + 65536);
        P.S.reserve(s_cap);
        uint64_t numq = 0;
        sss_build(text_x, n, tau, h->cfg.sss_base, h->cfg.sss_qmax, P.S.get(), sizeof(pos_t), s_cap, P.sss, s, &size_s, &numq);
        if (size_s > s_cap) {                          // rare: denser than expected, redo with the exact size
            s_cap = size_s;
            P.S.reserve(s_cap);
            sss_build(text_x, n, tau, h->cfg.sss_base, h->cfg.sss_qmax, P.S.get(), sizeof(pos_t), s_cap, P.sss, s, &size_s, &numq);
        }
        P.s = size_s;
        st.size_sss = size_s;
        st.num_q = numq;
        st.t_sss_ns += ns_since(t0);

        // K3
        t0 = clk::now();
        suffix_sort_sync<pos_t>(text_x, n, tau, h->cfg.sss_qmax, P.S.get(), size_s, P.suf, s);
        CUDA_CHECK(cudaStreamSynchronize(s));
        st.ctx_levels = P.suf.ctx_levels;
        st.doubling_rounds = P.suf.doubling_rounds;
        st.t_sort_ns += ns_since(t0);

        // range-min / nearest-smaller / successor structures, LPF candidates
        t0 = clk::now();
        build_lce_structures<pos_t>(text_x, n, tau, P.S.get(), size_s, P.suf, P.lce, s);
        CUDA_CHECK(cudaStreamSynchronize(s));
        st.t_lce_ns += ns_since(t0);
        t0 = clk::now();
        if (g_sync_every_kernel) fprintf(stderr, "candidates: mode %d, |S| = %llu, nblk = %u, levels = %d\n", cand_mode,
                                         (unsigned long long)size_s, P.lce.nblk, P.lce.sp_levels);
        if (cand_mode == 0) lpf_candidates<pos_t>(text_x, n, tau, P.S.get(), size_s, h->cfg.num_threads, P.suf, P.lce, s);
        else all_candidates<pos_t>(text_x, n, P.S.get(), size_s, cand_mode == 2,
                                   h->cfg.phr_mode == LZ77SSS_PHR_LPF_LNF_OPT ? LNF_LCAP : 0, P.suf, P.lce, s);
        candidate_sources<pos_t>(P.S.get(), size_s, P.lce, s);
        CUDA_CHECK(cudaStreamSynchronize(s));
        st.t_lpf_ns += ns_since(t0);

        // candidates to the host
        t0 = clk::now();
        if (size_s) {
            P.hS.reserve(size_s); P.h_psrc.reserve(size_s); P.h_nsrc.reserve(size_s);
            P.h_plce.reserve(size_s); P.h_nlce.reserve(size_s); P.h_plcel.reserve(size_s); P.h_nlcel.reserve(size_s);
            CUDA_CHECK(cudaMemcpyAsync(P.hS.p, P.S.get(), size_s * sizeof(pos_t), cudaMemcpyDeviceToHost, s));
            CUDA_CHECK(cudaMemcpyAsync(P.h_psrc.p, P.lce.psv_src.get(), size_s * sizeof(pos_t), cudaMemcpyDeviceToHost, s));
            CUDA_CHECK(cudaMemcpyAsync(P.h_nsrc.p, P.lce.nsv_src.get(), size_s * sizeof(pos_t), cudaMemcpyDeviceToHost, s));
            CUDA_CHECK(cudaMemcpyAsync(P.h_plce.p, P.lce.psv_lce.get(), size_s * sizeof(pos_t), cudaMemcpyDeviceToHost, s));
            CUDA_CHECK(cudaMemcpyAsync(P.h_nlce.p, P.lce.nsv_lce.get(), size_s * sizeof(pos_t), cudaMemcpyDeviceToHost, s));
            CUDA_CHECK(cudaMemcpyAsync(P.h_plcel.p, P.lce.psv_lcel.get(), size_s * sizeof(pos_t), cudaMemcpyDeviceToHost, s));
            CUDA_CHECK(cudaMemcpyAsync(P.h_nlcel.p, P.lce.nsv_lcel.get(), size_s * sizeof(pos_t), cudaMemcpyDeviceToHost, s));
            CUDA_CHECK(cudaStreamSynchronize(s));
            st.d2h_bytes += size_s * 7 * sizeof(pos_t);
        }
        st.t_d2h_ns += ns_since(t0);
    };
    const bool lnf_mode = h->cfg.phr_mode == LZ77SSS_PHR_LPF_LNF_NAIVE || h->cfg.phr_mode == LZ77SSS_PHR_LPF_LNF_OPT;
    const bool opt_mode = h->cfg.phr_mode == LZ77SSS_PHR_LPF_LNF_OPT;
    auto t0 = clk::now();
    if (lnf_mode) {
        // compute_approximation, lz77_sss.hpp:384-394: LNF phrases from the reversed text, then LPF phrases
        h->LPF.assign(h->cfg.num_threads, {});
        uint8_t* d_rev = alloc_text(h->rev_buf, n, s);
        reverse_text(d_text, d_rev, n, s);
        build_and_fetch(d_rev, 2);
        t0 = clk::now();
        all_chain<pos_t>(h, P, d_rev, true, opt_mode);
        st.t_chain_ns += ns_since(t0);
        build_and_fetch(d_text, 1);
        t0 = clk::now();
        all_chain<pos_t>(h, P, d_text, false, opt_mode);
        run_blocks(h->cfg.num_threads, [&](uint32_t i_p) { select_phrases(h->LPF[i_p]); });
        st.t_chain_ns += ns_since(t0);
    } else {
        build_and_fetch(d_text, 0);
    }

    // coverage chain, statistics, gap parameters (host scalars)
    t0 = clk::now();
    const bool naive = h->cfg.fact_mode == LZ77SSS_FACT_GREEDY_NAIVE;
    if (h->cfg.phr_mode == LZ77SSS_PHR_LPF_NAIVE) lpf_chain_naive<pos_t>(h, P);
    else if (!lnf_mode) lpf_chain<pos_t>(h, P);
    if (h->cfg.fact_mode == LZ77SSS_FACT_SKIP_PHRASES) {   // no statistics, no gap index (lz77_sss.hpp:417-418)
        st.t_chain_ns += ns_since(t0);
        h->have_result = true;
        t0 = clk::now();
        st.num_factors = skip_gaps_emit(n, merged_phrase_list(h), emit, ctx);
        st.t_emit_ns = ns_since(t0);
        return LZ77SSS_OK;
    }
    phrase_info(h);
    gap_params g = choose_gap_params(h);
    for (int x = 0; x < 5; x++) st.patt_lens[x] = g.patt_lens[x];
    st.roll_threshold = g.roll_threshold;
    st.table_entries = g.table_entries;
    st.t_chain_ns += ns_since(t0);
    h->have_result = true;
    if (g.patt_lens[0] == 0) return fail(LZ77SSS_E_UNSUPPORTED, "degenerate phrase statistics (no LPF phrase or no gap): pattern lengths undefined in the reference (SURVEY.md A7)");

    const phrase_list ML = merged_phrase_list(h);
    if (h->prepare_only) {                     // lz77sss_shard_prepare: the ranges are parsed by lz77sss_shard_parse
        h->table_entries = g.table_entries;
        h->host_text = host_text;
        h->prepared = true;
        return LZ77SSS_OK;
    }
    bucket_provider bp;
    bp.h = h;
    for (int x = 0; x < 5; x++) { bp.gp.patt_len[x] = g.patt_lens[x]; bp.gp.base[x] = h->cfg.gap_bases[x]; }
    bp.gp.mask = g.table_entries - 1;

    if (h->host_emit) {
        // debugging path (LZ77SSS_HOST_EMIT=1): the whole emission on the host, bucket ids from the device
        t0 = clk::now();
        {
            uint64_t cur = 0;
            for (auto& V : h->LPF)
                for (auto& ph : V) {
                    if (ph.end <= cur) continue;
                    const uint64_t b = std::max(ph.beg, cur);
                    bp.beg.push_back(cur);
                    bp.end.push_back(b + 1);
                    cur = ph.end;
                }
            if (cur < n) { bp.beg.push_back(cur); bp.end.push_back(n); }
        }
        bp.fetch_main();
        st.t_gap_hash_ns = ns_since(t0);
        st.h2d_bytes += 24 * bp.beg.size();
        t0 = clk::now();
        uint64_t z;
        if (h->cfg.pos_bits == 32) {
            dense_table<uint32_t> H(g.table_entries);
            gap_index<dense_table<uint32_t>> idx(bp, H, n, g, 0);
            z = naive ? greedy_naive_emit_from(host_text, n, ML, idx, 0, 0, emit, ctx)
                      : greedy_emit_from(host_text, n, ML, g, idx, 0, 0, emit, ctx);
        } else {
            dense_table<uint64_t> H(g.table_entries);
            gap_index<dense_table<uint64_t>> idx(bp, H, n, g, 0);
            z = naive ? greedy_naive_emit_from(host_text, n, ML, idx, 0, 0, emit, ctx)
                      : greedy_emit_from(host_text, n, ML, g, idx, 0, 0, emit, ctx);
        }
        st.num_factors = z;
        st.gap_positions_hashed = bp.hashed;
        st.d2h_bytes += 20 * bp.hashed;
        st.t_emit_ns = ns_since(t0);
        return LZ77SSS_OK;
    }

    // greedy emission on the device up to n_cut; the last 65 positions (end-of-text rules of the gap index) on the host
    t0 = clk::now();
    const uint64_t n_cut = n > 65 ? n - 65 : 0;
    h->f_src.clear();
    h->f_len.clear();
    uint64_t ho[4];
    gap_hash_params gpar = bp.gp;
    gap_parse_run<pos_t>(d_text, n, tau, P.S.get(), size_s, P.suf, P.lce, gpar, g.table_entries, ML.p, ML.size(), n_cut,
                         h->slice_len, h->block_len, h->gp, s, h->f_src, h->f_len, ho, naive);
    st.h2d_bytes += 24 * ML.size();
    st.d2h_bytes += 16 * h->f_src.size();
    st.gap_positions_hashed = h->gp.entries_sorted / 5;
    st.parse_rounds = h->gp.rounds;
    st.parse_slices = h->gp.slices;
    st.sparse_slices = h->gp.sparse_slices;
    st.sparse_fallbacks = h->gp.sparse_fallbacks;
    st.repair_entries = h->gp.repair_entries;
    st.prefetched_slices = h->gp.prefetched_slices;
    st.t_gap_hash_ns = ns_since(t0);

    t0 = clk::now();
    uint64_t z = h->f_src.size();
    emit_all(h->f_src.data(), h->f_len.data(), z, emit, ctx);
    if (ho[0] < n) z += emit_tail(h, host_text, n, ML, g, bp, ho, naive, emit, ctx);
    st.num_factors = z;
    st.t_emit_ns = ns_since(t0);
    return LZ77SSS_OK;
}

int factorize_dispatch(lz77sss_handle* h, const uint8_t* d_text, const uint8_t* host_text, uint64_t n,
                       lz77sss_emit_fn emit, void* ctx) {
    if (!h || !host_text || !emit) return fail(LZ77SSS_E_ARG, "null argument");
    if (n == 0) return fail(LZ77SSS_E_ARG, "empty input");
    if (h->cfg.phr_mode > LZ77SSS_PHR_LPF_LNF_OPT || h->cfg.fact_mode > LZ77SSS_FACT_SKIP_PHRASES)
        return fail(LZ77SSS_E_ARG, "unknown phrase or factorization mode");
    if (h->cfg.num_threads > n) return fail(LZ77SSS_E_ARG, "num_threads > n");
    if (h->cfg.pos_bits == 32 && n > 0xFFFFFFFFull) return fail(LZ77SSS_E_ARG, "pos_bits == 32 but n > UINT32_MAX");
    std::memset(&h->stats, 0, sizeof h->stats);
    h->have_result = false;
    const unsigned long long l0 = g_kernel_launches;
    auto t0 = clk::now();
    int rc;
    // internal position width; LZ77SSS_FORCE_POS64=1 takes the 64-bit kernels for small inputs too (tests)
    static const bool force64 = getenv("LZ77SSS_FORCE_POS64") && getenv("LZ77SSS_FORCE_POS64")[0] == '1';
    if (n < 0xFFFF0000ull && !force64) rc = factorize_impl<uint32_t>(h, d_text, host_text, n, emit, ctx);
    else rc = factorize_impl<uint64_t>(h, d_text, host_text, n, emit, ctx);
    h->stats.t_total_ns += ns_since(t0);
    h->stats.kernel_launches = g_kernel_launches - l0;
    if (h->cfg.log) {
        auto& s = h->stats;
        printf("lz77sss_b200: n=%llu |S|=%llu z=%llu  h2d=%.3f sss=%.3f sort=%.3f lce=%.3f lpf=%.3f d2h=%.3f chain=%.3f gap_hash=%.3f emit=%.3f total=%.3f ms\n",
               (unsigned long long)s.n, (unsigned long long)s.size_sss, (unsigned long long)s.num_factors, s.t_h2d_ns / 1e6,
               s.t_sss_ns / 1e6, s.t_sort_ns / 1e6, s.t_lce_ns / 1e6, s.t_lpf_ns / 1e6, s.t_d2h_ns / 1e6, s.t_chain_ns / 1e6,
               s.t_gap_hash_ns / 1e6, s.t_emit_ns / 1e6, s.t_total_ns / 1e6);
    }
    return rc;
}

gap_params stored_gap_params(const lz77sss_handle* h) {
    gap_params g;
    for (int x = 0; x < 5; x++) g.patt_lens[x] = h->stats.patt_lens[x];
    g.roll_threshold = h->stats.roll_threshold;
    g.table_entries = h->table_entries;
    return g;
}

// One range of a sharded run: the device emission over [beg, end) from the given entry state and table.
template <typename pos_t>
int shard_parse_impl(lz77sss_handle* h, uint64_t beg, uint64_t end, const uint64_t entry[4], const uint64_t* d_table_in,
                     uint64_t* d_table_out, uint64_t exit_state[4], uint64_t* num_factors, bool index_only = false) {
    auto& P = pipe<pos_t>(h);
    const uint64_t n = h->n;
    const gap_params g = stored_gap_params(h);
    gap_hash_params gpar;
    for (int x = 0; x < 5; x++) { gpar.patt_len[x] = g.patt_lens[x]; gpar.base[x] = h->cfg.gap_bases[x]; }
    gpar.mask = g.table_entries - 1;
    const size_t m = h->ml_count;
    h->f_src.clear();
    h->f_len.clear();
    gap_parse_range R;
    R.beg = beg; R.end = end;
    for (int k = 0; k < 4; k++) R.entry[k] = entry[k];
    R.d_table_in = d_table_in;
    R.d_table_out = d_table_out;
    R.retain = true;
    R.index_only = index_only;
    uint64_t ho[4];
    const uint64_t n_cut = n > 65 ? n - 65 : 0;
    auto t0 = clk::now();
    gap_parse_run<pos_t>(h->d_text, n, h->cfg.tau, P.S.get(), P.s, P.suf, P.lce, gpar, g.table_entries, h->ml_buf.p, m,
                         n_cut, h->slice_len, h->block_len, h->gp, h->st, h->f_src, h->f_len, ho,
                         h->cfg.fact_mode == LZ77SSS_FACT_GREEDY_NAIVE, &R);
    h->stats.t_gap_hash_ns += ns_since(t0);
    h->stats.parse_rounds += h->gp.rounds;
    h->stats.parse_slices += h->gp.slices;
    h->stats.gap_positions_hashed += h->gp.entries_sorted / 5;
    h->stats.d2h_bytes += 16 * h->f_src.size();
    for (int k = 0; k < 4; k++) exit_state[k] = ho[k];
    if (num_factors) *num_factors = h->f_src.size();
    return LZ77SSS_OK;
}

template <typename pos_t, typename out_t>
int copy_out(const pos_t* d, uint64_t cnt, out_t* out, uint64_t cap, uint64_t* count, cudaStream_t st) {
    if (count) *count = cnt;
    if (!out) return LZ77SSS_OK;
    uint64_t m = std::min(cnt, cap);
    std::vector<pos_t> tmp(m);
    if (m) {
        CUDA_CHECK(cudaMemcpyAsync(tmp.data(), d, m * sizeof(pos_t), cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
    }
    for (uint64_t i = 0; i < m; i++) out[i] = (out_t)tmp[i];
    return LZ77SSS_OK;
}

}  // namespace

extern "C" {
#pragma GCC visibility push(default)

void lz77sss_sink_emit(void* ctx, uint64_t src, uint64_t len) {
    auto* s = static_cast<lz77sss_factor_sink*>(ctx);
    if (s->count < s->cap) {
        s->src[s->count] = src;
        s->len[s->count] = len;
    }
    s->count++;
}

const char* lz77sss_last_error(void) { return g_last_error.c_str(); }
const char* lz77sss_version(void) { return "lz77sss_b200 0.1 (sm_100a)"; }

int lz77sss_device_count(void) {
    int c = 0;
    if (cudaGetDeviceCount(&c) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return c;
}

int lz77sss_create(const lz77sss_config* cfg, lz77sss_handle** out) {
    if (!cfg || !out) return fail(LZ77SSS_E_ARG, "null argument");
    if (cfg->struct_size != sizeof(lz77sss_config)) return fail(LZ77SSS_E_ARG, "lz77sss_config.struct_size mismatch");
    if (!valid_tau(cfg->tau)) return fail(LZ77SSS_E_ARG, "tau must be a power of two in [4, 4096]");
    if (cfg->pos_bits != 0 && cfg->pos_bits != 32 && cfg->pos_bits != 64) return fail(LZ77SSS_E_ARG, "pos_bits must be 0, 32 or 64");
    for (int x = 0; x < 5; x++)
        if (cfg->gap_bases[x] < 257 || cfg->gap_bases[x] >= (1ull << 20))
            return fail(LZ77SSS_E_ARG, "gap_bases must lie in [257, 2^20-1] (rolling_hash.hpp:31-36)");
    return guarded([&]() {
        int ndev = 0;
        CUDA_CHECK(cudaGetDeviceCount(&ndev));
        if (cfg->device < 0 || cfg->device >= ndev) return fail(LZ77SSS_E_CUDA, "no such CUDA device (there is no CPU fallback)");
        CUDA_CHECK(cudaSetDevice(cfg->device));
        auto h = std::make_unique<lz77sss_handle>();
        h->cfg = *cfg;
        if (h->cfg.num_threads == 0) h->cfg.num_threads = 1;
        if (h->cfg.sss_base == 0) h->cfg.sss_base = 0x9E3779B1u;
        if (h->cfg.sss_qmax == 0xFFFFFFFFu) h->cfg.sss_qmax = 32;
        if ((h->cfg.sss_base & 1) == 0) return fail(LZ77SSS_E_ARG, "sss_base must be odd");
        if (h->cfg.sss_qmax > 32) return fail(LZ77SSS_E_ARG, "sss_qmax must be <= 32 (one warp lane per candidate period)");
        std::memset(&h->stats, 0, sizeof h->stats);
        if (const char* e = getenv("LZ77SSS_HOST_EMIT")) h->host_emit = e[0] == '1';
        if (const char* e = getenv("LZ77SSS_SLICE_LEN")) h->slice_len = std::max<uint64_t>(64, strtoull(e, nullptr, 10));
        if (const char* e = getenv("LZ77SSS_BLOCK_LEN")) h->block_len = (uint32_t)std::max<uint64_t>(32, strtoull(e, nullptr, 10));
        {   // the main stream outranks the side stream that builds the next slice's index during a parse (gap_parse.cu)
            int least = 0, greatest = 0;
            CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&least, &greatest));
            CUDA_CHECK(cudaStreamCreateWithPriority(&h->st, cudaStreamNonBlocking, greatest));
        }
        *out = h.release();
        return LZ77SSS_OK;
    });
}

void lz77sss_destroy(lz77sss_handle* h) {
    if (!h) return;
    cudaSetDevice(h->cfg.device);
    for (auto* p : h->user_texts) cudaFree(p - FRONT_PAD);
    if (h->st) cudaStreamDestroy(h->st);
    delete h;
}

int lz77sss_factorize_approximate(lz77sss_handle* h, const uint8_t* text, uint64_t n, lz77sss_emit_fn emit, void* ctx) {
    if (!h || !text) return fail(LZ77SSS_E_ARG, "null argument");
    if (n == 0) return fail(LZ77SSS_E_ARG, "empty input");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        pos_bits_guard guard(h->cfg, n);
        auto t0 = clk::now();
        uint8_t* d = alloc_text(h->text_buf, n, h->st);
        CUDA_CHECK(cudaMemcpyAsync(d, text, n, cudaMemcpyHostToDevice, h->st));
        CUDA_CHECK(cudaStreamSynchronize(h->st));
        uint64_t t_h2d = ns_since(t0);
        int rc = factorize_dispatch(h, d, text, n, emit, ctx);
        h->stats.t_h2d_ns = t_h2d;
        h->stats.h2d_bytes += n;
        h->stats.t_total_ns += t_h2d;
        return rc;
    });
}

int lz77sss_factorize_exact(lz77sss_handle* h, const uint8_t* text, uint64_t n, lz77sss_emit_fn emit, void* ctx) {
    if (!h || !text || !emit) return fail(LZ77SSS_E_ARG, "null argument");
    if (n == 0) return fail(LZ77SSS_E_ARG, "empty input");
    if (h->cfg.fact_mode == LZ77SSS_FACT_SKIP_PHRASES) return fail(LZ77SSS_E_ARG, "factorize_exact: fact_mode must not be skip_phrases (lz77_sss.hpp:319)");
    if (n >= 0xFFFF0000ull) return fail(LZ77SSS_E_UNSUPPORTED, "factorize_exact: n must be below 2^32 - 2^16 (the suffix array of a longer text does not fit one GPU)");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        pos_bits_guard guard(h->cfg, n);
        auto t0 = clk::now();
        uint8_t* d = alloc_text(h->text_buf, n, h->st);
        CUDA_CHECK(cudaMemcpyAsync(d, text, n, cudaMemcpyHostToDevice, h->st));
        CUDA_CHECK(cudaStreamSynchronize(h->st));
        const uint64_t t_h2d = ns_since(t0);
        // the approximation (lz77_sss.hpp:320-325): its phrase boundaries place the sections, its LCE structure answers
        // the parse's longest-common-extension queries
        std::vector<uint64_t> aprx_len;
        auto collect = +[](void* c, uint64_t, uint64_t len) { static_cast<std::vector<uint64_t>*>(c)->push_back(len ? len : 1); };
        int rc = factorize_dispatch(h, d, text, n, collect, &aprx_len);
        if (rc) return rc;
        const uint64_t z = aprx_len.size();
        const uint32_t p = h->cfg.num_threads;
        const uint64_t nsect = p == 1 ? 1 : (uint64_t)p * 16;               // num_par_sect_per_thr (lz77_sss.hpp:94)
        std::vector<uint64_t> sect{0};
        if (nsect > 1 && z / nsect > 0) {       // section k starts where phrase k * (z / nsect) of the approximation starts
            const uint64_t step = z / nsect;
            uint64_t pos = 0, next = step, k = 1;
            for (uint64_t f = 0; f < z && k < nsect; f++) {
                if (f == next) {
                    sect.push_back(pos);
                    k++;
                    next = k * step;
                }
                pos += aprx_len[f];
            }
        }
        sect.push_back(n);
        // the emission buffers of the approximation (a slice's index, traces) are not needed any more; the suffix array
        // and its sort buffers (28 bytes per text byte) take their place
        h->gp = gap_parse_scratch();
        t0 = clk::now();
        h->f_src.clear();
        h->f_len.clear();
        if (h->pos_bytes == 4) {
            auto& P = h->p32;
            exact_run<uint32_t>(d, n, h->cfg.tau, P.S.get(), P.s, P.suf, P.lce, sect, h->ex, h->st, h->f_src, h->f_len);
        } else {
            auto& P = h->p64;
            exact_run<uint64_t>(d, n, h->cfg.tau, P.S.get(), P.s, P.suf, P.lce, sect, h->ex, h->st, h->f_src, h->f_len);
        }
        h->stats.t_exact_ns = ns_since(t0);
        h->stats.exact_doubling_rounds = h->ex.rounds;
        h->stats.exact_parse_rounds = h->ex.parse_rounds;
        const uint64_t ze = h->f_src.size();
        emit_all(h->f_src.data(), h->f_len.data(), ze, emit, ctx);
        h->stats.num_factors = ze;
        h->stats.t_h2d_ns = t_h2d;
        h->stats.h2d_bytes += n;
        h->stats.d2h_bytes += 16 * ze;
        h->stats.t_total_ns += t_h2d + h->stats.t_exact_ns;
        return LZ77SSS_OK;
    });
}

int lz77sss_device_text_alloc(lz77sss_handle* h, uint64_t n, uint8_t** d_text) {
    if (!h || !d_text) return fail(LZ77SSS_E_ARG, "null argument");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        uint8_t* p = nullptr;
        CUDA_CHECK(cudaMalloc(&p, FRONT_PAD + n + TEXT_PAD));
        CUDA_CHECK(cudaMemsetAsync(p, 0, FRONT_PAD, h->st));
        CUDA_CHECK(cudaMemsetAsync(p + FRONT_PAD + n, 0, TEXT_PAD, h->st));
        CUDA_CHECK(cudaStreamSynchronize(h->st));
        h->user_texts.push_back(p + FRONT_PAD);
        *d_text = p + FRONT_PAD;
        return LZ77SSS_OK;
    });
}

int lz77sss_device_text_upload(lz77sss_handle* h, uint8_t* d_text, const uint8_t* host_text, uint64_t n) {
    if (!h || !d_text || !host_text) return fail(LZ77SSS_E_ARG, "null argument");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        CUDA_CHECK(cudaMemcpyAsync(d_text, host_text, n, cudaMemcpyHostToDevice, h->st));
        CUDA_CHECK(cudaStreamSynchronize(h->st));
        return LZ77SSS_OK;
    });
}

int lz77sss_device_text_free(lz77sss_handle* h, uint8_t* d_text) {
    if (!h || !d_text) return fail(LZ77SSS_E_ARG, "null argument");
    auto it = std::find(h->user_texts.begin(), h->user_texts.end(), d_text);
    if (it == h->user_texts.end()) return fail(LZ77SSS_E_ARG, "not a buffer of this handle");
    h->user_texts.erase(it);
    cudaSetDevice(h->cfg.device);
    cudaFree(d_text - FRONT_PAD);
    return LZ77SSS_OK;
}

int lz77sss_factorize_approximate_device(lz77sss_handle* h, const uint8_t* d_text, const uint8_t* host_text, uint64_t n,
                                         lz77sss_emit_fn emit, void* ctx) {
    if (!h || !d_text) return fail(LZ77SSS_E_ARG, "null argument");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        pos_bits_guard guard(h->cfg, n);
        return factorize_dispatch(h, d_text, host_text, n, emit, ctx);
    });
}

#define NEED_PREPARED() \
    if (!h) return fail(LZ77SSS_E_ARG, "null handle"); \
    if (!h->prepared) return fail(LZ77SSS_E_STATE, "lz77sss_shard_prepare has not run on this handle")

int lz77sss_shard_prepare(lz77sss_handle* h, const uint8_t* d_text, const uint8_t* host_text, uint64_t n) {
    if (!h || !d_text || !host_text) return fail(LZ77SSS_E_ARG, "null argument");
    if (h->cfg.fact_mode == LZ77SSS_FACT_SKIP_PHRASES) return fail(LZ77SSS_E_UNSUPPORTED, "skip_phrases has no gap index to shard");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        pos_bits_guard guard(h->cfg, n);
        h->prepared = false;
        h->prepare_only = true;
        for (auto& K : h->gp.keep) K->sorted = K->parsed = false;     // the buffers stay, their contents are stale
        auto nop = +[](void*, uint64_t, uint64_t) {};
        int rc = factorize_dispatch(h, d_text, host_text, n, nop, nullptr);
        h->prepare_only = false;
        h->stats.t_gap_hash_ns = 0;
        return rc;
    });
}

int lz77sss_shard_table_entries(lz77sss_handle* h, uint64_t* entries) {
    NEED_PREPARED();
    if (!entries) return fail(LZ77SSS_E_ARG, "null argument");
    *entries = h->table_entries;
    return LZ77SSS_OK;
}

int lz77sss_shard_entry_guess(lz77sss_handle* h, uint64_t pos, uint64_t state[4]) {
    NEED_PREPARED();
    if (!state || pos >= h->n) return fail(LZ77SSS_E_ARG, "bad argument");
    // the speculation the blocks of a slice start from (gap_parse.cu, spec_entry_kernel)
    const phrase_list ML{h->ml_buf.p, h->ml_count};
    size_t lo = 0, hi = ML.size() - 1;              // first phrase with end > pos (the sentinel ends at n + 1)
    while (lo < hi) {
        const size_t mid = (lo + hi) / 2;
        if (ML[mid].end > pos) hi = mid;
        else lo = mid + 1;
    }
    state[0] = pos; state[1] = pos; state[3] = 1;
    if (pos > 0 && ML[lo].beg < pos && h->cfg.fact_mode != LZ77SSS_FACT_GREEDY_NAIVE) {
        state[0] = ML[lo].end;
        state[1] = ML[lo].beg + 1 < state[0] ? ML[lo].beg + 1 : ML[lo].beg;
        state[3] = 0;
        while (lo + 1 < ML.size() && ML[lo].end <= state[0]) lo++;
    }
    state[2] = lo;
    return LZ77SSS_OK;
}

int lz77sss_shard_split(lz77sss_handle* h, uint32_t parts, uint64_t align, uint64_t* bounds) {
    NEED_PREPARED();
    if (!bounds || !parts || !align) return fail(LZ77SSS_E_ARG, "bad argument");
    // The work of a range is its gap bytes (hashed, sorted, probed); the inside of a phrase costs a bitmap bit.  Cut where
    // the running weight (16 per gap byte, 1 per phrase byte) passes k / parts of the total, rounded to `align`.
    const phrase_list ML{h->ml_buf.p, h->ml_count};
    const uint64_t n = h->n, m = ML.size() - 1;        // (the last entry is the sentinel)
    auto weight_to = [&](uint64_t k, uint64_t& gaps, uint64_t& phr, uint64_t& at) {   // after phrase k: totals up to its end
        gaps += ML[k].beg > at ? ML[k].beg - at : 0;
        phr += ML[k].end - std::max(ML[k].beg, at);
        at = std::max(at, (uint64_t)ML[k].end);
    };
    uint64_t gaps = 0, phr = 0, at = 0;
    for (uint64_t k = 0; k < m; k++) weight_to(k, gaps, phr, at);
    gaps += n > at ? n - at : 0;
    const long double total = 16.0L * gaps + phr;
    bounds[0] = 0;
    uint64_t g2 = 0, p2 = 0, at2 = 0, k = 0;
    for (uint32_t r = 1; r < parts; r++) {
        const long double want = total * r / parts;
        while (k < m) {
            uint64_t g3 = g2, p3 = p2, a3 = at2;
            weight_to(k, g3, p3, a3);
            if (16.0L * g3 + p3 > want) break;
            g2 = g3; p2 = p3; at2 = a3; k++;
        }
        // inside the gap in front of phrase k, or inside phrase k: interpolate
        uint64_t cut = at2;
        const long double have = 16.0L * g2 + p2;
        const uint64_t gap_len = k < m && ML[k].beg > at2 ? ML[k].beg - at2 : (k == m ? n - at2 : 0);
        if (have + 16.0L * gap_len >= want) cut = at2 + (uint64_t)((want - have) / 16.0L);
        else cut = at2 + gap_len + (uint64_t)(want - have - 16.0L * gap_len);
        cut = std::min(n, (cut + align / 2) / align * align);
        bounds[r] = std::max(cut, bounds[r - 1]);
    }
    bounds[parts] = n;
    return LZ77SSS_OK;
}

int lz77sss_shard_parse(lz77sss_handle* h, uint64_t beg, uint64_t end, const uint64_t entry[4], const uint64_t* d_table_in,
                        uint64_t* d_table_out, uint64_t exit_state[4], uint64_t* num_factors) {
    NEED_PREPARED();
    if (!entry || !exit_state || beg >= end || end > h->n) return fail(LZ77SSS_E_ARG, "bad argument");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        const unsigned long long l0 = g_kernel_launches;
        int rc = h->pos_bytes == 4 ? shard_parse_impl<uint32_t>(h, beg, end, entry, d_table_in, d_table_out, exit_state, num_factors)
                                   : shard_parse_impl<uint64_t>(h, beg, end, entry, d_table_in, d_table_out, exit_state, num_factors);
        h->stats.kernel_launches += g_kernel_launches - l0;
        return rc;
    });
}

int lz77sss_shard_index(lz77sss_handle* h, uint64_t beg, uint64_t end, uint64_t* d_table_guess_out) {
    NEED_PREPARED();
    if (beg >= end || end > h->n) return fail(LZ77SSS_E_ARG, "bad argument");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        const unsigned long long l0 = g_kernel_launches;
        const uint64_t entry[4] = {beg, beg, 0, 1};
        uint64_t ex[4];
        int rc = h->pos_bytes == 4 ? shard_parse_impl<uint32_t>(h, beg, end, entry, nullptr, d_table_guess_out, ex, nullptr, true)
                                   : shard_parse_impl<uint64_t>(h, beg, end, entry, nullptr, d_table_guess_out, ex, nullptr, true);
        h->stats.kernel_launches += g_kernel_launches - l0;
        return rc;
    });
}

int lz77sss_shard_factors(lz77sss_handle* h, uint64_t* src, uint64_t* len, uint64_t cap, uint64_t* count) {
    NEED_PREPARED();
    const uint64_t z = h->f_src.size();
    if (count) *count = z;
    const uint64_t m = std::min(z, cap);
    if (src) std::memcpy(src, h->f_src.data(), m * 8);
    if (len) std::memcpy(len, h->f_len.data(), m * 8);
    return LZ77SSS_OK;
}

int lz77sss_shard_factors_view(lz77sss_handle* h, const uint64_t** src, const uint64_t** len, uint64_t* count) {
    NEED_PREPARED();
    if (!src || !len || !count) return fail(LZ77SSS_E_ARG, "null argument");
    *src = h->f_src.data();
    *len = h->f_len.data();
    *count = h->f_src.size();
    return LZ77SSS_OK;
}

int lz77sss_shard_finish(lz77sss_handle* h, const uint64_t state[4], lz77sss_emit_fn emit, void* ctx) {
    NEED_PREPARED();
    if (!state || !emit) return fail(LZ77SSS_E_ARG, "null argument");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        if (state[0] >= h->n) return LZ77SSS_OK;
        const gap_params g = stored_gap_params(h);
        bucket_provider bp;
        bp.h = h;
        for (int x = 0; x < 5; x++) { bp.gp.patt_len[x] = g.patt_lens[x]; bp.gp.base[x] = h->cfg.gap_bases[x]; }
        bp.gp.mask = g.table_entries - 1;
        h->stats.num_factors += emit_tail(h, h->host_text, h->n, phrase_list{h->ml_buf.p, h->ml_count}, g, bp, state,
                                          h->cfg.fact_mode == LZ77SSS_FACT_GREEDY_NAIVE, emit, ctx);
        return LZ77SSS_OK;
    });
}

// ---- rabin_karp_substring<31, pos_t> (rabin_karp_substring.hpp:35-240) ----
int lz77sss_rks_create(lz77sss_handle* h, const uint8_t* text, uint64_t n, uint64_t base, lz77sss_rks** out) {
    if (!h || !out || (!text && n)) return fail(LZ77SSS_E_ARG, "null argument");
    if (base < 257 || base > 0x7FFFFFFEull) return fail(LZ77SSS_E_ARG, "base must lie in [257, 2^31 - 2] (rabin_karp_substring.hpp:92-96)");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        auto R = std::make_unique<lz77sss_rks>();
        R->h = h;
        R->d_text = alloc_text(R->text, n, h->st);
        if (n) CUDA_CHECK(cudaMemcpyAsync(R->d_text, text, n, cudaMemcpyHostToDevice, h->st));
        const unsigned long long l0 = g_kernel_launches;
        rks_build(R->d_text, n, (uint32_t)base, R->st, h->st);
        CUDA_CHECK(cudaStreamSynchronize(h->st));
        R->launches = g_kernel_launches - l0;
        *out = R.release();
        return LZ77SSS_OK;
    });
}

int lz77sss_rks_table_noalloc(const uint8_t* d_text, uint64_t n, uint64_t base, uint32_t* d_table, void* stream) {
    if ((!d_text || !d_table) && n) return fail(LZ77SSS_E_ARG, "null argument");
    if (base < 257 || base > 0x7FFFFFFEull) return fail(LZ77SSS_E_ARG, "base must lie in [257, 2^31 - 2] (rabin_karp_substring.hpp:92-96)");
    if (n > RKS_NOALLOC_MAX_BYTES)
        return fail(LZ77SSS_E_UNSUPPORTED, "lz77sss_rks_table_noalloc: n exceeds LZ77SSS_RKS_NOALLOC_MAX_BYTES (2^30): the scratch is a "
                                            "static device array; use lz77sss_rks_create for longer texts");
    if ((reinterpret_cast<uintptr_t>(d_text) & 15) || (reinterpret_cast<uintptr_t>(d_table) & 15))
        return fail(LZ77SSS_E_ARG, "lz77sss_rks_table_noalloc: d_text and d_table must be 16-byte aligned");
    return guarded([&]() {
        rks_build_noalloc(d_text, n, (uint32_t)base, d_table, static_cast<cudaStream_t>(stream));   // (no allocation, no synchronisation)
        return LZ77SSS_OK;
    });
}

void lz77sss_rks_destroy(lz77sss_rks* r) {
    if (!r) return;
    cudaSetDevice(r->h->cfg.device);
    delete r;
}

int lz77sss_rks_table(lz77sss_rks* r, uint32_t* out, uint64_t cap, uint64_t* count) {
    if (!r) return fail(LZ77SSS_E_ARG, "null argument");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(r->h->cfg.device));
        return copy_out<uint32_t, uint32_t>(r->st.pf.get(), r->st.ngran, out, cap, count, r->h->st);
    });
}

int lz77sss_rks_fps(lz77sss_rks* r, uint64_t s, uint32_t* out, uint64_t cap, uint64_t* count) {
    if (!r || s == 0) return fail(LZ77SSS_E_ARG, "bad argument");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(r->h->cfg.device));
        const uint64_t cnt = (r->st.n + s - 1) / s + 1;          // div_ceil(n, s) + 1 entries (:97,131,154)
        if (count) *count = cnt;
        if (!out) return LZ77SSS_OK;
        const uint64_t m = std::min(cnt, cap);
        dbuf<uint32_t> d;
        d.reserve(m + 1);
        rks_fps(r->d_text, r->st, s, m, d, r->h->st);
        CUDA_CHECK(cudaMemcpyAsync(out, d.get(), m * 4, cudaMemcpyDeviceToHost, r->h->st));
        CUDA_CHECK(cudaStreamSynchronize(r->h->st));
        return LZ77SSS_OK;
    });
}

int lz77sss_rks_substring(lz77sss_rks* r, const uint64_t* pos, const uint64_t* len, uint64_t nq, uint32_t* out) {
    if (!r || ((!pos || !len || !out) && nq)) return fail(LZ77SSS_E_ARG, "null argument");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(r->h->cfg.device));
        if (!nq) return LZ77SSS_OK;
        for (uint64_t k = 0; k < nq; k++)
            if (pos[k] + len[k] > r->st.n || pos[k] + len[k] < pos[k]) return fail(LZ77SSS_E_ARG, "substring out of range");
        dbuf<uint64_t> dq;
        dbuf<uint32_t> dout;
        dq.reserve(2 * nq);
        dout.reserve(nq);
        CUDA_CHECK(cudaMemcpyAsync(dq.get(), pos, nq * 8, cudaMemcpyHostToDevice, r->h->st));
        CUDA_CHECK(cudaMemcpyAsync(dq.get() + nq, len, nq * 8, cudaMemcpyHostToDevice, r->h->st));
        rks_substring_batch(r->d_text, r->st, dq.get(), dq.get() + nq, nq, dout, r->h->st);
        CUDA_CHECK(cudaMemcpyAsync(out, dout.get(), nq * 4, cudaMemcpyDeviceToHost, r->h->st));
        CUDA_CHECK(cudaStreamSynchronize(r->h->st));
        return LZ77SSS_OK;
    });
}

int lz77sss_rks_time_build(lz77sss_rks* r, int iters, double* ms_out, uint64_t* launches_out) {
    if (!r || iters <= 0) return fail(LZ77SSS_E_ARG, "bad argument");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(r->h->cfg.device));
        cudaEvent_t e0, e1;
        CUDA_CHECK(cudaEventCreate(&e0));
        CUDA_CHECK(cudaEventCreate(&e1));
        double total = 0;
        for (int it = -1; it < iters; it++) {
            CUDA_CHECK(cudaEventRecord(e0, r->h->st));
            rks_build(r->d_text, r->st.n, r->st.base, r->st, r->h->st);
            CUDA_CHECK(cudaEventRecord(e1, r->h->st));
            CUDA_CHECK(cudaEventSynchronize(e1));
            float ms = 0;
            CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
            if (it >= 0) total += ms;
        }
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        if (ms_out) *ms_out = total / iters;
        if (launches_out) *launches_out = r->launches;
        return LZ77SSS_OK;
    });
}

namespace {
// Device -> host memory of the caller.  Pinned memory takes one copy at link speed.  Pageable memory would be staged by the
// driver at a fraction of it: chunks go through two pinned buffers instead, and several host threads copy chunk k - 1 out
// (first touch of the caller's pages included) while chunk k is on the link.
void copy_to_host(lz77sss_handle* h, uint8_t* out, const uint8_t* d, uint64_t n) {
    cudaPointerAttributes attr;
    const bool pinned = cudaPointerGetAttributes(&attr, out) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    cudaGetLastError();                             // (an unregistered pointer may leave an error behind on old drivers)
    constexpr uint64_t CH = 64ull << 20;
    if (pinned || n <= CH / 4) {
        CUDA_CHECK(cudaMemcpyAsync(out, d, n, cudaMemcpyDeviceToHost, h->st));
        CUDA_CHECK(cudaStreamSynchronize(h->st));
        return;
    }
    cudaEvent_t ev[2];
    for (auto& e : ev) CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    struct drop { cudaEvent_t* e; ~drop() { cudaEventDestroy(e[0]); cudaEventDestroy(e[1]); } } guard{ev};
    for (auto& b : h->d2h_stage) b.resize(CH);
    const uint64_t chunks = (n + CH - 1) / CH;
    auto drain = [&](uint64_t k) {                  // chunk k has arrived in its buffer: out to the caller's memory
        CUDA_CHECK(cudaEventSynchronize(ev[k & 1]));
        const uint64_t off = k * CH, len = std::min(CH, n - off);
        const uint8_t* from = h->d2h_stage[k & 1].data();
        const uint32_t parts = 8;
        run_blocks(parts, [&](uint32_t t) {
            const uint64_t b = len * t / parts, e = len * (t + 1) / parts;
            std::memcpy(out + off + b, from + b, e - b);
        });
    };
    for (uint64_t k = 0; k < chunks; k++) {
        const uint64_t off = k * CH, len = std::min(CH, n - off);
        CUDA_CHECK(cudaMemcpyAsync(h->d2h_stage[k & 1].data(), d + off, len, cudaMemcpyDeviceToHost, h->st));
        CUDA_CHECK(cudaEventRecord(ev[k & 1], h->st));
        if (k > 0) drain(k - 1);
    }
    drain(chunks - 1);
}
}  // namespace

int lz77sss_decode(lz77sss_handle* h, const uint64_t* src, const uint64_t* len, uint64_t z, uint8_t* out, uint64_t n) {
    if (!h || !src || !len || (!out && n)) return fail(LZ77SSS_E_ARG, "null argument");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        if (n == 0) return LZ77SSS_OK;
        if (z == 0) return fail(LZ77SSS_E_ARG, "no factors");
        h->dec_src.reserve(z); h->dec_len.reserve(z); h->dec_out.reserve(n);
        CUDA_CHECK(cudaMemcpyAsync(h->dec_src.get(), src, z * 8, cudaMemcpyHostToDevice, h->st));
        CUDA_CHECK(cudaMemcpyAsync(h->dec_len.get(), len, z * 8, cudaMemcpyHostToDevice, h->st));
        int rounds = 0;
        int rc = decode_factors(h->dec_src, h->dec_len, z, h->dec_out, n, h->st, &rounds, h->dec);
        if (rc == -4) return fail(LZ77SSS_E_INTERNAL, "decode: the copy chains did not resolve");
        if (rc) return fail(LZ77SSS_E_ARG, "decode: malformed factor stream (lengths do not add up to n or a source is not earlier text)");
        copy_to_host(h, out, h->dec_out.get(), n);
        return LZ77SSS_OK;
    });
}

int lz77sss_time_decode(lz77sss_handle* h, const uint64_t* src, const uint64_t* len, uint64_t z, uint64_t n, int iters,
                        double* ms_out, int* rounds_out, uint64_t* launches_out, const uint8_t* d_expect, int* equal_out) {
    if (!h || !src || !len || iters <= 0 || !n || !z) return fail(LZ77SSS_E_ARG, "bad argument");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        h->dec_src.reserve(z); h->dec_len.reserve(z); h->dec_out.reserve(n);
        CUDA_CHECK(cudaMemcpyAsync(h->dec_src.get(), src, z * 8, cudaMemcpyHostToDevice, h->st));
        CUDA_CHECK(cudaMemcpyAsync(h->dec_len.get(), len, z * 8, cudaMemcpyHostToDevice, h->st));
        cudaEvent_t e0, e1;
        CUDA_CHECK(cudaEventCreate(&e0));
        CUDA_CHECK(cudaEventCreate(&e1));
        double total = 0;
        int rounds = 0;
        unsigned long long l0 = 0;
        for (int it = -1; it < iters; it++) {     // one untimed warm-up
            l0 = g_kernel_launches;
            CUDA_CHECK(cudaEventRecord(e0, h->st));
            const int rc = decode_factors(h->dec_src, h->dec_len, z, h->dec_out, n, h->st, &rounds, h->dec);
            CUDA_CHECK(cudaEventRecord(e1, h->st));
            CUDA_CHECK(cudaEventSynchronize(e1));
            if (rc) return fail(rc == -4 ? LZ77SSS_E_INTERNAL : LZ77SSS_E_ARG, "decode failed");
            float ms = 0;
            CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
            if (it >= 0) total += ms;
        }
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        if (ms_out) *ms_out = total / iters;
        if (rounds_out) *rounds_out = rounds;
        if (launches_out) *launches_out = g_kernel_launches - l0;
        if (d_expect && equal_out) *equal_out = device_equal(h->dec_out.get(), d_expect, n, h->st) ? 1 : 0;
        return LZ77SSS_OK;
    });
}

int lz77sss_build_sss(lz77sss_handle* h, const uint8_t* text, uint64_t n, uint64_t* S_out, uint64_t cap, uint64_t* count_out,
                      uint64_t* num_q_out) {
    if (!h || (!text && n)) return fail(LZ77SSS_E_ARG, "null argument");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        uint8_t* d = alloc_text(h->text_buf, n, h->st);
        if (n) CUDA_CHECK(cudaMemcpyAsync(d, text, n, cudaMemcpyHostToDevice, h->st));
        auto& P = h->p64;
        uint64_t s_cap = n + 16;
        P.S.reserve(s_cap);
        uint64_t size_s = 0, numq = 0;
        sss_build(d, n, h->cfg.tau, h->cfg.sss_base, h->cfg.sss_qmax, P.S.get(), 8, s_cap, P.sss, h->st, &size_s, &numq);
        if (count_out) *count_out = size_s;
        if (num_q_out) *num_q_out = numq;
        return copy_out<uint64_t, uint64_t>(P.S.get(), size_s, S_out, cap, nullptr, h->st);
    });
}

int lz77sss_time_sss_kernel(lz77sss_handle* h, const uint8_t* d_text, uint64_t n, int iters, double* ms_out,
                            uint64_t* count_out) {
    if (!h || !d_text || iters <= 0) return fail(LZ77SSS_E_ARG, "bad argument");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        const bool narrow = n < 0xFFFF0000ull;
        uint64_t s_cap = std::min<uint64_t>(n, 8 * (n / h->cfg.tau) + 65536);
        if (h->cfg.tau < 16) s_cap = n;
        uint64_t size_s = 0, numq = 0;
        cudaEvent_t e0, e1;
        CUDA_CHECK(cudaEventCreate(&e0));
        CUDA_CHECK(cudaEventCreate(&e1));
        double total = 0;
        for (int it = -1; it < iters; it++) {     // one untimed warm-up
            CUDA_CHECK(cudaEventRecord(e0, h->st));
            if (narrow) {
                h->p32.S.reserve(s_cap);
                sss_build(d_text, n, h->cfg.tau, h->cfg.sss_base, h->cfg.sss_qmax, h->p32.S.get(), 4, s_cap, h->p32.sss, h->st, &size_s, &numq);
            } else {
                h->p64.S.reserve(s_cap);
                sss_build(d_text, n, h->cfg.tau, h->cfg.sss_base, h->cfg.sss_qmax, h->p64.S.get(), 8, s_cap, h->p64.sss, h->st, &size_s, &numq);
            }
            CUDA_CHECK(cudaEventRecord(e1, h->st));
            CUDA_CHECK(cudaEventSynchronize(e1));
            float ms = 0;
            CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
            if (it >= 0) total += ms;
        }
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        if (ms_out) *ms_out = total / iters;
        if (count_out) *count_out = size_s;
        return LZ77SSS_OK;
    });
}

int lz77sss_kernel_profile(lz77sss_handle* h, int enable) {
    if (!h) return fail(LZ77SSS_E_ARG, "null argument");
    g_prof_on = enable != 0;
    if (enable) g_prof.clear();
    return LZ77SSS_OK;
}

int lz77sss_kernel_profile_get(lz77sss_handle* h, uint32_t idx, char* name, uint32_t name_cap, double* ms, uint64_t* launches,
                               uint64_t* bytes) {
    if (!h || !name || !name_cap) return fail(LZ77SSS_E_ARG, "null argument");
    if (idx >= g_prof.size()) return fail(LZ77SSS_E_ARG, "no such profile row");
    const auto& r = g_prof[idx];
    snprintf(name, name_cap, "%s", r.name.c_str());
    if (ms) *ms = r.ms;
    if (launches) *launches = r.launches;
    if (bytes) *bytes = r.bytes;
    return LZ77SSS_OK;
}

int lz77sss_get_stats(lz77sss_handle* h, lz77sss_stats* out) {
    if (!h || !out) return fail(LZ77SSS_E_ARG, "null argument");
    *out = h->stats;
    return LZ77SSS_OK;
}

#define NEED_RESULT() \
    if (!h) return fail(LZ77SSS_E_ARG, "null handle"); \
    if (!h->have_result) return fail(LZ77SSS_E_STATE, "no factorization on this handle yet")

int lz77sss_get_sss(lz77sss_handle* h, uint64_t* out, uint64_t cap, uint64_t* count) {
    NEED_RESULT();
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        if (h->pos_bytes == 4) return copy_out<uint32_t, uint64_t>(h->p32.S.get(), h->p32.s, out, cap, count, h->st);
        return copy_out<uint64_t, uint64_t>(h->p64.S.get(), h->p64.s, out, cap, count, h->st);
    });
}

int lz77sss_get_sa_s(lz77sss_handle* h, uint32_t* out, uint64_t cap, uint64_t* count) {
    NEED_RESULT();
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        if (h->pos_bytes == 4) return copy_out<uint32_t, uint32_t>(h->p32.suf.SA.get(), h->p32.s, out, cap, count, h->st);
        return copy_out<uint32_t, uint32_t>(h->p64.suf.SA.get(), h->p64.s, out, cap, count, h->st);
    });
}

int lz77sss_get_lcp_s(lz77sss_handle* h, uint64_t* out, uint64_t cap, uint64_t* count) {
    NEED_RESULT();
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        if (h->pos_bytes == 4) return copy_out<uint32_t, uint64_t>(h->p32.suf.LCP.get(), h->p32.s, out, cap, count, h->st);
        return copy_out<uint64_t, uint64_t>(h->p64.suf.LCP.get(), h->p64.s, out, cap, count, h->st);
    });
}

int lz77sss_get_lpf(lz77sss_handle* h, uint64_t* beg, uint64_t* end, uint64_t* src, uint32_t* blk_out, uint64_t cap,
                    uint64_t* count) {
    NEED_RESULT();
    uint64_t k = 0;
    for (uint32_t b = 0; b < h->LPF.size(); b++)
        for (auto& ph : h->LPF[b]) {
            if (k < cap) {
                if (beg) beg[k] = ph.beg;
                if (end) end[k] = ph.end;
                if (src) src[k] = ph.src;
                if (blk_out) blk_out[k] = b;
            }
            k++;
        }
    if (count) *count = k;
    return LZ77SSS_OK;
}

int lz77sss_lce_batch(lz77sss_handle* h, const uint64_t* qi, const uint64_t* qj, uint64_t nq, uint64_t* out) {
    NEED_RESULT();
    if (!qi || !qj || !out) return fail(LZ77SSS_E_ARG, "null argument");
    return guarded([&]() {
        CUDA_CHECK(cudaSetDevice(h->cfg.device));
        if (nq == 0) return LZ77SSS_OK;
        for (uint64_t k = 0; k < nq; k++)
            if (qi[k] >= h->n || qj[k] >= h->n) return fail(LZ77SSS_E_ARG, "lce query out of range");
        dbuf<uint64_t> d_i, d_j, d_o;
        d_i.reserve(nq); d_j.reserve(nq); d_o.reserve(nq);
        CUDA_CHECK(cudaMemcpyAsync(d_i.get(), qi, nq * 8, cudaMemcpyHostToDevice, h->st));
        CUDA_CHECK(cudaMemcpyAsync(d_j.get(), qj, nq * 8, cudaMemcpyHostToDevice, h->st));
        if (h->pos_bytes == 4)
            lce_r_batch<uint32_t>(h->d_text, h->n, h->cfg.tau, h->p32.S.get(), h->p32.s, h->p32.suf, h->p32.lce, d_i, d_j, nq, d_o, h->st);
        else
            lce_r_batch<uint64_t>(h->d_text, h->n, h->cfg.tau, h->p64.S.get(), h->p64.s, h->p64.suf, h->p64.lce, d_i, d_j, nq, d_o, h->st);
        CUDA_CHECK(cudaMemcpyAsync(out, d_o.get(), nq * 8, cudaMemcpyDeviceToHost, h->st));
        CUDA_CHECK(cudaStreamSynchronize(h->st));
        return LZ77SSS_OK;
    });
}

#pragma GCC visibility pop
}  // extern "C"
