// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/core/sss_core.hpp header).
//
// Standalone, sequential CPU restatement of the lz77-sss 3-approximation hot path
//   lz77_sss<pos_t>::factorize_approximate<greedy, lpf_opt, tau>  and  lz77_sss<pos_t>::decode
// following the reference function by function (citations are relative to /root/reference/).
// It needs nothing from /root/reference at build or run time, so it travels to the GPU box.
// It is pinned against the reference's own code (oracle/_ref/libref.so = reference headers +
// oracle/shim) by tests/test_oracle_vs_ref.py and by the golden fixtures in tests/golden/.
// The SSS / suffix-order part is shared with the shim (oracle/core) and is PARITY UNPINNED
// against the missing upstream lce submodule.
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

#include "../core/sss_core.hpp"

namespace {
__extension__ typedef unsigned __int128 u128;

struct lpf_t { uint64_t beg, end, src; };           // lz77_sss.hpp:206-210
struct factor_t { uint64_t src, len; };             // lz77_sss.hpp:129-131

// ---- rk_prime<107> (patched-files/external/lce/include/rolling_hash/rolling_hash.hpp:24-156) ----
struct rk107 {
    static constexpr u128 P = (u128(1) << 107) - 1;
    u128 base = 0, fp = 0, neg_btau = 0;            // neg_btau = -(base^tau) mod P
    static u128 mod(u128 x) {
        x = (x & P) + (x >> 107);
        x = (x & P) + (x >> 107);
        return x >= P ? x - P : x;
    }
    static u128 mulmod(u128 a, u128 b) {             // general product (setup only)
        u128 r = 0;
        a = mod(a);
        while (b) {
            if (b & 1) r = mod(r + a);
            a = mod(a + a);
            b >>= 1;
        }
        return r;
    }
    void init(uint64_t tau, uint64_t b) {            // rolling_hash.hpp:30-47,133-155
        base = b;
        fp = 0;
        u128 pw = 1, x = base;
        uint64_t e = tau;
        while (e) {
            if (e & 1) pw = mulmod(pw, x);
            x = mulmod(x, x);
            e >>= 1;
        }
        neg_btau = pw == 0 ? 0 : P - pw;
    }
    // m_char_influence[out][in] = in - out*base^tau (rolling_hash.hpp:139-154)
    u128 influence(uint8_t out, uint8_t in) const { return mod(neg_btau * (u128)out + in); }  // < 2^115, no overflow
    void roll(uint8_t out, uint8_t in) {             // rolling_hash.hpp:82-86
        fp *= base;
        fp = mod(fp + influence(out, in));
    }
    void reset() { fp = 0; }
};

// ---- rolling_hash_index_107 (include/lz77_sss/data_structures/rolling_hash_index_107.hpp:33-173) --
struct gap_index {
    static constexpr uint64_t min_rh_index_size = 1 << 20;      // lz77_sss.hpp:91
    static constexpr uint64_t max_rh_index_size = 1 << 30;      // lz77_sss.hpp:92
    static constexpr double min_rel_rh_index_size = 0.1;        // lz77_sss.hpp:93
    static constexpr int64_t sizeof_rk_prime = 80;              // x86-64 libstdc++; checked in tests
    const uint8_t* input = nullptr;
    uint64_t input_size = 0;
    std::array<uint64_t, 5> patt_lens{};
    rk107 rh[5];
    std::vector<uint64_t> H;
    uint64_t mask = 0, none = 0;
    uint64_t cur_pos = 0;

    static uint64_t table_entries(uint64_t n, int64_t target_size_in_bytes, unsigned sizeof_pos) {
        // rolling_hash_index_107.hpp:59-67
        int64_t rolling_hash_size = (sizeof_rk_prime + 16 * 256 * 256) * 5;
        uint64_t a = min_rh_index_size, b;
        if (sizeof_pos == 4) b = (uint32_t)(n * min_rel_rh_index_size);
        else b = (uint64_t)(n * min_rel_rh_index_size);
        int64_t min_index_size = std::max<uint64_t>(a, b) / sizeof_pos;
        int64_t max_index_size = max_rh_index_size / sizeof_pos;
        int64_t target_index_size = std::max<int64_t>(0, target_size_in_bytes - rolling_hash_size) / sizeof_pos;
        uint64_t target_size_h = std::min<int64_t>(max_index_size, std::max<int64_t>(min_index_size, target_index_size));
        uint8_t log2_size_h = std::round(std::log2(target_size_h));
        return uint64_t(1) << log2_size_h;
    }
    void build(const uint8_t* in, uint64_t n, std::array<uint64_t, 5> pl, int64_t target, unsigned sizeof_pos,
               const uint64_t bases[5]) {
        input = in;
        input_size = n;
        patt_lens = pl;
        uint64_t size_h = table_entries(n, target, sizeof_pos);
        mask = size_h - 1;
        none = sizeof_pos == 4 ? 0xFFFFFFFFull : ~0ull;
        H.assign(size_h, none);
        for (int i = 0; i < 5; i++) rh[i].init(patt_lens[i], bases[i]);
        reinit(0);
    }
    void reinit(uint64_t pos) {                      // :80-99
        cur_pos = pos;
        for (int i = 0; i < 5; i++) {
            rh[i].reset();
            if (cur_pos + patt_lens[i] < input_size)
                for (uint64_t j = 0; j < patt_lens[i]; j++) rh[i].roll(0, input[cur_pos + j]);
        }
    }
    void roll_i(int i) { rh[i].roll(input[cur_pos], input[cur_pos + patt_lens[i]]); }   // :101-107
    void roll() {                                    // :109-118
        for (int i = 0; i < 5; i++)
            if (cur_pos + patt_lens[i] < input_size) roll_i(i);
        cur_pos++;
    }
    void advance_i(int i) {                          // :120-127
        if (cur_pos + patt_lens[i] < input_size) {
            H[(uint64_t)(rh[i].fp & mask)] = cur_pos;
            roll_i(i);
        }
    }
    void advance() {                                 // :129-136
        for (int i = 0; i < 5; i++) advance_i(i);
        cur_pos++;
    }
    uint64_t advance_and_get_occ(int i) {            // :138-150
        uint64_t h = (uint64_t)(rh[i].fp & mask);
        uint64_t occ = H[h];
        H[h] = cur_pos;
        if (cur_pos + patt_lens[i] < input_size) roll_i(i);
        return occ;
    }
};

struct restate {
    const uint8_t* T;
    uint64_t n, tau;
    uint32_t sss_base, qmax;
    unsigned sizeof_pos;
    uint16_t p;
    uint64_t bases[5];

    std::vector<uint64_t> S;
    bool has_runs = false;
    std::vector<uint32_t> SA, ISA;
    std::vector<uint64_t> LCP;
    oracle::block_rmq<uint64_t> rmq;
    oracle::succ_index succ;
    std::vector<uint32_t> PSV, NSV;
    std::vector<std::vector<lpf_t>> LPF;
    uint64_t num_gaps = 0, num_lpf = 0, len_lpf_phr = 0, len_gaps = 0, num_fact = 0, roll_threshold = 0;
    std::array<uint64_t, 5> patt_lens{};
    uint64_t target_index_size = 0;
    gap_index gap_idx;

    // ---- build_lce (algorithms/common.cpp:58-80) -> lce_sss ctor (lce_sss.hpp:44-96) ----
    void build_lce() {
        oracle::sss_result r = oracle::build_sss(T, n, tau, sss_base, qmax);         // lce_sss.hpp:53
        S.swap(r.S);
        has_runs = r.has_runs;
        unsigned shift = 0;
        while ((uint64_t(2) << shift) <= tau) shift++;                               // bit_width(tau)-1
        succ.build(S, shift);                                                         // lce_sss.hpp:68-69
        SA = oracle::sort_sync_suffixes(T, n, tau, qmax, S);                          // lce_sss.hpp:82-83 + lce_classic_for_sss.hpp:52
        uint64_t s = S.size();
        ISA.resize(s);
        for (uint64_t i = 0; i < s; i++) ISA[SA[i]] = (uint32_t)i;                    // lce_classic_for_sss.hpp:65-69
        LCP.assign(s, 0);                                                             // :81-120 (single slice)
        uint64_t cur = 0;
        for (uint64_t i = 0; i < s; i++) {
            uint64_t r_ = ISA[i];
            if (r_ == 0) continue;
            uint64_t prev = SA[r_ - 1];
            cur += oracle::naive_lce(T, n, S[i] + cur, S[prev] + cur);
            LCP[r_] = cur;
            if (i == s - 1) break;
            uint64_t diff = S[i + 1] - S[i];
            if (cur < 2 * tau + diff) cur = 0; else cur -= diff;
        }
        rmq.build(LCP.data(), s);                                                     // :133
    }

    // lce_sss::lce / lce_lr (lce_sss.hpp:102-177)
    uint64_t LCE_R(uint64_t i, uint64_t j) const {
        if (i == j) return n - i;
        uint64_t l = std::min(i, j), r = std::max(i, j);
        uint64_t lce_max = n - r;
        uint64_t lce_local_max = std::min<uint64_t>(3 * tau, lce_max);
        uint64_t lce_local = oracle::naive_lce_lr(T, r + lce_local_max, l, r);
        if (lce_local < lce_local_max || lce_local == lce_max) return lce_local;
        uint64_t l_ = succ.successor(l), r_ = succ.successor(r);
        if (S[l_] - l != S[r_] - r) return std::min(S[l_] - l, S[r_] - r) + 2 * tau - 1;
        uint64_t a = ISA[l_], b = ISA[r_];
        return (S[l_] - l) + LCP[rmq.argmin(std::min(a, b) + 1, std::max(a, b))];    // lce_classic_for_sss.hpp:153-155
    }

    // lce_l_64 (algorithms/lce_l.hpp:33-83): equal characters leftwards from T[i], T[j], capped
    uint64_t LCE_L(uint64_t i, uint64_t j, uint64_t lce_max) const {
        lce_max = std::min<uint64_t>(lce_max, std::min(i, j) + 1);
        if (i == j) return lce_max;
        uint64_t k = 0;
        while (k < lce_max && T[i - k] == T[j - k]) k++;
        return k;
    }

    // build_PSV_NSV_S (approximate/lpf_lnf/nxv_pxv.cpp:33-92)
    void build_PSV_NSV_S() {
        uint32_t s = (uint32_t)S.size();
        PSV.assign(s, s);
        NSV.assign(s, s);
        if (!s) return;
        for (uint32_t i = 1; i < s; i++) {
            uint32_t j = i - 1;
            while (j != s && SA[j] > SA[i]) {
                NSV[j] = i;
                j = PSV[j];
            }
            PSV[i] = j;
            if (j != s) NSV[j] = s;
        }
    }

    // build_LPF_opt (approximate/lpf_lnf/lpf_opt.cpp:33-157)
    void build_LPF_opt() {
        build_PSV_NSV_S();
        uint64_t s = S.size();
        LPF.assign(p, {});
        for (uint16_t i_p = 0; i_p < p; i_p++) {
            uint64_t b = i_p * (n / p);
            uint64_t e = i_p == p - 1 ? n : ((i_p + 1) * (n / p));
            uint64_t i_min = std::lower_bound(S.begin(), S.end(), b) - S.begin();
            uint64_t i_max = std::lower_bound(S.begin(), S.end(), e) - S.begin();
            uint64_t max_end = b;
            for (uint64_t i = i_min; i < i_max; i++) {
                while (i + 1 < i_max && S[i + 1] <= max_end) i++;
                uint64_t lst_end = max_end;
                lpf_t phr { 0, 0, 0 };
                auto candidate = [&](uint32_t xv, bool is_psv) {
                    uint64_t src = S[SA[xv]];
                    uint64_t end = S[i] + LCE_R(src, S[i]);
                    if (end > lst_end) {
                        uint64_t beg = S[i];
                        if (S[i] > lst_end && src != 0 && S[i] != 0) {
                            uint64_t l = LCE_L(src - 1, S[i] - 1, S[i] - lst_end);
                            beg -= l;
                            src -= l;
                        }
                        if (beg < lst_end) {
                            uint64_t exc = lst_end - beg;
                            beg += exc;
                            src += exc;
                        }
                        if (end > max_end) max_end = end;
                        if (is_psv) {
                            if (end - beg > 1) phr = { beg, end, src };                // :99-101
                        } else {
                            if (end - beg > phr.end - phr.beg) phr = { beg, end, src }; // :136-138
                        }
                    }
                };
                if (PSV[ISA[i]] != s) candidate(PSV[ISA[i]], true);
                if (NSV[ISA[i]] != s) {
                    candidate(NSV[ISA[i]], false);
                    if (phr.end - phr.beg > 1) LPF[i_p].push_back(phr);              // :141-143
                }
            }
        }
    }

    // get_phrase_info (approximate/common.cpp:100-157)
    void get_phrase_info() {
        for (uint16_t i_p = 0; i_p < p; i_p++) {
            uint64_t num_gaps_thr = 0, len_lpf_phr_thr = 0, num_lpf_thr = 0;
            uint64_t b = i_p * (n / p);
            uint64_t e = i_p == p - 1 ? n : ((i_p + 1) * (n / p));
            uint64_t i = 0;
            auto& L = LPF[i_p];
            if (!L.empty()) e = std::max(e, L.back().end);
            for (uint16_t j = 0; j < i_p; j++)
                if (!LPF[j].empty()) b = std::max(b, LPF[j].back().end);
            if (!L.empty())
                while (i < L.size() && L[i].end <= b) i++;
            num_lpf_thr = L.size() - i;
            if (num_lpf_thr > 0) {
                len_lpf_phr_thr += L[i].end - std::max(L[i].beg, b);
                if (L[i].beg > b) num_gaps_thr = 1;
                i++;
                while (i < L.size()) {
                    len_lpf_phr_thr += L[i].end - L[i].beg;
                    if (L[i].beg > L[i - 1].end) num_gaps_thr++;
                    i++;
                }
                if (L.back().end < e) num_gaps_thr++;
            } else {
                num_gaps_thr = 1;
            }
            num_gaps += num_gaps_thr;
            num_lpf += num_lpf_thr;
            len_lpf_phr += len_lpf_phr_thr;
        }
    }

    // compute_approximation after the LPF stage (lz77_sss.hpp:415-476)
    void choose_parameters() {
        using entry_t = std::pair<double, std::array<uint64_t, 5>>;
        static const double infty = std::numeric_limits<double>::max();
        static const std::array<entry_t, 10> patt_len_table {{                        // lz77_sss.hpp:99-110
            { 6, { 2, 3, 4, 5, 6 } }, { 8, { 2, 3, 4, 6, 8 } }, { 12, { 2, 3, 4, 8, 12 } },
            { 16, { 2, 4, 6, 9, 16 } }, { 32, { 2, 4, 6, 10, 20 } }, { 64, { 2, 4, 7, 12, 28 } },
            { 128, { 2, 4, 8, 16, 36 } }, { 256, { 2, 5, 10, 20, 42 } }, { 1024, { 2, 6, 12, 24, 48 } },
            { infty, { 2, 8, 16, 32, 64 } } }};
        get_phrase_info();
        LPF[p - 1].push_back(lpf_t { n, n + 1, 0 });
        // pos_t arithmetic: len_gaps is a pos_t (lz77_sss.hpp:242,425)
        len_gaps = n - len_lpf_phr;
        double rel_len_gaps = len_gaps / (double)n;
        double avg_gap_len = len_gaps / (double)num_gaps;
        double avg_lpf_phr_len = len_lpf_phr / (double)num_lpf;
        // get_target_gap_idx_size, malloc_count stubbed to 0 (lz77_sss.hpp:117-122)
        target_index_size = std::min<uint64_t>(uint64_t(1) << 30, std::max<uint64_t>({ uint64_t(1) << 20,
            uint64_t(0), (uint64_t)((n / 3.0) * rel_len_gaps) }));
        double patt_len_guess = std::min<double>({ avg_gap_len, avg_lpf_phr_len,
            8.0 * std::pow(128, 1.0 - rel_len_gaps) });                                // lz77_sss.hpp:112-115
        for (auto& [threshold, lens] : patt_len_table) {
            if (patt_len_guess <= threshold) { patt_lens = lens; break; }
        }
        roll_threshold = 0;
        for (int j = 0; j < 5; j++) roll_threshold += patt_lens[j];
        roll_threshold /= 5;
    }

    // longest_prev_occ (approximate/factorize/common.cpp:33-61)
    factor_t longest_prev_occ(uint64_t pos) {
        factor_t f { T[pos], 0 };
        for (int x = 4; x >= 0; x--) {
            if (f.len == 0) {
                uint64_t src = gap_idx.advance_and_get_occ(x);
                if (src < pos && T[src] == T[pos]) {
                    f.len = LCE_R(src, pos);
                    f.src = src;
                }
            } else {
                gap_idx.advance_i(x);
            }
        }
        gap_idx.cur_pos++;
        return f;
    }

    // factorize + next_lpf (approximate/factorize/common.cpp:66-111), factorize_greedy (greedy.cpp:34-140)
    template <typename out_t>
    void factorize_greedy(out_t output) {
        uint16_t i_p = 0;
        uint32_t li = 0;
        while (LPF[i_p].empty()) i_p++;
        auto next_lpf = [&]() {
            lpf_t phr = LPF[i_p][li++];
            if (li == LPF[i_p].size()) {
                if (i_p < p - 1 && p > 1) {
                    do {
                        i_p++;
                        li = 0;
                        while (li < LPF[i_p].size() && LPF[i_p][li].end <= phr.end) li++;
                        if (li < LPF[i_p].size() && LPF[i_p][li].beg < phr.end) {
                            uint64_t offs = phr.end - LPF[i_p][li].beg;
                            LPF[i_p][li].beg += offs;
                            LPF[i_p][li].src += offs;
                        }
                    } while (li == LPF[i_p].size() && i_p < p - 1);
                } else {
                    li--;
                }
            }
            return phr;
        };
        lpf_t ph = next_lpf();
        for (uint64_t i = 0; true;) {
            uint64_t gap_end = ph.beg;
            if (i < gap_end) {
                if (gap_idx.cur_pos < i) {
                    if (i - gap_idx.cur_pos <= roll_threshold) {
                        do { gap_idx.roll(); } while (gap_idx.cur_pos < i);
                    } else {
                        gap_idx.reinit(i);
                    }
                }
                do {
                    factor_t f = longest_prev_occ(i);
                    i += std::max<uint64_t>(1, f.len);
                    if (i > gap_end) {
                        if (i <= ph.end) {
                            f.len -= i - gap_end;
                            i = gap_end;
                        } else {
                            do { ph = next_lpf(); } while (ph.end <= i);
                            while (gap_idx.cur_pos < gap_end) gap_idx.advance();
                            gap_end = ph.beg;
                        }
                    }
                    output(f);
                    num_fact++;
                    while (gap_idx.cur_pos < i) gap_idx.advance();
                } while (i < gap_end);
            }
            if (i == n) break;
            uint64_t exc = i - gap_end;
            factor_t lp { ph.src + exc, (ph.end - ph.beg) - exc };
            if (gap_idx.cur_pos == i) {
                factor_t f = longest_prev_occ(i);
                if (f.len > lp.len) lp = f;
            }
            output(lp);
            num_fact++;
            i += lp.len;
            while (ph.end <= i) ph = next_lpf();
        }
    }

    template <typename out_t>
    void run(out_t output) {                         // compute_approximation<greedy, lpf_opt> (lz77_sss.hpp:374-491)
        build_lce();
        build_LPF_opt();
        choose_parameters();
        gap_idx.build(T, n, patt_lens, (int64_t)target_index_size, sizeof_pos, bases);
        factorize_greedy(output);
    }
};

std::vector<uint8_t> padded(const uint8_t* text, uint64_t n) {
    std::vector<uint8_t> buf(n + 4 * 4096, 0);
    std::memcpy(buf.data(), text, n);
    return buf;
}
}  // namespace

extern "C" {

int64_t oracle_sss(const uint8_t* text, uint64_t n, uint64_t tau, uint32_t base, uint32_t qmax,
                   uint64_t* S_out, uint64_t cap, uint64_t* num_q) {
    auto buf = padded(text, n);
    oracle::sss_result r = oracle::build_sss(buf.data(), n, tau, base, qmax);
    for (uint64_t i = 0; i < r.S.size() && i < cap; i++) S_out[i] = r.S[i];
    if (num_q) *num_q = r.num_q;
    return (int64_t)r.S.size();
}

// plain definition of Q, O(n*tau*P): q_out[j] = 1 iff per(T[j..j+tau)) <= P
void oracle_q_bruteforce(const uint8_t* T, uint64_t n, uint64_t tau, uint32_t qmax, uint8_t* q_out) {
    uint32_t P = oracle::q_period_bound(tau, qmax);
    for (uint64_t j = 0; j + tau <= n; j++) {
        q_out[j] = 0;
        for (uint32_t p = 1; p <= P && !q_out[j]; p++) {
            bool ok = true;
            for (uint64_t x = 0; x + p < tau; x++)
                if (T[j + x] != T[j + x + p]) { ok = false; break; }
            if (ok) q_out[j] = 1;
        }
    }
}

void oracle_q_anchored(const uint8_t* text, uint64_t n, uint64_t tau, uint32_t qmax, uint8_t* q_out) {
    auto buf = padded(text, n);
    oracle::flag_vec q;
    oracle::mark_q(buf.data(), n, tau, qmax, q);
    std::copy(q.begin(), q.end(), q_out);
}

void oracle_window_ids(const uint8_t* text, uint64_t n, uint64_t tau, uint32_t base, uint32_t qmax, uint32_t* id_out) {
    auto buf = padded(text, n);
    oracle::id_vec id;
    uint64_t nq;
    oracle::window_ids(buf.data(), n, tau, base, qmax, id, nq);
    std::copy(id.begin(), id.end(), id_out);
}

void oracle_sa_s(const uint8_t* text, uint64_t n, uint64_t tau, uint32_t qmax, const uint64_t* S, uint64_t s,
                 uint32_t* sa_out, int naive) {
    auto buf = padded(text, n);
    std::vector<uint64_t> Sv(S, S + s);
    std::vector<uint32_t> sa = naive ? oracle::sort_sync_suffixes_naive(buf.data(), n, Sv)
                                     : oracle::sort_sync_suffixes(buf.data(), n, tau, qmax, Sv);
    std::copy(sa.begin(), sa.end(), sa_out);
}

// LCP over SA_S in text characters (lce_classic_for_sss.hpp:81-120)
void oracle_lcp_s(const uint8_t* text, uint64_t n, const uint64_t* S, const uint32_t* sa, uint64_t s, uint64_t* lcp_out) {
    for (uint64_t r = 0; r < s; r++)
        lcp_out[r] = r == 0 ? 0 : oracle::naive_lce(text, n, S[sa[r - 1]], S[sa[r]]);
}

// structure-based LCE (lce_sss.hpp:102-177) and naive LCE for q query pairs
void oracle_lce_batch(const uint8_t* text, uint64_t n, uint64_t tau, uint32_t base, uint32_t qmax,
                      const uint64_t* qi, const uint64_t* qj, uint64_t q, uint64_t* out_struct, uint64_t* out_naive) {
    auto buf = padded(text, n);
    restate r { buf.data(), n, tau, base, qmax, 4, 1, {} };
    r.build_lce();
    for (uint64_t k = 0; k < q; k++) {
        out_struct[k] = r.LCE_R(qi[k], qj[k]);
        out_naive[k] = oracle::naive_lce(buf.data(), n, qi[k], qj[k]);
    }
}

// LPF phrases of build_LPF_opt, concatenated over the p_ref blocks (no sentinel); blk_out[k] = block
int64_t oracle_lpf_opt(const uint8_t* text, uint64_t n, uint64_t tau, uint32_t base, uint32_t qmax, int p_ref,
                       uint64_t* beg, uint64_t* end, uint64_t* src, uint32_t* blk_out, uint64_t cap) {
    auto buf = padded(text, n);
    restate r { buf.data(), n, tau, base, qmax, 4, (uint16_t)p_ref, {} };
    r.build_lce();
    r.build_LPF_opt();
    uint64_t k = 0;
    for (int b = 0; b < p_ref; b++)
        for (auto& ph : r.LPF[b]) {
            if (k < cap) { beg[k] = ph.beg; end[k] = ph.end; src[k] = ph.src; blk_out[k] = b; }
            k++;
        }
    return (int64_t)k;
}

struct oracle_stats { uint64_t size_sss, num_lpf, len_lpf_phr, num_gaps, patt_lens[5], roll_threshold, table_entries, has_runs; };

// factorize_approximate<greedy, lpf_opt, tau> with every hash parameter explicit.
// pos_bits 32|64 selects sizeof(pos_t) (only the gap table sizing depends on it).
int64_t oracle_factorize_approximate(const uint8_t* text, uint64_t n, uint64_t tau, uint32_t sss_base, uint32_t qmax,
                                     const uint64_t bases[5], int p_ref, int pos_bits, uint64_t* src, uint64_t* len,
                                     uint64_t cap, oracle_stats* st) {
    if (n == 0) return 0;
    auto buf = padded(text, n);
    restate r { buf.data(), n, tau, sss_base, qmax, (unsigned)(pos_bits / 8), (uint16_t)std::max(1, p_ref), {} };
    std::copy(bases, bases + 5, r.bases);
    uint64_t z = 0;
    r.run([&](factor_t f) {
        if (z < cap) { src[z] = f.src; len[z] = f.len; }
        z++;
    });
    if (st) {
        st->size_sss = r.S.size(); st->num_lpf = r.num_lpf; st->len_lpf_phr = r.len_lpf_phr; st->num_gaps = r.num_gaps;
        for (int i = 0; i < 5; i++) st->patt_lens[i] = r.patt_lens[i];
        st->roll_threshold = r.roll_threshold; st->table_entries = r.gap_idx.H.size(); st->has_runs = r.has_runs;
    }
    return (int64_t)z;
}

// lz77_sss<pos_t>::decode (algorithms/common.cpp:33-54)
// Independent exact LZ77 (row f4 of SURVEY.md §8; the role of include/lz77/kkp2.hpp:68-143 in lz77-sss-bench): suffix array by
// prefix doubling, then for every phrase start the nearest suffixes with a smaller text position on both sides in suffix
// order; the longer of the two common prefixes is the longest previous factor.  Literals as {byte, 0}.  For small inputs.
int64_t oracle_lz77_exact(const uint8_t* T, uint64_t n, uint64_t* src_out, uint64_t* len_out, uint64_t cap) {
    if (n == 0) return 0;
    std::vector<uint32_t> sa(n), rk(n), tmp(n);
    for (uint64_t i = 0; i < n; i++) { sa[i] = (uint32_t)i; rk[i] = T[i]; }
    for (uint64_t h = 1;; h *= 2) {
        auto key = [&](uint32_t i) { return std::make_pair(rk[i], i + h < n ? (int64_t)rk[i + h] : (int64_t)-1); };
        std::sort(sa.begin(), sa.end(), [&](uint32_t a, uint32_t b) { return key(a) < key(b); });
        tmp[sa[0]] = 0;
        for (uint64_t r = 1; r < n; r++) tmp[sa[r]] = tmp[sa[r - 1]] + (key(sa[r - 1]) < key(sa[r]) ? 1 : 0);
        rk.swap(tmp);
        if (rk[sa[n - 1]] == n - 1) break;
    }
    // psv[i] / nsv[i]: text position of the nearest suffix in front of / behind suffix i in suffix order that starts earlier
    const uint32_t NONE = 0xFFFFFFFFu;
    std::vector<uint32_t> psv(n, NONE), nsv(n, NONE), st;
    for (uint64_t r = 0; r < n; r++) {
        while (!st.empty() && st.back() > sa[r]) st.pop_back();
        if (!st.empty()) psv[sa[r]] = st.back();
        st.push_back(sa[r]);
    }
    st.clear();
    for (uint64_t r = n; r-- > 0;) {
        while (!st.empty() && st.back() > sa[r]) st.pop_back();
        if (!st.empty()) nsv[sa[r]] = st.back();
        st.push_back(sa[r]);
    }
    auto lce = [&](uint64_t a, uint64_t b) {
        uint64_t l = 0;
        while (b + l < n && T[a + l] == T[b + l]) l++;
        return l;
    };
    uint64_t z = 0;
    for (uint64_t i = 0; i < n;) {
        uint64_t best = 0, from = 0;
        if (psv[i] != NONE) { best = lce(psv[i], i); from = psv[i]; }
        if (nsv[i] != NONE) { const uint64_t l = lce(nsv[i], i); if (l > best) { best = l; from = nsv[i]; } }
        if (z < cap) {
            src_out[z] = best ? from : T[i];
            len_out[z] = best;
        }
        z++;
        i += best ? best : 1;
    }
    return (int64_t)z;
}

int oracle_decode(const uint64_t* src, const uint64_t* len, uint64_t z, uint8_t* out, uint64_t n) {
    uint64_t pos = 0, k = 0;
    while (pos < n) {
        if (k >= z) return -1;
        uint64_t s = src[k], l = len[k];
        k++;
        if (l == 0) {
            out[pos++] = (uint8_t)s;
        } else {
            if (s >= pos || pos + l > n) return -2;
            for (uint64_t i = 0; i < l; i++) out[pos + i] = out[s + i];
            pos += l;
        }
    }
    return 0;
}
}
