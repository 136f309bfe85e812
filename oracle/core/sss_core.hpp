// ORACLE — TEST INFRASTRUCTURE ONLY.  Never linked, imported or executed by the product
// path (lz77_sss_b200/); only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may use anything under oracle/.
//
// CPU restatement of the parts of the (missing) LukasNalbach/lce submodule
// (pinned @ 0359b1bb8e19574591cd695327a571878bb27af0, absent from /root/reference) that the
// lz77-sss hot path calls:  the tau-synchronizing set, the lexicographic order of the
// suffixes starting at synchronizing positions, naive LCE, successor index and RMQ.
//
// PARITY UNPINNED for this file: the reference holds no golden vectors or source for these
// functions; the contracts they must satisfy are taken from the call sites
//   patched-files/external/lce/include/ds/lce_sss.hpp:53,68-69,82-83,123-177
//   patched-files/external/lce/include/ds/lce_classic_for_sss.hpp:52,65-69,82-120,133,153-155
// and from the published construction (Kempa & Kociumaka, "String synchronizing sets", STOC'19,
// Construction 8.?; cited by the reference README.md:151).  Everything *downstream* of these
// (LPF, gap index, greedy emission) is the reference's own code in oracle/_ref.
//
// ---------------------------------------------------------------------------------------------
// Canonical definition used by the oracle AND the B200 kernels:
//   tau   : power of two, 4 <= tau <= 4096
//   base  : odd 32-bit multiplier
//   P     : min(floor(tau/3), qmax)       (period bound of the "density" set Q)
//   fp(j) = sum_{k<tau} T[j+k] * base^(tau-1-k)  (mod 2^32)          for 0 <= j <= n-tau
//   Q     = { j in [0,n-tau] : per(T[j..j+tau)) <= P }
//   id(j) = 0xFFFFFFFF                 if j in Q
//         = (fp(j) & 0x7FFFFFFE) | 1   otherwise
//   id(n-tau+1) = 0                    (virtual window holding the end-of-text terminator)
//   M(i)  = min{ id(j) : i <= j <= i+tau }
//   S     = { i in [0, n-2tau+1] : M(i) != 0xFFFFFFFF and (id(i) == M(i) or id(i+tau) == M(i)) }
// S is empty when n < 2tau-1.  n-2tau+1 is always the last element otherwise.
// SA_S  = the permutation of [0,|S|) that sorts the text suffixes T[S[k]..n) lexicographically
//         (a proper prefix is smaller).
// ---------------------------------------------------------------------------------------------
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <deque>
#include <memory>
#include <numeric>
#include <utility>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#include <parallel/algorithm>
#endif

namespace oracle {

// vectors of n ids / flags are written completely by parallel loops: skip the serial zero fill of resize()
template <typename T>
struct no_init_alloc : std::allocator<T> {
    template <typename U> struct rebind { using other = no_init_alloc<U>; };
    template <typename U> void construct(U* p) noexcept { ::new (static_cast<void*>(p)) U; }
    template <typename U, typename... A> void construct(U* p, A&&... a) { ::new (static_cast<void*>(p)) U(std::forward<A>(a)...); }
};
using id_vec = std::vector<uint32_t, no_init_alloc<uint32_t>>;
using flag_vec = std::vector<uint8_t, no_init_alloc<uint8_t>>;

static constexpr uint32_t ID_INF = 0xFFFFFFFFu;
static constexpr uint32_t DEFAULT_SSS_BASE = 0x9E3779B1u;  // odd
static constexpr uint32_t DEFAULT_QMAX = 32;

inline uint32_t q_period_bound(uint64_t tau, uint32_t qmax) {
    return (uint32_t)std::min<uint64_t>(tau / 3, qmax);
}

// largest power of two <= tau - 2P  (anchor spacing of the Q detection)
inline uint64_t q_anchor_spacing(uint64_t tau, uint32_t P) {
    uint64_t lim = tau - 2 * (uint64_t)P, a = 1;
    while (a * 2 <= lim) a *= 2;
    return a;
}

inline uint32_t pow_u32(uint32_t b, uint64_t e) {
    uint32_t r = 1;
    while (e) {
        if (e & 1) r *= b;
        b *= b;
        e >>= 1;
    }
    return r;
}

// Naive LCE of T[l..) and T[r..), bounded so that r + lce <= end (l < r).
// Contract of lce_naive_wordwise_xor::lce_lr at lce_sss.hpp:141,152.
inline uint64_t naive_lce_lr(const uint8_t* T, uint64_t end, uint64_t l, uint64_t r) {
    uint64_t maxl = end - r, k = 0;
    while (k + 8 <= maxl) {
        uint64_t a, b;
        std::memcpy(&a, T + l + k, 8);
        std::memcpy(&b, T + r + k, 8);
        if (a != b) return k + (__builtin_ctzll(a ^ b) >> 3);
        k += 8;
    }
    while (k < maxl && T[l + k] == T[r + k]) k++;
    return k;
}

inline uint64_t naive_lce(const uint8_t* T, uint64_t n, uint64_t i, uint64_t j) {
    if (i == j) return n - i;
    return naive_lce_lr(T, n, std::min(i, j), std::max(i, j));
}

// Marks Q (see header). q[j] = 1 iff j in Q.  Evaluated through anchors: every window j has a
// first anchor a = ceil(j/A)*A with T[a..a+2P) inside the window; the window has period <= P
// iff that segment has smallest period p <= P and the maximal p-periodic stretch around the
// segment covers the window.  tests/test_oracle_core.py checks this against the plain definition.
inline uint64_t mark_q(const uint8_t* T, uint64_t n, uint64_t tau, uint32_t qmax,
                       flag_vec& q) {
    uint64_t nw = n >= tau ? n - tau + 1 : 0;  // windows 0..nw-1
    q.resize(nw);
    if (nw == 0) return 0;
#pragma omp parallel for schedule(static)
    for (int64_t j = 0; j < (int64_t)nw; j++) q[j] = 0;
    const uint32_t P = q_period_bound(tau, qmax);
    const uint64_t A = q_anchor_spacing(tau, P);
    uint64_t marked = 0;
    int64_t n_anchor = (int64_t)((nw - 1 + A - 1) / A) + 1;
#pragma omp parallel for schedule(static) reduction(+ : marked)
    for (int64_t ai = 0; ai < n_anchor; ai++) {
        uint64_t a = (uint64_t)ai * A;
        if (a + 2 * (uint64_t)P > n) continue;
        uint32_t p = 0;
        for (uint32_t c = 1; c <= P && !p; c++) {
            bool ok = true;
            for (uint64_t x = 0; x + c < 2 * (uint64_t)P; x++)
                if (T[a + x] != T[a + x + c]) { ok = false; break; }
            if (ok) p = c;
        }
        if (!p) continue;
        // right end of the p-periodic stretch, capped at a+tau-1
        uint64_t x = a + 2 * (uint64_t)P - p;
        uint64_t capR = std::min<uint64_t>(n - 1, a + tau - 1);
        while (x + p <= capR && T[x] == T[x + p]) x++;
        uint64_t R = x + p - 1;
        // left end, capped at a-A+1
        uint64_t capL = a >= A - 1 ? a - (A - 1) : 0;
        uint64_t L = a;
        while (L > capL && T[L - 1] == T[L - 1 + p]) L--;
        if (R + 1 < tau) continue;
        uint64_t hi = std::min<uint64_t>(a, R + 1 - tau);
        if (hi >= nw) hi = nw - 1;
        for (uint64_t j = L; j <= hi && j + 1 != 0; j++) {
            q[j] = 1;
            marked++;
        }
    }
    return marked;
}

// id(j) for j in [0, n-tau+1] (last one is the virtual terminator window).
inline void window_ids(const uint8_t* T, uint64_t n, uint64_t tau, uint32_t base, uint32_t qmax,
                       id_vec& id, uint64_t& num_q) {
    id.clear();
    num_q = 0;
    if (n + 1 < tau) return;
    uint64_t nw = n - tau + 1;  // real windows (may be 0 when n == tau-1)
    id.resize(nw + 1);
    flag_vec q;
    num_q = mark_q(T, n, tau, qmax, q);
    const uint32_t btau = pow_u32(base, tau);
    if (nw) {
        int nthr = 1;
#ifdef _OPENMP
        nthr = omp_get_max_threads();
#endif
        uint64_t chunk = (nw + nthr - 1) / nthr;
#pragma omp parallel for schedule(static, 1)
        for (int t = 0; t < nthr; t++) {
            uint64_t b = (uint64_t)t * chunk, e = std::min(nw, b + chunk);
            if (b >= e) continue;
            uint32_t fp = 0;
            for (uint64_t k = 0; k < tau; k++) fp = fp * base + T[b + k];
            for (uint64_t j = b; j < e; j++) {
                id[j] = q[j] ? ID_INF : ((fp & 0x7FFFFFFEu) | 1u);
                if (j + 1 < e) fp = fp * base + T[j + tau] - btau * T[j];
            }
        }
    }
    id[nw] = 0;
}

struct sss_result {
    std::vector<uint64_t> S;
    uint64_t num_q = 0;   // |Q|
    bool has_runs = false;
};

inline sss_result build_sss(const uint8_t* T, uint64_t n, uint64_t tau,
                            uint32_t base = DEFAULT_SSS_BASE, uint32_t qmax = DEFAULT_QMAX) {
    sss_result r;
    if (n + 1 < 2 * tau) return r;
    id_vec id;
    window_ids(T, n, tau, base, qmax, id, r.num_q);
    r.has_runs = r.num_q > 0;
    uint64_t cnt = n - 2 * tau + 2;  // candidate positions 0..n-2tau+1
    // sliding minimum over [i, i+tau] with a monotone queue (a ring of tau + 2 window indices); the candidates are cut into
    // one chunk per thread, each chunk warms its queue up on the tau ids in front of its first window end
    int nthr = 1;
#ifdef _OPENMP
    nthr = omp_get_max_threads();
#endif
    if (cnt < (uint64_t)nthr * 4 * tau) nthr = 1;
    std::vector<std::vector<uint64_t>> part(nthr);
    const uint64_t chunk = (cnt + nthr - 1) / nthr;
#pragma omp parallel for schedule(static, 1) num_threads(nthr)
    for (int t = 0; t < nthr; t++) {
        const uint64_t cb = (uint64_t)t * chunk, ce = std::min(cnt, cb + chunk);
        if (cb >= ce) continue;
        const uint64_t cap = tau + 2;
        std::vector<uint64_t> ring(cap);
        uint64_t head = 0, tail = 0;          // queue = ring[head % cap .. tail % cap)
        auto push = [&](uint64_t j) {
            while (tail > head && id[ring[(tail - 1) % cap]] >= id[j]) tail--;
            ring[tail % cap] = j;
            tail++;
        };
        for (uint64_t j = cb; j < cb + tau; j++) push(j);
        auto& out = part[t];
        out.reserve(2 * (ce - cb) / tau + 16);
        for (uint64_t i = cb; i < ce; i++) {
            push(i + tau);
            while (ring[head % cap] < i) head++;
            const uint32_t m = id[ring[head % cap]];
            if (m != ID_INF && (id[i] == m || id[i + tau] == m)) out.push_back(i);
        }
    }
    size_t total = 0;
    for (auto& v : part) total += v.size();
    r.S.reserve(total);
    for (auto& v : part) r.S.insert(r.S.end(), v.begin(), v.end());
    return r;
}

// ---- order of the suffixes starting at synchronizing positions -----------------------------

// three-way compare of the 3tau-contexts of text positions x and y (clipped at n;
// a proper prefix is smaller)
inline int cmp_ctx(const uint8_t* T, uint64_t n, uint64_t tau, uint64_t x, uint64_t y) {
    uint64_t lx = std::min<uint64_t>(3 * tau, n - x), ly = std::min<uint64_t>(3 * tau, n - y);
    int c = std::memcmp(T + x, T + y, std::min(lx, ly));
    if (c) return c;
    return lx < ly ? -1 : (lx > ly ? 1 : 0);
}

// Reduced string: one symbol per synchronizing position such that the suffix order of the
// symbol string equals the lexicographic order of the text suffixes (role of
// reduce_fps_3tau_lexicographic, lce_sss.hpp:82-83).  Symbol = dense rank of the pair
// (rank of the 3tau-context, run key).  Run key (0 when the next synchronizing position is at
// most tau away): the gap d to the next synchronizing position encodes the length of the
// periodic run that follows; the suffix whose run breaks first is smaller iff the breaking
// character is smaller than the periodic continuation ("type L"), so type-L keys sort
// ascending in d below 0 and type-G keys descending in d above 0.
inline std::vector<uint32_t> reduce_contexts(const uint8_t* T, uint64_t n, uint64_t tau,
                                             uint32_t qmax, const std::vector<uint64_t>& S) {
    const uint64_t s = S.size();
    std::vector<uint32_t> red(s);
    if (!s) return red;
    std::vector<uint32_t> ord(s);
    std::iota(ord.begin(), ord.end(), 0u);
    auto less = [&](uint32_t a, uint32_t b) { return cmp_ctx(T, n, tau, S[a], S[b]) < 0; };
#ifdef _OPENMP
    __gnu_parallel::sort(ord.begin(), ord.end(), less);
#else
    std::sort(ord.begin(), ord.end(), less);
#endif
    std::vector<uint32_t> crank(s);
    uint32_t cr = 0;
    for (uint64_t i = 0; i < s; i++) {
        if (i && cmp_ctx(T, n, tau, S[ord[i - 1]], S[ord[i]]) != 0) cr = (uint32_t)i;
        crank[ord[i]] = cr;
    }
    const uint32_t P = q_period_bound(tau, qmax);
    const int64_t BIG = int64_t(1) << 41;
    std::vector<int64_t> rkey(s, 0);
    const int64_t s1 = (int64_t)s - 1;
#pragma omp parallel for schedule(static)
    for (int64_t k = 0; k < s1; k++) {
        uint64_t d = S[k + 1] - S[k];
        if (d <= tau) continue;
        uint64_t a = S[k + 1];  // window a lies in Q; its period is that of T[a..a+2P)
        uint32_t p = 0;
        for (uint32_t c = 1; c <= P && !p; c++) {
            bool ok = true;
            for (uint64_t x = 0; x + c < 2 * (uint64_t)P; x++)
                if (T[a + x] != T[a + x + c]) { ok = false; break; }
            if (ok) p = c;
        }
        uint64_t brk = a + 2 * tau - 1;  // first position after the run
        bool typeL = (brk >= n) || (T[brk] < T[brk - p]);
        rkey[k] = typeL ? -(BIG - (int64_t)d) : (BIG - (int64_t)d);
    }
    auto less2 = [&](uint32_t a, uint32_t b) {
        return crank[a] != crank[b] ? crank[a] < crank[b] : rkey[a] < rkey[b];
    };
#ifdef _OPENMP
    __gnu_parallel::sort(ord.begin(), ord.end(), less2);
#else
    std::sort(ord.begin(), ord.end(), less2);
#endif
    uint32_t r = 0;
    for (uint64_t i = 0; i < s; i++) {
        if (i && (crank[ord[i - 1]] != crank[ord[i]] || rkey[ord[i - 1]] != rkey[ord[i]])) r = (uint32_t)i;
        red[ord[i]] = r;
    }
    return red;
}

// Suffix array of an integer string whose last symbol is unique (role of gsaca_for_lce,
// lce_classic_for_sss.hpp:52).  Prefix doubling.
template <typename sym_t>
inline void suffix_sort_ints(const sym_t* text, uint32_t* sa, uint64_t s) {
    if (!s) return;
    std::vector<uint64_t> rank(s), tmp(s);
    std::iota(sa, sa + s, 0u);
    auto psort = [&](auto cmp) {
#ifdef _OPENMP
        __gnu_parallel::sort(sa, sa + s, cmp);
#else
        std::sort(sa, sa + s, cmp);
#endif
    };
    psort([&](uint32_t a, uint32_t b) { return text[a] < text[b]; });
    bool all_distinct = true;
    rank[sa[0]] = 0;
    for (uint64_t i = 1; i < s; i++) {
        bool same = text[sa[i]] == text[sa[i - 1]];
        rank[sa[i]] = same ? rank[sa[i - 1]] : i;
        all_distinct &= !same;
    }
    for (uint64_t h = 1; !all_distinct && h < s; h *= 2) {
        auto key2 = [&](uint32_t a) -> int64_t { return a + h < s ? (int64_t)rank[a + h] : -1; };
        psort([&](uint32_t a, uint32_t b) {
            return rank[a] != rank[b] ? rank[a] < rank[b] : key2(a) < key2(b);
        });
        all_distinct = true;
        tmp[sa[0]] = 0;
        for (uint64_t i = 1; i < s; i++) {
            bool same = rank[sa[i]] == rank[sa[i - 1]] && key2(sa[i]) == key2(sa[i - 1]);
            tmp[sa[i]] = same ? tmp[sa[i - 1]] : i;
            all_distinct &= !same;
        }
        rank.swap(tmp);
    }
}

inline std::vector<uint32_t> sort_sync_suffixes(const uint8_t* T, uint64_t n, uint64_t tau,
                                                uint32_t qmax, const std::vector<uint64_t>& S) {
    std::vector<uint32_t> red = reduce_contexts(T, n, tau, qmax, S);
    std::vector<uint32_t> sa(S.size());
    suffix_sort_ints(red.data(), sa.data(), S.size());
    return sa;
}

// brute force, for tests
inline std::vector<uint32_t> sort_sync_suffixes_naive(const uint8_t* T, uint64_t n,
                                                      const std::vector<uint64_t>& S) {
    std::vector<uint32_t> sa(S.size());
    std::iota(sa.begin(), sa.end(), 0u);
    std::sort(sa.begin(), sa.end(), [&](uint32_t a, uint32_t b) {
        uint64_t x = S[a], y = S[b];
        uint64_t l = naive_lce(T, n, x, y);
        if (x + l == n) return true;   // x is a proper prefix of y
        if (y + l == n) return false;
        return T[x + l] < T[y + l];
    });
    return sa;
}

// successor index (role of lce::pred::pred_index, lce_sss.hpp:68-69,132-134)
struct succ_index {
    const std::vector<uint64_t>* S = nullptr;
    unsigned shift = 0;
    std::vector<uint64_t> first;  // first[b] = smallest idx with S[idx] >> shift >= b
    void build(const std::vector<uint64_t>& s_, unsigned shift_) {
        S = &s_;
        shift = shift_;
        uint64_t nb = s_.empty() ? 1 : (s_.back() >> shift) + 2;
        first.assign(nb + 1, s_.size());
        uint64_t idx = 0;
        for (uint64_t b = 0; b <= nb; b++) {
            while (idx < s_.size() && (s_[idx] >> shift) < b) idx++;
            first[b] = idx;
        }
    }
    // returns idx of smallest S[idx] >= x, or |S|
    uint64_t successor(uint64_t x) const {
        uint64_t b = x >> shift;
        if (b + 1 >= first.size()) return S->size();
        uint64_t idx = first[b];
        while (idx < S->size() && (*S)[idx] < x) idx++;
        return idx;
    }
};

// range-minimum over an array, block decomposition (role of lce::rmq::rmq_n)
template <typename val_t>
struct block_rmq {
    const val_t* v = nullptr;
    uint64_t n = 0;
    static constexpr uint64_t B = 64;
    std::vector<std::vector<uint32_t>> tab;  // argmin per block range
    void build(const val_t* v_, uint64_t n_) {
        v = v_;
        n = n_;
        tab.clear();
        uint64_t nb = (n + B - 1) / B;
        if (!nb) return;
        tab.emplace_back(nb);
        for (uint64_t b = 0; b < nb; b++) {
            uint64_t best = b * B;
            for (uint64_t i = b * B; i < std::min(n, (b + 1) * B); i++)
                if (v[i] < v[best]) best = i;
            tab[0][b] = (uint32_t)best;
        }
        for (uint64_t l = 1; (uint64_t(1) << l) <= nb; l++) {
            uint64_t len = nb - (uint64_t(1) << l) + 1;
            tab.emplace_back(len);
            for (uint64_t b = 0; b < len; b++) {
                uint32_t x = tab[l - 1][b], y = tab[l - 1][b + (uint64_t(1) << (l - 1))];
                tab[l][b] = v[y] < v[x] ? y : x;
            }
        }
    }
    // argmin over [lo, hi] inclusive
    uint64_t argmin(uint64_t lo, uint64_t hi) const {
        uint64_t bl = lo / B, bh = hi / B, best = lo;
        if (bl == bh) {
            for (uint64_t i = lo; i <= hi; i++)
                if (v[i] < v[best]) best = i;
            return best;
        }
        for (uint64_t i = lo; i < (bl + 1) * B; i++)
            if (v[i] < v[best]) best = i;
        for (uint64_t i = bh * B; i <= hi; i++)
            if (v[i] < v[best]) best = i;
        if (bl + 1 < bh) {
            uint64_t a = bl + 1, b = bh - 1, len = b - a + 1;
            unsigned l = 63 - __builtin_clzll(len);
            uint32_t x = tab[l][a], y = tab[l][b + 1 - (uint64_t(1) << l)];
            if (v[x] < v[best]) best = x;
            if (v[y] < v[best]) best = y;
        }
        return best;
    }
};

}  // namespace oracle
