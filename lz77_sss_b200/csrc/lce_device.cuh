// Device-side exact LCE_R on the structures of lce_state / suffix_state (lce_sss::lce_lr,
// /root/reference/patched-files/external/lce/include/ds/lce_sss.hpp:123-177): naive compare for up to 3*tau
// characters, then successor synchronizing positions + range minimum over LCP_S.
#pragma once
#include "common.cuh"
#include "device_utils.cuh"
#include "stages.cuh"

namespace lz77sss {

constexpr int RMQ_BLK = 32;

// min of the block minima over blocks [lo, hi] (inclusive, lo <= hi)
template <typename val_t>
__device__ __forceinline__ val_t blk_range_min(const val_t* __restrict__ tab, uint32_t nblk, uint64_t lo, uint64_t hi) {
    const int l = 63 - __clzll((long long)(hi - lo + 1));
    val_t x = tab[(uint64_t)l * nblk + lo], y = tab[(uint64_t)l * nblk + hi + 1 - (1ull << l)];
    return y < x ? y : x;
}

// min v[lo..hi] (inclusive, lo <= hi)
template <typename val_t>
__device__ __forceinline__ val_t range_min(const val_t* __restrict__ v, const val_t* __restrict__ tab, uint32_t nblk,
                                           uint64_t lo, uint64_t hi) {
    const uint64_t bl = lo / RMQ_BLK, bh = hi / RMQ_BLK;
    val_t m = v[lo];
    if (bl == bh) {
        for (uint64_t i = lo + 1; i <= hi; i++) m = v[i] < m ? v[i] : m;
        return m;
    }
    for (uint64_t i = lo + 1; i < (bl + 1) * RMQ_BLK; i++) m = v[i] < m ? v[i] : m;
    for (uint64_t i = bh * RMQ_BLK; i <= hi; i++) m = v[i] < m ? v[i] : m;
    if (bl + 1 < bh) {
        val_t x = blk_range_min(tab, nblk, bl + 1, bh - 1);
        m = x < m ? x : m;
    }
    return m;
}


// ---- nearest smaller values over an array SA with a sparse table `tab` of its block minima (blocks of RMQ_BLK) -------------
// largest r' < r with SA[r'] < v, or NONE32
__device__ __forceinline__ uint32_t prev_smaller(const uint32_t* __restrict__ SA, const uint32_t* __restrict__ tab,
                                                 uint32_t nblk, uint64_t r, uint32_t v) {
    const uint64_t b = r / RMQ_BLK;
    for (uint64_t j = r; j > b * RMQ_BLK;) {
        j--;
        if (SA[j] < v) return (uint32_t)j;
    }
    if (b == 0) return NONE32;
    uint64_t e = b, w = 1, lo;
    while (true) {
        lo = e >= w ? e - w : 0;
        if (blk_range_min(tab, nblk, lo, e - 1) < v) break;
        if (lo == 0) return NONE32;
        w *= 2;
    }
    uint64_t hi = e - 1;
    while (lo < hi) {
        uint64_t mid = (lo + hi + 1) / 2;
        if (blk_range_min(tab, nblk, mid, hi) < v) lo = mid;
        else hi = mid - 1;
    }
    for (uint64_t j = (lo + 1) * RMQ_BLK; j > lo * RMQ_BLK;) {
        j--;
        if (SA[j] < v) return (uint32_t)j;
    }
    return NONE32;  // unreachable
}

// smallest r' > r with SA[r'] < v, or NONE32
__device__ __forceinline__ uint32_t next_smaller(const uint32_t* __restrict__ SA, const uint32_t* __restrict__ tab,
                                                 uint32_t nblk, uint64_t s, uint64_t r, uint32_t v) {
    const uint64_t b = r / RMQ_BLK;
    const uint64_t bend = min(s, (b + 1) * RMQ_BLK);
    for (uint64_t j = r + 1; j < bend; j++)
        if (SA[j] < v) return (uint32_t)j;
    if (b + 1 >= nblk) return NONE32;
    uint64_t e = b + 1, w = 1, hi;   // search blocks [e, ...]
    while (true) {
        hi = e + w - 1 < nblk ? e + w - 1 : nblk - 1;
        if (blk_range_min(tab, nblk, e, hi) < v) break;
        if (hi == nblk - 1) return NONE32;
        w *= 2;
    }
    uint64_t lo = e;
    while (lo < hi) {
        uint64_t mid = (lo + hi) / 2;
        if (blk_range_min(tab, nblk, lo, mid) < v) hi = mid;
        else lo = mid + 1;
    }
    const uint64_t be = min(s, (lo + 1) * RMQ_BLK);
    for (uint64_t j = lo * RMQ_BLK; j < be; j++)
        if (SA[j] < v) return (uint32_t)j;
    return NONE32;  // unreachable
}

template <typename pos_t>
struct lce_view {
    const uint8_t* T;
    uint64_t n;
    uint32_t tau;
    const pos_t* S;
    uint64_t s;
    const uint32_t* ISA;
    const pos_t* LCP;
    const pos_t* lcp_tab;
    uint32_t nblk;
    const uint32_t* first;
    uint32_t shift;
    uint64_t nb;
};

template <typename pos_t>
inline lce_view<pos_t> make_lce_view(const uint8_t* T, uint64_t n, uint32_t tau, const pos_t* S, uint64_t s,
                                     suffix_state<pos_t>& ss, lce_state<pos_t>& ls) {
    lce_view<pos_t> v;
    v.T = T; v.n = n; v.tau = tau; v.S = S; v.s = s;
    v.ISA = ss.ISA.get(); v.LCP = ss.LCP.get(); v.lcp_tab = ls.lcp_blkmin.get(); v.nblk = ls.nblk;
    v.first = ls.succ_first.get(); v.shift = ls.succ_shift; v.nb = ls.succ_nb;
    return v;
}

// All 32 lanes call it with the same (i, j); returns LCE of T[i..) and T[j..) (n - i for i == j).
template <typename pos_t>
__device__ __forceinline__ uint64_t warp_lce_r(const lce_view<pos_t>& V, uint64_t i, uint64_t j) {
    if (i == j) return V.n - i;
    const uint64_t l = min(i, j), r = max(i, j);
    const uint64_t lce_max = V.n - r;
    const uint64_t local_max = min((uint64_t)3 * V.tau, lce_max);
    const uint64_t local = warp_lce(V.T, l, r, 0, local_max);
    if (local < local_max || local == lce_max) return local;
    auto succ = [&](uint64_t x) -> uint64_t {
        uint64_t b = x >> V.shift;
        uint64_t idx = b < V.nb ? V.first[b] : V.s;
        while (idx < V.s && (uint64_t)V.S[idx] < x) idx++;
        return idx;
    };
    const uint64_t l_ = succ(l), r_ = succ(r);
    const uint64_t dl = (uint64_t)V.S[l_] - l, dr = (uint64_t)V.S[r_] - r;
    if (dl != dr) return min(dl, dr) + 2 * (uint64_t)V.tau - 1;
    uint64_t a = V.ISA[l_], b = V.ISA[r_];
    if (a > b) { uint64_t t = a; a = b; b = t; }
    return dl + (uint64_t)range_min(V.LCP, V.lcp_tab, V.nblk, a + 1, b);
}

// warp_lce_r for a group of G lanes (all lanes of the group call it with the same (i, j))
template <typename pos_t, int G>
__device__ __forceinline__ uint64_t group_lce_r(const lce_view<pos_t>& V, uint64_t i, uint64_t j, unsigned gmask, unsigned glane) {
    if (i == j) return V.n - i;
    const uint64_t l = min(i, j), r = max(i, j);
    const uint64_t lce_max = V.n - r;
    const uint64_t local_max = min((uint64_t)3 * V.tau, lce_max);
    const uint64_t local = group_lce<G>(V.T, l, r, 0, local_max, gmask, glane);
    if (local < local_max || local == lce_max) return local;
    auto succ = [&](uint64_t x) -> uint64_t {
        uint64_t b = x >> V.shift;
        uint64_t idx = b < V.nb ? V.first[b] : V.s;
        while (idx < V.s && (uint64_t)V.S[idx] < x) idx++;
        return idx;
    };
    const uint64_t l_ = succ(l), r_ = succ(r);
    const uint64_t dl = (uint64_t)V.S[l_] - l, dr = (uint64_t)V.S[r_] - r;
    if (dl != dr) return min(dl, dr) + 2 * (uint64_t)V.tau - 1;
    uint64_t a = V.ISA[l_], b = V.ISA[r_];
    if (a > b) { uint64_t t = a; a = b; b = t; }
    return dl + (uint64_t)range_min(V.LCP, V.lcp_tab, V.nblk, a + 1, b);
}

}  // namespace lz77sss
