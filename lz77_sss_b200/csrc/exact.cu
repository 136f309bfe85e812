// Exact LZ77 on the device: lz77_sss<pos_t>::factorize_exact
// (/root/reference/include/lz77_sss/lz77_sss.hpp:188-200,318-341,608-674).
//
// The reference turns its approximation into the exact factorization with a sample index over the approximation's phrase
// boundaries (sample_index/*, transform_to_exact/*) and 2-D range queries; what it emits is, per section, the greedy
// longest-previous-factor parse (lengths are canonical, sources depend on the range structure's scan order).
// Here the same parse is computed the way KKP-style factorizers do, with every step data-parallel:
//   1. suffix array of the WHOLE text by prefix doubling: 8-byte big-endian keys, then (rank[i], rank[i+h]) pairs, each
//      round one stable LSD radix sort of all n pairs (util.cu) and a regrouping scan,
//   2. block minima of SA + sparse table: "nearest suffix in suffix order that starts earlier in the text" on either side
//      of a suffix is a nearest-smaller-value query (the same device functions the LPF stage uses, lce_device.cuh),
//   3. the longest previous factor at a position = the longer of the two common prefixes, computed with the LCE structure
//      of the approximation stage (3 tau bytes compared, then synchronizing positions + range minimum over LCP_S), so a
//      factor of a megabyte costs as much as one of a kilobyte,
//   4. the parse itself: the text is cut into blocks, one warp parses a block from a speculative entry (its first
//      position); the chain of blocks is followed by pointer doubling, blocks whose true entry differs are parsed again
//      until nothing changes (LZ parses re-synchronise, so this takes two or three rounds),
//   sections (num_threads > 1: 16 per thread, cut where the approximation's phrases are, transform_to_exact/common.cpp:
//   48-74) only truncate factors at their ends (with_samples.cpp:212-214).
// Positions are 32 bit: n < 2^32 - 2^16 (the suffix array of a longer text does not fit one GPU next to its sort buffers).
#include <algorithm>
#include <vector>

#include "common.cuh"
#include "device_utils.cuh"
#include "lce_device.cuh"
#include "stages.cuh"

namespace lz77sss {
namespace {

constexpr uint64_t XNONE = ~0ull;
constexpr uint32_t XBL = 1024;               // positions per parse block

__global__ void ex_key0_kernel(const uint8_t* __restrict__ T, uint64_t n, uint64_t* __restrict__ keys,
                               uint32_t* __restrict__ vals) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t le = load_u64_unaligned(T + i);        // zero bytes behind n (TEXT_PAD)
    const uint32_t lo = (uint32_t)le, hi = (uint32_t)(le >> 32);
    keys[i] = ((uint64_t)__byte_perm(lo, 0, 0x0123) << 32) | __byte_perm(hi, 0, 0x0123);   // first byte most significant
    vals[i] = (uint32_t)i;
}

__global__ void ex_flags_kernel(const uint64_t* __restrict__ keys, uint64_t n, uint32_t* __restrict__ flag,
                                uint32_t* __restrict__ acc) {
    const uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t f = (k == 0 || keys[k] != keys[k - 1]) ? 1u : 0u;
    flag[k] = f;
    acc[k] = f;
}

// rank of a suffix = number of groups in front of its group (dense, order preserving)
__global__ void ex_rank_kernel(const uint32_t* __restrict__ vals, const uint32_t* __restrict__ excl,
                               const uint32_t* __restrict__ flag, uint64_t n, uint32_t* __restrict__ rank) {
    const uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) rank[vals[k]] = excl[k] + flag[k] - 1u;
}

__global__ void ex_dkey_kernel(const uint32_t* __restrict__ rank, uint64_t n, uint64_t h, uint64_t* __restrict__ keys,
                               uint32_t* __restrict__ vals) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t lo = i + h < n ? (uint64_t)rank[i + h] + 1 : 0;     // a suffix that ends first is smaller
    keys[i] = ((uint64_t)rank[i] << 32) | lo;
    vals[i] = (uint32_t)i;
}

__global__ void ex_inverse_kernel(const uint32_t* __restrict__ isa, uint64_t n, uint32_t* __restrict__ sa) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) sa[isa[i]] = (uint32_t)i;
}

__global__ void ex_blkmin_kernel(const uint32_t* __restrict__ v, uint64_t n, uint32_t* __restrict__ tab, uint32_t nblk) {
    const uint64_t b = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    if (b >= nblk) return;
    const uint64_t i = b * RMQ_BLK + lane;
    uint32_t x = i < n ? v[i] : 0xFFFFFFFFu;
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) x = min(x, __shfl_xor_sync(0xffffffffu, x, d));
    if (lane == 0) tab[b] = x;
}

__global__ void ex_level_kernel(uint32_t* __restrict__ tab, uint32_t nblk, int l) {
    const uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t half = 1ull << (l - 1);
    if (b + (1ull << l) > nblk) return;
    tab[(uint64_t)l * nblk + b] = min(tab[(uint64_t)(l - 1) * nblk + b], tab[(uint64_t)(l - 1) * nblk + b + half]);
}

struct xblock {
    uint64_t entry, exit_;
};

struct xctx {
    uint64_t n;
    uint32_t nb;
    const uint32_t* SA;
    const uint32_t* ISA;
    const uint32_t* tab;
    uint32_t nblk;
    const uint64_t* sect;        // section starts, ascending, sect[0] = 0, sect[nsect] = n
    uint32_t nsect;
    xblock* blk;
    uint8_t* marked;
    uint8_t* dirty;
    uint32_t* tr_cnt;
    uint32_t* tr_src;
    uint32_t* tr_len;
    uint32_t* changed;
};

// one warp per block: parse from the block's entry to the first position behind the block
template <typename pos_t>
__global__ void __launch_bounds__(128) ex_parse_kernel(const xctx C, const lce_view<pos_t> V) {
    const uint32_t b = (uint32_t)(((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const unsigned lane = threadIdx.x & 31;
    if (b >= C.nb || !C.marked[b] || !C.dirty[b]) return;
    const uint64_t bend = min(C.n, (uint64_t)(b + 1) * XBL);
    uint64_t i = C.blk[b].entry;
    // the section of i: its end cuts factors
    uint32_t lo = 0, hi = C.nsect;                  // last section start <= i
    while (lo + 1 < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (C.sect[mid] <= i) lo = mid;
        else hi = mid;
    }
    uint64_t send = C.sect[lo + 1];
    uint32_t sidx = lo + 1;
    const uint64_t slot0 = (uint64_t)b * XBL;
    uint32_t t = 0;
    while (i < bend) {
        while (i >= send) send = C.sect[++sidx];
        uint64_t cand = XNONE;
        if (lane < 2) {
            const uint64_t r = C.ISA[i];
            const uint32_t q = lane == 0 ? prev_smaller(C.SA, C.tab, C.nblk, r, (uint32_t)i)
                                         : next_smaller(C.SA, C.tab, C.nblk, C.n, r, (uint32_t)i);
            if (q != NONE32) cand = C.SA[q];
        }
        const uint64_t j1 = __shfl_sync(0xffffffffu, cand, 0), j2 = __shfl_sync(0xffffffffu, cand, 1);
        uint64_t len = 0, src = 0;
        if (j1 != XNONE) { len = warp_lce_r(V, j1, i); src = j1; }
        if (j2 != XNONE) {
            const uint64_t l2 = warp_lce_r(V, j2, i);
            if (l2 > len) { len = l2; src = j2; }
        }
        if (i + len > send) len = send - i;         // a section ends here (with_samples.cpp:212-214)
        if (len == 0) src = V.T[i];                 // literal (also: a copy cut down to nothing cannot happen, send > i)
        if (lane == 0) {
            C.tr_src[slot0 + t] = (uint32_t)src;
            C.tr_len[slot0 + t] = (uint32_t)len;
        }
        t++;
        i += len ? len : 1;
    }
    if (lane == 0) {
        if (C.blk[b].exit_ != i) atomicOr(C.changed, 1u);
        C.blk[b].exit_ = i;
        C.tr_cnt[b] = t;
        C.dirty[b] = 0;
    }
}

__global__ void ex_spec_kernel(const xctx C) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= C.nb) return;
    C.blk[b].entry = (uint64_t)b * XBL;
    C.blk[b].exit_ = XNONE;
    C.marked[b] = 1;
    C.dirty[b] = 1;
    C.tr_cnt[b] = 0;
}

// chain of blocks from block 0 by pointer doubling (phases as in gap_parse.cu, link_phase_kernel)
__global__ void __launch_bounds__(256) ex_link_kernel(const xctx C, uint32_t* __restrict__ J, int l, int phase) {
    const uint32_t nb = C.nb, N = nb + 1;
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= N) return;
    if (phase == 0) {
        uint32_t j = nb;
        if (b < nb) {
            const uint64_t x = C.blk[b].exit_;
            if (x != XNONE && x < C.n) j = (uint32_t)(x / XBL);
            C.marked[b] = b == 0;
        }
        J[b] = j;
    } else if (phase == 1) {
        const uint32_t* src = J + (size_t)(l - 1) * N;
        J[(size_t)l * N + b] = src[src[b]];
    } else if (phase == 2) {
        if (b < nb && C.marked[b]) {
            const uint32_t j = J[(size_t)l * N + b];
            if (j < nb) C.marked[j] = 1;
        }
    } else {
        if (b >= nb) return;
        if (b == 0 && C.blk[0].entry != 0) { C.blk[0].entry = 0; C.dirty[0] = 1; }
        if (C.marked[b]) {
            const uint32_t j = J[b];
            if (j < nb && C.blk[j].entry != C.blk[b].exit_) {
                C.blk[j].entry = C.blk[b].exit_;
                C.dirty[j] = 1;
            }
        }
    }
}

__global__ void ex_count_kernel(const xctx C, uint32_t* __restrict__ cnt, uint32_t* __restrict__ pending) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= C.nb) return;
    cnt[b] = C.marked[b] ? C.tr_cnt[b] : 0u;
    if (C.marked[b] && C.dirty[b]) atomicOr(pending, 1u);
}

__global__ void __launch_bounds__(128) ex_gather_kernel(const xctx C, const uint32_t* __restrict__ off,
                                                        uint64_t* __restrict__ fsrc, uint64_t* __restrict__ flen) {
    const uint32_t b = (uint32_t)(((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const unsigned lane = threadIdx.x & 31;
    if (b >= C.nb || !C.marked[b]) return;
    const uint32_t cnt = C.tr_cnt[b], o = off[b];
    const uint64_t slot0 = (uint64_t)b * XBL;
    for (uint32_t t = lane; t < cnt; t += 32) {
        fsrc[o + t] = C.tr_src[slot0 + t];
        flen[o + t] = C.tr_len[slot0 + t];
    }
}

int xbits_for(uint64_t maxval) {
    int b = 0;
    while (maxval >> b) b++;
    return b ? b : 1;
}

}  // namespace

template <typename pos_t>
void exact_run(const uint8_t* d_text, uint64_t n, uint32_t tau, const pos_t* d_S, uint64_t s, suffix_state<pos_t>& ss,
               lce_state<pos_t>& ls, const std::vector<uint64_t>& sect, exact_scratch& sc, cudaStream_t st,
               pinned_vec<uint64_t>& out_src, pinned_vec<uint64_t>& out_len) {
    if (n >= 0xFFFF0000ull) throw std::invalid_argument("factorize_exact: n must be below 2^32 - 2^16 in this version");
    const unsigned gb = ceil_div_u(n, 256);
    sc.k0.reserve(n + 8); sc.k1.reserve(n + 8); sc.v0.reserve(n + 8); sc.v1.reserve(n + 8);
    sc.rank.reserve(n + 8); sc.flag.reserve(n + 8); sc.acc.reserve(n + 8);
    sc.counter.reserve(8);
    uint64_t *keys = sc.k0, *kalt = sc.k1;
    uint32_t *vals = sc.v0, *valt = sc.v1;
    sc.rounds = 0;

    // ---- 1. suffix array by prefix doubling ----
    auto regroup = [&]() -> uint32_t {
        ex_flags_kernel<<<gb, 256, 0, st>>>(keys, n, sc.flag, sc.acc);
        KERNEL_CHECK();
        exclusive_scan_u32(sc.acc, n, sc.scan, st, sc.counter.get());
        ex_rank_kernel<<<gb, 256, 0, st>>>(vals, sc.acc, sc.flag, n, sc.rank);
        KERNEL_CHECK();
        uint32_t ng = 0;
        CUDA_CHECK(cudaMemcpyAsync(&ng, sc.counter.get(), 4, cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        return ng;
    };
    ex_key0_kernel<<<gb, 256, 0, st>>>(d_text, n, keys, vals);
    KERNEL_CHECK();
    radix_sort_pairs(keys, kalt, vals, valt, n, 0, 64, sc.sort, st);
    uint32_t ng = regroup();
    for (uint64_t h = 8; ng < n; h *= 2) {
        if (sc.rounds++ > 40) throw std::runtime_error("factorize_exact: prefix doubling did not converge");
        ex_dkey_kernel<<<gb, 256, 0, st>>>(sc.rank, n, h, keys, vals);
        KERNEL_CHECK();
        const int rb = xbits_for(ng);
        radix_sort_pairs(keys, kalt, vals, valt, n, 0, rb, sc.sort, st);
        radix_sort_pairs(keys, kalt, vals, valt, n, 32, 32 + rb, sc.sort, st);
        ng = regroup();
    }
    // rank is the inverse suffix array now; the suffix array goes where the sort payload was
    uint32_t* ISA = sc.rank;
    uint32_t* SA = vals;
    ex_inverse_kernel<<<gb, 256, 0, st>>>(ISA, n, SA);
    KERNEL_CHECK();

    // ---- 2. block minima of SA and their sparse table ----
    const uint32_t nblk = ceil_div_u(n, RMQ_BLK);
    int levels = 1;
    while ((2ull << (levels - 1)) <= nblk) levels++;
    sc.tab.reserve((uint64_t)levels * nblk);
    ex_blkmin_kernel<<<ceil_div_u((uint64_t)nblk * 32, 256), 256, 0, st>>>(SA, n, sc.tab, nblk);
    KERNEL_CHECK();
    for (int l = 1; l < levels; l++) {
        ex_level_kernel<<<ceil_div_u(nblk, 256), 256, 0, st>>>(sc.tab, nblk, l);
        KERNEL_CHECK();
    }

    // ---- 3 + 4. block-speculative greedy parse ----
    const uint32_t nb = ceil_div_u(n, XBL);
    int jl = 1;
    while ((1ull << (jl - 1)) < (uint64_t)nb + 1) jl++;
    sc.blk.reserve((uint64_t)nb * sizeof(xblock));
    sc.marked.reserve(nb + 1); sc.dirty.reserve(nb + 1);
    sc.tr_cnt.reserve(nb + 2); sc.cnt.reserve(nb + 2);
    // the key buffers are free now: the traces (two 32-bit words per position) live there
    uint32_t* tr_src = reinterpret_cast<uint32_t*>(keys);
    uint32_t* tr_len = reinterpret_cast<uint32_t*>(kalt);
    sc.J.reserve((uint64_t)jl * (nb + 1));
    sc.sect.reserve(sect.size());
    CUDA_CHECK(cudaMemcpyAsync(sc.sect.get(), sect.data(), sect.size() * 8, cudaMemcpyHostToDevice, st));
    xctx C;
    C.n = n; C.nb = nb; C.SA = SA; C.ISA = ISA; C.tab = sc.tab; C.nblk = nblk;
    C.sect = sc.sect; C.nsect = (uint32_t)sect.size() - 1;
    C.blk = reinterpret_cast<xblock*>(sc.blk.get());
    C.marked = sc.marked; C.dirty = sc.dirty; C.tr_cnt = sc.tr_cnt; C.tr_src = tr_src; C.tr_len = tr_len;
    C.changed = sc.counter.get() + 2;
    const lce_view<pos_t> V = make_lce_view<pos_t>(d_text, n, tau, d_S, s, ss, ls);
    const unsigned gw = ceil_div_u((uint64_t)nb * 32, 128), gl = ceil_div_u((uint64_t)nb + 1, 256);
    ex_spec_kernel<<<ceil_div_u(nb, 256), 256, 0, st>>>(C);
    KERNEL_CHECK();
    ex_parse_kernel<pos_t><<<gw, 128, 0, st>>>(C, V);
    KERNEL_CHECK();
    sc.parse_rounds = 1;
    for (uint32_t round = 1;; round++) {
        if (round > nb + 4) throw std::runtime_error("factorize_exact: the block chain did not reach its fixed point");
        ex_link_kernel<<<gl, 256, 0, st>>>(C, sc.J, 0, 0);
        KERNEL_CHECK();
        for (int l = 1; l < jl; l++) { ex_link_kernel<<<gl, 256, 0, st>>>(C, sc.J, l, 1); KERNEL_CHECK(); }
        for (int l = jl - 1; l >= 0; l--) { ex_link_kernel<<<gl, 256, 0, st>>>(C, sc.J, l, 2); KERNEL_CHECK(); }
        ex_link_kernel<<<gl, 256, 0, st>>>(C, sc.J, 0, 3);
        KERNEL_CHECK();
        CUDA_CHECK(cudaMemsetAsync(sc.counter.get() + 2, 0, 8, st));
        ex_count_kernel<<<ceil_div_u(nb, 256), 256, 0, st>>>(C, sc.cnt, sc.counter.get() + 3);
        KERNEL_CHECK();
        uint32_t pending = 0;
        CUDA_CHECK(cudaMemcpyAsync(&pending, sc.counter.get() + 3, 4, cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        if (!pending) break;                        // every block on the chain was parsed from its true entry
        ex_parse_kernel<pos_t><<<gw, 128, 0, st>>>(C, V);
        KERNEL_CHECK();
        sc.parse_rounds++;
    }
    // factors of the blocks on the chain, in order (sc.cnt holds their counts)
    exclusive_scan_u32(sc.cnt, nb, sc.scan, st, sc.cnt.get() + nb);
    uint32_t z = 0;
    CUDA_CHECK(cudaMemcpyAsync(&z, sc.cnt.get() + nb, 4, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    out_src.resize(z);
    out_len.resize(z);
    if (z) {
        sc.fsrc.reserve(z); sc.flen.reserve(z);
        ex_gather_kernel<<<gw, 128, 0, st>>>(C, sc.cnt, sc.fsrc, sc.flen);
        KERNEL_CHECK();
        CUDA_CHECK(cudaMemcpyAsync(out_src.data(), sc.fsrc.get(), (size_t)z * 8, cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaMemcpyAsync(out_len.data(), sc.flen.get(), (size_t)z * 8, cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
    }
}

template void exact_run<uint32_t>(const uint8_t*, uint64_t, uint32_t, const uint32_t*, uint64_t, suffix_state<uint32_t>&,
                                  lce_state<uint32_t>&, const std::vector<uint64_t>&, exact_scratch&, cudaStream_t,
                                  pinned_vec<uint64_t>&, pinned_vec<uint64_t>&);
template void exact_run<uint64_t>(const uint8_t*, uint64_t, uint32_t, const uint64_t*, uint64_t, suffix_state<uint64_t>&,
                                  lce_state<uint64_t>&, const std::vector<uint64_t>&, exact_scratch&, cudaStream_t,
                                  pinned_vec<uint64_t>&, pinned_vec<uint64_t>&);

}  // namespace lz77sss
