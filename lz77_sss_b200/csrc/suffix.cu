// K3: lexicographic order of the text suffixes that start at synchronizing positions (SA_S, ISA_S)
// and their LCP array in text characters.
//
// Replaces, for the path  lce_sss ctor -> reduce_fps_3tau_lexicographic -> lce_classic_for_sss ctor
// (/root/reference/patched-files/external/lce/include/ds/lce_sss.hpp:82-83,
//  .../ds/lce_classic_for_sss.hpp:52 (gsaca_for_lce), :65-69 (ISA), :82-120 (sparse LCP)).
//
//   1. context ranking: the 3tau-contexts T[S[i]..S[i]+3tau) are ranked exactly by a radix sort on their
//      first 7 bytes followed by rounds of "batched LCE against the group pivot + radix sort on
//      (group, side, mismatch offset, mismatch character)" — a fingerprint-free, deterministic
//      multikey refinement; every round reads each unresolved context at most once past the
//      prefix it already shares with its group.
//   2. run keys for positions followed by a periodic run (gap to the next position > tau).
//   3. prefix doubling over the reduced string with our radix sort; every level's rank array is kept:
//      together they answer "how many reduced symbols do suffixes a and b share" in O(levels).
//   4. LCP_S[r] = (S[a+k]-S[a]) + naive LCE from the first differing symbol (warp-cooperative).
#include "common.cuh"
#include "device_utils.cuh"
#include "stages.cuh"

namespace lz77sss {
namespace {

constexpr uint32_t MAXD = 16383;   // 14-bit mismatch-offset field (3*tau <= 12288)

template <typename pos_t>
__global__ void ctx_key0_kernel(const uint8_t* __restrict__ T, const pos_t* __restrict__ S, uint64_t s,
                                uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= s) return;
    uint64_t v = load_u64_unaligned(T + (uint64_t)S[k]);
    // big-endian order of the first 7 bytes; every context is at least 2tau-1 >= 7 bytes long
    uint64_t be = __byte_perm((uint32_t)(v >> 32), 0, 0x0123) | ((uint64_t)__byte_perm((uint32_t)v, 0, 0x0123) << 32);
    keys[k] = (be & ~0xFFull) | 7ull;
    vals[k] = (uint32_t)k;
}

__global__ void head_flags_kernel(const uint64_t* __restrict__ keys, uint64_t s, uint32_t* __restrict__ flags,
                                  uint32_t* __restrict__ gid) {
    uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= s) return;
    uint32_t f = (k == 0 || keys[k] != keys[k - 1]) ? 1u : 0u;
    flags[k] = f;
    gid[k] = f;
}

// gid holds the exclusive scan of the flags; make it the dense group id and record group starts
__global__ void group_ids_kernel(const uint32_t* __restrict__ flags, uint32_t* __restrict__ gid, uint64_t s,
                                 uint32_t* __restrict__ gstart, const uint32_t* __restrict__ total) {
    uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k == 0) gstart[*total] = (uint32_t)s;
    if (k >= s) return;
    uint32_t g = gid[k] + flags[k] - 1;
    gid[k] = g;
    if (flags[k]) gstart[g] = (uint32_t)k;
}

// LCE of every unresolved context with the context of its group's first element (pivot).  A warp takes 32 consecutive
// elements: the lanes classify their own element (coalesced loads; singletons, finished classes and pivots are done at once),
// then the warp walks over the elements that need a comparison, one warp-wide LCE each.  (One warp per element spent most of
// the later rounds, where few elements are unresolved, on launching 4.2 M warps that had nothing to compare.)
template <typename pos_t>
__global__ void __launch_bounds__(256) ctx_pivot_kernel(const uint8_t* __restrict__ T, uint64_t n, uint32_t tau,
                                                        const pos_t* __restrict__ S, uint64_t s,
                                                        const uint64_t* __restrict__ keys,
                                                        const uint32_t* __restrict__ ord,
                                                        const uint32_t* __restrict__ gid,
                                                        const uint32_t* __restrict__ gstart, int first,
                                                        uint64_t* __restrict__ newkeys, uint32_t* __restrict__ counter) {
    const uint64_t k0 = (((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) << 5;
    const unsigned lane = threadIdx.x & 31;
    if (k0 >= s) return;
    const uint64_t k = k0 + lane;
    bool work = false, counted = false;
    uint64_t x = 0, p = 0, off = 0, lim = 0, lenx = 0, lenp = 0, ghi = 0;
    if (k < s) {
        const uint32_t g = gid[k];
        const uint32_t gs = gstart[g], ge = gstart[g + 1];
        const uint64_t prev = keys[k];
        ghi = (uint64_t)g << 32;
        if (ge - gs == 1) {
            newkeys[k] = ghi;
        } else if (!first && ((prev >> 23) & 3) == 1) {      // class of contexts equal to an earlier pivot: final
            newkeys[k] = ghi | (1ull << 23);
        } else if (k == gs) {
            newkeys[k] = ghi | (1ull << 23);
            counted = true;
        } else {
            work = true;
            off = 7;
            if (!first) {
                const uint32_t side = (uint32_t)(prev >> 23) & 3, df = (uint32_t)(prev >> 9) & 0x3FFF;
                off = (side == 0 ? df : MAXD - df) + 1;
            }
            x = S[ord[k]];
            p = S[ord[gs]];
            lenx = min((uint64_t)3 * tau, n - x);
            lenp = min((uint64_t)3 * tau, n - p);
            lim = min(lenx, lenp);
            if (off > lim) off = lim;
        }
    }
    unsigned todo = __ballot_sync(0xffffffffu, work);
    uint64_t d = 0;
    while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const uint64_t xs = __shfl_sync(0xffffffffu, x, src), ps = __shfl_sync(0xffffffffu, p, src);
        const uint64_t os = __shfl_sync(0xffffffffu, off, src), ls = __shfl_sync(0xffffffffu, lim, src);
        const uint64_t ds = warp_lce(T, xs, ps, os, ls);
        if ((int)lane == src) d = ds;
    }
    if (work) {
        uint32_t side, cf;
        if (d == lim) {
            if (lenx == lenp) { side = 1; cf = 0; }
            else if (lenx < lenp) { side = 0; cf = 0; }                   // proper prefix of the pivot: smaller
            else { side = 2; cf = (uint32_t)T[x + d] + 1; }
        } else {
            const uint32_t cx = T[x + d], cp = T[p + d];
            side = cx < cp ? 0 : 2;
            cf = cx + 1;
        }
        const uint32_t df = side == 1 ? 0 : (side == 0 ? (uint32_t)d : MAXD - (uint32_t)d);
        newkeys[k] = ghi | ((uint64_t)side << 23) | ((uint64_t)df << 9) | cf;
        counted = true;
    }
    const unsigned c = __popc(__ballot_sync(0xffffffffu, counted));
    if (lane == 0 && c) atomicAdd(counter, c);
}

__global__ void scatter_rank_kernel(const uint32_t* __restrict__ ord, const uint32_t* __restrict__ gid, uint64_t s,
                                    uint32_t* __restrict__ out) {
    uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k < s) out[ord[k]] = gid[k];
}

// run keys (oracle/core/sss_core.hpp::reduce_contexts): only for positions whose successor is > tau away
template <typename pos_t>
__global__ void run_key_kernel(const uint8_t* __restrict__ T, uint64_t n, uint32_t tau, uint32_t P,
                               const pos_t* __restrict__ S, uint64_t s, uint64_t* __restrict__ enc,
                               uint32_t* __restrict__ counter) {
    uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= s) return;
    uint64_t e = 0;
    if (k + 1 < s) {
        uint64_t d = (uint64_t)S[k + 1] - (uint64_t)S[k];
        if (d > tau) {
            const uint64_t a = S[k + 1];
            uint32_t p = 0;
            for (uint32_t c = 1; c <= P && !p; c++) {
                bool ok = true;
                for (uint32_t x = 0; x + c < 2 * P; x++)
                    if (T[a + x] != T[a + x + c]) { ok = false; break; }
                if (ok) p = c;
            }
            const uint64_t brk = a + 2 * (uint64_t)tau - 1;
            const bool typeL = (brk >= n) || (T[brk] < T[brk - p]);
            e = typeL ? d : ((1ull << 41) - d);
            atomicAdd(counter, 1u);
        }
    }
    enc[k] = e;
}

__global__ void gather_key_kernel(const uint32_t* __restrict__ src, const uint32_t* __restrict__ idx, uint64_t s,
                                  uint64_t* __restrict__ keys) {
    uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k < s) keys[k] = src[idx[k]];
}
__global__ void enc_key_kernel(const uint64_t* __restrict__ enc, uint64_t s, uint64_t* __restrict__ keys,
                               uint32_t* __restrict__ vals) {
    uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k < s) { keys[k] = enc[k]; vals[k] = (uint32_t)k; }
}
__global__ void head_flags2_kernel(const uint32_t* __restrict__ crank, const uint64_t* __restrict__ enc,
                                   const uint32_t* __restrict__ ord, uint64_t s, uint32_t* __restrict__ flags,
                                   uint32_t* __restrict__ gid) {
    uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= s) return;
    uint32_t f = 1;
    if (k) {
        uint32_t a = ord[k - 1], b = ord[k];
        f = (crank[a] != crank[b] || enc[a] != enc[b]) ? 1u : 0u;
    }
    flags[k] = f;
    gid[k] = f;
}

__global__ void doubling_key_kernel(const uint32_t* __restrict__ rank, uint64_t s, uint64_t h,
                                    uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s) return;
    uint64_t lo = i + h < s ? (uint64_t)rank[i + h] + 1 : 0;
    keys[i] = ((uint64_t)rank[i] << 32) | lo;
    vals[i] = (uint32_t)i;
}

__global__ void inverse_kernel(const uint32_t* __restrict__ isa, uint64_t s, uint32_t* __restrict__ sa) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < s) sa[isa[i]] = (uint32_t)i;
}

struct level_ptrs {
    const uint32_t* r[40];
};

// number of leading reduced symbols shared by the suffixes of the reduced string at a and b (a != b)
__device__ __forceinline__ uint64_t sym_lce(const level_ptrs& lv, int levels, uint64_t s, uint64_t a, uint64_t b) {
    uint64_t k = 0;
    for (int l = levels - 1; l >= 0; l--) {
        if (a + k < s && b + k < s && lv.r[l][a + k] == lv.r[l][b + k]) k += 1ull << l;
    }
    return k;
}

template <typename pos_t>
__global__ void __launch_bounds__(256) lcp_kernel(const uint8_t* __restrict__ T, uint64_t n,
                                                  const pos_t* __restrict__ S, uint64_t s,
                                                  const uint32_t* __restrict__ SA, const level_ptrs lv, int levels,
                                                  pos_t* __restrict__ LCP) {
    const uint64_t r = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    if (r >= s) return;
    if (r == 0) {
        if (lane == 0) LCP[0] = 0;
        return;
    }
    const uint64_t a = SA[r - 1], b = SA[r];
    const uint64_t k = sym_lce(lv, levels, s, a, b);
    // both a+k and b+k exist: the last reduced symbol is unique
    const uint64_t pa = S[a + k], pb = S[b + k];
    const uint64_t lim = n - max(pa, pb);
    const uint64_t l = warp_lce(T, pa, pb, 0, lim);
    if (lane == 0) LCP[r] = (pos_t)((pa - (uint64_t)S[a]) + l);
}

int bits_for(uint64_t maxval) {
    int b = 0;
    while (maxval >> b) b++;
    return b ? b : 1;
}

template <typename pos_t>
uint32_t regroup(suffix_state<pos_t>& ss, const uint64_t* keys, uint64_t s, cudaStream_t st) {
    unsigned nb = ceil_div_u(s, 256);
    head_flags_kernel<<<nb, 256, 0, st>>>(keys, s, ss.flags, ss.gid);
    KERNEL_CHECK();
    exclusive_scan_u32(ss.gid, s, ss.scan, st, ss.counter.get() + 1);
    group_ids_kernel<<<nb, 256, 0, st>>>(ss.flags, ss.gid, s, ss.gstart, ss.counter.get() + 1);
    KERNEL_CHECK();
    uint32_t ng;
    CUDA_CHECK(cudaMemcpyAsync(&ng, ss.counter.get() + 1, 4, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    return ng;
}

}  // namespace

template <typename pos_t>
void suffix_sort_sync(const uint8_t* T, uint64_t n, uint32_t tau, uint32_t qmax, const pos_t* S, uint64_t s,
                      suffix_state<pos_t>& ss, cudaStream_t st) {
    ss.levels = 0;
    ss.ctx_levels = 0;
    ss.doubling_rounds = 0;
    ss.num_run_syms = 0;
    if (s == 0) return;
    if (s >= 0xFFFFFFF0ull) throw std::invalid_argument("suffix_sort_sync: |S| must be < 2^32");
    ss.SA.reserve(s); ss.ISA.reserve(s); ss.LCP.reserve(s);
    ss.k0.reserve(s); ss.k1.reserve(s); ss.v0.reserve(s); ss.v1.reserve(s);
    ss.flags.reserve(s); ss.gid.reserve(s); ss.gstart.reserve(s + 1); ss.crank.reserve(s);
    ss.counter.reserve(4);
    const unsigned nb = ceil_div_u(s, 256);
    const unsigned nbw = ceil_div_u(s * 32, 256);
    uint64_t *keys = ss.k0, *kalt = ss.k1;
    uint32_t *vals = ss.v0, *valt = ss.v1;

    // ---- 1. context ranking ----
    ctx_key0_kernel<pos_t><<<nb, 256, 0, st>>>(T, S, s, keys, vals);
    KERNEL_CHECK();
    radix_sort_pairs(keys, kalt, vals, valt, s, 8, 64, ss.sort, st);
    uint32_t ng = regroup(ss, keys, s, st);
    for (int level = 0;; level++) {
        if (ng == s) break;
        CUDA_CHECK(cudaMemsetAsync(ss.counter.get(), 0, 4, st));
        ctx_pivot_kernel<pos_t><<<nb, 256, 0, st>>>(T, n, tau, S, s, keys, vals, ss.gid, ss.gstart, level == 0, kalt,
                                                    ss.counter.get());
        KERNEL_CHECK();
        uint32_t active;
        CUDA_CHECK(cudaMemcpyAsync(&active, ss.counter.get(), 4, cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        if (active == 0) break;
        ss.ctx_levels++;
        // stable LSD: low field first, then the group id; vals ride along
        { uint64_t* t = keys; keys = kalt; kalt = t; }
        radix_sort_pairs(keys, kalt, vals, valt, s, 0, 25, ss.sort, st);
        radix_sort_pairs(keys, kalt, vals, valt, s, 32, 32 + bits_for(ng), ss.sort, st);
        ng = regroup(ss, keys, s, st);
        if (level > 3 * (int)MAXD) throw std::runtime_error("context ranking did not converge");
    }
    scatter_rank_kernel<<<nb, 256, 0, st>>>(vals, ss.gid, s, ss.crank);
    KERNEL_CHECK();

    // ---- 2. run keys ----
    auto new_level = [&]() -> uint32_t* {
        if ((int)ss.rank_lvl.size() <= ss.levels) ss.rank_lvl.emplace_back();
        ss.rank_lvl[ss.levels].reserve(s);
        return ss.rank_lvl[ss.levels++].get();
    };
    uint32_t* rank0 = new_level();
    {
        ss.enc.reserve(s);
        CUDA_CHECK(cudaMemsetAsync(ss.counter.get(), 0, 4, st));
        run_key_kernel<pos_t><<<nb, 256, 0, st>>>(T, n, tau, sss_period_bound(tau, qmax), S, s, ss.enc, ss.counter.get());
        KERNEL_CHECK();
        uint32_t nrun;
        CUDA_CHECK(cudaMemcpyAsync(&nrun, ss.counter.get(), 4, cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        ss.num_run_syms = nrun;
        if (nrun == 0) {
            CUDA_CHECK(cudaMemcpyAsync(rank0, ss.crank.get(), s * 4, cudaMemcpyDeviceToDevice, st));
        } else {
            enc_key_kernel<<<nb, 256, 0, st>>>(ss.enc, s, keys, vals);
            KERNEL_CHECK();
            radix_sort_pairs(keys, kalt, vals, valt, s, 0, 42, ss.sort, st);
            gather_key_kernel<<<nb, 256, 0, st>>>(ss.crank, vals, s, keys);
            KERNEL_CHECK();
            radix_sort_pairs(keys, kalt, vals, valt, s, 0, bits_for(ng), ss.sort, st);
            head_flags2_kernel<<<nb, 256, 0, st>>>(ss.crank, ss.enc, vals, s, ss.flags, ss.gid);
            KERNEL_CHECK();
            exclusive_scan_u32(ss.gid, s, ss.scan, st, ss.counter.get() + 1);
            group_ids_kernel<<<nb, 256, 0, st>>>(ss.flags, ss.gid, s, ss.gstart, ss.counter.get() + 1);
            KERNEL_CHECK();
            CUDA_CHECK(cudaMemcpyAsync(&ng, ss.counter.get() + 1, 4, cudaMemcpyDeviceToHost, st));
            CUDA_CHECK(cudaStreamSynchronize(st));
            scatter_rank_kernel<<<nb, 256, 0, st>>>(vals, ss.gid, s, rank0);
            KERNEL_CHECK();
        }
    }

    // ---- 3. prefix doubling over the reduced string ----
    uint32_t* rank = rank0;
    for (uint64_t h = 1; ng < s; h *= 2) {
        if (ss.levels >= 39) throw std::runtime_error("prefix doubling did not converge");
        doubling_key_kernel<<<nb, 256, 0, st>>>(rank, s, h, keys, vals);
        KERNEL_CHECK();
        int rb = bits_for(ng);
        radix_sort_pairs(keys, kalt, vals, valt, s, 0, rb, ss.sort, st);
        radix_sort_pairs(keys, kalt, vals, valt, s, 32, 32 + rb, ss.sort, st);
        ng = regroup(ss, keys, s, st);
        uint32_t* nr = new_level();
        scatter_rank_kernel<<<nb, 256, 0, st>>>(vals, ss.gid, s, nr);
        KERNEL_CHECK();
        rank = nr;
        ss.doubling_rounds++;
    }
    CUDA_CHECK(cudaMemcpyAsync(ss.ISA.get(), rank, s * 4, cudaMemcpyDeviceToDevice, st));
    inverse_kernel<<<nb, 256, 0, st>>>(ss.ISA, s, ss.SA);
    KERNEL_CHECK();

    // ---- 4. LCP in text characters ----
    level_ptrs lv;
    for (int l = 0; l < ss.levels; l++) lv.r[l] = ss.rank_lvl[l].get();
    lcp_kernel<pos_t><<<nbw, 256, 0, st>>>(T, n, S, s, ss.SA, lv, ss.levels, ss.LCP);
    KERNEL_CHECK();
}

template void suffix_sort_sync<uint32_t>(const uint8_t*, uint64_t, uint32_t, uint32_t, const uint32_t*, uint64_t,
                                         suffix_state<uint32_t>&, cudaStream_t);
template void suffix_sort_sync<uint64_t>(const uint8_t*, uint64_t, uint32_t, uint32_t, const uint64_t*, uint64_t,
                                         suffix_state<uint64_t>&, cudaStream_t);

}  // namespace lz77sss
