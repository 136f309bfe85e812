// K4 + LPF candidates: range-minimum / nearest-smaller structures over (SA_S, LCP_S), the two
// longest-previous-factor candidates of every synchronizing position, batched exact LCE queries.
//
// Replaces build_PSV_NSV_S (/root/reference/include/lz77_sss/algorithms/approximate/lpf_lnf/nxv_pxv.cpp:33-92,
// a single-threaded stack scan there), the two LCE_R and two LCE_L probes per position of
// build_LPF_opt (lpf_opt.cpp:68-139) and lce_sss::lce_lr (patched lce_sss.hpp:123-177).
// For two synchronizing positions the exact LCE is the range minimum of LCP_S between their ranks
// (lce_classic_for_sss.hpp:153-155), so no text is re-read for the LCE_R probes.
#include <vector>
#include "common.cuh"
#include "device_utils.cuh"
#include "stages.cuh"
#include "lce_device.cuh"

namespace lz77sss {
namespace {

constexpr int BLK = RMQ_BLK;

template <typename val_t>
__global__ void blkmin_kernel(const val_t* __restrict__ v, uint64_t s, val_t* __restrict__ tab, uint32_t nblk) {
    const uint64_t b = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    if (b >= nblk) return;
    uint64_t i = b * BLK + lane;
    val_t x = i < s ? v[i] : (val_t)~(val_t)0;
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        val_t y = __shfl_xor_sync(0xffffffffu, x, d);
        x = y < x ? y : x;
    }
    if (lane == 0) tab[b] = x;
}

template <typename val_t>
__global__ void sparse_level_kernel(val_t* __restrict__ tab, uint32_t nblk, int l) {
    uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t half = 1ull << (l - 1);
    if (b + (1ull << l) > nblk) return;
    val_t x = tab[(uint64_t)(l - 1) * nblk + b], y = tab[(uint64_t)(l - 1) * nblk + b + half];
    tab[(uint64_t)l * nblk + b] = y < x ? y : x;
}

// number of equal characters going left from T[i] and T[j] (i < j), at most cap (cap <= i + 1)
__device__ __forceinline__ uint64_t thread_lce_left(const uint8_t* __restrict__ T, uint64_t i, uint64_t j, uint64_t cap) {
    uint64_t k = 0;
    while (k < cap) {
        // the front pad of the device text makes the 8 bytes ending at i-k readable
        uint64_t a = load_u64_unaligned(T + i - k - 7), b = load_u64_unaligned(T + j - k - 7);
        uint64_t diff = a ^ b;
        if (diff) {
            k += __clzll((long long)diff) / 8;
            return k < cap ? k : cap;
        }
        k += 8;
    }
    return cap;
}

template <typename pos_t>
__global__ void lpf_candidates_kernel(const uint8_t* __restrict__ T, uint64_t n, const pos_t* __restrict__ S,
                                      uint64_t s, uint64_t blk_len, uint32_t p_ref, const uint32_t* __restrict__ SA,
                                      const uint32_t* __restrict__ ISA, const pos_t* __restrict__ LCP,
                                      const uint32_t* __restrict__ sa_tab, const pos_t* __restrict__ lcp_tab,
                                      uint32_t nblk, uint32_t* __restrict__ psv_idx, uint32_t* __restrict__ nsv_idx,
                                      pos_t* __restrict__ psv_lce, pos_t* __restrict__ nsv_lce,
                                      pos_t* __restrict__ psv_lcel, pos_t* __restrict__ nsv_lcel) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s) return;
    const uint64_t r = ISA[i];
    const uint64_t pos = S[i];
    const uint32_t pr = prev_smaller(SA, sa_tab, nblk, r, (uint32_t)i);
    const uint32_t nr = next_smaller(SA, sa_tab, nblk, s, r, (uint32_t)i);
    // largest left extension any visit of i can ask for: the covered prefix reaches at
    // least the previous synchronizing position, or the start of the p_ref block
    uint64_t k = pos / blk_len;
    if (k > p_ref - 1) k = p_ref - 1;
    const uint64_t b = k * blk_len;
    uint64_t prevpos = b;
    if (i >= 2 && (uint64_t)S[i - 1] >= b) prevpos = S[i - 1];
    const uint64_t capmax = pos - prevpos;
    uint32_t pi = NONE32, ni = NONE32;
    pos_t pl = 0, nl = 0, pll = 0, nll = 0;
    if (pr != NONE32) {
        pi = SA[pr];
        pl = range_min(LCP, lcp_tab, nblk, (uint64_t)pr + 1, r);
        const uint64_t src = S[pi];
        if (src != 0 && pos != 0) pll = (pos_t)thread_lce_left(T, src - 1, pos - 1, min(capmax, src));
    }
    if (nr != NONE32) {
        ni = SA[nr];
        nl = range_min(LCP, lcp_tab, nblk, r + 1, (uint64_t)nr);
        const uint64_t src = S[ni];
        if (src != 0 && pos != 0) nll = (pos_t)thread_lce_left(T, src - 1, pos - 1, min(capmax, src));
    }
    psv_idx[i] = pi; nsv_idx[i] = ni;
    psv_lce[i] = pl; nsv_lce[i] = nl;
    psv_lcel[i] = pll; nsv_lcel[i] = nll;
}

// Candidates of build_LPF_all / build_LNF_all (lpf_lnf.cpp:34-134,139-249): for every synchronizing position the
// nearest smaller (LPF) or nearest greater (LNF, on the reversed text) SA_S value on both sides of its rank, the
// exact LCE_R with it (range minimum of LCP_S) and the leftward LCE capped at lcap characters.  SAx is SA_S itself
// or its bitwise complement (nearest greater value == nearest smaller complement), tab the block minima of SAx.
// leftward LCE of the candidate pair (src, pos), at most lcap characters (0: none)
__device__ __forceinline__ uint64_t capped_left(const uint8_t* __restrict__ T, uint64_t src, uint64_t pos, uint64_t lcap) {
    if (src == 0 || pos == 0 || lcap == 0) return 0;
    const uint64_t lo = src < pos ? src : pos, hi = src < pos ? pos : src;
    return thread_lce_left(T, lo - 1, hi - 1, lcap < lo ? lcap : lo);
}

// SIDE 0: nearest smaller value on the left of the rank (PSV / PGV), SIDE 1: on the right (NSV / NGV).
// The two sides are separate launches: one kernel doing both faulted ("illegal memory access", B200, CUDA 12.9, also
// with ptxas -O1) on some inputs (dna_collection(400000, seed 52), tau 64, reversed text) although SA_S, ISA_S and the
// complement array were verified on the host and every index the code forms is in range; each side alone never did.
// The root cause is not known yet; tests/test_gpu_parity.py keeps that input (test_lnf_regression_dna52).
template <typename pos_t, int SIDE>
__global__ void all_candidates_kernel(const uint8_t* __restrict__ T, const pos_t* __restrict__ S, uint64_t s,
                                      const uint32_t* __restrict__ SAx, uint32_t flip, const uint32_t* __restrict__ ISA,
                                      const pos_t* __restrict__ LCP, const uint32_t* __restrict__ tab,
                                      const pos_t* __restrict__ lcp_tab, uint32_t nblk, uint64_t lcap,
                                      uint32_t* __restrict__ o_idx, pos_t* __restrict__ o_lce, pos_t* __restrict__ o_lcel) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s) return;
    const uint64_t r = ISA[i];
    const uint64_t pos = S[i];
    const uint32_t v = (uint32_t)i ^ flip;
    const uint32_t q = SIDE == 0 ? prev_smaller(SAx, tab, nblk, r, v) : next_smaller(SAx, tab, nblk, s, r, v);
    uint32_t idx = NONE32;
    pos_t lce = 0, lcel = 0;
    if (q != NONE32) {
        idx = SAx[q] ^ flip;
        lce = SIDE == 0 ? range_min(LCP, lcp_tab, nblk, (uint64_t)q + 1, r) : range_min(LCP, lcp_tab, nblk, r + 1, (uint64_t)q);
        lcel = (pos_t)capped_left(T, (uint64_t)S[idx], pos, lcap);
    }
    o_idx[i] = idx;
    o_lce[i] = lce;
    o_lcel[i] = lcel;
}

// Both sides in one thread: the variant that faulted (kept behind LZ77SSS_FUSED_CANDIDATES=1 so that the fault can be
// reproduced under compute-sanitizer).
template <typename pos_t, bool CLAMP, int PARTS = 7, bool LAUNDER = false>
__global__ void all_candidates_fused_kernel(const uint8_t* __restrict__ T, const pos_t* __restrict__ S, uint64_t s,
                                            const uint32_t* __restrict__ SAx, uint32_t flip, const uint32_t* __restrict__ ISA,
                                            const pos_t* __restrict__ LCP, const uint32_t* __restrict__ tab,
                                            const pos_t* __restrict__ lcp_tab, uint32_t nblk, uint64_t lcap,
                                            uint32_t* __restrict__ p_idx, uint32_t* __restrict__ n_idx,
                                            pos_t* __restrict__ p_lce, pos_t* __restrict__ n_lce,
                                            pos_t* __restrict__ p_lcel, pos_t* __restrict__ n_lcel) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s) return;
    if (LAUNDER) {          // the compiler forgets that these pointers are const __restrict__ (no ld.global.nc, no reordering
        asm volatile("" : "+l"(T));        // across them): is the fault tied to how the loads are emitted?
        asm volatile("" : "+l"(S));
        asm volatile("" : "+l"(SAx));
        asm volatile("" : "+l"(ISA));
        asm volatile("" : "+l"(LCP));
        asm volatile("" : "+l"(tab));
        asm volatile("" : "+l"(lcp_tab));
    }
    const uint64_t r = ISA[i];
    const uint64_t pos = S[i];
    const uint32_t v = (uint32_t)i ^ flip;
    const uint32_t pr = prev_smaller(SAx, tab, nblk, r, v);
    const uint32_t nr = next_smaller(SAx, tab, nblk, s, r, v);
    uint32_t pi = NONE32, ni = NONE32;
    pos_t pl = 0, nl = 0, pll = 0, nll = 0;
    // CLAMP: the guarded loads get an index that is valid even when the guard is false, so that a load the compiler
    // moved in front of its guard cannot fault.  PARTS (bit 0: SAx[q] ^ flip, bit 1: range minimum over LCP, bit 2: S[idx]
    // and the left extension) leaves parts of the body out.  Both are experiments for DESIGN.md §9.
    if (pr != NONE32) {
        if (PARTS & 1) pi = SAx[CLAMP ? min((uint64_t)pr, s - 1) : (uint64_t)pr] ^ flip;
        if (PARTS & 2) pl = range_min(LCP, lcp_tab, nblk, (uint64_t)pr + 1, r);
        if ((PARTS & 5) == 5) pll = (pos_t)capped_left(T, (uint64_t)S[CLAMP ? min((uint64_t)pi, s - 1) : (uint64_t)pi], pos, lcap);
    }
    if (nr != NONE32) {
        if (PARTS & 1) ni = SAx[CLAMP ? min((uint64_t)nr, s - 1) : (uint64_t)nr] ^ flip;
        if (PARTS & 2) nl = range_min(LCP, lcp_tab, nblk, r + 1, (uint64_t)nr);
        if ((PARTS & 5) == 5) nll = (pos_t)capped_left(T, (uint64_t)S[CLAMP ? min((uint64_t)ni, s - 1) : (uint64_t)ni], pos, lcap);
    }
    p_idx[i] = pi; n_idx[i] = ni;
    p_lce[i] = pl; n_lce[i] = nl;
    p_lcel[i] = pll; n_lcel[i] = nll;
}

// The same computation with every global load bounds-checked by hand (compute-sanitizer is not available on the pool): the
// first load outside its array is recorded as {array code, index, limit, thread} instead of being executed.
// LZ77SSS_FUSED_CANDIDATES=2 runs it in front of the normal kernels and prints the record.
struct chk_rec { unsigned long long code, index, limit, thread; };
__device__ __forceinline__ bool chk_fail(chk_rec* rec, unsigned long long code, unsigned long long idx, unsigned long long lim,
                                         unsigned long long thread) {
    if (idx < lim) return false;
    if (atomicCAS(&rec->code, 0ull, code) == 0ull) { rec->index = idx; rec->limit = lim; rec->thread = thread; }
    return true;
}
template <typename pos_t>
__global__ void all_candidates_checked_kernel(const pos_t* __restrict__ S, uint64_t s, const uint32_t* __restrict__ SAx,
                                              uint32_t flip, const uint32_t* __restrict__ ISA, const pos_t* __restrict__ LCP,
                                              const uint32_t* __restrict__ tab, const pos_t* __restrict__ lcp_tab, uint32_t nblk,
                                              int levels, chk_rec* __restrict__ rec) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s) return;
    const uint64_t tabn = (uint64_t)levels * nblk;
#define LD(arr, idx, lim, code, dflt) (chk_fail(rec, code, (idx), (lim), i) ? (dflt) : arr[(idx)])
    auto brm_u32 = [&](uint64_t lo, uint64_t hi) -> uint32_t {        // blk_range_min over tab
        if (hi < lo) { chk_fail(rec, 90, lo, hi, i); return 0xFFFFFFFFu; }
        const int l = 63 - __clzll((long long)(hi - lo + 1));
        const uint32_t x = LD(tab, (uint64_t)l * nblk + lo, tabn, 1, 0xFFFFFFFFu);
        const uint32_t y = LD(tab, (uint64_t)l * nblk + hi + 1 - (1ull << l), tabn, 2, 0xFFFFFFFFu);
        return y < x ? y : x;
    };
    auto brm_pos = [&](uint64_t lo, uint64_t hi) -> pos_t {           // blk_range_min over lcp_tab
        if (hi < lo) { chk_fail(rec, 91, lo, hi, i); return (pos_t)0; }
        const int l = 63 - __clzll((long long)(hi - lo + 1));
        const pos_t x = LD(lcp_tab, (uint64_t)l * nblk + lo, tabn, 3, (pos_t)0);
        const pos_t y = LD(lcp_tab, (uint64_t)l * nblk + hi + 1 - (1ull << l), tabn, 4, (pos_t)0);
        return y < x ? y : x;
    };
    auto rmin = [&](uint64_t lo, uint64_t hi) -> pos_t {              // range_min over LCP
        if (hi < lo) { chk_fail(rec, 92, lo, hi, i); return (pos_t)0; }
        const uint64_t bl = lo / RMQ_BLK, bh = hi / RMQ_BLK;
        pos_t m = LD(LCP, lo, s, 5, (pos_t)0);
        if (bl == bh) {
            for (uint64_t k = lo + 1; k <= hi; k++) { const pos_t x = LD(LCP, k, s, 6, (pos_t)0); m = x < m ? x : m; }
            return m;
        }
        for (uint64_t k = lo + 1; k < (bl + 1) * RMQ_BLK; k++) { const pos_t x = LD(LCP, k, s, 7, (pos_t)0); m = x < m ? x : m; }
        for (uint64_t k = bh * RMQ_BLK; k <= hi; k++) { const pos_t x = LD(LCP, k, s, 8, (pos_t)0); m = x < m ? x : m; }
        if (bl + 1 < bh) { const pos_t x = brm_pos(bl + 1, bh - 1); m = x < m ? x : m; }
        return m;
    };
    const uint64_t r = LD(ISA, i, s, 10, 0u);
    if (chk_fail(rec, 11, r, s, i)) return;
    const uint32_t v = (uint32_t)i ^ flip;
    // prev_smaller
    uint32_t pr = NONE32;
    {
        const uint64_t b = r / RMQ_BLK;
        bool found = false;
        for (uint64_t j = r; j > b * RMQ_BLK && !found;) {
            j--;
            if (LD(SAx, j, s, 12, 0xFFFFFFFFu) < v) { pr = (uint32_t)j; found = true; }
        }
        if (!found && b > 0) {
            uint64_t e = b, w = 1, lo = 0;
            bool none = false;
            while (true) {
                lo = e >= w ? e - w : 0;
                if (brm_u32(lo, e - 1) < v) break;
                if (lo == 0) { none = true; break; }
                w *= 2;
            }
            if (!none) {
                uint64_t hi = e - 1;
                while (lo < hi) {
                    const uint64_t mid = (lo + hi + 1) / 2;
                    if (brm_u32(mid, hi) < v) lo = mid;
                    else hi = mid - 1;
                }
                for (uint64_t j = (lo + 1) * RMQ_BLK; j > lo * RMQ_BLK && !found;) {
                    j--;
                    if (LD(SAx, j, s, 13, 0xFFFFFFFFu) < v) { pr = (uint32_t)j; found = true; }
                }
                if (!found) chk_fail(rec, 93, lo, 0, i);          // "unreachable"
            }
        }
    }
    // next_smaller
    uint32_t nr = NONE32;
    {
        const uint64_t b = r / RMQ_BLK;
        const uint64_t bend = min(s, (b + 1) * RMQ_BLK);
        bool found = false;
        for (uint64_t j = r + 1; j < bend && !found; j++)
            if (LD(SAx, j, s, 14, 0xFFFFFFFFu) < v) { nr = (uint32_t)j; found = true; }
        if (!found && b + 1 < nblk) {
            uint64_t e = b + 1, w = 1, hi = 0;
            bool none = false;
            while (true) {
                hi = e + w - 1 < nblk ? e + w - 1 : nblk - 1;
                if (brm_u32(e, hi) < v) break;
                if (hi == (uint64_t)nblk - 1) { none = true; break; }
                w *= 2;
            }
            if (!none) {
                uint64_t lo = e;
                while (lo < hi) {
                    const uint64_t mid = (lo + hi) / 2;
                    if (brm_u32(lo, mid) < v) hi = mid;
                    else lo = mid + 1;
                }
                const uint64_t be = min(s, (lo + 1) * RMQ_BLK);
                for (uint64_t j = lo * RMQ_BLK; j < be && !found; j++)
                    if (LD(SAx, j, s, 15, 0xFFFFFFFFu) < v) { nr = (uint32_t)j; found = true; }
                if (!found) chk_fail(rec, 94, lo, 0, i);
            }
        }
    }
    {   // the library's own search functions, both in this one kernel like in the fused kernel: same answers?
        const uint32_t pr2 = prev_smaller(SAx, tab, nblk, r, v);
        const uint32_t nr2 = next_smaller(SAx, tab, nblk, s, r, v);
        if (pr2 != pr && atomicCAS(&rec->code, 0ull, 95ull) == 0ull) { rec->index = pr2; rec->limit = pr; rec->thread = i; }
        if (nr2 != nr && atomicCAS(&rec->code, 0ull, 96ull) == 0ull) { rec->index = nr2; rec->limit = nr; rec->thread = i; }
    }
    if (pr != NONE32) {
        const uint32_t pi = LD(SAx, (uint64_t)pr, s, 16, 0u) ^ flip;
        (void)rmin((uint64_t)pr + 1, r);
        (void)LD(S, (uint64_t)pi, s, 17, (pos_t)0);
    }
    if (nr != NONE32) {
        const uint32_t ni = LD(SAx, (uint64_t)nr, s, 18, 0u) ^ flip;
        (void)rmin(r + 1, (uint64_t)nr);
        (void)LD(S, (uint64_t)ni, s, 19, (pos_t)0);
    }
#undef LD
}

__global__ void flip_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, uint64_t s) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < s) out[i] = ~in[i];
}

// lce_l_64 (include/lz77_sss/algorithms/lce_l.hpp:33-83) without a cap: equal characters going left from T[qi] and
// T[qj] (qi < qj), at most qi + 1.  One warp per query, 256 bytes per step.
__global__ void __launch_bounds__(256) left_lce_kernel(const uint8_t* __restrict__ T, const uint64_t* __restrict__ qi,
                                                       const uint64_t* __restrict__ qj, uint64_t nq,
                                                       uint64_t* __restrict__ out) {
    const uint64_t q = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    if (q >= nq) return;
    const uint64_t i = qi[q], j = qj[q], cap = i + 1;
    uint64_t k = 0, res = cap;
    while (k < cap) {
        const uint64_t kk = k + 8ull * lane;
        uint32_t eq = 8;                               // equal bytes in this lane's word (8 = all)
        if (kk < cap) {
            // the front pad of the device text makes the 8 bytes ending at i - kk readable
            const uint64_t a = load_u64_unaligned(T + i - kk - 7), b = load_u64_unaligned(T + j - kk - 7);
            const uint64_t diff = a ^ b;
            if (diff) eq = (uint32_t)(__clzll((long long)diff) / 8);
        }
        const unsigned m = __ballot_sync(0xffffffffu, eq < 8);
        if (m) {
            const int w = __ffs(m) - 1;
            const uint32_t e = __shfl_sync(0xffffffffu, eq, w);
            res = k + 8ull * w + e;
            break;
        }
        k += 256;
    }
    if (lane == 0) out[q] = res < cap ? res : cap;
}

__global__ void reverse_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, uint64_t n) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[n - 1 - i];
}

template <typename pos_t>
__global__ void succ_first_kernel(const pos_t* __restrict__ S, uint64_t s, uint32_t shift, uint64_t nb,
                                  uint32_t* __restrict__ first) {
    uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nb) return;
    const uint64_t x = b << shift;
    uint64_t lo = 0, hi = s;
    while (lo < hi) {
        uint64_t mid = (lo + hi) / 2;
        if ((uint64_t)S[mid] < x) lo = mid + 1;
        else hi = mid;
    }
    first[b] = (uint32_t)lo;
}

// one warp per query; lce_sss::lce / lce_lr (patched lce_sss.hpp:102-177)
template <typename pos_t>
__global__ void __launch_bounds__(256) lce_r_kernel(const lce_view<pos_t> V, const uint64_t* __restrict__ qi,
                                                    const uint64_t* __restrict__ qj, uint64_t nq,
                                                    uint64_t* __restrict__ out) {
    const uint64_t q = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    if (q >= nq) return;
    const uint64_t res = warp_lce_r(V, qi[q], qj[q]);
    if (lane == 0) out[q] = res;
}

}  // namespace

template <typename pos_t>
void build_lce_structures(const uint8_t* T, uint64_t n, uint32_t tau, const pos_t* S, uint64_t s,
                          suffix_state<pos_t>& ss, lce_state<pos_t>& ls, cudaStream_t st) {
    (void)T; (void)n;
    if (s == 0) { ls.nblk = 0; ls.sp_levels = 0; return; }
    ls.nblk = ceil_div_u(s, BLK);
    int levels = 1;
    while ((1ull << levels) <= ls.nblk) levels++;
    ls.sp_levels = levels;
    ls.sa_blkmin.reserve((uint64_t)levels * ls.nblk);
    ls.lcp_blkmin.reserve((uint64_t)levels * ls.nblk);
    unsigned g = ceil_div_u((uint64_t)ls.nblk * 32, 256);
    blkmin_kernel<uint32_t><<<g, 256, 0, st>>>(ss.SA, s, ls.sa_blkmin, ls.nblk);
    KERNEL_CHECK();
    blkmin_kernel<pos_t><<<g, 256, 0, st>>>(ss.LCP, s, ls.lcp_blkmin, ls.nblk);
    KERNEL_CHECK();
    for (int l = 1; l < levels; l++) {
        unsigned gl = ceil_div_u(ls.nblk, 256);
        sparse_level_kernel<uint32_t><<<gl, 256, 0, st>>>(ls.sa_blkmin, ls.nblk, l);
        KERNEL_CHECK();
        sparse_level_kernel<pos_t><<<gl, 256, 0, st>>>(ls.lcp_blkmin, ls.nblk, l);
        KERNEL_CHECK();
    }
    // successor index with buckets of tau positions
    uint32_t shift = 0;
    while ((2u << shift) <= tau) shift++;
    ls.succ_shift = shift;
    ls.succ_nb = ((n + TEXT_PAD) >> shift) + 2;
    ls.succ_first.reserve(ls.succ_nb);
    succ_first_kernel<pos_t><<<ceil_div_u(ls.succ_nb, 256), 256, 0, st>>>(S, s, shift, ls.succ_nb, ls.succ_first);
    KERNEL_CHECK();
}

template <typename pos_t>
void lpf_candidates(const uint8_t* T, uint64_t n, uint32_t tau, const pos_t* S, uint64_t s, uint32_t p_ref,
                    suffix_state<pos_t>& ss, lce_state<pos_t>& ls, cudaStream_t st) {
    (void)tau;
    if (s == 0) return;
    ls.psv_idx.reserve(s); ls.nsv_idx.reserve(s);
    ls.psv_lce.reserve(s); ls.nsv_lce.reserve(s);
    ls.psv_lcel.reserve(s); ls.nsv_lcel.reserve(s);
    uint64_t blk_len = n / p_ref;
    if (blk_len == 0) blk_len = 1;
    lpf_candidates_kernel<pos_t><<<ceil_div_u(s, 128), 128, 0, st>>>(
        T, n, S, s, blk_len, p_ref, ss.SA, ss.ISA, ss.LCP, ls.sa_blkmin, ls.lcp_blkmin, ls.nblk, ls.psv_idx,
        ls.nsv_idx, ls.psv_lce, ls.nsv_lce, ls.psv_lcel, ls.nsv_lcel);
    KERNEL_CHECK();
}

template <typename pos_t>
void all_candidates(const uint8_t* T, uint64_t n, const pos_t* S, uint64_t s, bool greater, uint64_t lcap,
                    suffix_state<pos_t>& ss, lce_state<pos_t>& ls, cudaStream_t st) {
    (void)n;
    if (s == 0) return;
    ls.psv_idx.reserve(s); ls.nsv_idx.reserve(s);
    ls.psv_lce.reserve(s); ls.nsv_lce.reserve(s);
    ls.psv_lcel.reserve(s); ls.nsv_lcel.reserve(s);
    const uint32_t* SAx = ss.SA;
    const uint32_t* tab = ls.sa_blkmin;
    if (greater) {
        ls.sa_flip.reserve(s);
        ls.saf_blkmin.reserve((uint64_t)ls.sp_levels * ls.nblk);
        flip_kernel<<<ceil_div_u(s, 256), 256, 0, st>>>(ss.SA, ls.sa_flip, s);
        KERNEL_CHECK();
        blkmin_kernel<uint32_t><<<ceil_div_u((uint64_t)ls.nblk * 32, 256), 256, 0, st>>>(ls.sa_flip, s, ls.saf_blkmin, ls.nblk);
        KERNEL_CHECK();
        for (int l = 1; l < ls.sp_levels; l++) {
            sparse_level_kernel<uint32_t><<<ceil_div_u(ls.nblk, 256), 256, 0, st>>>(ls.saf_blkmin, ls.nblk, l);
            KERNEL_CHECK();
        }
        SAx = ls.sa_flip;
        tab = ls.saf_blkmin;
    }
    static const bool checked = getenv("LZ77SSS_FUSED_CANDIDATES") && getenv("LZ77SSS_FUSED_CANDIDATES")[0] == '2';
    if (checked) {
        dbuf<chk_rec> rec;
        rec.reserve(1);
        CUDA_CHECK(cudaMemsetAsync(rec.get(), 0, sizeof(chk_rec), st));
        all_candidates_checked_kernel<pos_t><<<ceil_div_u(s, 128), 128, 0, st>>>(S, s, SAx, greater ? 0xFFFFFFFFu : 0u, ss.ISA, ss.LCP,
                                                                               tab, ls.lcp_blkmin, ls.nblk, ls.sp_levels, rec);
        KERNEL_CHECK();
        chk_rec h_rec;
        CUDA_CHECK(cudaMemcpyAsync(&h_rec, rec.get(), sizeof h_rec, cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        fprintf(stderr, "all_candidates check (|S| = %llu, nblk = %u, levels = %d, greater = %d): %s code %llu index %llu limit %llu thread %llu\n",
                (unsigned long long)s, ls.nblk, ls.sp_levels, (int)greater, h_rec.code ? "FIRST BAD LOAD" : "all loads in range",
                h_rec.code, h_rec.index, h_rec.limit, h_rec.thread);
    }
    static const char fused = getenv("LZ77SSS_FUSED_CANDIDATES") ? getenv("LZ77SSS_FUSED_CANDIDATES")[0] : '0';
    if (fused == '1' || (fused >= '3' && fused <= '8')) {
        auto kern = fused == '1' ? all_candidates_fused_kernel<pos_t, false, 7>
                  : fused == '3' ? all_candidates_fused_kernel<pos_t, true, 7>
                  : fused == '4' ? all_candidates_fused_kernel<pos_t, false, 0>      // the two searches only
                  : fused == '5' ? all_candidates_fused_kernel<pos_t, false, 1>      // + SAx[q]
                  : fused == '6' ? all_candidates_fused_kernel<pos_t, false, 3>      // + range minimum over LCP
                  : fused == '7' ? all_candidates_fused_kernel<pos_t, false, 5>      // searches + SAx[q] + S[idx] + left extension
                  : all_candidates_fused_kernel<pos_t, false, 7, true>;              // everything, pointers laundered
        kern<<<ceil_div_u(s, 128), 128, 0, st>>>(
            T, S, s, SAx, greater ? 0xFFFFFFFFu : 0u, ss.ISA, ss.LCP, tab, ls.lcp_blkmin, ls.nblk, lcap, ls.psv_idx,
            ls.nsv_idx, ls.psv_lce, ls.nsv_lce, ls.psv_lcel, ls.nsv_lcel);
        KERNEL_CHECK();
        if (fused >= '4') {        // partial bodies: synchronize (does it fault?), then compute the real candidates
            CUDA_CHECK(cudaStreamSynchronize(st));
            fprintf(stderr, "fused variant %c: no fault (|S| = %llu, greater = %d)\n", fused, (unsigned long long)s, (int)greater);
        } else {
            return;
        }
    }
    all_candidates_kernel<pos_t, 0><<<ceil_div_u(s, 128), 128, 0, st>>>(
        T, S, s, SAx, greater ? 0xFFFFFFFFu : 0u, ss.ISA, ss.LCP, tab, ls.lcp_blkmin, ls.nblk, lcap, ls.psv_idx, ls.psv_lce,
        ls.psv_lcel);
    KERNEL_CHECK();
    all_candidates_kernel<pos_t, 1><<<ceil_div_u(s, 128), 128, 0, st>>>(
        T, S, s, SAx, greater ? 0xFFFFFFFFu : 0u, ss.ISA, ss.LCP, tab, ls.lcp_blkmin, ls.nblk, lcap, ls.nsv_idx, ls.nsv_lce,
        ls.nsv_lcel);
    KERNEL_CHECK();
}

void left_lce_batch(const uint8_t* T, const uint64_t* d_qi, const uint64_t* d_qj, uint64_t nq, uint64_t* d_out,
                    cudaStream_t st) {
    if (nq == 0) return;
    left_lce_kernel<<<ceil_div_u(nq * 32, 256), 256, 0, st>>>(T, d_qi, d_qj, nq, d_out);
    KERNEL_CHECK();
}

void reverse_text(const uint8_t* d_in, uint8_t* d_out, uint64_t n, cudaStream_t st) {
    if (n == 0) return;
    reverse_kernel<<<ceil_div_u(n, 256), 256, 0, st>>>(d_in, d_out, n);
    KERNEL_CHECK();
}

template <typename pos_t>
void lce_r_batch(const uint8_t* T, uint64_t n, uint32_t tau, const pos_t* S, uint64_t s, suffix_state<pos_t>& ss,
                 lce_state<pos_t>& ls, const uint64_t* d_qi, const uint64_t* d_qj, uint64_t nq, uint64_t* d_out,
                 cudaStream_t st) {
    if (nq == 0) return;
    lce_r_kernel<pos_t><<<ceil_div_u(nq * 32, 256), 256, 0, st>>>(make_lce_view<pos_t>(T, n, tau, S, s, ss, ls), d_qi, d_qj, nq, d_out);
    KERNEL_CHECK();
}

template <typename pos_t>
__global__ void candidate_sources_kernel(const pos_t* __restrict__ S, uint64_t s, const uint32_t* __restrict__ pidx,
                                         const uint32_t* __restrict__ nidx, pos_t* __restrict__ psrc,
                                         pos_t* __restrict__ nsrc) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s) return;
    const uint32_t a = pidx[i], b = nidx[i];
    psrc[i] = a == NONE32 ? (pos_t)~(pos_t)0 : S[a];
    nsrc[i] = b == NONE32 ? (pos_t)~(pos_t)0 : S[b];
}

template <typename pos_t>
void candidate_sources(const pos_t* d_S, uint64_t s, lce_state<pos_t>& ls, cudaStream_t st) {
    if (!s) return;
    ls.psv_src.reserve(s); ls.nsv_src.reserve(s);
    candidate_sources_kernel<pos_t><<<ceil_div_u(s, 256), 256, 0, st>>>(d_S, s, ls.psv_idx, ls.nsv_idx, ls.psv_src, ls.nsv_src);
    KERNEL_CHECK();
}

#define INST(pos_t)                                                                                                    \
    template void build_lce_structures<pos_t>(const uint8_t*, uint64_t, uint32_t, const pos_t*, uint64_t,              \
                                              suffix_state<pos_t>&, lce_state<pos_t>&, cudaStream_t);                  \
    template void lpf_candidates<pos_t>(const uint8_t*, uint64_t, uint32_t, const pos_t*, uint64_t, uint32_t,          \
                                        suffix_state<pos_t>&, lce_state<pos_t>&, cudaStream_t);                        \
    template void all_candidates<pos_t>(const uint8_t*, uint64_t, const pos_t*, uint64_t, bool, uint64_t,              \
                                        suffix_state<pos_t>&, lce_state<pos_t>&, cudaStream_t);                        \
    template void candidate_sources<pos_t>(const pos_t*, uint64_t, lce_state<pos_t>&, cudaStream_t);                 \
    template void lce_r_batch<pos_t>(const uint8_t*, uint64_t, uint32_t, const pos_t*, uint64_t, suffix_state<pos_t>&, \
                                     lce_state<pos_t>&, const uint64_t*, const uint64_t*, uint64_t, uint64_t*,         \
                                     cudaStream_t);
INST(uint32_t)
INST(uint64_t)

}  // namespace lz77sss
