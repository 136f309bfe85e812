// Device-wide primitives used by every stage: exclusive scan and a stable LSD radix sort of
// (64-bit key, 32-bit payload) pairs.  Hand-written (no CUB/thrust): the suffix-sorting stage
//  is on the hot path, so its sort is ours.
#include "common.cuh"

namespace lz77sss {

namespace {

constexpr int SCAN_NT = 256;
constexpr int SCAN_IPT = 8;
constexpr int SCAN_TILE = SCAN_NT * SCAN_IPT;

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
    const unsigned lane = threadIdx.x & 31;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= (unsigned)d) v += y;
    }
    return v;
}

// exclusive scan of one value per thread over a block of SCAN_NT threads; returns exclusive prefix,
// *total = block sum
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* total) {
    __shared__ uint32_t warp_sums[SCAN_NT / 32];
    __shared__ uint32_t s_total;
    const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = warp_incl_scan(v);
    if (lane == 31) warp_sums[w] = inc;
    __syncthreads();
    if (w == 0) {
        uint32_t x = lane < SCAN_NT / 32 ? warp_sums[lane] : 0;
        uint32_t xi = warp_incl_scan(x);
        if (lane < SCAN_NT / 32) warp_sums[lane] = xi - x;
        if (lane == SCAN_NT / 32 - 1) s_total = xi;
    }
    __syncthreads();
    uint32_t r = warp_sums[w] + inc - v;
    *total = s_total;
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(SCAN_NT) scan_reduce_kernel(const uint32_t* __restrict__ in, uint64_t n,
                                                              uint32_t* __restrict__ block_sums) {
    uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE;
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) {
        uint64_t i = base + (uint64_t)k * SCAN_NT + threadIdx.x;
        if (i < n) s += in[i];
    }
    uint32_t tot;
    block_excl_scan(s, &tot);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(SCAN_NT) scan_apply_kernel(uint32_t* __restrict__ d, uint64_t n,
                                                             const uint32_t* __restrict__ block_offs,
                                                             uint32_t* __restrict__ total_out) {
    __shared__ uint32_t tile[SCAN_TILE];
    uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE;
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) {
        uint64_t i = base + (uint64_t)k * SCAN_NT + threadIdx.x;
        tile[k * SCAN_NT + threadIdx.x] = i < n ? d[i] : 0;
    }
    __syncthreads();
    uint32_t loc[SCAN_IPT];
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) {
        loc[k] = s;
        s += tile[threadIdx.x * SCAN_IPT + k];
    }
    uint32_t tot;
    uint32_t ex = block_excl_scan(s, &tot);
    uint32_t off = block_offs ? block_offs[blockIdx.x] : 0;
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) tile[threadIdx.x * SCAN_IPT + k] = loc[k] + ex + off;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) {
        uint64_t i = base + (uint64_t)k * SCAN_NT + threadIdx.x;
        if (i < n) d[i] = tile[k * SCAN_NT + threadIdx.x];
    }
    if (total_out && blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) *total_out = off + tot;
}

void scan_rec(uint32_t* d, uint64_t n, uint32_t* l1, uint32_t* l2, cudaStream_t st, uint32_t* d_total) {
    if (n == 0) {
        if (d_total) CUDA_CHECK(cudaMemsetAsync(d_total, 0, 4, st));
        return;
    }
    unsigned nb = ceil_div_u(n, SCAN_TILE);
    if (nb == 1) {
        scan_apply_kernel<<<1, SCAN_NT, 0, st>>>(d, n, nullptr, d_total);
        KERNEL_CHECK();
        return;
    }
    scan_reduce_kernel<<<nb, SCAN_NT, 0, st>>>(d, n, l1);
    KERNEL_CHECK();
    // scan the block sums (second level uses l2 as its own first-level scratch; a third level
    // would need nb > 2048*2048, i.e. n > 2^33 — out of range for uint32 sums anyway)
    unsigned nb2 = ceil_div_u(nb, SCAN_TILE);
    if (nb2 == 1) {
        scan_apply_kernel<<<1, SCAN_NT, 0, st>>>(l1, nb, nullptr, nullptr);
        KERNEL_CHECK();
    } else {
        scan_reduce_kernel<<<nb2, SCAN_NT, 0, st>>>(l1, nb, l2);
        KERNEL_CHECK();
        if (nb2 > SCAN_TILE) throw std::runtime_error("exclusive_scan_u32: input too large");
        scan_apply_kernel<<<1, SCAN_NT, 0, st>>>(l2, nb2, nullptr, nullptr);
        KERNEL_CHECK();
        scan_apply_kernel<<<nb2, SCAN_NT, 0, st>>>(l1, nb, l2, nullptr);
        KERNEL_CHECK();
    }
    scan_apply_kernel<<<nb, SCAN_NT, 0, st>>>(d, n, l1, d_total);
    KERNEL_CHECK();
}

__global__ void fill_u32_kernel(uint32_t* d, uint64_t n, uint32_t v) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) d[i] = v;
}
__global__ void iota_u32_kernel(uint32_t* d, uint64_t n) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) d[i] = (uint32_t)i;
}

// ---------------- radix sort ----------------------------------------------------------------
// One LSD pass = histogram per tile (hist[tile][digit], rows written and read coalesced), a column scan over the tiles
// (chunk sums, one CTA over the chunks and the digits, chunk apply), scatter.  The scatter ranks its tile
// with per-warp match_any counting (stable), reorders the tile in shared memory by digit and writes every
// digit's run with consecutive threads, so global stores are coalesced runs instead of single scattered words.
// A pass takes a digit of up to 9 bits (NB = 256 or 512 bins): the bits of a sort are split into the fewest passes.
constexpr int RS_NT = 256;
constexpr int RS_NW = RS_NT / 32;
#ifndef RS_IPT32
#define RS_IPT32 16
#endif
#ifndef RS_IPT64
#define RS_IPT64 8
#endif
#ifndef RS_MIN_CTAS
#define RS_MIN_CTAS 4
#endif
template <typename K> struct rs_cfg { static constexpr int IPT = sizeof(K) == 4 ? RS_IPT32 : RS_IPT64; };

template <typename K, int NB>
__global__ void __launch_bounds__(RS_NT) rs_hist_kernel(const K* __restrict__ keys, uint64_t n, int shift,
                                                        uint32_t mask, uint32_t* __restrict__ hist, uint32_t ntiles) {
    constexpr int TILE = RS_NT * rs_cfg<K>::IPT;          // 16 KiB of keys for both key types
    constexpr int VPT = TILE * (int)sizeof(K) / 16 / RS_NT;   // 16-byte vectors per thread
    constexpr int KPV = 16 / (int)sizeof(K);              // keys per vector
    __shared__ uint32_t cnt[NB];
    for (int d = threadIdx.x; d < NB; d += RS_NT) cnt[d] = 0;
    __syncthreads();
    const uint64_t base = (uint64_t)blockIdx.x * TILE;
    if (base + TILE <= n) {                               // full tile: coalesced 16-byte loads, all issued up front
        const uint4* src = reinterpret_cast<const uint4*>(keys + base);
        uint4 v[VPT];
#pragma unroll
        for (int k = 0; k < VPT; k++) v[k] = src[k * RS_NT + threadIdx.x];
#pragma unroll
        for (int k = 0; k < VPT; k++) {
            K kk[KPV];
            memcpy(kk, &v[k], 16);
#pragma unroll
            for (int j = 0; j < KPV; j++) atomicAdd(&cnt[(uint32_t)(kk[j] >> shift) & mask], 1u);
        }
    } else {
#pragma unroll
        for (int k = 0; k < rs_cfg<K>::IPT; k++) {
            uint64_t i = base + (uint64_t)k * RS_NT + threadIdx.x;
            if (i < n) atomicAdd(&cnt[(uint32_t)(keys[i] >> shift) & mask], 1u);
        }
    }
    __syncthreads();
    for (int d = threadIdx.x; d < NB; d += RS_NT) hist[(uint64_t)blockIdx.x * NB + d] = cnt[d];
}

// (Measured and dropped, profiles/r02_scatter_*_experiment.*: other tile sizes and CTAs per SM, 14.2-16.1 ms per GiB for the gap
// index; a loop over tiles with the next tile's keys and payloads in flight, 17.6-24 ms; all MATCH instructions of a thread issued in
// front of the counter updates, 15.3 against 15.1 ms: neither residency, nor the latency of a tile's own loads, nor the
// latency of the matches is what holds the pass at 0.3 of the HBM peak.)
template <typename K, int NB, bool IOTA>
__global__ void __launch_bounds__(RS_NT, RS_MIN_CTAS) rs_scatter_kernel(const K* __restrict__ keys,
                                                           const uint32_t* __restrict__ vals,
                                                           K* __restrict__ okeys, uint32_t* __restrict__ ovals,
                                                           uint64_t n, int shift, uint32_t mask,
                                                           const uint32_t* __restrict__ offs, uint32_t ntiles) {
    constexpr int IPT = rs_cfg<K>::IPT;
    constexpr int TILE = RS_NT * IPT;
    constexpr int DPT = NB / RS_NT;       // digits per thread in the scan of the digit totals
    __shared__ uint16_t cnt[RS_NW][NB];   // a warp holds at most 32 * IPT items
    __shared__ uint16_t dstart[NB];       // tile-local start of every digit
    __shared__ uint32_t goff[NB];         // global offset of the tile's first item of every digit
    __shared__ K skeys[TILE];
    __shared__ uint32_t svals[TILE];
    __shared__ uint32_t wsum[RS_NW];
    const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < RS_NW * NB / 2; i += RS_NT) reinterpret_cast<uint32_t*>(&cnt[0][0])[i] = 0;
    __syncthreads();
    const uint64_t tbase = (uint64_t)blockIdx.x * TILE;
    const uint64_t base = tbase + (uint64_t)w * (32 * IPT);
    const uint32_t in_tile = n - tbase < (uint64_t)TILE ? (uint32_t)(n - tbase) : (uint32_t)TILE;
    const uint32_t in_warp = in_tile > w * (32 * IPT) ? in_tile - w * (32 * IPT) : 0;      // items of this warp
    K key[IPT];
    uint32_t val[IPT];
    uint32_t rank2[IPT / 2];              // rank inside the warp's items of the same digit, two per word
#pragma unroll
    for (int k = 0; k < IPT; k++) {
        const bool valid = (uint32_t)(k * 32) + lane < in_warp;
        key[k] = valid ? keys[base + (uint64_t)k * 32 + lane] : (K)~(K)0;
    }
#pragma unroll
    for (int k = 0; k < IPT; k++) {
        const bool valid = (uint32_t)(k * 32) + lane < in_warp;
        const uint32_t d = valid ? ((uint32_t)(key[k] >> shift) & mask) : (uint32_t)NB;   // NB = "no item"
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        uint32_t old = 0;
        if (valid) old = cnt[w][d];
        __syncwarp();
        if (valid && lane == (unsigned)(__ffs(peers) - 1)) cnt[w][d] = (uint16_t)(old + __popc(peers));
        __syncwarp();
        const uint32_t r = old + __popc(peers & ((1u << lane) - 1));
        if (k & 1) rank2[k / 2] |= r << 16;
        else rank2[k / 2] = r;
    }
#pragma unroll
    for (int k = 0; k < IPT; k++)          // in flight during the scan of the digit totals
        val[k] = (uint32_t)(k * 32) + lane >= in_warp || !ovals ? 0u : IOTA ? (uint32_t)(base + (uint64_t)k * 32 + lane) : vals[base + (uint64_t)k * 32 + lane];
    __syncthreads();
    {   // thread t, digits DPT*t .. DPT*t+DPT-1: per-warp counts -> exclusive offsets over the warps; digit totals
        uint32_t tot[DPT];
        uint32_t acc = 0;
#pragma unroll
        for (int j = 0; j < DPT; j++) {
            const uint32_t d = threadIdx.x * DPT + j;
            uint32_t a = 0;
#pragma unroll
            for (int ww = 0; ww < RS_NW; ww++) {
                const uint32_t t = cnt[ww][d];
                cnt[ww][d] = (uint16_t)a;
                a += t;
            }
            tot[j] = a;
            acc += a;
        }
        // exclusive scan of the totals over the 256 threads (digit order)
        uint32_t inc = acc;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, inc, s);
            if (lane >= (unsigned)s) inc += y;
        }
        if (lane == 31) wsum[w] = inc;
        __syncthreads();
        uint32_t wb = 0;
#pragma unroll
        for (int ww = 0; ww < RS_NW; ww++) wb += ww < (int)w ? wsum[ww] : 0u;
        uint32_t start = wb + inc - acc;
#pragma unroll
        for (int j = 0; j < DPT; j++) {
            const uint32_t d = threadIdx.x * DPT + j;
            dstart[d] = (uint16_t)start;
            goff[d] = offs[(uint64_t)blockIdx.x * NB + d];
            start += tot[j];
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < IPT; k++) {
        if ((uint32_t)(k * 32) + lane < in_warp) {
            const uint32_t d = (uint32_t)(key[k] >> shift) & mask;
            const uint32_t r = (k & 1) ? rank2[k / 2] >> 16 : rank2[k / 2] & 0xFFFFu;
            const uint32_t lp = (uint32_t)dstart[d] + cnt[w][d] + r;
            skeys[lp] = key[k];
            svals[lp] = val[k];
        }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < in_tile; i += RS_NT) {
        const K kk = skeys[i];
        const uint32_t d = (uint32_t)(kk >> shift) & mask;
        const uint32_t pos = goff[d] + (i - dstart[d]);
        okeys[pos] = kk;
        if (ovals) ovals[pos] = svals[i];       // (keys-only sorts pass no payload arrays)
    }
}

}  // namespace

void exclusive_scan_u32(uint32_t* d, uint64_t n, scan_scratch& s, cudaStream_t st, uint32_t* d_total) {
    uint64_t nb = (n + SCAN_TILE - 1) / SCAN_TILE;
    s.lvl1.reserve(nb + 1);
    s.lvl2.reserve((nb + SCAN_TILE - 1) / SCAN_TILE + 1);
    scan_rec(d, n, s.lvl1, s.lvl2, st, d_total);
}

// hist[tile][digit] -> global start of (digit, tile): digits major, tiles minor.  Thread d owns column d.
constexpr int RS_ROWS = 128;             // tiles per chunk
template <int NB>
__global__ void __launch_bounds__(NB) rs_colsum_kernel(const uint32_t* __restrict__ hist, uint32_t ntiles,
                                                      uint32_t* __restrict__ csum) {
    const uint32_t r0 = blockIdx.x * RS_ROWS, r1 = min(r0 + RS_ROWS, ntiles);
    uint32_t acc = 0;
#pragma unroll 8
    for (uint32_t r = r0; r < r1; r++) acc += hist[(uint64_t)r * NB + threadIdx.x];
    csum[(uint64_t)blockIdx.x * NB + threadIdx.x] = acc;
}
template <int NB>
__global__ void __launch_bounds__(NB) rs_colbase_kernel(const uint32_t* __restrict__ csum, uint32_t nchunks,
                                                       uint32_t* __restrict__ cpre) {
    __shared__ uint32_t wsum[NB / 32];
    uint32_t run = 0;
#pragma unroll 16
    for (uint32_t c = 0; c < nchunks; c++) {
        const uint32_t t = csum[(uint64_t)c * NB + threadIdx.x];
        cpre[(uint64_t)c * NB + threadIdx.x] = run;
        run += t;
    }
    // exclusive scan of the digit totals
    const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = run;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, inc, s);
        if (lane >= (unsigned)s) inc += y;
    }
    if (lane == 31) wsum[w] = inc;
    __syncthreads();
    uint32_t wb = 0;
    for (unsigned ww = 0; ww < w; ww++) wb += wsum[ww];
    cpre[(uint64_t)nchunks * NB + threadIdx.x] = wb + inc - run;
}
template <int NB>
__global__ void __launch_bounds__(NB) rs_colapply_kernel(const uint32_t* __restrict__ hist, uint32_t ntiles,
                                                        const uint32_t* __restrict__ cpre, uint32_t nchunks,
                                                        uint32_t* __restrict__ offs) {
    const uint32_t r0 = blockIdx.x * RS_ROWS, r1 = min(r0 + RS_ROWS, ntiles);
    uint32_t run = cpre[(uint64_t)nchunks * NB + threadIdx.x] + cpre[(uint64_t)blockIdx.x * NB + threadIdx.x];
#pragma unroll 16
    for (uint32_t r = r0; r < r1; r++) {
        const uint32_t t = hist[(uint64_t)r * NB + threadIdx.x];
        offs[(uint64_t)r * NB + threadIdx.x] = run;
        run += t;
    }
}

void iota_u32(uint32_t* d, uint64_t n, cudaStream_t st);

template <typename K, int NB>
void radix_pass(K* keys, K* kalt, uint32_t* vals, uint32_t* valt, uint64_t n, int shift, int bits, uint32_t ntiles,
                sort_scratch& s, cudaStream_t st, bool iota) {
    const uint32_t mask = (1u << bits) - 1;
    rs_hist_kernel<K, NB><<<ntiles, RS_NT, 0, st>>>(keys, n, shift, mask, s.hist, ntiles);
    KERNEL_CHECK();
    const uint32_t nchunks = ceil_div_u(ntiles, RS_ROWS);
    s.csum.reserve((uint64_t)nchunks * NB);
    s.cpre.reserve((uint64_t)(nchunks + 1) * NB);
    rs_colsum_kernel<NB><<<nchunks, NB, 0, st>>>(s.hist, ntiles, s.csum);
    KERNEL_CHECK();
    rs_colbase_kernel<NB><<<1, NB, 0, st>>>(s.csum, nchunks, s.cpre);
    KERNEL_CHECK();
    rs_colapply_kernel<NB><<<nchunks, NB, 0, st>>>(s.hist, ntiles, s.cpre, nchunks, s.offs);
    KERNEL_CHECK();
    PROF_BEGIN(st);
    if (iota) rs_scatter_kernel<K, NB, true><<<ntiles, RS_NT, 0, st>>>(keys, vals, kalt, valt, n, shift, mask, s.offs, ntiles);
    else rs_scatter_kernel<K, NB, false><<<ntiles, RS_NT, 0, st>>>(keys, vals, kalt, valt, n, shift, mask, s.offs, ntiles);
    KERNEL_CHECK();
    // algorithmic bytes: keys and payloads read once and written once (generated payloads are not read)
    PROF_END(st, sizeof(K) == 4 ? "rs_scatter_kernel<u32> (radix pass of the gap index)" : "rs_scatter_kernel<u64> (radix pass of the SSS suffix sort)",
             n * (2 * sizeof(K) + (!valt ? 0 : iota ? 4 : 8)));
}

template <typename K>
void radix_sort_pairs_t(K*& keys, K*& kalt, uint32_t*& vals, uint32_t*& valt, uint64_t n, int bit_lo,
                        int bit_hi, sort_scratch& s, cudaStream_t st, bool iota, const K* first_keys = nullptr) {
    if (n <= 1 || bit_hi <= bit_lo) {
        if (iota && vals) iota_u32(vals, n, st);
        if (first_keys && n) CUDA_CHECK(cudaMemcpyAsync(keys, first_keys, n * sizeof(K), cudaMemcpyDeviceToDevice, st));
        return;
    }
    uint32_t ntiles = ceil_div_u(n, RS_NT * rs_cfg<K>::IPT);
    s.hist.reserve((uint64_t)512 * ntiles);
    s.offs.reserve((uint64_t)512 * ntiles);
    // the fewest passes of at most 9 bits, widths as equal as possible (wider digits first)
    const int total = bit_hi - bit_lo, passes = (total + 8) / 9;
    int shift = bit_lo;
    for (int p = 0; p < passes; p++) {
        const int left = bit_hi - shift, bits = (left + (passes - p) - 1) / (passes - p);
        K* in = p == 0 && first_keys ? const_cast<K*>(first_keys) : keys;     // only read
        if (bits > 8) radix_pass<K, 512>(in, kalt, vals, valt, n, shift, bits, ntiles, s, st, iota && p == 0);
        else radix_pass<K, 256>(in, kalt, vals, valt, n, shift, bits, ntiles, s, st, iota && p == 0);
        K* tk = keys; keys = kalt; kalt = tk;
        uint32_t* tv = vals; vals = valt; valt = tv;
        shift += bits;
    }
}

void radix_sort_pairs(uint64_t*& keys, uint64_t*& kalt, uint32_t*& vals, uint32_t*& valt, uint64_t n, int bit_lo,
                      int bit_hi, sort_scratch& s, cudaStream_t st) {
    radix_sort_pairs_t<uint64_t>(keys, kalt, vals, valt, n, bit_lo, bit_hi, s, st, false);
}
void radix_sort_pairs32(uint32_t*& keys, uint32_t*& kalt, uint32_t*& vals, uint32_t*& valt, uint64_t n, int bit_lo,
                        int bit_hi, sort_scratch& s, cudaStream_t st, bool vals_are_iota, const uint32_t* first_keys) {
    radix_sort_pairs_t<uint32_t>(keys, kalt, vals, valt, n, bit_lo, bit_hi, s, st, vals_are_iota, first_keys);
}

void fill_u32(uint32_t* d, uint64_t n, uint32_t v, cudaStream_t st) {
    if (!n) return;
    fill_u32_kernel<<<ceil_div_u(n, 256), 256, 0, st>>>(d, n, v);
    KERNEL_CHECK();
}
void iota_u32(uint32_t* d, uint64_t n, cudaStream_t st) {
    if (!n) return;
    iota_u32_kernel<<<ceil_div_u(n, 256), 256, 0, st>>>(d, n);
    KERNEL_CHECK();
}

}  // namespace lz77sss

namespace lz77sss {
// keys only (the repair lists of the gap index: the entry id is part of the key)
void radix_sort_keys(uint64_t*& keys, uint64_t*& kalt, uint64_t n, int bit_lo, int bit_hi, sort_scratch& s, cudaStream_t st) {
    uint32_t *v = nullptr, *va = nullptr;
    radix_sort_pairs_t<uint64_t>(keys, kalt, v, va, n, bit_lo, bit_hi, s, st, false);
}
}  // namespace lz77sss
