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
// One LSD pass = histogram per tile, device scan of hist[digit][tile], scatter.  The scatter ranks its tile
// with per-warp match_any counting (stable), reorders the tile in shared memory by digit and writes every
// digit's run with consecutive threads, so global stores are coalesced runs instead of single scattered words.
constexpr int RS_NT = 256;
constexpr int RS_NW = RS_NT / 32;
template <typename K> struct rs_cfg { static constexpr int IPT = sizeof(K) == 4 ? 16 : 8; };

template <typename K>
__global__ void __launch_bounds__(RS_NT) rs_hist_kernel(const K* __restrict__ keys, uint64_t n, int shift,
                                                        uint32_t mask, uint32_t* __restrict__ hist, uint32_t ntiles) {
    constexpr int TILE = RS_NT * rs_cfg<K>::IPT;
    __shared__ uint32_t cnt[256];
    cnt[threadIdx.x] = 0;
    __syncthreads();
    uint64_t base = (uint64_t)blockIdx.x * TILE;
#pragma unroll
    for (int k = 0; k < rs_cfg<K>::IPT; k++) {
        uint64_t i = base + (uint64_t)k * RS_NT + threadIdx.x;
        if (i < n) atomicAdd(&cnt[(uint32_t)(keys[i] >> shift) & mask], 1u);
    }
    __syncthreads();
    hist[(uint64_t)threadIdx.x * ntiles + blockIdx.x] = cnt[threadIdx.x];
}

template <typename K>
__global__ void __launch_bounds__(RS_NT) rs_scatter_kernel(const K* __restrict__ keys,
                                                           const uint32_t* __restrict__ vals,
                                                           K* __restrict__ okeys, uint32_t* __restrict__ ovals,
                                                           uint64_t n, int shift, uint32_t mask,
                                                           const uint32_t* __restrict__ offs, uint32_t ntiles) {
    constexpr int IPT = rs_cfg<K>::IPT;
    constexpr int TILE = RS_NT * IPT;
    __shared__ uint32_t cnt[RS_NW][256];
    __shared__ uint32_t dstart[256];      // tile-local start of every digit
    __shared__ uint32_t goff[256];        // global offset of the tile's first item of every digit
    __shared__ K skeys[TILE];
    __shared__ uint32_t svals[TILE];
    __shared__ uint32_t wsum[RS_NW];
    const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < RS_NW * 256; i += RS_NT) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const uint64_t tbase = (uint64_t)blockIdx.x * TILE;
    const uint64_t base = tbase + (uint64_t)w * (32 * IPT);
    const uint32_t in_tile = n - tbase < (uint64_t)TILE ? (uint32_t)(n - tbase) : (uint32_t)TILE;
    K key[IPT];
    uint32_t val[IPT], rank[IPT], dig[IPT];
#pragma unroll
    for (int k = 0; k < IPT; k++) {
        uint64_t i = base + (uint64_t)k * 32 + lane;
        bool valid = i < n;
        key[k] = valid ? keys[i] : (K)~(K)0;
        val[k] = valid ? vals[i] : 0;
        uint32_t d = valid ? ((uint32_t)(key[k] >> shift) & mask) : 256u;   // 256 = "no item"
        dig[k] = d;
        unsigned peers = __match_any_sync(0xffffffffu, d);
        uint32_t old = 0;
        if (valid) old = cnt[w][d];
        __syncwarp();
        if (valid && lane == (unsigned)(__ffs(peers) - 1)) cnt[w][d] = old + __popc(peers);
        __syncwarp();
        rank[k] = old + __popc(peers & ((1u << lane) - 1));
    }
    __syncthreads();
    {   // thread d: per-warp counts of digit d -> exclusive offsets over the warps; digit total
        const uint32_t d = threadIdx.x;
        uint32_t acc = 0;
#pragma unroll
        for (int ww = 0; ww < RS_NW; ww++) {
            uint32_t t = cnt[ww][d];
            cnt[ww][d] = acc;
            acc += t;
        }
        // exclusive scan of the digit totals over the 256 threads
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
        dstart[d] = wb + inc - acc;
        goff[d] = offs[(uint64_t)d * ntiles + blockIdx.x];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < IPT; k++) {
        if (dig[k] < 256u) {
            const uint32_t lp = dstart[dig[k]] + cnt[w][dig[k]] + rank[k];
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
        ovals[pos] = svals[i];
    }
}

}  // namespace

void exclusive_scan_u32(uint32_t* d, uint64_t n, scan_scratch& s, cudaStream_t st, uint32_t* d_total) {
    uint64_t nb = (n + SCAN_TILE - 1) / SCAN_TILE;
    s.lvl1.reserve(nb + 1);
    s.lvl2.reserve((nb + SCAN_TILE - 1) / SCAN_TILE + 1);
    scan_rec(d, n, s.lvl1, s.lvl2, st, d_total);
}

template <typename K>
void radix_sort_pairs_t(K*& keys, K*& kalt, uint32_t*& vals, uint32_t*& valt, uint64_t n, int bit_lo,
                        int bit_hi, sort_scratch& s, cudaStream_t st) {
    if (n <= 1 || bit_hi <= bit_lo) return;
    uint32_t ntiles = ceil_div_u(n, RS_NT * rs_cfg<K>::IPT);
    s.hist.reserve((uint64_t)256 * ntiles);
    for (int shift = bit_lo; shift < bit_hi; shift += 8) {
        int bits = bit_hi - shift < 8 ? bit_hi - shift : 8;
        uint32_t mask = (1u << bits) - 1;
        rs_hist_kernel<K><<<ntiles, RS_NT, 0, st>>>(keys, n, shift, mask, s.hist, ntiles);
        KERNEL_CHECK();
        exclusive_scan_u32(s.hist, (uint64_t)256 * ntiles, s.scan, st);
        rs_scatter_kernel<K><<<ntiles, RS_NT, 0, st>>>(keys, vals, kalt, valt, n, shift, mask, s.hist, ntiles);
        KERNEL_CHECK();
        K* tk = keys; keys = kalt; kalt = tk;
        uint32_t* tv = vals; vals = valt; valt = tv;
    }
}

void radix_sort_pairs(uint64_t*& keys, uint64_t*& kalt, uint32_t*& vals, uint32_t*& valt, uint64_t n, int bit_lo,
                      int bit_hi, sort_scratch& s, cudaStream_t st) {
    radix_sort_pairs_t<uint64_t>(keys, kalt, vals, valt, n, bit_lo, bit_hi, s, st);
}
void radix_sort_pairs32(uint32_t*& keys, uint32_t*& kalt, uint32_t*& vals, uint32_t*& valt, uint64_t n, int bit_lo,
                        int bit_hi, sort_scratch& s, cudaStream_t st) {
    radix_sort_pairs_t<uint32_t>(keys, kalt, vals, valt, n, bit_lo, bit_hi, s, st);
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
