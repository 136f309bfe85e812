// K6: parallel LZ77 decode.  Replaces lz77_sss<pos_t>::decode
// (/root/reference/include/lz77_sss/algorithms/common.cpp:33-54), a sequential byte-by-byte copy loop whose
// every output byte may depend on an earlier one.
//
//   1. output offsets = exclusive scan of max(1,len) over the factors (64-bit sums),
//   2. every output byte gets a pointer to the byte it copies (or its literal value); a factor that overlaps its own
//      source (period d = start - src < len) points every byte into the period in front of it, so a run costs one hop,
//   3. pointer jumping: one word per byte holds either the byte's value (resolved) or the position it copies from; an
//      unresolved byte reads its source's word and adopts it (the value, or the source's own source: the chain length
//      halves every round).  Tiles of 1024 bytes that are completely resolved are skipped.
//      n < 2^32 - 256: 32-bit words, resolved = 0xFFFFFF00 | value; larger n: 64-bit words, bit 63 = resolved.
//      The word is read and written whole, so "resolved" and the value are seen together.
// Algorithmic bytes: 2n + z*sizeof(factor); a hop moves one gathered word per unresolved byte, a round
// reads and writes the byte's own word once.
#include "common.cuh"
#include "stages.cuh"

namespace lz77sss {
namespace {

constexpr int DT = 1024;                       // bytes per tile of the jumping rounds
constexpr uint64_t RES64 = 1ull << 63;
constexpr uint32_t RES32 = 0xFFFFFF00u;         // 32-bit words: values at and above it are resolved bytes
template <typename P> __device__ __forceinline__ bool is_res(P v) { return sizeof(P) == 4 ? (uint32_t)v >= RES32 : ((uint64_t)v & RES64) != 0; }
template <typename P> __device__ __forceinline__ P res_word(uint64_t byte) { return sizeof(P) == 4 ? (P)(RES32 | (uint32_t)(byte & 0xFF)) : (P)(RES64 | (byte & 0xFF)); }

// ---- 64-bit exclusive scan of max(1, len) -----------------------------------------------------------------------------
constexpr int SC_NT = 256, SC_IPT = 8, SC_TILE = SC_NT * SC_IPT;
__device__ __forceinline__ uint64_t block_sum_u64(uint64_t v, uint64_t* sh) {
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    uint64_t t = 0;
    for (int w = 0; w < SC_NT / 32; w++) t += sh[w];
    __syncthreads();
    return t;
}
__global__ void __launch_bounds__(SC_NT) flen_tile_sum_kernel(const uint64_t* __restrict__ len, uint64_t z,
                                                              uint64_t* __restrict__ tsum) {
    __shared__ uint64_t sh[SC_NT / 32];
    const uint64_t base = (uint64_t)blockIdx.x * SC_TILE;
    uint64_t s = 0;
    for (int k = 0; k < SC_IPT; k++) {
        const uint64_t i = base + (uint64_t)k * SC_NT + threadIdx.x;
        if (i < z) s += len[i] ? len[i] : 1;
    }
    const uint64_t t = block_sum_u64(s, sh);
    if (threadIdx.x == 0) tsum[blockIdx.x] = t;
}
// one CTA: exclusive scan of the tile sums in chunks of 1024 with a carry; total behind the last entry
__global__ void __launch_bounds__(1024) tile_scan_kernel(uint64_t* __restrict__ tsum, uint64_t nt) {
    __shared__ uint64_t wsum[32];
    __shared__ uint64_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (uint64_t c = 0; c < nt; c += 1024) {
        const uint64_t i = c + threadIdx.x;
        const uint64_t v = i < nt ? tsum[i] : 0;
        uint64_t inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint64_t y = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= (unsigned)d) inc += y;
        }
        if (lane == 31) wsum[w] = inc;
        __syncthreads();
        uint64_t wb = 0;
        for (unsigned ww = 0; ww < w; ww++) wb += wsum[ww];
        const uint64_t base = carry;
        if (i < nt) tsum[i] = base + wb + inc - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = base + wb + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) tsum[nt] = carry;
}
template <typename P>
__global__ void __launch_bounds__(SC_NT) flen_apply_kernel(const uint64_t* __restrict__ len, uint64_t z,
                                                           const uint64_t* __restrict__ tsum, P* __restrict__ start) {
    __shared__ uint64_t wsum[SC_NT / 32];
    const uint64_t base = (uint64_t)blockIdx.x * SC_TILE + (uint64_t)threadIdx.x * SC_IPT;
    uint64_t loc[SC_IPT];
    uint64_t s = 0;
#pragma unroll
    for (int k = 0; k < SC_IPT; k++) {
        loc[k] = s;
        if (base + k < z) s += len[base + k] ? len[base + k] : 1;
    }
    const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint64_t inc = s;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint64_t y = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= (unsigned)d) inc += y;
    }
    if (lane == 31) wsum[w] = inc;
    __syncthreads();
    uint64_t wb = 0;
    for (unsigned ww = 0; ww < w; ww++) wb += wsum[ww];
    const uint64_t off = tsum[blockIdx.x] + wb + inc - s;
#pragma unroll
    for (int k = 0; k < SC_IPT; k++)
        if (base + k < z) start[base + k] = (P)(off + loc[k]);
}

// ---- pointers --------------------------------------------------------------------------------------------------------
// one warp per factor: initialise the pointers / literals of its output range
template <typename P>
__global__ void __launch_bounds__(256) decode_init_kernel(const uint64_t* __restrict__ src, const uint64_t* __restrict__ len,
                                                          const P* __restrict__ start, uint64_t z, uint64_t n,
                                                          P* __restrict__ ptr, int* __restrict__ err) {
    const uint64_t k = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    if (k >= z) return;
    const uint64_t st = start[k], l = len[k], s = src[k];
    if (l == 0) {
        if (lane == 0) {
            if (st >= n) *err = 2;
            else ptr[st] = res_word<P>(s);
        }
        return;
    }
    if (s >= st || st + l > n || st + l < st) {
        if (lane == 0) *err = 2;
        return;
    }
    const uint64_t d = st - s;                    // a self-overlapping factor repeats T[s .. st) with period d
    if (l <= d) {
        for (uint64_t i = lane; i < l; i += 32) ptr[st + i] = (P)(s + i);
    } else {
        for (uint64_t i = lane; i < l; i += 32) ptr[st + i] = (P)(s + i % d);
    }
}

template <typename P, int HOPS>
__global__ void __launch_bounds__(256) decode_round_kernel(uint64_t n, P* __restrict__ ptr, uint8_t* __restrict__ tile_done,
                                                           unsigned long long* __restrict__ unresolved) {
    if (tile_done[blockIdx.x]) return;
    const uint64_t x0 = (uint64_t)blockIdx.x * DT;
    P v[DT / 256];
    int pending = 0;
#pragma unroll
    for (int k = 0; k < DT / 256; k++) {          // own words, then the first hops, all in flight together
        const uint64_t x = x0 + (uint64_t)k * 256 + threadIdx.x;
        v[k] = x < n ? ptr[x] : res_word<P>(0);
    }
    P w[DT / 256];
#pragma unroll
    for (int k = 0; k < DT / 256; k++) w[k] = is_res<P>(v[k]) ? v[k] : *((volatile P*)&ptr[v[k]]);
    if (HOPS > 1) {
#pragma unroll
        for (int k = 0; k < DT / 256; k++)        // second hops
            if (!is_res<P>(w[k])) w[k] = *((volatile P*)&ptr[w[k]]);
    }
#pragma unroll
    for (int k = 0; k < DT / 256; k++) {
        const uint64_t x = x0 + (uint64_t)k * 256 + threadIdx.x;
        if (x < n && !is_res<P>(v[k])) {
            ptr[x] = w[k];
            if (!is_res<P>(w[k])) pending++;
        }
    }
    const int any = __syncthreads_count(pending);
    if (threadIdx.x == 0) {
        if (any) atomicAdd(unresolved, (unsigned long long)any);
        else tile_done[blockIdx.x] = 1;
    }
}

template <typename P>
__global__ void decode_pack_kernel(const P* __restrict__ ptr, uint64_t n, uint8_t* __restrict__ out) {
    const uint64_t x = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (x + 3 < n) {
        uchar4 o;
        o.x = (uint8_t)(ptr[x] & 0xFF); o.y = (uint8_t)(ptr[x + 1] & 0xFF); o.z = (uint8_t)(ptr[x + 2] & 0xFF); o.w = (uint8_t)(ptr[x + 3] & 0xFF);
        *reinterpret_cast<uchar4*>(out + x) = o;
    } else {
        for (uint64_t k = x; k < n; k++) out[k] = (uint8_t)(ptr[k] & 0xFF);
    }
}

__global__ void __launch_bounds__(256) bytes_differ_kernel(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b,
                                                           uint64_t n, unsigned int* __restrict__ differ) {
    const uint64_t x = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
    if (x >= n) return;
    bool d = false;
    for (uint64_t k = x; k < x + 8 && k < n; k++) d |= a[k] != b[k];
    if (d) atomicOr(differ, 1u);
}

template <typename P>
int decode_impl(const uint64_t* d_src, const uint64_t* d_len, uint64_t z, uint8_t* d_out, uint64_t n, cudaStream_t st,
                int* rounds, decode_scratch& sc) {
    const uint64_t nt = (z + SC_TILE - 1) / SC_TILE;
    const uint64_t ntile = (n + DT - 1) / DT;
    sc.start.reserve((z + 1) * sizeof(P));
    sc.ptr.reserve(n * sizeof(P));
    sc.tsum.reserve(nt + 2);
    sc.tile_done.reserve(ntile + 1);
    sc.cnt.reserve(2);
    P* start = reinterpret_cast<P*>(sc.start.get());
    P* ptr = reinterpret_cast<P*>(sc.ptr.get());
    int* err = reinterpret_cast<int*>(sc.cnt.get() + 1);
    CUDA_CHECK(cudaMemsetAsync(sc.cnt.get(), 0, 16, st));
    CUDA_CHECK(cudaMemsetAsync(sc.tile_done.get(), 0, ntile, st));
    flen_tile_sum_kernel<<<(unsigned)nt, SC_NT, 0, st>>>(d_len, z, sc.tsum);
    KERNEL_CHECK();
    tile_scan_kernel<<<1, 1024, 0, st>>>(sc.tsum, nt);
    KERNEL_CHECK();
    flen_apply_kernel<P><<<(unsigned)nt, SC_NT, 0, st>>>(d_len, z, sc.tsum, start);
    KERNEL_CHECK();
    uint64_t h_total = 0;
    CUDA_CHECK(cudaMemcpyAsync(&h_total, sc.tsum.get() + nt, 8, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    if (h_total != n) return -1;   // factor lengths do not add up to n
    decode_init_kernel<P><<<ceil_div_u(z * 32, 256), 256, 0, st>>>(d_src, d_len, start, z, n, ptr, err);
    KERNEL_CHECK();
    int h_err = 0;
    CUDA_CHECK(cudaMemcpyAsync(&h_err, err, 4, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    if (h_err) return -2;
    // the longest chain has fewer than n hops and halves every round
    // (two hops per round, LZ77SSS_DECODE_HOPS=2, save a round but gather more words in total: measured slower)
    static const bool two_hops = getenv("LZ77SSS_DECODE_HOPS") && getenv("LZ77SSS_DECODE_HOPS")[0] == '2';
    bool done = false;
    for (int r = 0; r < 72 && !done; r++) {
        CUDA_CHECK(cudaMemsetAsync(sc.cnt.get(), 0, 8, st));
        if (two_hops) decode_round_kernel<P, 2><<<(unsigned)ntile, 256, 0, st>>>(n, ptr, sc.tile_done, sc.cnt);
        else decode_round_kernel<P, 1><<<(unsigned)ntile, 256, 0, st>>>(n, ptr, sc.tile_done, sc.cnt);
        KERNEL_CHECK();
        unsigned long long h_cnt = 0;
        CUDA_CHECK(cudaMemcpyAsync(&h_cnt, sc.cnt.get(), 8, cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        if (rounds) *rounds = r + 1;
        done = h_cnt == 0;
    }
    if (!done) return -4;          // did not converge: never emit unresolved bytes as zeros
    decode_pack_kernel<P><<<ceil_div_u((n + 3) / 4, 256), 256, 0, st>>>(ptr, n, d_out);
    KERNEL_CHECK();
    return 0;
}

}  // namespace

bool device_equal(const uint8_t* d_a, const uint8_t* d_b, uint64_t n, cudaStream_t st) {
    if (!n) return true;
    dbuf<unsigned int> flag;
    flag.reserve(1);
    CUDA_CHECK(cudaMemsetAsync(flag.get(), 0, 4, st));
    bytes_differ_kernel<<<ceil_div_u((n + 7) / 8, 256), 256, 0, st>>>(d_a, d_b, n, flag);
    KERNEL_CHECK();
    unsigned int h = 1;
    CUDA_CHECK(cudaMemcpyAsync(&h, flag.get(), 4, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    return h == 0;
}

int decode_factors(const uint64_t* d_src, const uint64_t* d_len, uint64_t z, uint8_t* d_out, uint64_t n,
                   cudaStream_t st, int* rounds, decode_scratch& sc) {
    if (rounds) *rounds = 0;
    if (n == 0) return 0;
    static const bool force64 = getenv("LZ77SSS_FORCE_POS64") && getenv("LZ77SSS_FORCE_POS64")[0] == '1';
    if (n < (uint64_t)RES32 && !force64) return decode_impl<uint32_t>(d_src, d_len, z, d_out, n, st, rounds, sc);
    return decode_impl<uint64_t>(d_src, d_len, z, d_out, n, st, rounds, sc);
}

}  // namespace lz77sss
