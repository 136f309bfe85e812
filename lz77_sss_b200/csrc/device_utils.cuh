// Small device helpers shared by the kernels.
#pragma once
#include <cstdint>

namespace lz77sss {

// 8 text bytes starting at an arbitrary address, little-endian (byte 0 in bits 0..7).
// Reads the two enclosing aligned words: up to 15 bytes past p (covered by TEXT_PAD).
__device__ __forceinline__ uint64_t load_u64_unaligned(const uint8_t* p) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const uint64_t* q = reinterpret_cast<const uint64_t*>(a & ~uintptr_t(7));
    const unsigned sh = (unsigned)(a & 7) * 8;
    uint64_t lo = q[0];
    if (sh == 0) return lo;
    uint64_t hi = q[1];
    return (lo >> sh) | (hi << (64 - sh));
}

// 8 text bytes ENDING at address p (inclusive): bytes p-7..p, byte p in bits 56..63.
// Caller guarantees p-7 >= text base - 8 is readable (we only call it with p >= 7).
__device__ __forceinline__ uint64_t load_u64_ending_at(const uint8_t* p) { return load_u64_unaligned(p - 7); }

// Warp-cooperative LCE of T[x..) and T[y..), at most lim characters, starting at offset off.
// All 32 lanes call it with the same arguments; 256 bytes per iteration per string.
__device__ __forceinline__ uint64_t warp_lce(const uint8_t* __restrict__ T, uint64_t x, uint64_t y, uint64_t off,
                                             uint64_t lim) {
    const unsigned lane = threadIdx.x & 31;
    while (off < lim) {
        const uint64_t o = off + 8 * lane;
        uint64_t diff = 0;
        if (o < lim) diff = load_u64_unaligned(T + x + o) ^ load_u64_unaligned(T + y + o);
        const unsigned m = __ballot_sync(0xffffffffu, diff != 0);
        if (m) {
            const int fl = __ffs(m) - 1;
            const uint64_t dl = __shfl_sync(0xffffffffu, diff, fl);
            const uint64_t r = off + 8 * (uint64_t)fl + (__ffsll((long long)dl) - 1) / 8;
            return r < lim ? r : lim;
        }
        off += 256;
    }
    return lim;
}

// The same for a GROUP of G consecutive lanes of a warp (G = 8, 16 or 32; several groups of one warp work on different
// queries at the same time).  gmask = the group's lanes, glane = lane index inside the group.  256 bytes per iteration per
// string whatever G is: a lane compares 256 / G consecutive bytes.
template <int G>
__device__ __forceinline__ uint64_t group_lce(const uint8_t* __restrict__ T, uint64_t x, uint64_t y, uint64_t off, uint64_t lim,
                                              unsigned gmask, unsigned glane) {
    constexpr int PER = 256 / G;                        // bytes per lane and iteration
    const unsigned gbase = __ffs(gmask) - 1;
    while (off < lim) {
        const uint64_t o0 = off + (uint64_t)PER * glane;
        uint32_t first = PER;                           // first differing byte of the lane's chunk, PER = none
#pragma unroll
        for (int k = PER / 8 - 1; k >= 0; k--) {
            const uint64_t o = o0 + 8 * k;
            if (o < lim) {
                const uint64_t diff = load_u64_unaligned(T + x + o) ^ load_u64_unaligned(T + y + o);
                if (diff) first = 8 * k + (__ffsll((long long)diff) - 1) / 8;
            }
        }
        const unsigned m = __ballot_sync(gmask, first < (uint32_t)PER);
        if (m) {
            const int fl = __ffs(m) - 1;                // warp lane of the first group lane with a difference
            const uint32_t f = __shfl_sync(gmask, first, fl);
            const uint64_t r = off + (uint64_t)PER * (fl - gbase) + f;
            return r < lim ? r : lim;
        }
        off += 256;
    }
    return lim;
}

// single-thread LCE (8 bytes per step)
__device__ __forceinline__ uint64_t thread_lce(const uint8_t* __restrict__ T, uint64_t x, uint64_t y, uint64_t lim) {
    uint64_t k = 0;
    while (k < lim) {
        uint64_t diff = load_u64_unaligned(T + x + k) ^ load_u64_unaligned(T + y + k);
        if (diff) {
            k += (__ffsll((long long)diff) - 1) / 8;
            return k < lim ? k : lim;
        }
        k += 8;
    }
    return lim;
}

__device__ __forceinline__ int bit_width_u64(uint64_t x) { return x ? 64 - __clzll((long long)x) : 0; }

}  // namespace lz77sss
