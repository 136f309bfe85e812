// K5: table bucket ids of the gap index — five Rabin-Karp fingerprints modulo the Mersenne prime
// 2^107-1 per text position, one per pattern length.
//
// Replaces the per-byte work of rolling_hash_index_107
// (/root/reference/include/lz77_sss/data_structures/rolling_hash_index_107.hpp:80-150) and rk_prime<107>
// (/root/reference/patched-files/external/lce/include/rolling_hash/rolling_hash.hpp:75-86,133-155):
//   fp(pos) = sum_{k<L} T[pos+k] * base^(L-1-k)  mod 2^107-1,   bucket = fp & (table_entries-1).
// The fingerprint of a window is a pure function of its L characters (init-by-rolling from 0 and
// roll give the same value), so all positions are independent; each thread initialises once and
// then rolls over CHUNK consecutive positions.
#include "common.cuh"
#include "stages.cuh"

namespace lz77sss {
namespace {

__extension__ typedef unsigned __int128 u128;
constexpr int CHUNK = 32;

struct ghp {
    uint64_t L[5];
    uint64_t base[5];
    uint64_t nb_lo[5], nb_hi[5];   // -(base^L) mod P as two 64-bit halves
    uint64_t mask;
};

__device__ __forceinline__ u128 mers_mod(u128 x) {
    const u128 P = (((u128)1) << 107) - 1;
    x = (x & P) + (x >> 107);
    x = (x & P) + (x >> 107);
    return x >= P ? x - P : x;
}

__global__ void __launch_bounds__(128) gap_bucket_kernel(const uint8_t* __restrict__ T, uint64_t n, const ghp hp,
                                                         const uint64_t* __restrict__ rbeg,
                                                         const uint64_t* __restrict__ rend,
                                                         const uint64_t* __restrict__ roff, uint64_t nranges,
                                                         uint64_t total, uint32_t* __restrict__ out) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t flat = t * CHUNK;
    if (flat >= total) return;
    uint64_t flat_end = flat + CHUNK < total ? flat + CHUNK : total;
    // range containing flat: last k with roff[k] <= flat
    uint64_t lo = 0, hi = nranges - 1;
    while (lo < hi) {
        uint64_t mid = (lo + hi + 1) / 2;
        if (roff[mid] <= flat) lo = mid;
        else hi = mid - 1;
    }
    uint64_t k = lo;
    while (flat < flat_end) {
        while (k + 1 < nranges && roff[k + 1] <= flat) k++;        // skip empty ranges
        const uint64_t pos0 = rbeg[k] + (flat - roff[k]);
        uint64_t cnt = rend[k] - pos0;
        if (cnt > flat_end - flat) cnt = flat_end - flat;
#pragma unroll 1
        for (int x = 0; x < 5; x++) {
            const uint64_t L = hp.L[x];
            const u128 base = hp.base[x];
            const u128 nb = ((u128)hp.nb_hi[x] << 64) | hp.nb_lo[x];
            u128 fp = 0;
            bool valid = pos0 + L <= n;
            if (valid)
                for (uint64_t j = 0; j < L; j++) fp = mers_mod(fp * base + T[pos0 + j]);
            for (uint64_t c = 0; c < cnt; c++) {
                const uint64_t pos = pos0 + c;
                uint32_t v = NONE32;
                if (pos + L <= n) {
                    v = (uint32_t)((uint64_t)fp & hp.mask);
                    if (pos + L < n) {   // roll to pos+1 (reads T[pos+L] < n)
                        u128 infl = mers_mod(nb * (u128)T[pos] + T[pos + L]);
                        fp = mers_mod(fp * base + infl);
                    }
                }
                out[5 * (flat + c) + x] = v;
            }
        }
        flat += cnt;
    }
}

u128 h_mod(u128 x) {
    const u128 P = (((u128)1) << 107) - 1;
    x = (x & P) + (x >> 107);
    x = (x & P) + (x >> 107);
    return x >= P ? x - P : x;
}
u128 h_mulmod(u128 a, u128 b) {
    u128 r = 0;
    a = h_mod(a);
    while (b) {
        if (b & 1) r = h_mod(r + a);
        a = h_mod(a + a);
        b >>= 1;
    }
    return r;
}

}  // namespace

void gap_bucket_ids(const uint8_t* d_text, uint64_t n, const gap_hash_params& gp, const uint64_t* d_beg,
                    const uint64_t* d_end, const uint64_t* d_off, uint64_t nranges, uint64_t total, uint32_t* d_out,
                    cudaStream_t st) {
    if (total == 0 || nranges == 0) return;
    ghp hp;
    const u128 P = (((u128)1) << 107) - 1;
    for (int x = 0; x < 5; x++) {
        hp.L[x] = gp.patt_len[x];
        hp.base[x] = gp.base[x];
        u128 pw = 1, b = gp.base[x];
        uint64_t e = gp.patt_len[x];
        while (e) {
            if (e & 1) pw = h_mulmod(pw, b);
            b = h_mulmod(b, b);
            e >>= 1;
        }
        u128 nb = pw == 0 ? 0 : P - pw;
        hp.nb_lo[x] = (uint64_t)nb;
        hp.nb_hi[x] = (uint64_t)(nb >> 64);
    }
    hp.mask = gp.mask;
    uint64_t nthreads = (total + CHUNK - 1) / CHUNK;
    gap_bucket_kernel<<<ceil_div_u(nthreads, 128), 128, 0, st>>>(d_text, n, hp, d_beg, d_end, d_off, nranges, total, d_out);
    KERNEL_CHECK();
}

}  // namespace lz77sss
