// rabin_karp_substring<31, pos_t> (/root/reference/include/lz77_sss/data_structures/rabin_karp_substring.hpp:35-240):
// fingerprints of all prefixes T[0..16k) modulo the Mersenne prime 2^31 - 1, and batched substring fingerprints
//   substring(pos, len) = fp(T[0..pos+len)) - fp(T[0..pos)) * b^len        (:223-239)
// The reference builds its table `fps` with one sequential pass per thread and a sequential stitch (:136-171).
// Here the pass is ONE kernel, a prefix scan of hash powers:
//   a lane hashes 64 consecutive bytes (16 per step, as four byte-wise dot products, see poly16), the warp combines its
//   lanes by a shuffle scan with the multipliers b^(64 * 2^k), the CTA its 8 warps: every 16-byte granule has its
//   fingerprint RELATIVE to the 16 KiB tile and the tile has its aggregate A.  A tile publishes A,
//   warp 0 looks back over its predecessors (decoupled look-back, 64 tiles per step: A_{t-1} + W A_{t-2} + ... with
//   W = b^16384, until a tile that already knows its inclusive value), publishes the inclusive value, and every lane
//   finishes: value = carry(tile) * b^(16 * (granule in tile + 1)) + relative value.
// Stored residues are canonical (in [0, 2^31 - 1)), like the reference's mod() (:57-64), so every value is bit-identical
// to the reference's for the same base.  The table has one entry per 16 bytes whatever sampling rate the caller names;
// the reference's `fps` vector for a rate s is derived from it (rks_fps_kernel) for the parity tests.
// HBM bytes of the build: n read + 4 * n / 16 written = the algorithmic bytes.
#include <algorithm>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "stages.cuh"

namespace lz77sss {
namespace {

constexpr uint32_t P31 = 0x7FFFFFFFu;
constexpr int RK_G = 16;                 // bytes per granule (table entry)
constexpr int RK_LANE_BYTES = 64;        // bytes per lane
constexpr int RK_NW = 8;                 // warps per CTA
constexpr int RK_TILE = RK_NW * 32 * RK_LANE_BYTES;       // 16384 bytes
constexpr int RK_TILE_G = RK_TILE / RK_G;                 // 1024 granules
constexpr int RK_CTAS_PER_SM = 8;        // resident CTAs per SM of the persistent build kernel

__host__ __device__ __forceinline__ uint32_t mod31(uint64_t x) {      // any 64-bit x
    x = (x >> 31) + (x & P31);          // < 2^34
    x = (x >> 31) + (x & P31);          // < 2^31 + 8
    uint32_t r = (uint32_t)x;
    if (x >= P31) r = (uint32_t)(x - P31);
    return r;
}
__host__ __device__ __forceinline__ uint32_t mulmod31(uint32_t a, uint32_t b) { return mod31((uint64_t)a * b); }

struct rk_consts {
    uint32_t wv[4][4];       // wv[m][q] byte j = byte m of b^(15 - 4 q - j): the weights of a granule's word q for __dp4a
    uint32_t b16, b64;
    uint32_t scan_w[5];      // b^(64 * 2^k): lanes of a warp
    uint32_t warp_w;         // b^2048: one warp's bytes
    uint32_t lane_w[32];     // b^(64 * lane)
    uint32_t gran_w[4];      // b^(16 * (g + 1)), g = granule inside a lane
    uint32_t warp_pow[RK_NW];    // b^(2048 * w)
    uint32_t tile_pow[64];   // W^l, W = b^RK_TILE
    uint32_t tile_w64;       // W^64
};

// sum_{j<16} byte_j * b^(15-j) for the 16 bytes of a granule (byte 0 of q.x is the first character), exact, < 2^45.
// Each weight is split into its four bytes, so the sum is four byte-wise dot products (IDP.4A, no byte extraction):
// sum_j c_j w_j = sum_m 2^(8m) sum_j c_j w_j[m], every inner sum < 16 * 255 * 255 < 2^20.
__device__ __forceinline__ uint64_t poly16(const uint4 q, const rk_consts& c) {
    uint32_t a[4];
#pragma unroll
    for (int m = 0; m < 4; m++)
        a[m] = __dp4a(q.x, c.wv[m][0], __dp4a(q.y, c.wv[m][1], __dp4a(q.z, c.wv[m][2], __dp4a(q.w, c.wv[m][3], 0u))));
    const uint32_t lo = a[0] + (a[1] << 8), hi = a[2] + (a[3] << 8);            // < 2^29 each
    return (uint64_t)hi * 65536u + lo;
}
// Inside the kernel values are kept "wide": < 2^32 and congruent to the fingerprint.  wide * canonical + wide is below
// 2^63 + 2^32, and for such x (2^31 = 1 and 2^62 = 1 modulo the prime) the three-term sum below is again < 2^32.
__device__ __forceinline__ uint32_t fold31(uint64_t x) {
    return ((uint32_t)x & P31) + ((uint32_t)(x >> 31) & P31) + (uint32_t)(x >> 62);
}
// the canonical residue of such an x, in 32-bit steps
__device__ __forceinline__ uint32_t canon31(uint64_t x) {
    const uint32_t t = fold31(x);
    uint32_t r = (t & P31) + (t >> 31);          // <= 2^31
    if (r >= P31) r -= P31;
    return r;
}
__device__ __forceinline__ uint32_t mulw(uint32_t wide, uint32_t canonical) { return fold31((uint64_t)wide * canonical); }

constexpr unsigned long long RK_AGG = 1ull << 32, RK_INC = 2ull << 32;

// Persistent CTAs: CTA c takes the tiles c, c + gridDim.x, ... in order, and every CTA of the grid is resident, so the
// tile in front of any tile is being worked on (or done): the look-back cannot wait for a CTA that is not running.
// The loads of the CTA's next tile are issued before the look-back of the current one, so they overlap its latency.
__global__ void __launch_bounds__(RK_NW * 32, RK_CTAS_PER_SM) rks_build_kernel(const uint8_t* __restrict__ T, uint64_t ngran,
                                                                              uint32_t ntiles, const rk_consts c,
                                                                              uint32_t* __restrict__ pf,
                                                                              unsigned long long* status) {
    __shared__ uint32_t s_warp[RK_NW];
    __shared__ uint32_t s_carry;
    const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint64_t lane_g = (uint64_t)(w * 32 + lane) * 4;                       // first granule of the lane inside a tile
    const uint32_t lane_pow = canon31((uint64_t)c.warp_pow[w] * c.lane_w[lane]);           // b^(bytes of the tile in front of the lane)
    uint32_t tile = blockIdx.x;
    uint4 q[4];
    {
        const uint64_t g0 = (uint64_t)tile * RK_TILE_G + lane_g;
        const uint4* src = reinterpret_cast<const uint4*>(T + g0 * RK_G);        // T is 16-byte aligned, zero padded
#pragma unroll
        for (int g = 0; g < 4; g++) {
            q[g] = make_uint4(0, 0, 0, 0);
            if (tile < ntiles && g0 + g < ngran) q[g] = src[g];
        }
    }
    for (; tile < ntiles; tile += gridDim.x) {
        const uint64_t g0 = (uint64_t)tile * RK_TILE_G + lane_g;
        uint32_t r[4];
        uint32_t fp = 0;                 // r[] and fp are congruent to the fingerprints but only < 2^32 until the last step
#pragma unroll
        for (int g = 0; g < 4; g++) {
            fp = fold31((uint64_t)fp * c.b16 + poly16(q[g], c));
            r[g] = fp;
        }
        {   // the next tile of this CTA
            const uint64_t nt = (uint64_t)tile + gridDim.x;
            const uint64_t n0 = nt * RK_TILE_G + lane_g;
            const uint4* src = reinterpret_cast<const uint4*>(T + n0 * RK_G);
#pragma unroll
            for (int g = 0; g < 4; g++) {
                q[g] = make_uint4(0, 0, 0, 0);
                if (nt < ntiles && n0 + g < ngran) q[g] = src[g];
            }
        }
        // inclusive scan over the lanes: x = sum_{u <= lane} L_u * b^(64 * (lane - u))
        uint32_t x = fp;
#pragma unroll
        for (int k = 0; k < 5; k++) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, x, 1 << k);
            if (lane >= (1u << k)) x = fold31((uint64_t)y * c.scan_w[k] + x);
        }
        uint32_t ex = __shfl_up_sync(0xffffffffu, x, 1);
        if (lane == 0) ex = 0;
        if (lane == 31) s_warp[w] = x;
        __syncthreads();
        // fingerprint of the tile's bytes in front of this warp: lanes u < w hold warp u's share
        uint64_t part = 0;
        if (lane < w) part = mulw(s_warp[lane], c.warp_pow[w - 1 - lane]);
#pragma unroll
        for (int d = RK_NW / 2; d >= 1; d >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d);
        const uint32_t wprev = fold31(__shfl_sync(0xffffffffu, part, 0));
        const uint32_t in_tile = fold31((uint64_t)wprev * c.lane_w[lane] + ex);   // ... in front of this lane
        if (w == 0) {
            uint64_t ap = 0;             // the whole tile
            if (lane < RK_NW) ap = mulw(s_warp[lane], c.warp_pow[RK_NW - 1 - lane]);
#pragma unroll
            for (int d = RK_NW / 2; d >= 1; d >>= 1) ap += __shfl_xor_sync(0xffffffffu, ap, d);
            const uint32_t A = fold31(__shfl_sync(0xffffffffu, ap, 0));
            if (lane == 0 && tile > 0) atomicExch(status + tile, RK_AGG | A);
            uint32_t E = 0;              // fingerprint of everything in front of the tile
            if (tile > 0) {
                uint32_t m = 1;          // W^(tiles already folded in)
                int64_t first = (int64_t)tile - 1;
                while (true) {          // 64 tiles per step: a lane reads the states of first - 2 lane and the one in front
                    const int64_t i0 = first - 2 * (int64_t)lane, i1 = i0 - 1;
                    unsigned long long s0 = RK_INC, s1 = RK_INC;                  // in front of the text: inclusive, value 0
                    while (true) {
                        if (i0 >= 0) s0 = *((volatile unsigned long long*)(status + i0));
                        if (i1 >= 0) s1 = *((volatile unsigned long long*)(status + i1));
                        const unsigned f0 = (unsigned)(s0 >> 32), f1 = (unsigned)(s1 >> 32);
                        if (f0 == 2 || (f0 == 1 && f1 != 0)) break;
                    }
                    const bool inc0 = (s0 >> 32) == 2;
                    const unsigned inc = __ballot_sync(0xffffffffu, inc0 || (s1 >> 32) == 2);
                    const int stop = inc ? __ffs(inc) - 1 : 31;                   // lane of the nearest inclusive predecessor
                    uint64_t pt = 0;
                    if ((int)lane <= stop) {
                        pt = mulw((uint32_t)s0, c.tile_pow[2 * lane]);
                        if (!inc0) pt += mulw((uint32_t)s1, c.tile_pow[2 * lane + 1]);
                        pt = mulw(fold31(pt), m);
                    }
#pragma unroll
                    for (int d = 16; d >= 1; d >>= 1) pt += __shfl_xor_sync(0xffffffffu, pt, d);
                    E = fold31((uint64_t)E + pt);
                    if (inc) break;
                    m = canon31((uint64_t)m * c.tile_w64);
                    first -= 64;
                }
            }
            if (lane == 0) {
                atomicExch(status + tile, RK_INC | fold31((uint64_t)E * c.tile_pow[1] + A));
                s_carry = E;
            }
        }
        __syncthreads();
        const uint32_t cin = s_carry;
        const uint32_t carry = fold31((uint64_t)cin * lane_pow + in_tile);        // in front of this lane, whole text
        uint4 out;
        out.x = canon31((uint64_t)carry * c.gran_w[0] + r[0]);
        out.y = canon31((uint64_t)carry * c.gran_w[1] + r[1]);
        out.z = canon31((uint64_t)carry * c.gran_w[2] + r[2]);
        out.w = canon31((uint64_t)carry * c.gran_w[3] + r[3]);
        if (g0 + 3 < ngran) {
            *reinterpret_cast<uint4*>(pf + g0) = out;
        } else {
            if (g0 < ngran) pf[g0] = out.x;
            if (g0 + 1 < ngran) pf[g0 + 1] = out.y;
            if (g0 + 2 < ngran) pf[g0 + 2] = out.z;
        }
    }
}

struct rk_query_consts {
    uint32_t b;
    uint32_t sq[40];         // b^(2^k)
};

__device__ __forceinline__ uint32_t rk_pow(const rk_query_consts& q, uint64_t e) {
    uint32_t r = 1;
    for (int k = 0; e; k++, e >>= 1)
        if (e & 1) r = mulmod31(r, q.sq[k]);
    return r;
}
// substring_up_to (:212-219) on the 16-byte table
__device__ __forceinline__ uint32_t rk_up_to(const uint8_t* __restrict__ T, const uint32_t* __restrict__ pf, uint32_t b,
                                             uint64_t pos) {
    const uint64_t g = pos / RK_G;
    uint32_t fp = g ? pf[g - 1] : 0u;
    for (uint64_t i = g * RK_G; i < pos; i++) fp = mod31((uint64_t)fp * b + T[i]);
    return fp;
}

__global__ void rks_substring_kernel(const uint8_t* __restrict__ T, const uint32_t* __restrict__ pf, const rk_query_consts q,
                                     const uint64_t* __restrict__ pos, const uint64_t* __restrict__ len, uint64_t nq,
                                     uint32_t* __restrict__ out) {
    const uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nq) return;
    const uint32_t beg = mulmod31(rk_up_to(T, pf, q.b, pos[k]), rk_pow(q, len[k]));
    const uint32_t end = rk_up_to(T, pf, q.b, pos[k] + len[k]);
    out[k] = end >= beg ? end - beg : P31 - (beg - end);
}

// the reference's table for sampling rate s: fps[k] = fingerprint of T[0 .. min(k * s, n)), k = 0 .. ceil(n / s)
__global__ void rks_fps_kernel(const uint8_t* __restrict__ T, const uint32_t* __restrict__ pf, uint32_t b, uint64_t n, uint64_t s,
                               uint64_t cnt, uint32_t* __restrict__ out) {
    const uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= cnt) return;
    uint64_t pos = k * s;
    if (pos > n) pos = n;
    out[k] = rk_up_to(T, pf, b, pos);
}

uint32_t h_pow(uint32_t b, uint64_t e) {
    uint32_t r = 1, x = b;
    while (e) {
        if (e & 1) r = mulmod31(r, x);
        x = mulmod31(x, x);
        e >>= 1;
    }
    return r;
}

}  // namespace

void rks_build(const uint8_t* d_text, uint64_t n, uint32_t base, rks_state& R, cudaStream_t st) {
    R.n = n;
    R.base = base;
    R.ngran = (n + RK_G - 1) / RK_G;
    if (R.ngran == 0) return;
    const uint32_t ntiles = ceil_div_u(R.ngran, RK_TILE_G);
    R.pf.reserve(R.ngran + 4);
    R.status.reserve(ntiles);
    CUDA_CHECK(cudaMemsetAsync(R.status.get(), 0, (uint64_t)ntiles * 8, st));
    rk_consts c;
    uint32_t bp[16];
    bp[0] = 1;
    for (int k = 1; k < 16; k++) bp[k] = mulmod31(bp[k - 1], base);
    for (int m = 0; m < 4; m++)
        for (int q = 0; q < 4; q++) {
            c.wv[m][q] = 0;
            for (int j = 0; j < 4; j++) c.wv[m][q] |= ((bp[15 - 4 * q - j] >> (8 * m)) & 0xFFu) << (8 * j);
        }
    c.b16 = h_pow(base, 16);
    c.b64 = h_pow(base, 64);
    for (int k = 0; k < 5; k++) c.scan_w[k] = h_pow(base, 64ull << k);
    c.warp_w = h_pow(base, 2048);
    for (int l = 0; l < 32; l++) c.lane_w[l] = h_pow(base, 64ull * l);
    for (int g = 0; g < 4; g++) c.gran_w[g] = h_pow(base, 16ull * (g + 1));
    for (int w = 0; w < RK_NW; w++) c.warp_pow[w] = h_pow(base, 2048ull * w);
    const uint32_t W = h_pow(base, RK_TILE);
    c.tile_pow[0] = 1;
    for (int l = 1; l < 64; l++) c.tile_pow[l] = mulmod31(c.tile_pow[l - 1], W);
    c.tile_w64 = mulmod31(c.tile_pow[63], W);
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        CUDA_CHECK(cudaGetDevice(&dev));
        CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    }
    int per_sm = 0;                      // every CTA of the grid has to be resident (see the kernel)
    CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rks_build_kernel, RK_NW * 32, 0));
    if (per_sm < 1) throw std::runtime_error("rks_build: kernel does not fit an SM");
    if (per_sm > RK_CTAS_PER_SM) per_sm = RK_CTAS_PER_SM;
    const uint32_t grid = std::min<uint64_t>(ntiles, (uint64_t)sms * per_sm);
    rks_build_kernel<<<grid, RK_NW * 32, 0, st>>>(d_text, R.ngran, ntiles, c, R.pf, R.status.get());
    KERNEL_CHECK();
}

// ---- the build without any allocation (lz77sss_rks_table_noalloc) ---------------------------------------------------------
// The look-back states of the tiles live in a file-scope device array instead of a cudaMalloc'ed buffer: every GPU that loads
// the library holds these RKS_NOALLOC_MAX_TILES * 8 bytes (512 KiB) until the process ends.  All calls share the array and it
// keeps the last call's contents: it is cleared on the caller's stream in front of every launch, and two calls that overlap
// on the GPU (different streams, no ordering between them) race on it.
static __device__ __align__(256) unsigned long long g_rks_noalloc_status[RKS_NOALLOC_MAX_BYTES / RK_TILE];

void rks_build_noalloc(const uint8_t* d_text, uint64_t n, uint32_t base, uint32_t* d_out, cudaStream_t st) {
    if (n > RKS_NOALLOC_MAX_BYTES)
        throw std::invalid_argument("rks_build_noalloc: n exceeds the static scratch (RKS_NOALLOC_MAX_BYTES = 2^30 bytes of text)");
    const uint64_t ngran = (n + RK_G - 1) / RK_G;
    if (ngran == 0) return;
    const uint32_t ntiles = ceil_div_u(ngran, RK_TILE_G);
    // In host code the array's name is only a stand-in: its device address is asked for once per device, with that device
    // current, and kept.
    constexpr int MAX_DEV = 64;
    static unsigned long long* addr[MAX_DEV] = {};
    static int sms_of[MAX_DEV] = {};
    static int per_sm_of[MAX_DEV] = {};
    static std::mutex mu;
    int dev = 0;
    CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEV) throw std::invalid_argument("rks_build_noalloc: device ordinal out of range");
    unsigned long long* status = nullptr;
    int sms = 0, per_sm = 0;
    {
        std::lock_guard<std::mutex> g(mu);
        if (!addr[dev]) {
            void* p = nullptr;
            CUDA_CHECK(cudaGetSymbolAddress(&p, g_rks_noalloc_status));
            CUDA_CHECK(cudaDeviceGetAttribute(&sms_of[dev], cudaDevAttrMultiProcessorCount, dev));
            CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_of[dev], rks_build_kernel, RK_NW * 32, 0));
            addr[dev] = static_cast<unsigned long long*>(p);
        }
        status = addr[dev];
        sms = sms_of[dev];
        per_sm = per_sm_of[dev];
    }
    if (per_sm < 1) throw std::runtime_error("rks_build_noalloc: kernel does not fit an SM");
    if (per_sm > RK_CTAS_PER_SM) per_sm = RK_CTAS_PER_SM;
    CUDA_CHECK(cudaMemsetAsync(status, 0, (uint64_t)ntiles * 8, st));
    rk_consts c;
    uint32_t bp[16];
    bp[0] = 1;
    for (int k = 1; k < 16; k++) bp[k] = mulmod31(bp[k - 1], base);
    for (int m = 0; m < 4; m++)
        for (int q = 0; q < 4; q++) {
            c.wv[m][q] = 0;
            for (int j = 0; j < 4; j++) c.wv[m][q] |= ((bp[15 - 4 * q - j] >> (8 * m)) & 0xFFu) << (8 * j);
        }
    c.b16 = h_pow(base, 16);
    c.b64 = h_pow(base, 64);
    for (int k = 0; k < 5; k++) c.scan_w[k] = h_pow(base, 64ull << k);
    c.warp_w = h_pow(base, 2048);
    for (int l = 0; l < 32; l++) c.lane_w[l] = h_pow(base, 64ull * l);
    for (int g = 0; g < 4; g++) c.gran_w[g] = h_pow(base, 16ull * (g + 1));
    for (int w = 0; w < RK_NW; w++) c.warp_pow[w] = h_pow(base, 2048ull * w);
    const uint32_t W = h_pow(base, RK_TILE);
    c.tile_pow[0] = 1;
    for (int l = 1; l < 64; l++) c.tile_pow[l] = mulmod31(c.tile_pow[l - 1], W);
    c.tile_w64 = mulmod31(c.tile_pow[63], W);
    const uint32_t grid = std::min<uint64_t>(ntiles, (uint64_t)sms * per_sm);
    rks_build_kernel<<<grid, RK_NW * 32, 0, st>>>(d_text, ngran, ntiles, c, d_out, status);
    KERNEL_CHECK();
}

void rks_substring_batch(const uint8_t* d_text, const rks_state& R, const uint64_t* d_pos, const uint64_t* d_len, uint64_t nq,
                         uint32_t* d_out, cudaStream_t st) {
    if (!nq) return;
    rk_query_consts q;
    q.b = R.base;
    uint32_t x = R.base;
    for (int k = 0; k < 40; k++) { q.sq[k] = x; x = mulmod31(x, x); }
    rks_substring_kernel<<<ceil_div_u(nq, 256), 256, 0, st>>>(d_text, R.pf, q, d_pos, d_len, nq, d_out);
    KERNEL_CHECK();
}

void rks_fps(const uint8_t* d_text, const rks_state& R, uint64_t s, uint64_t cnt, uint32_t* d_out, cudaStream_t st) {
    if (!cnt) return;
    rks_fps_kernel<<<ceil_div_u(cnt, 256), 256, 0, st>>>(d_text, R.pf, R.base, R.n, s, cnt, d_out);
    KERNEL_CHECK();
}

}  // namespace lz77sss
