// Greedy phrase emission on the device: factorize_greedy + longest_prev_occ + rolling_hash_index_107
// (/root/reference/include/lz77_sss/algorithms/approximate/factorize/greedy.cpp:34-140,
//  .../factorize/common.cpp:33-111, include/lz77_sss/data_structures/rolling_hash_index_107.hpp:80-150).
//
// The reference walks the text once, and at every gap position performs 5 read-modify-writes of one shared hash
// table H[fp & mask] = position.  Everything here rests on one observation: positions are inserted in increasing
// order, so "H[b] at the time position i is probed" is the LARGEST INSERTED position <= i among the entries that
// fall into bucket b.  That is an offline predecessor problem:
//   1. hash:   bucket ids of all 5 pattern lengths for every position of a slice (gap_entries_kernel),
//   2. sort:   stable radix sort of the entries by bucket  ->  every bucket is a list in insertion order,
//   3. parse:  the text is cut into blocks; one warp parses one block exactly like the reference does
//              (gap steps, phrase steps, truncation, phrase-start probes), answering each table probe by walking
//              its bucket list backwards to the nearest entry whose position is marked "inserted";
//              blocks start speculatively at their first position, then
//   4. link:   the true chain of blocks (exit state of one = entry state of the next) is followed by pointer
//              doubling, entry states are propagated, the inserted-bitmap is rebuilt from the traces,
//   and 3-4 repeat until no trace changes.  The fixed point is unique and equals the sequential parse
//   (the prefix that is already right grows by at least one block per round; in practice 3-5 rounds suffice,
//   because LZ parses re-synchronise within a few factors).
// Slices bound the memory (5 entries per position) and carry the table as "last inserted position per bucket".
// The last 65 positions of the text, where the reference's end-of-text rules for stale fingerprints apply,
// are finished by the host with the table entries it asks for (api.cu).
#include <algorithm>
#include <vector>

#include "common.cuh"
#include "device_utils.cuh"
#include "lce_device.cuh"
#include "stages.cuh"

namespace lz77sss {
namespace {

__extension__ typedef unsigned __int128 u128;
constexpr uint64_t NONE64 = ~0ull;
constexpr uint32_t KIND_PHR = 1u << 30, KIND_PROBED = 1u << 31, POS_MASK = (1u << 30) - 1;

// residues modulo the Mersenne prime P = 2^107 - 1 as (lo: 64 bits, hi: 43 bits)
struct m107 {
    uint64_t lo, hi;
};
constexpr uint64_t HI_MASK = (1ull << 43) - 1;

struct ghp5 {
    uint64_t L[5];
    uint64_t base[5];
    uint64_t mask;
    uint32_t invalid_key;          // table_entries (sorts behind every bucket)
    const m107* nbtab;             // [5][256]: (-(base^L) * c) mod P for every character c
};

// (hi : lo) with hi < 2^63  ->  value mod P
__device__ __forceinline__ m107 fold107(uint64_t lo, uint64_t hi) {
    uint64_t c = hi >> 43;        // x = c * 2^107 + rest == c + rest (mod P)
    hi &= HI_MASK;
    lo += c;
    hi += lo < c;
    if (hi >> 43) {               // at most one more wrap
        hi &= HI_MASK;
        lo += 1;
        hi += lo == 0;
    }
    if (hi == HI_MASK && lo == ~0ull) { hi = 0; lo = 0; }   // x == P
    m107 r;
    r.lo = lo; r.hi = hi;
    return r;
}
// (fp * base + add) mod P, base < 2^20, fp and add reduced
__device__ __forceinline__ m107 muladd107(m107 fp, uint64_t base, m107 add) {
    uint64_t lo = fp.lo * base;
    uint64_t hi = fp.hi * base + __umul64hi(fp.lo, base);    // < 2^63 + 2^20
    lo += add.lo;
    hi += add.hi + (lo < add.lo);
    return fold107(lo, hi);
}

// keys[5*rel + (4-x)] = bucket of the window of pattern x at position a0+rel (insertion order of the reference:
// positions ascending, patterns 4..0), invalid_key where the reference does not hash (pos + L >= n).
// The same value goes to tkeys (kept in time order; keys is consumed by the sort) and vals[e] = e.
// One CTA = 2048 positions: the text tile is staged in shared memory, warp x rolls pattern x (a lane owns 64
// consecutive positions), the 5 x 2048 bucket ids are transposed through shared memory and leave coalesced.
constexpr int GE_TILE = 2048, GE_SEG = 64, GE_HALO = 64;
__device__ __forceinline__ int ge_txt(int p) { return p + ((p >> 6) << 2); }          // 68-byte pitch per 64-byte segment
constexpr uint32_t NOENT = 0xFFFFFFFFu;       // bucket chains (below): no entry in front
constexpr uint32_t NOLINK = 0xFFFFFFFEu;      // bucket chains: the link of this entry has not been looked up yet
__device__ __forceinline__ int ge_out(int o) { return o + o / (5 * GE_SEG); }          // 321-word pitch per lane row
__global__ void __launch_bounds__(160) gap_entries_kernel(const uint8_t* __restrict__ T, uint64_t n, const ghp5 hp,
                                                          uint64_t a0, uint64_t cnt, uint32_t* __restrict__ keys,
                                                          uint32_t* __restrict__ tkeys, uint32_t* __restrict__ vals,
                                                          const uint32_t* __restrict__ skip,
                                                          const uint32_t* __restrict__ tile_off, uint32_t* __restrict__ prevs) {
    __shared__ __align__(16) uint8_t sT[(GE_TILE + GE_HALO) / 64 * 68 + 16];
    __shared__ uint32_t sOut[5 * GE_TILE + GE_TILE / GE_SEG + 8];
    __shared__ uint32_t sNeed[GE_TILE / 32], sRank[GE_TILE / 32];
    const int tid = threadIdx.x, lane = tid & 31, x = tid >> 5;          // warp x <-> pattern x
    const uint64_t t0 = (uint64_t)blockIdx.x * GE_TILE;                   // tile start, slice-relative
    const uint64_t g0 = a0 + t0;                                          // text position of the tile
    const uint32_t in_tile = cnt - t0 < (uint64_t)GE_TILE ? (uint32_t)(cnt - t0) : (uint32_t)GE_TILE;
    if (skip && tile_off[blockIdx.x + 1] == tile_off[blockIdx.x]) return;   // sparse slice: no position of this tile needs entries
    // stage GE_TILE + GE_HALO text bytes (the device text is zero-padded behind n); g0 is 16-byte aligned
    // whenever a0 is (slices start at multiples of the slice length); fall back to byte loads otherwise
    if ((g0 & 15) == 0) {
        const uint4* src = reinterpret_cast<const uint4*>(T + g0);
        for (int v = tid; v < (GE_TILE + GE_HALO) / 16; v += 160) {
            const uint4 q = src[v];
            const int p = v * 16;                                         // 16 bytes never straddle a 64-byte segment
            uint32_t* d = reinterpret_cast<uint32_t*>(sT + ge_txt(p));
            d[0] = q.x; d[1] = q.y; d[2] = q.z; d[3] = q.w;
        }
    } else {
        for (int p = tid; p < GE_TILE + GE_HALO; p += 160) sT[ge_txt(p)] = T[g0 + p];
    }
    __syncthreads();
    {
        const uint64_t L = hp.L[x];
        const uint64_t base = hp.base[x];
        const m107* __restrict__ nbt = hp.nbtab + x * 256;
        const int r0 = lane * GE_SEG;
        const uint64_t pos0 = g0 + r0;
        m107 fp;
        fp.lo = fp.hi = 0;
        if (pos0 + L < n)
            for (int j = 0; j < (int)L; j++) {
                m107 c;
                c.lo = sT[ge_txt(r0 + j)]; c.hi = 0;
                fp = muladd107(fp, base, c);
            }
        for (int j = 0; j < GE_SEG; j++) {
            const int r = r0 + j;
            const uint64_t pos = g0 + r;
            uint32_t key = hp.invalid_key;
            if (pos + L < n) {
                key = (uint32_t)(fp.lo & hp.mask);
                m107 infl = nbt[sT[ge_txt(r)]];                           // -(base^L) * T[pos]
                const uint64_t in = sT[ge_txt(r + (int)L)];               // T[pos + L], pos + L < n
                infl.lo += in;
                infl.hi += infl.lo < in;
                infl = fold107(infl.lo, infl.hi);
                fp = muladd107(fp, base, infl);
            }
            sOut[ge_out(5 * r + (4 - x))] = key;
        }
    }
    __syncthreads();
    const uint64_t e0 = 5 * t0;
    if (!skip) {
        for (uint32_t o = tid; o < 5 * in_tile; o += 160) {
            const uint32_t key = sOut[ge_out(o)];
            if (keys) keys[e0 + o] = key;
            tkeys[e0 + o] = key;
            prevs[e0 + o] = NOLINK;
            if (vals) vals[e0 + o] = (uint32_t)(e0 + o);
        }
        return;
    }
    // sparse slice: only the positions whose skip bit is clear get entries; the sort sees them compacted (in time
    // order), the time-order copy keeps its dense index so that a probe finds the buckets of its position directly
    if (tid < GE_TILE / 32) {
        const uint64_t wi = t0 / 32 + tid;
        uint32_t need = (uint64_t)tid * 32 < in_tile ? ~skip[wi] : 0u;
        if (((uint64_t)tid + 1) * 32 > in_tile && (uint64_t)tid * 32 < in_tile) need &= (1u << (in_tile & 31)) - 1u;
        sNeed[tid] = need;
    }
    __syncthreads();
    if (tid == 0) {
        uint32_t acc = 0;
        for (int k = 0; k < GE_TILE / 32; k++) { sRank[k] = acc; acc += __popc(sNeed[k]); }
    }
    __syncthreads();
    const uint64_t c0 = 5ull * tile_off[blockIdx.x];
    for (uint32_t o = tid; o < 5 * in_tile; o += 160) {
        const uint32_t r = o / 5u, j = o - 5u * r;
        const uint32_t need = sNeed[r >> 5];
        if (!((need >> (r & 31)) & 1u)) continue;
        const uint32_t rank = sRank[r >> 5] + __popc(need & ((1u << (r & 31)) - 1u));
        const uint32_t key = sOut[ge_out(o)];
        keys[c0 + 5ull * rank + j] = key;
        vals[c0 + 5ull * rank + j] = (uint32_t)(e0 + o);
        tkeys[e0 + o] = key;
        prevs[e0 + o] = NOLINK;
    }
}

// Sparse slice, few positions need entries (a few percent on repetitive texts): the dense kernel above rolls every window of
// every tile.  Here one CTA takes SP_TILE = 16384 positions, lists the 64-position segments that hold a needed position, and
// a lane rolls one listed segment (warp x <-> pattern x as above) from its first to its last needed position only.  Output
// as above: keys / vals compacted in time order at 5 * (needed positions in front), tkey at the dense index.
constexpr int SP_TILE = 16384, SP_SEGS = SP_TILE / GE_SEG, SP_WORDS = SP_TILE / 32;
__global__ void __launch_bounds__(160) gap_entries_sparse_kernel(const uint8_t* __restrict__ T, uint64_t n, const ghp5 hp,
                                                                 uint64_t a0, uint64_t cnt, uint32_t* __restrict__ keys,
                                                                 uint32_t* __restrict__ tkeys, uint32_t* __restrict__ vals,
                                                                 const uint32_t* __restrict__ skip,
                                                                 const uint32_t* __restrict__ tile_off, uint32_t* __restrict__ prevs) {
    __shared__ uint32_t sNeed[SP_WORDS], sRank[SP_WORDS + 1];
    __shared__ uint16_t sAct[SP_SEGS];
    __shared__ uint32_t s_nact;
    // bucket ids of the 32 segments of one iteration, [segment][5 * position + entry slot] (pitch 321: no bank conflicts):
    // the five pattern warps fill it, then all threads write a segment's entries as consecutive words
    __shared__ uint32_t sK[32][5 * GE_SEG + 1];
    const int tid = threadIdx.x, lane = tid & 31, x = tid >> 5;
    const uint64_t t0 = (uint64_t)blockIdx.x * SP_TILE;                   // tile start, slice-relative
    const uint32_t in_tile = cnt - t0 < (uint64_t)SP_TILE ? (uint32_t)(cnt - t0) : (uint32_t)SP_TILE;
    constexpr int SUB = SP_TILE / GE_TILE;                                // tile_off counts per GE_TILE positions
    const uint32_t first = tile_off[(uint64_t)blockIdx.x * SUB];
    const uint64_t lastsub = min((uint64_t)(blockIdx.x + 1) * SUB, (cnt + GE_TILE - 1) / GE_TILE);
    if (tile_off[lastsub] == first) return;                               // nothing needed in this tile
    if (tid == 0) s_nact = 0;
    for (int k = tid; k < SP_WORDS; k += 160) {
        const uint64_t p0 = (uint64_t)k * 32;
        uint32_t need = p0 < in_tile ? ~skip[t0 / 32 + k] : 0u;
        if (p0 < in_tile && p0 + 32 > in_tile) need &= (1u << (in_tile & 31)) - 1u;
        sNeed[k] = need;
    }
    __syncthreads();
    if (x == 0) {                                                         // exclusive prefix of the popcounts, one warp
        uint32_t run = 0;
        for (int c = 0; c < SP_WORDS; c += 32) {
            const uint32_t v = __popc(sNeed[c + lane]);
            uint32_t inc = v;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d) inc += y;
            }
            sRank[c + lane] = run + inc - v;
            run += __shfl_sync(0xffffffffu, inc, 31);
        }
    } else if (x == 1) {                                                  // list of the segments with a needed position
        for (int sg = lane; sg < SP_SEGS; sg += 32)
            if (sNeed[2 * sg] | sNeed[2 * sg + 1]) sAct[atomicAdd(&s_nact, 1u)] = (uint16_t)sg;
    }
    __syncthreads();
    const uint32_t nact = s_nact;
    const uint64_t L = hp.L[x], base = hp.base[x];
    const m107* __restrict__ nbt = hp.nbtab + x * 256;
    const uint64_t c0 = 5ull * first, e0 = 5 * t0;
    for (uint32_t abase = 0; abase < nact; abase += 32) {
      const uint32_t ai = abase + lane;
      if (ai < nact) {
        const int sg = sAct[ai];
        const uint64_t need = (uint64_t)sNeed[2 * sg] | ((uint64_t)sNeed[2 * sg + 1] << 32);
        const int lo = __ffsll((long long)need) - 1, hi = 63 - __clzll((long long)need);
        const uint8_t* __restrict__ tp = T + a0 + t0 + (uint64_t)sg * GE_SEG;          // the device text is zero-padded behind n
        const uint64_t pos0 = a0 + t0 + (uint64_t)sg * GE_SEG;
        m107 fp;
        fp.lo = fp.hi = 0;
        bool live = false;                                                // fp holds the window at the current position
        for (int j = lo; j <= hi; j++) {
            const uint64_t pos = pos0 + j;
            const bool fits = pos + L < n;
            if (fits && !live) {
                fp.lo = fp.hi = 0;
                for (uint64_t q = 0; q < L; q++) {
                    m107 c;
                    c.lo = tp[j + q]; c.hi = 0;
                    fp = muladd107(fp, base, c);
                }
                live = true;
            }
            if ((need >> j) & 1ull) sK[lane][5 * j + (4 - x)] = fits ? (uint32_t)(fp.lo & hp.mask) : hp.invalid_key;
            if (fits && j < hi) {
                if (!((need >> (j + 1)) & 1ull) && __popcll(need >> (j + 1)) && __ffsll((long long)(need >> (j + 1))) - 1 > (int)L) {
                    live = false;                                         // a long stretch without needed positions: start over
                    j += __ffsll((long long)(need >> (j + 1))) - 1;       // (the loop's j++ lands on the next needed one)
                    continue;
                }
                m107 infl = nbt[tp[j]];                                   // -(base^L) * T[pos]
                const uint64_t in = tp[j + L];
                infl.lo += in;
                infl.hi += infl.lo < in;
                infl = fold107(infl.lo, infl.hi);
                fp = muladd107(fp, base, infl);
            }
        }
      }
      __syncthreads();
      // write-out, segment by segment: thread t takes the slots t, t + 160 of the segment's 5 x 64; the entries of a run of
      // needed positions are consecutive words in all four arrays
      const uint32_t nseg = nact - abase < 32u ? nact - abase : 32u;
      for (uint32_t sl = 0; sl < nseg; sl++) {
          const int sg = sAct[abase + sl];
          const uint64_t need = (uint64_t)sNeed[2 * sg] | ((uint64_t)sNeed[2 * sg + 1] << 32);
          const uint32_t rank0 = sRank[2 * sg];
          for (int o = tid; o < 5 * GE_SEG; o += 160) {
              const int j = o / 5, y = o - 5 * j;
              if (!((need >> j) & 1ull)) continue;
              const uint32_t r = rank0 + (uint32_t)__popcll(need & ((1ull << j) - 1ull));
              const uint32_t key = sK[sl][o];
              const uint64_t od = 5ull * (uint64_t)(sg * GE_SEG) + (uint64_t)o;
              keys[c0 + 5ull * r + y] = key;
              vals[c0 + 5ull * r + y] = (uint32_t)(e0 + od);
              tkeys[e0 + od] = key;
              prevs[e0 + od] = NOLINK;
          }
      }
      __syncthreads();
    }
}

// positions of the slice that need no entry when no gap factor runs over a whole phrase: the inside (beg, end) of
// every phrase (the phrase start is probed, greedy.cpp:111-117).  One warp per phrase k0 + w.
__global__ void __launch_bounds__(128) skip_bits_kernel(const uint64_t* __restrict__ ml_beg, const uint64_t* __restrict__ ml_end,
                                                        uint64_t k0, uint64_t k1, uint64_t a0, uint64_t a1,
                                                        uint32_t* __restrict__ bm, int adjacent_starts);

// needed positions per tile of GE_TILE positions
__global__ void tile_need_kernel(const uint32_t* __restrict__ skip, uint64_t cnt, uint32_t ntiles, uint32_t* __restrict__ out) {
    const uint32_t t = (uint32_t)(((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const unsigned lane = threadIdx.x & 31;
    if (t >= ntiles) return;
    uint32_t c = 0;
    for (int k = lane; k < GE_TILE / 32; k += 32) {
        const uint64_t p0 = (uint64_t)t * GE_TILE + (uint64_t)k * 32;
        if (p0 >= cnt) continue;
        uint32_t need = ~skip[p0 / 32];
        if (p0 + 32 > cnt) need &= (1u << (cnt - p0)) - 1u;
        c += __popc(need);
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    if (lane == 0) out[t] = c;
}

// bstart[b] = first sorted index whose key is >= b, for b in [0, nkeys]  (nkeys = table_entries + 1 keys incl. invalid)
__global__ void __launch_bounds__(256) bucket_start_kernel(const uint32_t* __restrict__ skey, uint64_t E, uint32_t nkeys,
                                                           uint32_t* __restrict__ bstart) {
    // a thread takes 4 consecutive keys (one 16-byte load; the sort buffers are 16-byte aligned); the key in front of
    // them comes from the lane below
    const uint64_t k0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const unsigned lane = threadIdx.x & 31;
    uint32_t key[4] = {0, 0, 0, 0};
    if (k0 + 3 < E) {
        const uint4 v = *reinterpret_cast<const uint4*>(skey + k0);
        key[0] = v.x; key[1] = v.y; key[2] = v.z; key[3] = v.w;
    } else {
        for (int j = 0; j < 4; j++)
            if (k0 + j < E) key[j] = skey[k0 + j];
    }
    uint32_t below = __shfl_up_sync(0xffffffffu, key[3], 1);
    if (lane == 0 && k0 > 0 && k0 < E) below = skey[k0 - 1];
    int64_t prev = k0 ? (int64_t)below : -1;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const uint64_t k = k0 + j;
        if (k >= E) break;
        for (int64_t b = prev + 1; b <= (int64_t)key[j]; b++) bstart[b] = (uint32_t)k;
        if (k == E - 1)
            for (uint32_t b = key[j] + 1; b <= nkeys; b++) bstart[b] = (uint32_t)E;
        prev = key[j];
    }
}

// rstart[b] = first index of the (sorted) big repair list whose bucket is >= b, for b in [0, nkeys]
__global__ void __launch_bounds__(256) repair_start_kernel(const uint64_t* __restrict__ rk, uint32_t R, uint32_t nkeys,
                                                           uint32_t* __restrict__ rstart) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= R) return;
    const uint32_t cur = (uint32_t)(rk[k] >> 32);
    const int64_t prev = k ? (int64_t)(uint32_t)(rk[k - 1] >> 32) : -1;
    for (int64_t b = prev + 1; b <= (int64_t)cur && b <= (int64_t)nkeys; b++) rstart[b] = k;
    if (k == R - 1)
        for (uint64_t b = (uint64_t)cur + 1; b <= nkeys; b++) rstart[b] = R;
}

// Bucket chains: prev[e] = the entry in front of e in its bucket's list (time order = entry id order) over the sorted
// index and both repair lists, NOENT for the first, NOLINK while nobody has asked.  A probe walks the chain from its own
// entry; a missing link is looked up by bisecting the three lists (raw_pred below) and stored, so that the later rounds,
// which probe nearly the same positions, find it with one load.  The buckets of the short patterns (2 and 5 bytes) hold
// millions of entries, and every level of a bisection is a dependent load on the critical path of a block.
// (Linking every entry up front, one scattered 4-byte store per sorted entry, was measured: 24.5 ms per GiB for 660 M
// entries, as much as the probes gained; a round looks up some 20 M links.)
// Splices the sorted repair list L (bucket << 32 | entry id; all of it, the step is idempotent) into the chains.  The chains
// hold the sorted index and, when L is the small list, the big list (big, rstart).  One thread per entry: its own link, and
// the link of the entry behind it if no other entry of L lies in between.
// (One entry per thread.  Threads that take 8 consecutive entries and search from where their last search ended were measured:
// 11.4 against 7.0 ms per GiB - the lanes of a warp bisect towards neighbouring targets and share all but the last loads.)
__global__ void __launch_bounds__(256) splice_repair_kernel(const uint32_t* __restrict__ sval, const uint32_t* __restrict__ bstart,
                                                            const uint64_t* __restrict__ L, uint32_t cnt,
                                                            const uint64_t* __restrict__ big, const uint32_t* __restrict__ rstart,
                                                            uint32_t* __restrict__ prev) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= cnt) return;
    const uint64_t x = L[k];
    const uint32_t b = (uint32_t)(x >> 32), e = (uint32_t)x;
    int64_t pred = -1;
    uint32_t succ = NOENT;
    {
        const uint32_t lo = bstart[b], hi = bstart[b + 1];
        uint32_t l = lo, h = hi;
        while (l < h) {
            const uint32_t mid = (l + h) >> 1;
            if (sval[mid] < e) l = mid + 1;
            else h = mid;
        }
        if (l > lo) pred = sval[l - 1];
        if (l < hi) succ = sval[l];
    }
    if (big) {
        const uint32_t lo = rstart[b], hi = rstart[b + 1];
        uint32_t l = lo, h = hi;
        while (l < h) {
            const uint32_t mid = (l + h) >> 1;
            if (big[mid] < x) l = mid + 1;
            else h = mid;
        }
        if (l > lo) pred = max(pred, (int64_t)(uint32_t)big[l - 1]);
        if (l < hi) succ = min(succ, (uint32_t)big[l]);
    }
    if (k > 0 && (uint32_t)(L[k - 1] >> 32) == b) pred = max(pred, (int64_t)(uint32_t)L[k - 1]);
    uint32_t own_next = NOENT;
    if (k + 1 < cnt && (uint32_t)(L[k + 1] >> 32) == b) own_next = (uint32_t)L[k + 1];
    prev[e] = pred < 0 ? NOENT : (uint32_t)pred;
    if (succ != NOENT && own_next > succ) prev[succ] = e;
}

struct block_state {
    uint64_t i, cur, k;
    uint32_t gap;      // last step before i was a gap step (positions up to i are inserted)
    uint32_t pad;
};

struct parse_ctx {
    uint64_t n, n_cut, a0, a1;
    uint32_t BL, nb;
    uint32_t bmin;             // blocks below it cannot change in this round (nothing before them changed in the last one)
    uint32_t naive;            // factorize_greedy_naive (greedy_naive.cpp:34-113): every position is inserted, gap
                               // factors stop at the next phrase, phrase-start probes stop at the phrase after it
    const uint32_t* tkey;      // bucket of every entry, time order
    const uint32_t* sval;      // entries sorted by bucket, time order inside a bucket
    const uint32_t* bstart;    // bucket b = sval[bstart[b] .. bstart[b+1])
    uint32_t* prev;            // bucket chains: the entry in front of an entry in its bucket over the index and the repair lists
                               // (NOLINK until a probe has looked it up)
    const uint64_t* Hc;        // last inserted position + 1 per bucket (0 = none), state before the slice
    const uint32_t* notins;    // bit = 1: position of the slice is NOT inserted
    const uint32_t* skip;      // sparse slice: bit = 1: the position has no entries (nullptr: every position has)
    const uint64_t* rk;        // sparse slice: entries added after the sort (repair list), (bucket << 32 | entry id) ascending
    uint32_t R;                // number of entries in the repair list
    const uint32_t* rstart;    // bucket b of the big repair list = rk[rstart[b] .. rstart[b+1])
    const uint64_t* rk2;       // second, small repair list: what later rounds add (re-sorted alone; merged when it grows)
    uint32_t R2;
    uint32_t* req;             // sparse slice: req[0] = number of repair requests of this round, then (beg, end) pairs
    uint32_t req_cap;          //   (slice-relative positions that a parse needs entries for)
    uint8_t* ovr;              // per block: its parse needed entries that a sparse slice does not have (yet)
    uint8_t* dirty;            // per block: an input of its parse (entry state, a bitmap bit or a table entry that one
                               // of its probes can have read) changed since it was parsed
    const uint64_t* ml_beg;
    const uint64_t* ml_end;
    const uint64_t* ml_src;
    uint64_t m;
    block_state* entry;
    block_state* exit_;
    uint8_t* marked;
    uint32_t* tr_cnt;
    uint32_t* tr_pos;
    uint64_t* tr_src;
    uint64_t* tr_len;
    uint32_t* changed;
    uint32_t dbg;              // count link lookups and chain hops in changed[8..11]
};

__device__ __forceinline__ bool not_inserted(const uint32_t* __restrict__ bm, uint32_t rel) {
    return (bm[rel >> 5] >> (rel & 31)) & 1u;
}

// sparse slice: a parse needs entries for the positions [s, e) of the text (clipped to the slice)
__device__ __forceinline__ void file_request(const parse_ctx& C, uint64_t s, uint64_t e) {
    if (e > C.a1) e = C.a1;
    if (s >= e) return;
    const uint32_t k = atomicAdd(C.req, 1u);
    if (k < C.req_cap) {
        C.req[2 + 2 * k] = (uint32_t)(s - C.a0);
        C.req[3 + 2 * k] = (uint32_t)(e - C.a0);
    }
}
// does any position of [s, e) (text positions inside the slice) lack entries?  All lanes of the group.
template <int G>
__device__ __forceinline__ bool group_any_skip(const parse_ctx& C, uint64_t s, uint64_t e, unsigned gmask, unsigned glane) {
    if (e > C.a1) e = C.a1;
    if (s >= e) return false;
    const uint64_t rs = s - C.a0, re = e - C.a0;
    const uint64_t w0 = rs >> 5, w1 = (re - 1) >> 5;
    bool any = false;
    for (uint64_t w = w0 + glane; w <= w1 && !any; w += G) {
        uint32_t m = 0xFFFFFFFFu;
        if (w == w0) m &= 0xFFFFFFFFu << (rs & 31);
        if (w == w1) m &= 0xFFFFFFFFu >> (31 - ((re - 1) & 31));
        any = (C.skip[w] & m) != 0;
    }
    return __ballot_sync(gmask, any) != 0;
}

// lower bound of key in the ascending arr[lo, hi), searched outwards from g, the result of this lane's last search in the
// same list (any value is safe): consecutive probes of a block ask for neighbouring entries, so the doubling steps find
// the bracket within a few dependent loads where a bisection of a 2-byte pattern's bucket takes 23.
template <typename T>
__device__ __forceinline__ uint32_t hinted_lower_bound(const T* __restrict__ arr, uint32_t lo, uint32_t hi, T key, uint32_t g) {
    uint32_t l = lo, h = hi;
    if (g >= lo && g <= hi && lo < hi) {
        if (g < hi && arr[g] < key) {                 // the answer lies behind g
            uint32_t step = 1;
            l = g + 1;
            while (l + step <= hi && arr[l + step - 1] < key) { l += step; step <<= 1; }
            if (l + step <= hi) h = l + step - 1;
        } else {                                      // arr[g] >= key (or g == hi): the answer is g or lies in front of it
            uint32_t step = 1;
            h = g;
            while (h >= lo + step && arr[h - step] >= key) { h -= step; step <<= 1; }
            if (h >= lo + step) l = h - step + 1;
        }
    }
    while (l < h) {
        const uint32_t mid = (l + h) >> 1;
        if (arr[mid] < key) l = mid + 1;
        else h = mid;
    }
    return l;
}

// the largest entry id below e in bucket b over the sorted index and both repair lists (NOENT: none).
// hint: this lane's last positions in the three lists (shared memory), or nullptr.
__device__ __forceinline__ uint32_t raw_pred(const parse_ctx& C, uint32_t e, uint32_t b, uint32_t* hint) {
    const uint64_t key = ((uint64_t)b << 32) | e;
    int64_t pred = -1;
    {
        const uint32_t lo = C.bstart[b], hi = C.bstart[b + 1];
        const uint32_t l = hinted_lower_bound<uint32_t>(C.sval, lo, hi, e, hint ? hint[0] : NOENT);
        if (hint) hint[0] = l;
        if (l > lo) pred = C.sval[l - 1];
    }
    if (C.R) {
        const uint32_t lo = C.rstart[b], hi = C.rstart[b + 1];
        const uint32_t l = hinted_lower_bound<uint64_t>(C.rk, lo, hi, key, hint ? hint[1] : NOENT);
        if (hint) hint[1] = l;
        if (l > lo) pred = max(pred, (int64_t)(uint32_t)C.rk[l - 1]);
    }
    if (C.rk2 && C.R2) {
        const uint32_t l = hinted_lower_bound<uint64_t>(C.rk2, 0, C.R2, key, hint ? hint[2] : NOENT);
        if (hint) hint[2] = l;
        if (l > 0) {
            const uint64_t v = C.rk2[l - 1];
            if ((uint32_t)(v >> 32) == b) pred = max(pred, (int64_t)(uint32_t)v);
        }
    }
    return pred < 0 ? NOENT : (uint32_t)pred;
}

// longest_prev_occ (factorize/common.cpp:33-61) at position i, all lanes of the group; its lanes 0..4 own patterns 4..0.
// "H[b] when i is probed" = the largest inserted position among the entries of bucket b in front of the probe's own
// entry, found along the bucket's chain (sorted index and repair lists in one list, time order), else the carried table.
template <typename pos_t, int G>
__device__ __forceinline__ void probe(const parse_ctx& C, const lce_view<pos_t>& V, uint64_t i, uint64_t& f_src,
                                      uint64_t& f_len, bool& bad, unsigned gmask, unsigned lane, uint32_t* hint) {
    const uint32_t rel = (uint32_t)(i - C.a0);
    if (C.skip && ((C.skip[rel >> 5] >> (rel & 31)) & 1u)) {       // a wrong speculation, or the position a gap factor that
        bad = true;                                                 // ran over whole phrases lands on: ask for its entries
        if (lane == 0) file_request(C, i, i + 1);
        f_src = V.T[i];
        f_len = 0;
        return;
    }
    uint64_t occ = NONE64;
    if (lane < 5) {
        const uint32_t e = 5u * rel + lane;
        const uint32_t b = C.tkey[e];                 // (the lookup of a missing link and the fallback need it)
        uint32_t e1 = C.prev[e];
        if (e1 == NOLINK) {
            e1 = raw_pred(C, e, b, hint);
            C.prev[e] = e1;
            if (C.dbg) atomicAdd(C.changed + 8, 1u);
        }
        if (C.dbg) atomicAdd(C.changed + 11, 1u);
        bool found = false;
        while (e1 != NOENT) {
            const uint32_t q = e1 / 5u;
            if (q == rel) { occ = i; found = true; break; }        // an earlier pattern of this position wrote the bucket
            uint32_t nxt = C.prev[e1];                // next link, in flight together with the bitmap word
            if (!not_inserted(C.notins, q)) { occ = C.a0 + q; found = true; break; }
            if (C.dbg) atomicAdd(C.changed + 10, 1u);
            if (nxt == NOLINK) {
                nxt = raw_pred(C, e1, b, hint);
                C.prev[e1] = nxt;
                if (C.dbg) atomicAdd(C.changed + 9, 1u);
            }
            e1 = nxt;
        }
        if (!found) occ = C.Hc[b] - 1;
    }
    const uint8_t ci = V.T[i];
    const bool valid = lane < 5 && occ != NONE64 && occ < i && V.T[occ] == ci;
    const unsigned m = __ballot_sync(gmask, valid);
    if (!m) {
        f_src = ci;
        f_len = 0;
        return;
    }
    const int win = __ffs(m) - 1;                       // (a warp lane: the group's lowest lane with a match)
    f_src = __shfl_sync(gmask, occ, win);
    f_len = group_lce_r<pos_t, G>(V, f_src, i, gmask, lane);
}

// G lanes of a warp parse one block (G = 32: a warp per block, the default; G = 8: four blocks per warp, an experiment: more
// blocks in flight per SM did not shorten the rounds).
template <typename pos_t, int G>
__global__ void __launch_bounds__(128, G == 32 ? 16 : 12) parse_kernel(const parse_ctx C, const lce_view<pos_t> V) {
    const uint32_t b = (uint32_t)(((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) / G);
    const unsigned lane = threadIdx.x & (G - 1);                       // lane inside the group
    const unsigned gmask = G == 32 ? 0xffffffffu : (((1u << G) - 1u) << (threadIdx.x & 31 & ~(G - 1)));
    __shared__ uint32_t s_hint[128 / G][5][3];      // per pattern lane: where its last lookups ended in the three lists
    if (lane < 5) s_hint[(threadIdx.x & 127) / G][lane][0] = s_hint[(threadIdx.x & 127) / G][lane][1] = s_hint[(threadIdx.x & 127) / G][lane][2] = NOENT;
    uint32_t* hint = s_hint[(threadIdx.x & 127) / G][lane < 5 ? lane : 0];
    if (b >= C.nb || b < C.bmin || !C.marked[b] || !C.dirty[b]) return;
    const block_state st = C.entry[b];
    uint64_t i = st.i, cur = st.cur, k = st.k;
    uint32_t gap = st.gap;
    uint64_t bend = C.a0 + (uint64_t)(b + 1) * C.BL;
    if (bend > C.a1) bend = C.a1;
    const uint64_t lim = bend < C.n_cut ? bend : C.n_cut;
    const uint32_t oldlen = C.tr_cnt[b];
    const uint64_t slot0 = (uint64_t)b * C.BL;
    uint32_t t = 0;
    bool chg = false, bad = false;
    while (i < lim) {
        const uint64_t pb = C.ml_beg[k], pe = C.ml_end[k];
        uint64_t o_src, o_len;
        uint32_t kind;
        const uint64_t at = i;
        if (i < pb) {   // gap step (greedy.cpp:60-97, greedy_naive.cpp:56-67)
            probe<pos_t, G>(C, V, i, o_src, o_len, bad, gmask, lane, hint);
            if (C.naive && o_len > pb - i) o_len = pb - i;
            uint64_t i2 = i + (o_len ? o_len : 1);
            if (i2 > pb) {
                if (i2 <= pe) {
                    o_len -= i2 - pb;
                    i2 = pb;
                } else {
                    // the positions of the phrases the factor ran over get inserted, and the position it lands on is probed
                    // if it lies inside a phrase (greedy.cpp:68-77,111-117): a sparse slice may lack their entries
                    if (C.skip && group_any_skip<G>(C, pb, i2 + 1, gmask, lane)) {
                        bad = true;
                        if (lane == 0) file_request(C, pb, i2 + 1);
                    }
                    while (C.ml_end[k] <= i2) k++;
                }
            }
            kind = 0;
            cur = i2;
            i = i2;
            gap = 1;
        } else {        // phrase step (greedy.cpp:100-131)
            const uint64_t exc = i - pb;
            o_src = C.ml_src[k] + exc;
            o_len = (pe - pb) - exc;
            kind = KIND_PHR;
            if (cur == i) {
                uint64_t fs, fl;
                probe<pos_t, G>(C, V, i, fs, fl, bad, gmask, lane, hint);
                cur = i + 1;
                kind |= KIND_PROBED;
                if (C.naive && k + 1 < C.m) {            // greedy_naive.cpp:82-85: clipped to the next phrase start
                    const uint64_t room = C.ml_beg[k + 1] > i ? C.ml_beg[k + 1] - i : 0;
                    if (fl > room) fl = room;
                }
                if (fl > o_len) { o_src = fs; o_len = fl; }
            }
            i += o_len;
            while (k + 1 < C.m && C.ml_end[k] <= i) k++;
            gap = 0;
        }
        if (lane == 0) {
            const uint64_t sl = slot0 + t;
            const uint32_t np = (uint32_t)(at - C.a0) | kind;
            if (t >= oldlen || C.tr_pos[sl] != np || C.tr_src[sl] != o_src || C.tr_len[sl] != o_len) chg = true;
            C.tr_pos[sl] = np;
            C.tr_src[sl] = o_src;
            C.tr_len[sl] = o_len;
        }
        t++;
    }
    if (lane == 0) {
        const block_state old = C.exit_[b];
        if (t != oldlen || old.i != i || old.cur != cur || old.k != k || old.gap != gap) chg = true;
        block_state ex;
        ex.i = i; ex.cur = cur; ex.k = k; ex.gap = gap; ex.pad = 0;
        C.exit_[b] = ex;
        C.tr_cnt[b] = t;
        C.ovr[b] = bad ? 1 : 0;
        C.dirty[b] = bad ? 1 : 0;          // a block that lacked entries is parsed again once they exist
        if (chg) {
            atomicOr(C.changed, 1u);
            atomicAdd(C.changed + 3, 1u);
            atomicMin(C.changed + 4, b);
        }
    }
}

// speculative entry of every block: a factor starts at its first position, the table cursor stands there
__global__ void spec_entry_kernel(const parse_ctx C, block_state first) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= C.nb) return;
    const uint64_t bs = C.a0 + (uint64_t)b * C.BL;
    uint64_t bend = bs + C.BL;
    if (bend > C.a1) bend = C.a1;
    block_state st;
    st.pad = 0;
    C.tr_cnt[b] = 0;
    C.dirty[b] = 1;
    block_state z;
    z.i = z.cur = z.k = NONE64; z.gap = 0; z.pad = 0;
    C.exit_[b] = z;
    if (first.i >= bend) {          // the chain enters the slice behind this block
        C.marked[b] = 0;
        return;
    }
    C.marked[b] = 1;
    if (first.i >= bs) {
        st = first;
    } else {
        st.i = bs; st.cur = bs; st.gap = 1;
        uint64_t lo = 0, hi = C.m - 1;   // first phrase with end > bs (the sentinel ends at n+1)
        while (lo < hi) {
            const uint64_t mid = (lo + hi) / 2;
            if (C.ml_end[mid] > bs) hi = mid;
            else lo = mid + 1;
        }
        // inside a phrase the true parse most likely arrives by taking that phrase to its end without touching
        // the table there (greedy.cpp:100-131)
        if (C.ml_beg[lo] < bs) {
            st.i = C.ml_end[lo];
            st.cur = C.ml_beg[lo] + 1 < st.i ? C.ml_beg[lo] + 1 : C.ml_beg[lo];
            st.gap = 0;
            while (lo + 1 < C.m && C.ml_end[lo] <= st.i) lo++;
        }
        st.k = lo;
    }
    C.entry[b] = st;
}

// Follows the chain of blocks from the block that holds `first.i` by pointer doubling, marks the blocks on it, hands
// every exit state to the block it lands in, and reports the state that leaves the slice.  One grid-wide launch per
// step (phase 0: successor of every block + reset of the marks; 1: J_l = J_{l-1} o J_{l-1}; 2: marked blocks mark
// their 2^l-th successor, levels descending; 3: entries): ~2 levels launches of a few microseconds instead of one CTA
// looping over all blocks (1.5 ms per round at 131072 blocks).
__global__ void __launch_bounds__(256) link_phase_kernel(const parse_ctx C, block_state first, uint32_t* __restrict__ J,
                                                         int l, int phase, block_state* __restrict__ slice_exit) {
    const uint32_t nb = C.nb, N = nb + 1;
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= N) return;
    const uint64_t stop = C.a1 < C.n_cut ? C.a1 : C.n_cut;
    const uint32_t b0 = first.i < stop ? (uint32_t)((first.i - C.a0) / C.BL) : nb;
    if (phase == 0) {
        uint32_t j = nb;
        if (b < nb) {
            const uint64_t x = C.exit_[b].i;
            if (x != NONE64 && x < stop) j = (uint32_t)((x - C.a0) / C.BL);
            C.marked[b] = (b == b0);
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
        auto assign = [&](uint32_t j, const block_state& v) {
            const block_state o = C.entry[j];
            if (o.i != v.i || o.cur != v.cur || o.k != v.k || o.gap != v.gap) {
                C.entry[j] = v;
                C.dirty[j] = 1;
            }
        };
        if (b == nb) {
            if (b0 == nb) *slice_exit = first;     // nothing to parse in this slice
            return;
        }
        if (b == b0) assign(b, first);
        if (C.marked[b]) {
            const uint32_t j = J[b];
            if (j < nb) assign(j, C.exit_[b]);
            else *slice_exit = C.exit_[b];
        }
    }
}

// Which blocks can see a changed bitmap bit: a probe walks its bucket list backwards until the first inserted
// entry, so the entries BEHIND a flipped position, up to and including the first one that is inserted before and
// after the change, are the only ones whose walk can reach it; of those, only the ones that are inserted in one of
// the two parses can be probes (advance_and_get_occ inserts the position it probes).
constexpr unsigned DB_WPW = 4;
__global__ void __launch_bounds__(256) dirty_from_bitmap_kernel(const parse_ctx C, const uint32_t* __restrict__ old_bm,
                                                                uint64_t nwords) {
    // one warp per DB_WPW bitmap words (flips cluster: few words per warp spread them over more warps); a changed word
    // is handled by the whole warp, one lane per flipped bit
    const uint64_t w0 = (((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * DB_WPW;
    const unsigned lane = threadIdx.x & 31;
    if (w0 >= nwords) return;
    const uint32_t* __restrict__ new_bm = C.notins;
    const uint64_t cnt = C.a1 - C.a0;
    uint32_t mine = 0;
    if (lane < DB_WPW && w0 + lane < nwords) mine = old_bm[w0 + lane] ^ new_bm[w0 + lane];
    unsigned todo = __ballot_sync(0xffffffffu, mine != 0);
    while (todo) {
        const int wl = __ffs(todo) - 1;
        todo &= todo - 1;
        const uint32_t x = __shfl_sync(0xffffffffu, mine, wl);
        if (!((x >> lane) & 1u)) continue;
        const uint64_t rel = (w0 + wl) * 32 + lane;
        if (rel >= cnt) continue;
        if (C.skip && ((C.skip[rel >> 5] >> (rel & 31)) & 1u)) continue;
        for (uint32_t pat = 0; pat < 5; pat++) {
            const uint32_t e = 5u * (uint32_t)rel + pat;
            const uint32_t b = C.tkey[e];
            uint32_t l = C.bstart[b], h = C.bstart[b + 1];
            if (l >= h) continue;
            const uint32_t hi = h;
            // e is in the list and buckets are hash values, so the entries of a list are spread evenly over time:
            // interpolate a few times, then bisect what is left
            {
                uint32_t lo = l, up = h - 1;
                uint32_t vlo = C.sval[lo], vup = C.sval[up];
                for (int it = 0; it < 4 && up - lo > 8 && vlo < e && e < vup; it++) {
                    uint32_t mid = lo + (uint32_t)(((uint64_t)(e - vlo) * (up - lo)) / (vup - vlo));
                    if (mid <= lo) mid = lo + 1;
                    if (mid >= up) mid = up - 1;
                    const uint32_t vm = C.sval[mid];
                    if (vm <= e) { lo = mid; vlo = vm; }
                    else { up = mid; vup = vm; }
                }
                if (vup <= e) lo = up;
                l = lo; h = up + 1;
            }
            while (l + 1 < h) {
                const uint32_t mid = (l + h) >> 1;
                if (C.sval[mid] <= e) l = mid;
                else h = mid;
            }
            // (the entries of a position of the repair list are not in this list: start at the first entry behind e)
            for (uint32_t kk = C.sval[l] <= e ? l + 1 : l; kk < hi; kk++) {
                const uint32_t q = C.sval[kk] / 5u;
                // a probe inserts its position: a position that neither parse inserts is not probed by either
                const bool oi = !not_inserted(old_bm, q), ni = !not_inserted(new_bm, q);
                if (oi || ni) C.dirty[q / C.BL] = 1;
                if (oi && ni) break;
            }
        }
        for (int which = 0; which < 2; which++) {        // the same walk over the repair lists
            const uint64_t* __restrict__ rk = which ? C.rk2 : C.rk;
            const uint32_t R = which ? C.R2 : C.R;
            if (!R) continue;
            for (uint32_t pat = 0; pat < 5; pat++) {
                const uint32_t e = 5u * (uint32_t)rel + pat;
                const uint32_t b = C.tkey[e];
                const uint64_t key = ((uint64_t)b << 32) | e;
                uint32_t rl = 0, rh = R;
                if (!which) { rl = C.rstart[b]; rh = C.rstart[b + 1]; }
                while (rl < rh) {
                    const uint32_t mid = (rl + rh) >> 1;
                    if (rk[mid] <= key) rl = mid + 1;
                    else rh = mid;
                }
                for (; rl < R; rl++) {
                    const uint64_t v = rk[rl];
                    if ((uint32_t)(v >> 32) != b) break;
                    const uint32_t q = (uint32_t)v / 5u;
                    const bool oi = !not_inserted(old_bm, q), ni = !not_inserted(new_bm, q);
                    if (oi || ni) C.dirty[q / C.BL] = 1;
                    if (oi && ni) break;
                }
            }
        }
    }
}

// The bitmap only reflects the probes of blocks that are on the chain: a block that joins the chain is parsed again
// whatever its flag says.
__global__ void joined_chain_kernel(const parse_ctx C, const uint8_t* __restrict__ marked_prev) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < C.nb && C.marked[b] && !marked_prev[b]) C.dirty[b] = 1;
}

// Same for a changed table entry (a resumed slice of a sharded run): the walks that fall through the whole list of
// the bucket start at its first entries.
__global__ void dirty_from_table_kernel(const parse_ctx C, const unsigned long long* __restrict__ old_tab,
                                        const uint32_t* __restrict__ old_bm, uint32_t nbuckets) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nbuckets || old_tab[b] == C.Hc[b]) return;
    const uint32_t hi = C.bstart[b + 1];
    for (uint32_t kk = C.bstart[b]; kk < hi; kk++) {
        const uint32_t q = C.sval[kk] / 5u;
        const bool oi = !not_inserted(old_bm, q), ni = !not_inserted(C.notins, q);
        if (oi || ni) C.dirty[q / C.BL] = 1;
        if (oi && ni) break;
    }
    for (int which = 0; which < 2; which++) {            // the bucket's entries in the repair lists, from their start
        const uint64_t* __restrict__ rk = which ? C.rk2 : C.rk;
        const uint32_t R = which ? C.R2 : C.R;
        if (!R) continue;
        uint32_t rl = 0, rh = R;
        if (!which) {
            rl = C.rstart[b];
        } else {
            const uint64_t key = (uint64_t)b << 32;
            while (rl < rh) {
                const uint32_t mid = (rl + rh) >> 1;
                if (rk[mid] < key) rl = mid + 1;
                else rh = mid;
            }
        }
        for (; rl < R; rl++) {
            const uint64_t v = rk[rl];
            if ((uint32_t)(v >> 32) != b) break;
            const uint32_t q = (uint32_t)v / 5u;
            const bool oi = !not_inserted(old_bm, q), ni = !not_inserted(C.notins, q);
            if (oi || ni) C.dirty[q / C.BL] = 1;
            if (oi && ni) break;
        }
    }
}

__device__ __forceinline__ void set_bits(uint32_t* __restrict__ bm, uint64_t s, uint64_t e, unsigned lane) {   // [s, e), warp
    if (s >= e) return;
    const uint64_t w0 = s >> 5, w1 = (e - 1) >> 5;
    for (uint64_t w = w0 + lane; w <= w1; w += 32) {
        uint32_t m = 0xFFFFFFFFu;
        if (w == w0) m &= 0xFFFFFFFFu << (s & 31);
        if (w == w1) m &= 0xFFFFFFFFu >> (31 - ((e - 1) & 31));
        if (w == w0 || w == w1) atomicOr(&bm[w], m);
        else bm[w] = m;
    }
}

// not-inserted bitmap of the slice from the traces of the blocks on the chain: a phrase step at i of length l leaves
// (i, i+l) untouched, and i itself unless the table was probed there (greedy.cpp:111-117)
__global__ void __launch_bounds__(128) notins_kernel(const parse_ctx C, uint32_t* __restrict__ bm, block_state first) {
    const uint32_t b = (uint32_t)(((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const unsigned lane = threadIdx.x & 31;
    if (b >= C.nb) return;
    const uint64_t len = C.a1 - C.a0;
    if (b == 0 && !first.gap && first.i > C.a0) set_bits(bm, 0, (first.i < C.a1 ? first.i : C.a1) - C.a0, lane);
    if (!C.marked[b]) return;
    const uint32_t cnt = C.tr_cnt[b];
    const uint64_t slot0 = (uint64_t)b * C.BL;
    for (uint32_t t = 0; t < cnt; t++) {
        const uint32_t p = C.tr_pos[slot0 + t];
        if (!(p & KIND_PHR)) continue;
        const uint64_t s = (uint64_t)(p & POS_MASK) + ((p & KIND_PROBED) ? 1 : 0);
        uint64_t e = (uint64_t)(p & POS_MASK) + C.tr_len[slot0 + t];
        if (e > len) e = len;
        set_bits(bm, s, e, lane);
    }
}

__global__ void set_range_kernel(uint32_t* __restrict__ bm, uint64_t s, uint64_t e) { set_bits(bm, s, e, threadIdx.x & 31); }

__global__ void __launch_bounds__(128) skip_bits_kernel(const uint64_t* __restrict__ ml_beg, const uint64_t* __restrict__ ml_end,
                                                        uint64_t k0, uint64_t k1, uint64_t a0, uint64_t a1,
                                                        uint32_t* __restrict__ bm, int adjacent_starts) {
    const uint64_t k = k0 + (((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (k >= k1) return;
    uint64_t s = ml_beg[k] + 1, e = ml_end[k];
    // a phrase that follows its predecessor without a gap is (most likely) not probed at its start either
    if (adjacent_starts && k > 0 && ml_end[k - 1] >= ml_beg[k]) s = ml_beg[k];
    if (s < a0) s = a0;
    if (e > a1) e = a1;
    if (s < e) set_bits(bm, s - a0, e - a0, threadIdx.x & 31);
}

// clear [s, e) (one warp)
__global__ void clear_range_kernel(uint32_t* __restrict__ bm, uint64_t s, uint64_t e) {
    if (s >= e) return;
    const unsigned lane = threadIdx.x & 31;
    const uint64_t w0 = s >> 5, w1 = (e - 1) >> 5;
    for (uint64_t w = w0 + lane; w <= w1; w += 32) {
        uint32_t m = 0xFFFFFFFFu;
        if (w == w0) m &= 0xFFFFFFFFu << (s & 31);
        if (w == w1) m &= 0xFFFFFFFFu >> (31 - ((e - 1) & 31));
        bm[w] &= ~m;
    }
}

// Sparse slice: entries for the positions the parses asked for (one warp per request).  A position is claimed by clearing
// its skip bit; its five fingerprints are computed from scratch (L steps each: requests are few and short), the bucket ids
// go to the time-order array like those of the sorted index, and (bucket << 32 | entry id) is appended to the repair list.
__global__ void __launch_bounds__(128) repair_entries_kernel(const uint8_t* __restrict__ T, uint64_t n, const ghp5 hp, uint64_t a0,
                                                             const uint32_t* __restrict__ req, uint32_t nreq,
                                                             uint32_t* __restrict__ skip, uint32_t* __restrict__ tkey,
                                                             uint64_t* __restrict__ rk, uint32_t* __restrict__ rcount,
                                                             uint32_t rcap, uint32_t* __restrict__ first_pos) {
    const uint32_t w = (uint32_t)(((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const unsigned lane = threadIdx.x & 31;
    if (w >= nreq) return;
    const uint32_t s = req[2 + 2 * w], e = req[3 + 2 * w];
    if (lane == 0) atomicMin(first_pos, s);
    // a lane takes RCH consecutive positions: it claims those that still lack entries, then rolls every pattern from the
    // first to the last claimed one (a fresh window per position would cost 79 multiply-adds instead of 5 roll steps)
    constexpr uint32_t RCH = 8;             // (32 positions per lane start fewer windows but measured 2 ms per GiB slower)
    for (uint32_t c0 = s + lane * RCH; c0 < e; c0 += 32 * RCH) {
        const uint32_t c1 = c0 + RCH < e ? c0 + RCH : e;
        uint32_t mine = 0;
        for (uint32_t p = c0; p < c1; p++) {
            const uint32_t bit = 1u << (p & 31);
            if (!(skip[p >> 5] & bit)) continue;
            if (atomicAnd(&skip[p >> 5], ~bit) & bit) mine |= 1u << (p - c0);      // (else another request claimed it)
        }
        if (!mine) continue;
        const uint32_t slot = atomicAdd(rcount, 5u * (uint32_t)__popc(mine));
        const int jlo = __ffs(mine) - 1, jhi = 31 - __clz(mine);
#pragma unroll 1
        for (int x = 0; x < 5; x++) {
            const uint64_t L = hp.L[x], base = hp.base[x];
            const m107* __restrict__ nbt = hp.nbtab + x * 256;
            m107 fp;
            fp.lo = fp.hi = 0;
            bool live = false;
            uint32_t rank = 0;
            for (int j = jlo; j <= jhi; j++) {
                const uint32_t p = c0 + j;
                const uint64_t pos = a0 + p;
                const bool fits = pos + L < n;
                if (fits && !live) {
                    for (uint64_t q = 0; q < L; q++) {
                        m107 c;
                        c.lo = T[pos + q]; c.hi = 0;
                        fp = muladd107(fp, base, c);
                    }
                    live = true;
                }
                if ((mine >> j) & 1u) {
                    const uint32_t key = fits ? (uint32_t)(fp.lo & hp.mask) : hp.invalid_key;
                    const uint32_t en = 5u * p + (4 - x);
                    tkey[en] = key;
                    const uint32_t at = slot + 5u * rank + (4 - x);
                    if (at < rcap) rk[at] = ((uint64_t)key << 32) | en;
                    rank++;
                }
                if (fits && j < jhi) {
                    m107 infl = nbt[T[pos]];                               // -(base^L) * T[pos]
                    const uint64_t in = T[pos + L];
                    infl.lo += in;
                    infl.hi += infl.lo < in;
                    infl = fold107(infl.lo, infl.hi);
                    fp = muladd107(fp, base, infl);
                }
            }
        }
    }
}

// blocks from fb on are parsed again (their probes may see the new entries of the repair list)
__global__ void dirty_from_block_kernel(uint8_t* __restrict__ dirty, uint32_t fb, uint32_t nb) {
    const uint32_t b = fb + blockIdx.x * blockDim.x + threadIdx.x;
    if (b < nb) dirty[b] = 1;
}

// Which blocks can see NEW entries of the repair lists (change tracking): like a position whose bit flipped to "inserted",
// a new entry can become the answer of the probes behind it in its bucket, up to and including the first entry that is
// inserted (a probe inserts its position, so an entry that is not inserted is not a probe).  newent: the entries as the
// repair kernel wrote them, (bucket << 32 | entry id); the lists are sorted again by now.
__global__ void __launch_bounds__(256) dirty_from_new_entries_kernel(const parse_ctx C, const uint64_t* __restrict__ newent,
                                                                    uint32_t cnt, uint32_t nbuckets) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= cnt) return;
    const uint64_t key = newent[k];
    const uint32_t b = (uint32_t)(key >> 32), e = (uint32_t)key;
    if (b >= nbuckets) return;
    {
        uint32_t l = C.bstart[b], h = C.bstart[b + 1];
        const uint32_t hi = h;
        while (l < h) {                              // first entry of the sorted index behind e
            const uint32_t mid = (l + h) >> 1;
            if (C.sval[mid] <= e) l = mid + 1;
            else h = mid;
        }
        for (uint32_t kk = l; kk < hi; kk++) {
            const uint32_t q = C.sval[kk] / 5u;
            if (!not_inserted(C.notins, q)) { C.dirty[q / C.BL] = 1; break; }
        }
    }
    for (int which = 0; which < 2; which++) {
        const uint64_t* __restrict__ rk = which ? C.rk2 : C.rk;
        const uint32_t R = which ? C.R2 : C.R;
        if (!R) continue;
        uint32_t rl = 0, rh = R;
        if (!which) { rl = C.rstart[b]; rh = C.rstart[b + 1]; }
        while (rl < rh) {
            const uint32_t mid = (rl + rh) >> 1;
            if (rk[mid] <= key) rl = mid + 1;
            else rh = mid;
        }
        for (; rl < R; rl++) {
            const uint64_t v = rk[rl];
            if ((uint32_t)(v >> 32) != b) break;
            const uint32_t q = (uint32_t)v / 5u;
            if (!not_inserted(C.notins, q)) { C.dirty[q / C.BL] = 1; break; }
        }
    }
}

// the big repair list's share of the carried table: one thread per bucket walks the bucket's entries backwards (the bucket
// starts make this a coalesced sweep; an atomic per entry into a table of hundreds of megabytes was 14 ms per GiB)
__global__ void carry_repair_buckets_kernel(const parse_ctx C, uint32_t nbuckets, unsigned long long* __restrict__ Hc) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nbuckets) return;
    const uint32_t lo = C.rstart[b];
    for (uint32_t kk = C.rstart[b + 1]; kk > lo;) {
        kk--;
        const uint32_t q = (uint32_t)C.rk[kk] / 5u;
        if (!not_inserted(C.notins, q)) {
            const unsigned long long cand = C.a0 + q + 1;
            if (cand > Hc[b]) Hc[b] = cand;
            return;
        }
    }
}

// a small repair list's share of the carried table
__global__ void carry_repair_kernel(const parse_ctx C, const uint64_t* __restrict__ rk, uint32_t R, uint32_t nbuckets,
                                    unsigned long long* __restrict__ Hc) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= R) return;
    const uint64_t v = rk[k];
    const uint32_t b = (uint32_t)(v >> 32);
    if (b >= nbuckets) return;
    const uint32_t q = (uint32_t)v / 5u;
    if (!not_inserted(C.notins, q)) atomicMax(&Hc[b], (unsigned long long)(C.a0 + q + 1));
}

// carry the table over the slice: last inserted position per bucket.  One thread per bucket walks its list
// backwards; lists that do not settle within 64 entries are handed to a warp each (carry_big_kernel).
__global__ void carry_kernel(const parse_ctx C, uint32_t nbuckets, unsigned long long* __restrict__ Hc,
                             uint32_t* __restrict__ big, uint32_t* __restrict__ nbig) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nbuckets) return;
    const uint32_t lo = C.bstart[b];
    uint32_t kk = C.bstart[b + 1];
    int steps = 0;
    while (kk > lo) {
        kk--;
        const uint32_t q = C.sval[kk] / 5u;
        if (!not_inserted(C.notins, q)) {
            Hc[b] = C.a0 + q + 1;
            return;
        }
        if (++steps == 64) {
            big[atomicAdd(nbig, 1u)] = b;
            return;
        }
    }
}
__global__ void __launch_bounds__(128) carry_big_kernel(const parse_ctx C, unsigned long long* __restrict__ Hc,
                                                        const uint32_t* __restrict__ big, const uint32_t* __restrict__ nbig) {
    const unsigned lane = threadIdx.x & 31;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t w = (uint32_t)(((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5); w < *nbig; w += nwarps) {
        const uint32_t b = big[w];
        const uint32_t lo = C.bstart[b];
        uint32_t hi = C.bstart[b + 1];
        while (hi > lo) {
            bool ins = false;
            uint32_t q = 0;
            if (hi - lo > lane) {                            // lane 0 looks at the last entry
                q = C.sval[hi - 1 - lane] / 5u;
                ins = !not_inserted(C.notins, q);
            }
            const unsigned m = __ballot_sync(0xffffffffu, ins);
            if (m) {
                const uint32_t qq = __shfl_sync(0xffffffffu, q, __ffs(m) - 1);
                if (lane == 0) Hc[b] = C.a0 + qq + 1;
                break;
            }
            if (hi - lo <= 32) break;
            hi -= 32;
        }
    }
}

__global__ void count_kernel(const parse_ctx C, uint32_t* __restrict__ cnt) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < C.nb) {
        cnt[b] = C.marked[b] ? C.tr_cnt[b] : 0u;
        if (C.marked[b] && C.ovr[b]) atomicOr(C.changed + 2, 1u);
    }
}

__global__ void __launch_bounds__(128) gather_kernel(const parse_ctx C, const uint32_t* __restrict__ off,
                                                     uint64_t* __restrict__ fsrc, uint64_t* __restrict__ flen) {
    const uint32_t b = (uint32_t)(((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const unsigned lane = threadIdx.x & 31;
    if (b >= C.nb || !C.marked[b]) return;
    const uint32_t cnt = C.tr_cnt[b], o = off[b];
    const uint64_t slot0 = (uint64_t)b * C.BL;
    for (uint32_t t = lane; t < cnt; t += 32) {
        fsrc[o + t] = C.tr_src[slot0 + t];
        flen[o + t] = C.tr_len[slot0 + t];
    }
}

__global__ void hquery_kernel(const unsigned long long* __restrict__ Hc, const uint32_t* __restrict__ buckets, uint32_t cnt,
                              uint64_t* __restrict__ out) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < cnt) out[k] = Hc[buckets[k]];
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

template <typename pos_t>
void gap_parse_run(const uint8_t* d_text, uint64_t n, uint32_t tau, const pos_t* d_S, uint64_t s, suffix_state<pos_t>& ss,
                   lce_state<pos_t>& ls, const gap_hash_params& gp, uint64_t table_entries, const uint64_t* h_ml,
                   uint64_t m, uint64_t n_cut, uint64_t slice_len, uint32_t block_len, gap_parse_scratch& sc,
                   cudaStream_t st, pinned_vec<uint64_t>& out_src, pinned_vec<uint64_t>& out_len, uint64_t handover[4],
                   bool naive, const gap_parse_range* range) {
    static_assert(sizeof(block_state) == 32, "block_state layout");
    if (slice_len > (1ull << 28)) slice_len = 1ull << 28;
    if (block_len < 32) block_len = 32;
    const lce_view<pos_t> V = make_lce_view<pos_t>(d_text, n, tau, d_S, s, ss, ls);
    ghp5 hp;
    std::vector<m107> h_nbtab(5 * 256);
    {
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
            const u128 nb = pw == 0 ? 0 : P - pw;
            for (int c = 0; c < 256; c++) {
                const u128 v = h_mulmod(nb, (u128)c);
                h_nbtab[x * 256 + c].lo = (uint64_t)v;
                h_nbtab[x * 256 + c].hi = (uint64_t)(v >> 64);
            }
        }
        hp.mask = gp.mask;
        hp.invalid_key = (uint32_t)table_entries;
        sc.nbtab.reserve(5 * 256 * sizeof(m107));
        CUDA_CHECK(cudaMemcpyAsync(sc.nbtab.get(), h_nbtab.data(), 5 * 256 * sizeof(m107), cudaMemcpyHostToDevice, st));
        CUDA_CHECK(cudaStreamSynchronize(st));     // h_nbtab is pageable
        hp.nbtab = reinterpret_cast<const m107*>(sc.nbtab.get());
    }
    int key_bits = 1;
    while ((1ull << key_bits) <= table_entries) key_bits++;      // buckets 0..table_entries-1 and the invalid key
    sc.rounds = 0;
    sc.slices = 0;
    sc.entries_sorted = 0;
    sc.sparse_slices = 0;
    sc.sparse_fallbacks = 0;
    sc.repair_entries = 0;
    // Sparse slices (entries only for the positions the reference can insert when no gap factor runs over a whole
    // phrase; the positions such factors do run over are added to a small repair list while the parse iterates) are the
    // default; LZ77SSS_SPARSE_GAP_INDEX=0 indexes every position of every slice.
    const bool sparse_ok = !(getenv("LZ77SSS_SPARSE_GAP_INDEX") && getenv("LZ77SSS_SPARSE_GAP_INDEX")[0] == '0');

    sc.ml.reserve(3 * m);
    CUDA_CHECK(cudaMemcpyAsync(sc.ml.get(), h_ml, 3 * m * 8, cudaMemcpyHostToDevice, st));
    sc.Hc.reserve(table_entries);
    if (range && range->d_table_in)
        CUDA_CHECK(cudaMemcpyAsync(sc.Hc.get(), range->d_table_in, table_entries * 8, cudaMemcpyDeviceToDevice, st));
    else
        CUDA_CHECK(cudaMemsetAsync(sc.Hc.get(), 0, table_entries * 8, st));
    sc.changed.reserve(16);       // [8..11]: debug counters of the probes (LZ77SSS_DEBUG_ROUNDS)
    const bool dbg_rounds = getenv("LZ77SSS_DEBUG_ROUNDS") != nullptr;
    // lanes per parsed block (LZ77SSS_PARSE_GROUP = 8, 16 or 32).  Measured at 1 GiB: 65.9 / 74.5 / 80.2 ms of parse kernel
    // for 32 / 16 / 8 lanes: four times the blocks in flight do not help, a round is as long as its heaviest block
    const int parse_group = getenv("LZ77SSS_PARSE_GROUP") ? atoi(getenv("LZ77SSS_PARSE_GROUP")) : 32;
    const uint64_t track_div = getenv("LZ77SSS_TRACK_DIV") ? strtoull(getenv("LZ77SSS_TRACK_DIV"), nullptr, 10) : 64;
    sc.slice_exit.reserve(sizeof(block_state));

    block_state first;
    first.i = 0; first.cur = 0; first.k = 0; first.gap = 1; first.pad = 0;
    // A sharded run parses only [r_beg, r_end) from a given entry state and table, keeps the sorted
    // entries and the traces of every slice, and resumes from them when the entry state or the table changed.
    const uint64_t r_beg = range ? range->beg : 0, r_end = range ? range->end : n;
    const bool retain = range && range->retain;
    if (range) { first.i = range->entry[0]; first.cur = range->entry[1]; first.k = range->entry[2]; first.gap = (uint32_t)range->entry[3]; }
    if (!retain) sc.keep.clear();
    auto swap_keep = [&](slice_keep& K) {
        std::swap(K.v0, sc.v0); std::swap(K.v1, sc.v1); std::swap(K.tkey, sc.tkey); std::swap(K.bstart, sc.bstart);
        std::swap(K.entry, sc.entry); std::swap(K.exit_, sc.exit_); std::swap(K.marked, sc.marked); std::swap(K.ovr, sc.ovr);
        std::swap(K.dirty, sc.dirty); std::swap(K.notins_prev, sc.notins_prev); std::swap(K.Hc_prev, sc.Hc_prev);
        std::swap(K.tr_cnt, sc.tr_cnt); std::swap(K.tr_pos, sc.tr_pos); std::swap(K.tr_src, sc.tr_src); std::swap(K.tr_len, sc.tr_len);
        std::swap(K.skip, sc.skip); std::swap(K.rstart, sc.rstart); std::swap(K.prev, sc.prev);
        std::swap(K.rk0, sc.rk0); std::swap(K.rk1, sc.rk1); std::swap(K.rkb0, sc.rkb0); std::swap(K.rkb1, sc.rkb1);
    };

    // 1. hash + 2. sort of the slice [a0, a1) into sc.{skip, tkey, v0 / v1, bstart}, on the stream bs.  Sparse (entries
    // only for the positions that can be inserted or probed when no gap factor runs over a whole phrase, i.e. the gaps
    // and the phrase starts) unless the gaps are most of the slice anyway.  head: that many leading positions were
    // inserted by a gap factor that reaches into the slice (known only to a build in step with the parse).
    // LZ77SSS_DEBUG_ROUNDS: device timeline of the builds (either stream) and the parses
    struct tl_mark { const char* what; uint64_t a0; cudaEvent_t ev; };
    std::vector<tl_mark> tl;
    auto mark = [&](const char* what, uint64_t a0, cudaStream_t s) {
        if (!dbg_rounds) return;
        cudaEvent_t e;
        CUDA_CHECK(cudaEventCreate(&e));
        CUDA_CHECK(cudaEventRecord(e, s));
        tl.push_back({what, a0, e});
    };
    struct built_index { uint64_t E = 0; bool sparse = false, vals_in_v0 = true; };
    auto build_index = [&](uint64_t a0, uint64_t a1, bool sparse, uint64_t head, sort_scratch& srt, scan_scratch& scn,
                           cudaStream_t bs) -> built_index {
        const uint64_t cnt = a1 - a0;
        const uint64_t nwords = (cnt + 31) / 32 + 1;
        const uint32_t ntiles = ceil_div_u(cnt, GE_TILE);
        uint64_t E = 5 * cnt;
        mark(bs == st ? "build begin (main stream)" : "build begin (side stream)", a0, bs);
        if (sparse) {
            // phrases that overlap the slice (the sentinel is not one of them)
            const uint64_t k0 = std::upper_bound(h_ml + m, h_ml + m + (m - 1), a0) - (h_ml + m);
            const uint64_t k1 = std::lower_bound(h_ml, h_ml + (m - 1), a1) - h_ml;
            sc.skip.reserve(nwords);
            sc.tile_off.reserve(ntiles + 2);
            CUDA_CHECK(cudaMemsetAsync(sc.skip.get(), 0, nwords * 4, bs));
            if (k1 > k0) {
                skip_bits_kernel<<<ceil_div_u((k1 - k0) * 32, 128), 128, 0, bs>>>(sc.ml.get(), sc.ml.get() + m, k0, k1, a0, a1, sc.skip, 0);
                KERNEL_CHECK();
            }
            if (head) {
                clear_range_kernel<<<1, 32, 0, bs>>>(sc.skip, 0, head);
                KERNEL_CHECK();
            }
            tile_need_kernel<<<ceil_div_u((uint64_t)ntiles * 32, 256), 256, 0, bs>>>(sc.skip, cnt, ntiles, sc.tile_off);
            KERNEL_CHECK();
            exclusive_scan_u32(sc.tile_off, ntiles, scn, bs, sc.tile_off.get() + ntiles);
            uint32_t needed = 0;
            CUDA_CHECK(cudaMemcpyAsync(&needed, sc.tile_off.get() + ntiles, 4, cudaMemcpyDeviceToHost, bs));
            CUDA_CHECK(cudaStreamSynchronize(bs));
            if ((uint64_t)needed * 2 > cnt) sparse = false;   // mostly gaps: the dense slice costs about the same and cannot fail
            else E = 5ull * needed;
        }
        sc.k0.reserve(E + 8); sc.k1.reserve(E + 8); sc.v0.reserve(E + 8); sc.v1.reserve(E + 8); sc.tkey.reserve(5 * cnt);
        sc.bstart.reserve(table_entries + 4);
        sc.prev.reserve(5 * cnt);
        uint32_t *keys = sc.k0, *kalt = sc.k1, *vals = sc.v0, *valt = sc.v1;
        // a dense slice has vals[e] = e (the sort generates them in its first pass) and keys = tkey (its first pass
        // reads tkey)
        PROF_BEGIN(bs);
        // sparse slice with few needed positions: only the 64-position segments that hold one are rolled
        if (sparse && E * 8 < 5 * cnt)
            gap_entries_sparse_kernel<<<ceil_div_u(cnt, SP_TILE), 160, 0, bs>>>(d_text, n, hp, a0, cnt, keys, sc.tkey, vals,
                                                                              sc.skip.get(), sc.tile_off.get(), sc.prev.get());
        else
            gap_entries_kernel<<<ntiles, 160, 0, bs>>>(d_text, n, hp, a0, cnt, sparse ? keys : nullptr, sc.tkey, sparse ? vals : nullptr,
                                                       sparse ? sc.skip.get() : nullptr, sparse ? sc.tile_off.get() : nullptr, sc.prev.get());
        KERNEL_CHECK();
        // algorithmic bytes: the slice's text once, the bucket ids of its indexed positions in time order, and on a
        // sparse slice the compacted (bucket, entry id) pairs for the sort
        PROF_END(bs, "gap_entries_kernel (5 rolling fingerprints mod 2^107-1)", cnt + (sparse ? E * 12 : 5 * cnt * 4));
        if (E == 0) {                                  // no entry at all: every bucket list is empty
            CUDA_CHECK(cudaMemsetAsync(sc.bstart.get(), 0, (table_entries + 2) * 4, bs));
        } else {
            radix_sort_pairs32(keys, kalt, vals, valt, E, 0, key_bits, srt, bs, !sparse, sparse ? nullptr : sc.tkey.get());
            bucket_start_kernel<<<ceil_div_u(E, 1024), 256, 0, bs>>>(keys, E, (uint32_t)table_entries + 1, sc.bstart);
            KERNEL_CHECK();
        }
        mark("build end", a0, bs);
        built_index r;
        r.E = E; r.sparse = sparse; r.vals_in_v0 = vals == sc.v0.get();
        return r;
    };
    // second index set <-> the one the parse reads (host-side pointer swap; enqueued kernels keep their pointers)
    auto swap_sets = [&]() {
        std::swap(sc.next.v0, sc.v0); std::swap(sc.next.v1, sc.v1); std::swap(sc.next.tkey, sc.tkey);
        std::swap(sc.next.bstart, sc.bstart); std::swap(sc.next.skip, sc.skip); std::swap(sc.next.prev, sc.prev);
    };
    // The next slice's index is built while this one is parsed unless the run keeps its slices (sharded passes), the
    // kernels are being timed one by one, or the second set does not fit.
    bool prefetch = !range && !g_prof_on && r_end - r_beg > slice_len &&
                    !(getenv("LZ77SSS_PREFETCH_INDEX") && getenv("LZ77SSS_PREFETCH_INDEX")[0] == '0');
    if (prefetch && !sc.next.tkey.cap) {
        size_t free_b = 0, total_b = 0;
        CUDA_CHECK(cudaMemGetInfo(&free_b, &total_b));
        if ((double)free_b < 3.0 * 20.0 * (double)slice_len + 4e9) prefetch = false;
    }
    if (prefetch && !sc.side.st) {
        int least = 0, greatest = 0;
        CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&least, &greatest));
        CUDA_CHECK(cudaStreamCreateWithPriority(&sc.side.st, cudaStreamNonBlocking, least));
        CUDA_CHECK(cudaEventCreateWithFlags(&sc.side.ev_parse, cudaEventDisableTiming));
        CUDA_CHECK(cudaEventCreateWithFlags(&sc.side.ev_build, cudaEventDisableTiming));
    }
    bool pf_valid = false;
    uint64_t pf_a0 = 0;
    int pf_attempt = 0;
    built_index pf;
    sc.prefetched_slices = 0;
    // (a build that is still in flight when the parse ends or throws must not outlive the buffers' next use)
    struct drain { cudaStream_t s; ~drain() { if (s) cudaStreamSynchronize(s); } } drain_guard{prefetch ? sc.side.st : nullptr};

    size_t slice_no = 0;
    for (uint64_t a0 = r_beg; a0 < r_end; a0 += slice_len, slice_no++) {
        const uint64_t a1 = a0 + slice_len < r_end ? a0 + slice_len : r_end;
        const bool index_only = range && range->index_only;
        if (first.i >= a1 && !first.gap && !naive && !index_only) continue; // a phrase covers the whole slice: nothing is inserted
        slice_keep* K = nullptr;
        if (retain) {
            while (sc.keep.size() <= slice_no) sc.keep.emplace_back(new slice_keep());
            K = sc.keep[slice_no].get();
            swap_keep(*K);
        }
        struct swap_back { slice_keep* K; decltype(swap_keep)& f; ~swap_back() { if (K) f(*K); } } guard{K, swap_keep};
        bool have_sorted = K && K->sorted;
        bool have_parse = K && K->parsed;
        const uint64_t cnt = a1 - a0;
        const uint32_t BL = block_len;
        const uint32_t nb = (uint32_t)((cnt + BL - 1) / BL);
        const uint64_t nwords = (cnt + 31) / 32 + 1;
        const uint32_t ntiles = ceil_div_u(cnt, GE_TILE);
        sc.slices++;
        // Attempt 0 (sparse): entries only for the positions that can be inserted or probed when no gap factor runs
        // over a whole phrase (greedy.cpp:68-77), i.e. the gaps and the phrase starts.  A parse that does run over a
        // phrase asks for the entries of the positions it covers; they are hashed into the repair list, which the
        // probes search next to the sorted index, and the blocks behind them are parsed again.  Attempt 1 (every
        // position) is taken when the gaps are most of the slice anyway, or when the repair list outgrows its bound.
        // phrases that overlap the slice (the sentinel is not one of them)
        const uint64_t k0 = std::upper_bound(h_ml + m, h_ml + m + (m - 1), a0) - (h_ml + m);
        const uint64_t k1 = std::lower_bound(h_ml, h_ml + (m - 1), a1) - h_ml;
        // a gap factor reaches into the slice: those positions are inserted and need entries
        const uint64_t sparse_head = (first.gap && first.i > a0) ? (first.i < a1 ? first.i : a1) - a0 : 0;
        uint64_t head_pending = 0;
      for (int attempt = have_sorted ? (K->sparse ? 0 : 1) : (naive || !sparse_ok) ? 1 : 0; attempt < 2; attempt++) {
        bool sparse = attempt == 0;
        uint64_t E = 5 * cnt;
        uint32_t* vals = nullptr;
        if (have_sorted) {
            if (sparse) E = K->E;                     // a retained sparse slice: its skip bitmap, index and repair lists stand
            vals = K->vals_in_v0 ? sc.v0.get() : sc.v1.get();
            // an index that an index-only pass built (sharded runs) did not know the entry state of this slice: the
            // positions a gap factor reaching into it has inserted get their entries in front of the first parse
            // (a slice that was parsed before asks for them where it resumes, below)
            if (sparse && !have_parse) head_pending = sparse_head;
        } else {
            built_index bi;
            if (pf_valid && pf_a0 == a0 && attempt == pf_attempt) {
                // built on the second stream while the slice before was parsed
                CUDA_CHECK(cudaStreamWaitEvent(st, sc.side.ev_build, 0));
                swap_sets();
                bi = pf;
                pf_valid = false;
                head_pending = sparse_head;           // (the build did not know the entry state)
                sc.prefetched_slices++;
            } else {
                if (pf_valid) {                       // a build is in flight and shares the key buffers: after it
                    CUDA_CHECK(cudaStreamWaitEvent(st, sc.side.ev_build, 0));
                    if (pf_a0 <= a0) pf_valid = false;
                }
                bi = build_index(a0, a1, sparse, sparse_head, sc.sort, sc.scan, st);
            }
            E = bi.E;
            if (sparse && !bi.sparse) { sparse = false; attempt = 1; }   // mostly gaps: every position is indexed
            vals = bi.vals_in_v0 ? sc.v0.get() : sc.v1.get();
            sc.entries_sorted += E;
            if (K) { K->sorted = true; K->vals_in_v0 = bi.vals_in_v0; K->sparse = sparse; K->E = E; K->R = K->R2 = 0; }
        }
        // the index of the next slice, on the second stream, while this one is parsed
        if (prefetch && !pf_valid && a0 + slice_len < r_end) {
            const uint64_t b0 = a0 + slice_len, b1 = b0 + slice_len < r_end ? b0 + slice_len : r_end;
            CUDA_CHECK(cudaEventRecord(sc.side.ev_parse, st));      // (the carry kernels of the slice before still read the other set)
            CUDA_CHECK(cudaStreamWaitEvent(sc.side.st, sc.side.ev_parse, 0));
            swap_sets();
            try {
                pf = build_index(b0, b1, !naive && sparse_ok, 0, sc.sort_b, sc.scan_b, sc.side.st);
            } catch (...) { swap_sets(); throw; }
            swap_sets();
            CUDA_CHECK(cudaEventRecord(sc.side.ev_build, sc.side.st));
            pf_valid = true;
            pf_a0 = b0;
            pf_attempt = (naive || !sparse_ok) ? 1 : 0;
        }
        sc.big.reserve(table_entries + 4);
        if (index_only) {
            // speculative table of this range: last position per bucket among the positions that the phrase list
            // alone says are inserted (everything but the inside of the phrases)
            sc.notins.reserve(nwords);
            CUDA_CHECK(cudaMemsetAsync(sc.notins.get(), 0, nwords * 4, st));
            if (!naive && k1 > k0) {
                skip_bits_kernel<<<ceil_div_u((k1 - k0) * 32, 128), 128, 0, st>>>(sc.ml.get(), sc.ml.get() + m, k0, k1, a0, a1, sc.notins, 1);
                KERNEL_CHECK();
            }
            parse_ctx Ci{};
            Ci.a0 = a0; Ci.a1 = a1; Ci.sval = vals; Ci.bstart = sc.bstart; Ci.notins = sc.notins;
            CUDA_CHECK(cudaMemsetAsync(sc.changed.get() + 1, 0, 4, st));
            carry_kernel<<<ceil_div_u(table_entries, 256), 256, 0, st>>>(Ci, (uint32_t)table_entries, sc.Hc.get(), sc.big, sc.changed.get() + 1);
            KERNEL_CHECK();
            carry_big_kernel<<<1184, 128, 0, st>>>(Ci, sc.Hc.get(), sc.big, sc.changed.get() + 1);
            KERNEL_CHECK();
            break;
        }
        // per-block state
        sc.notins.reserve(nwords);
        sc.notins_prev.reserve(nwords);
        sc.dirty.reserve(nb + 1);
        if (K) sc.Hc_prev.reserve(table_entries);
        sc.ovr.reserve(nb + 1);
        sc.entry.reserve((uint64_t)nb * sizeof(block_state));
        sc.exit_.reserve((uint64_t)nb * sizeof(block_state));
        sc.marked.reserve(nb + 1);
        sc.marked_prev.reserve(nb + 1);
        sc.tr_cnt.reserve(nb);
        sc.tr_pos.reserve((uint64_t)nb * BL);
        sc.tr_src.reserve((uint64_t)nb * BL);
        sc.tr_len.reserve((uint64_t)nb * BL);
        int levels = 1;
        while ((1ull << (levels - 1)) < (uint64_t)nb + 1) levels++;
        sc.J.reserve((uint64_t)levels * (nb + 1));
        sc.cnt.reserve(nb + 2);

        mark("parse begin", a0, st);
        parse_ctx C;
        C.n = n; C.n_cut = n_cut; C.a0 = a0; C.a1 = a1; C.BL = BL; C.nb = nb; C.naive = naive ? 1u : 0u; C.bmin = 0;
        C.tkey = sc.tkey; C.sval = vals; C.bstart = sc.bstart; C.prev = sc.prev;
        C.Hc = reinterpret_cast<const uint64_t*>(sc.Hc.get());
        C.notins = sc.notins;
        C.skip = sparse ? sc.skip.get() : nullptr;
        C.rk = nullptr; C.R = 0; C.rk2 = nullptr; C.R2 = 0; C.req = nullptr; C.req_cap = 0; C.rstart = nullptr;
        uint32_t rcap = 0, rcap2 = 0;
        uint64_t *rk_cur = nullptr, *rk_alt = nullptr, *rk2_cur = nullptr, *rk2_alt = nullptr;
        if (sparse) {
            C.req_cap = std::max<uint32_t>(1u << 16, 4 * nb);
            sc.req.reserve(2 + 2 * (uint64_t)C.req_cap);
            C.req = sc.req;
            rcap = (uint32_t)std::min<uint64_t>(1ull << 30, std::max<uint64_t>(1ull << 20, E / 2));
            rcap2 = std::max<uint32_t>(1u << 20, rcap / 2);
            sc.rk0.reserve(rcap + 8); sc.rk1.reserve(rcap + 8);
            sc.rkb0.reserve(rcap2 + 8); sc.rkb1.reserve(rcap2 + 8);
            sc.rstart.reserve(table_entries + 4);
            C.rstart = sc.rstart;
            rk_cur = sc.rk0; rk_alt = sc.rk1; rk2_cur = sc.rkb0; rk2_alt = sc.rkb1;
            if (have_sorted) {                        // the repair lists of the slice's last parse
                if (!K->rk_in_0) std::swap(rk_cur, rk_alt);
                if (!K->rk2_in_0) std::swap(rk2_cur, rk2_alt);
                C.R = K->R; C.R2 = K->R2;
                C.rk = C.R ? rk_cur : nullptr;
                C.rk2 = C.R2 ? rk2_cur : nullptr;
            }
            CUDA_CHECK(cudaMemsetAsync(sc.req.get(), 0, 8, st));
        }
        bool give_up_sparse = false, track_now = false;
        bool big_open = false;                      // the big list still takes batches (up to the one of round 0)
        uint64_t repair_total = 0;
        uint32_t pending_fb = 0xFFFFFFFFu;          // blocks from here on see new repair entries in the next round
        C.ovr = sc.ovr;
        C.dirty = sc.dirty;
        C.ml_beg = sc.ml.get(); C.ml_end = sc.ml.get() + m; C.ml_src = sc.ml.get() + 2 * m; C.m = m;
        C.entry = reinterpret_cast<block_state*>(sc.entry.get());
        C.exit_ = reinterpret_cast<block_state*>(sc.exit_.get());
        C.marked = sc.marked;
        C.tr_cnt = sc.tr_cnt; C.tr_pos = sc.tr_pos; C.tr_src = sc.tr_src; C.tr_len = sc.tr_len;
        C.changed = sc.changed;
        C.dbg = dbg_rounds ? 1u : 0u;
        if (dbg_rounds) CUDA_CHECK(cudaMemsetAsync(sc.changed.get() + 8, 0, 16, st));
        block_state* d_exit = reinterpret_cast<block_state*>(sc.slice_exit.get());
        const unsigned gw = ceil_div_u((uint64_t)nb * 32, 128);
        auto launch_parse = [&](const parse_ctx& CC) {
            if (parse_group == 8) parse_kernel<pos_t, 8><<<ceil_div_u((uint64_t)nb * 8, 128), 128, 0, st>>>(CC, V);
            else if (parse_group == 16) parse_kernel<pos_t, 16><<<ceil_div_u((uint64_t)nb * 16, 128), 128, 0, st>>>(CC, V);
            else parse_kernel<pos_t, 32><<<gw, 128, 0, st>>>(CC, V);
        };

        // sparse slice, after a parse round: hash the requested positions into the repair list and sort it
        auto repair = [&]() -> bool {               // true: entries were added
            uint32_t nreq = 0;
            CUDA_CHECK(cudaMemcpyAsync(&nreq, sc.req.get(), 4, cudaMemcpyDeviceToHost, st));
            CUDA_CHECK(cudaStreamSynchronize(st));
            if (!nreq) return false;
            if (nreq > C.req_cap) { give_up_sparse = true; return false; }
            // the first batch of a slice (what round 0 found: most of it) becomes the big list; later batches go to the
            // small list, which is sorted alone and folded into the big one when it outgrows its buffer
            const bool to_big = C.R == 0 || big_open;
            const uint32_t before = to_big ? C.R : C.R2;
            const uint32_t cap_here = to_big ? rcap : rcap2;
            CUDA_CHECK(cudaMemsetAsync(sc.changed.get() + 5, 0xFF, 4, st));
            CUDA_CHECK(cudaMemcpyAsync(sc.changed.get() + 6, &before, 4, cudaMemcpyHostToDevice, st));
            repair_entries_kernel<<<ceil_div_u((uint64_t)nreq * 32, 128), 128, 0, st>>>(d_text, n, hp, a0, sc.req, nreq, sc.skip, sc.tkey,
                                                                                      to_big ? rk_cur : rk2_cur, sc.changed.get() + 6,
                                                                                      cap_here, sc.changed.get() + 5);
            KERNEL_CHECK();
            uint32_t r2[2];
            CUDA_CHECK(cudaMemcpyAsync(r2, sc.changed.get() + 5, 8, cudaMemcpyDeviceToHost, st));
            CUDA_CHECK(cudaMemsetAsync(sc.req.get(), 0, 4, st));
            CUDA_CHECK(cudaStreamSynchronize(st));
            if (r2[1] == before) return false;
            repair_total += r2[1] - before;
            // the new entries, kept aside for the change tracking below (the sorts permute the lists)
            const uint32_t nnew = std::min(r2[1], cap_here) - before;
            if (track_now) {
                sc.newent.reserve(nnew + 8);
                CUDA_CHECK(cudaMemcpyAsync(sc.newent.get(), (to_big ? rk_cur : rk2_cur) + before, (size_t)nnew * 8, cudaMemcpyDeviceToDevice, st));
            }
            if (to_big) {
                if (r2[1] > rcap) { give_up_sparse = true; return false; }
                C.R = r2[1];
                radix_sort_keys(rk_cur, rk_alt, C.R, 0, 32 + key_bits, sc.sort, st);
                C.rk = rk_cur;
                repair_start_kernel<<<ceil_div_u(C.R, 256), 256, 0, st>>>(rk_cur, C.R, (uint32_t)table_entries + 1, sc.rstart);
                KERNEL_CHECK();
                splice_repair_kernel<<<ceil_div_u(C.R, 256), 256, 0, st>>>(C.sval, C.bstart, rk_cur, C.R, nullptr, nullptr, sc.prev);
                KERNEL_CHECK();
            } else if (r2[1] <= rcap2) {
                C.R2 = r2[1];
                radix_sort_keys(rk2_cur, rk2_alt, C.R2, 0, 32 + key_bits, sc.sort, st);
                C.rk2 = rk2_cur;
                splice_repair_kernel<<<ceil_div_u(C.R2, 256), 256, 0, st>>>(C.sval, C.bstart, rk2_cur, C.R2, C.R ? C.rk : nullptr,
                                                                            sc.rstart, sc.prev);
                KERNEL_CHECK();
            } else {
                // the small list overflowed: some positions lost their skip bit without getting entries, so the slice is
                // redone with every position
                give_up_sparse = true;
                return false;
            }
            if (C.R2 > rcap2 / 2 && (uint64_t)C.R + C.R2 <= rcap) {       // fold the small list into the big one
                CUDA_CHECK(cudaMemcpyAsync(rk_cur + C.R, rk2_cur, (size_t)C.R2 * 8, cudaMemcpyDeviceToDevice, st));
                C.R += C.R2;
                C.R2 = 0;
                C.rk2 = nullptr;
                radix_sort_keys(rk_cur, rk_alt, C.R, 0, 32 + key_bits, sc.sort, st);
                C.rk = rk_cur;
                repair_start_kernel<<<ceil_div_u(C.R, 256), 256, 0, st>>>(rk_cur, C.R, (uint32_t)table_entries + 1, sc.rstart);
                KERNEL_CHECK();
            }
            if (track_now) {
                // following single changes: only the blocks whose probes can see one of the new entries (the block that
                // asked for them kept its dirty flag)
                dirty_from_new_entries_kernel<<<ceil_div_u(nnew, 256), 256, 0, st>>>(C, sc.newent.get(), nnew, (uint32_t)table_entries);
                KERNEL_CHECK();
            } else {
                const uint32_t fb = r2[0] / BL;     // every block from the first request on is parsed again
                if (fb < pending_fb) pending_fb = fb;
            }
            return true;
        };
        const uint64_t stop = a1 < n_cut ? a1 : n_cut;
        bool converged = false;
        if (first.i < stop) {
          if (!have_parse) {     // (a resumed slice starts from the traces and exit states of its last parse)
            // round 0: speculative entries
            spec_entry_kernel<<<ceil_div_u(nb, 256), 256, 0, st>>>(C, first);
            KERNEL_CHECK();
            CUDA_CHECK(cudaMemsetAsync(sc.notins.get(), 0, nwords * 4, st));
            if (!naive) {                                   // greedy_naive inserts every position: the bitmap stays zero
                // speculative bitmap of round 0: the inside of every phrase is not inserted
                if (k1 > k0) {
                    skip_bits_kernel<<<ceil_div_u((k1 - k0) * 32, 128), 128, 0, st>>>(sc.ml.get(), sc.ml.get() + m, k0, k1, a0, a1, sc.notins, 1);
                    KERNEL_CHECK();
                }
                if (first.gap && first.i > a0) {
                    clear_range_kernel<<<1, 32, 0, st>>>(sc.notins, 0, (first.i < a1 ? first.i : a1) - a0);
                    KERNEL_CHECK();
                }
                notins_kernel<<<gw, 128, 0, st>>>(C, sc.notins, first);
                KERNEL_CHECK();
            }
            if (sparse && head_pending) {
                // the index was built ahead of the entry state: the positions a gap factor reaching into the slice has
                // inserted get their entries now (first entries of the big repair list)
                const uint32_t rq[4] = {1u, 0u, 0u, (uint32_t)head_pending};
                CUDA_CHECK(cudaMemcpyAsync(sc.req.get(), rq, sizeof rq, cudaMemcpyHostToDevice, st));
                CUDA_CHECK(cudaStreamSynchronize(st));      // (rq lives on this frame)
                repair();
                head_pending = 0;
                pending_fb = 0xFFFFFFFFu;                   // (round 0 parses every block anyway)
                if (give_up_sparse) { sc.sparse_fallbacks++; continue; }
            }
            big_open = true;
            PROF_BEGIN(st);
            launch_parse(C);
            KERNEL_CHECK();
            PROF_END(st, "parse_kernel (block-speculative greedy parse, latency-bound)", 0);
            sc.rounds++;
            if (sparse) repair();
            big_open = false;
          }
            if (K) K->parsed = true;
            bool track = have_parse;    // follow single changes (resumed slices from the start, fresh ones once few are left)
            if (sparse && have_parse && first.gap && first.i > a0) {
                // resumed sparse slice whose entry state says that a gap factor reaches into it: those positions are
                // inserted and need entries (a fresh slice clears their skip bits before it hashes)
                const uint32_t rq[4] = {1u, 0u, 0u, (uint32_t)((first.i < a1 ? first.i : a1) - a0)};
                CUDA_CHECK(cudaMemcpyAsync(sc.req.get(), rq, sizeof rq, cudaMemcpyHostToDevice, st));
                track_now = true;
                repair();
            }
            for (uint64_t round = 1; round <= (uint64_t)nb + 4; round++) {
                if (track) CUDA_CHECK(cudaMemcpyAsync(sc.marked_prev.get(), sc.marked.get(), nb, cudaMemcpyDeviceToDevice, st));
                {
                    const unsigned gl = ceil_div_u((uint64_t)nb + 1, 256);
                    link_phase_kernel<<<gl, 256, 0, st>>>(C, first, sc.J, 0, 0, d_exit);
                    KERNEL_CHECK();
                    for (int l = 1; l < levels; l++) {
                        link_phase_kernel<<<gl, 256, 0, st>>>(C, first, sc.J, l, 1, d_exit);
                        KERNEL_CHECK();
                    }
                    for (int l = levels - 1; l >= 0; l--) {
                        link_phase_kernel<<<gl, 256, 0, st>>>(C, first, sc.J, l, 2, d_exit);
                        KERNEL_CHECK();
                    }
                    link_phase_kernel<<<gl, 256, 0, st>>>(C, first, sc.J, 0, 3, d_exit);
                    KERNEL_CHECK();
                }
                mark("  round: chain linked", a0, st);
                CUDA_CHECK(cudaMemsetAsync(sc.notins.get(), 0, nwords * 4, st));
                if (!naive) {
                    notins_kernel<<<gw, 128, 0, st>>>(C, sc.notins, first);
                    KERNEL_CHECK();
                }
                mark("  round: bitmap built", a0, st);
                // A resumed slice (sharded run, later passes) starts from a converged parse: few bits flip, and only
                // the blocks whose probes can see a flipped bit (or a changed table entry, or a new entry state) are
                // parsed again.  A fresh slice flips bits everywhere in its first rounds; following them costs more
                // than it saves (measured), so there every block on the chain behind the first change is parsed
                // until a round changes fewer than 1/64 of the blocks; from then on the changes are followed.
                if (track) {
                    joined_chain_kernel<<<ceil_div_u(nb, 256), 256, 0, st>>>(C, sc.marked_prev.get());
                    KERNEL_CHECK();
                    if (!naive) {
                        PROF_BEGIN(st);
                        dirty_from_bitmap_kernel<<<ceil_div_u(nwords, 8 * DB_WPW), 256, 0, st>>>(C, sc.notins_prev, nwords);   // 8 warps x DB_WPW words per CTA
                        KERNEL_CHECK();
                        PROF_END(st, "dirty_from_bitmap_kernel (change tracking, latency-bound)", 0);
                    }
                } else {
                    CUDA_CHECK(cudaMemsetAsync(sc.dirty.get(), 1, nb, st));
                }
                if (pending_fb != 0xFFFFFFFFu) {    // new repair entries: every block behind the first of them
                    if (pending_fb < nb) {
                        dirty_from_block_kernel<<<ceil_div_u(nb - pending_fb, 256), 256, 0, st>>>(sc.dirty, pending_fb, nb);
                        KERNEL_CHECK();
                    }
                    if (pending_fb < C.bmin) C.bmin = pending_fb;
                    pending_fb = 0xFFFFFFFFu;
                }
                if (have_parse && round == 1) {     // resumed slice: the table in front of it may have changed too
                    dirty_from_table_kernel<<<ceil_div_u(table_entries, 256), 256, 0, st>>>(C, sc.Hc_prev.get(), sc.notins_prev,
                                                                                            (uint32_t)table_entries);
                    KERNEL_CHECK();
                }
                mark("  round: changes marked", a0, st);
                CUDA_CHECK(cudaMemsetAsync(sc.changed.get(), 0, 4, st));
                CUDA_CHECK(cudaMemsetAsync(sc.changed.get() + 3, 0, 4, st));
                CUDA_CHECK(cudaMemsetAsync(sc.changed.get() + 4, 0xFF, 4, st));
                PROF_BEGIN(st);
                launch_parse(C);
                KERNEL_CHECK();
                PROF_END(st, "parse_kernel (block-speculative greedy parse, latency-bound)", 0);
                if (g_prof_on && dbg_rounds) {       // per-round view: the same launch once more under its round number
                    char nm[64];
                    snprintf(nm, sizeof nm, "parse round %02llu%s", (unsigned long long)round, track ? " (tracking)" : "");
                    prof_end(st, nm, 0);
                }
                mark("  round: parsed", a0, st);
                sc.rounds++;
                uint32_t chg = 0;
                CUDA_CHECK(cudaMemcpyAsync(&chg, sc.changed.get(), 4, cudaMemcpyDeviceToHost, st));
                CUDA_CHECK(cudaStreamSynchronize(st));
                if (dbg_rounds) {
                    uint32_t d[2];
                    CUDA_CHECK(cudaMemcpy(d, sc.changed.get() + 3, 8, cudaMemcpyDeviceToHost));
                    uint32_t c4[4];
                    CUDA_CHECK(cudaMemcpy(c4, sc.changed.get() + 8, 16, cudaMemcpyDeviceToHost));
                    CUDA_CHECK(cudaMemset(sc.changed.get() + 8, 0, 16));
                    fprintf(stderr, "slice %llu round %llu: %u of %u blocks changed, first %u; lane probes %u, links looked up %u, chain hops %u, "
                            "links looked up on a hop %u (incl. round 0 if this is round 1)\n", (unsigned long long)a0,
                            (unsigned long long)round, d[0], nb, d[1], c4[3], c4[0], c4[2], c4[1]);
                }
                CUDA_CHECK(cudaMemcpyAsync(sc.notins_prev.get(), sc.notins.get(), nwords * 4, cudaMemcpyDeviceToDevice, st));
                uint32_t d[2] = {0, 0};                 // blocks whose trace or exit changed, the first of them
                if (!track) CUDA_CHECK(cudaMemcpy(d, sc.changed.get() + 3, 8, cudaMemcpyDeviceToHost));
                // the next round follows single changes if this one did, or if few blocks changed (decided before the
                // repair step, which marks the blocks that see its new entries only for a round that follows changes)
                track_now = track || (uint64_t)d[0] * track_div <= nb;
                if (!track && track_now && C.bmin > 0)      // the blocks in front of the last round's first change were not
                    CUDA_CHECK(cudaMemsetAsync(sc.dirty.get(), 0, C.bmin, st));   // parsed and need not be: drop their flags
                const bool repaired = sparse && repair();
                mark("  round: repaired", a0, st);
                if (give_up_sparse) break;
                if (!chg && !repaired) { converged = true; break; }
                if (!track) {
                    if ((uint64_t)d[0] * track_div <= nb) {    // few changes left: follow them
                        track = true;
                        C.bmin = 0;
                    } else {
                        C.bmin = d[1];
                    }
                }
            }
            if (give_up_sparse) {                       // the repair list outgrew its bound: index every position
                sc.sparse_fallbacks++;
                if (K) K->sorted = K->parsed = false;
                have_sorted = have_parse = false;
                continue;
            }
            mark("converged", a0, st);
            if (!converged) throw std::runtime_error("gap_parse: the block chain did not reach its fixed point");
            if (K) CUDA_CHECK(cudaMemcpyAsync(sc.Hc_prev.get(), sc.Hc.get(), table_entries * 8, cudaMemcpyDeviceToDevice, st));
            // factors of the blocks on the chain, in order
            CUDA_CHECK(cudaMemsetAsync(sc.changed.get() + 2, 0, 4, st));
            count_kernel<<<ceil_div_u(nb, 256), 256, 0, st>>>(C, sc.cnt);
            KERNEL_CHECK();
            exclusive_scan_u32(sc.cnt, nb, sc.scan, st, sc.cnt.get() + nb);
            uint32_t zs = 0, overrun = 0;
            CUDA_CHECK(cudaMemcpyAsync(&overrun, sc.changed.get() + 2, 4, cudaMemcpyDeviceToHost, st));
            CUDA_CHECK(cudaMemcpyAsync(&zs, sc.cnt.get() + nb, 4, cudaMemcpyDeviceToHost, st));
            block_state ex;
            CUDA_CHECK(cudaMemcpyAsync(&ex, d_exit, sizeof ex, cudaMemcpyDeviceToHost, st));
            CUDA_CHECK(cudaStreamSynchronize(st));
            if (sparse && overrun)                      // a converged parse has every entry it asked for
                throw std::logic_error("gap_parse: a block of the converged parse still lacks index entries");
            if (zs) {
                sc.fsrc.reserve(zs); sc.flen.reserve(zs);
                gather_kernel<<<gw, 128, 0, st>>>(C, sc.cnt, sc.fsrc, sc.flen);
                KERNEL_CHECK();
                const size_t z0 = out_src.size();
                out_src.resize(z0 + zs);
                out_len.resize(z0 + zs);
                CUDA_CHECK(cudaMemcpyAsync(out_src.data() + z0, sc.fsrc.get(), (size_t)zs * 8, cudaMemcpyDeviceToHost, st));
                CUDA_CHECK(cudaMemcpyAsync(out_len.data() + z0, sc.flen.get(), (size_t)zs * 8, cudaMemcpyDeviceToHost, st));
                CUDA_CHECK(cudaStreamSynchronize(st));
            }
            // the bitmap of the last round is the final one (it was built from the converged traces);
            // positions from the exit state on have not been reached yet
            // (greedy_naive: every position below the exit state is inserted before the next probe)
            const uint64_t unreached = naive ? std::max<uint64_t>({n_cut, ex.i, a0}) : ex.i;
            if (unreached < a1) {
                set_range_kernel<<<1, 32, 0, st>>>(sc.notins, unreached - a0, cnt);
                KERNEL_CHECK();
            }
            CUDA_CHECK(cudaMemsetAsync(sc.changed.get() + 1, 0, 4, st));
            carry_kernel<<<ceil_div_u(table_entries, 256), 256, 0, st>>>(C, (uint32_t)table_entries, sc.Hc.get(), sc.big, sc.changed.get() + 1);
            KERNEL_CHECK();
            carry_big_kernel<<<1184, 128, 0, st>>>(C, sc.Hc.get(), sc.big, sc.changed.get() + 1);
            KERNEL_CHECK();
            if (C.R) {
                carry_repair_buckets_kernel<<<ceil_div_u(table_entries, 256), 256, 0, st>>>(C, (uint32_t)table_entries, sc.Hc.get());
                KERNEL_CHECK();
            }
            if (C.R2) {
                carry_repair_kernel<<<ceil_div_u(C.R2, 256), 256, 0, st>>>(C, C.rk2, C.R2, (uint32_t)table_entries, sc.Hc.get());
                KERNEL_CHECK();
            }
            if (K && sparse) {
                K->R = C.R; K->R2 = C.R2;
                K->rk_in_0 = rk_cur == sc.rk0.get();
                K->rk2_in_0 = rk2_cur == sc.rkb0.get();
            }
            first = ex;
        } else {
            // nothing to parse here, but positions up to first.i may have been inserted by a gap factor
            if (K) K->parsed = false;
            if (sparse && head_pending) {               // (see round 0 above)
                const uint32_t rq[4] = {1u, 0u, 0u, (uint32_t)head_pending};
                CUDA_CHECK(cudaMemcpyAsync(sc.req.get(), rq, sizeof rq, cudaMemcpyHostToDevice, st));
                CUDA_CHECK(cudaStreamSynchronize(st));
                repair();
                head_pending = 0;
                if (give_up_sparse) { sc.sparse_fallbacks++; continue; }
            }
            CUDA_CHECK(cudaMemsetAsync(sc.marked.get(), 0, nb + 1, st));
            CUDA_CHECK(cudaMemsetAsync(sc.tr_cnt.get(), 0, (size_t)nb * 4, st));
            CUDA_CHECK(cudaMemsetAsync(sc.notins.get(), 0, nwords * 4, st));
            block_state f2 = first;
            if (!naive) {
                notins_kernel<<<gw, 128, 0, st>>>(C, sc.notins, f2);
                KERNEL_CHECK();
            }
            // positions at and behind first.i (greedy_naive: behind n_cut) are not inserted yet
            const uint64_t unreached = naive ? std::max<uint64_t>(n_cut, first.i) : first.i;
            if (unreached < a1) {
                set_range_kernel<<<1, 32, 0, st>>>(sc.notins, (unreached > a0 ? unreached : a0) - a0, cnt);
                KERNEL_CHECK();
            }
            CUDA_CHECK(cudaMemsetAsync(sc.changed.get() + 1, 0, 4, st));
            carry_kernel<<<ceil_div_u(table_entries, 256), 256, 0, st>>>(C, (uint32_t)table_entries, sc.Hc.get(), sc.big, sc.changed.get() + 1);
            KERNEL_CHECK();
            carry_big_kernel<<<1184, 128, 0, st>>>(C, sc.Hc.get(), sc.big, sc.changed.get() + 1);
            KERNEL_CHECK();
            if (C.R) {
                carry_repair_buckets_kernel<<<ceil_div_u(table_entries, 256), 256, 0, st>>>(C, (uint32_t)table_entries, sc.Hc.get());
                KERNEL_CHECK();
            }
        }
        mark("parse end", a0, st);
        if (sparse) { sc.sparse_slices++; sc.repair_entries += repair_total; }
        break;
      }
    }
    if (range && range->d_table_out)
        CUDA_CHECK(cudaMemcpyAsync(range->d_table_out, sc.Hc.get(), table_entries * 8, cudaMemcpyDeviceToDevice, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    if (!tl.empty()) {
        if (sc.side.st) CUDA_CHECK(cudaStreamSynchronize(sc.side.st));
        for (auto& t : tl) {
            float ms = 0;
            cudaEventElapsedTime(&ms, tl[0].ev, t.ev);
            fprintf(stderr, "timeline %9.3f ms  slice %llu  %s\n", ms, (unsigned long long)t.a0, t.what);
        }
        for (auto& t : tl) cudaEventDestroy(t.ev);
    }
    handover[0] = first.i;
    handover[1] = first.cur;
    handover[2] = first.k;
    handover[3] = first.gap;
}

void gap_parse_query(gap_parse_scratch& sc, const uint32_t* h_buckets, uint32_t cnt, uint64_t* h_out, cudaStream_t st) {
    if (!cnt) return;
    dbuf<uint32_t> db;
    dbuf<uint64_t> dout;
    db.reserve(cnt);
    dout.reserve(cnt);
    CUDA_CHECK(cudaMemcpyAsync(db.get(), h_buckets, (size_t)cnt * 4, cudaMemcpyHostToDevice, st));
    hquery_kernel<<<ceil_div_u(cnt, 256), 256, 0, st>>>(sc.Hc.get(), db, cnt, dout);
    KERNEL_CHECK();
    CUDA_CHECK(cudaMemcpyAsync(h_out, dout.get(), (size_t)cnt * 8, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
}

#define INST(pos_t)                                                                                                     \
    template void gap_parse_run<pos_t>(const uint8_t*, uint64_t, uint32_t, const pos_t*, uint64_t, suffix_state<pos_t>&, \
                                       lce_state<pos_t>&, const gap_hash_params&, uint64_t, const uint64_t*, uint64_t,  \
                                       uint64_t, uint64_t, uint32_t, gap_parse_scratch&, cudaStream_t,                  \
                                       pinned_vec<uint64_t>&, pinned_vec<uint64_t>&, uint64_t[4], bool,                   \
                                       const gap_parse_range*);
INST(uint32_t)
INST(uint64_t)

}  // namespace lz77sss
