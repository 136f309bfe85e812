// K1+K2 for tau >= 128: rolling window fingerprints + exact sliding-window minimum by "cell thresholds".
//
// Same definition as sss.cu; replaces lce::rolling_hash::sss<idx,tau>(text,size,false) at
// /root/reference/patched-files/external/lce/include/ds/lce_sss.hpp:53.
//
// One CTA (512 threads) = one tile of NP = 16384 window ids = B = NP - tau candidate positions.
// A thread owns a CELL of 32 consecutive windows and keeps their ids in registers:
//   stage   one elected thread arms an mbarrier and issues ONE bulk-async copy (cp.async.bulk, SASS UBLKCP: the TMA unit) of
//           the tile's NP + tau + 16 text bytes into shared memory (the region that later holds the ids); every byte is
//           fetched from HBM once, and with 2 CTAs per SM the copy of one tile overlaps the arithmetic of the other,
//   pass 1  roll the fingerprint of its cell from zero (2 IMAD + 2 byte extracts per window, the 32 outgoing and the 32
//           incoming bytes as 16-byte shared-memory loads);  the missing start value of every cell comes from one
//           block-wide scan of affine maps (fingerprint algebra: fp' = fp * b^32 + c),
//   pass 2  ids = F * b^k + partial_k, cell minimum, ids to (swizzled) shared memory,
//   cells   F[t] = min of the tau/32 - 1 cells that every window starting in cell t fully contains (doubling),
//   pass 3  a window id can only be the minimum of [i, i+tau] if it is <= F (or, as the right end, <= F[t - tau/32]):
//           one compare per window; the few survivors (~4 per tau windows) are checked exactly against the two
//           partial cells by one thread each,
//   output  selected positions are set in a shared bitmap, counted, ordered across tiles by decoupled
//           look-back (tiles take tickets) and written.
// Low-period windows (Q) and the end of the text are handled by a block-uniform "careful" variant of pass 2,
// taken only by tiles that contain such windows.
// Algorithmic HBM bytes: n + sizeof(pos_t) * |S|.
#include "common.cuh"
#include "stages.cuh"

namespace lz77sss {
namespace {

constexpr uint32_t INFI = 0x80000000u;   // internal "no id" (real ids are odd and < 2^31, the terminator is 0)
constexpr int NT = 512;
constexpr int CH = 32;
constexpr int NP = NT * CH;              // ids per tile
constexpr int CMPAD = 136;

struct cell_consts {
    uint32_t base, btau;
    uint32_t bk[32];        // base^k
    uint32_t lanepow[32];   // base^(32*lane)
    uint32_t pw32[5];       // base^(32 * 2^k)      warp scan of cell maps
    uint32_t pwW[5];        // base^(1024 * 2^k)    scan over warps
    uint32_t pwH[5];        // base^(128 * 2^k)     combining the first-window hash
    uint32_t B32;           // base^32
    uint32_t P, A;
    uint32_t idmask, idone; // 0x7FFFFFFE, 1 (kept in registers so that the id is one LOP3)
};

__device__ __forceinline__ uint32_t byte_of(const uint32_t (&w)[8], int k) { return __byte_perm(w[k >> 2], 0u, 0x4440u + (k & 3)); }

// (fp & m) | o in ONE LOP3 (operands in registers; with immediates the compiler emits two)
__device__ __forceinline__ uint32_t make_id(uint32_t fp, uint32_t m, uint32_t o) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(fp), "r"(m), "r"(o));
    return r;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ int id_slot(int p) {   // word index of id p in the swizzled shared array
    const int t = p >> 5, m = (p >> 2) & 7, e = p & 3;
    return (t << 5) + ((m ^ (t & 7)) << 2) + e;
}

template <int TAU, typename pos_t>
__global__ void __launch_bounds__(NT, 2) sss_cell_kernel(const uint8_t* __restrict__ T, uint64_t n, const cell_consts c,
                                                         uint16_t* __restrict__ stage, uint32_t cap,
                                                         uint32_t* __restrict__ tile_cnt,
                                                         unsigned long long* __restrict__ counters) {
    constexpr int B = NP - TAU;
    constexpr int NC = TAU / CH;          // cells per tau
    constexpr int W = NC - 1;             // full cells inside every window
    constexpr int NW = NT / 32;
    constexpr int PW = W >= 64 ? 64 : W >= 32 ? 32 : W >= 16 ? 16 : W >= 8 ? 8 : W >= 4 ? 4 : 2;   // largest 2^k <= W
    extern __shared__ __align__(16) uint8_t smem[];
    uint32_t* sId = reinterpret_cast<uint32_t*>(smem);                  // NP ids (swizzled)
    uint32_t* sCM0 = sId + NP;                                          // NT + CMPAD
    uint32_t* sCM1 = sCM0 + NT + CMPAD;                                 // NT + CMPAD
    uint32_t* sF = sCM1 + NT + CMPAD;                                   // NT
    uint32_t* sBits = sF + NT;                                          // NT words (B/32 used)
    uint32_t* sH = sBits + NT;                                          // 128
    uint16_t* sQlist = reinterpret_cast<uint16_t*>(sH + 128);           // NT
    uint8_t* sQ = reinterpret_cast<uint8_t*>(sQlist + NT);              // NP flags
    __shared__ uint32_t s_warp[NW + 1];
    __shared__ uint32_t s_cnt[NW];
    __shared__ uint32_t s_nqpass, s_anyq, s_q;
    __shared__ __align__(8) uint64_t s_mbar;
    constexpr uint32_t TXT = NP + TAU + 16;                            // text bytes of the tile; fits the id array (4 NP bytes)
    const uint8_t* sT = reinterpret_cast<const uint8_t*>(sId);         // staged text, dead before the ids are written

    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) {
        s_nqpass = 0;
        s_anyq = 0;
        s_q = 0;
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&s_mbar)), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    sBits[tid] = 0;
    if (tid < CMPAD) { sCM0[NT + tid] = INFI; sCM1[NT + tid] = INFI; }
    if (tid < 128) sH[tid] = 0;
    __syncthreads();
    const uint32_t tile = blockIdx.x;
    const uint64_t i0 = (uint64_t)tile * B;
    const uint8_t* __restrict__ gT = T + i0;
    if (tid == 0) {
        // i0 is a multiple of 32, the text buffer is 256-byte aligned and zero-padded 64 KiB behind n: the copy is 16-byte
        // aligned at both ends and stays inside the allocation
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&s_mbar)), "r"(TXT) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(sId)), "l"(gT), "r"(TXT), "r"(smem_u32(&s_mbar)) : "memory");
    }
    const uint64_t nw = n - TAU + 1;                                    // real windows 0..nw-1; id(nw) = 0
    const uint64_t cnt_pos = n - 2 * (uint64_t)TAU + 2;                 // candidates 0..cnt_pos-1
    const bool last_tile = tile == gridDim.x - 1;
    // tile-relative limits (clamped)
    const int rel_nw = nw - i0 >= (uint64_t)NP ? NP + 1 : (int)(nw - i0);           // first non-real window
    const int rel_cnt = cnt_pos - i0 >= (uint64_t)B ? B : (int)(cnt_pos - i0);      // candidates in this tile
    const bool tail = rel_nw <= NP;

    // ---- pass 1: roll the cell from zero ----
    uint32_t val[CH];   // partial_k, later id_k
    uint32_t cell_c, cell_min;
    {
        uint32_t ow[8], iw[8];
        {   // the staged text has arrived
            uint32_t done = 0;
            while (!done)
                asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                             : "=r"(done) : "r"(smem_u32(&s_mbar)), "r"(0) : "memory");
        }
        const uint4* po = reinterpret_cast<const uint4*>(sT + (size_t)tid * CH);
        const uint4* pi = reinterpret_cast<const uint4*>(sT + (size_t)tid * CH + TAU);
        uint4 a0 = po[0], a1 = po[1], b0 = pi[0], b1 = pi[1];
        ow[0] = a0.x; ow[1] = a0.y; ow[2] = a0.z; ow[3] = a0.w; ow[4] = a1.x; ow[5] = a1.y; ow[6] = a1.z; ow[7] = a1.w;
        iw[0] = b0.x; iw[1] = b0.y; iw[2] = b0.z; iw[3] = b0.w; iw[4] = b1.x; iw[5] = b1.y; iw[6] = b1.z; iw[7] = b1.w;
        if (tid < NC) {   // hash of the own 32 bytes: the first window of the tile is made of the first NC cells
            uint32_t h = 0;
#pragma unroll
            for (int k = 0; k < CH; k++) h = h * c.base + byte_of(ow, k);
            sH[128 - NC + tid] = h;
        }
        uint32_t fp = 0;
#pragma unroll
        for (int k = 0; k < CH; k++) {
            val[k] = fp;
            fp = fp * c.base + byte_of(iw, k);
            fp = fp - c.btau * byte_of(ow, k);
        }
        cell_c = fp;
    }
    // inclusive scan of the cell maps inside the warp: x = sum_{u<=lane} c_u * b^(32*(lane-u))
    uint32_t x = cell_c;
#pragma unroll
    for (int k = 0; k < 5; k++) {
        uint32_t y = __shfl_up_sync(0xffffffffu, x, 1 << k);
        if (lane >= (1 << k)) x = y * c.pw32[k] + x;
    }
    uint32_t ex = __shfl_up_sync(0xffffffffu, x, 1);
    if (lane == 0) ex = 0;
    if (lane == 31) s_warp[w + 1] = x;

    // ---- Q prefilter: a window of period <= P needs T[a..a+2P) periodic at its anchor a (anchors are A apart).
    //      8 threads per anchor, 4 candidate periods each: does the first 8-byte word of the anchor recur at +c ? ----
    {
        const int A = (int)c.A, P = (int)c.P;
        const int nanch = (NP + A - 2) / A + 1;
        for (int slot = tid; slot < nanch * 8; slot += NT) {
            const int ai = slot >> 3, sub = slot & 7;
            const int la = ai * A;
            const uint64_t a = i0 + (uint64_t)la;
            bool pass = false;
            if (a + 2 * (uint64_t)P <= n) {
                if (P >= 8) {
                    const uint32_t* q = reinterpret_cast<const uint32_t*>(gT + la);   // la is a multiple of A >= 4
                    const uint32_t w0 = q[0], w1 = q[1];
                    // periods c = 4*sub+1 .. 4*sub+4 read the words sub .. sub+3
                    const uint32_t x0 = q[sub], x1 = q[sub + 1], x2 = q[sub + 2], x3 = q[sub + 3];
                    const int c0 = 4 * sub;
                    pass |= (c0 + 1 <= P) && __funnelshift_r(x0, x1, 8) == w0 && __funnelshift_r(x1, x2, 8) == w1;
                    pass |= (c0 + 2 <= P) && __funnelshift_r(x0, x1, 16) == w0 && __funnelshift_r(x1, x2, 16) == w1;
                    pass |= (c0 + 3 <= P) && __funnelshift_r(x0, x1, 24) == w0 && __funnelshift_r(x1, x2, 24) == w1;
                    pass |= (c0 + 4 <= P) && x1 == w0 && x2 == w1;
                    (void)x3;
                } else {
                    pass = sub == 0;
                }
            }
            // one list entry per anchor
            const unsigned grp = __ballot_sync(__activemask(), pass);
            const unsigned mine = (grp >> (lane & 24)) & 0xFFu;
            if (mine && sub == (__ffs(mine) - 1)) sQlist[atomicAdd(&s_nqpass, 1u)] = (uint16_t)ai;
        }
    }
    __syncthreads();   // #1: s_warp totals, sH, Q list

    // ---- warp 0: first-window hash F0 and the scan over warps ----
    if (w == 0) {
        // F0 = sum_{u<NC} H_u * B32^(NC-1-u); sH is right-aligned in 128 slots, lane owns 4 of them
        uint32_t g = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) g = g * c.B32 + sH[4 * lane + k];
#pragma unroll
        for (int k = 0; k < 5; k++) {
            uint32_t y = __shfl_up_sync(0xffffffffu, g, 1 << k);
            if (lane >= (1 << k)) g = y * c.pwH[k] + g;
        }
        const uint32_t F0 = __shfl_sync(0xffffffffu, g, 31);
        // elements: lane 0 = F0, lane l = total of warp l-1
        uint32_t xw = lane == 0 ? F0 : (lane <= NW ? s_warp[lane] : 0);
#pragma unroll
        for (int k = 0; k < 5; k++) {
            uint32_t y = __shfl_up_sync(0xffffffffu, xw, 1 << k);
            if (lane >= (1 << k)) xw = y * c.pwW[k] + xw;
        }
        if (lane < NW) s_warp[lane] = xw;     // value entering warp `lane`
    }
    // ---- Q marking (rare): zero the flags, then one warp per anchor that passed the prefilter ----
    const uint32_t nqpass = s_nqpass;
    if (nqpass) {
        for (int v = tid; v < NP / 16; v += NT) reinterpret_cast<uint4*>(sQ)[v] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();   // #2
    if (nqpass) {
        const uint32_t A = c.A, P = c.P;
        for (uint32_t qi = w; qi < nqpass; qi += NW) {
            const int ai = sQlist[qi];
            const int la = ai * (int)A;
            const uint64_t a = i0 + (uint64_t)la;
            const uint32_t cnd = lane + 1;
            bool ok = cnd <= P;
            for (uint32_t xx = 0; ok && xx + cnd < 2 * P; xx++) ok = gT[la + xx] == gT[la + xx + cnd];
            const unsigned pm = __ballot_sync(0xffffffffu, ok);
            if (!pm) continue;
            const uint32_t p = __ffs(pm);
            uint64_t capR = a + TAU - 1;
            if (capR > n - 1) capR = n - 1;
            uint64_t xr = a + 2 * (uint64_t)P - p;
            while (true) {
                uint64_t xi = xr + lane;
                bool good = (xi + p <= capR) && T[xi] == T[xi + p];
                unsigned m = __ballot_sync(0xffffffffu, good);
                if (m == 0xffffffffu) { xr += 32; continue; }
                xr += __ffs(~m) - 1;
                break;
            }
            const uint64_t R = xr + p - 1;
            uint64_t capL = a >= (uint64_t)(A - 1) ? a - (A - 1) : 0;
            if (capL < i0) capL = i0;
            uint64_t L = a;
            while (true) {
                bool good = (L >= capL + 1 + lane);
                if (good) {
                    uint64_t li = L - 1 - lane;
                    good = T[li] == T[li + p];
                }
                unsigned m = __ballot_sync(0xffffffffu, good);
                if (m == 0xffffffffu) { L -= 32; continue; }
                L -= __ffs(~m) - 1;
                break;
            }
            if (R + 1 < (uint64_t)TAU) continue;
            uint64_t hi = R + 1 - TAU;
            if (hi > a) hi = a;
            if (hi >= nw) hi = nw - 1;
            bool any = false;
            for (uint64_t j = L + lane; j <= hi; j += 32)
                if (j - i0 < (uint64_t)NP) { sQ[j - i0] = 1; any = true; }
            if (any) s_anyq = 1;
        }
        __syncthreads();   // #2b (block-uniform)
    }
    const bool careful = tail || (nqpass && s_anyq);

    // ---- pass 2: ids of the own cell ----
    {
        const uint32_t F = s_warp[w] * c.lanepow[lane] + ex;   // fingerprint of the first window of the cell
        uint32_t cm = INFI;
        if (!careful) {
#pragma unroll
            for (int k = 0; k < CH; k++) {
                const uint32_t fp = F * c.bk[k] + val[k];
                const uint32_t id = make_id(fp, c.idmask, c.idone);
                val[k] = id;
                cm = min(cm, id);
            }
        } else {
            uint32_t qcount = 0;
            const bool use_q = nqpass && s_anyq;
#pragma unroll
            for (int k = 0; k < CH; k++) {
                const int p = tid * CH + k;
                const uint32_t fp = F * c.bk[k] + val[k];
                uint32_t id = make_id(fp, c.idmask, c.idone);
                if (use_q && sQ[p]) {
                    id = INFI;
                    if (p < rel_nw && (p < B || last_tile)) qcount++;
                }
                if (p >= rel_nw) id = (p == rel_nw) ? 0u : INFI;
                val[k] = id;
                cm = min(cm, id);
            }
            if (qcount) atomicAdd(&s_q, qcount);
        }
#pragma unroll
        for (int m = 0; m < 8; m++)
            reinterpret_cast<uint4*>(sId)[(tid << 3) + (m ^ (tid & 7))] =
                make_uint4(val[4 * m], val[4 * m + 1], val[4 * m + 2], val[4 * m + 3]);
        sCM0[tid] = cm;
        cell_min = cm;
    }
    __syncthreads();   // #3

    // ---- cells: F[t] = min cm[t+1 .. t+W] by doubling ----
    {
        uint32_t* cur = sCM0;
        uint32_t* nxt = sCM1;
#pragma unroll
        for (int d = 1; d < PW; d <<= 1) {
            nxt[tid] = min(cur[tid], cur[tid + d]);
            __syncthreads();
            uint32_t* tsw = cur; cur = nxt; nxt = tsw;
        }
        // cur[t] = min cm[t .. t+PW-1]
        sF[tid] = min(cur[tid + 1], cur[tid + 1 + W - PW]);
    }
    __syncthreads();

    // ---- pass 3: cells whose minimum can be a window minimum (cm <= F[t] or cm <= F[t-NC], ~1 in 8) are
    //      examined by the whole warp: lane = window of the cell, ids from shared memory ----
    {
        const bool canL = tid < B / CH, canR = tid >= NC;
        uint32_t thL = canL ? sF[tid] : 0u, thR = canR ? sF[tid - NC] : 0u;
        if (thL > 0x7FFFFFFFu) thL = 0x7FFFFFFFu;       // INFI never passes
        if (thR > 0x7FFFFFFFu) thR = 0x7FFFFFFFu;
        const bool fl = (canL && cell_min <= thL) || (canR && cell_min <= thR);
        unsigned todo = __ballot_sync(0xffffffffu, fl);
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const int t = (w << 5) + src;
            const uint32_t tL = __shfl_sync(0xffffffffu, thL, src), tR = __shfl_sync(0xffffffffu, thR, src);
            const bool cL = t < B / CH, cR = t >= NC;
            const int p = (t << 5) + lane;
            const uint32_t own = sId[id_slot(p)];
            unsigned mL = __ballot_sync(0xffffffffu, cL && own <= tL && p < rel_cnt);
            unsigned mR = __ballot_sync(0xffffffffu, cR && own <= tR && p - TAU < rel_cnt);
            if (mL) {
                const uint32_t right = sId[id_slot(((t + NC) << 5) + lane)];
                while (mL) {
                    const int o = __ffs(mL) - 1;
                    mL &= mL - 1;
                    const uint32_t v = __shfl_sync(0xffffffffu, own, o);
                    const bool bad = lane > o ? own < v : right < v;      // window [p, p+tau]: own cell after o, far cell up to o
                    if (!__any_sync(0xffffffffu, bad) && lane == 0) {
                        const int ps = (t << 5) + o;
                        atomicOr(&sBits[ps >> 5], 1u << (ps & 31));
                    }
                }
            }
            if (mR) {
                const uint32_t left = sId[id_slot(((t - NC) << 5) + lane)];
                while (mR) {
                    const int o = __ffs(mR) - 1;
                    mR &= mR - 1;
                    const uint32_t v = __shfl_sync(0xffffffffu, own, o);
                    const bool bad = lane < o ? own < v : left < v;       // window [p-tau, p]: own cell before o, far cell from o
                    if (!__any_sync(0xffffffffu, bad) && lane == 0) {
                        const int ps = (t << 5) + o - TAU;
                        atomicOr(&sBits[ps >> 5], 1u << (ps & 31));
                    }
                }
            }
        }
    }
    __syncthreads();

    // ---- output: the tile's positions go, in order, to its slice of the staging buffer (tile-relative, 16 bit);
    //      sss_gather_kernel places them once the tile counts have been scanned (no inter-tile waiting) ----
    uint32_t bits = tid < B / 32 ? sBits[tid] : 0;
    const uint32_t mycnt = __popc(bits);
    uint32_t inc = mycnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += y;
    }
    if (lane == 31) s_cnt[w] = inc;
    __syncthreads();
    uint32_t wbase = (lane < NW && lane < w) ? s_cnt[lane] : 0u;     // sum of the counts of the warps before this one
#pragma unroll
    for (int d = 8; d >= 1; d >>= 1) wbase += __shfl_xor_sync(0xffffffffu, wbase, d);
    wbase = __shfl_sync(0xffffffffu, wbase, 0);
    if (tid == NT - 1) {
        const uint32_t total = wbase + inc;
        tile_cnt[tile] = total;
        if (total > cap) atomicOr(&counters[2], 1ull);
        if (s_q) atomicAdd(&counters[1], (unsigned long long)s_q);
    }
    if (bits) {
        uint32_t o = wbase + (inc - mycnt);
        uint16_t* dst = stage + (size_t)tile * cap;
        while (bits) {
            const int k = __ffs(bits) - 1;
            bits &= bits - 1;
            if (o < cap) dst[o] = (uint16_t)(tid * 32 + k);
            o++;
        }
    }
}

// one warp per tile: S[off[tile] + k] = tile * B + stage[tile][k]
template <typename pos_t>
__global__ void __launch_bounds__(256) sss_gather_kernel(const uint16_t* __restrict__ stage, uint32_t cap, uint32_t B,
                                                         const uint32_t* __restrict__ off, uint32_t ntiles,
                                                         pos_t* __restrict__ S_out, uint64_t s_cap,
                                                         unsigned long long* __restrict__ counters) {
    const uint32_t tile = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    if (tile >= ntiles) return;
    const uint32_t o0 = off[tile], cnt = off[tile + 1] - o0;
    if (tile == 0 && lane == 0) counters[0] = off[ntiles];
    const uint16_t* src = stage + (size_t)tile * cap;
    const uint64_t base = (uint64_t)tile * B;
    for (uint32_t k = lane; k < cnt && k < cap; k += 32)
        if ((uint64_t)o0 + k < s_cap) S_out[o0 + k] = (pos_t)(base + src[k]);
}

uint32_t pow_u32c(uint32_t b, uint64_t e) {
    uint32_t r = 1;
    while (e) {
        if (e & 1) r *= b;
        b *= b;
        e >>= 1;
    }
    return r;
}

template <int TAU, typename pos_t>
void launch_cell(const uint8_t* d_text, uint64_t n, const cell_consts& c, pos_t* d_S, uint64_t s_cap, sss_scratch& sc,
                 cudaStream_t st) {
    constexpr int B = NP - TAU;
    const uint64_t cnt_pos = n - 2 * (uint64_t)TAU + 2;
    const uint64_t ntiles = (cnt_pos + B - 1) / B;
    if (ntiles >= 0xFFFFFFF0ull) throw std::invalid_argument("sss_cell: text too long");
    const size_t smem = (size_t)NP * 4 + 2 * (size_t)(NT + CMPAD) * 4 + (size_t)NT * 4 * 2 + 128 * 4 + (size_t)NT * 2 +
                        (size_t)NP;
    auto k = sss_cell_kernel<TAU, pos_t>;
    CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // expected 2B/tau positions per tile; a tile that selects more than `cap` raises counters[2] and the
    // build is repeated with room for every position
    uint32_t cap = 16u * (uint32_t)(B / TAU) + 256u;
    for (int attempt = 0; attempt < 2; attempt++) {
        sc.blk.reserve(ntiles + 2);
        sc.stage.reserve(ntiles * (uint64_t)cap);
        CUDA_CHECK(cudaMemsetAsync(sc.counters.get(), 0, 32, st));
        k<<<(unsigned)ntiles, NT, smem, st>>>(d_text, n, c, sc.stage.get(), cap, sc.blk.get(), sc.counters.get());
        KERNEL_CHECK();
        exclusive_scan_u32(sc.blk.get(), ntiles, sc.scan, st, sc.blk.get() + ntiles);
        sss_gather_kernel<pos_t><<<ceil_div_u(ntiles * 32, 256), 256, 0, st>>>(sc.stage.get(), cap, (uint32_t)B, sc.blk.get(),
                                                                               (uint32_t)ntiles, d_S, s_cap, sc.counters.get());
        KERNEL_CHECK();
        unsigned long long ovf = 0;
        CUDA_CHECK(cudaMemcpyAsync(&ovf, sc.counters.get() + 2, 8, cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        if (!ovf) return;
        cap = (uint32_t)B;
    }
    throw std::runtime_error("sss_cell: staging overflow with full capacity");
}

template <typename pos_t>
void dispatch_cell(const uint8_t* d_text, uint64_t n, uint32_t tau, const cell_consts& c, pos_t* d_S, uint64_t s_cap,
                   sss_scratch& sc, cudaStream_t st) {
    switch (tau) {
        case 128: launch_cell<128, pos_t>(d_text, n, c, d_S, s_cap, sc, st); break;
        case 256: launch_cell<256, pos_t>(d_text, n, c, d_S, s_cap, sc, st); break;
        case 512: launch_cell<512, pos_t>(d_text, n, c, d_S, s_cap, sc, st); break;
        case 1024: launch_cell<1024, pos_t>(d_text, n, c, d_S, s_cap, sc, st); break;
        case 2048: launch_cell<2048, pos_t>(d_text, n, c, d_S, s_cap, sc, st); break;
        case 4096: launch_cell<4096, pos_t>(d_text, n, c, d_S, s_cap, sc, st); break;
        default: throw std::invalid_argument("sss_cell: tau must be a power of two in [128,4096]");
    }
}

}  // namespace

// tau >= 128; d_text readable (zero-filled) up to n + TEXT_PAD bytes; counters/ticket reserved by the caller
void sss_build_cell(const uint8_t* d_text, uint64_t n, uint32_t tau, uint32_t base, uint32_t qmax, void* d_S, int pos_bytes,
                    uint64_t s_cap, sss_scratch& sc, cudaStream_t st) {
    cell_consts c;
    c.base = base;
    c.btau = pow_u32c(base, tau);
    for (int k = 0; k < 32; k++) {
        c.bk[k] = pow_u32c(base, k);
        c.lanepow[k] = pow_u32c(base, 32ull * k);
    }
    for (int k = 0; k < 5; k++) {
        c.pw32[k] = pow_u32c(base, 32ull << k);
        c.pwW[k] = pow_u32c(base, 1024ull << k);
        c.pwH[k] = pow_u32c(base, 128ull << k);
    }
    c.B32 = pow_u32c(base, 32);
    c.idmask = 0x7FFFFFFEu;
    c.idone = 1u;
    c.P = sss_period_bound(tau, qmax);
    c.A = sss_anchor_spacing(tau, c.P);
    if (pos_bytes == 4) dispatch_cell<uint32_t>(d_text, n, tau, c, (uint32_t*)d_S, s_cap, sc, st);
    else dispatch_cell<uint64_t>(d_text, n, tau, c, (uint64_t*)d_S, s_cap, sc, st);
}

}  // namespace lz77sss
