// K1+K2 fused: rolling fingerprints of all length-tau windows + low-period (Q) exclusion +
// sliding-window minimum + selection + ordered compaction  ->  the tau-synchronizing set S.
//
// Replaces lce::rolling_hash::sss<idx,tau>(text,size,false) (call site
// /root/reference/patched-files/external/lce/include/ds/lce_sss.hpp:53; the constructor itself is
// in the missing lce submodule).  Canonical definition: DESIGN.md §SSS (same as the oracle's):
//   fp(j) = sum_k T[j+k]*base^(tau-1-k) mod 2^32;  id(j) = INF if per(T[j..j+tau)) <= P else (fp&0x7FFFFFFE)|1;
//   id(n-tau+1) = 0;  i in S  <=>  i <= n-2tau+1,  M(i) = min id[i..i+tau] != INF,  id(i)==M(i) or id(i+tau)==M(i).
//
// One CTA = one tile of B = TILE-2*tau candidate positions; the text tile is staged once in shared
// memory (TILE+A+16 bytes), every thread owns 16 consecutive windows:
//   1. warp-per-anchor period detection marks Q windows in shared memory,
//   2. 16-byte Horner hash per thread, block prefix-scan of (hash, base^16) pairs  -> window
//      fingerprint of the first own window = E[t+tau/16] - E[t]*base^tau, then 15 rolls,
//   3. per-thread prefix/suffix minima + segmented scan over the tau-aligned block (van Herk),
//   4. M(i) = min(suffixmin_b(i), prefixmin_{b+1}(i+tau)), selection, block scan of the counts,
//      decoupled look-back across tiles (tiles take tickets, so S comes out sorted).
// Algorithmic HBM bytes: n (text read once) + sizeof(pos_t)*|S|.
#include "common.cuh"
#include "stages.cuh"

namespace lz77sss {
namespace {

constexpr uint32_t ID_INF = 0xFFFFFFFFu;
constexpr int CH = 16;  // windows per thread

struct sss_consts {
    uint32_t base, btau;      // base^tau
    uint32_t pw16[5];         // base^(16*2^k)
    uint32_t pw512[5];        // base^(512*2^k)
    uint32_t lanepow[32];     // base^(16*lane)
    uint32_t P, A;            // period bound, anchor spacing
};

constexpr uint64_t ST_AGG = 1ull << 62, ST_INC = 2ull << 62, ST_MASK = (1ull << 62) - 1;

template <int TAU, int NT, typename pos_t>
__global__ void __launch_bounds__(NT) sss_kernel(const uint8_t* __restrict__ T, uint64_t n, const sss_consts c,
                                                 pos_t* __restrict__ S_out, uint64_t s_cap,
                                                 unsigned long long* __restrict__ status,   // [ntiles], zeroed
                                                 unsigned int* __restrict__ ticket,          // zeroed
                                                 unsigned long long* __restrict__ counters)  // [0]=|S| [1]=|Q|
{
    constexpr int TILE = NT * CH;            // windows hashed per tile
    constexpr int B = TILE - 2 * TAU;        // candidate positions per tile
    constexpr int NID = B + TAU;             // ids per tile
    constexpr int SEG = TAU / CH;            // threads per tau-aligned block
    constexpr int NW = NT / 32;
    static_assert(TAU >= CH && B > 0 && B % TAU == 0, "tile shape");
    extern __shared__ __align__(16) uint8_t smem[];
    const uint32_t A = c.A, P = c.P;
    const int TXT = TILE + (int)A + 16;
    uint8_t* sT = smem;                                        // TXT bytes (rounded to 16)
    uint8_t* sQ = sT + ((TXT + 15) & ~15);                     // NID flags
    uint32_t* sPre = (uint32_t*)(sQ + NID);                    // NID prefix minima (NID % 16 == 0)
    uint32_t* sId = sPre + NID;                                // NID ids
    uint32_t* sE = sId + NID;                                  // NT exclusive prefix hashes
    __shared__ uint32_t s_warp[NW];
    __shared__ uint32_t s_wmin_pre[NW];
    __shared__ uint32_t s_tile, s_cnt[NW];
    __shared__ unsigned long long s_base;
    __shared__ uint32_t s_q;

    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) {
        s_tile = atomicAdd(ticket, 1u);
        s_q = 0;
    }
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint64_t i0 = (uint64_t)tile * B;                    // multiple of tau (and of A)
    const uint64_t nw = n - TAU + 1;                           // real windows 0..nw-1; id(nw) = 0
    const uint64_t cnt_pos = n - 2 * (uint64_t)TAU + 2;        // candidates 0..cnt_pos-1

    // ---- stage the text tile (the device copy is zero-padded far beyond n) ----
    {
        const uint4* g = reinterpret_cast<const uint4*>(T + i0);   // i0 % 16 == 0, T 256-B aligned
        uint4* s = reinterpret_cast<uint4*>(sT);
        const int nv = (TXT + 15) / 16;
        for (int v = tid; v < nv; v += NT) s[v] = g[v];
        for (int v = tid; v < NID / 16; v += NT) reinterpret_cast<uint4*>(sQ)[v] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();

    // ---- 1. Q marking, one warp per anchor ----
    {
        const int nanch = (NID + (int)A - 2) / (int)A + 1;
        for (int ai = w; ai < nanch; ai += NW) {
            const uint64_t a = i0 + (uint64_t)ai * A;
            if (a + 2 * (uint64_t)P > n) continue;
            const int la = ai * (int)A;                        // local offset of the anchor
            const uint32_t cnd = lane + 1;
            bool ok = cnd <= P;
            for (uint32_t x = 0; ok && x + cnd < 2 * P; x++) ok = sT[la + x] == sT[la + x + cnd];
            const unsigned pm = __ballot_sync(0xffffffffu, ok);
            if (!pm) continue;
            const uint32_t p = __ffs(pm);                      // smallest period of T[a..a+2P)
            // right end of the p-periodic stretch, capped at a+tau-1 and n-1
            uint64_t capR = a + TAU - 1;
            if (capR > n - 1) capR = n - 1;
            uint64_t x = a + 2 * (uint64_t)P - p;
            while (true) {
                uint64_t xi = x + lane;
                bool good = (xi + p <= capR) && sT[xi - i0] == sT[xi - i0 + p];
                unsigned m = __ballot_sync(0xffffffffu, good);
                if (m == 0xffffffffu) { x += 32; continue; }
                x += __ffs(~m) - 1;
                break;
            }
            const uint64_t R = x + p - 1;
            // left end, capped at a-(A-1), 0 and the tile start (earlier windows belong to the previous tile)
            uint64_t capL = a >= (uint64_t)(A - 1) ? a - (A - 1) : 0;
            if (capL < i0) capL = i0;
            uint64_t L = a;
            while (true) {
                bool good = (L >= capL + 1 + lane);
                if (good) {
                    uint64_t li = L - 1 - lane;
                    good = sT[li - i0] == sT[li - i0 + p];
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
            for (uint64_t j = L + lane; j <= hi; j += 32)
                if (j - i0 < (uint64_t)NID) sQ[j - i0] = 1;
        }
    }

    // ---- 2. prefix hashes: E[t] = hash of the 16*t tile bytes before thread t's chunk ----
    uint32_t own[4];
    {
        uint4 v = reinterpret_cast<const uint4*>(sT)[tid];
        own[0] = v.x; own[1] = v.y; own[2] = v.z; own[3] = v.w;
        uint32_t h = 0;
#pragma unroll
        for (int k = 0; k < 16; k++) h = h * c.base + ((own[k >> 2] >> ((k & 3) * 8)) & 0xFFu);
        uint32_t x = h;
#pragma unroll
        for (int k = 0; k < 5; k++) {
            uint32_t y = __shfl_up_sync(0xffffffffu, x, 1 << k);
            if (lane >= (1 << k)) x = y * c.pw16[k] + x;
        }
        uint32_t e = __shfl_up_sync(0xffffffffu, x, 1);
        if (lane == 0) e = 0;
        if (lane == 31) s_warp[w] = x;
        __syncthreads();   // also publishes sQ
        if (w == 0) {
            uint32_t xw = lane < NW ? s_warp[lane] : 0;
            uint32_t orig = xw;
#pragma unroll
            for (int k = 0; k < 5; k++) {
                uint32_t y = __shfl_up_sync(0xffffffffu, xw, 1 << k);
                if (lane >= (1 << k)) xw = y * c.pw512[k] + xw;
            }
            // exclusive: hash of all warps before this one
            uint32_t ex = __shfl_up_sync(0xffffffffu, xw, 1);
            if (lane == 0) ex = 0;
            (void)orig;
            if (lane < NW) s_warp[lane] = ex;
        }
        __syncthreads();
        sE[tid] = s_warp[w] * c.lanepow[lane] + e;
    }
    __syncthreads();

    // ---- ids of the 16 own windows (threads < NID/16) ----
    const bool has_ids = tid < NID / CH;
    uint32_t id[CH];
    uint32_t qcount = 0;
    if (has_ids) {
        uint32_t fp = sE[tid + SEG] - sE[tid] * c.btau;
        uint4 v = reinterpret_cast<const uint4*>(sT + TAU)[tid];  // bytes entering the window
        uint32_t in4[4] = { v.x, v.y, v.z, v.w };
        uint4 qv = reinterpret_cast<const uint4*>(sQ)[tid];
        uint32_t q4[4] = { qv.x, qv.y, qv.z, qv.w };
        const uint64_t j0 = i0 + (uint64_t)tid * CH;
#pragma unroll
        for (int k = 0; k < CH; k++) {
            const uint64_t j = j0 + k;
            uint32_t q = (q4[k >> 2] >> ((k & 3) * 8)) & 0xFFu;
            uint32_t v_id = q ? ID_INF : ((fp & 0x7FFFFFFEu) | 1u);
            if (j >= nw) v_id = (j == nw) ? 0u : ID_INF;
            id[k] = v_id;
            if (q && (j - i0 < (uint64_t)B || tile == gridDim.x - 1) && j < nw) qcount++;   // the last tile also owns the windows past its candidates
            uint32_t bo = (own[k >> 2] >> ((k & 3) * 8)) & 0xFFu;
            uint32_t bi = (in4[k >> 2] >> ((k & 3) * 8)) & 0xFFu;
            fp = fp * c.base + bi - c.btau * bo;
        }
    } else {
#pragma unroll
        for (int k = 0; k < CH; k++) id[k] = ID_INF;
    }

    // ---- 3. prefix / suffix minima inside the tau-aligned block (SEG threads) ----
    uint32_t pre[CH], suf[CH];
    {
        uint32_t m = ID_INF;
#pragma unroll
        for (int k = 0; k < CH; k++) { m = min(m, id[k]); pre[k] = m; }
        m = ID_INF;
#pragma unroll
        for (int k = CH - 1; k >= 0; k--) { m = min(m, id[k]); suf[k] = m; }
    }
    uint32_t pre_ex = ID_INF, suf_ex = ID_INF;   // minima of the other threads of the block before / after
    if (SEG > 1) {
        constexpr int WSEG = SEG < 32 ? SEG : 32;     // segment width inside a warp
        const int sl = lane & (WSEG - 1);
        uint32_t a = pre[CH - 1];                      // thread aggregate
        uint32_t x = a;
#pragma unroll
        for (int d = 1; d < WSEG; d <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, x, d);
            if (sl >= d) x = min(x, y);
        }
        uint32_t ex = __shfl_up_sync(0xffffffffu, x, 1);
        pre_ex = sl == 0 ? ID_INF : ex;
        uint32_t z = a;
#pragma unroll
        for (int d = 1; d < WSEG; d <<= 1) {
            uint32_t y = __shfl_down_sync(0xffffffffu, z, d);
            if (sl + d < WSEG) z = min(z, y);
        }
        uint32_t exs = __shfl_down_sync(0xffffffffu, z, 1);
        suf_ex = sl == WSEG - 1 ? ID_INF : exs;
        if (SEG > 32) {   // block spans SEG/32 warps
            constexpr int WPS = SEG > 32 ? SEG / 32 : 1;
            if (lane == 31) s_wmin_pre[w] = x;     // warp minimum
            __syncthreads();
            const int w0 = w & ~(WPS - 1);
            for (int ww = w0; ww < w; ww++) pre_ex = min(pre_ex, s_wmin_pre[ww]);
            for (int ww = w + 1; ww < w0 + WPS; ww++) suf_ex = min(suf_ex, s_wmin_pre[ww]);
        }
    }
    // ---- exchange: prefix minima and ids of the own windows ----
    if (has_ids) {
#pragma unroll
        for (int k = 0; k < CH; k += 4) {
            reinterpret_cast<uint4*>(sPre)[(tid * CH + k) >> 2] =
                make_uint4(min(pre[k], pre_ex), min(pre[k + 1], pre_ex), min(pre[k + 2], pre_ex), min(pre[k + 3], pre_ex));
            reinterpret_cast<uint4*>(sId)[(tid * CH + k) >> 2] = make_uint4(id[k], id[k + 1], id[k + 2], id[k + 3]);
        }
    }
    __syncthreads();

    // ---- 4. selection ----
    uint32_t flags = 0;
    if (tid < B / CH) {
        const uint64_t p0 = i0 + (uint64_t)tid * CH;
#pragma unroll
        for (int k = 0; k < CH; k++) {
            uint32_t pr = sPre[tid * CH + k + TAU];
            uint32_t idr = sId[tid * CH + k + TAU];
            uint32_t sf = min(suf[k], suf_ex);
            uint32_t M = min(sf, pr);
            bool sel = (M != ID_INF) && (id[k] == M || idr == M) && (p0 + k < cnt_pos);
            flags |= (uint32_t)sel << k;
        }
    }
    const uint32_t mycnt = __popc(flags);
    // block exclusive scan of the counts
    uint32_t inc = mycnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += y;
    }
    if (lane == 31) s_cnt[w] = inc;
    if (qcount) atomicAdd(&s_q, qcount);
    __syncthreads();
    if (w == 0) {
        uint32_t x = lane < NW ? s_cnt[lane] : 0, xi = x;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, xi, d);
            if (lane >= d) xi += y;
        }
        if (lane < NW) s_cnt[lane] = xi - x;
        const uint32_t total = __shfl_sync(0xffffffffu, xi, 31);
        if (lane == 0) {
            // decoupled look-back over the tiles with smaller tickets
            unsigned long long prefix = 0;
            if (tile == 0) {
                atomicExch(&status[0], ST_INC | total);
            } else {
                atomicExch(&status[tile], ST_AGG | total);
                long long j = (long long)tile - 1;
                while (true) {
                    unsigned long long v = *((volatile unsigned long long*)&status[j]);
                    if ((v >> 62) == 0) continue;
                    prefix += v & ST_MASK;
                    if ((v >> 62) == 2) break;
                    j--;
                }
                atomicExch(&status[tile], ST_INC | (prefix + total));
            }
            s_base = prefix;
            if (tile == gridDim.x - 1) counters[0] = prefix + total;
            if (s_q) atomicAdd(&counters[1], (unsigned long long)s_q);
        }
    }
    __syncthreads();
    if (flags) {
        unsigned long long o = s_base + s_cnt[w] + (inc - mycnt);
        const uint64_t p0 = i0 + (uint64_t)tid * CH;
        while (flags) {
            int k = __ffs(flags) - 1;
            flags &= flags - 1;
            if (o < s_cap) S_out[o] = (pos_t)(p0 + k);
            o++;
        }
    }
}

uint32_t pow_u32(uint32_t b, uint64_t e) {
    uint32_t r = 1;
    while (e) {
        if (e & 1) r *= b;
        b *= b;
        e >>= 1;
    }
    return r;
}

template <int TAU, int NT, typename pos_t>
void launch(const uint8_t* d_text, uint64_t n, const sss_consts& c, pos_t* d_S, uint64_t s_cap, sss_scratch& sc,
            cudaStream_t st) {
    constexpr int TILE = NT * CH, B = TILE - 2 * TAU, NID = B + TAU;
    uint64_t cnt_pos = n - 2 * (uint64_t)TAU + 2;
    uint64_t ntiles = (cnt_pos + B - 1) / B;
    sc.status.reserve(ntiles + 4);
    CUDA_CHECK(cudaMemsetAsync(sc.status.get(), 0, (ntiles + 4) * 8, st));
    CUDA_CHECK(cudaMemsetAsync(sc.ticket.get(), 0, 4, st));
    CUDA_CHECK(cudaMemsetAsync(sc.counters.get(), 0, 16, st));
    size_t smem = ((TILE + c.A + 16 + 15) & ~15) + NID + (size_t)NID * 8 + (size_t)NT * 4;
    auto k = sss_kernel<TAU, NT, pos_t>;
    CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<(unsigned)ntiles, NT, smem, st>>>(d_text, n, c, d_S, s_cap, sc.status.get(), sc.ticket.get(), sc.counters.get());
    KERNEL_CHECK();
}

template <typename pos_t>
void dispatch(const uint8_t* d_text, uint64_t n, uint32_t tau, const sss_consts& c, pos_t* d_S, uint64_t s_cap,
              sss_scratch& sc, cudaStream_t st) {
    switch (tau) {
        case 16: launch<16, 512, pos_t>(d_text, n, c, d_S, s_cap, sc, st); break;
        case 32: launch<32, 512, pos_t>(d_text, n, c, d_S, s_cap, sc, st); break;
        case 64: launch<64, 512, pos_t>(d_text, n, c, d_S, s_cap, sc, st); break;
        case 128: launch<128, 512, pos_t>(d_text, n, c, d_S, s_cap, sc, st); break;
        case 256: launch<256, 512, pos_t>(d_text, n, c, d_S, s_cap, sc, st); break;
        case 512: launch<512, 512, pos_t>(d_text, n, c, d_S, s_cap, sc, st); break;
        case 1024: launch<1024, 1024, pos_t>(d_text, n, c, d_S, s_cap, sc, st); break;
        case 2048: launch<2048, 1024, pos_t>(d_text, n, c, d_S, s_cap, sc, st); break;
        case 4096: launch<4096, 1024, pos_t>(d_text, n, c, d_S, s_cap, sc, st); break;
        default: throw std::invalid_argument("sss: tau must be a power of two in [4,4096]");
    }
}

}  // namespace

uint32_t sss_period_bound(uint32_t tau, uint32_t qmax) { return tau / 3 < qmax ? tau / 3 : qmax; }
uint32_t sss_anchor_spacing(uint32_t tau, uint32_t P) {
    uint32_t lim = tau - 2 * P, a = 1;
    while (a * 2 <= lim) a *= 2;
    return a;
}

// d_text must be readable (zero-filled) up to n + TEXT_PAD bytes.
void sss_build(const uint8_t* d_text, uint64_t n, uint32_t tau, uint32_t base, uint32_t qmax, void* d_S, int pos_bytes,
               uint64_t s_cap, sss_scratch& sc, cudaStream_t st, uint64_t* h_size, uint64_t* h_numq) {
    *h_size = 0;
    *h_numq = 0;
    if (n + 1 < 2 * (uint64_t)tau) return;
    sc.ticket.reserve(1);
    sc.counters.reserve(4);
    if (tau < 16) {
        sss_build_small_tau(d_text, n, tau, base, qmax, d_S, pos_bytes, s_cap, sc, st);
    } else if (tau >= 128) {
        sss_build_cell(d_text, n, tau, base, qmax, d_S, pos_bytes, s_cap, sc, st);
    } else {
        sss_consts c;
        c.base = base;
        c.btau = pow_u32(base, tau);
        for (int k = 0; k < 5; k++) {
            c.pw16[k] = pow_u32(base, 16ull << k);
            c.pw512[k] = pow_u32(base, 512ull << k);
        }
        for (int l = 0; l < 32; l++) c.lanepow[l] = pow_u32(base, 16ull * l);
        c.P = sss_period_bound(tau, qmax);
        c.A = sss_anchor_spacing(tau, c.P);
        if (pos_bytes == 4) dispatch<uint32_t>(d_text, n, tau, c, (uint32_t*)d_S, s_cap, sc, st);
        else dispatch<uint64_t>(d_text, n, tau, c, (uint64_t*)d_S, s_cap, sc, st);
    }
    unsigned long long h[2];
    CUDA_CHECK(cudaMemcpyAsync(h, sc.counters.get(), 16, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    *h_size = h[0];
    *h_numq = h[1];
}

}  // namespace lz77sss
