// tau-synchronizing set for tau in {4, 8}: every quantity a position needs fits in 3*tau bytes, so one
// thread evaluates the definition directly from registers (same definition as sss.cu / DESIGN.md §SSS).
// Two passes (count per block, scan, write) — S is then as large as the text itself, the text read is
// not the dominant traffic.
#include "common.cuh"
#include "stages.cuh"

namespace lz77sss {
namespace {

constexpr uint32_t ID_INF = 0xFFFFFFFFu;
constexpr int NT = 256;

template <int TAU>
__device__ __forceinline__ bool select_pos(const uint8_t* __restrict__ T, uint64_t n, uint64_t i, uint32_t base,
                                           uint32_t P, bool* q_here) {
    const uint64_t nw = n - TAU + 1;
    uint8_t b[3 * TAU];
#pragma unroll
    for (int k = 0; k < 3 * TAU; k++) b[k] = T[i + k];    // padded text: always readable
    uint32_t id[TAU + 1];
#pragma unroll
    for (int d = 0; d <= TAU; d++) {
        const uint64_t j = i + d;
        uint32_t fp = 0;
#pragma unroll
        for (int k = 0; k < TAU; k++) fp = fp * base + b[d + k];
        bool q = false;
        for (uint32_t p = 1; p <= P; p++) {
            bool ok = true;
#pragma unroll
            for (int x = 0; x < TAU; x++)
                if (x + (int)p < TAU && b[d + x] != b[d + x + p]) ok = false;
            q |= ok;
        }
        uint32_t v = q ? ID_INF : ((fp & 0x7FFFFFFEu) | 1u);
        if (j >= nw) v = (j == nw) ? 0u : ID_INF;
        if (d == 0) *q_here = q && j < nw;
        id[d] = v;
    }
    uint32_t M = ID_INF;
#pragma unroll
    for (int d = 0; d <= TAU; d++) M = min(M, id[d]);
    return M != ID_INF && (id[0] == M || id[TAU] == M);
}

template <int TAU, typename pos_t, bool WRITE>
__global__ void __launch_bounds__(NT) small_kernel(const uint8_t* __restrict__ T, uint64_t n, uint32_t base, uint32_t P,
                                                   uint32_t* __restrict__ blk, pos_t* __restrict__ S_out, uint64_t s_cap,
                                                   unsigned long long* __restrict__ counters) {
    const uint64_t cnt_pos = n - 2 * (uint64_t)TAU + 2;
    const uint64_t i = (uint64_t)blockIdx.x * NT + threadIdx.x;
    bool q = false;
    bool sel = false;
    if (i < n - TAU + 1 || i < cnt_pos) sel = select_pos<TAU>(T, n, i, base, P, &q) && i < cnt_pos;
    __shared__ uint32_t wcnt[NT / 32];
    __shared__ uint32_t qc;
    if (threadIdx.x == 0) qc = 0;
    __syncthreads();
    const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    unsigned m = __ballot_sync(0xffffffffu, sel);
    if (lane == 0) wcnt[w] = __popc(m);
    if (!WRITE) {
        unsigned mq = __ballot_sync(0xffffffffu, q && i < n - TAU + 1);
        if (lane == 0 && mq) atomicAdd(&qc, __popc(mq));
    }
    __syncthreads();
    if (!WRITE) {
        if (threadIdx.x == 0) {
            uint32_t t = 0;
            for (int k = 0; k < NT / 32; k++) t += wcnt[k];
            blk[blockIdx.x] = t;
            if (qc) atomicAdd(&counters[1], (unsigned long long)qc);
        }
    } else {
        uint32_t off = blk[blockIdx.x];
        for (unsigned k = 0; k < w; k++) off += wcnt[k];
        if (sel) {
            uint64_t o = (uint64_t)off + __popc(m & ((1u << lane) - 1));
            if (o < s_cap) S_out[o] = (pos_t)i;
        }
    }
}

__global__ void store_total(const uint32_t* total, unsigned long long* counters) { counters[0] = *total; }

template <int TAU, typename pos_t>
void run(const uint8_t* d_text, uint64_t n, uint32_t base, uint32_t P, pos_t* d_S, uint64_t s_cap, sss_scratch& sc,
         cudaStream_t st) {
    uint64_t cnt_pos = n - 2 * (uint64_t)TAU + 2;
    // Q windows beyond the last candidate still count for |Q|: cover all windows
    uint64_t npos = n - TAU + 1;
    if (npos < cnt_pos) npos = cnt_pos;
    unsigned nb = ceil_div_u(npos, NT);
    sc.blk.reserve(nb + 2);
    CUDA_CHECK(cudaMemsetAsync(sc.counters.get(), 0, 16, st));
    small_kernel<TAU, pos_t, false><<<nb, NT, 0, st>>>(d_text, n, base, P, sc.blk.get(), d_S, s_cap, sc.counters.get());
    KERNEL_CHECK();
    exclusive_scan_u32(sc.blk.get(), nb, sc.scan, st, sc.blk.get() + nb + 1);
    store_total<<<1, 1, 0, st>>>(sc.blk.get() + nb + 1, sc.counters.get());
    KERNEL_CHECK();
    small_kernel<TAU, pos_t, true><<<nb, NT, 0, st>>>(d_text, n, base, P, sc.blk.get(), d_S, s_cap, sc.counters.get());
    KERNEL_CHECK();
}

}  // namespace

void sss_build_small_tau(const uint8_t* d_text, uint64_t n, uint32_t tau, uint32_t base, uint32_t qmax, void* d_S,
                         int pos_bytes, uint64_t s_cap, sss_scratch& sc, cudaStream_t st) {
    uint32_t P = sss_period_bound(tau, qmax);
    if (tau == 4) {
        if (pos_bytes == 4) run<4, uint32_t>(d_text, n, base, P, (uint32_t*)d_S, s_cap, sc, st);
        else run<4, uint64_t>(d_text, n, base, P, (uint64_t*)d_S, s_cap, sc, st);
    } else if (tau == 8) {
        if (pos_bytes == 4) run<8, uint32_t>(d_text, n, base, P, (uint32_t*)d_S, s_cap, sc, st);
        else run<8, uint64_t>(d_text, n, base, P, (uint64_t*)d_S, s_cap, sc, st);
    } else {
        throw std::invalid_argument("sss: tau must be a power of two in [4,4096]");
    }
}

}  // namespace lz77sss
