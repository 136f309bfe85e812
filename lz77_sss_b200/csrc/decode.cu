// K6: parallel LZ77 decode.  Replaces lz77_sss<pos_t>::decode
// (/root/reference/include/lz77_sss/algorithms/common.cpp:33-54), a sequential byte-by-byte copy loop whose
// every output byte may depend on an earlier one.
//
//   1. output offsets = exclusive scan of max(1,len) over the factors,
//   2. every output byte gets a pointer to the byte it copies (or its literal value),
//   3. pointer jumping: a byte whose source is resolved takes its value, otherwise it adopts the source's
//      pointer; O(log depth) rounds.  A resolved byte is published as one 16-bit word (0x100 | value) so that
//      "resolved" and the value are read atomically.
// Algorithmic bytes: 2n + z*sizeof(factor); the first version spends ~10 B per byte and round.
#include "common.cuh"
#include "stages.cuh"

namespace lz77sss {
namespace {

__global__ void factor_len_kernel(const uint64_t* __restrict__ len, uint64_t z, uint32_t* __restrict__ out) {
    uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k < z) out[k] = len[k] ? (uint32_t)len[k] : 1u;
}

// one warp per factor: initialise the pointers / literals of its output range
__global__ void __launch_bounds__(256) decode_init_kernel(const uint64_t* __restrict__ src,
                                                          const uint64_t* __restrict__ len,
                                                          const uint32_t* __restrict__ start, uint64_t z, uint64_t n,
                                                          uint32_t* __restrict__ ptr, uint16_t* __restrict__ res,
                                                          int* __restrict__ err) {
    const uint64_t k = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    if (k >= z) return;
    const uint64_t st = start[k], l = len[k], s = src[k];
    if (l == 0) {
        if (lane == 0) {
            if (st < n) res[st] = (uint16_t)(0x100u | (uint32_t)(s & 0xFF));
            else *err = 2;
        }
        return;
    }
    if (s >= st || st + l > n) {
        if (lane == 0) *err = 2;
        return;
    }
    for (uint64_t i = lane; i < l; i += 32) {
        ptr[st + i] = (uint32_t)(s + i);
        res[st + i] = 0;
    }
}

__global__ void __launch_bounds__(256) decode_round_kernel(uint64_t n, uint32_t* __restrict__ ptr,
                                                           uint16_t* __restrict__ res,
                                                           unsigned long long* __restrict__ unresolved) {
    const uint64_t x = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool pending = false;
    if (x < n) {
        uint16_t r = res[x];
        if (r == 0) {
            const uint32_t y = ptr[x];
            const uint16_t ry = *((volatile uint16_t*)&res[y]);
            if (ry) {
                res[x] = ry;
            } else {
                ptr[x] = *((volatile uint32_t*)&ptr[y]);
                pending = true;
            }
        }
    }
    unsigned m = __ballot_sync(0xffffffffu, pending);
    if (m && (threadIdx.x & 31) == 0) atomicAdd(unresolved, (unsigned long long)__popc(m));
}

__global__ void decode_pack_kernel(const uint16_t* __restrict__ res, uint64_t n, uint8_t* __restrict__ out) {
    const uint64_t x = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (x < n) out[x] = (uint8_t)(res[x] & 0xFF);
}

}  // namespace

int decode_factors(const uint64_t* d_src, const uint64_t* d_len, uint64_t z, uint8_t* d_out, uint64_t n,
                   cudaStream_t st, int* rounds) {
    if (rounds) *rounds = 0;
    if (n == 0) return 0;
    if (n >= 0xFFFFFFF0ull) return -3;   // 32-bit byte pointers in this version
    static thread_local scan_scratch scan;
    dbuf<uint32_t> start, ptr, total;
    dbuf<uint16_t> res;
    dbuf<int> err;
    dbuf<unsigned long long> cnt;
    start.reserve(z + 1); ptr.reserve(n); res.reserve(n); total.reserve(1); err.reserve(1); cnt.reserve(1);
    CUDA_CHECK(cudaMemsetAsync(err.get(), 0, 4, st));
    factor_len_kernel<<<ceil_div_u(z, 256), 256, 0, st>>>(d_len, z, start);
    KERNEL_CHECK();
    exclusive_scan_u32(start, z, scan, st, total.get());
    uint32_t h_total;
    CUDA_CHECK(cudaMemcpyAsync(&h_total, total.get(), 4, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    if (h_total != n) return -1;   // factor lengths do not add up to n
    decode_init_kernel<<<ceil_div_u(z * 32, 256), 256, 0, st>>>(d_src, d_len, start, z, n, ptr, res, err);
    KERNEL_CHECK();
    int h_err;
    CUDA_CHECK(cudaMemcpyAsync(&h_err, err.get(), 4, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    if (h_err) return -2;
    for (int r = 0; r < 80; r++) {
        CUDA_CHECK(cudaMemsetAsync(cnt.get(), 0, 8, st));
        decode_round_kernel<<<ceil_div_u(n, 256), 256, 0, st>>>(n, ptr, res, cnt);
        KERNEL_CHECK();
        unsigned long long h_cnt;
        CUDA_CHECK(cudaMemcpyAsync(&h_cnt, cnt.get(), 8, cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        if (rounds) *rounds = r + 1;
        if (h_cnt == 0) break;
    }
    decode_pack_kernel<<<ceil_div_u(n, 256), 256, 0, st>>>(res, n, d_out);
    KERNEL_CHECK();
    return 0;
}

}  // namespace lz77sss
