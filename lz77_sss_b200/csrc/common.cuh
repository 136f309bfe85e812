// Shared declarations of the lz77-sss B200 device library (sm_100a only).
// Everything under csrc/ is built into liblz77sss_b200.so; the only exported symbols are
// the extern "C" entry points declared in include/lz77sss_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>

namespace lz77sss {

struct cuda_error : std::runtime_error {
    using std::runtime_error::runtime_error;
};

inline void cuda_check(cudaError_t e, const char* what, const char* file, int line) {
    if (e != cudaSuccess) {
        char buf[512];
        snprintf(buf, sizeof buf, "%s failed at %s:%d: %s", what, file, line, cudaGetErrorString(e));
        throw cuda_error(buf);
    }
}
#define CUDA_CHECK(x) ::lz77sss::cuda_check((x), #x, __FILE__, __LINE__)
// every kernel launch of the library goes through KERNEL_CHECK, which also counts it (bench.py's gpu_launches)
extern thread_local unsigned long long g_kernel_launches;   // per host thread: handles driven from different threads count separately
extern bool g_sync_every_kernel;   // LZ77SSS_SYNC_KERNELS=1: device-synchronize after every launch, so a fault names its kernel
#define KERNEL_CHECK()                                                                    \
    do {                                                                                  \
        ::lz77sss::g_kernel_launches++;                                                   \
        ::lz77sss::cuda_check(cudaGetLastError(), "kernel launch", __FILE__, __LINE__);   \
        if (::lz77sss::g_sync_every_kernel)                                               \
            ::lz77sss::cuda_check(cudaDeviceSynchronize(), "kernel execution", __FILE__, __LINE__); \
    } while (0)

// Optional per-kernel timing of the heavy kernels (lz77sss_kernel_profile): CUDA events around the launch on its own
// stream, one synchronization per launch while enabled (so a profiled run is not a timed run).
extern bool g_prof_on;
void prof_begin(cudaStream_t st);
void prof_end(cudaStream_t st, const char* name, uint64_t algorithmic_bytes);
#define PROF_BEGIN(st) do { if (::lz77sss::g_prof_on) ::lz77sss::prof_begin(st); } while (0)
#define PROF_END(st, name, bytes) do { if (::lz77sss::g_prof_on) ::lz77sss::prof_end(st, name, bytes); } while (0)

// Owning device buffer (cudaMallocAsync on the handle's stream would also do; plain cudaMalloc keeps
// the allocation pattern visible and the buffers are reused across calls of one handle).
template <typename T>
struct dbuf {
    T* p = nullptr;
    size_t cap = 0;
    dbuf() = default;
    dbuf(const dbuf&) = delete;
    dbuf& operator=(const dbuf&) = delete;
    dbuf(dbuf&& o) noexcept : p(o.p), cap(o.cap) { o.p = nullptr; o.cap = 0; }
    dbuf& operator=(dbuf&& o) noexcept {
        if (this != &o) { release(); p = o.p; cap = o.cap; o.p = nullptr; o.cap = 0; }
        return *this;
    }
    ~dbuf() { release(); }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    // grow-only; contents are NOT preserved
    void reserve(size_t n) {
        if (n <= cap) return;
        release();
        size_t want = n + n / 8 + 64;
        CUDA_CHECK(cudaMalloc(&p, want * sizeof(T)));
        cap = want;
    }
    T* get() const { return p; }
    operator T*() const { return p; }
};

// Pinned host array that grows and keeps its contents (factor output: device-to-host copies into pageable memory are
// staged by the driver and run at a fraction of the link rate).
template <typename T>
struct pinned_vec {
    T* p = nullptr;
    size_t n = 0, cap = 0;
    pinned_vec() = default;
    pinned_vec(const pinned_vec&) = delete;
    pinned_vec& operator=(const pinned_vec&) = delete;
    ~pinned_vec() { if (p) cudaFreeHost(p); }
    size_t size() const { return n; }
    void clear() { n = 0; }
    T* data() { return p; }
    const T* data() const { return p; }
    T& operator[](size_t i) { return p[i]; }
    const T& operator[](size_t i) const { return p[i]; }
    void resize(size_t m) {
        if (m > cap) {
            const size_t want = m + m / 2 + 4096;
            T* q = nullptr;
            CUDA_CHECK(cudaMallocHost(&q, want * sizeof(T)));
            if (n) memcpy(q, p, n * sizeof(T));
            if (p) cudaFreeHost(p);
            p = q;
            cap = want;
        }
        n = m;
    }
};

static inline unsigned ceil_div_u(uint64_t a, uint64_t b) { return (unsigned)((a + b - 1) / b); }

// ---- utility device primitives implemented in util.cu ------------------------------------------
// exclusive prefix sum (uint32 values, uint32 sums) in place; n up to 2^31. tmp is scratch.
struct scan_scratch {
    dbuf<uint32_t> lvl1, lvl2;
};
void exclusive_scan_u32(uint32_t* d, uint64_t n, scan_scratch& s, cudaStream_t st, uint32_t* d_total = nullptr);

// stable LSD radix sort of (key64, val32) pairs on bits [bit_lo, bit_hi) of the key.
// Result ends in (keys, vals) (ping-pong handled inside; kalt/valt are scratch of the same size).
struct sort_scratch {
    dbuf<uint32_t> hist, offs;  // hist[tile][digit] and the global offsets made of it
    dbuf<uint32_t> csum, cpre;  // per chunk of tiles: column sums; column prefixes and, behind them, the digit bases
    scan_scratch scan;
};
void radix_sort_pairs(uint64_t*& keys, uint64_t*& kalt, uint32_t*& vals, uint32_t*& valt, uint64_t n, int bit_lo,
                      int bit_hi, sort_scratch& s, cudaStream_t st);

void radix_sort_keys(uint64_t*& keys, uint64_t*& kalt, uint64_t n, int bit_lo, int bit_hi, sort_scratch& s, cudaStream_t st);

// same with 32-bit keys (gap index entries: bucket id -> entry id)
// vals_are_iota: the values are 0, 1, 2, ... and need not be read (nor written by the caller) before the first pass
// first_keys: the unsorted keys live in this read-only array (keys is then only a ping-pong buffer)
void radix_sort_pairs32(uint32_t*& keys, uint32_t*& kalt, uint32_t*& vals, uint32_t*& valt, uint64_t n, int bit_lo,
                        int bit_hi, sort_scratch& s, cudaStream_t st, bool vals_are_iota = false,
                        const uint32_t* first_keys = nullptr);

void fill_u32(uint32_t* d, uint64_t n, uint32_t v, cudaStream_t st);
void iota_u32(uint32_t* d, uint64_t n, cudaStream_t st);

}  // namespace lz77sss
