/* lz77sss_b200.h — C ABI of the B200-native lz77-sss hot path (liblz77sss_b200.so).
 *
 * The reference (LukasNalbach/lz77-sss) has no FFI layer: its boundary is the header-only template API
 *   lz77_sss<pos_t>::factorize_approximate<fact_mode, phr_mode, tau>(input, n, output, {num_threads, log})
 *   lz77_sss<pos_t>::factorize_exact<...>(...)      lz77_sss<pos_t>::decode(fact_it, out_it, n)
 * (/root/reference/include/lz77_sss/lz77_sss.hpp:176-203).  lz77_sss_b200/include/lz77_sss/lz77_sss.hpp keeps those
 * templates and forwards to the entry points below; INTEGRATION.md shows the reference-side binding.
 *
 * Plain C types only; no CUDA or torch types.  All functions return 0 on success or a negative
 * LZ77SSS_E_* code; lz77sss_last_error() gives the message.  There is no CPU fallback: without a
 * CUDA device every compute call fails with LZ77SSS_E_CUDA.
 */
#ifndef LZ77SSS_B200_H
#define LZ77SSS_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define LZ77SSS_OK 0
#define LZ77SSS_E_ARG (-1)       /* invalid argument (tau not a power of two in [4,4096], n == 0, ...) */
#define LZ77SSS_E_CUDA (-2)      /* CUDA runtime error / no device */
#define LZ77SSS_E_UNSUPPORTED (-3) /* mode not built yet (see DESIGN.md scope table) */
#define LZ77SSS_E_INTERNAL (-4)  /* an internal invariant failed; the result must not be used */
#define LZ77SSS_E_STATE (-5)     /* introspection call without a preceding factorization */

/* enum values of the reference: lz77_sss.hpp:47-58 */
#define LZ77SSS_PHR_LPF_NAIVE 0
#define LZ77SSS_PHR_LPF_LNF_NAIVE 1
#define LZ77SSS_PHR_LPF_OPT 2
#define LZ77SSS_PHR_LPF_LNF_OPT 3
#define LZ77SSS_FACT_GREEDY_NAIVE 0
#define LZ77SSS_FACT_GREEDY 1
#define LZ77SSS_FACT_SKIP_PHRASES 2

typedef struct lz77sss_handle lz77sss_handle;

/* Everything the factor stream is a function of.  The reference draws its
 * rolling-hash bases from std::random_device (rolling_hash.hpp:30-39,127-130); here they are explicit. */
typedef struct lz77sss_config {
    uint32_t struct_size;   /* = sizeof(lz77sss_config) */
    uint32_t tau;           /* template parameter tau (lz77_sss.hpp:179); power of two, 4..4096 */
    uint32_t num_threads;   /* parameters::num_threads (lz77_sss.hpp:67-70): the LOGICAL partition count p of
                               build_LPF_opt (lpf_opt.cpp:50-56); it changes the factor stream, so it is kept as a
                               parameter and is independent of the number of GPUs.  0 -> 1 */
    uint32_t pos_bits;      /* 32 or 64: sizeof(pos_t)*8 of the replaced lz77_sss<pos_t> (gap table sizing,
                               rolling_hash_index_107.hpp:59-67); 0 -> 32 iff n <= UINT32_MAX (cli/lz77_sss_3_aprx.cpp:73) */
    uint32_t phr_mode;      /* LZ77SSS_PHR_* (all four phrase modes) */
    uint32_t fact_mode;     /* LZ77SSS_FACT_* (all three; skip_phrases emits the gapped stream of skip_gaps.cpp:34-61) */
    uint32_t sss_base;      /* multiplier of the window fingerprint of the synchronizing set; 0 -> 0x9E3779B1 */
    uint32_t sss_qmax;      /* period bound cap of the low-period set Q; 0xFFFFFFFF -> 32 */
    uint64_t gap_bases[5];  /* bases of the five rk_prime<107> of the gap index (rolling_hash_index_107.hpp:72-75),
                               each in [257, 2^20-1] (rolling_hash.hpp:31-36) */
    int32_t device;         /* CUDA device ordinal */
    uint32_t log;           /* parameters::log: print phase times to stdout */
} lz77sss_config;

/* phase times of the last factorization, nanoseconds, keyed like the reference's log_phase names
 *  plus the copies. */
typedef struct lz77sss_stats {
    uint64_t n, size_sss, num_q, num_factors;
    uint64_t num_lpf, len_lpf_phr, num_gaps, patt_lens[5], roll_threshold, table_entries;
    uint64_t ctx_levels, doubling_rounds, gap_positions_hashed;
    uint64_t t_h2d_ns, t_sss_ns, t_sort_ns, t_lce_ns, t_lpf_ns, t_d2h_ns, t_chain_ns, t_gap_hash_ns, t_emit_ns, t_total_ns;
    uint64_t kernel_launches;
    uint64_t h2d_bytes, d2h_bytes;   /* bytes copied host->device / device->host inside the call */
    uint64_t parse_rounds, parse_slices;   /* device emission: block-parse rounds over all slices, slices */
    uint64_t sparse_slices, sparse_fallbacks;  /* slices parsed with gap-only index entries / redone with all positions */
    uint64_t repair_entries;               /* index entries added while parsing (positions under gap factors that ran over phrases) */
    uint64_t t_exact_ns, exact_doubling_rounds, exact_parse_rounds;   /* lz77sss_factorize_exact: suffix array + parse */
    uint64_t prefetched_slices;            /* slices whose index was built on the second stream while the slice before was parsed */
} lz77sss_stats;

/* factor callback: lz77_sss<pos_t>::factor {src,len} by value, in text order, on the calling thread
 * (lz77_sss.hpp:129-131; len == 0 means a literal whose byte is src). */
typedef void (*lz77sss_emit_fn)(void* ctx, uint64_t src, uint64_t len);

/* Ready-made callback for callers that just want the factors in two arrays (bindings from languages where a
 * per-factor callback is expensive): pass lz77sss_sink_emit as emit and a lz77sss_factor_sink* as ctx.
 * count keeps counting past cap; only the first cap factors are stored. */
typedef struct lz77sss_factor_sink {
    uint64_t* src;
    uint64_t* len;
    uint64_t cap;
    uint64_t count;
} lz77sss_factor_sink;
void lz77sss_sink_emit(void* ctx, uint64_t src, uint64_t len);

const char* lz77sss_last_error(void);
const char* lz77sss_version(void);
int lz77sss_device_count(void);

int lz77sss_create(const lz77sss_config* cfg, lz77sss_handle** out);
void lz77sss_destroy(lz77sss_handle* h);

/* lz77_sss<pos_t>::factorize_approximate<greedy, lpf_opt, tau> (lz77_sss.hpp:176-186).
 * text: HOST pointer to n bytes (no padding required); copied to the device inside the call. */
int lz77sss_factorize_approximate(lz77sss_handle* h, const uint8_t* text, uint64_t n, lz77sss_emit_fn emit, void* ctx);

/* lz77_sss<pos_t>::factorize_exact<fact_mode, phr_mode, transf_mode, range_ds_t, tau> (lz77_sss.hpp:188-200): the exact LZ77
 * factorization; with num_threads > 1 exact per section (16 sections per thread, cut at phrase boundaries of the
 * approximation, transform_to_exact/common.cpp:48-74).  The approximation runs first with the handle's phr_mode / fact_mode
 * (fact_mode must not be skip_phrases, lz77_sss.hpp:319); then the exact parse is computed from the suffix array of the text
 * (csrc/exact.cu).  Factor LENGTHS equal the reference's for every transform_mode and range structure; sources are valid
 * earlier occurrences (the reference's own choice depends on its range structure's scan order).  n < 2^32 - 2^16. */
int lz77sss_factorize_exact(lz77sss_handle* h, const uint8_t* text, uint64_t n, lz77sss_emit_fn emit, void* ctx);

/* Same, with the text already resident in device memory: d_text is a DEVICE pointer to a buffer obtained from
 * lz77sss_device_text_alloc (which provides the zero padding the kernels rely on).  host_text must hold the
 * same n bytes: the sequential greedy emission runs on the host and reads the text there. */
int lz77sss_device_text_alloc(lz77sss_handle* h, uint64_t n, uint8_t** d_text);
int lz77sss_device_text_upload(lz77sss_handle* h, uint8_t* d_text, const uint8_t* host_text, uint64_t n);
int lz77sss_device_text_free(lz77sss_handle* h, uint8_t* d_text);
int lz77sss_factorize_approximate_device(lz77sss_handle* h, const uint8_t* d_text, const uint8_t* host_text, uint64_t n,
                                         lz77sss_emit_fn emit, void* ctx);

/* ---- one text sharded over several GPUs ------------------------------------------------------------
 * One process per GPU; every rank holds the whole text in HBM (the host layer all-gathers the shards over NVLink) and
 * runs lz77sss_shard_prepare (synchronizing set, SA_S, LPF phrases, gap parameters: replicated, 1/5 of the work).
 * The gap index and the greedy emission (factorize_greedy, greedy.cpp:34-140; rolling_hash_index_107.hpp:33-173) are
 * split by text range: a rank parses [beg, end) from an entry state {i, table cursor, phrase index, gap flag} and
 * the table "last inserted position + 1 per bucket before beg" (device array of lz77sss_shard_table_entries u64).
 * Entry states and tables of ranges behind the first are guesses at first; the host layer exchanges exit states and
 * tables (NCCL all-gather) and re-parses the ranges whose inputs changed until nothing changes.  A re-parse resumes
 * from the sorted entries and traces the handle kept.  The result is the sequential factor stream.
 * lz77_sss_b200/sharded.py is that host layer. */
int lz77sss_shard_prepare(lz77sss_handle* h, const uint8_t* d_text, const uint8_t* host_text, uint64_t n);
int lz77sss_shard_table_entries(lz77sss_handle* h, uint64_t* entries);
int lz77sss_shard_entry_guess(lz77sss_handle* h, uint64_t pos, uint64_t state[4]);
/* ranges of about equal work for `parts` ranks: bounds[0] = 0 <= bounds[1] <= ... <= bounds[parts] = n, multiples of `align`.
 * The work of a range is dominated by its gap bytes (the bytes outside the phrases of build_LPF_*), so equal byte counts are
 * not equal work: the first copy of a repeated sequence is one long gap. */
int lz77sss_shard_split(lz77sss_handle* h, uint32_t parts, uint64_t align, uint64_t* bounds);
/* optional, before the first lz77sss_shard_parse of a range: hash and sort its gap-index entries (kept for the parses) and
 * write a GUESS of the range's table (last position per bucket among the positions the phrase list alone says are
 * inserted); exchanged once, it gives the later ranges a nearly right table for their first parse */
int lz77sss_shard_index(lz77sss_handle* h, uint64_t beg, uint64_t end, uint64_t* d_table_guess_out);
int lz77sss_shard_parse(lz77sss_handle* h, uint64_t beg, uint64_t end, const uint64_t entry[4], const uint64_t* d_table_in,
                        uint64_t* d_table_out, uint64_t exit_state[4], uint64_t* num_factors);
/* factors of the last lz77sss_shard_parse on this handle (HOST arrays) */
int lz77sss_shard_factors(lz77sss_handle* h, uint64_t* src, uint64_t* len, uint64_t cap, uint64_t* count);
/* the same factors without a copy: pointers to the library's PINNED host arrays (valid until the next call on the handle),
 * so that a host layer can send them to a device or a peer at link speed */
int lz77sss_shard_factors_view(lz77sss_handle* h, const uint64_t** src, const uint64_t** len, uint64_t* count);
/* the last 65 positions of the text (end-of-text rules of the gap index), on the rank that parsed the last range,
 * right after its final lz77sss_shard_parse: emits the remaining factors from `state` */
int lz77sss_shard_finish(lz77sss_handle* h, const uint64_t state[4], lz77sss_emit_fn emit, void* ctx);

/* ---- rabin_karp_substring<31, pos_t> (include/lz77_sss/data_structures/rabin_karp_substring.hpp:35-240), the fingerprint
 * structure of the exact path's sample index (sample_index.hpp:139,373).  lz77sss_rks_create copies the text to the
 * device and builds the fingerprints of all prefixes T[0..16k) modulo 2^31 - 1 for the given base (the reference draws
 * it from std::random_device in [257, 2^31 - 2], :89-96).  Residues are canonical, so every value equals the
 * reference's for the same base. */
typedef struct lz77sss_rks lz77sss_rks;
int lz77sss_rks_create(lz77sss_handle* h, const uint8_t* text, uint64_t n, uint64_t base, lz77sss_rks** out);
void lz77sss_rks_destroy(lz77sss_rks* r);
/* table[g] = fingerprint of T[0 .. 16 (g + 1)), g < ceil(n / 16) */
int lz77sss_rks_table(lz77sss_rks* r, uint32_t* out, uint64_t cap, uint64_t* count);
/* the reference's `fps` vector for sampling rate s (:131-171): fps[k] = fingerprint of T[0 .. min(k s, n)) */
int lz77sss_rks_fps(lz77sss_rks* r, uint64_t s, uint32_t* out, uint64_t cap, uint64_t* count);
/* batched rabin_karp_substring::substring(pos, len) (:223-239) */
int lz77sss_rks_substring(lz77sss_rks* r, const uint64_t* pos, const uint64_t* len, uint64_t nq, uint32_t* out);
/* The table of lz77sss_rks_table WITHOUT any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * d_text: DEVICE pointer to T[0..n), 16-byte aligned, zero bytes behind the text up to the next multiple of 16;
 * d_table: DEVICE array of ceil(n / 16) u32, 16-byte aligned; stream: a cudaStream_t (NULL = the default stream) of the
 * CURRENT device.  The call needs no handle, enqueues a memset and one kernel on `stream` and returns without waiting;
 * it calls no cudaMalloc / cudaFree.  The kernel's look-back scratch is a static device array of
 * LZ77SSS_RKS_NOALLOC_MAX_BYTES / 16384 words of 8 bytes (512 KiB, held by every GPU that loads the library until the
 * process ends): n > LZ77SSS_RKS_NOALLOC_MAX_BYTES is rejected with LZ77SSS_E_UNSUPPORTED.  All calls share that array
 * (cleared in front of every launch): calls that overlap on one GPU (two streams without ordering) race on it. */
#define LZ77SSS_RKS_NOALLOC_MAX_BYTES (1ull << 30)
int lz77sss_rks_table_noalloc(const uint8_t* d_text, uint64_t n, uint64_t base, uint32_t* d_table, void* stream);
/* re-run the build `iters` times on the resident text; mean time (CUDA events on the library's stream) */
int lz77sss_rks_time_build(lz77sss_rks* r, int iters, double* ms_out, uint64_t* launches_out);

/* lz77_sss<pos_t>::decode (algorithms/common.cpp:33-54): z factors -> n bytes into the HOST buffer out; any n < 2^63
 * that fits the device (6 bytes of state per output byte below 4 GiB, 9 bytes above).  A malformed stream is
 * LZ77SSS_E_ARG; chains that do not resolve are LZ77SSS_E_INTERNAL (never silently wrong output). */
int lz77sss_decode(lz77sss_handle* h, const uint64_t* src, const uint64_t* len, uint64_t z, uint8_t* out, uint64_t n);

/* Decode timing for bench.py: the factors are uploaded once, the decode kernels run `iters` times on the device (mean time
 * by CUDA events on the library's stream, the output stays in HBM); d_expect (DEVICE pointer from
 * lz77sss_device_text_alloc, or NULL) is compared with the decoded bytes on the device: *equal_out = 1 iff identical. */
int lz77sss_time_decode(lz77sss_handle* h, const uint64_t* src, const uint64_t* len, uint64_t z, uint64_t n, int iters,
                        double* ms_out, int* rounds_out, uint64_t* launches_out, const uint8_t* d_expect, int* equal_out);

/* ---- stage-level entry points (used by the parity tests and by bench.py's kernel timing) ---- */
/* Only the synchronizing set of a host text (stage K1+K2). count_out = |S|; S_out may be NULL. */
int lz77sss_build_sss(lz77sss_handle* h, const uint8_t* text, uint64_t n, uint64_t* S_out, uint64_t cap, uint64_t* count_out,
                      uint64_t* num_q_out);
/* Re-run the K1+K2 kernel `iters` times on a device-resident text and return the mean kernel time (CUDA events). */
int lz77sss_time_sss_kernel(lz77sss_handle* h, const uint8_t* d_text, uint64_t n, int iters, double* ms_out,
                            uint64_t* count_out);
/* Per-kernel timing of the kernels that make up a step (bench.py's `other_kernels`): while enabled, the radix scatter,
 * the gap-index hash kernel and the parse rounds are bracketed by CUDA events on the library's stream (one
 * synchronization per launch, so a profiled call is not a timed call).  Rows accumulate over calls:
 * name, total milliseconds, launches, algorithmic bytes (0 where the kernel is latency-bound by design). */
int lz77sss_kernel_profile(lz77sss_handle* h, int enable);
int lz77sss_kernel_profile_get(lz77sss_handle* h, uint32_t idx, char* name, uint32_t name_cap, double* ms, uint64_t* launches,
                               uint64_t* bytes);
/* Introspection of the last factorization on this handle (arrays of |S| entries unless noted). */
int lz77sss_get_stats(lz77sss_handle* h, lz77sss_stats* out);
int lz77sss_get_sss(lz77sss_handle* h, uint64_t* out, uint64_t cap, uint64_t* count);
int lz77sss_get_sa_s(lz77sss_handle* h, uint32_t* out, uint64_t cap, uint64_t* count);
int lz77sss_get_lcp_s(lz77sss_handle* h, uint64_t* out, uint64_t cap, uint64_t* count);
/* LPF phrases of build_LPF_opt, concatenated over the logical blocks; blk_out[k] = block index */
int lz77sss_get_lpf(lz77sss_handle* h, uint64_t* beg, uint64_t* end, uint64_t* src, uint32_t* blk_out, uint64_t cap,
                    uint64_t* count);
/* batched exact LCE_R on the text of the last factorization (lce_sss.hpp:102-177) */
int lz77sss_lce_batch(lz77sss_handle* h, const uint64_t* qi, const uint64_t* qj, uint64_t nq, uint64_t* out);

#ifdef __cplusplus
}
#endif
#endif
