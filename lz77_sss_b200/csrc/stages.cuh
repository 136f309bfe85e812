// Stage interfaces of the device pipeline (host-callable wrappers around the kernels).
#pragma once
#include <memory>
#include <vector>
#include "common.cuh"

namespace lz77sss {

// The device copy of the text is zero-filled for TEXT_PAD bytes after T[n-1] (the reference's
// callers pad with 4*512 zero bytes, cli/lz77_sss_3_aprx.cpp:66; we never trust the host tail).
constexpr uint64_t TEXT_PAD = 65536;
constexpr uint32_t NONE32 = 0xFFFFFFFFu;

// ---- sss.cu / sss_small.cu --------------------------------------------------------------------
struct sss_scratch {
    dbuf<unsigned long long> status;
    dbuf<unsigned int> ticket;
    dbuf<unsigned long long> counters;
    dbuf<uint32_t> blk;
    dbuf<uint16_t> stage;   // per-tile staging of selected positions (sss_cell.cu)
    scan_scratch scan;
};
uint32_t sss_period_bound(uint32_t tau, uint32_t qmax);
uint32_t sss_anchor_spacing(uint32_t tau, uint32_t P);
void sss_build(const uint8_t* d_text, uint64_t n, uint32_t tau, uint32_t base, uint32_t qmax, void* d_S, int pos_bytes,
               uint64_t s_cap, sss_scratch& sc, cudaStream_t st, uint64_t* h_size, uint64_t* h_numq);
void sss_build_cell(const uint8_t* d_text, uint64_t n, uint32_t tau, uint32_t base, uint32_t qmax, void* d_S, int pos_bytes,
                    uint64_t s_cap, sss_scratch& sc, cudaStream_t st);
void sss_build_small_tau(const uint8_t* d_text, uint64_t n, uint32_t tau, uint32_t base, uint32_t qmax, void* d_S,
                         int pos_bytes, uint64_t s_cap, sss_scratch& sc, cudaStream_t st);

// ---- suffix.cu: order of the suffixes starting at synchronizing positions ------------------------
template <typename pos_t>
struct suffix_state {
    // outputs
    dbuf<uint32_t> SA, ISA;          // |S| each
    dbuf<pos_t> LCP;                 // LCP[r] = lce(S[SA[r-1]], S[SA[r]]) in text characters, LCP[0] = 0
    std::vector<dbuf<uint32_t>> rank_lvl;   // rank_lvl[l][i] = rank of sync suffix i by its first 2^l reduced symbols
    int levels = 0;
    // scratch
    dbuf<uint64_t> k0, k1, enc;
    dbuf<uint32_t> v0, v1, flags, gid, gstart, crank;
    dbuf<uint32_t> counter;
    sort_scratch sort;
    scan_scratch scan;
    // statistics
    int ctx_levels = 0, doubling_rounds = 0;
    uint64_t num_run_syms = 0;
};
template <typename pos_t>
void suffix_sort_sync(const uint8_t* d_text, uint64_t n, uint32_t tau, uint32_t qmax, const pos_t* d_S, uint64_t s,
                      suffix_state<pos_t>& ss, cudaStream_t st);

// ---- lpf.cu: range-min / nearest-smaller structures, LPF candidates, batched LCE -------------------
template <typename pos_t>
struct lce_state {
    dbuf<uint32_t> sa_blkmin;              // sparse table over block minima of SA (ANSV)
    dbuf<pos_t> lcp_blkmin;                // sparse table over block minima of LCP (RMQ)
    uint32_t nblk = 0;
    int sp_levels = 0;
    dbuf<uint32_t> succ_first;             // successor index: first[b] = smallest idx with S[idx] >> shift >= b
    uint32_t succ_shift = 0;
    uint64_t succ_nb = 0;
    // per synchronizing position: the two LPF candidates
    dbuf<uint32_t> psv_idx, nsv_idx;       // sync index of the source or NONE32
    dbuf<pos_t> psv_src, nsv_src;          // its text position S[idx], or ~0 (candidate_sources)
    dbuf<pos_t> psv_lce, nsv_lce, psv_lcel, nsv_lcel;
    dbuf<uint32_t> sa_flip, saf_blkmin;    // ~SA and its block minima (nearest GREATER values for build_LNF_all)
    scan_scratch scan;
};
template <typename pos_t>
void build_lce_structures(const uint8_t* d_text, uint64_t n, uint32_t tau, const pos_t* d_S, uint64_t s,
                          suffix_state<pos_t>& ss, lce_state<pos_t>& ls, cudaStream_t st);
// candidates for build_LPF_opt (lpf_opt.cpp:68-139); blk_b[i] = start of the p_ref block of S[i]
template <typename pos_t>
void lpf_candidates(const uint8_t* d_text, uint64_t n, uint32_t tau, const pos_t* d_S, uint64_t s, uint32_t p_ref,
                    suffix_state<pos_t>& ss, lce_state<pos_t>& ls, cudaStream_t st);
// psv_src / nsv_src from psv_idx / nsv_idx: the host chains read positions, not indices into S
template <typename pos_t>
void candidate_sources(const pos_t* d_S, uint64_t s, lce_state<pos_t>& ls, cudaStream_t st);
// candidates for build_LPF_all (greater = false) / build_LNF_all (greater = true, text reversed), lpf_lnf.cpp:34-249:
// nearest smaller / greater SA_S value on both sides, exact LCE_R, leftward LCE capped at lcap
template <typename pos_t>
void all_candidates(const uint8_t* d_text, uint64_t n, const pos_t* d_S, uint64_t s, bool greater, uint64_t lcap,
                    suffix_state<pos_t>& ss, lce_state<pos_t>& ls, cudaStream_t st);
// uncapped leftward LCE (lce_l.hpp:33-83) of T[..qi] and T[..qj], qi < qj
void left_lce_batch(const uint8_t* d_text, const uint64_t* d_qi, const uint64_t* d_qj, uint64_t nq, uint64_t* d_out,
                    cudaStream_t st);
void reverse_text(const uint8_t* d_in, uint8_t* d_out, uint64_t n, cudaStream_t st);
// batched exact LCE_R(i,j) (lce_sss.hpp:102-177) for arbitrary text positions
template <typename pos_t>
void lce_r_batch(const uint8_t* d_text, uint64_t n, uint32_t tau, const pos_t* d_S, uint64_t s,
                 suffix_state<pos_t>& ss, lce_state<pos_t>& ls, const uint64_t* d_qi, const uint64_t* d_qj,
                 uint64_t nq, uint64_t* d_out, cudaStream_t st);

// ---- gap_hash.cu: bucket ids of the 5 rolling hashes (mod 2^107-1) for ranges of text positions ------
struct gap_hash_params {
    uint64_t patt_len[5];
    uint64_t base[5];
    uint64_t mask;       // table entries - 1
};
// ranges [beg[k], end[k]) of text positions; out[5 * (off[k] + (pos - beg[k])) + x] = fp_x(pos) & mask
// for every window that fits (pos + patt_len[x] <= n), NONE32 otherwise.  The reference only hashes at
// pos + patt_len[x] < n (rolling_hash_index_107.hpp:93,112,123) and keeps the stale fingerprint of the
// window at n - patt_len[x] afterwards; the host emission applies that rule.
void gap_bucket_ids(const uint8_t* d_text, uint64_t n, const gap_hash_params& gp, const uint64_t* d_beg,
                    const uint64_t* d_end, const uint64_t* d_off, uint64_t nranges, uint64_t total, uint32_t* d_out,
                    cudaStream_t st);

// ---- gap_parse.cu: greedy phrase emission on the device ------------------------------------------------
// per-slice buffers a sharded run keeps between its passes (sorted entries, traces, block states)
struct slice_keep {
    dbuf<uint32_t> v0, v1, tkey, bstart, tr_cnt, tr_pos;
    dbuf<uint64_t> tr_src, tr_len;
    dbuf<uint8_t> entry, exit_, marked, ovr, dirty;
    dbuf<uint32_t> notins_prev;            // the bitmap the blocks were last parsed under
    dbuf<unsigned long long> Hc_prev;      // the table in front of the slice at its last parse
    dbuf<uint32_t> skip, rstart;           // sparse slice: positions without entries, bucket starts of the big repair list
    dbuf<uint32_t> prev;                   // bucket chains
    dbuf<uint64_t> rk0, rk1, rkb0, rkb1;   // sparse slice: the repair lists
    uint32_t R = 0, R2 = 0;                // their sizes, and which buffer of each ping-pong pair holds them
    bool rk_in_0 = true, rk2_in_0 = true, sparse = false;
    uint64_t E = 0;                        // entries of the sorted index
    bool sorted = false, parsed = false, vals_in_v0 = true;
};
// sharded run: parse only [beg, end) from the entry state {i, table cursor, phrase index, gap flag} and the table
// d_table_in (last inserted position + 1 per bucket before beg; nullptr = empty), write the table behind end
struct gap_parse_range {
    uint64_t beg, end;
    uint64_t entry[4];
    const uint64_t* d_table_in;
    uint64_t* d_table_out;
    bool retain;               // keep the slices' sorted entries and traces; later calls with the same range resume
    bool index_only;           // only hash + sort the slices and write the speculative table of the range (no parse)
};
struct gap_parse_scratch {
    dbuf<uint32_t> k0, k1, v0, v1, tkey, bstart, big, notins, tr_cnt, tr_pos, changed, J, cnt;
    dbuf<uint64_t> tr_src, tr_len, ml, fsrc, flen;
    dbuf<unsigned long long> Hc;
    dbuf<uint8_t> entry, exit_, marked, slice_exit, nbtab, ovr, dirty;
    dbuf<uint8_t> marked_prev;             // the chain of the round before (a block that joins the chain is parsed again)
    dbuf<uint32_t> notins_prev;
    dbuf<unsigned long long> Hc_prev;
    dbuf<uint32_t> skip, tile_off;         // sparse slices: positions without entries, compaction offsets per tile
    dbuf<uint32_t> prev;                   // bucket chains: per entry id, the entry in front of it in its bucket
    dbuf<uint64_t> rk0, rk1;               // sparse slices: repair list (bucket << 32 | entry id), ping-pong for its sort
    dbuf<uint64_t> rkb0, rkb1;             // the small second repair list
    dbuf<uint32_t> rstart;                 // bucket starts of the big repair list
    dbuf<uint64_t> newent;                 // the entries a repair step added (for the change tracking)
    dbuf<uint32_t> req;                    // repair requests of a parse round
    uint64_t repair_entries = 0;           // entries added to repair lists over all slices
    sort_scratch sort;
    scan_scratch scan;
    uint64_t rounds = 0, slices = 0, entries_sorted = 0, sparse_slices = 0, sparse_fallbacks = 0;
    std::vector<std::unique_ptr<slice_keep>> keep;
    // The index (hash + sort) of slice s+1 is built on a second stream while slice s is parsed: the parse rounds are
    // bound by the latency of dependent loads and leave the memory system idle, the sort is bandwidth work.
    struct index_set { dbuf<uint32_t> v0, v1, tkey, bstart, skip, prev; } next;
    sort_scratch sort_b;                   // scratch of the builds on the second stream
    scan_scratch scan_b;
    struct side_stream {                   // (movable, so that the scratch can be reset by assignment)
        cudaStream_t st = nullptr;
        cudaEvent_t ev_parse = nullptr, ev_build = nullptr;
        side_stream() = default;
        side_stream(side_stream&& o) noexcept : st(o.st), ev_parse(o.ev_parse), ev_build(o.ev_build) { o.st = nullptr; o.ev_parse = o.ev_build = nullptr; }
        side_stream& operator=(side_stream&& o) noexcept {
            if (this != &o) { std::swap(st, o.st); std::swap(ev_parse, o.ev_parse); std::swap(ev_build, o.ev_build); }
            return *this;
        }
        ~side_stream() {
            if (st) cudaStreamDestroy(st);
            if (ev_parse) cudaEventDestroy(ev_parse);
            if (ev_build) cudaEventDestroy(ev_build);
        }
    } side;
    uint64_t prefetched_slices = 0;
};
// h_ml: three planes (beg, end, src) of the merged phrase list incl. the sentinel {n, n+1, 0} (m entries).
// Parses [0, n_cut); appends the factors to out_src/out_len; handover = {i, cur_pos, phrase index, gap flag}.
template <typename pos_t>
void gap_parse_run(const uint8_t* d_text, uint64_t n, uint32_t tau, const pos_t* d_S, uint64_t s, suffix_state<pos_t>& ss,
                   lce_state<pos_t>& ls, const gap_hash_params& gp, uint64_t table_entries, const uint64_t* h_ml,
                   uint64_t m, uint64_t n_cut, uint64_t slice_len, uint32_t block_len, gap_parse_scratch& sc,
                   cudaStream_t st, pinned_vec<uint64_t>& out_src, pinned_vec<uint64_t>& out_len, uint64_t handover[4],
                   bool naive = false, const gap_parse_range* range = nullptr);
// table entries (last inserted position + 1, 0 = none) of the given buckets after gap_parse_run
void gap_parse_query(gap_parse_scratch& sc, const uint32_t* h_buckets, uint32_t cnt, uint64_t* h_out, cudaStream_t st);

// ---- exact.cu: exact LZ77 (factorize_exact) from the full suffix array -------------------------------------------------------
struct exact_scratch {
    dbuf<uint64_t> k0, k1, fsrc, flen, sect;
    dbuf<uint32_t> v0, v1, rank, flag, acc, counter, tab, tr_cnt, cnt, J;
    dbuf<uint8_t> blk, marked, dirty;
    sort_scratch sort;
    scan_scratch scan;
    int rounds = 0, parse_rounds = 0;      // prefix-doubling rounds, parse rounds of the last run
};
// sect: section starts, ascending, sect.front() == 0, sect.back() == n (factors are cut at section ends).
// Needs the LCE structures of a preceding approximation on the same text (S, ss, ls).
template <typename pos_t>
void exact_run(const uint8_t* d_text, uint64_t n, uint32_t tau, const pos_t* d_S, uint64_t s, suffix_state<pos_t>& ss,
               lce_state<pos_t>& ls, const std::vector<uint64_t>& sect, exact_scratch& sc, cudaStream_t st,
               pinned_vec<uint64_t>& out_src, pinned_vec<uint64_t>& out_len);

// ---- rks.cu: rabin_karp_substring<31> (prefix fingerprints modulo 2^31 - 1, one per 16 bytes) ---------------------------
struct rks_state {
    dbuf<uint32_t> pf;           // pf[g] = fingerprint of T[0 .. 16 (g + 1))
    dbuf<unsigned long long> status;   // per tile: (1 << 32 | aggregate) or (2 << 32 | inclusive value)
    uint64_t n = 0, ngran = 0;
    uint32_t base = 0;
};
void rks_build(const uint8_t* d_text, uint64_t n, uint32_t base, rks_state& R, cudaStream_t st);
// the same table (pf[g] for g < ceil(n / 16)) into the caller's d_out, look-back states in a static device array: no allocation.
// n <= RKS_NOALLOC_MAX_BYTES; d_text and d_out 16-byte aligned, zero bytes behind the text up to the next multiple of 16.
constexpr uint64_t RKS_NOALLOC_MAX_BYTES = 1ull << 30;
void rks_build_noalloc(const uint8_t* d_text, uint64_t n, uint32_t base, uint32_t* d_out, cudaStream_t st);
void rks_substring_batch(const uint8_t* d_text, const rks_state& R, const uint64_t* d_pos, const uint64_t* d_len, uint64_t nq,
                         uint32_t* d_out, cudaStream_t st);
void rks_fps(const uint8_t* d_text, const rks_state& R, uint64_t s, uint64_t cnt, uint32_t* d_out, cudaStream_t st);

// ---- decode.cu ------------------------------------------------------------------------------------
// parallel LZ77 decode: factors (src,len) with len == 0 meaning literal src; returns 0 or an error code
// returns 0, -1 (lengths do not add up to n), -2 (a source is not earlier text), -4 (no convergence)
struct decode_scratch {
    dbuf<uint8_t> start, ptr, tile_done;   // start / ptr hold 32- or 64-bit words depending on n
    dbuf<uint64_t> tsum;
    dbuf<unsigned long long> cnt;
};
int decode_factors(const uint64_t* d_src, const uint64_t* d_len, uint64_t z, uint8_t* d_out, uint64_t n,
                   cudaStream_t st, int* rounds, decode_scratch& sc);
bool device_equal(const uint8_t* d_a, const uint8_t* d_b, uint64_t n, cudaStream_t st);

}  // namespace lz77sss
