"""ctypes binding of the C ABI in include/lz77sss_b200.h (liblz77sss_b200.so).

This is harness plumbing for tests/ and bench.py: the product is the shared library and the C++
header lz77_sss_b200/include/lz77_sss/lz77_sss.hpp.  Nothing here computes; there is no fallback:
if the library is missing or no CUDA device is present the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liblz77sss_b200.so")

PHR_LPF_NAIVE, PHR_LPF_LNF_NAIVE, PHR_LPF_OPT, PHR_LPF_LNF_OPT = 0, 1, 2, 3
FACT_GREEDY_NAIVE, FACT_GREEDY, FACT_SKIP_PHRASES = 0, 1, 2
DEFAULT_SSS_BASE = 0x9E3779B1
DEFAULT_GAP_BASES = (65599, 131071, 262147, 524309, 1048573)

EXPORTS = [
    "lz77sss_last_error", "lz77sss_version", "lz77sss_device_count", "lz77sss_create", "lz77sss_destroy",
    "lz77sss_factorize_approximate", "lz77sss_factorize_exact", "lz77sss_device_text_alloc", "lz77sss_device_text_upload",
    "lz77sss_device_text_free", "lz77sss_factorize_approximate_device", "lz77sss_decode", "lz77sss_build_sss",
    "lz77sss_time_sss_kernel", "lz77sss_get_stats", "lz77sss_get_sss", "lz77sss_get_sa_s", "lz77sss_get_lcp_s",
    "lz77sss_get_lpf", "lz77sss_lce_batch", "lz77sss_sink_emit",
    "lz77sss_shard_prepare", "lz77sss_shard_table_entries", "lz77sss_shard_entry_guess", "lz77sss_shard_split", "lz77sss_shard_parse", "lz77sss_shard_factors_view",
    "lz77sss_shard_factors", "lz77sss_shard_finish", "lz77sss_shard_index",
    "lz77sss_rks_create", "lz77sss_rks_destroy", "lz77sss_rks_table", "lz77sss_rks_table_noalloc", "lz77sss_rks_fps", "lz77sss_rks_substring",
    "lz77sss_rks_time_build", "lz77sss_time_decode", "lz77sss_kernel_profile", "lz77sss_kernel_profile_get",
]


class Config(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("tau", C.c_uint32), ("num_threads", C.c_uint32), ("pos_bits", C.c_uint32),
                ("phr_mode", C.c_uint32), ("fact_mode", C.c_uint32), ("sss_base", C.c_uint32), ("sss_qmax", C.c_uint32),
                ("gap_bases", C.c_uint64 * 5), ("device", C.c_int32), ("log", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("n", "size_sss", "num_q", "num_factors", "num_lpf", "len_lpf_phr", "num_gaps")] + \
               [("patt_lens", C.c_uint64 * 5)] + \
               [(k, C.c_uint64) for k in ("roll_threshold", "table_entries", "ctx_levels", "doubling_rounds",
                                          "gap_positions_hashed", "t_h2d_ns", "t_sss_ns", "t_sort_ns", "t_lce_ns",
                                          "t_lpf_ns", "t_d2h_ns", "t_chain_ns", "t_gap_hash_ns", "t_emit_ns", "t_total_ns",
                                          "kernel_launches", "h2d_bytes", "d2h_bytes", "parse_rounds", "parse_slices", "sparse_slices", "sparse_fallbacks", "repair_entries",
                                          "t_exact_ns", "exact_doubling_rounds", "exact_parse_rounds", "prefetched_slices")]

    def as_dict(self):
        d = {}
        for k, _ in self._fields_:
            v = getattr(self, k)
            d[k] = list(v) if k == "patt_lens" else int(v)
        return d


class Sink(C.Structure):
    _fields_ = [("src", C.POINTER(C.c_uint64)), ("len", C.POINTER(C.c_uint64)), ("cap", C.c_uint64), ("count", C.c_uint64)]


class Lz77SssError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"lz77sss error {code}: {msg}")
        self.code = code


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                                    "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.lz77sss_last_error.restype = C.c_char_p
        L.lz77sss_version.restype = C.c_char_p
        L.lz77sss_create.argtypes = [C.POINTER(Config), C.POINTER(C.c_void_p)]
        L.lz77sss_destroy.argtypes = [C.c_void_p]
        L.lz77sss_destroy.restype = None
        L.lz77sss_factorize_approximate.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
        L.lz77sss_factorize_exact.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
        L.lz77sss_device_text_alloc.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.c_void_p)]
        L.lz77sss_device_text_upload.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
        L.lz77sss_device_text_free.argtypes = [C.c_void_p, C.c_void_p]
        L.lz77sss_factorize_approximate_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
        L.lz77sss_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64]
        L.lz77sss_build_sss.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.lz77sss_time_sss_kernel.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
        L.lz77sss_get_stats.argtypes = [C.c_void_p, C.POINTER(Stats)]
        L.lz77sss_get_sss.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
        L.lz77sss_get_sa_s.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
        L.lz77sss_get_lcp_s.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
        L.lz77sss_get_lpf.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
        L.lz77sss_lce_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
        L.lz77sss_shard_prepare.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]
        L.lz77sss_shard_table_entries.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
        L.lz77sss_shard_entry_guess.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
        L.lz77sss_shard_split.argtypes = [C.c_void_p, C.c_uint32, C.c_uint64, C.POINTER(C.c_uint64)]
        L.lz77sss_shard_parse.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64), C.c_void_p, C.c_void_p,
                                          C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.lz77sss_shard_index.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p]
        L.lz77sss_shard_factors.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
        L.lz77sss_shard_factors_view.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
        L.lz77sss_shard_finish.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.c_void_p, C.c_void_p]
        L.lz77sss_rks_create.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(C.c_void_p)]
        L.lz77sss_rks_destroy.argtypes = [C.c_void_p]
        L.lz77sss_rks_destroy.restype = None
        L.lz77sss_rks_table.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
        L.lz77sss_rks_table_noalloc.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p]
        L.lz77sss_rks_fps.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
        L.lz77sss_rks_substring.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
        L.lz77sss_rks_time_build.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
        L.lz77sss_time_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.POINTER(C.c_double),
                                          C.POINTER(C.c_int), C.POINTER(C.c_uint64), C.c_void_p, C.POINTER(C.c_int)]
        L.lz77sss_kernel_profile.argtypes = [C.c_void_p, C.c_int]
        L.lz77sss_kernel_profile_get.argtypes = [C.c_void_p, C.c_uint32, C.c_char_p, C.c_uint32, C.POINTER(C.c_double),
                                                 C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise Lz77SssError(rc, lib().lz77sss_last_error().decode())


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


def _u8(text) -> np.ndarray:
    if isinstance(text, (bytes, bytearray, memoryview)):
        text = np.frombuffer(bytes(text), dtype=np.uint8)
    return np.ascontiguousarray(text, dtype=np.uint8)


class Factorizer:
    """One lz77sss_handle: factorize_approximate<greedy, lpf_opt, tau> with every parameter explicit."""

    def __init__(self, tau=512, num_threads=1, gap_bases=DEFAULT_GAP_BASES, pos_bits=0, sss_base=DEFAULT_SSS_BASE,
                 sss_qmax=32, device=0, phr_mode=PHR_LPF_OPT, fact_mode=FACT_GREEDY, log=False):
        cfg = Config()
        cfg.struct_size = C.sizeof(Config)
        cfg.tau, cfg.num_threads, cfg.pos_bits = tau, num_threads, pos_bits
        cfg.phr_mode, cfg.fact_mode = phr_mode, fact_mode
        cfg.sss_base, cfg.sss_qmax = sss_base, sss_qmax
        for i in range(5):
            cfg.gap_bases[i] = int(gap_bases[i])
        cfg.device, cfg.log = device, int(log)
        self._h = C.c_void_p()
        _check(lib().lz77sss_create(C.byref(cfg), C.byref(self._h)))
        self.tau = tau

    def close(self):
        if self._h:
            lib().lz77sss_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # --- the public path -------------------------------------------------------------------------
    def _sink(self, cap):
        src = np.empty(cap, dtype=np.uint64)
        ln = np.empty(cap, dtype=np.uint64)
        s = Sink(src.ctypes.data_as(C.POINTER(C.c_uint64)), ln.ctypes.data_as(C.POINTER(C.c_uint64)), cap, 0)
        return src, ln, s

    def factorize(self, text, cap=None):
        """Host text -> (src, len) arrays (len == 0: literal byte in src)."""
        t = _u8(text)
        cap = cap if cap is not None else t.size + 1
        src, ln, s = self._sink(cap)
        emit = C.cast(lib().lz77sss_sink_emit, C.c_void_p)
        _check(lib().lz77sss_factorize_approximate(self._h, _ptr(t), t.size, emit, C.byref(s)))
        z = int(s.count)
        if z > cap:
            raise RuntimeError(f"factor sink too small: {z} > {cap}")
        return src[:z], ln[:z]

    def factorize_exact(self, text, cap=None):
        """Exact LZ77 (per section when num_threads > 1): host text -> (src, len)."""
        t = _u8(text)
        cap = cap if cap is not None else t.size + 1
        src, ln, s = self._sink(cap)
        emit = C.cast(lib().lz77sss_sink_emit, C.c_void_p)
        _check(lib().lz77sss_factorize_exact(self._h, _ptr(t), t.size, emit, C.byref(s)))
        z = int(s.count)
        if z > cap:
            raise RuntimeError(f"factor sink too small: {z} > {cap}")
        return src[:z], ln[:z]

    def factorize_into(self, text_ptr, n, sink, src_keepalive=None):
        """Zero-copy variant for bench.py: text_ptr is a host address (e.g. pinned), sink a prepared Sink."""
        emit = C.cast(lib().lz77sss_sink_emit, C.c_void_p)
        sink.count = 0
        _check(lib().lz77sss_factorize_approximate(self._h, C.c_void_p(text_ptr), n, emit, C.byref(sink)))
        return int(sink.count)

    def device_text(self, text):
        t = _u8(text)
        d = C.c_void_p()
        _check(lib().lz77sss_device_text_alloc(self._h, t.size, C.byref(d)))
        _check(lib().lz77sss_device_text_upload(self._h, d, _ptr(t), t.size))
        return d

    def free_device_text(self, d):
        _check(lib().lz77sss_device_text_free(self._h, d))

    def factorize_device_into(self, d_text, host_ptr, n, sink):
        emit = C.cast(lib().lz77sss_sink_emit, C.c_void_p)
        sink.count = 0
        _check(lib().lz77sss_factorize_approximate_device(self._h, d_text, C.c_void_p(host_ptr), n, emit, C.byref(sink)))
        return int(sink.count)

    def decode(self, src, ln, n):
        src = np.ascontiguousarray(src, dtype=np.uint64)
        ln = np.ascontiguousarray(ln, dtype=np.uint64)
        out = np.empty(n, dtype=np.uint8)
        _check(lib().lz77sss_decode(self._h, _ptr(src), _ptr(ln), src.size, _ptr(out), n))
        return out

    def time_decode(self, src, ln, n, iters=5, d_expect=None):
        """Device-resident decode timing: (mean ms, pointer-jumping rounds, kernel launches per decode, output == d_expect)."""
        src = np.ascontiguousarray(src, dtype=np.uint64)
        ln = np.ascontiguousarray(ln, dtype=np.uint64)
        ms, rounds, launches, eq = C.c_double(0), C.c_int(0), C.c_uint64(0), C.c_int(-1)
        _check(lib().lz77sss_time_decode(self._h, _ptr(src), _ptr(ln), src.size, n, iters, C.byref(ms), C.byref(rounds),
                                         C.byref(launches), d_expect, C.byref(eq)))
        return ms.value, rounds.value, int(launches.value), (None if d_expect is None else bool(eq.value))

    # --- stage-level -----------------------------------------------------------------------------
    def build_sss(self, text):
        t = _u8(text)
        cap = max(1, t.size)
        S = np.empty(cap, dtype=np.uint64)
        cnt, nq = C.c_uint64(0), C.c_uint64(0)
        _check(lib().lz77sss_build_sss(self._h, _ptr(t), t.size, _ptr(S), cap, C.byref(cnt), C.byref(nq)))
        return S[:cnt.value].copy(), int(nq.value)

    def time_sss_kernel(self, d_text, n, iters=5):
        ms, cnt = C.c_double(0), C.c_uint64(0)
        _check(lib().lz77sss_time_sss_kernel(self._h, d_text, n, iters, C.byref(ms), C.byref(cnt)))
        return ms.value, int(cnt.value)

    def kernel_profile(self, enable):
        _check(lib().lz77sss_kernel_profile(self._h, int(enable)))

    def kernel_profile_rows(self):
        rows = []
        idx = 0
        while True:
            name = C.create_string_buffer(128)
            ms, launches, nbytes = C.c_double(0), C.c_uint64(0), C.c_uint64(0)
            if lib().lz77sss_kernel_profile_get(self._h, idx, name, 128, C.byref(ms), C.byref(launches), C.byref(nbytes)) != 0:
                break
            rows.append({"kernel": name.value.decode(), "ms": ms.value, "launches": int(launches.value), "bytes": int(nbytes.value)})
            idx += 1
        return rows

    def stats(self):
        st = Stats()
        _check(lib().lz77sss_get_stats(self._h, C.byref(st)))
        return st.as_dict()

    def _get(self, fn, dtype):
        cnt = C.c_uint64(0)
        _check(fn(self._h, None, 0, C.byref(cnt)))
        out = np.empty(max(1, cnt.value), dtype=dtype)
        _check(fn(self._h, _ptr(out), cnt.value, C.byref(cnt)))
        return out[:cnt.value]

    def sss(self):
        return self._get(lib().lz77sss_get_sss, np.uint64)

    def sa_s(self):
        return self._get(lib().lz77sss_get_sa_s, np.uint32)

    def lcp_s(self):
        return self._get(lib().lz77sss_get_lcp_s, np.uint64)

    def lpf(self):
        cnt = C.c_uint64(0)
        _check(lib().lz77sss_get_lpf(self._h, None, None, None, None, 0, C.byref(cnt)))
        k = max(1, cnt.value)
        beg = np.empty(k, dtype=np.uint64); end = np.empty(k, dtype=np.uint64)
        src = np.empty(k, dtype=np.uint64); blk = np.empty(k, dtype=np.uint32)
        _check(lib().lz77sss_get_lpf(self._h, _ptr(beg), _ptr(end), _ptr(src), _ptr(blk), cnt.value, C.byref(cnt)))
        k = cnt.value
        return beg[:k], end[:k], src[:k], blk[:k]

    def lce_batch(self, qi, qj):
        qi = np.ascontiguousarray(qi, dtype=np.uint64)
        qj = np.ascontiguousarray(qj, dtype=np.uint64)
        out = np.empty(max(1, qi.size), dtype=np.uint64)
        _check(lib().lz77sss_lce_batch(self._h, _ptr(qi), _ptr(qj), qi.size, _ptr(out)))
        return out[:qi.size]

    # --- one text sharded over several GPUs (include/lz77sss_b200.h, lz77sss_shard_*) -------------
    def shard_prepare(self, d_text_ptr, host_ptr, n):
        _check(lib().lz77sss_shard_prepare(self._h, C.c_void_p(d_text_ptr), C.c_void_p(host_ptr), n))

    def shard_table_entries(self):
        e = C.c_uint64(0)
        _check(lib().lz77sss_shard_table_entries(self._h, C.byref(e)))
        return int(e.value)

    def shard_split(self, parts, align=2048):
        """Ranges of about equal work (gap bytes weigh 16x phrase bytes): [(beg, end)] * parts."""
        b = (C.c_uint64 * (parts + 1))()
        _check(lib().lz77sss_shard_split(self._h, parts, align, b))
        return [(int(b[r]), int(b[r + 1])) for r in range(parts)]

    def shard_entry_guess(self, pos):
        st = (C.c_uint64 * 4)()
        _check(lib().lz77sss_shard_entry_guess(self._h, pos, st))
        return tuple(int(x) for x in st)

    def shard_parse(self, beg, end, entry, table_in_ptr, table_out_ptr):
        """Parses [beg, end); returns (exit_state, number of factors). Table pointers are device addresses (or 0)."""
        e = (C.c_uint64 * 4)(*[int(x) for x in entry])
        x = (C.c_uint64 * 4)()
        z = C.c_uint64(0)
        _check(lib().lz77sss_shard_parse(self._h, beg, end, e, C.c_void_p(table_in_ptr or None), C.c_void_p(table_out_ptr or None),
                                         x, C.byref(z)))
        return tuple(int(v) for v in x), int(z.value)

    def shard_index(self, beg, end, table_guess_ptr):
        _check(lib().lz77sss_shard_index(self._h, beg, end, C.c_void_p(table_guess_ptr)))

    def shard_factors(self):
        cnt = C.c_uint64(0)
        _check(lib().lz77sss_shard_factors(self._h, None, None, 0, C.byref(cnt)))
        z = int(cnt.value)
        src = np.empty(max(1, z), dtype=np.uint64)
        ln = np.empty(max(1, z), dtype=np.uint64)
        _check(lib().lz77sss_shard_factors(self._h, _ptr(src), _ptr(ln), z, C.byref(cnt)))
        return src[:z], ln[:z]

    def shard_factors_view(self):
        """The factors of the last shard_parse as views of the library's pinned arrays (no copy; valid until the next call)."""
        ps, pl, cnt = C.c_void_p(), C.c_void_p(), C.c_uint64(0)
        _check(lib().lz77sss_shard_factors_view(self._h, C.byref(ps), C.byref(pl), C.byref(cnt)))
        z = int(cnt.value)
        if z == 0:
            return np.empty(0, np.uint64), np.empty(0, np.uint64)
        return (np.ctypeslib.as_array((C.c_uint64 * z).from_address(ps.value)),
                np.ctypeslib.as_array((C.c_uint64 * z).from_address(pl.value)))

    def shard_finish(self, state, cap=4096):
        src, ln, s = self._sink(cap)
        st = (C.c_uint64 * 4)(*[int(x) for x in state])
        emit = C.cast(lib().lz77sss_sink_emit, C.c_void_p)
        _check(lib().lz77sss_shard_finish(self._h, st, emit, C.byref(s)))
        z = int(s.count)
        return src[:z].copy(), ln[:z].copy()


class RabinKarpSubstring:
    """rabin_karp_substring<31, pos_t> (rabin_karp_substring.hpp:35-240) built on the device of a Factorizer's handle."""

    def __init__(self, factorizer, text, base):
        self._F = factorizer
        t = _u8(text)
        self.n = t.size
        self._r = C.c_void_p()
        _check(lib().lz77sss_rks_create(factorizer._h, _ptr(t), t.size, int(base), C.byref(self._r)))

    def close(self):
        if self._r:
            lib().lz77sss_rks_destroy(self._r)
            self._r = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def table(self):
        cnt = C.c_uint64(0)
        _check(lib().lz77sss_rks_table(self._r, None, 0, C.byref(cnt)))
        out = np.empty(max(1, cnt.value), dtype=np.uint32)
        _check(lib().lz77sss_rks_table(self._r, _ptr(out), cnt.value, C.byref(cnt)))
        return out[:cnt.value]

    def fps(self, s):
        cnt = C.c_uint64(0)
        _check(lib().lz77sss_rks_fps(self._r, s, None, 0, C.byref(cnt)))
        out = np.empty(max(1, cnt.value), dtype=np.uint32)
        _check(lib().lz77sss_rks_fps(self._r, s, _ptr(out), cnt.value, C.byref(cnt)))
        return out[:cnt.value]

    def substring(self, pos, length):
        pos = np.ascontiguousarray(pos, dtype=np.uint64)
        length = np.ascontiguousarray(length, dtype=np.uint64)
        out = np.empty(max(1, pos.size), dtype=np.uint32)
        _check(lib().lz77sss_rks_substring(self._r, _ptr(pos), _ptr(length), pos.size, _ptr(out)))
        return out[:pos.size]

    def time_build(self, iters=10):
        ms, launches = C.c_double(0), C.c_uint64(0)
        _check(lib().lz77sss_rks_time_build(self._r, iters, C.byref(ms), C.byref(launches)))
        return ms.value, int(launches.value)
