"""ctypes bindings of the ORACLE libraries (test infrastructure; never imported by the product).

liboracle.so : oracle/restate/lz77_sss_restate.cpp — standalone CPU restatement (travels to the GPU box)
libref_tau*.so: oracle/ref/ref_driver.cpp — the reference's own headers + oracle/shim (built only
                where /root/reference exists; the prebuilt files travel in oracle/_ref/)
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
DEFAULT_SSS_BASE = 0x9E3779B1
DEFAULT_QMAX = 32
LPF_NAIVE, LPF_LNF_NAIVE, LPF_OPT, LPF_LNF_OPT = 0, 1, 2, 3      # lz77_sss.hpp:47-52
GREEDY_NAIVE, GREEDY, SKIP_PHRASES = 0, 1, 2                      # lz77_sss.hpp:54-58

_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")


class OracleStats(C.Structure):
    _fields_ = [("size_sss", C.c_uint64), ("num_lpf", C.c_uint64), ("len_lpf_phr", C.c_uint64),
                ("num_gaps", C.c_uint64), ("patt_lens", C.c_uint64 * 5), ("roll_threshold", C.c_uint64),
                ("table_entries", C.c_uint64), ("has_runs", C.c_uint64)]


def build_core():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "core"], check=True)


def build_ref():
    if os.path.isdir("/root/reference/include"):
        subprocess.run(["make", "-s", "-j8", "-C", ORACLE_DIR, "ref"], check=True)


_core = None


def core():
    global _core
    if _core is None:
        path = os.path.join(ORACLE_DIR, "_build", "liboracle.so")
        if not os.path.exists(path):
            build_core()
        L = C.CDLL(path)
        L.oracle_sss.restype = C.c_int64
        L.oracle_sss.argtypes = [_u8p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, _u64p, C.c_uint64, C.POINTER(C.c_uint64)]
        L.oracle_q_bruteforce.argtypes = [_u8p, C.c_uint64, C.c_uint64, C.c_uint32, _u8p]
        L.oracle_q_anchored.argtypes = [_u8p, C.c_uint64, C.c_uint64, C.c_uint32, _u8p]
        L.oracle_window_ids.argtypes = [_u8p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, _u32p]
        L.oracle_sa_s.argtypes = [_u8p, C.c_uint64, C.c_uint64, C.c_uint32, _u64p, C.c_uint64, _u32p, C.c_int]
        L.oracle_lcp_s.argtypes = [_u8p, C.c_uint64, _u64p, _u32p, C.c_uint64, _u64p]
        L.oracle_lce_batch.argtypes = [_u8p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, _u64p, _u64p, C.c_uint64, _u64p, _u64p]
        L.oracle_lpf_opt.restype = C.c_int64
        L.oracle_lpf_opt.argtypes = [_u8p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_int, _u64p, _u64p, _u64p, _u32p, C.c_uint64]
        L.oracle_factorize_approximate.restype = C.c_int64
        L.oracle_factorize_approximate.argtypes = [_u8p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, _u64p, C.c_int, C.c_int,
                                                   _u64p, _u64p, C.c_uint64, C.POINTER(OracleStats)]
        L.oracle_decode.restype = C.c_int
        L.oracle_decode.argtypes = [_u64p, _u64p, C.c_uint64, _u8p, C.c_uint64]
        _core = L
    return _core


def _text(t) -> np.ndarray:
    if isinstance(t, (bytes, bytearray)):
        t = np.frombuffer(bytes(t), dtype=np.uint8)
    return np.ascontiguousarray(t, dtype=np.uint8)


def _pad1(a):  # ctypes ndpointer dislikes zero-size arrays on some numpy versions
    return a if a.size else np.zeros(1, dtype=a.dtype)


def sss(text, tau, base=DEFAULT_SSS_BASE, qmax=DEFAULT_QMAX):
    t = _text(text)
    n = t.size
    cap = max(1, n)
    S = np.zeros(cap, dtype=np.uint64)
    nq = C.c_uint64(0)
    s = core().oracle_sss(_pad1(t), n, tau, base, qmax, S, cap, C.byref(nq))
    return S[:s].copy(), int(nq.value)


def q_bruteforce(text, tau, qmax=DEFAULT_QMAX):
    t = _text(text)
    q = np.zeros(max(1, t.size), dtype=np.uint8)
    core().oracle_q_bruteforce(_pad1(t), t.size, tau, qmax, q)
    return q[:max(0, t.size - tau + 1)]


def q_anchored(text, tau, qmax=DEFAULT_QMAX):
    t = _text(text)
    q = np.zeros(max(1, t.size), dtype=np.uint8)
    core().oracle_q_anchored(_pad1(t), t.size, tau, qmax, q)
    return q[:max(0, t.size - tau + 1)]


def window_ids(text, tau, base=DEFAULT_SSS_BASE, qmax=DEFAULT_QMAX):
    t = _text(text)
    ids = np.zeros(t.size + 2, dtype=np.uint32)
    core().oracle_window_ids(_pad1(t), t.size, tau, base, qmax, ids)
    return ids[:max(0, t.size - tau + 2)]


def sa_s(text, tau, S, qmax=DEFAULT_QMAX, naive=False):
    t = _text(text)
    S = np.ascontiguousarray(S, dtype=np.uint64)
    sa = np.zeros(max(1, S.size), dtype=np.uint32)
    core().oracle_sa_s(_pad1(t), t.size, tau, qmax, _pad1(S), S.size, sa, int(naive))
    return sa[:S.size]


def lcp_s(text, S, sa):
    t = _text(text)
    S = np.ascontiguousarray(S, dtype=np.uint64)
    sa = np.ascontiguousarray(sa, dtype=np.uint32)
    out = np.zeros(max(1, S.size), dtype=np.uint64)
    core().oracle_lcp_s(_pad1(t), t.size, _pad1(S), _pad1(sa), S.size, out)
    return out[:S.size]


def lce_batch(text, tau, qi, qj, base=DEFAULT_SSS_BASE, qmax=DEFAULT_QMAX):
    t = _text(text)
    qi = np.ascontiguousarray(qi, dtype=np.uint64)
    qj = np.ascontiguousarray(qj, dtype=np.uint64)
    a = np.zeros(qi.size, dtype=np.uint64)
    b = np.zeros(qi.size, dtype=np.uint64)
    core().oracle_lce_batch(t, t.size, tau, base, qmax, qi, qj, qi.size, a, b)
    return a, b


def lpf_opt(text, tau, p_ref=1, base=DEFAULT_SSS_BASE, qmax=DEFAULT_QMAX):
    t = _text(text)
    cap = t.size + 1
    beg = np.zeros(cap, dtype=np.uint64); end = np.zeros(cap, dtype=np.uint64)
    src = np.zeros(cap, dtype=np.uint64); blk = np.zeros(cap, dtype=np.uint32)
    k = core().oracle_lpf_opt(_pad1(t), t.size, tau, base, qmax, p_ref, beg, end, src, blk, cap)
    return beg[:k].copy(), end[:k].copy(), src[:k].copy(), blk[:k].copy()


def factorize(text, tau, bases, p_ref=1, pos_bits=32, sss_base=DEFAULT_SSS_BASE, qmax=DEFAULT_QMAX):
    """CPU restatement of factorize_approximate<greedy, lpf_opt, tau>. Returns (src, len, stats)."""
    t = _text(text)
    cap = t.size + 1
    src = np.zeros(cap, dtype=np.uint64); ln = np.zeros(cap, dtype=np.uint64)
    st = OracleStats()
    b = np.ascontiguousarray(bases, dtype=np.uint64)
    z = core().oracle_factorize_approximate(_pad1(t), t.size, tau, sss_base, qmax, b, p_ref, pos_bits, src, ln, cap, C.byref(st))
    assert 0 <= z <= cap
    return src[:z].copy(), ln[:z].copy(), st


def decode(src, ln, n):
    src = np.ascontiguousarray(src, dtype=np.uint64)
    ln = np.ascontiguousarray(ln, dtype=np.uint64)
    out = np.zeros(max(1, n), dtype=np.uint8)
    rc = core().oracle_decode(_pad1(src), _pad1(ln), src.size, out, n)
    assert rc == 0, rc
    return out[:n]


_ref = {}


def ref_available(tau=512):
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", f"libref_tau{tau}.so"))


def ref(tau):
    if tau not in _ref:
        path = os.path.join(ORACLE_DIR, "_ref", f"libref_tau{tau}.so")
        L = C.CDLL(path)
        L.refdrv_tau.restype = C.c_uint64
        L.refdrv_sizeof_rk_prime.restype = C.c_uint64
        L.refdrv_set_sss_params.argtypes = [C.c_uint32, C.c_uint32]
        L.refdrv_peek_next_bases.argtypes = [_u64p]
        L.refdrv_factorize_approximate.restype = C.c_int64
        L.refdrv_factorize_approximate.argtypes = [_u8p, C.c_uint64, C.c_int, C.c_int, C.c_int, _u64p, _u64p, C.c_uint64,
                                                   C.POINTER(C.c_double)]
        L.refdrv_factorize_approximate_logged.restype = C.c_int64
        L.refdrv_factorize_approximate_logged.argtypes = [_u8p, C.c_uint64, C.c_int, C.c_int, C.c_int, _u64p, _u64p, C.c_uint64,
                                                          C.POINTER(C.c_double), C.c_char_p, C.c_uint64]
        L.refdrv_decode.argtypes = [_u64p, _u64p, C.c_uint64, _u8p, C.c_uint64]
        _u32p_ = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
        L.refdrv_rks.restype = C.c_int64
        L.refdrv_rks.argtypes = [_u8p, C.c_uint64, C.c_uint64, C.c_int, C.POINTER(C.c_uint64), _u32p_, C.c_uint64, _u64p, _u64p,
                                 C.c_uint64, _u32p_]
        L.refdrv_factorize_exact.restype = C.c_int64
        L.refdrv_factorize_exact.argtypes = [_u8p, C.c_uint64, C.c_int, C.c_int, _u64p, _u64p, C.c_uint64, C.POINTER(C.c_double)]
        L.refdrv_huff_write.argtypes = [_u64p, _u64p, C.c_uint64, C.c_uint64, C.c_char_p]
        L.refdrv_huff_read.restype = C.c_int64
        L.refdrv_huff_read.argtypes = [C.c_char_p, _u64p, _u64p, C.c_uint64, C.POINTER(C.c_uint64)]
        L.refdrv_vbyte_write.argtypes = [_u64p, C.c_uint64, C.c_int, C.c_char_p]
        assert L.refdrv_tau() == tau
        _ref[tau] = L
    return _ref[tau]


def ref_factorize(text, tau, p=1, phr_mode=LPF_OPT, fact_mode=GREEDY, sss_base=DEFAULT_SSS_BASE, qmax=DEFAULT_QMAX):
    """The reference's own factorize_approximate. Returns (src, len, bases_used, seconds)."""
    t = _text(text)
    L = ref(tau)
    L.refdrv_set_sss_params(sss_base, qmax)
    bases = np.zeros(5, dtype=np.uint64)
    L.refdrv_peek_next_bases(bases)
    cap = t.size + 1
    src = np.zeros(cap, dtype=np.uint64); ln = np.zeros(cap, dtype=np.uint64)
    sec = C.c_double(0)
    z = L.refdrv_factorize_approximate(t, t.size, phr_mode, fact_mode, p, src, ln, cap, C.byref(sec))
    assert 0 <= z <= cap, z
    return src[:z].copy(), ln[:z].copy(), bases, sec.value


def ref_require():
    """LZ77SSS_REQUIRE_REF=1: a missing oracle/_ref build is a failure, not a skip (the driver's GPU box gets the prebuilt
    files with the snapshot)."""
    return os.environ.get("LZ77SSS_REQUIRE_REF", "") == "1"


def ref_factorize_logged(text, tau, p=1, phr_mode=LPF_OPT, fact_mode=GREEDY, sss_base=DEFAULT_SSS_BASE, qmax=DEFAULT_QMAX):
    """The reference's factorize_approximate with parameters::log = true.  Returns (src, len, bases, seconds, phases):
    phases = the key=value pairs of the reference's own RESULT line (times in ns; lz77_sss.hpp:298-311,357-365)."""
    t = _text(text)
    L = ref(tau)
    L.refdrv_set_sss_params(sss_base, qmax)
    bases = np.zeros(5, dtype=np.uint64)
    L.refdrv_peek_next_bases(bases)
    cap = t.size + 1
    src = np.zeros(cap, dtype=np.uint64); ln = np.zeros(cap, dtype=np.uint64)
    sec = C.c_double(0)
    buf = C.create_string_buffer(4096)
    z = L.refdrv_factorize_approximate_logged(t, t.size, phr_mode, fact_mode, p, src, ln, cap, C.byref(sec), buf, 4096)
    assert 0 <= z <= cap, z
    phases = {}
    for tok in buf.value.decode().split():
        if "=" in tok:
            k, v = tok.split("=", 1)
            try:
                phases[k] = int(v)
            except ValueError:
                try:
                    phases[k] = float(v)
                except ValueError:
                    phases[k] = v
    return src[:z].copy(), ln[:z].copy(), bases, sec.value, phases


def ref_decode(src, ln, n, tau=512):
    out = np.zeros(max(1, n), dtype=np.uint8)
    ref(tau).refdrv_decode(np.ascontiguousarray(src, dtype=np.uint64), np.ascontiguousarray(ln, dtype=np.uint64), len(src), out, n)
    return out[:n]


def ref_rks(text, s, qpos, qlen, p=1, tau=512):
    """The reference's rabin_karp_substring<31, uint32_t>: (base it drew, its fps table for rate s, substring(qpos, qlen))."""
    t = _text(text)
    qpos = np.ascontiguousarray(qpos, dtype=np.uint64)
    qlen = np.ascontiguousarray(qlen, dtype=np.uint64)
    cap = t.size // s + 3
    fps = np.zeros(cap, dtype=np.uint32)
    out = np.zeros(max(1, qpos.size), dtype=np.uint32)
    base = C.c_uint64(0)
    cnt = ref(tau).refdrv_rks(_pad1(t), t.size, s, p, C.byref(base), fps, cap, _pad1(qpos), _pad1(qlen), qpos.size, out)
    assert 0 < cnt <= cap
    return int(base.value), fps[:cnt].copy(), out[:qpos.size].copy()


def ref_huff_write(src, ln, n, path, tau=512):
    """The reference's huff_writer (misc/huffman.hpp:321-383) on a factor stream."""
    src = np.ascontiguousarray(src, dtype=np.uint64)
    ln = np.ascontiguousarray(ln, dtype=np.uint64)
    assert ref(tau).refdrv_huff_write(_pad1(src), _pad1(ln), src.size, n, str(path).encode()) == 0


def ref_huff_read(path, cap, tau=512):
    """The reference's huff_factor_iterator (misc/huffman.hpp:385-436). Returns (src, len, n)."""
    src = np.zeros(max(1, cap), dtype=np.uint64); ln = np.zeros(max(1, cap), dtype=np.uint64)
    n = C.c_uint64(0)
    z = ref(tau).refdrv_huff_read(str(path).encode(), src, ln, cap, C.byref(n))
    assert 0 <= z <= cap, z
    return src[:z].copy(), ln[:z].copy(), int(n.value)


def ref_vbyte_write(vals, bits, path, tau=512):
    vals = np.ascontiguousarray(vals, dtype=np.uint64)
    assert ref(tau).refdrv_vbyte_write(_pad1(vals), vals.size, bits, str(path).encode()) == 0


NAIVE, WITH_SAMPLES, WITHOUT_SAMPLES = 0, 1, 2      # lz77_sss.hpp:60-64


def ref_factorize_exact(text, tau=512, p=1, transf_mode=WITH_SAMPLES):
    """The reference's own factorize_exact<greedy, lpf_opt, transf_mode> (tau 64 or 512). Returns (src, len, seconds)."""
    t = _text(text)
    cap = t.size + 1
    src = np.zeros(cap, dtype=np.uint64); ln = np.zeros(cap, dtype=np.uint64)
    sec = C.c_double(0)
    z = ref(tau).refdrv_factorize_exact(t, t.size, transf_mode, p, src, ln, cap, C.byref(sec))
    assert 0 <= z <= cap, z
    return src[:z].copy(), ln[:z].copy(), sec.value


def lz77_exact(text):
    """Independent exact LZ77 (suffix array + nearest earlier suffixes in suffix order, KKP-style): (src, len)."""
    t = _text(text)
    cap = t.size + 1
    src = np.zeros(cap, dtype=np.uint64); ln = np.zeros(cap, dtype=np.uint64)
    L = core()
    L.oracle_lz77_exact.restype = C.c_int64
    L.oracle_lz77_exact.argtypes = [_u8p, C.c_uint64, _u64p, _u64p, C.c_uint64]
    z = L.oracle_lz77_exact(_pad1(t), t.size, src, ln, cap)
    assert 0 <= z <= cap
    return src[:z].copy(), ln[:z].copy()
