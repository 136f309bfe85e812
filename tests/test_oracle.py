"""CPU tests of the oracle (test infrastructure): the standalone restatement against the committed golden vectors
(produced by the reference's own headers, tests/golden/make_golden.py), against the reference library itself
when it is present (oracle/_ref, build container only), and against plain definitions."""
import glob
import os

import numpy as np
import pytest

from lz77_sss_b200 import synth

GOLDEN = [p for p in sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))
          if not os.path.basename(p).startswith("modes_")]   # modes_*: reference outputs of the other modes (GPU tests)


def _cases(g):
    for k in g.files:
        if k.endswith("_src"):
            tau, p = k[:-4].split("_")
            yield int(tau[3:]), int(p[1:]), k[:-4]


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_restatement_matches_reference_golden(oracle, path):
    g = np.load(path)
    text = g["text"]
    n_cases = 0
    for tau, p, key in _cases(g):
        src, ln, st = oracle.factorize(text, tau, g[key + "_bases"], p_ref=p)
        assert src.size == g[key + "_src"].size
        assert (src == g[key + "_src"]).all() and (ln == g[key + "_len"]).all()
        dec = oracle.decode(src, ln, text.size)
        assert (dec == text).all()
        b, e, s, blk = oracle.lpf_opt(text, tau, p_ref=p)
        lpf = g[key + "_lpf"]
        assert (b == lpf[0]).all() and (e == lpf[1]).all() and (s == lpf[2]).all() and (blk == lpf[3]).all()
        n_cases += 1
    assert n_cases >= 3
    for tau in (4, 64, 512):
        S, nq = oracle.sss(text, tau)
        assert (S == g[f"tau{tau}_S"]).all() and nq == int(g[f"tau{tau}_nq"])
        assert (oracle.sa_s(text, tau, S) == g[f"tau{tau}_SA"]).all()


@pytest.mark.parametrize("tau", [4, 8, 16, 64, 256])
def test_q_anchored_equals_definition(oracle, tau):
    rng = np.random.default_rng(tau)
    parts = [rng.integers(0, 3, size=300).astype(np.uint8), np.zeros(3 * tau + 5, np.uint8),
             np.tile(np.array([1, 2], np.uint8), 2 * tau), np.tile(rng.integers(0, 4, size=7).astype(np.uint8), tau),
             rng.integers(0, 2, size=500).astype(np.uint8), np.tile(rng.integers(0, 4, size=33).astype(np.uint8), tau)]
    text = np.concatenate(parts)
    assert (oracle.q_anchored(text, tau) == oracle.q_bruteforce(text, tau)).all()


@pytest.mark.parametrize("tau", [4, 16, 64, 512])
def test_sync_set_properties(oracle, tau):
    """Consistency (membership depends on T[i..i+2tau) only) and density (gaps > tau only inside low-period runs)."""
    text = synth.random_repetitive(60_000, seed=100 + tau, max_rep_len=3000, max_run_len=2000)
    S, _ = oracle.sss(text, tau)
    assert S.size and (np.diff(S.astype(np.int64)) > 0).all()
    assert S[-1] == text.size - 2 * tau + 1
    members = set(int(x) for x in S)
    ctx = {}
    for i in range(0, text.size - 2 * tau + 1):
        key = text[i:i + 2 * tau].tobytes()
        m = i in members
        if i + 2 * tau < text.size - 1:          # away from the end-of-text terminator
            assert ctx.setdefault(key, m) == m
    q = oracle.q_bruteforce(text, tau)
    for a, b in zip(S[:-1], S[1:]):
        if b - a > tau:
            assert q[int(a) + 1:int(b)].all()


@pytest.mark.parametrize("tau", [4, 64, 512])
def test_suffix_order_and_lce(oracle, tau):
    text = synth.random_repetitive(40_000, seed=7 + tau, max_rep_len=6000, max_run_len=3000)
    S, _ = oracle.sss(text, tau)
    assert (oracle.sa_s(text, tau, S) == oracle.sa_s(text, tau, S, naive=True)).all()
    rng = np.random.default_rng(3)
    qi = rng.integers(0, text.size, size=3000).astype(np.uint64)
    qj = rng.integers(0, text.size, size=3000).astype(np.uint64)
    qj[:200] = qi[:200]
    a, b = oracle.lce_batch(text, tau, qi, qj)
    assert (a == b).all()


def test_restatement_matches_reference_library(oracle):
    """Only where /root/reference was compiled (oracle/_ref): fresh seeds beyond the committed fixtures."""
    if not oracle.ref_available(64):
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    assert oracle.ref(512).refdrv_sizeof_rk_prime() == 80
    for seed, n, tau in ((31, 150_000, 64), (32, 400_000, 512), (33, 90_000, 16)):
        text = synth.random_repetitive(n, seed=seed)
        rs, rl, bases, _ = oracle.ref_factorize(text, tau, p=1)
        os_, ol, _ = oracle.factorize(text, tau, bases, p_ref=1)
        assert rs.size == os_.size and (rs == os_).all() and (rl == ol).all()
        assert (oracle.ref_decode(rs, rl, n, tau) == text).all()


def test_decode_rejects_bad_stream(oracle):
    src = np.array([65, 5, 0], dtype=np.uint64)
    ln = np.array([0, 3, 2], dtype=np.uint64)
    out = np.zeros(6, dtype=np.uint8)
    assert oracle.core().oracle_decode(src, ln, 3, out, 6) != 0


REF_SRC = "/root/reference/include"


def test_reference_own_gtest_passes_against_the_shim():
    """the reference's own tests/test_lz77_sss.cpp (8 approximate, 4 naive-exact, 2 with-samples,
    2 without-samples mode combinations; decode(factorize(T)) == T), compiled where it lies against oracle/shim with the
    reference's debug asserts ON (`make -C oracle reftests`).  Only in the container that holds /root/reference."""
    import subprocess
    if not os.path.isdir(REF_SRC):
        pytest.skip("no /root/reference on this machine")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.run(["make", "-s", "-C", os.path.join(root, "oracle"), "reftests"], check=True)
    env = dict(os.environ, LZ77_SSS_TEST_ITERATIONS="150")
    r = subprocess.run([os.path.join(root, "oracle", "_ref", "test-lz77-sss")], capture_output=True, text=True, env=env, timeout=900)
    assert r.returncode == 0 and "[  PASSED  ] 4 tests." in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def test_reference_debug_asserts_hold_on_the_golden_inputs():
    """The reference build WITHOUT -DNDEBUG (`make -C oracle refdbg`): its own asserts execute on the golden texts in every
    phrase and factorize mode — lce_sss.hpp:167,173 (lce_lr == naive LCE: the shim's S is tau-synchronizing),
    lpf_opt.cpp:91-97 (LPF phrases are real matches), greedy.cpp:81-89,119-125 (every emitted factor matches the text) —
    and the streams equal the -DNDEBUG build's.  A failed assert aborts the subprocess."""
    import subprocess
    import sys
    if not os.path.isdir(REF_SRC):
        pytest.skip("no /root/reference on this machine")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.run(["make", "-s", "-j4", "-C", os.path.join(root, "oracle"), "refdbg"], check=True)
    code = r"""
import ctypes as C, glob, os, sys, numpy as np
sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, "tests"))
import oracle_lib as O
u8 = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS"); u64 = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
runs = 0
for tau in (16, 64, 512):
    L = C.CDLL(os.path.join(O.ORACLE_DIR, "_ref", f"libref_dbg_tau{tau}.so"))
    L.refdrv_factorize_approximate.restype = C.c_int64
    L.refdrv_factorize_approximate.argtypes = [u8, C.c_uint64, C.c_int, C.c_int, C.c_int, u64, u64, C.c_uint64, C.POINTER(C.c_double)]
    for path in sorted(glob.glob(os.path.join(%r, "tests", "golden", "*.npz"))):
        if os.path.basename(path).startswith("modes_"):
            continue
        text = np.ascontiguousarray(np.load(path)["text"])[:120_000]
        for fm in (0, 1, 2):
            for pm in (0, 1, 2, 3):
                for p in (1, 3):
                    cap = text.size + 1
                    s = np.zeros(cap, np.uint64); l = np.zeros(cap, np.uint64)
                    z = L.refdrv_factorize_approximate(text, text.size, pm, fm, p, s, l, cap, None)
                    assert z > 0
                    # same call order on both libraries: their pinned base generators advance in lockstep
                    rs, rl, _, _ = O.ref_factorize(text, tau, p=p, phr_mode=pm, fact_mode=fm)
                    assert z == rs.size and (s[:z] == rs).all() and (l[:z] == rl).all(), (tau, path, fm, pm, p)
                    runs += 1
print("asserts held in", runs, "runs")
""" % (root, root, root)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0 and "asserts held" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]
