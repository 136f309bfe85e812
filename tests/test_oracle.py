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
