"""GPU parity tests (run with -m gpu on a B200): the CUDA path, called through the C ABI, against
 (1) the committed golden vectors produced by the reference's own code,
 (2) the oracle on fresh seeded inputs (every stage: S, SA_S, LCP_S, LPF phrases, factor stream, LCE batch),
 (3) size-independent properties at bench-scale sizes (decode round trip, factors well formed)."""
import glob
import os
import subprocess

import numpy as np
import pytest

from lz77_sss_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_ALL_NPZ = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))
GOLDEN = [p for p in _ALL_NPZ if not os.path.basename(p).startswith("modes_")]
GOLDEN_MODES = [p for p in _ALL_NPZ if os.path.basename(p).startswith("modes_")]
TAUS = [4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096]


def _eq(a, b):
    return a.size == b.size and (a == b).all()


def _need_ref(oracle, tau):
    """oracle/_ref (the reference's own headers, built in the CPU container) travels with the snapshot; without it the
    test is skipped, or fails when LZ77SSS_REQUIRE_REF=1."""
    if oracle.ref_available(tau):
        return
    if oracle.ref_require():
        pytest.fail(f"oracle/_ref/libref_tau{tau}.so is missing and LZ77SSS_REQUIRE_REF=1")
    pytest.skip("oracle/_ref was not prebuilt")


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_golden_vectors(capi, path):
    g = np.load(path)
    text = g["text"]
    for k in [k for k in g.files if k.endswith("_src")]:
        tau, p = k[:-4].split("_")
        tau, p, key = int(tau[3:]), int(p[1:]), k[:-4]
        with capi.Factorizer(tau=tau, num_threads=p, gap_bases=g[key + "_bases"]) as F:
            src, ln = F.factorize(text)
            assert _eq(src, g[key + "_src"]) and _eq(ln, g[key + "_len"]), (key, "factor stream")
            assert _eq(F.sss(), g[f"tau{tau}_S"]), (key, "S")
            assert _eq(F.sa_s(), g[f"tau{tau}_SA"]), (key, "SA_S")
            b, e, s, blk = F.lpf()
            lpf = g[key + "_lpf"]
            assert _eq(b, lpf[0]) and _eq(e, lpf[1]) and _eq(s, lpf[2]) and _eq(blk.astype(np.uint64), lpf[3]), (key, "LPF")
            assert (F.decode(src, ln, text.size) == text).all(), (key, "decode")


@pytest.mark.parametrize("path", GOLDEN_MODES, ids=[os.path.basename(p)[:-4] for p in GOLDEN_MODES])
def test_golden_vectors_all_modes(capi, path):
    """Every factorize_mode x phrase_mode (lz77_sss.hpp:47-58) against the reference's own output."""
    g = np.load(path)
    text = g["text"]
    keys = [k[:-4] for k in g.files if k.endswith("_src")]
    assert keys
    for key in keys:
        tau, p, fm, pm = key.split("_")
        tau, p, fm, pm = int(tau[3:]), int(p[1:]), int(fm[2:]), int(pm[2:])
        with capi.Factorizer(tau=tau, num_threads=p, gap_bases=g[key + "_bases"], phr_mode=pm, fact_mode=fm) as F:
            src, ln = F.factorize(text)
            assert _eq(src, g[key + "_src"]) and _eq(ln, g[key + "_len"]), (key, "factor stream")
            if fm != capi.FACT_SKIP_PHRASES:
                assert (F.decode(src, ln, text.size) == text).all(), (key, "decode")


@pytest.mark.parametrize("fm,pm", [(fm, pm) for fm in (0, 1, 2) for pm in (0, 1, 2, 3)])
def test_modes_against_reference_build(capi, oracle, fm, pm):
    """Fresh inputs against oracle/_ref (the reference's own headers), where that build travelled to this box."""
    tau = 64
    _need_ref(oracle, tau)
    for name, text in _texts().items():
        if name in ("zeros", "period3"):
            continue      # no LPF phrase or no gap: pattern lengths undefined in the reference
        for p in (1, 2):
            rs, rl, bases, _ = oracle.ref_factorize(text, tau, p=p, phr_mode=pm, fact_mode=fm)
            with capi.Factorizer(tau=tau, num_threads=p, gap_bases=bases, phr_mode=pm, fact_mode=fm) as F:
                src, ln = F.factorize(text)
                assert _eq(src, rs) and _eq(ln, rl), (name, p, fm, pm)


def _texts():
    rng = np.random.default_rng(5)
    return {
        "rep": synth.random_repetitive(400_000, seed=41),
        "rep_runs": synth.random_repetitive(300_000, seed=42, rep=0.5, run=0.9),
        "dna": synth.dna_collection(500_000, seed=43, chrom_len=60_000),
        "markov": synth.markov_text(200_000, seed=44),
        "random": rng.integers(0, 256, size=80_000).astype(np.uint8),
        "binary": rng.integers(0, 2, size=60_000).astype(np.uint8),
        "zeros": np.zeros(50_000, dtype=np.uint8),
        "period3": np.tile(np.frombuffer(b"abc", dtype=np.uint8), 20_000),
        "period100": np.tile(rng.integers(0, 4, size=100).astype(np.uint8), 900),
        "xx": np.tile(synth.markov_text(70_000, seed=45), 2),
    }


@pytest.mark.parametrize("tau", TAUS)
def test_sync_set_all_taus(capi, oracle, tau):
    for name, text in _texts().items():
        with capi.Factorizer(tau=tau) as F:
            S, nq = F.build_sss(text)
        So, nqo = oracle.sss(text, tau)
        assert _eq(S, So), (name, tau)
        assert nq == nqo, (name, tau)


@pytest.mark.parametrize("n", [0, 1, 2, 6, 7, 8, 15, 100, 1022, 1023, 1024, 1025, 2047, 8191, 8192, 8193, 16385])
def test_sync_set_tiny_and_ragged(capi, oracle, n):
    text = synth.random_repetitive(max(n, 1), seed=n + 1)[:n]
    for tau in (4, 16, 512, 4096):
        with capi.Factorizer(tau=tau) as F:
            S, nq = F.build_sss(text)
        So, nqo = oracle.sss(text, tau)
        assert _eq(S, So) and nq == nqo, (n, tau)


@pytest.mark.parametrize("tau,p", [(4, 1), (8, 2), (16, 1), (32, 3), (64, 1), (128, 4), (256, 1), (512, 1), (512, 5),
                                   (1024, 1), (2048, 2), (4096, 1)])
def test_every_stage_against_oracle(capi, oracle, tau, p):
    bases = (300 + tau, 70_001, 123_457, 999_983, 1_048_575)
    for name, text in _texts().items():
        n = text.size
        osrc, olen, ost = oracle.factorize(text, tau, bases, p_ref=p)
        with capi.Factorizer(tau=tau, num_threads=p, gap_bases=bases) as F:
            src, ln = F.factorize(text)
            S = F.sss()
            So, _ = oracle.sss(text, tau)
            assert _eq(S, So), (name, "S")
            if S.size:
                sao = oracle.sa_s(text, tau, So)
                assert _eq(F.sa_s(), sao), (name, "SA_S")
                assert _eq(F.lcp_s(), oracle.lcp_s(text, So, sao)), (name, "LCP_S")
            b, e, s, blk = F.lpf()
            bo, eo, so, blko = oracle.lpf_opt(text, tau, p_ref=p)
            assert _eq(b, bo) and _eq(e, eo) and _eq(s, so) and _eq(blk, blko), (name, "LPF")
            st = F.stats()
            assert st["patt_lens"] == list(ost.patt_lens) and st["table_entries"] == ost.table_entries, (name, "gap parameters")
            assert st["num_lpf"] == ost.num_lpf and st["num_gaps"] == ost.num_gaps and st["len_lpf_phr"] == ost.len_lpf_phr
            assert _eq(src, osrc) and _eq(ln, olen), (name, "factor stream")
            rng = np.random.default_rng(9)
            qi = rng.integers(0, n, size=1500).astype(np.uint64)
            qj = rng.integers(0, n, size=1500).astype(np.uint64)
            if S.size > 1:
                qi[:400] = S[rng.integers(0, S.size, size=400)]
                qj[:400] = S[rng.integers(0, S.size, size=400)]
            qj[400:450] = qi[400:450]
            _, lce = oracle.lce_batch(text, tau, qi, qj)
            assert _eq(F.lce_batch(qi, qj), lce), (name, "LCE")
            assert (F.decode(src, ln, n) == text).all(), (name, "decode")


@pytest.mark.parametrize("n", [1, 2, 3, 5, 9, 17, 33, 63, 64, 65, 66, 130, 700, 1500, 3000])
def test_short_inputs_end_of_text_rules(capi, oracle, n):
    """Stale-fingerprint and window-fit rules at the end of the text (rolling_hash_index_107.hpp:93,112,123,145)."""
    bases = (257, 65_599, 131_071, 524_287, 1_048_573)
    for seed in range(4):
        text = np.random.default_rng(1000 * n + seed).integers(0, 3, size=n).astype(np.uint8)
        for tau in (4, 512):
            osrc, olen, _ = oracle.factorize(text, tau, bases, p_ref=1)
            with capi.Factorizer(tau=tau, num_threads=1, gap_bases=bases) as F:
                src, ln = F.factorize(text)
                assert _eq(src, osrc) and _eq(ln, olen), (n, seed, tau)
                assert (F.decode(src, ln, n) == text).all()


@pytest.mark.parametrize("env", [{}, {"LZ77SSS_SPARSE_GAP_INDEX": "0"}, {"LZ77SSS_SLICE_LEN": "40000", "LZ77SSS_BLOCK_LEN": "64"},
                                 {"LZ77SSS_SPARSE_GAP_INDEX": "0", "LZ77SSS_SLICE_LEN": "100000"},
                                 {"LZ77SSS_TRACK_DIV": "1", "LZ77SSS_BLOCK_LEN": "256"},
                                 {"LZ77SSS_TRACK_DIV": "1000000000000", "LZ77SSS_SLICE_LEN": "150000"}],
                         ids=["default_sparse", "dense", "sparse_small_slices", "dense_small_slices", "sparse_track_from_round_1",
                              "sparse_never_track"])
def test_emission_variants_give_the_same_factors(capi, oracle, env, monkeypatch):
    """Slicing, block length, change tracking and the sparse gap index (entries only for gaps and phrase starts; the
    positions under gap factors that run over whole phrases are added to a repair list while the parse iterates) are
    internal choices: the factor stream must not depend on them."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    bases = (401, 70_001, 123_457, 999_983, 1_048_575)
    repaired = sparse = 0
    for name, text in _texts().items():
        if name in ("zeros", "period3"):
            continue
        for tau, fm in ((16, capi.FACT_GREEDY), (512, capi.FACT_GREEDY), (64, capi.FACT_GREEDY_NAIVE)):
            if fm == capi.FACT_GREEDY:
                osrc, olen, _ = oracle.factorize(text, tau, bases, p_ref=1)
            elif oracle.ref_available(tau):
                osrc, olen, bases, _ = oracle.ref_factorize(text, tau, p=1, fact_mode=oracle.GREEDY_NAIVE)
            else:
                continue
            with capi.Factorizer(tau=tau, num_threads=1, gap_bases=bases, fact_mode=fm) as F:
                src, ln = F.factorize(text)
                assert _eq(src, osrc) and _eq(ln, olen), (name, tau, fm, env)
                st = F.stats()
                repaired += st["repair_entries"]
                sparse += st["sparse_slices"]
                assert st["sparse_fallbacks"] == 0, (name, tau, fm, env)
    if env.get("LZ77SSS_SPARSE_GAP_INDEX") == "0":
        assert sparse == 0 and repaired == 0
    else:
        assert sparse > 0 and repaired > 0, "the sparse index and its repair list were not exercised"


def _sharded_virtual(capi, text, world, tau, bases, fact_mode=None, env_slice=None):
    """world virtual ranks (threads, one handle each) on cuda:0 through lz77_sss_b200.sharded."""
    import threading
    import torch
    from lz77_sss_b200 import sharded
    fact_mode = capi.FACT_GREEDY if fact_mode is None else fact_mode
    n = text.size
    dev = torch.device("cuda", 0)
    shared = sharded.ThreadComm.make_shared(world)
    res, err = [None] * world, []
    host = np.ascontiguousarray(text)

    def run(r):
        try:
            with capi.Factorizer(tau=tau, num_threads=1, gap_bases=bases, fact_mode=fact_mode) as F:
                be = sharded.CudaBackend(F, dev)
                d_ptr = be.distribute_text(host, n)          # world == 1 inside: every virtual rank uploads the whole text
                entries = be.prepare(d_ptr, host.ctypes.data, n)
                comm = sharded.ThreadComm(r, world, dev, shared)
                res[r] = sharded.factorize_sharded(be, n, entries, r, world, dev, comm=comm)
        except Exception as e:  # noqa: BLE001
            err.append(e)
            shared["barrier"].abort()

    th = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    for t in th:
        t.start()
    for t in th:
        t.join(timeout=600)
    if err:
        raise err[0]
    return res[0]


@pytest.mark.parametrize("world", [1, 2, 3, 5])
def test_sharded_single_text_equals_sequential(capi, oracle, world):
    """ranges of one text parsed by different ranks, states and tables exchanged until nothing changes."""
    bases = (401, 70_001, 123_457, 999_983, 1_048_575)
    for name, text in _texts().items():
        if name in ("zeros", "period3"):
            continue
        for tau in (16, 512):
            osrc, olen, _ = oracle.factorize(text, tau, bases, p_ref=1)
            src, ln, passes = _sharded_virtual(capi, text, world, tau, bases)
            assert _eq(src, osrc) and _eq(ln, olen), (name, tau, world, passes)


def test_sharded_greedy_naive_and_small_slices(capi, oracle, monkeypatch):
    monkeypatch.setenv("LZ77SSS_SLICE_LEN", "30000")      # several slices per range: the resumed passes walk them all
    tau = 64
    _need_ref(oracle, tau)
    for name in ("rep", "dna", "xx"):
        text = _texts()[name]
        for fm in (oracle.GREEDY, oracle.GREEDY_NAIVE):
            rs, rl, bases, _ = oracle.ref_factorize(text, tau, p=1, fact_mode=fm)
            src, ln, _ = _sharded_virtual(capi, text, 3, tau, bases, fact_mode=fm)
            assert _eq(src, rs) and _eq(ln, rl), (name, fm)


@pytest.mark.parametrize("slice_len", [4096, 7777, 12345, 20011])
def test_sharded_gap_factor_into_a_slice_indexed_ahead(capi, oracle, monkeypatch, slice_len):
    """A sharded run indexes its slices before it knows their entry states (index-only pass).  A gap factor that reaches
    into such a slice inserts positions the sparse index has no entries for: they are added in front of the slice's first
    parse.  (At 1 GiB with ranges of equal work, a later factor copied from an older occurrence without this.)"""
    monkeypatch.setenv("LZ77SSS_SLICE_LEN", str(slice_len))
    tau = 64
    _need_ref(oracle, tau)
    for name in ("rep", "dna"):
        text = _texts()[name]
        rs, rl, bases, _ = oracle.ref_factorize(text, tau, p=1)
        for world in (2, 3):
            src, ln, _ = _sharded_virtual(capi, text, world, tau, bases)
            assert _eq(src, rs) and _eq(ln, rl), (name, world, slice_len)


def test_lnf_regression_dna52(capi, oracle):
    """The input on which a fused two-sided candidate kernel faulted (see lpf.cu, all_candidates_kernel)."""
    tau = 64
    _need_ref(oracle, tau)
    text = synth.dna_collection(400_000, seed=52, chrom_len=50_000)
    for pm in (oracle.LPF_LNF_NAIVE, oracle.LPF_LNF_OPT):
        rs, rl, bases, _ = oracle.ref_factorize(text, tau, p=1, phr_mode=pm)
        with capi.Factorizer(tau=tau, num_threads=1, gap_bases=bases, phr_mode=pm) as F:
            src, ln = F.factorize(text)
            assert _eq(src, rs) and _eq(ln, rl), pm


def test_64_bit_position_kernels_on_small_inputs(capi, oracle):
    """Texts of 4 GiB and more run the pos_t = uint64_t instantiation of every kernel; LZ77SSS_FORCE_POS64 runs it on
    inputs the oracle can check (subprocess: the switch is read once per process)."""
    code = r"""
import sys, numpy as np
sys.path.insert(0, %r); sys.path.insert(0, %r)
import oracle_lib as O
from lz77_sss_b200 import capi, synth
bases = (401, 70_001, 123_457, 999_983, 1_048_575)
texts = {"rep": synth.random_repetitive(300_000, seed=51), "dna": synth.dna_collection(400_000, seed=52, chrom_len=50_000),
         "markov": synth.markov_text(100_000, seed=53)}
for name, text in texts.items():
    for tau, p, pm in ((16, 1, 2), (512, 3, 2), (64, 1, 3), (64, 2, 0)):
        if pm == 2:
            osrc, olen, _ = O.factorize(text, tau, bases, p_ref=p)
            b = bases
        elif O.ref_available(tau):
            osrc, olen, b, _ = O.ref_factorize(text, tau, p=p, phr_mode=pm)
        else:
            continue
        with capi.Factorizer(tau=tau, num_threads=p, gap_bases=b, phr_mode=pm) as F:
            src, ln = F.factorize(text)
            assert src.size == osrc.size and (src == osrc).all() and (ln == olen).all(), (name, tau, p, pm)
            So, _ = O.sss(text, tau)
            assert (F.sss() == So).all() and (F.sa_s() == O.sa_s(text, tau, So)).all(), (name, tau)
            assert (F.decode(src, ln, text.size) == text).all()
print("pos64 ok")
""" % (ROOT, os.path.join(ROOT, "tests"))
    env = dict(os.environ, LZ77SSS_FORCE_POS64="1")
    r = subprocess.run([os.sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=900)
    assert r.returncode == 0 and "pos64 ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def _poly_prefix(text, base, upto):
    """fingerprint of text[:upto] modulo 2^31 - 1, numpy (8 characters per step)."""
    P = (1 << 31) - 1
    h = 0
    for c in text[:upto].tolist():
        h = (h * base + c) % P
    return h


@pytest.mark.parametrize("n", [1, 2, 15, 16, 17, 31, 33, 1000, 16383, 16384, 16385, 70_000, 300_001])
def test_rabin_karp_substring_against_reference(capi, oracle, n):
    """rabin_karp_substring<31> (rabin_karp_substring.hpp:35-240): the reference's own table `fps` for several sampling
    rates and its substring() on random queries, same base (drawn by the reference, pinned generator)."""
    _need_ref(oracle, 512)
    rng = np.random.default_rng(n)
    text = synth.random_repetitive(max(n, 64), seed=n)[:n] if n > 40 else rng.integers(0, 256, size=n).astype(np.uint8)
    nq = 400
    qpos = rng.integers(0, n, size=nq).astype(np.uint64)
    qlen = np.minimum(rng.integers(0, max(2, n), size=nq).astype(np.uint64), np.uint64(n) - qpos)
    qlen[:20] = np.minimum(rng.integers(1, 20, size=20).astype(np.uint64), np.uint64(n) - qpos[:20])
    with capi.Factorizer(tau=512) as F:
        for s in (16, 1, 7, 1000, 20_000):
            base, fps, want = oracle.ref_rks(text, s, qpos, qlen, p=1 + (n % 3))
            with capi.RabinKarpSubstring(F, text, base) as R:
                got = R.fps(s)
                assert _eq(got, fps), (n, s, "fps")
                assert _eq(R.substring(qpos, qlen), want), (n, s, "substring")


def test_rabin_karp_substring_large_prefix_scan(capi):
    """48 MiB: more than 1024 tiles, so the single-CTA scan of the tile aggregates works on chunks; spot checks against
    a direct evaluation and the algebra of the reference's own test (tests/test_rabin_karp_substring.cpp:74: concat)."""
    n = 48 << 20
    text = synth.dna_collection(n, seed=9, chrom_len=1 << 20)
    base = 1_291_279_556
    P = (1 << 31) - 1
    with capi.Factorizer(tau=512) as F, capi.RabinKarpSubstring(F, text, base) as R:
        tab = R.table()
        assert tab.size == n // 16
        for g in (0, 1, 1023, 1024, 5000, 70_000):
            assert int(tab[g]) == _poly_prefix(text, base, 16 * (g + 1)), g
        rng = np.random.default_rng(1)
        # fp(T[a..c)) = fp(T[a..b)) * base^(c-b) + fp(T[b..c))
        a = rng.integers(0, n - 3000, size=300).astype(np.uint64)
        l1 = rng.integers(1, 1500, size=300).astype(np.uint64)
        l2 = rng.integers(1, 1500, size=300).astype(np.uint64)
        f1, f2, f12 = R.substring(a, l1), R.substring(a + l1, l2), R.substring(a, l1 + l2)
        for k in range(300):
            assert (int(f1[k]) * pow(base, int(l2[k]), P) + int(f2[k])) % P == int(f12[k])
        # equal substrings of the collection have equal fingerprints: copies of the chromosome
        far = R.substring(np.array([100, 100 + (1 << 20)], dtype=np.uint64), np.array([5, 5], dtype=np.uint64))
        assert (text[100:105] == text[100 + (1 << 20):105 + (1 << 20)]).all() == (far[0] == far[1])
        whole = R.substring(np.array([0], dtype=np.uint64), np.array([n], dtype=np.uint64))
        assert int(whole[0]) == int(tab[-1])


def _check_factors(text, src, ln):
    n = text.size
    lens = np.maximum(ln, 1).astype(np.int64)
    pos = np.concatenate([[0], np.cumsum(lens)])
    assert pos[-1] == n
    cp = ln > 0
    assert (src[cp].astype(np.int64) < pos[:-1][cp]).all()
    assert (text[pos[:-1][~cp]] == src[~cp].astype(np.uint8)).all()
    # spot-check the copies
    idx = np.nonzero(cp)[0]
    for k in idx[:: max(1, idx.size // 2000)]:
        s, l, p = int(src[k]), int(ln[k]), int(pos[k])
        l = min(l, 4096)
        assert (text[s:s + l] == text[p:p + l]).all()


def _stream_hash(src, ln):
    import hashlib
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(src, dtype=np.uint64).tobytes())
    h.update(np.ascontiguousarray(ln, dtype=np.uint64).tobytes())
    return h.hexdigest()


def _bit_exact_against_reference(capi, oracle, text, tau, p):
    """factor stream of the CUDA path == factor stream of the reference's own code (oracle/_ref), same tau, same
    num_threads, same gap-index bases.  p > 1 only where the reference itself factorizes sequentially: its parallel gap
    path (lz77_sss.hpp:470-474, taken for rel_len_gaps > 0.2) is not run-to-run deterministic."""
    _need_ref(oracle, tau)
    rs, rl, bases, sec = oracle.ref_factorize(text, tau, p=p)
    with capi.Factorizer(tau=tau, num_threads=p, gap_bases=bases) as F:
        src, ln = F.factorize(text)
        st = F.stats()
        if p > 1:
            assert (text.size - st["len_lpf_phr"]) / text.size <= 0.2, "the reference takes its racy parallel gap path here"
        assert src.size == rs.size, (src.size, rs.size)
        assert _eq(src, rs) and _eq(ln, rl)
        assert _stream_hash(src, ln) == _stream_hash(rs, rl)
        assert (F.decode(src, ln, text.size) == text).all()
    return st, sec


def test_c1_64mib_bit_exact_against_reference(capi, oracle):
    """BASELINE config C1 at its full size: 64 MiB random_repetitive_string, tau = 512, the factor STREAM (not only its
    properties) against the reference build."""
    text = synth.random_repetitive(64 << 20, seed=1)
    _bit_exact_against_reference(capi, oracle, text, 512, 1)


def test_c2_first_256mib_bit_exact_against_reference(capi, oracle):
    """The first 256 MiB of the bench text (BASELINE config C2), num_threads = 1: one full emission slice with carried
    state, 2^24 table entries, 131072 parse blocks."""
    text = synth.dna_collection(256 << 20, seed=2, chrom_len=16 << 20)
    st, _ = _bit_exact_against_reference(capi, oracle, text, 512, 1)
    assert st["parse_slices"] >= 1


def test_c2_1gib_bit_exact_against_reference(capi, oracle):
    """The whole bench text (1 GiB, 4 emission slices with the table carried between them), at the thread count the
    reference arm of bench.py uses: the reference factorizes sequentially there (rel_len_gaps = 0.12)."""
    text = synth.dna_collection(1 << 30, seed=2, chrom_len=16 << 20)
    p = min(os.cpu_count() or 1, 64)
    st, sec = _bit_exact_against_reference(capi, oracle, text, 512, p)
    assert st["parse_slices"] == 4
    print(f"reference: {sec:.2f} s at {p} threads")


def test_c1_64mib_properties(capi):
    """BASELINE config C1 at full size: decode(factorize(T)) == T, factors well formed."""
    text = synth.random_repetitive(64 << 20, seed=1)
    with capi.Factorizer(tau=512, num_threads=1) as F:
        src, ln = F.factorize(text)
        _check_factors(text, src, ln)
        assert (F.decode(src, ln, text.size) == text).all()
        S = F.sss()
        assert (np.diff(S.astype(np.int64)) > 0).all() and S[-1] == text.size - 2 * 512 + 1


def test_c2_quarter_gib_dna_properties(capi):
    text = synth.dna_collection(256 << 20, seed=2, chrom_len=4 << 20)
    with capi.Factorizer(tau=512, num_threads=1) as F:
        src, ln = F.factorize(text)
        _check_factors(text, src, ln)
        assert (F.decode(src, ln, text.size) == text).all()


def test_decode_handmade_streams(capi, oracle):
    """decode (algorithms/common.cpp:33-54) on streams the factorizer would not produce: self-overlapping copies (runs and
    longer periods), deep chains (every factor copies the one before it), and malformed streams, which must be refused."""
    rng = np.random.default_rng(77)
    cases = []
    cases.append((np.array([97, 0], np.uint64), np.array([0, 999], np.uint64)))                         # a^1000
    cases.append((np.array([97, 98, 99, 0], np.uint64), np.array([0, 0, 0, 5000], np.uint64)))          # (abc)^k
    src, ln = [65, 66, 67, 68], [0, 0, 0, 0]                                                            # chain of depth 3000
    pos = 4
    for _ in range(3000):
        src.append(pos - 4); ln.append(4); pos += 4
    cases.append((np.array(src, np.uint64), np.array(ln, np.uint64)))
    src, ln, pos = [], [], 0                                                                            # random mixture
    for _ in range(20000):
        if pos < 10 or rng.random() < 0.2:
            src.append(int(rng.integers(0, 256))); ln.append(0); pos += 1
        else:
            l = int(rng.integers(1, 300)); s = int(rng.integers(0, pos))
            src.append(s); ln.append(l); pos += l
    cases.append((np.array(src, np.uint64), np.array(ln, np.uint64)))
    with capi.Factorizer(tau=512) as F:
        for src, ln in cases:
            n = int(np.maximum(ln, 1).sum())
            want = oracle.decode(src, ln, n)
            assert (F.decode(src, ln, n) == want).all()
            ms, rounds, launches, eq = F.time_decode(src, ln, n, iters=1)
            assert rounds <= 16 and launches > 0
        for src, ln, n in ((np.array([65, 5], np.uint64), np.array([0, 3], np.uint64), 4),      # source not earlier text
                           (np.array([65, 0], np.uint64), np.array([0, 3], np.uint64), 5),      # lengths do not add up
                           (np.array([65, 1], np.uint64), np.array([0, 3], np.uint64), 4)):     # source == own position
            with pytest.raises(capi.Lz77SssError):
                F.decode(src, ln, n)


def test_change_tracking_is_invisible(capi, monkeypatch):
    """A fresh slice re-parses every block behind the first change until few blocks change, then follows single changes
    (gap_parse.cu, dirty_from_bitmap_kernel).  When it switches is an internal choice: never, early and the default
    must give the same factors, here on 32768 blocks."""
    text = synth.dna_collection(64 << 20, seed=12, chrom_len=4 << 20)
    bases = (401, 70_001, 123_457, 999_983, 1_048_575)
    got = []
    for div in ("1000000000000", "64", "1"):
        monkeypatch.setenv("LZ77SSS_TRACK_DIV", div)
        with capi.Factorizer(tau=512, num_threads=1, gap_bases=bases) as F:
            got.append(F.factorize(text))
    for src, ln in got[1:]:
        assert _eq(src, got[0][0]) and _eq(ln, got[0][1])
    _check_factors(text, *got[0])


def test_cpp_header_example_runs(capi, tmp_path):
    exe = tmp_path / "example"
    cmd = ["/usr/bin/g++", "-std=c++20", "-O1", "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "lz77_sss_b200", "include"), os.path.join(ROOT, "examples", "lz77_sss_example.cpp"),
           "-L" + os.path.join(ROOT, "lz77_sss_b200"), "-llz77sss_b200", "-o", str(exe)]
    subprocess.run(cmd, check=True)
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(ROOT, "lz77_sss_b200"))
    r = subprocess.run([str(exe)], capture_output=True, text=True, env=env)
    assert r.returncode == 0 and "decode ok" in r.stdout and "rabin_karp_substring ok" in r.stdout, r.stdout + r.stderr


def _build_tool(name, tmp_path):
    exe = tmp_path / name
    cmd = ["/usr/bin/g++", "-std=c++20", "-O1", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "lz77_sss_b200", "include"),
           os.path.join(ROOT, "tools", "cpp", name + ".cpp"), "-L" + os.path.join(ROOT, "lz77_sss_b200"), "-llz77sss_b200", "-o", str(exe)]
    subprocess.run(cmd, check=True)
    return str(exe)


def test_file_level_tools_round_trip(capi, oracle, tmp_path):
    """tools/cpp: the file-level drop-ins for lz77-sss-3-aprx / lz77-sss-decode (Huffman factor files, misc/huffman.hpp),
    the tau-sweep harness with the reference's RESULT lines (bench/lz77_sss_bench_tau.cpp) and ssszip's gapped container
    (skip_phrases + vbyte, cli/ssszip.cpp:120-180)."""
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(ROOT, "lz77_sss_b200"))
    text = synth.dna_collection(300_000, seed=61, chrom_len=40_000)
    src_file = tmp_path / "text.bin"
    text.tofile(src_file)
    aprx, dec = _build_tool("lz77_sss_3_aprx", tmp_path), _build_tool("lz77_sss_decode", tmp_path)
    r = subprocess.run([aprx, str(src_file), str(tmp_path / "text.lz77"), "1"], capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    r = subprocess.run([dec, str(tmp_path / "text.lz77"), str(tmp_path / "text.out")], capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    assert (np.fromfile(tmp_path / "text.out", dtype=np.uint8) == text).all()
    if oracle.ref_available(512):          # the reference's own reader and decoder accept the file
        fs, fl, n = oracle.ref_huff_read(tmp_path / "text.lz77", text.size + 1)
        assert n == text.size and (oracle.ref_decode(fs, fl, n) == text).all()
    elif oracle.ref_require():
        pytest.fail("oracle/_ref missing and LZ77SSS_REQUIRE_REF=1")
    bench = _build_tool("lz77_sss_bench_tau", tmp_path)
    r = subprocess.run([bench, str(src_file), str(tmp_path / "results.txt"), "16", "512"], capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = open(tmp_path / "results.txt").read().strip().splitlines()
    assert len(lines) == 6                                       # tau = 16, 32, 64, 128, 256, 512
    for line, tau in zip(lines, (16, 32, 64, 128, 256, 512)):
        kv = dict(tok.split("=", 1) for tok in line.split()[1:])
        assert line.startswith("RESULT ") and kv["tau"] == str(tau) and kv["n"] == str(text.size) and kv["num_threads"] == "1"
        for key in ("lce_t", "nsv_psv", "lpf", "phrase_info", "init_gap_idx", "factorize", "num_factors", "comp_ratio", "time", "throughput"):
            assert key in kv
    zipper = _build_tool("ssszip_gapped", tmp_path)
    for cmd, a, b in (("encode", src_file, tmp_path / "text.gap"), ("decode", tmp_path / "text.gap", tmp_path / "text.gap.out")):
        r = subprocess.run([zipper, cmd, str(a), str(b)], capture_output=True, text=True, env=env)
        assert r.returncode == 0, r.stdout + r.stderr
    assert (np.fromfile(tmp_path / "text.gap.out", dtype=np.uint8) == text).all()
    assert os.path.getsize(tmp_path / "text.gap") < text.size // 2


def _valid_copies(text, src, ln):
    lens = np.maximum(ln, 1).astype(np.int64)
    pos = np.concatenate([[0], np.cumsum(lens)])[:-1]
    for k in np.nonzero(ln > 0)[0]:
        s_, l_, p_ = int(src[k]), int(ln[k]), int(pos[k])
        if not (s_ < p_ and (text[s_:s_ + l_] == text[p_:p_ + l_]).all()):
            return False
    lit = ln == 0
    return bool((text[pos[lit]] == src[lit].astype(np.uint8)).all())


def test_factorize_exact_equals_exact_lz77(capi, oracle):
    """factorize_exact (lz77_sss.hpp:188-200) at num_threads = 1: the factor lengths are those of the greedy longest-previous-
    factor parse (independent suffix-array checker in the oracle; the reference's three transform modes give the same
    lengths, tests/test_oracle.py), every source is a real earlier occurrence, and decode reproduces the text."""
    for name, text in _texts().items():
        if name in ("zeros", "period3"):
            continue
        for tau in (16, 512):
            es, el = oracle.lz77_exact(text)
            with capi.Factorizer(tau=tau, num_threads=1) as F:
                src, ln = F.factorize_exact(text)
                assert _eq(ln, el), (name, tau, src.size, es.size)
                assert _valid_copies(text, src, ln), (name, tau)
                assert (F.decode(src, ln, text.size) == text).all()
                st = F.stats()
                assert st["exact_parse_rounds"] >= 1 and st["t_exact_ns"] > 0


def test_factorize_exact_sections_against_reference_build(capi, oracle):
    """num_threads > 1: exact per section (16 sections per thread, cut at the approximation's phrase boundaries,
    transform_to_exact/common.cpp:48-74): lengths equal the reference's own factorize_exact<greedy, lpf_opt, with_samples>."""
    tau = 64
    _need_ref(oracle, tau)
    for name in ("rep", "dna", "markov", "xx"):
        text = _texts()[name]
        for p, tm in ((1, oracle.WITH_SAMPLES), (2, oracle.WITH_SAMPLES), (3, oracle.WITHOUT_SAMPLES)):
            # the reference draws its gap-index bases while it runs; peek them first so that both approximations agree
            bases = np.zeros(5, dtype=np.uint64)
            oracle.ref(tau).refdrv_peek_next_bases(bases)
            rs, rl, _ = oracle.ref_factorize_exact(text, tau, p=p, transf_mode=tm)
            with capi.Factorizer(tau=tau, num_threads=p, gap_bases=bases) as F:
                src, ln = F.factorize_exact(text)
                assert _eq(ln, rl), (name, p, tm, src.size, rs.size)
                assert _valid_copies(text, src, ln), (name, p)
