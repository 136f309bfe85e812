#!/usr/bin/env python
"""Large single-GPU runs of the hot path (BASELINE.json configs C3-C5 at sizes one B200 holds): texts of 4 GiB and more
take the pos_t = uint64_t instantiation of every kernel.  Prints one JSON line.
  python tools/big_run.py --size-mib 4160 --workload genomes --phr lpf_opt --ref          (C3's text, 3-aprx, vs reference)
  python tools/big_run.py --size-mib 16384 --workload markov --phr lpf_lnf_opt            (C4: lpf-lnf at 16 GiB)
Checks: factor lengths add up to n, every source is earlier text, sampled copies equal their sources, GPU decode of the
factors reproduces the text (round trip); with --ref the factor STREAM equals the reference build's (oracle/_ref)."""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402

MIB = 1 << 20


def make_text(args):
    from lz77_sss_b200 import synth
    n = args.size_mib * MIB
    if args.workload == "genomes":
        return synth.genome_collection(n, seed=3)
    if args.workload == "dna":
        return synth.dna_collection(n, seed=2, chrom_len=16 * MIB)
    if args.workload == "markov":      # dewiki-like; generated once for 256 MiB and extended by reuse of earlier spans
        base = synth.markov_text(min(n, 256 * MIB), seed=4)
        if base.size == n:
            return base
        out = np.empty(n, dtype=np.uint8)
        out[:base.size] = base
        rng = np.random.default_rng(44)
        pos = base.size
        while pos < n:
            ln = int(min(n - pos, rng.integers(64 * 1024, 8 * MIB)))
            src = int(rng.integers(0, pos - ln))
            out[pos:pos + ln] = out[src:src + ln]
            k = rng.binomial(ln, 2e-4)                       # edits inside the reused span
            out[pos + rng.integers(0, ln, size=k)] = rng.integers(32, 127, size=k).astype(np.uint8)
            pos += ln
        return out
    return synth.random_repetitive(n, seed=1)


def check_factors(text, src, ln, samples=4000):
    n = text.size
    lens = np.maximum(ln, 1).astype(np.int64)
    pos = np.concatenate([[0], np.cumsum(lens)])
    assert int(pos[-1]) == n, (int(pos[-1]), n)
    cp = ln > 0
    assert (src[cp].astype(np.int64) < pos[:-1][cp]).all(), "a source is not earlier text"
    assert (text[pos[:-1][~cp]] == src[~cp].astype(np.uint8)).all(), "a literal differs"
    idx = np.nonzero(cp)[0]
    for k in idx[:: max(1, idx.size // samples)]:
        s, l, p = int(src[k]), min(int(ln[k]), 1 << 16), int(pos[k])
        assert (text[s:s + l] == text[p:p + l]).all(), k
    return int((pos[:-1] >= (1 << 32)).sum())


def stream_hash(src, ln):
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(src, dtype=np.uint64).tobytes())
    h.update(np.ascontiguousarray(ln, dtype=np.uint64).tobytes())
    return h.hexdigest()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size-mib", type=int, default=4160)
    ap.add_argument("--workload", default="genomes", choices=["genomes", "dna", "markov", "rep"])
    ap.add_argument("--phr", default="lpf_opt", choices=["lpf_naive", "lpf_lnf_naive", "lpf_opt", "lpf_lnf_opt"])
    ap.add_argument("--tau", type=int, default=512)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--ref", action="store_true", help="compare the factor stream with the reference build (CPU, minutes)")
    ap.add_argument("--no-decode", action="store_true")
    ap.add_argument("--exact", action="store_true", help="factorize_exact instead of factorize_approximate (n < 4 GiB)")
    args = ap.parse_args()
    from lz77_sss_b200 import capi
    threads = args.threads or max(1, min(os.cpu_count() or 1, 64))
    t0 = time.perf_counter()
    text = make_text(args)
    n = text.size
    gen_s = time.perf_counter() - t0
    phr = {"lpf_naive": 0, "lpf_lnf_naive": 1, "lpf_opt": 2, "lpf_lnf_opt": 3}[args.phr]
    bases = capi.DEFAULT_GAP_BASES
    ref = None
    if args.ref:
        import oracle_lib as O
        rs, rl, bases, sec, phases = O.ref_factorize_logged(text, args.tau, p=threads, phr_mode=phr)
        keep = ("lce_t", "lce_rev_t", "nsv_psv", "lpf", "lnf", "select_phrases", "phrase_info", "init_gap_idx", "factorize")
        ref = {"seconds": sec, "factors": int(rs.size), "hash": stream_hash(rs, rl),
               "phases_ms": {k: v / 1e6 for k, v in phases.items() if k in keep}}
        del rs, rl
    out = {"workload": args.workload, "n_bytes": int(n), "tau": args.tau, "num_threads": threads, "phr_mode": args.phr,
           "pos_bits": 64 if n > 0xFFFFFFFF else 32, "generate_s": gen_s, "mode": "factorize_exact" if args.exact else "factorize_approximate"}
    F = capi.Factorizer(tau=args.tau, num_threads=threads, gap_bases=bases, phr_mode=phr)
    times = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        src, ln = (F.factorize_exact if args.exact else F.factorize)(text, cap=min(n // 4, 1 << 28) + 1024)
        times.append(time.perf_counter() - t0)
    st = F.stats()
    F.close()
    out.update(factors=int(src.size), size_sss=int(st["size_sss"]), seconds=min(times), mib_per_s=(n / MIB) / min(times),
               phases_ms={k[2:-3]: st[k] / 1e6 for k in st if k.startswith("t_")},
               index={k: int(st[k]) for k in ("parse_slices", "sparse_slices", "sparse_fallbacks", "repair_entries", "parse_rounds")},
               kernel_launches=int(st["kernel_launches"]),
               exact={k: int(st[k]) for k in ("exact_doubling_rounds", "exact_parse_rounds")})
    out["factors_starting_beyond_4gib"] = check_factors(text, src, ln)
    out["factors_well_formed"] = True
    if ref:
        out["reference"] = {k: ref[k] for k in ("seconds", "factors", "phases_ms")}
        out["reference"]["mib_per_s"] = (n / MIB) / ref["seconds"]
        out["stream_equals_reference"] = bool(int(src.size) == ref["factors"] and stream_hash(src, ln) == ref["hash"])
    if not args.no_decode:
        with capi.Factorizer(tau=args.tau) as D:
            if n <= (8 << 30):      # kernel time by CUDA events, compared with the text on the device
                d_text = D.device_text(text)
                try:
                    ms, rounds, launches, eq = D.time_decode(src, ln, n, iters=1, d_expect=d_text)
                finally:
                    D.free_device_text(d_text)
                out["decode"] = {"ms": ms, "mib_per_s": (n / MIB) / (ms / 1e3), "rounds": rounds, "gpu_launches": launches,
                                 "equals_text": bool(eq), "algorithmic_gb_per_s": (2 * n + 16 * int(src.size)) / (ms / 1e3) / 1e9}
            else:                   # 9 bytes of state per output byte leave no room for a second copy of the text in HBM
                t0 = time.perf_counter()
                dec = D.decode(src, ln, n)
                sec = time.perf_counter() - t0
                out["decode"] = {"e2e_ms": 1e3 * sec, "e2e_mib_per_s": (n / MIB) / sec, "equals_text": bool(np.array_equal(dec, text))}
                del dec
    print(json.dumps(out))
    ok = out["factors_well_formed"] and out.get("stream_equals_reference", True) and out.get("decode", {}).get("equals_text", True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
