"""Stage-by-stage comparison of the CUDA path with the oracle (run on a GPU box; diagnostic output)."""
import sys, os, time, traceback
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import oracle_lib as O
from lz77_sss_b200 import synth, capi

BASES = capi.DEFAULT_GAP_BASES

def first_diff(a, b):
    m = min(len(a), len(b))
    d = np.nonzero(a[:m] != b[:m])[0]
    return int(d[0]) if d.size else (m if len(a) != len(b) else -1)

def check_sss(t, tau, tag):
    with capi.Factorizer(tau=tau) as F:
        S, nq = F.build_sss(t)
    So, nqo = O.sss(t, tau)
    ok = len(S) == len(So) and (S == So).all() and nq == nqo
    print(f"[sss] {tag} tau={tau} n={len(t)} |S|={len(S)}/{len(So)} nq={nq}/{nqo} {'OK' if ok else 'FAIL first_diff=%d' % first_diff(S, So)}", flush=True)
    if not ok:
        k = first_diff(S, So)
        print("   gpu:", S[max(0,k-2):k+4], " oracle:", So[max(0,k-2):k+4])
    return ok

def check_full(t, tau, p, tag):
    n = len(t)
    ok_all = True
    with capi.Factorizer(tau=tau, num_threads=p, gap_bases=BASES) as F:
        t0 = time.time()
        try:
            src, ln = F.factorize(t)
        except Exception as e:
            print(f"[full] {tag} tau={tau} p={p} n={n} EXC {e}", flush=True)
            src = ln = None
        dt = time.time() - t0
        st = F.stats()
        S = F.sss()
        So, _ = O.sss(t, tau)
        ok = len(S) == len(So) and (S == So).all()
        ok_all &= ok
        msg = [f"S:{'ok' if ok else 'FAIL'}"]
        if ok and len(S):
            sa = F.sa_s()
            sao = O.sa_s(t, tau, So)
            ok = (sa == sao).all()
            ok_all &= ok
            msg.append(f"SA:{'ok' if ok else 'FAIL@%d' % first_diff(sa, sao)}")
            if ok:
                lcp = F.lcp_s()
                lcpo = O.lcp_s(t, So, sao)
                ok = (lcp == lcpo).all()
                ok_all &= ok
                msg.append(f"LCP:{'ok' if ok else 'FAIL@%d' % first_diff(lcp, lcpo)}")
            b, e, s_, blk = F.lpf()
            bo, eo, so, blko = O.lpf_opt(t, tau, p_ref=p)
            ok = len(b) == len(bo) and (b == bo).all() and (e == eo).all() and (s_ == so).all() and (blk == blko).all()
            ok_all &= ok
            msg.append(f"LPF({len(b)}/{len(bo)}):{'ok' if ok else 'FAIL@%d' % first_diff(b, bo)}")
            # LCE batch
            rng = np.random.default_rng(7)
            qi = rng.integers(0, n, size=2000).astype(np.uint64); qj = rng.integers(0, n, size=2000).astype(np.uint64)
            # bias some queries towards sync positions with long matches
            if len(S) > 2:
                qi[:500] = S[rng.integers(0, len(S), size=500)]; qj[:500] = S[rng.integers(0, len(S), size=500)]
            lg = F.lce_batch(qi, qj)
            _, ln_naive = O.lce_batch(t, tau, qi, qj)
            ok = (lg == ln_naive).all()
            ok_all &= ok
            msg.append(f"LCE:{'ok' if ok else 'FAIL@%d' % first_diff(lg, ln_naive)}")
        if src is not None:
            os_, ol, ost = O.factorize(t, tau, BASES, p_ref=p)
            ok = len(src) == len(os_) and (src == os_).all() and (ln == ol).all()
            ok_all &= ok
            msg.append(f"FACT({len(src)}/{len(os_)}):{'ok' if ok else 'FAIL@%d' % first_diff(src*(1<<20)+ln, os_*(1<<20)+ol)}")
            if not ok:
                k = first_diff(src*(1<<20)+ln, os_*(1<<20)+ol)
                print("   gpu:", list(zip(src[max(0,k-2):k+3], ln[max(0,k-2):k+3])), " oracle:", list(zip(os_[max(0,k-2):k+3], ol[max(0,k-2):k+3])))
                print("   stats gpu:", {k_: st[k_] for k_ in ('num_lpf','len_lpf_phr','num_gaps','patt_lens','roll_threshold','table_entries')},
                      " oracle:", ost.num_lpf, ost.len_lpf_phr, ost.num_gaps, list(ost.patt_lens), ost.roll_threshold, ost.table_entries)
            try:
                dec = F.decode(src, ln, n)
                okd = (dec == t).all()
            except Exception as ex:
                okd = False
                print("   decode EXC", ex)
            ok_all &= okd
            msg.append(f"DEC:{'ok' if okd else 'FAIL'}")
        else:
            ok_all = False
        print(f"[full] {tag} tau={tau} p={p} n={n} |S|={st['size_sss']} z={st['num_factors']} ctx_lv={st['ctx_levels']} dbl={st['doubling_rounds']} "
              f"t={dt*1e3:.1f}ms " + " ".join(msg), flush=True)
    return ok_all

def main():
    O.build_core()
    print(capi.lib().lz77sss_version().decode(), "devices:", capi.lib().lz77sss_device_count(), flush=True)
    fails = 0
    texts = {
        "rep20k": synth.random_repetitive(20000, seed=1),
        "rep200k": synth.random_repetitive(200000, seed=2),
        "rep1m": synth.random_repetitive(1000000, seed=3),
        "dna300k": synth.dna_collection(300000, seed=2, chrom_len=40000),
        "markov150k": synth.markov_text(150000, seed=4),
        "rand50k": np.random.default_rng(5).integers(0, 256, size=50000).astype(np.uint8),
        "zeros30k": np.zeros(30000, dtype=np.uint8),
        "per7": np.tile(np.frombuffer(b"abcdefg", dtype=np.uint8), 5000),
        "per100": np.tile(np.random.default_rng(6).integers(0, 4, size=100).astype(np.uint8), 700),
    }
    for tag, t in texts.items():
        for tau in (4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096):
            try:
                fails += not check_sss(t, tau, tag)
            except Exception:
                traceback.print_exc(); fails += 1
    for tag, t in texts.items():
        for tau in (4, 16, 64, 512):
            for p in (1, 3):
                try:
                    fails += not check_full(t, tau, p, tag)
                except Exception:
                    traceback.print_exc(); fails += 1
    print("FAILS:", fails)
    return 1 if fails else 0

if __name__ == "__main__":
    sys.exit(main())
