"""Run factorizations of the bench text on the GPU box and print the library's statistics.
  python tools/run_stats.py <MiB> <dna|rep> <num_threads> [profile] [decode]
profile: one more run with the per-kernel event timing on (with LZ77SSS_DEBUG_ROUNDS=1 also per parse round)
decode : decode the factors on the device once (so that an ncu capture of this command sees the decode kernels)"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from lz77_sss_b200 import capi, synth
mib = int(sys.argv[1]) if len(sys.argv) > 1 else 256
workload = sys.argv[2] if len(sys.argv) > 2 else "dna"
threads = int(sys.argv[3]) if len(sys.argv) > 3 else 1
flags = sys.argv[4:]
n = mib << 20
text = synth.dna_collection(n, seed=2, chrom_len=min(16 << 20, max(n // 64, 1 << 16))) if workload == "dna" else synth.random_repetitive(n, seed=1)
with capi.Factorizer(tau=512, num_threads=threads) as F:
    for rep in range(2):
        t0 = time.perf_counter()
        src, ln = F.factorize(text)
        dt = time.perf_counter() - t0
        st = F.stats()
        keep = {k: v for k, v in st.items() if not k.startswith("t_")}
        keep.update({k: round(v / 1e6, 2) for k, v in st.items() if k.startswith("t_")})
        print(json.dumps({"wall_ms": round(dt * 1e3, 1), **keep}))
    if "profile" in flags:
        F.kernel_profile(True)
        F.factorize(text)
        F.kernel_profile(False)
        for r in F.kernel_profile_rows():
            print(f"{r['ms']:10.3f} ms {r['launches']:5d}  {r['kernel']}")
    if "decode" in flags:
        ms, rounds, launches, eq = F.time_decode(src, ln, n, iters=1)
        print(json.dumps({"decode_ms": ms, "rounds": rounds, "launches": launches}))
