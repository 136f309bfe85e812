"""Time the rabin_karp_substring<31> build (prefix fingerprints) on the bench text; algorithmic bytes n + 4 n / 16."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lz77_sss_b200 import capi, synth
mib = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
n = mib << 20
text = synth.dna_collection(n, seed=2, chrom_len=min(16 << 20, max(n // 64, 1 << 16)))
with capi.Factorizer(tau=512) as F, capi.RabinKarpSubstring(F, text, 1_291_279_556) as R:
    ms, launches = R.time_build(10)
    alg = n + 4 * (n // 16)
    print(json.dumps({"n": n, "build_ms": ms, "launches": launches, "algorithmic_bytes": alg, "achieved_GBs": alg / ms / 1e6}))
