"""Sharded emission against the single-GPU path at bench sizes, on ONE GPU: `world` virtual ranks (threads, one handle each).
  python tools/sharded_check.py <MiB> <world> [num_threads] [equal]
equal: ranges of equal length instead of equal work; genomes: the 4 GiB configuration's text generator (64-bit positions above 4 GiB).  Prints where the streams first differ (exit 1) or "equal" (exit 0)."""
import os, sys, threading
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from lz77_sss_b200 import capi, sharded, synth

mib = int(sys.argv[1]) if len(sys.argv) > 1 else 256
world = int(sys.argv[2]) if len(sys.argv) > 2 else 2
threads = int(sys.argv[3]) if len(sys.argv) > 3 else 16
equal = "equal" in sys.argv[4:]
profile = "profile" in sys.argv[4:]      # per-kernel event times of the LAST rank (all its passes)
n = mib << 20
text = (synth.genome_collection(n, seed=3) if "genomes" in sys.argv[4:] else
        synth.dna_collection(n, seed=2, chrom_len=min(16 << 20, max(n // 64, 1 << 16))))
dev = torch.device("cuda", 0)
shared = sharded.ThreadComm.make_shared(world)
res, err, logs = [None] * world, [], [[] for _ in range(world)]


class Backend(sharded.CudaBackend):
    def split(self, w):
        return sharded.split_ranges(n, w) if equal else self.F.shard_split(w)


def run(r):
    try:
        with capi.Factorizer(tau=512, num_threads=threads, gap_bases=capi.DEFAULT_GAP_BASES) as F:
            be = Backend(F, dev)
            d_ptr = be.distribute_text(text, n)
            entries = be.prepare(d_ptr, text.ctypes.data, n)
            comm = sharded.ThreadComm(r, world, dev, shared)
            if profile and r == world - 1:
                F.kernel_profile(True)
            res[r] = sharded.factorize_sharded(be, n, entries, r, world, dev, comm=comm, log=logs[r])
            if profile and r == world - 1:
                F.kernel_profile(False)
                for row in F.kernel_profile_rows():
                    print(f"rank {r}: {row['ms']:10.3f} ms {row['launches']:5d}  {row['kernel']}")
            if r == 0:
                res[r] = (res[r][0].copy(), res[r][1].copy(), res[r][2])
    except Exception as e:  # noqa: BLE001
        err.append(e)
        shared["barrier"].abort()


th = [threading.Thread(target=run, args=(r,)) for r in range(world)]
for t in th:
    t.start()
for t in th:
    t.join(timeout=900)
if err:
    raise err[0]
src, ln, passes = res[0]
with capi.Factorizer(tau=512, num_threads=threads, gap_bases=capi.DEFAULT_GAP_BASES) as G:
    s1, l1 = G.factorize(text)
for r in range(world):
    print("rank", r, [e for e in logs[r] if "pass" in e])
print("sharded", src.size, "single", s1.size, "passes", passes)
m = min(src.size, s1.size)
d = np.nonzero((src[:m] != s1[:m]) | (ln[:m] != l1[:m]))[0]
if d.size == 0 and src.size == s1.size:
    print("equal")
    sys.exit(0)
k = int(d[0]) if d.size else m
pos_a = int(np.maximum(ln[:k], 1).sum())
print("first difference at factor", k, "text position", pos_a)
for j in range(max(0, k - 2), min(m, k + 4)):
    print(j, "sharded", int(src[j]), int(ln[j]), "single", int(s1[j]), int(l1[j]))
sys.exit(1)
