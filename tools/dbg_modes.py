"""Debug helper (GPU box): per mode combination, where the factor stream first differs from the golden reference output."""
import glob, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lz77_sss_b200 import capi

only = sys.argv[1] if len(sys.argv) > 1 else ""
for path in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "modes_*.npz"))):
    g = np.load(path)
    text = g["text"]
    for key in [k[:-4] for k in g.files if k.endswith("_src")]:
        if only and only not in key:
            continue
        tau, p, fm, pm = key.split("_")
        tau, p, fm, pm = int(tau[3:]), int(p[1:]), int(fm[2:]), int(pm[2:])
        try:
            with capi.Factorizer(tau=tau, num_threads=p, gap_bases=g[key + "_bases"], phr_mode=pm, fact_mode=fm) as F:
                src, ln = F.factorize(text)
        except Exception as e:  # noqa: BLE001
            print(os.path.basename(path), key, "ERROR", e)
            continue
        rs, rl = g[key + "_src"].astype(np.uint64), g[key + "_len"].astype(np.uint64)
        m = min(src.size, rs.size)
        bad = np.nonzero((src[:m] != rs[:m]) | (ln[:m] != rl[:m]))[0]
        if src.size == rs.size and bad.size == 0:
            print(os.path.basename(path), key, "ok", src.size)
        else:
            k = int(bad[0]) if bad.size else m
            pos = int(np.maximum(rl[:k], 1).sum())
            print(os.path.basename(path), key, "DIFF z", src.size, rs.size, "first at factor", k, "pos", pos,
                  "got", src[k:k + 3].tolist(), ln[k:k + 3].tolist(), "want", rs[k:k + 3].tolist(), rl[k:k + 3].tolist())
