"""Print the interesting parts of a bench.py JSON line."""
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
if len(sys.argv) > 2:
    print(d["config"]["tau"], "MiB/s", round(d["value"]), "e2e", round(d["e2e"]["value"]), "parity", d.get("parity_checked"), "|S|", d.get("size_sss"),
          "z", d.get("factors"), "sss frac", round(d.get("roofline_k1k2", d["roofline"])["frac"], 3), "cpu", round((d.get("cpu_baseline") or {}).get("value", 0), 1),
          "fallbacks", d.get("index", {}).get("sparse_fallbacks"))
    sys.exit(0)
print({k: d.get(k) for k in ("value", "ms_per_step", "parity_checked", "gpu_launches", "index")})
print("e2e", d["e2e"])
print("phases", {k: round(v, 1) for k, v in d["phases_ms"].items()})
print("roofline", {k: d["roofline"][k] for k in ("achieved", "frac", "kernel_ms")})
print("decode", d.get("decode"))
print("cpu", d.get("cpu_baseline"))
for r in d.get("other_kernels", []):
    print(f"  {r['kernel'][:60]:60s} {r.get('launches_per_step', '')!s:5s} {r.get('ms_per_step', r.get('kernel_ms', 0)):8.2f} ms  frac {r.get('frac')}")
