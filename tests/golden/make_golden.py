"""Generates tests/golden/*.npz: small inputs with the outputs of the REFERENCE's own code
(oracle/_ref/libref_tau*.so = /root/reference headers + oracle/shim, pinned hash parameters) and the
intermediate structures of the oracle restatement.  Run in the build container (needs /root/reference):
    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O  # noqa: E402
from lz77_sss_b200 import synth  # noqa: E402

CASES = {
    "rep60k": lambda: synth.random_repetitive(60_000, seed=21),
    "rep250k": lambda: synth.random_repetitive(250_000, seed=22),
    "dna200k": lambda: synth.dna_collection(200_000, seed=23, chrom_len=25_000),
    "markov100k": lambda: synth.markov_text(100_000, seed=24),
    "runs40k": lambda: np.concatenate([np.full(9000, 65, np.uint8), synth.random_repetitive(22_000, seed=25),
                                       np.tile(np.frombuffer(b"xyz", np.uint8), 3000)]),
}


def main():
    O.build_core()
    O.build_ref()
    for name, gen in CASES.items():
        text = gen()
        out = {"text": text}
        for tau in (4, 64, 512):
            for p in (1, 4):
                rs, rl, bases, _ = O.ref_factorize(text, tau, p=p)
                # the reference's parallel gap path is not deterministic; keep p>1 only when the
                # restatement (which always factorizes sequentially) reproduces the reference exactly
                os_, ol, st = O.factorize(text, tau, bases, p_ref=p)
                same = rs.size == os_.size and (rs == os_).all() and (rl == ol).all()
                if p == 1:
                    assert same, (name, tau, p)
                if not same:
                    continue
                key = f"tau{tau}_p{p}"
                out[key + "_bases"] = bases
                out[key + "_src"] = rs
                out[key + "_len"] = rl
                b, e, s, blk = O.lpf_opt(text, tau, p_ref=p)
                out[key + "_lpf"] = np.stack([b, e, s, blk.astype(np.uint64)])
            S, nq = O.sss(text, tau)
            out[f"tau{tau}_S"] = S
            out[f"tau{tau}_nq"] = np.uint64(nq)
            out[f"tau{tau}_SA"] = O.sa_s(text, tau, S)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, text.size, {k: v.shape for k, v in out.items() if k.endswith("_src")})


if __name__ == "__main__":
    main()
