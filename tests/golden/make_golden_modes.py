"""Generates tests/golden/modes_*.npz: the factor streams of the REFERENCE's own code (oracle/_ref) for every
combination of factorize_mode x phrase_mode (lz77_sss.hpp:47-58) on small inputs.  Run in the build
container (needs /root/reference):
    python tests/golden/make_golden_modes.py
Keys: f"tau{tau}_p{p}_fm{fact_mode}_pm{phr_mode}_{src|len|bases}".
p > 1 is kept only for inputs below the reference's min_par_input_size (lz77_sss.hpp:86), where it
factorizes sequentially and is therefore deterministic.
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
    "modes_rep120k": lambda: synth.random_repetitive(120_000, seed=31),
    "modes_dna150k": lambda: synth.dna_collection(150_000, seed=32, chrom_len=20_000),
    "modes_markov12k": lambda: synth.markov_text(12_000, seed=33),
    "modes_xx16k": lambda: np.tile(synth.markov_text(8_000, seed=34), 2),
}


def main():
    O.build_core()
    O.build_ref()
    for name, gen in CASES.items():
        text = gen()
        out = {"text": text}
        for tau in (8, 64, 512):
            for p in (1, 3):
                for fm in (O.GREEDY_NAIVE, O.GREEDY, O.SKIP_PHRASES):
                    for pm in (O.LPF_NAIVE, O.LPF_LNF_NAIVE, O.LPF_OPT, O.LPF_LNF_OPT):
                        rs, rl, bases, _ = O.ref_factorize(text, tau, p=p, phr_mode=pm, fact_mode=fm)
                        if fm != O.SKIP_PHRASES:
                            assert (O.decode(rs, rl, text.size) == text).all(), (name, tau, p, fm, pm)
                        key = f"tau{tau}_p{p}_fm{fm}_pm{pm}"
                        out[key + "_bases"] = bases
                        out[key + "_src"] = rs.astype(np.uint32)
                        out[key + "_len"] = rl.astype(np.uint32)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, text.size, len(out))


if __name__ == "__main__":
    main()
