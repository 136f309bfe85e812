"""Fixtures of the reference's file formats, written by the reference's OWN headers (oracle/_ref):
tests/golden/formats/<name>.factors.bin (z pairs of u64 src, len), <name>.lz77 (misc/huffman.hpp huff_writer output),
vbyte32.bin / vbyte64.bin (misc/vbyte.hpp encode_vbyte of vbyte_values.bin).  Run in the container that holds
/root/reference:  python tests/golden/make_golden_formats.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
import oracle_lib as O  # noqa: E402
from lz77_sss_b200 import synth  # noqa: E402

OUT = os.path.join(HERE, "formats")


def streams():
    rng = np.random.default_rng(2024)
    # real factor streams of the reference (3-aprx), small enough to commit
    for name, text in (("rep", synth.random_repetitive(60_000, seed=5)), ("dna", synth.dna_collection(90_000, seed=6, chrom_len=20_000)),
                       ("markov", synth.markov_text(50_000, seed=7))):
        src, ln, _, _ = O.ref_factorize(text, 64, p=1)
        yield name, src, ln, text.size
    # synthetic streams that exercise wide classes (64-bit positions), ties in the frequency table, one-symbol and
    # two-symbol alphabets, more than one block (2^14 factors), and code lengths at the cap of 15
    src, ln, pos = [], [], 0
    for k in range(20_000):
        if pos < 3 or k % 7 == 0:
            src.append(int(rng.integers(0, 256))); ln.append(0); pos += 1
        else:
            l = int(2 ** rng.integers(0, 37 if k % 5000 == 1 else 27)) + int(rng.integers(0, 3)); s = int(rng.integers(0, pos))
            src.append(s); ln.append(l); pos += l
    yield "wide", np.array(src, np.uint64), np.array(ln, np.uint64), pos
    yield "literals", np.arange(256, dtype=np.uint64), np.zeros(256, np.uint64), 256
    yield "run", np.array([97, 0], np.uint64), np.array([0, 99_999], np.uint64), 100_000
    # Fibonacci-like class frequencies: natural Huffman depths far beyond 15
    src, ln, pos = [65], [0], 1
    a, b = 1, 1
    for c in range(1, 30):
        for _ in range(a):
            src.append(pos - 1); ln.append(1 << (c - 1)); pos += 1 << (c - 1)
        a, b = b, a + b
        if len(src) > 18_000:
            break
    yield "skewed", np.array(src, np.uint64), np.array(ln, np.uint64), pos


def main():
    os.makedirs(OUT, exist_ok=True)
    for name, src, ln, n in streams():
        pairs = np.empty(2 * src.size, dtype=np.uint64)
        pairs[0::2] = src
        pairs[1::2] = ln
        pairs.tofile(os.path.join(OUT, f"{name}.factors.bin"))
        with open(os.path.join(OUT, f"{name}.n"), "w") as f:
            f.write(str(int(n)))
        O.ref_huff_write(src, ln, int(n), os.path.join(OUT, f"{name}.lz77"))
        s2, l2, n2 = O.ref_huff_read(os.path.join(OUT, f"{name}.lz77"), src.size + 1)
        assert n2 == n and (s2 == src).all() and (l2 == ln).all(), name
    rng = np.random.default_rng(7)
    vals = np.concatenate([np.array([0, 1, 127, 128, 16383, 16384, 2**32 - 1], np.uint64),
                           (2 ** rng.integers(0, 32, size=500)).astype(np.uint64) + rng.integers(0, 1000, size=500).astype(np.uint64)])
    vals32 = vals & np.uint64(0xFFFFFFFF)
    vals64 = np.concatenate([vals, np.array([2**40 + 5, 2**63 + 1, 2**64 - 1], np.uint64)])
    vals32.tofile(os.path.join(OUT, "vbyte_values32.bin"))
    vals64.tofile(os.path.join(OUT, "vbyte_values64.bin"))
    O.ref_vbyte_write(vals32, 32, os.path.join(OUT, "vbyte32.bin"))
    O.ref_vbyte_write(vals64, 64, os.path.join(OUT, "vbyte64.bin"))
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
