"""Synthetic workloads (fixed seeds) for tests and bench — SURVEY.md §8(d), BASELINE.md §4.

`random_repetitive` restates the reference's test generator `random_repetitive_string`
(/root/reference/include/lz77_sss/misc/utils.hpp:579-640): a stream of three operations —
new random character, copy of an earlier substring (may run into itself), single-character
run — drawn with weights {2-(rep+run), rep, run}; here with explicit parameters and a numpy
generator instead of std::random_device.
"""
from __future__ import annotations

import numpy as np


def random_repetitive(n: int, seed: int = 1, rep: float = 0.8, run: float = 0.3,
                      sigma: int = 256, max_rep_len: int | None = None,
                      max_run_len: int | None = None) -> np.ndarray:
    rng = np.random.default_rng(seed)
    out = np.empty(n, dtype=np.uint8)
    if n == 0:
        return out
    max_rep = max(int(rep * n / 100), 1) if max_rep_len is None else max_rep_len
    max_run = max(int(run * n / 200), 1) if max_run_len is None else max_run_len
    w = np.array([2 - (rep + run), rep, run], dtype=np.float64)
    w /= w.sum()
    out[0] = rng.integers(0, sigma)
    size = 1
    # draw operations in batches to keep the python loop short
    while size < n:
        ops = rng.choice(3, size=4096, p=w)
        chars = rng.integers(0, sigma, size=4096, dtype=np.uint8)
        lens_rep = rng.integers(1, max_rep + 1, size=4096)
        lens_run = rng.integers(1, max_run + 1, size=4096)
        srcs = rng.random(4096)
        for k in range(4096):
            if size >= n:
                break
            op = ops[k]
            if op == 0:
                out[size] = chars[k]
                size += 1
            elif op == 1:
                ln = min(n - size, int(lens_rep[k]))
                src = int(srcs[k] * size)
                if src + ln <= size:
                    out[size:size + ln] = out[src:src + ln]
                else:  # self-overlapping copy => periodic extension
                    per = size - src
                    reps = -(-ln // per)
                    out[size:size + ln] = np.tile(out[src:size], reps)[:ln]
                size += ln
            else:
                ln = min(n - size, int(lens_run[k]))
                out[size:size + ln] = chars[k]
                size += ln
    return out


def dna_collection(n: int, seed: int = 2, chrom_len: int = 16 << 20, mut_rate: float = 1e-3) -> np.ndarray:
    """chr19-like: one i.i.d. {A,C,G,T} chromosome concatenated n/chrom_len times, each copy
    with `mut_rate` point mutations."""
    rng = np.random.default_rng(seed)
    alpha = np.frombuffer(b"ACGT", dtype=np.uint8)
    chrom_len = min(chrom_len, n)
    base = alpha[rng.integers(0, 4, size=chrom_len)]
    out = np.empty(n, dtype=np.uint8)
    pos = 0
    while pos < n:
        ln = min(chrom_len, n - pos)
        out[pos:pos + ln] = base[:ln]
        if pos > 0:
            k = rng.binomial(ln, mut_rate)
            idx = rng.integers(0, ln, size=k)
            out[pos + idx] = alpha[rng.integers(0, 4, size=k)]
        pos += ln
    return out


def markov_text(n: int, seed: int = 4, reuse: float = 0.3) -> np.ndarray:
    """dewiki-like: printable random text with `reuse` of the content copied verbatim from
    earlier 100 B - 10 kB spans."""
    rng = np.random.default_rng(seed)
    out = np.empty(n, dtype=np.uint8)
    pos = 0
    while pos < n:
        if pos > 20000 and rng.random() < reuse:
            ln = int(min(n - pos, rng.integers(100, 10000)))
            src = int(rng.integers(0, pos - ln)) if pos > ln else 0
            out[pos:pos + ln] = out[src:src + ln]
        else:
            ln = int(min(n - pos, rng.integers(100, 10000)))
            # skewed letter distribution over 64 printable symbols
            out[pos:pos + ln] = (32 + np.minimum(rng.geometric(0.08, size=ln), 90)).astype(np.uint8)
        pos += ln
    return out


def genome_collection(n: int, seed: int = 3, genome_len: int = 30_000, edits: int = 10) -> np.ndarray:
    """sars2-like: one random genome over {A,C,G,T,N}, then copies with `edits` random substitutions,
    insertions or deletions each (against the previous copy, so the collection drifts), separated by '\\n'."""
    rng = np.random.default_rng(seed)
    alpha = np.frombuffer(b"ACGTN", dtype=np.uint8)
    cur = alpha[rng.choice(5, size=genome_len, p=[0.3, 0.2, 0.2, 0.29, 0.01])]
    out = np.empty(n, dtype=np.uint8)
    pos = 0
    while pos < n:
        ln = min(cur.size, n - pos)
        out[pos:pos + ln] = cur[:ln]
        pos += ln
        if pos < n:
            out[pos] = 10
            pos += 1
        # next copy: `edits` edit operations on the current genome
        kinds = rng.integers(0, 3, size=edits)
        at = np.sort(rng.integers(0, cur.size, size=edits))[::-1]
        for kind, a in zip(kinds, at):
            if kind == 0:
                cur[a] = alpha[rng.integers(0, 4)]
            elif kind == 1:
                cur = np.insert(cur, a, alpha[rng.integers(0, 4)])
            elif cur.size > genome_len // 2:
                cur = np.delete(cur, a)
    return out
