"""Host logic of the sharded factorization (lz77_sss_b200/sharded.py) on CPU: world_size 2 and 3 over gloo.

The exchange layer is exercised with a small stand-in backend (a sequential greedy parser with one hash table, the
same interface as the C ABI: parse a range from an entry state and a table, report the exit state and the table
behind the range).  The real backend is covered by the -m gpu tests.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from lz77_sss_b200 import sharded

E = 1 << 12


def _bucket(text, i):
    return (int(text[i]) * 257 + int(text[i + 1]) * 31 + int(text[i + 2])) % E if i + 2 < text.size else None


class ToyBackend:
    """Greedy parse with one table H[bucket of T[i..i+3)] = last inserted position + 1; every position a factor covers
    is inserted (the role of gap positions in factorize_greedy)."""

    def __init__(self, text):
        self.text = text
        self.out = ([], [])

    def new_table(self, entries):
        return torch.zeros(entries, dtype=torch.int64)

    def entry_guess(self, pos):
        return (pos, pos, 0, 1)

    def parse(self, beg, end, entry, table_in, table_out):
        T, n = self.text, self.text.size
        H = table_in.clone()
        i = entry[0]
        for q in range(beg, min(i, end)):          # a factor of the previous range reaches in here
            b = _bucket(T, q)
            if b is not None:
                H[b] = q + 1
        src, ln = [], []
        while i < end:
            b = _bucket(T, i)
            l, s = 0, int(T[i])
            if b is not None:
                occ = int(H[b])
                H[b] = i + 1
                if occ and T[occ - 1] == T[i]:
                    o = occ - 1
                    while i + l < n and T[o + l] == T[i + l]:
                        l += 1
                    s = o
            src.append(s)
            ln.append(l)
            nxt = i + max(1, l)
            for q in range(i + 1, min(nxt, end)):
                bq = _bucket(T, q)
                if bq is not None:
                    H[bq] = q + 1
            i = nxt
        table_out.copy_(H)
        self.out = (src, ln)
        return (i, i, 0, 1), len(src)

    def factors(self):
        return np.array(self.out[0], dtype=np.uint64), np.array(self.out[1], dtype=np.uint64)

    def finish(self, state):
        return np.empty(0, np.uint64), np.empty(0, np.uint64)


def _text(n, seed):
    rng = np.random.default_rng(seed)
    base = rng.integers(0, 4, size=n // 5).astype(np.uint8)
    t = np.concatenate([base, base[: n // 7], rng.integers(0, 4, size=n // 9).astype(np.uint8), base, base])
    return t[:n].copy()


def _worker(rank, world, port, n, seed, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        text = _text(n, seed)
        n = text.size
        log = []
        src, ln, passes = sharded.factorize_sharded(ToyBackend(text), n, E, rank, world, torch.device("cpu"), log=log)
        if rank == 0:
            q.put((src.tolist(), ln.tolist(), passes))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_matches_sequential_parse_over_gloo(world):
    n, seed = 20_000, 7 + world
    text = _text(n, seed)
    n = text.size
    one = ToyBackend(text)
    one.parse(0, n, (0, 0, 0, 1), torch.zeros(E, dtype=torch.int64), torch.zeros(E, dtype=torch.int64))
    want_src, want_len = one.factors()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 20_000, seed, q)) for r in range(world)]
    for p in procs:
        p.start()
    src, ln, passes = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert src == want_src.tolist() and ln == want_len.tolist()
    assert 2 <= passes <= world + 2


def test_split_and_chain():
    r = sharded.split_ranges(10_000, 4)
    assert r[0][0] == 0 and r[-1][1] == 10_000 and all(a[1] == b[0] for a, b in zip(r, r[1:]))
    assert all(a[0] % 2048 == 0 or a[0] == 10_000 for a in r)
    assert sharded.split_ranges(100, 4) == [(0, 100), (100, 100), (100, 100), (100, 100)]
    ranges = [(0, 100), (100, 200), (200, 300), (300, 400)]
    exits = [(250, 250, 3, 1), (9, 9, 9, 9), (310, 305, 5, 0), (400, 400, 7, 1)]
    ent = sharded.chain_entries(exits, ranges, (0, 0, 0, 1))
    # range 1 is passed through by the factor that ends at 250, so range 2 starts from that state too
    assert ent == [(0, 0, 0, 1), (250, 250, 3, 1), (250, 250, 3, 1), (310, 305, 5, 0)]


def test_thread_comm_virtual_ranks():
    """The in-process communicator the GPU tests use for several virtual ranks on one device."""
    import threading
    world = 3
    text = _text(9_000, 3)
    n = text.size
    one = ToyBackend(text)
    one.parse(0, n, (0, 0, 0, 1), torch.zeros(E, dtype=torch.int64), torch.zeros(E, dtype=torch.int64))
    want_src, want_len = one.factors()
    shared = sharded.ThreadComm.make_shared(world)
    res = [None] * world

    def run(r):
        comm = sharded.ThreadComm(r, world, torch.device("cpu"), shared)
        res[r] = sharded.factorize_sharded(ToyBackend(text), n, E, r, world, torch.device("cpu"), comm=comm)

    th = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    for t in th:
        t.start()
    for t in th:
        t.join(timeout=120)
    src, ln, passes = res[0]
    assert src.tolist() == want_src.tolist() and ln.tolist() == want_len.tolist()


def test_weighted_split_covers_the_text_and_gives_the_first_range_more():
    """CudaBackend.split: pieces of equal work from the library (here: a stand-in with equal lengths) are dealt out so that
    the first range, which parses once, gets FIRST_RANGE_SHARE of a later range's work."""
    n = 1_000_003

    class F:
        def shard_split(self, parts, align=1):
            cuts = [n * k // parts for k in range(parts + 1)]
            return [(cuts[k], cuts[k + 1]) for k in range(parts)]

    be = sharded.CudaBackend(F(), torch.device("cpu"))
    assert be.split(1) == [(0, n)]
    for world in (2, 3, 4, 8):
        r = be.split(world)
        assert len(r) == world and r[0][0] == 0 and r[-1][1] == n
        assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
        assert all(b <= e for b, e in r)
        first, later = r[0][1] - r[0][0], r[1][1] - r[1][0]
        assert abs(first / later - sharded.CudaBackend.FIRST_RANGE_SHARE) < 0.1
