"""One text factorized by several GPUs: the host layer above the lz77sss_shard_* C ABI.

One process per GPU (torch.distributed, NCCL over NVLink; gloo in the CPU tests of this file's logic).
  1. every rank uploads its shard of the text and the shards are all-gathered, so each GPU holds the whole text;
  2. every rank runs the replicated front of the pipeline (synchronizing set, SA_S, LPF phrases, gap parameters:
     lz77sss_shard_prepare);
  3. the gap index and the greedy emission (factorize_greedy, greedy.cpp:34-140) are split by text range.  A range is
     parsed from an entry state and the table "last inserted position per bucket before the range".  Both are guesses
     for every range but the first; after each pass the ranks exchange exit states (all-gather of 4 words) and the
     per-range tables (all-gather of table_entries words), and the ranges whose inputs changed are parsed again
     (resuming from the sorted entries and traces the handle kept).  When no input changes any more, every range was
     parsed from the true state of the sequential algorithm, so the concatenated factors are the sequential stream;
  4. the rank of the last range finishes the last 65 positions on the host; the factors are gathered on rank 0.
Nothing here computes on text: it moves states and tables between ranks.  `backend` is the object that does the work
(CudaBackend = the C ABI; the CPU tests plug in a small reference parser to exercise the exchange logic).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

FRONT_PAD = 256          # api.cu: readable zero bytes in front of the text
TEXT_PAD = 65536         # stages.cuh: zero bytes behind it


def split_ranges(n: int, world: int, align: int = 2048):
    """world contiguous ranges covering [0, n), equal up to alignment; empty ranges at the end for tiny texts."""
    per = -(-n // world)
    per = -(-per // align) * align
    out = []
    for r in range(world):
        b = min(n, r * per)
        e = min(n, (r + 1) * per)
        out.append((b, e))
    return out


def chain_entries(exits, ranges, first_entry):
    """Entry state of every range from the exit states of the last pass: a range whose predecessor's parse already
    ran past its end is passed through (its exit will be its entry), so its successor gets the same state now."""
    entries = [tuple(first_entry)]
    for r in range(1, len(ranges)):
        prev_entry = entries[r - 1]
        pb, pe = ranges[r - 1]
        if pb >= pe or prev_entry[0] >= pe:        # empty or passed-through range: nothing starts inside it
            entries.append(prev_entry)
        else:
            entries.append(tuple(int(x) for x in exits[r - 1]))
    return entries


class CudaBackend:
    """The C ABI on this rank's GPU."""

    def __init__(self, factorizer, device):
        self.F = factorizer
        self.device = device
        self.text = None

    def distribute_text(self, host_shard: np.ndarray, n: int, group=None):
        """host_shard: this rank's bytes of the text (ranges of split_ranges(n, world, 1<<16)).  Returns the address of
        T[0] in a padded device buffer that holds the whole text."""
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        per = split_ranges(n, world, 1 << 16)[0][1]
        buf = torch.zeros(FRONT_PAD + per * world + TEXT_PAD, dtype=torch.uint8, device=self.device)
        body = buf[FRONT_PAD:FRONT_PAD + per * world]
        mine = torch.zeros(per, dtype=torch.uint8, device=self.device)
        if host_shard.size:
            mine[:host_shard.size].copy_(torch.from_numpy(host_shard), non_blocking=True)
        if world > 1:
            dist.all_gather_into_tensor(body, mine, group=group)
        else:
            body.copy_(mine)
        buf[FRONT_PAD + n:].zero_()
        torch.cuda.synchronize(self.device)
        self.text = buf
        return buf.data_ptr() + FRONT_PAD

    def prepare(self, d_text_ptr, host_ptr, n):
        self.F.shard_prepare(d_text_ptr, host_ptr, n)
        return self.F.shard_table_entries()

    def new_table(self, entries):
        return torch.zeros(entries, dtype=torch.int64, device=self.device)

    def entry_guess(self, pos):
        return self.F.shard_entry_guess(pos)

    # Work share of the first range relative to each later one.  A later range parses twice (its first parse runs on a
    # guessed table and entry state, the second one follows the changes: 25 ms after 50 ms at N = 2 on the bench text), the
    # first range only once, so it takes more of the text.
    FIRST_RANGE_SHARE = 1.35

    def split(self, world):
        if world == 1:
            return self.F.shard_split(1)
        fine = 64 * world
        pieces = self.F.shard_split(fine)                  # pieces of equal work
        total = self.FIRST_RANGE_SHARE + (world - 1)
        cuts, acc = [0], 0.0
        for r in range(world - 1):
            acc += (self.FIRST_RANGE_SHARE if r == 0 else 1.0) / total
            cuts.append(max(cuts[-1], min(fine, int(round(acc * fine)))))
        cuts.append(fine)
        n_end = pieces[-1][1]
        bounds = [pieces[c][0] if c < fine else n_end for c in cuts]
        return [(bounds[r], bounds[r + 1]) for r in range(world)]

    def index(self, beg, end, table_guess):
        torch.cuda.synchronize(self.device)
        self.F.shard_index(beg, end, table_guess.data_ptr())

    def parse(self, beg, end, entry, table_in, table_out):
        torch.cuda.synchronize(self.device)       # the tables were written on torch's stream
        return self.F.shard_parse(beg, end, entry, table_in.data_ptr(), table_out.data_ptr())

    def factors(self):
        return self.F.shard_factors_view()

    def finish(self, state):
        return self.F.shard_finish(state)


class TorchComm:
    """The collectives of step 3/4 over torch.distributed (NCCL on the GPUs, gloo in the CPU tests)."""

    def __init__(self, rank, world, device, group=None):
        self.rank, self.world, self.device, self.group = rank, world, device, group

    def all_gather(self, t):
        if self.world == 1:
            return [t]
        out = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(out, t, group=self.group)
        return out

    def any(self, flag: bool) -> bool:
        t = torch.tensor([1 if flag else 0], dtype=torch.int64, device=self.device)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return bool(int(t.item()))


class ThreadComm:
    """Same interface for `world` virtual ranks running as threads of one process on one GPU (tests)."""

    def __init__(self, rank, world, device, shared):
        self.rank, self.world, self.device, self.sh = rank, world, device, shared

    @staticmethod
    def make_shared(world):
        import threading
        return {"barrier": threading.Barrier(world), "slots": [None] * world}

    def all_gather(self, t):
        if t.is_cuda:
            torch.cuda.synchronize(t.device)
        self.sh["slots"][self.rank] = t.clone()
        self.sh["barrier"].wait()
        out = [x.clone() for x in self.sh["slots"]]
        self.sh["barrier"].wait()
        return out

    def any(self, flag: bool) -> bool:
        got = self.all_gather(torch.tensor([1 if flag else 0], dtype=torch.int64))
        return any(int(g.item()) for g in got)


_PINNED = {}


def _pinned(key, numel):
    """Grow-only pinned int64 staging buffers (allocating pinned memory costs milliseconds per call); keyed by rank because
    the tests run virtual ranks as threads of one process."""
    t = _PINNED.get(key)
    if t is None or t.numel() < numel:
        t = torch.empty(max(int(numel * 1.25), 1024), dtype=torch.int64, pin_memory=True)
        _PINNED[key] = t
    return t


def factorize_sharded(backend, n: int, table_entries: int, rank: int, world: int, device, group=None, max_passes=None,
                      log=None, comm=None):
    """Steps 3 and 4 of the module docstring.  Returns (src, len, passes) on rank 0 and (None, None, passes) elsewhere."""
    comm = comm or TorchComm(rank, world, device, group)
    # ranges of equal work where the backend can tell (gap bytes are the work: equal byte counts leave the rank that holds
    # the first copy of a repeated sequence as the straggler), else of equal length
    ranges = backend.split(world) if world > 1 and hasattr(backend, "split") else split_ranges(n, world)
    beg, end = ranges[rank]
    first_entry = (0, 0, 0, 1)
    entry = first_entry if rank == 0 or beg >= n else backend.entry_guess(beg)
    table_in = backend.new_table(table_entries)
    table_out = backend.new_table(table_entries)
    local = backend.new_table(table_entries)
    # pass 0 (backends that can): index the own range and exchange guessed tables, so that the first parse of the
    # later ranges already sees nearly the table it will end up with
    import time
    t_i = time.perf_counter()
    if world > 1 and hasattr(backend, "index"):
        if beg < end:
            backend.index(beg, end, local)
        t_j = time.perf_counter()
        guesses = comm.all_gather(local)
        for r in range(rank):
            torch.maximum(table_in, guesses[r], out=table_in)
        if log is not None:
            log.append({"index_ms": round((t_j - t_i) * 1e3, 2), "guess_exchange_ms": round((time.perf_counter() - t_j) * 1e3, 2)})
    exit_state = entry
    passes = 0
    need_parse = True
    cache = prev_local = new_in = None          # tables of the ranks in front, own table at the last exchange, their maximum
    limit = max_passes if max_passes is not None else world + 2
    while True:
        t_a = time.perf_counter()
        if need_parse:
            if beg < end:
                exit_state, _ = backend.parse(beg, end, entry, table_in, table_out)
                torch.where(table_out != table_in, table_out, torch.zeros_like(table_out), out=local)
            else:
                exit_state = entry
                local.zero_()
        passes += 1
        t_b = time.perf_counter()
        # exchange: exit states and the per-range tables (last inserted position + 1 per bucket inside the range)
        exits = [tuple(t.tolist()) for t in
                 comm.all_gather(torch.tensor([int(x) for x in exit_state], dtype=torch.int64, device=device))]
        if cache is None:
            # first exchange: whole tables (one per rank; a rank keeps those of the ranks in front of it)
            gathered = comm.all_gather(local)
            cache = [gathered[r] if r < rank else None for r in range(world)]
            prev_local = local.clone()
            new_in = torch.zeros_like(table_in)
            for r in range(rank):                  # positions grow with the rank: the maximum is the latest writer
                torch.maximum(new_in, cache[r], out=new_in)
            table_changed = not torch.equal(new_in, table_in)
        else:
            # later exchanges: only the entries that changed since the last one travel (index, value), not 8 bytes per bucket
            idx = torch.nonzero(local != prev_local).squeeze(1)
            vals = local[idx]
            prev_local[idx] = vals
            counts = [int(c.item()) for c in comm.all_gather(torch.tensor([idx.numel()], dtype=torch.int64, device=device))]
            cmax = max(counts)
            table_changed = False
            if cmax:
                buf = torch.zeros(2 * cmax, dtype=torch.int64, device=device)
                buf[:idx.numel()] = idx
                buf[cmax:cmax + idx.numel()] = vals
                parts = comm.all_gather(buf)
                touched = []
                for r in range(rank):
                    if counts[r]:
                        i = parts[r][:counts[r]]
                        cache[r][i] = parts[r][cmax:cmax + counts[r]]
                        touched.append(i)
                if touched:
                    u = torch.cat(touched)
                    m = torch.zeros_like(u)
                    for r in range(rank):
                        torch.maximum(m, cache[r][u], out=m)
                    table_changed = not torch.equal(m, table_in[u])
                    new_in[u] = m
        new_entry = chain_entries(exits, ranges, first_entry)[rank]
        changed = tuple(new_entry) != tuple(entry) or table_changed
        any_changed = comm.any(changed)
        if log is not None:
            log.append({"pass": passes, "changed_here": bool(changed), "parsed": bool(need_parse), "exit": exit_state,
                        "parse_ms": round((t_b - t_a) * 1e3, 2), "exchange_ms": round((time.perf_counter() - t_b) * 1e3, 2)})
        if not any_changed:
            break
        if passes >= limit:
            raise RuntimeError("sharded factorization did not reach its fixed point")   # cannot happen: rank r is final after pass r+1
        need_parse = changed
        if changed:
            entry = tuple(new_entry)
            table_in.copy_(new_in)
    # step 4: factors of the own range (+ the tail on the last non-empty range), gathered on rank 0
    t_c = time.perf_counter()
    if beg < end:
        src, ln = backend.factors()         # (the CUDA backend hands out views of the library's pinned arrays)
    else:
        src, ln = np.empty(0, np.uint64), np.empty(0, np.uint64)
    ts, tl = np.empty(0, np.uint64), np.empty(0, np.uint64)
    last_rank = max(r for r in range(world) if ranges[r][0] < ranges[r][1])
    if rank == last_rank and exit_state[0] < n:
        ts, tl = backend.finish(exit_state)
    if world == 1:
        return np.concatenate([src, ts]), np.concatenate([ln, tl]), passes
    mine_z = src.size + ts.size
    counts = [int(c.item()) for c in comm.all_gather(torch.tensor([mine_z], dtype=torch.int64, device=device))]
    zmax = max(max(counts), 1)
    cuda = torch.device(device).type == "cuda"
    mine = torch.empty(2 * zmax, dtype=torch.int64, device=device)
    # (uint64 -> int64 is a view; pinned sources go to the GPU without staging, the few tail factors are pageable)
    for k, (a, t) in enumerate(((src, ts), (ln, tl))):
        if a.size:
            mine[k * zmax:k * zmax + a.size].copy_(torch.from_numpy(a.view(np.int64)), non_blocking=cuda)
        if t.size:
            mine[k * zmax + a.size:k * zmax + mine_z].copy_(torch.from_numpy(np.ascontiguousarray(t).view(np.int64)))
    if cuda:
        torch.cuda.synchronize(device)      # the views may be overwritten by the handle's next call
    parts = comm.all_gather(mine)
    if rank != 0:
        if log is not None:
            log.append({"gather_ms": round((time.perf_counter() - t_c) * 1e3, 2)})
        return None, None, passes
    total = sum(counts)
    if cuda:   # one pinned buffer per output array, kept between calls: the returned arrays are views of it
        outs = [_pinned((rank, "down", k), total) for k in range(2)]
    else:
        outs = [torch.empty(max(total, 1), dtype=torch.int64) for _ in range(2)]
    at = 0
    for r in range(world):
        for k in range(2):
            outs[k][at:at + counts[r]].copy_(parts[r][k * zmax:k * zmax + counts[r]], non_blocking=cuda)
        at += counts[r]
    if cuda:
        torch.cuda.synchronize(device)
    if log is not None:
        log.append({"gather_ms": round((time.perf_counter() - t_c) * 1e3, 2)})
    return outs[0][:total].numpy().view(np.uint64), outs[1][:total].numpy().view(np.uint64), passes
