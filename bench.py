#!/usr/bin/env python
"""bench.py — throughput of the lz77-sss 3-approximation hot path on B200.

Workload (BASELINE.json configs[1]): lz77-sss-3-aprx (factorize_approximate<greedy, lpf_opt>) on a 1 GiB synthetic
chr19-like DNA text (16 MiB i.i.d. {A,C,G,T} chromosome x 64 copies with 0.1 % point mutations, seed 2),
num_threads = 1 as in lz77-sss-bench-tau, tau = 512 by default (--tau for the other points of the sweep).
A "step" = one complete factorization of that text.  Metric: input MiB/s factorized (higher is better).

  value : text already resident in HBM (lz77sss_factorize_approximate_device), timed region = whole call
  e2e   : the public call on a HOST buffer (lz77sss_factorize_approximate): H2D of the text and D2H of the
          device results inside the timed region; factors delivered through the C callback into host arrays
  roofline : the fused fingerprint+SSS-selection kernel (K1+K2), algorithmic bytes n + 4|S| over its
             CUDA-event time on the library's stream, against the measured HBM copy peak
  cpu_baseline : the reference's own CPU code (oracle/_ref) on a bounded prefix of the same text

  parity_checked : outside the timed region the factor stream (count and SHA-256 of the src/len arrays) is compared with
          the reference arm's output on the SAME text, tau, num_threads and gap-index bases

num_threads is the reference's LOGICAL partition count of build_LPF_opt (it changes the factor stream, not the device
work); both arms use the same value, the number of host threads the reference arm runs on, so the two lines describe
the same configuration (`--threads` overrides it for both).

N > 1 (torchrun): the SAME text sharded over the N GPUs (lz77_sss_b200/sharded.py; strong scaling, NCCL all-gathers of
shards, tables and exit states); `--independent` keeps one text per GPU (weak scaling, no collective).
`--impl reference` times the reference's own CPU implementation (oracle/_ref) on the whole text (rank 0 only); its
line carries the reference's own phase split (`reference_phases_ms`, from its RESULT line).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

MIB = 1 << 20
METRIC = "input MiB/s factorized (3-aprx)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--tau", type=int, default=512)
    ap.add_argument("--size-mib", type=int, default=1024)
    ap.add_argument("--workload", default="dna", choices=["dna", "rep"])
    ap.add_argument("--cpu-sample-mib", type=int, default=0, help="0 = the whole text (same configuration as the GPU arm)")
    ap.add_argument("--threads", type=int, default=0,
                    help="num_threads of BOTH arms (the reference's thread count = the logical LPF partition count); "
                         "0 = min(host cores, 64)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--independent", action="store_true",
                    help="N > 1: every rank factorizes its own text (weak scaling) instead of sharding one text")
    a = ap.parse_args()
    if a.threads <= 0:
        a.threads = max(1, min(os.cpu_count() or 1, 64))
    if a.cpu_sample_mib <= 0 or a.cpu_sample_mib > a.size_mib:
        a.cpu_sample_mib = a.size_mib
    return a


def make_text(args, rank):
    from lz77_sss_b200 import synth
    n = args.size_mib * MIB
    if args.workload == "dna":
        return synth.dna_collection(n, seed=2 + rank, chrom_len=min(16 * MIB, max(n // 64, 1 << 16)))
    return synth.random_repetitive(n, seed=1 + rank)


def workload_name(args):
    if args.workload == "dna":
        return (f"C2: lz77-sss-3-aprx on {args.size_mib} MiB synthetic chr19-like DNA (64 mutated copies of one "
                f"chromosome), tau={args.tau}, num_threads={args.threads}")
    return f"C1: lz77-sss-3-aprx on {args.size_mib} MiB random_repetitive_string, tau={args.tau}, num_threads={args.threads}"


def config_dict(args):
    """Identical for both arms: what is factorized and with which parameters."""
    return {"workload": workload_name(args), "n_bytes": args.size_mib * MIB, "tau": args.tau, "num_threads": args.threads,
            "l2": f"input ({args.size_mib} MiB) is larger than L2 (126 MB); no explicit flush"}


def sss_traffic(n, tau):
    """dram bytes of one K1+K2 launch from the committed ncu --set full capture of this workload (null if none)."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "sss_traffic.json")))
        return int(t[f"n{n}_tau{tau}"]["traffic_bytes"])
    except (OSError, KeyError, ValueError):
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for k, nm in enumerate(names):
                if len(r) > 4 + k and r[4 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def scatter_traffic():
    """DRAM bytes of one captured launch of the gap index scatter (profiles/scatter_traffic.json, from ncu --set full)."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "scatter_traffic.json")))
    except (OSError, ValueError):
        return None


def stream_hash(src, ln):
    import hashlib
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(src, dtype=np.uint64).tobytes())
    h.update(np.ascontiguousarray(ln, dtype=np.uint64).tobytes())
    return h.hexdigest()


REF_PHASES = ("lce_t", "nsv_psv", "lpf", "phrase_info", "init_gap_idx", "factorize")


def cpu_reference_run(text, tau, threads, keep=False):
    """The reference's own factorize_approximate<greedy, lpf_opt, tau> (oracle/_ref), timed as lz77_sss.hpp:314-353.
    Returns a dict: seconds, kind, cores, phases_ms (the reference's own RESULT line), and with keep the factor
    stream's count / hash / bases for the parity check."""
    import oracle_lib as O
    if O.ref_available(tau):
        src, ln, bases, sec, ph = O.ref_factorize_logged(text, tau, p=threads)
        out = {"sec": sec, "kind": "reference", "cores": threads,
               "phases_ms": {k: ph[k] / 1e6 for k in REF_PHASES if k in ph}}
        if keep:
            out.update(z=int(src.size), hash=stream_hash(src, ln), bases=[int(b) for b in bases])
        return out
    # the reference library was not prebuilt for this tau: time the oracle's sequential restatement instead
    from lz77_sss_b200 import capi
    t0 = time.perf_counter()
    src, ln, _ = O.factorize(text, tau, capi.DEFAULT_GAP_BASES, p_ref=threads)
    out = {"sec": time.perf_counter() - t0, "kind": "port", "cores": 1, "phases_ms": {}}
    if keep:
        out.update(z=int(src.size), hash=stream_hash(src, ln), bases=list(capi.DEFAULT_GAP_BASES))
    return out


def shim_share(phases_ms):
    """lce_t (build_lce: synchronizing set, its suffix order, LCP, RMQ) runs oracle/shim + oracle/core code, because the
    `lce` submodule that holds it is absent from the reference tree; every other phase is the reference's own code."""
    tot = sum(phases_ms.values())
    return (phases_ms.get("lce_t", 0.0) / tot) if tot else None


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    text = make_text(args, 0)
    sample = text[: args.cpu_sample_mib * MIB]
    times, phases, r = [], {}, None
    for it in range(args.warmup + args.steps):
        r = cpu_reference_run(sample, args.tau, args.threads)
        if it >= args.warmup:
            times.append(r["sec"])
            for k, v in r["phases_ms"].items():
                phases[k] = phases.get(k, 0.0) + v / args.steps
    ms = 1e3 * sum(times) / len(times)
    val = (sample.size / MIB) / (ms / 1e3)
    what = "the whole text" if sample.size == text.size else f"first {sample.size // MIB} MiB of the text"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "MiB/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic", "config": config_dict(args),
        "cpu_baseline": {"value": val, "unit": "MiB/s", "cores": r["cores"], "kind": r["kind"],
                         "sample": f"{what}, num_threads={args.threads}, {args.steps} timed runs"},
        "reference_phases_ms": phases, "reference_shim_share": shim_share(phases),
        "e2e": {"value": val, "unit": "MiB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_sharded(args, world, rank, local, barrier):
    """N > 1: ONE text, sharded (lz77_sss_b200/sharded.py).  Strong scaling: the text is BASELINE's
    configs[1] for every N.  value: shards resident in HBM, timed = all-gather of the shards + replicated front +
    range parses with their exchanges + gather of the factors on rank 0.  e2e: the shards start in pinned host memory."""
    import torch
    import torch.distributed as dist
    from lz77_sss_b200 import capi, sharded
    dev = torch.device("cuda", local)
    # the ranks of one box share its cores: the host side of a rank (coverage chains, copies) takes its share of them
    os.environ.setdefault("LZ77SSS_HOST_THREADS", str(max(1, (os.cpu_count() or 1) // world)))
    text = make_text(args, 0)                      # the same text on every rank; a rank only uploads its shard
    n = text.size
    sh_ranges = sharded.split_ranges(n, world, 1 << 16)
    sb, se = sh_ranges[rank]
    pinned = torch.empty(max(1, se - sb), dtype=torch.uint8).pin_memory()
    pinned.numpy()[:se - sb] = text[sb:se]
    per = sh_ranges[0][1]
    F = capi.Factorizer(tau=args.tau, num_threads=args.threads, gap_bases=capi.DEFAULT_GAP_BASES, device=local)
    be = sharded.CudaBackend(F, dev)
    buf = torch.zeros(sharded.FRONT_PAD + per * world + sharded.TEXT_PAD, dtype=torch.uint8, device=dev)
    body = buf[sharded.FRONT_PAD:sharded.FRONT_PAD + per * world]
    mine = torch.zeros(per, dtype=torch.uint8, device=dev)
    d_ptr = buf.data_ptr() + sharded.FRONT_PAD
    stats = {"launches": 0, "passes": 0, "z": 0, "h2d": 0, "d2h": 0}

    def step(from_host):
        if from_host:
            mine[:se - sb].copy_(pinned[:se - sb], non_blocking=True)
        dist.all_gather_into_tensor(body, mine)
        buf[sharded.FRONT_PAD + n:].zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        entries = be.prepare(d_ptr, text.ctypes.data, n)
        t1 = time.perf_counter()
        log = []
        src, ln, passes = sharded.factorize_sharded(be, n, entries, rank, world, dev, log=log)
        if os.environ.get("LZ77SSS_BENCH_VERBOSE"):
            print(f"rank {rank}: prepare {1e3 * (t1 - t0):.1f} ms, sharded {1e3 * (time.perf_counter() - t1):.1f} ms, {log}",
                  file=sys.stderr, flush=True)
        st = F.stats()
        stats["launches"] += st["kernel_launches"]
        stats["passes"] = passes
        stats["h2d"] = (se - sb if from_host else 0) + st["h2d_bytes"]
        stats["d2h"] = st["d2h_bytes"]
        if rank == 0:
            stats["z"] = int(src.size)
            stats["hash"] = None if not stats.get("want_hash") else stream_hash(src, ln)

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        stats["launches"] = 0
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / steps

    mine[:se - sb].copy_(pinned[:se - sb])
    sampler = ClockSampler(local)
    sampler.start()
    ms_res = timed(lambda: step(False), args.steps, args.warmup)
    lt = torch.tensor([stats["launches"]], device="cuda", dtype=torch.int64)
    dist.all_reduce(lt)
    ms_e2e = timed(lambda: step(True), args.steps, max(1, args.warmup - 2))
    clocks = sampler.stop()
    # outside the timed region: the sharded factor stream against the single-GPU path on rank 0's GPU (same text and parameters)
    stats["want_hash"] = True
    keep = (stats["h2d"], stats["d2h"])              # (of the last end-to-end step)
    step(False)
    stats["h2d"], stats["d2h"] = keep
    sharded_ok = None
    if rank == 0:
        with capi.Factorizer(tau=args.tau, num_threads=args.threads, gap_bases=capi.DEFAULT_GAP_BASES, device=local) as G:
            s1, l1 = G.factorize(text)
        sharded_ok = bool(int(s1.size) == stats["z"] and stream_hash(s1, l1) == stats["hash"])
        if not sharded_ok:
            print(f"sharded stream: {stats['z']} factors, single GPU: {int(s1.size)}", file=sys.stderr, flush=True)
        del s1, l1
    barrier()
    if rank == 0 and not sharded_ok:
        raise RuntimeError("the sharded factor stream differs from the single-GPU path: no number is reported")
    st = F.stats()
    k_ms, s_size = F.time_sss_kernel(C.c_void_p(d_ptr), n, iters=10)
    hb = torch.tensor([stats["h2d"], stats["d2h"]], device="cuda", dtype=torch.int64)
    dist.all_reduce(hb)
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        alg_bytes = n + 4 * s_size
        achieved = alg_bytes / (k_ms / 1e3) / 1e9
        out = {
            "metric": METRIC, "value": (n / MIB) / (ms_res / 1e3), "unit": "MiB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_res, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": config_dict(args),
            "parallelism": f"one text sharded over {world} GPUs: shards all-gathered (NCCL), replicated SSS/SA_S/LPF "
                           f"front, gap index + greedy emission split by text range, tables and exit states "
                           f"all-gathered between passes ({stats['passes']} passes); every rank also holds the whole text in host memory",
            "factors": stats["z"], "size_sss": int(st["size_sss"]),
            "parity_checked": sharded_ok, "parity": {"against": "the single-GPU path on rank 0 (itself compared with oracle/_ref at N = 1)",
                                                     "sha256_equal": sharded_ok},
            "cpu_baseline": None, "cpu_baseline_note": "timed at N = 1 only (bench.py --gpus 1 and --impl reference)",
            "e2e": {"value": (n / MIB) / (ms_e2e / 1e3), "unit": "MiB/s", "h2d_bytes_per_step": int(hb[0].item()),
                    "d2h_bytes_per_step": int(hb[1].item()), "ms_per_step": ms_e2e},
            "gpu_launches": int(lt.item()),
            "roofline": {"kernel": "sss_kernel (fused window fingerprints + SSS selection)", "bound": "hbm",
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s",
                         "algorithmic_bytes": int(alg_bytes), "kernel_ms": k_ms, "traffic": sss_traffic(n, args.tau)},
            "phases_ms": {k[2:-3]: st[k] / 1e6 for k in st if k.startswith("t_")},
            "clocks": clocks,
        }
        print(json.dumps(out))
    barrier()
    F.close()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
        return
    import torch
    import torch.distributed as dist
    from lz77_sss_b200 import capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if world > 1 and not args.independent:
        run_sharded(args, world, rank, local, barrier)
        dist.destroy_process_group()
        return

    text = make_text(args, rank)
    n = text.size
    pinned = torch.empty(n, dtype=torch.uint8).pin_memory()          # plumbing: pinned host memory for the e2e leg
    pinned.numpy()[:] = text
    host_ptr = pinned.data_ptr()
    F = capi.Factorizer(tau=args.tau, num_threads=args.threads, gap_bases=capi.DEFAULT_GAP_BASES, device=local)
    cap = n // 8 + 1024
    src = np.empty(cap, dtype=np.uint64)
    ln = np.empty(cap, dtype=np.uint64)
    sink = capi.Sink(src.ctypes.data_as(C.POINTER(C.c_uint64)), ln.ctypes.data_as(C.POINTER(C.c_uint64)), cap, 0)
    d_text = F.device_text(text)

    launches = {"n": 0}

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        launches["n"] = 0
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)       # CUDA events; the host part of a step lies between them
        if world > 1:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms / steps

    def step_resident():
        z = F.factorize_device_into(d_text, host_ptr, n, sink)
        launches["n"] += F.stats()["kernel_launches"]
        return z

    def step_e2e():
        return F.factorize_into(host_ptr, n, sink)

    sampler = ClockSampler(local)
    sampler.start()
    ms_res = timed(step_resident, args.steps, args.warmup)
    launches_timed = launches["n"]
    st_res = F.stats()
    z = st_res["num_factors"]
    ms_e2e = timed(step_e2e, args.steps, max(1, args.warmup - 2))
    st_e2e = F.stats()
    clocks = sampler.stop()

    # one more, untimed, step with the per-kernel event timing on: the kernels that make up a step
    F.kernel_profile(True)
    step_resident()
    F.kernel_profile(False)
    prof_rows = F.kernel_profile_rows()
    st_prof = F.stats()

    # decode (lz77_sss<>::decode, algorithms/common.cpp:33-54) of the factors just produced: device-resident kernel time
    # (CUDA events on the library's stream, output compared with the text in HBM) and the public call with host buffers
    zz = int(sink.count)
    dec_ms, dec_rounds, dec_launches, dec_ok = F.time_decode(src[:zz], ln[:zz], n, iters=5, d_expect=d_text)
    t0 = time.perf_counter()
    dec_host = F.decode(src[:zz], ln[:zz], n)
    dec_e2e_ms = 1e3 * (time.perf_counter() - t0)
    dec_ok = bool(dec_ok) and bool((dec_host == text).all())
    del dec_host

    # the dominant streaming kernel (K1+K2), CUDA events on the library's own stream
    k_ms, s_size = F.time_sss_kernel(d_text, n, iters=10)
    # the other streaming kernel of the path: the prefix-fingerprint table of rabin_karp_substring<31> (row a19)
    rks_ms = None
    if rank == 0:
        with capi.RabinKarpSubstring(F, text, 1_291_279_556) as R:
            rks_ms, _ = R.time_build(10)

    value = world * (n / MIB) / (ms_res / 1e3)
    e2e = world * (n / MIB) / (ms_e2e / 1e3)
    out = None
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        alg_bytes = n + 4 * s_size
        achieved = alg_bytes / (k_ms / 1e3) / 1e9
        out = {
            "metric": METRIC, "value": value, "unit": "MiB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_res, "higher_is_better": True, "scaling": "strong" if world == 1 else "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": config_dict(args),
            "parallelism": "single GPU" if world == 1 else f"{world} independent texts, one per GPU (--independent)",
            "factors": int(z), "size_sss": int(st_res["size_sss"]),
            "e2e": {"value": e2e, "unit": "MiB/s", "h2d_bytes_per_step": int(st_e2e["h2d_bytes"]),
                    "d2h_bytes_per_step": int(st_e2e["d2h_bytes"]), "ms_per_step": ms_e2e},
            "gpu_launches": int(launches_timed),
            "roofline": {"kernel": "sss_kernel (fused window fingerprints + SSS selection)", "bound": "hbm",
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s",
                         "algorithmic_bytes": int(alg_bytes), "kernel_ms": k_ms, "traffic": sss_traffic(n, args.tau)},
            "phases_ms": {k[2:-3]: st_res[k] / 1e6 for k in st_res if k.startswith("t_")},
            "clocks": clocks,
        }
        if rks_ms:
            rks_bytes = n + 4 * (n // 16)
            rks_gbs = rks_bytes / (rks_ms / 1e3) / 1e9
            out["other_kernels"] = [{"kernel": "rks_build_kernel (rabin_karp_substring<31> prefix fingerprints)", "bound": "hbm",
                                     "achieved": rks_gbs, "peak": peak, "unit": "GB/s", "frac": rks_gbs / peak,
                                     "algorithmic_bytes": int(rks_bytes), "kernel_ms": rks_ms}]
        dec_bytes = 2 * n + 16 * zz
        out["decode"] = {"metric": "output MiB/s decoded", "value": (n / MIB) / (dec_ms / 1e3), "unit": "MiB/s", "ms": dec_ms,
                         "rounds": dec_rounds, "gpu_launches": dec_launches, "round_trip_ok": dec_ok,
                         "e2e": {"value": (n / MIB) / (dec_e2e_ms / 1e3), "unit": "MiB/s", "ms": dec_e2e_ms,
                                 "h2d_bytes": 16 * zz, "d2h_bytes": int(n)},
                         "roofline": {"bound": "hbm", "algorithmic_bytes": int(dec_bytes), "achieved": dec_bytes / (dec_ms / 1e3) / 1e9,
                                      "peak": peak, "unit": "GB/s", "frac": dec_bytes / (dec_ms / 1e3) / 1e9 / peak}}
        out.setdefault("other_kernels", [])
        for r in prof_rows:         # live CUDA-event times of one profiled step; frac only where the kernel has a byte count
            row = {"kernel": r["kernel"], "launches_per_step": r["launches"], "ms_per_step": r["ms"],
                   "share_of_step": r["ms"] / (st_prof["t_total_ns"] / 1e6) if st_prof["t_total_ns"] else None}
            if r["bytes"]:
                gbs = r["bytes"] / (r["ms"] / 1e3) / 1e9
                row.update(bound="hbm", achieved=gbs, peak=peak, unit="GB/s", frac=gbs / peak, algorithmic_bytes=r["bytes"])
            else:
                row.update(bound="latency (dependent loads)", achieved=None, frac=None)
            out["other_kernels"].append(row)
        # `roofline` names the kernel that is the largest share of the step among those a byte roofline applies to (the parse
        # rounds are a larger share but wait on dependent loads: no byte count describes them); K1+K2, the kernel north_star
        # names, keeps its line as `roofline_k1k2`
        hbm_rows = [r for r in out["other_kernels"] if r.get("frac") is not None and r.get("ms_per_step")]
        if hbm_rows:
            top = max(hbm_rows, key=lambda r: r["ms_per_step"])
            out["roofline_k1k2"] = out["roofline"]
            trf = scatter_traffic() if top["kernel"].startswith("rs_scatter_kernel<u32>") else None
            out["roofline"] = {"kernel": top["kernel"], "bound": "hbm", "achieved": top["achieved"], "peak": peak, "unit": "GB/s",
                               "frac": top["frac"], "peak_source": out["roofline_k1k2"]["peak_source"],
                               "algorithmic_bytes": int(top["algorithmic_bytes"] / top["launches_per_step"]),
                               "kernel_ms": top["ms_per_step"] / top["launches_per_step"], "launches_per_step": top["launches_per_step"],
                               "share_of_step": top["share_of_step"],
                               "traffic": trf["dram_bytes"] if trf else None,
                               "traffic_note": (f"ncu --set full of one launch with {trf['algorithmic_bytes']} algorithmic bytes "
                                                f"({trf['source']})") if trf else None}
        out["index"] = {k: int(st_res[k]) for k in ("parse_slices", "sparse_slices", "sparse_fallbacks", "repair_entries",
                                                    "parse_rounds", "gap_positions_hashed", "prefetched_slices")}
        if not args.no_cpu_baseline:
            # the reference on the same text, tau and num_threads: its time (cpu_baseline) and, outside every timed
            # region, its factor stream against the CUDA path's for the bases the reference drew (parity_checked)
            sample = text[: args.cpu_sample_mib * MIB]
            r = cpu_reference_run(sample, args.tau, args.threads, keep=True)
            what = "the whole text" if sample.size == n else f"first {sample.size // MIB} MiB of the workload text"
            out["cpu_baseline"] = {"value": (sample.size / MIB) / r["sec"], "unit": "MiB/s", "cores": r["cores"], "kind": r["kind"],
                                   "sample": f"{what}, num_threads={args.threads}, one run",
                                   "phases_ms": r["phases_ms"], "shim_share": shim_share(r["phases_ms"])}
            with capi.Factorizer(tau=args.tau, num_threads=args.threads, gap_bases=r["bases"], device=local) as F2:
                gsrc, gln = F2.factorize(sample, cap=cap)
            ok = int(gsrc.size) == r["z"] and stream_hash(gsrc, gln) == r["hash"]
            out["parity_checked"] = bool(ok)
            out["parity"] = {"against": f"oracle/_ref ({r['kind']}) on {what}", "factors": int(gsrc.size), "reference_factors": r["z"],
                             "sha256_equal": bool(ok)}
            if not ok:
                print(json.dumps(out))
                raise SystemExit("bench.py: the factor stream differs from the reference's")
    barrier()
    if rank == 0:
        print(json.dumps(out))
    F.free_device_text(d_text)
    F.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
