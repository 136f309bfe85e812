"""CPU tests of the drop-in boundary: the C-ABI library builds for sm_100a, loads, exports every symbol the header
declares, validates arguments, and fails loudly (no CPU fallback) when no CUDA device is present."""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_exports_match_header(capi):
    hdr = open(os.path.join(ROOT, "include", "lz77sss_b200.h")).read()
    declared = set(re.findall(r"\b(lz77sss_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"lz77sss_emit_fn"}
    assert declared == set(capi.EXPORTS)
    L = capi.lib()
    for name in declared:
        assert getattr(L, name) is not None
    out = subprocess.run(["nm", "-D", "--defined-only", capi.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (lz77sss_[a-z0-9_]+)", out))
    assert exported == declared
    assert b"sm_100a" in L.lz77sss_version()


def test_struct_layout(capi):
    import ctypes as C
    assert C.sizeof(capi.Config) == 4 * 8 + 8 * 5 + 4 + 4
    assert C.sizeof(capi.Stats) == 8 * 39


def test_argument_validation(capi):
    import ctypes as C
    L = capi.lib()
    h = C.c_void_p()
    cfg = capi.Config()
    cfg.struct_size = C.sizeof(capi.Config)
    cfg.tau = 100                              # not a power of two
    for i in range(5):
        cfg.gap_bases[i] = 1000 + i
    assert L.lz77sss_create(C.byref(cfg), C.byref(h)) == -1
    assert b"tau" in L.lz77sss_last_error()
    cfg.tau = 512
    cfg.gap_bases[2] = 5                       # below 257
    assert L.lz77sss_create(C.byref(cfg), C.byref(h)) == -1
    cfg.gap_bases[2] = 1 << 20                 # above 2^20-1
    assert L.lz77sss_create(C.byref(cfg), C.byref(h)) == -1
    cfg.struct_size = 8
    assert L.lz77sss_create(C.byref(cfg), C.byref(h)) == -1


def test_no_cpu_fallback(capi):
    if capi.lib().lz77sss_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(capi.Lz77SssError) as e:
        capi.Factorizer(tau=512)
    assert e.value.code == -2


def test_product_does_not_touch_the_oracle():
    """The product path (package + csrc + headers + example) never imports, links or names oracle/."""
    bad = []
    for base in ("lz77_sss_b200", "include", "examples"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            if "_obj" in dp or "__pycache__" in dp:
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                    s = open(os.path.join(dp, f), errors="ignore").read()
                    if re.search(r"oracle_lib|liboracle|oracle/_|#include\s*[\"<].*oracle", s):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_cpp_header_compiles_against_the_abi(capi, tmp_path):
    exe = tmp_path / "example"
    cmd = ["/usr/bin/g++", "-std=c++20", "-O1", "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "lz77_sss_b200", "include"), os.path.join(ROOT, "examples", "lz77_sss_example.cpp"),
           "-L" + os.path.join(ROOT, "lz77_sss_b200"), "-llz77sss_b200", "-o", str(exe)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_synthetic_workloads_are_deterministic():
    from lz77_sss_b200 import synth
    a = synth.dna_collection(1 << 16, seed=2, chrom_len=1 << 12)
    b = synth.dna_collection(1 << 16, seed=2, chrom_len=1 << 12)
    assert (a == b).all() and set(np.unique(a)) <= set(b"ACGT")
    r = synth.random_repetitive(50_000, seed=1)
    assert r.size == 50_000 and (r == synth.random_repetitive(50_000, seed=1)).all()
