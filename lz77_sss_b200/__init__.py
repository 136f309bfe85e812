"""lz77_sss_b200 — B200-native (sm_100a) implementation of the lz77-sss factorization hot path.

The product is `liblz77sss_b200.so` (hand-written CUDA + the C ABI of include/lz77sss_b200.h) and the C++
header `lz77_sss_b200/include/lz77_sss/lz77_sss.hpp` that keeps the reference's template API.
`capi` is the ctypes plumbing tests and bench use; `synth` generates the synthetic workloads.
"""
from . import capi, synth  # noqa: F401
from .capi import Factorizer, Lz77SssError  # noqa: F401
