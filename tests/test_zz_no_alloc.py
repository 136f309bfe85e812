"""lz77sss_rks_table_noalloc: the prefix-fingerprint table of rabin_karp_substring<31> built without any allocation (the
kernel's look-back scratch is a static device array).  Checked against lz77sss_rks_table, bit for bit (integer work), on two
calls in a row: the second call finds the first call's contents in the shared scratch.

Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from lz77_sss_b200 import synth

pytestmark = pytest.mark.gpu


def _noalloc_table(capi, torch, text, base, check_free=False):
    n = text.size
    dev = torch.device("cuda", 0)
    d_text = torch.zeros((n + 15) // 16 * 16 + 16, dtype=torch.uint8, device=dev)       # zero bytes behind the text
    d_text[:n].copy_(torch.from_numpy(text))
    d_out = torch.full(((n + 15) // 16 + 4,), 0xDEAD, dtype=torch.int32, device=dev)
    torch.cuda.synchronize(dev)
    before = torch.cuda.memory_allocated(dev)
    free_before = torch.cuda.mem_get_info(dev)[0]       # the driver's count: sees a cudaMalloc of the library too
    capi._check(capi.lib().lz77sss_rks_table_noalloc(C.c_void_p(d_text.data_ptr()), n, int(base), C.c_void_p(d_out.data_ptr()), None))
    torch.cuda.synchronize(dev)
    assert torch.cuda.memory_allocated(dev) == before
    if check_free:      # (not on a first call: loading the kernel's module may take device memory once)
        assert torch.cuda.mem_get_info(dev)[0] == free_before, "device memory was allocated during the call"
    out = d_out.cpu().numpy().view(np.uint32)
    g = (n + 15) // 16
    assert (out[g:] == 0xDEAD).all(), "wrote behind the table"
    return out[:g]


def test_noalloc_table_equals_the_allocating_entry_point_twice(capi):
    import torch
    if capi.lib().lz77sss_device_count() < 1:
        pytest.fail("needs a CUDA device: there is no CPU fallback")
    with capi.Factorizer(tau=64, num_threads=1, gap_bases=capi.DEFAULT_GAP_BASES) as F:
        # two calls in a row with different texts, lengths (ragged last granule, several tiles) and bases
        for n, seed, base in ((1_000_003, 5, 257), (300_001, 6, 0x7FFFFFFE), (16384 * 70 + 5, 7, 1_234_567), (15, 8, 999_983)):
            text = synth.random_repetitive(n, seed=seed)
            with capi.RabinKarpSubstring(F, text, base) as R:
                want = R.table()
            got = _noalloc_table(capi, torch, text, base)
            assert got.size == want.size and (got == want).all(), (n, base)
            again = _noalloc_table(capi, torch, text, base, check_free=True)
            assert (again == want).all(), (n, base, "second call")


def test_noalloc_table_rejects_what_it_cannot_hold(capi):
    import torch
    lib = capi.lib()
    d = torch.zeros(64, dtype=torch.uint8, device="cuda:0")
    p = C.c_void_p(d.data_ptr())
    assert lib.lz77sss_rks_table_noalloc(p, (1 << 30) + 1, 257, p, None) == -3          # LZ77SSS_E_UNSUPPORTED
    assert b"LZ77SSS_RKS_NOALLOC_MAX_BYTES" in lib.lz77sss_last_error()
    assert lib.lz77sss_rks_table_noalloc(p, 32, 256, p, None) == -1                       # base out of range
    assert lib.lz77sss_rks_table_noalloc(C.c_void_p(d.data_ptr() + 1), 32, 257, p, None) == -1     # misaligned
    assert lib.lz77sss_rks_table_noalloc(p, 0, 257, p, None) == 0                                  # empty text: nothing to do
