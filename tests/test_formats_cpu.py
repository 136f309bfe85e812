"""CPU tests of the file formats: the B200 build's host headers
lz77_sss_b200/include/lz77_sss/misc/{huffman,vbyte}.hpp against files written by the reference's own headers
(committed fixtures from tests/golden/make_golden_formats.py, and live against oracle/_ref where it exists)."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FIX = os.path.join(ROOT, "tests", "golden", "formats")
NAMES = sorted(f[:-5] for f in os.listdir(FIX) if f.endswith(".lz77"))


@pytest.fixture(scope="session")
def tool(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("fmt") / "factor_file")
    subprocess.run(["/usr/bin/g++", "-std=c++20", "-O2", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "lz77_sss_b200", "include"),
                    os.path.join(ROOT, "tools", "cpp", "factor_file.cpp"), "-o", exe], check=True)
    return exe


def _read(path):
    with open(path, "rb") as f:
        return f.read()


@pytest.mark.parametrize("name", NAMES)
def test_huffman_file_is_byte_identical_to_the_reference(tool, tmp_path, name):
    n = int(_read(os.path.join(FIX, name + ".n")))
    out = tmp_path / "mine.lz77"
    subprocess.run([tool, "huff-write", os.path.join(FIX, name + ".factors.bin"), str(n), str(out)], check=True)
    assert _read(out) == _read(os.path.join(FIX, name + ".lz77"))


@pytest.mark.parametrize("name", NAMES)
def test_huffman_file_of_the_reference_is_read_back(tool, tmp_path, name):
    n = int(_read(os.path.join(FIX, name + ".n")))
    out = tmp_path / "factors.bin"
    r = subprocess.run([tool, "huff-read", os.path.join(FIX, name + ".lz77"), str(out)], check=True, capture_output=True, text=True)
    want = np.fromfile(os.path.join(FIX, name + ".factors.bin"), dtype=np.uint64)
    got = np.fromfile(out, dtype=np.uint64)
    assert r.stdout.split() == [str(n), str(want.size // 2)]
    assert got.size == want.size and (got == want).all()


@pytest.mark.parametrize("bits", [32, 64])
def test_vbyte_matches_the_reference(tool, tmp_path, bits):
    vals = os.path.join(FIX, f"vbyte_values{bits}.bin")
    enc, dec = tmp_path / "enc", tmp_path / "dec"
    subprocess.run([tool, "vbyte-write", vals, str(bits), str(enc)], check=True)
    assert _read(enc) == _read(os.path.join(FIX, f"vbyte{bits}.bin"))
    subprocess.run([tool, "vbyte-read", os.path.join(FIX, f"vbyte{bits}.bin"), str(bits), str(dec)], check=True)
    assert _read(dec) == _read(vals)


def test_huffman_live_against_the_reference_build(tool, tmp_path, oracle):
    """Fresh random streams (many block shapes) where oracle/_ref exists: both directions."""
    if not oracle.ref_available(512):
        if oracle.ref_require():
            pytest.fail("oracle/_ref missing and LZ77SSS_REQUIRE_REF=1")
        pytest.skip("oracle/_ref was not prebuilt")
    rng = np.random.default_rng(99)
    for it in range(12):
        z = int(rng.integers(1, 40_000))
        src, ln, pos = np.zeros(z, np.uint64), np.zeros(z, np.uint64), 0
        maxexp = int(rng.integers(1, 30))
        lit = rng.random()
        for k in range(z):
            if pos == 0 or rng.random() < lit:
                src[k] = rng.integers(0, 256); pos += 1
            else:
                l = int(rng.integers(1, 2 ** int(rng.integers(1, maxexp + 1)) + 1))
                src[k] = rng.integers(0, pos); ln[k] = l; pos += l
        pairs = np.empty(2 * z, np.uint64)
        pairs[0::2], pairs[1::2] = src, ln
        fb, mine, theirs, back = tmp_path / "f.bin", tmp_path / "mine", tmp_path / "theirs", tmp_path / "back.bin"
        pairs.tofile(fb)
        subprocess.run([tool, "huff-write", str(fb), str(pos), str(mine)], check=True)
        oracle.ref_huff_write(src, ln, pos, theirs)
        assert _read(mine) == _read(theirs), it
        s2, l2, n2 = oracle.ref_huff_read(mine, z + 1)
        assert n2 == pos and (s2 == src).all() and (l2 == ln).all()
        subprocess.run([tool, "huff-read", str(theirs), str(back)], check=True, capture_output=True)
        assert (np.fromfile(back, dtype=np.uint64) == pairs).all()
