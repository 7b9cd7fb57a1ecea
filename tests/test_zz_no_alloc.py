"""GPU test (-m gpu) of hs_gpu_shufti_noalloc (include/hs_gpu.h): the byte-class search with its scratch in a fixed file-scope
device array instead of cudaMalloc'ed memory.  Collected LAST on purpose.  Checked through
the C ABI against the existing entry point hs_gpu_shufti, bit-exact (offsets are integers), on calls in a row -- the shared
array keeps the previous call's contents, so a longer search followed by a shorter one and a forward one followed by a
reverse one are what would show stale scratch."""
import ctypes as C
import random

import pytest

from vectorscan_b200 import gpu

pytestmark = pytest.mark.gpu
MAX_BYTES = 1 << 20          # HS_GPU_NOALLOC_MAX_BYTES
HS_INSUFFICIENT_SPACE = -12


def masks_of(chars):
    lo, hi = bytearray(16), bytearray(16)
    his = sorted({c >> 4 for c in chars})
    for c in chars:
        bit = 1 << his.index(c >> 4)
        lo[c & 15] |= bit
        hi[c >> 4] |= bit
    return bytes(lo), bytes(hi)


def lib():
    L = gpu.lib()
    L.hs_gpu_shufti.restype = C.c_longlong
    L.hs_gpu_shufti.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t, C.c_int]
    L.hs_gpu_shufti_noalloc.restype = C.c_int
    L.hs_gpu_shufti_noalloc.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t, C.c_int, C.POINTER(C.c_longlong)]
    return L


def noalloc(L, lo, hi, buf, reverse):
    out = C.c_longlong(-12345)
    rv = L.hs_gpu_shufti_noalloc(lo, hi, buf, len(buf), reverse, C.byref(out))
    return rv, out.value


def test_no_alloc_shufti_equals_the_allocating_entry_point_call_after_call():
    L = lib()
    rng = random.Random(11)
    lo, hi = masks_of({ord("X"), ord("q"), 0x00, 0xf3})     # 0x00 is a member: the zero padding must not count as a hit
    fill = bytes(c for c in range(1, 256) if c not in (ord("X"), ord("q"), 0xf3))
    # two calls in a row on the same data, then sizes going up and down, hits anywhere / nowhere, both directions
    base = bytes(rng.choice(fill) for _ in range(70001))
    hit = bytearray(base)
    hit[40000] = ord("X")
    hit[65000] = 0x00
    for reverse in (0, 1):
        want = L.hs_gpu_shufti(lo, hi, bytes(hit), len(hit), reverse)
        assert want == (65000 if reverse else 40000)
        assert noalloc(L, lo, hi, bytes(hit), reverse) == (0, want)
        assert noalloc(L, lo, hi, bytes(hit), reverse) == (0, want)       # again: same answer from the reused scratch
    checked = 0
    for n in (MAX_BYTES, 1, 17, 4095, 300000, 16, 2048, 2049, 33, MAX_BYTES - 5, 100):
        data = bytearray(rng.choice(fill) for _ in range(min(n, 4096))) * (n // min(n, 4096) + 1)
        data = data[:n]
        for k in range(rng.choice([0, 0, 1, 3])):
            data[rng.randrange(n)] = rng.choice([ord("X"), ord("q"), 0x00, 0xf3])
        data = bytes(data)
        for reverse in (0, 1):
            want = L.hs_gpu_shufti(lo, hi, data, n, reverse)
            assert noalloc(L, lo, hi, data, reverse) == (0, want), (n, reverse)
            checked += 1
    assert checked == 22
    # other masks right after: the class table in the scratch is rewritten by every call
    lo2, hi2 = masks_of({ord("Z")})
    data = b"abcXdefZghi" * 1000
    assert noalloc(L, lo2, hi2, data, 0) == (0, L.hs_gpu_shufti(lo2, hi2, data, len(data), 0)) == (0, 7)
    assert noalloc(L, lo, hi, data, 1) == (0, L.hs_gpu_shufti(lo, hi, data, len(data), 1))


def test_no_alloc_limits_and_arguments():
    L = lib()
    lo, hi = masks_of({ord("X")})
    assert noalloc(L, lo, hi, b"", 0) == (0, 0) and noalloc(L, lo, hi, b"", 1) == (0, -1)
    big = bytes(MAX_BYTES + 1)
    rv, out = noalloc(L, lo, hi, big, 0)
    assert rv == HS_INSUFFICIENT_SPACE and out == -12345          # rejected, result untouched
    assert L.hs_gpu_shufti_noalloc(lo, hi, b"abc", 3, 0, None) == gpu.HS_INVALID
    assert L.hs_gpu_shufti_noalloc(None, hi, b"abc", 3, 0, C.byref(C.c_longlong())) == gpu.HS_INVALID
    # and the entry point still works after a rejected call
    assert noalloc(L, lo, hi, b"..X..", 0) == (0, 2)
