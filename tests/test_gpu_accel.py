"""GPU tests: shufti / truffle / vermicelli byte-class searches against the reference's own routines
(oracle/_ref exports shuftiExec, rshuftiExec, truffleExec, rtruffleExec, vermicelliExec, ...) with the
vectors' shape of unit/internal/shufti.cpp, truffle.cpp, vermicelli.cpp: every start alignment, hit / no hit."""
import ctypes as C
import random

import pytest

from vectorscan_b200 import gpu

pytestmark = pytest.mark.gpu


def shufti_masks(chars):
    """One-bucket shufti masks for a small class (src/nfa/shufticompile.cpp:54-110 with a single bucket per
    distinct high nibble)."""
    lo, hi = bytearray(16), bytearray(16)
    his = sorted({c >> 4 for c in chars})
    assert len(his) <= 8
    for c in chars:
        bit = 1 << his.index(c >> 4)
        lo[c & 15] |= bit
        hi[c >> 4] |= bit
    return bytes(lo), bytes(hi)


def truffle_masks(chars):
    m1, m2 = bytearray(16), bytearray(16)
    for c in chars:
        (m2 if c & 0x80 else m1)[c & 15] |= 1 << ((c >> 4) & 7)
    return bytes(m1), bytes(m2)


def in_shufti(lo, hi, c):
    return (lo[c & 15] & hi[c >> 4]) != 0


def first(pred, buf):
    for i, c in enumerate(buf):
        if pred(c):
            return i
    return len(buf)


def last(pred, buf):
    for i in range(len(buf) - 1, -1, -1):
        if pred(buf[i]):
            return i
    return -1


def test_shufti_truffle_vermicelli_all_alignments(ref):
    L = gpu.lib()
    L.hs_gpu_shufti.restype = L.hs_gpu_truffle.restype = L.hs_gpu_vermicelli.restype = C.c_longlong
    L.hs_gpu_vermicelli_double.restype = C.c_longlong
    rng = random.Random(3)
    for trial in range(40):
        chars = set(rng.sample(range(256), rng.randint(1, 6)))
        while len({c >> 4 for c in chars}) > 8:
            chars.pop()
        lo, hi = shufti_masks(chars)
        m1, m2 = truffle_masks(chars)
        filler = [c for c in range(256) if not in_shufti(lo, hi, c) and c not in chars]
        n = rng.choice([1, 15, 16, 17, 64, 100, 1000, 70000])
        base = bytearray(rng.choice(filler) for _ in range(n + 40))
        for _ in range(rng.choice([0, 1, 3])):
            base[rng.randrange(len(base))] = rng.choice(sorted(chars))
        for align in range(0, 33, rng.choice([1, 3, 7])):
            buf = bytes(base[align:align + n])
            assert L.hs_gpu_shufti(lo, hi, buf, len(buf), 0) == first(lambda c: in_shufti(lo, hi, c), buf)
            assert L.hs_gpu_shufti(lo, hi, buf, len(buf), 1) == last(lambda c: in_shufti(lo, hi, c), buf)
            assert L.hs_gpu_truffle(m1, m2, buf, len(buf), 0) == first(lambda c: c in chars, buf)
            assert L.hs_gpu_truffle(m1, m2, buf, len(buf), 1) == last(lambda c: c in chars, buf)
            ch = sorted(chars)[0]
            assert L.hs_gpu_vermicelli(ch, 0, buf, len(buf), 0, 0) == first(lambda c: c == ch, buf)
            assert L.hs_gpu_vermicelli(ch, 0, buf, len(buf), 1, 0) == last(lambda c: c == ch, buf)
            assert L.hs_gpu_vermicelli(ch, 0, buf, len(buf), 0, 1) == first(lambda c: c != ch, buf)
            # against the reference's own SIMD routines (need >= 1 byte)
            b = C.create_string_buffer(buf, len(buf))
            p0 = C.addressof(b)
            lo_t, hi_t = (C.c_ubyte * 16)(*lo), (C.c_ubyte * 16)(*hi)
            ref.L.ref_shim_shufti.restype = C.c_longlong
            ref.L.ref_shim_shufti.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
            ref.L.ref_shim_truffle.restype = C.c_longlong
            ref.L.ref_shim_truffle.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
            for rev in (0, 1):
                assert L.hs_gpu_shufti(lo, hi, buf, len(buf), rev) == ref.L.ref_shim_shufti(lo_t, hi_t, p0, len(buf), rev)
                m1_t, m2_t = (C.c_ubyte * 16)(*m1), (C.c_ubyte * 16)(*m2)
                got_t, ref_t = L.hs_gpu_truffle(m1, m2, buf, len(buf), rev), ref.L.ref_shim_truffle(m1_t, m2_t, p0, len(buf), rev)
                if rev and 0 in chars and len(buf) < 64 and got_t != ref_t:
                    # reference defect in its 256/512-bit builds (src/nfa/truffle_simd.hpp:184-195): a reverse scan of a
                    # buffer shorter than the vector zero-pads it, NUL is in the class, so the padding "matches" past the
                    # end and the routine answers "not found".  The scalar definition (checked above) stands.
                    assert ref_t == -1
                    continue
                assert got_t == ref_t


def test_vermicelli_double_and_nocase(ref):
    # the reference's 512-bit forward vermicelli misses a match on the very last byte of the buffer (seen with n = 200
    # and 5000 in this very loop; scalar definition src/nfa/vermicelli.hpp): the comparison uses its 256-bit build
    import refrt
    ref = refrt.RefRuntime("avx2")
    L = gpu.lib()
    L.hs_gpu_vermicelli.restype = L.hs_gpu_vermicelli_double.restype = C.c_longlong
    ref.L.ref_shim_verm.restype = C.c_longlong
    ref.L.ref_shim_verm.argtypes = [C.c_ubyte, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int]
    ref.L.ref_shim_dverm.restype = C.c_longlong
    ref.L.ref_shim_dverm.argtypes = [C.c_ubyte, C.c_ubyte, C.c_int, C.c_void_p, C.c_size_t]
    rng = random.Random(9)
    for trial in range(60):
        n = rng.choice([1, 2, 16, 31, 33, 200, 5000])
        buf = bytearray(rng.choice(b"xyzXYZ.") for _ in range(n))
        if rng.random() < 0.7 and n >= 2:
            p = rng.randrange(n - 1)
            buf[p:p + 2] = rng.choice([b"ab", b"Ab", b"aB", b"AB"])
        if rng.random() < 0.3:
            buf[-1] = ord("a")
        buf = bytes(buf)
        b = C.create_string_buffer(buf, len(buf))
        for nocase in (0, 1):
            c1, c2 = (ord("A"), ord("B")) if nocase else (ord("a"), ord("b"))
            for rev in (0, 1):
                for neg in (0, 1):
                    assert L.hs_gpu_vermicelli(c1, nocase, buf, n, rev, neg) == \
                        ref.L.ref_shim_verm(c1, nocase, C.addressof(b), n, rev, neg), (buf, nocase, rev, neg)
            assert L.hs_gpu_vermicelli_double(c1, c2, nocase, buf, n) == ref.L.ref_shim_dverm(c1, c2, nocase, C.addressof(b), n)


def dshufti_masks(pairs):
    """Double-shufti masks for up to 8 byte pairs, one bucket per pair (src/nfa/shufticompile.cpp:112-200): bit k of
    lo1/hi1 (lo2/hi2) is CLEAR where pair k's first (second) byte has that nibble."""
    lo1, hi1, lo2, hi2 = (bytearray([0xff] * 16) for _ in range(4))
    for k, (a, b) in enumerate(pairs):
        lo1[a & 15] &= ~(1 << k) & 0xff
        hi1[a >> 4] &= ~(1 << k) & 0xff
        lo2[b & 15] &= ~(1 << k) & 0xff
        hi2[b >> 4] &= ~(1 << k) & 0xff
    return bytes(lo1), bytes(hi1), bytes(lo2), bytes(hi2)


def test_shufti_double(ref):
    """shuftiDoubleExec: first position whose byte pair is one of the listed pairs (a trailing first byte counts).  The
    filler never contains a first byte of any pair, so the reference's conservative answers at the ends of its vector
    lanes (x86/shufti.hpp:69-70 shifts zeros in) cannot differ from the scalar definition."""
    L = gpu.lib()
    L.hs_gpu_shufti_double.restype = C.c_longlong
    ref.L.ref_shim_dshufti.restype = C.c_longlong
    ref.L.ref_shim_dshufti.argtypes = [C.c_void_p] * 5 + [C.c_size_t]
    rng = random.Random(17)
    for trial in range(60):
        pairs = [(rng.randrange(97, 110), rng.randrange(32, 127)) for _ in range(rng.randint(1, 8))]
        lo1, hi1, lo2, hi2 = dshufti_masks(pairs)

        def m1(c):
            return (lo1[c & 15] | hi1[c >> 4]) & 0xff

        def m2(c):
            return (lo2[c & 15] | hi2[c >> 4]) & 0xff
        filler = [c for c in range(32, 127) if m1(c) == 0xff]
        n = rng.choice([16, 17, 64, 100, 1000, 70000])
        buf = bytearray(rng.choice(filler) for _ in range(n))
        for _ in range(rng.choice([0, 1, 2])):
            a, b = rng.choice(pairs)
            p = rng.randrange(n - 1)
            buf[p:p + 2] = bytes([a, b])
        if rng.random() < 0.3:
            buf[-1] = pairs[0][0]
        buf = bytes(buf)
        exp = n
        for i in range(n):
            if (m1(buf[i]) | (m2(buf[i + 1]) if i + 1 < n else 0)) != 0xff:
                exp = i
                break
        got = L.hs_gpu_shufti_double(lo1, hi1, lo2, hi2, buf, n)
        assert got == exp, (pairs, n)
        b = C.create_string_buffer(buf, n)
        t = [(C.c_ubyte * 16)(*m) for m in (lo1, hi1, lo2, hi2)]
        assert ref.L.ref_shim_dshufti(t[0], t[1], t[2], t[3], C.addressof(b), n) == exp
