"""GPU tests (-m gpu) of hs_gpu_scan_vector on HS_MODE_VECTORED pure-literal databases: the buffers
are consecutive blocks of one stream (src/runtime.c:1106-1174) -- a literal may straddle any number of blocks, offsets run
on, single-match ids stay exhausted to the end of the call.  Expected values: the brute force on the concatenation and
the oracle port's block-by-block restatement (tests/test_vectored.py pins both on the unmodified reference runtime)."""
import random

import numpy as np
import pytest

import litgen
import oracle_port
from test_vectored import HS_MODE_VECTORED, cases, cut
from vectorscan_b200 import gpu, litdb

pytestmark = pytest.mark.gpu


def gpu_vectored(db, bufs, frontend=0):
    sc = gpu.alloc_scratch(db)
    sc.set_option("frontend", frontend)
    got = [(int(b), int(i), int(t)) for b, i, t in gpu.scan_vector_ex(db, bufs, sc)]
    sc.close()
    assert all(b == 0 for b, _, _ in got)      # one stream: every match belongs to logical buffer 0
    return [(i, t) for _, i, t in got]


@pytest.mark.parametrize("frontend", [0, 1, 2, 3])
def test_vectored_scan_equals_brute_force_and_port(frontend):
    rng = random.Random(40 + frontend)
    total = 0
    for lits, flags, ids, text, pieces in cases(3 + frontend):
        vdb = litdb.compile_lit_multi(lits, flags, ids, mode=HS_MODE_VECTORED)
        bufs = cut(rng, text, pieces)
        got = gpu_vectored(vdb, bufs, frontend)
        assert sorted(got) == litgen.brute_force(lits, flags, ids, text), (len(lits), len(text), pieces)
        prv, pgot = oracle_port.scan_vector(vdb, bufs)
        assert prv == 0 and sorted(pgot) == sorted(got)
        assert got == sorted(got, key=lambda m: (m[1], m[0]))     # delivered in (to, id) order
        total += len(got)
    assert total > 1000


def test_vectored_callbacks_entry_points_and_modes():
    lits = [b"needle", b"hay", b"zz"]
    flags = [0, 0, 8]
    vdb = litdb.compile_lit_multi(lits, flags, [1, 2, 3], mode=HS_MODE_VECTORED)
    bdb = litdb.compile_lit_multi(lits, flags, [1, 2, 3])
    bufs = [b"a nee", b"dle in the h", b"", b"ayzz", b"zzz"]
    sc = gpu.alloc_scratch(vdb)
    seen = []
    assert gpu.scan_vector(vdb, bufs, sc, lambda i, f, t, fl, ctx: seen.append((i, t)) or 0) == gpu.HS_SUCCESS
    assert seen == [(1, 8), (2, 19), (3, 21)]
    # the callback stops the scan: HS_SCAN_TERMINATED
    seen = []
    assert gpu.scan_vector(vdb, bufs, sc, lambda i, f, t, fl, ctx: seen.append((i, t)) or 1) == gpu.HS_SCAN_TERMINATED
    assert seen == [(1, 8)]
    # hs_scan on a vectored database: HS_DB_MODE_ERROR, as in the reference (src/runtime.c:335-337)
    assert gpu.scan(vdb, b"needle", sc) == gpu.HS_DB_MODE_ERROR
    # a resident corpus is one stream for a vectored database, independent buffers for a block-mode one
    flat = b"".join(bufs)
    offs = np.cumsum([0] + [len(b) for b in bufs]).astype(np.uint64)
    corpus = gpu.Corpus(flat, offs)
    assert gpu.scan_corpus(vdb, corpus, sc) == 3
    assert [(int(b), int(i), int(t)) for b, i, t in gpu.fetch_matches(sc)] == [(0, 1, 8), (0, 2, 19), (0, 3, 21)]
    sc.grow(bdb)
    assert gpu.scan_corpus(bdb, corpus, sc) == 2
    assert [(int(b), int(i), int(t)) for b, i, t in gpu.fetch_matches(sc)] == [(3, 3, 4), (4, 3, 2)]
    assert gpu.scan_corpus(vdb, corpus, sc) == 3     # and again: the stream view is kept with the corpus
    sc.close()


def test_vectored_large_stream_of_packets():
    """16 MiB of 1500-byte packets as ONE stream (the pcapscan shape with cross-packet matches) against the same text as one
    block-mode buffer, and against planted literals that straddle packet boundaries."""
    rng = np.random.default_rng(9)
    n_lits = 2000
    lits = sorted({bytes(rng.integers(97, 123, size=int(rng.integers(4, 12))).astype(np.uint8)) for _ in range(n_lits)})
    ids = list(range(len(lits)))
    total = 16 << 20
    text = rng.integers(97, 123, size=total).astype(np.uint8)
    planted = []
    for pk in rng.choice(np.arange(1, total // 1500 - 1), size=3000, replace=False):   # distinct packet boundaries
        lit = lits[int(rng.integers(0, len(lits)))]
        p = 1500 * int(pk) - int(rng.integers(1, len(lit)))   # straddles the boundary
        text[p:p + len(lit)] = np.frombuffer(lit, dtype=np.uint8)
        planted.append(p + len(lit))
    data = text.tobytes()
    bufs = [data[o:o + 1500] for o in range(0, total, 1500)]
    vdb = litdb.compile_lit_multi(lits, [0] * len(lits), ids, mode=HS_MODE_VECTORED)
    bdb = litdb.compile_lit_multi(lits, [0] * len(lits), ids)
    sc = gpu.alloc_scratch(vdb)
    got = [(int(i), int(t)) for _, i, t in gpu.scan_vector_ex(vdb, bufs, sc)]
    sc.grow(bdb)
    whole = [(int(i), int(t)) for _, i, t in gpu.scan_vector_ex(bdb, [data], sc)]
    sc.close()
    assert got == whole and len(got) >= 2500
    ends = {t for _, t in got}
    assert all(e in ends for e in planted)
    prv, pgot = oracle_port.scan_vector(vdb, bufs[:700])
    n_prefix = sum(len(b) for b in bufs[:700])
    assert prv == 0 and sorted(pgot) == sorted(m for m in got if m[1] <= n_prefix)


def test_reference_named_hs_scan_vector_through_the_shim():
    """integration/hs_gpu_shim.c: hs_scan_vector by the reference's own name and prototype (src/hs_runtime.h:522-527)."""
    import ctypes as C
    import os
    from vectorscan_b200.gpu import MATCH_CB
    shim = C.CDLL(os.path.join(os.path.dirname(gpu.__file__), "libhs_gpushim.so"))
    shim.hs_scan_vector.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.POINTER(C.c_uint), C.c_uint, C.c_uint, C.c_void_p,
                                    MATCH_CB, C.c_void_p]
    shim.hs_alloc_scratch.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
    shim.hs_free_scratch.argtypes = [C.c_void_p]
    lits = [b"straddle", b"dd"]
    vdb = litdb.compile_lit_multi(lits, [0, 0], [10, 11], mode=HS_MODE_VECTORED)
    bdb = litdb.compile_lit_multi(lits, [0, 0], [10, 11])
    bufs = [b"xxstr", b"ad", b"dle", b"", b"d", b"d"]
    arr = (C.c_char_p * len(bufs))(*bufs)
    lens = (C.c_uint * len(bufs))(*[len(b) for b in bufs])
    sc = C.c_void_p()
    assert shim.hs_alloc_scratch(vdb.ptr, C.byref(sc)) == 0
    seen = []
    cb = MATCH_CB(lambda i, f, t, fl, ctx: seen.append((i, t)) or 0)
    assert shim.hs_scan_vector(vdb.ptr, arr, lens, len(bufs), 0, sc, cb, None) == 0
    assert seen == [(11, 8), (10, 10), (11, 12)]
    assert shim.hs_alloc_scratch(bdb.ptr, C.byref(sc)) == 0
    assert shim.hs_scan_vector(bdb.ptr, arr, lens, len(bufs), 0, sc, cb, None) == gpu.HS_DB_MODE_ERROR
    shim.hs_free_scratch(sc)
