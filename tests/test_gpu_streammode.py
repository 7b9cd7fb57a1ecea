"""GPU tests (-m gpu) of stream mode on pure-literal HS_MODE_STREAM databases: hs_gpu_open_stream /
hs_gpu_scan_stream / hs_gpu_close_stream / reset / copy, and the batch form hs_gpu_scan_streams (one block for each of
many streams per device pass).  Expected values: the brute force on each stream's concatenation and the oracle port's
block-by-block restatement (tests/test_streammode.py pins both on the unmodified reference runtime's stream API)."""
import random

import pytest

import litgen
import oracle_port
from test_vectored import cases, cut
from vectorscan_b200 import gpu, litdb

pytestmark = pytest.mark.gpu
HS_MODE_STREAM = litdb.HS_MODE_STREAM


def collect(out):
    return lambda i, f, t, fl, ctx: out.append((int(i), int(t))) or 0


@pytest.mark.parametrize("frontend", [0, 1, 2, 3])
def test_one_stream_block_by_block(frontend):
    rng = random.Random(60 + frontend)
    total = 0
    for lits, flags, ids, text, pieces in cases(7 + frontend):
        sdb = litdb.compile_lit_multi(lits, flags, ids, mode=HS_MODE_STREAM)
        bufs = cut(rng, text, min(pieces, 16))
        sc = gpu.alloc_scratch(sdb)
        sc.set_option("frontend", frontend)
        st = gpu.Stream(sdb)
        got = []
        for b in bufs:
            assert st.scan(b, sc, collect(got)) == gpu.HS_SUCCESS
        st.close()
        sc.close()
        assert sorted(got) == litgen.brute_force(lits, flags, ids, text), (len(lits), len(text), pieces)
        prv, pgot = oracle_port.scan_stream_blocks(sdb, bufs)
        assert prv == 0 and sorted(pgot) == sorted(got)
        assert [t for _, t in got] == sorted(t for _, t in got)        # offsets never go back
        total += len(got)
    assert total > 1000


def test_many_streams_per_device_pass():
    """120 flows fed round by round, one packet of every live flow per hs_gpu_scan_streams call (pcapscan's stream mode)."""
    rng = random.Random(17)
    lits, flags, ids, _ = litgen.random_case(rng, None, 300, bytes(range(97, 123)), 10)
    sdb = litdb.compile_lit_multi(lits, flags, ids, mode=HS_MODE_STREAM)
    flows = []
    for k in range(120):
        text = bytearray(rng.choice(b"abcdefghijklmnopqrstuvwxyz") for _ in range(rng.choice([5, 300, 2500])))
        for _ in range(len(text) // 40):
            lit = rng.choice(lits)
            if len(lit) <= len(text):
                p = rng.randrange(0, len(text) - len(lit) + 1)
                text[p:p + len(lit)] = lit
        flows.append(cut(rng, bytes(text), rng.choice([1, 2, 5, 9])))
    sc = gpu.alloc_scratch(sdb)
    streams = [gpu.Stream(sdb) for _ in flows]
    got = [[] for _ in flows]

    def on_event(i, f, t, fl, ctx):
        got[int(ctx or 0)].append((int(i), int(t)))
        return 0
    for rnd in range(9):
        live = [k for k, fl in enumerate(flows) if rnd < len(fl)]
        if not live:
            break
        rv = gpu.scan_streams([streams[k] for k in live], [flows[k][rnd] for k in live], sc, on_event, contexts=live)
        assert rv == gpu.HS_SUCCESS
    n = 0
    for k, fl in enumerate(flows):
        assert sorted(got[k]) == litgen.brute_force(lits, flags, ids, b"".join(fl)), k
        n += len(got[k])
    assert n > 300
    # the same stream twice in one call is refused (its blocks would not be consecutive)
    assert gpu.scan_streams([streams[0], streams[0]], [b"ab", b"cd"], sc) == gpu.HS_INVALID
    for s in streams:
        s.close()
    sc.close()


def test_stream_termination_reset_copy_and_modes():
    lits, flags, ids = [b"abc", b"bcd", b"zz"], [8, 0, 8], [5, 6, 7]
    sdb = litdb.compile_lit_multi(lits, flags, ids, mode=HS_MODE_STREAM)
    bdb = litdb.compile_lit_multi(lits, flags, ids)
    bufs = [b"xabcd", b"abcd", b"ab", b"cdzzz", b"zzabc"]
    want = oracle_port.scan_stream_blocks(sdb, bufs)[1]
    sc = gpu.alloc_scratch(sdb)
    st = gpu.Stream(sdb)
    got = []
    for b in bufs[:3]:
        assert st.scan(b, sc, collect(got)) == 0
    fork = st.copy()                       # hs_copy_stream: both continue from the same point
    got2 = list(got)
    for b in bufs[3:]:
        assert st.scan(b, sc, collect(got)) == 0
        assert fork.scan(b, sc, collect(got2)) == 0
    assert got == want and got2 == want
    assert [g for g in got if g[0] == 5] == [(5, 4)] and [g for g in got if g[0] == 7] == [(7, 15)]   # single-match ids once
    # reset: offsets and exhaustion start again
    assert st.reset() == 0
    again = []
    for b in bufs:
        assert st.scan(b, sc, collect(again)) == 0
    assert again == want
    # a callback that asks to stop terminates the stream for good
    assert st.reset() == 0
    part = []
    assert st.scan(bufs[0], sc, lambda i, f, t, fl, c: part.append((int(i), int(t))) or (1 if len(part) >= 2 else 0)) == gpu.HS_SCAN_TERMINATED
    assert part == want[:2]
    assert st.scan(bufs[1], sc, collect(part)) == gpu.HS_SCAN_TERMINATED and part == want[:2]
    # mode checks, as the reference's entry points make them
    with pytest.raises(gpu.HsGpuError):
        gpu.Stream(bdb)                                                  # hs_open_stream on a block database
    assert gpu.scan(sdb, b"abc", sc) == gpu.HS_DB_MODE_ERROR             # hs_scan on a stream database
    assert gpu.scan_vector(sdb, [b"abc"], sc) == gpu.HS_DB_MODE_ERROR
    for s in (st, fork):
        s.close()
    sc.close()


def test_stream_api_through_the_shim_by_the_reference_names():
    import ctypes as C
    import os
    from vectorscan_b200.gpu import MATCH_CB
    shim = C.CDLL(os.path.join(os.path.dirname(gpu.__file__), "libhs_gpushim.so"))
    vp = C.c_void_p
    shim.hs_alloc_scratch.argtypes = [vp, C.POINTER(vp)]
    shim.hs_free_scratch.argtypes = [vp]
    shim.hs_open_stream.argtypes = [vp, C.c_uint, C.POINTER(vp)]
    shim.hs_scan_stream.argtypes = [vp, C.c_char_p, C.c_uint, C.c_uint, vp, MATCH_CB, vp]
    shim.hs_close_stream.argtypes = [vp, vp, MATCH_CB, vp]
    sdb = litdb.compile_lit_multi([b"straddle", b"dd"], [0, 0], [10, 11], mode=HS_MODE_STREAM)
    sc, st = vp(), vp()
    assert shim.hs_alloc_scratch(sdb.ptr, C.byref(sc)) == 0
    assert shim.hs_open_stream(sdb.ptr, 0, C.byref(st)) == 0
    seen = []
    cb = MATCH_CB(lambda i, f, t, fl, ctx: seen.append((i, t)) or 0)
    for b in [b"xxstr", b"ad", b"dle", b"", b"d", b"d"]:
        assert shim.hs_scan_stream(st, b, len(b), 0, sc, cb, None) == 0
    assert shim.hs_close_stream(st, sc, cb, None) == 0
    assert seen == [(11, 8), (10, 10), (11, 12)]
    shim.hs_free_scratch(sc)


# ---- stream mode of ONE DFA engine: the engine's state id goes into and comes out of every block scan -------------------
def test_single_dfa_streams_block_by_block(ref):
    import regen
    from test_streammode import dfa_stream_cases
    n = 0
    for db, ex, fl, ids, bufs, text in dfa_stream_cases(77, eod=True):
        sc = gpu.alloc_scratch(db)
        st = gpu.Stream(db)
        got = []
        for b in bufs:
            assert st.scan(b, sc, collect(got)) == gpu.HS_SUCCESS
        assert st.close(sc, collect(got)) == gpu.HS_SUCCESS      # \z patterns report here, at the stream's length
        sc.close()
        assert sorted(got) == sorted(regen.brute_force(ex, fl, ids, text)), (ex, bufs)
        prv, pgot = oracle_port.scan_stream_blocks(db, bufs)     # callback order: by offset; within one offset the device
        assert prv == 0 and sorted(pgot) == sorted(got)          # delivers by id, the reference by engine accept order
        assert [t for _, t in got] == [t for _, t in pgot]
        n += len(got)
    assert n > 100


def test_many_single_dfa_streams_per_pass_and_refusals():
    import regen
    rng = random.Random(23)
    ex = [b"passwd\\s+[A-Za-z0-9_]{4,9}=", b"(login|admin)[^&]*&id=", b"User-Agent: [^\\r\\n]{8}", b"^GET /[^ ]*\\.php", b"abc"]
    fl, ids = [1, 0, 0, 0, 0], [3, 1, 4, 5, 9]
    db = litdb.compile_regex_multi(ex, fl, ids, engine=litdb.ENGINE_MCCLELLAN, mode=HS_MODE_STREAM)
    flows = []
    for k in range(200):
        text = regen.random_text(rng, ex, rng.choice([3, 200, 1500]))
        if k % 3 == 0:
            text = b"GET /a/b.php " + text
        flows.append(cut(rng, text, rng.choice([1, 2, 4, 7])))
    sc = gpu.alloc_scratch(db)
    streams = [gpu.Stream(db) for _ in flows]
    got = [[] for _ in flows]

    def on_event(i, f, t, fl_, ctx):
        got[int(ctx or 0)].append((int(i), int(t)))
        return 0
    for rnd in range(7):
        live = [k for k, fl_ in enumerate(flows) if rnd < len(fl_)]
        assert gpu.scan_streams([streams[k] for k in live], [flows[k][rnd] for k in live], sc, on_event, contexts=live) == 0
    n = 0
    for k, blocks in enumerate(flows):
        assert got[k] == sorted(regen.brute_force(ex, fl, ids, b"".join(blocks)), key=lambda m: (m[1], m[0])), k
        n += len(got[k])
    assert n > 50
    # anchored pattern: only at stream offset 0, and not after a reset point in the middle
    st = gpu.Stream(db)
    one = []
    for b in (b"GET /x", b".php GET /y.php"):
        assert st.scan(b, sc, collect(one)) == 0
    assert one == [(5, 10)]
    assert st.reset() == 0
    one = []
    assert st.scan(b"GET /z.php", sc, collect(one)) == 0 and one == [(5, 10)]
    st.close()
    for s in streams:
        s.close()
    sc.close()
    # end-of-data accepts come with hs_close_stream -- when it is given a handler -- and only if the stream ends there
    edb = litdb.compile_regex_multi([b"abc\\z", b"de"], [0, 0], [1, 2], engine=litdb.ENGINE_MCCLELLAN, mode=HS_MODE_STREAM)
    esc = gpu.alloc_scratch(edb)
    for blocks, want in (([b"xde ab", b"c"], [(2, 3), (1, 7)]), ([b"abc", b"de"], [(2, 5)]), ([b"abc"], [(1, 3)])):
        st, seen = gpu.Stream(edb), []
        for b in blocks:
            assert st.scan(b, esc, collect(seen)) == 0
        assert st.close(esc, collect(seen)) == 0 and seen == want, (blocks, seen)
    st = gpu.Stream(edb)
    assert st.scan(b"abc", esc) == 0 and st.close() == 0       # no handler: nothing to report, as in the reference
    esc.close()
    # refused in stream mode: LimEx engines (the writer refuses those itself)
    with pytest.raises(litdb.CompileError):
        litdb.compile_regex_multi([b"abc"], [0], [1], engine=litdb.ENGINE_LIMEX, mode=HS_MODE_STREAM)
