"""CPU tests of HS_MODE_STREAM pure-literal databases: the writer's stream-mode databases are accepted
by the UNMODIFIED reference runtime's hs_open_stream / hs_scan_stream / hs_close_stream (src/runtime.c:543-1060), which
report -- however the text is cut into blocks -- what the brute force finds in the concatenation and what the reference's
block-mode hs_scan reports for the same literals on the concatenation, callback order included; the oracle port's
block-by-block restatement agrees."""
import random

import litgen
import oracle_port
import pytest
from test_vectored import cases, cut, ref_block
from vectorscan_b200 import litdb

HS_MODE_STREAM = litdb.HS_MODE_STREAM


def ref_stream(R, db, bufs, stop_after=0):
    rdb = R.deserialize(db.serialize())
    assert R.L.ref_shim_db_is_valid(rdb) == 0
    sc = R.alloc_scratch(rdb)
    rvs, got = R.scan_stream_blocks(rdb, sc, bufs, stop_after=stop_after)
    R.L.hs_free_scratch(sc)
    R.L.hs_free_database(rdb)
    return rvs, got


@pytest.mark.parametrize("seed", [3, 4])
def test_reference_stream_api_on_written_databases(ref, seed):
    rng = random.Random(200 + seed)
    total = 0
    for lits, flags, ids, text, pieces in cases(seed):
        sdb = litdb.compile_lit_multi(lits, flags, ids, mode=HS_MODE_STREAM)
        bufs = cut(rng, text, pieces)
        rvs, got = ref_stream(ref, sdb, bufs)
        assert all(rv == 0 for rv in rvs)
        assert sorted(set(got)) == litgen.brute_force(lits, flags, ids, text)
        assert got == ref_block(ref, litdb.compile_lit_multi(lits, flags, ids), text)
        assert oracle_port.scan_stream_blocks(sdb, bufs) == (0, got)
        total += len(got)
    assert total > 2000


def test_stream_database_is_refused_by_block_and_vectored_entry_points(ref):
    sdb = litdb.compile_lit_multi([b"needle", b"hay"], [0, 0], [1, 2], mode=HS_MODE_STREAM)
    rdb = ref.deserialize(sdb.serialize())
    sc = ref.alloc_scratch(rdb)
    assert ref.scan(rdb, sc, b"needle")[0] == -7
    assert ref.scan_vector(rdb, sc, [b"needle"])[0] == -7
    rvs, got = ref.scan_stream_blocks(rdb, sc, [b"a nee", b"dle in the h", b"", b"ay"])
    assert rvs == [0, 0, 0, 0, 0] and got == [(1, 8), (2, 19)]


def test_stream_termination_and_single_match(ref):
    lits, flags, ids = [b"abc", b"bcd", b"zz"], [8, 0, 8], [5, 6, 7]
    sdb = litdb.compile_lit_multi(lits, flags, ids, mode=HS_MODE_STREAM)
    bufs = [b"xabcd", b"abcd", b"ab", b"cdzzz", b"zzabc"]
    rvs, got = ref_stream(ref, sdb, bufs)
    assert all(rv == 0 for rv in rvs)
    assert sorted(got) == litgen.brute_force(lits, flags, ids, b"".join(bufs))
    assert oracle_port.scan_stream_blocks(sdb, bufs) == (0, got)
    # stop at the second match (inside block 0): that call and every later hs_scan_stream return HS_SCAN_TERMINATED
    rvs, part = ref_stream(ref, sdb, bufs, stop_after=2)
    assert part == got[:2] and rvs[:5] == [-3] * 5


# ---- stream mode of ONE DFA engine (single-outfix databases): the engine's state is what crosses block boundaries ----
DFA_ENGINES = [litdb.ENGINE_MCCLELLAN, litdb.ENGINE_MCCLELLAN_SHERMAN, litdb.ENGINE_MCSHENG, litdb.ENGINE_SHENG,
               litdb.ENGINE_MCCLELLAN_WIDE]


def dfa_stream_cases(seed, eod):
    """(database, patterns, flags, ids, blocks, whole text) for seeded sets over every DFA engine the writer knows."""
    import regen
    rng = random.Random(seed)
    for trial in range(40):
        ex, fl, ids, smp = regen.random_set(rng, rng.randint(1, 6), True)
        if eod and trial % 2:
            ex = regen.with_eod(rng, ex, 0.4)
        try:
            db = litdb.compile_regex_multi(ex, fl, ids, engine=DFA_ENGINES[trial % len(DFA_ENGINES)], mode=HS_MODE_STREAM)
        except litdb.CompileError:
            continue
        for k in range(4):
            text = regen.random_text(rng, ex, rng.choice([0, 5, 60, 300]), smp)
            if smp and k % 2 == 0:
                x = smp[k % len(smp)]
                text = text[:max(0, len(text) - len(x))] + x
            yield db, ex, fl, ids, cut(rng, text, rng.choice([1, 2, 5, 12])), text


def test_reference_stream_api_on_single_dfa_databases(ref):
    import regen
    n = 0
    for db, ex, fl, ids, bufs, text in dfa_stream_cases(31, eod=True):
        rvs, got = ref_stream(ref, db, bufs)
        assert all(rv == 0 for rv in rvs)
        assert sorted(set(got), key=lambda x: (x[1], x[0])) == regen.brute_force(ex, fl, ids, text)
        assert oracle_port.scan_stream_blocks(db, bufs) == (0, got)
        n += len(got)
    assert n > 150
