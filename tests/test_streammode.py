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
