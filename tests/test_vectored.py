"""CPU tests of HS_MODE_VECTORED pure-literal databases: the writer's vectored databases are accepted
by the UNMODIFIED reference runtime's hs_scan_vector (src/runtime.c:1106-1174), which reports -- whatever way the text is
cut into blocks -- exactly what the brute force finds in the concatenation and exactly what the reference's block-mode
hs_scan reports for the same literal set on the concatenation, callback order included; the oracle port's block-by-block
restatement (history in front of every block) agrees with both."""
import random

import pytest

import litgen
import oracle_port
from vectorscan_b200 import litdb

HS_MODE_VECTORED = litdb.HS_MODE_VECTORED


def cut(rng, text, pieces):
    """text as `pieces` consecutive blocks (some possibly empty)."""
    cuts = sorted(rng.randrange(0, len(text) + 1) for _ in range(max(0, pieces - 1)))
    return [text[a:b] for a, b in zip([0] + cuts, cuts + [len(text)])]


def cases(seed):
    rng = random.Random(seed)
    for n in (1, 3, 20, 70, 300, 1500):
        for alpha in (b"abcd", bytes(range(97, 123)), bytes(range(256))):
            for tl in (7, 200, 3000):
                lits, flags, ids, text = litgen.random_case(rng, None, n, alpha, tl)
                yield lits, flags, ids, text, rng.choice([1, 2, 4, 16, 64])


def ref_vectored(R, db, bufs, stop_after=0):
    rdb = R.deserialize(db.serialize())
    assert R.L.ref_shim_db_is_valid(rdb) == 0
    sc = R.alloc_scratch(rdb)
    rv, got = R.scan_vector(rdb, sc, bufs, stop_after=stop_after)
    R.L.hs_free_scratch(sc)
    R.L.hs_free_database(rdb)
    return rv, got


def ref_block(R, db, text):
    rdb = R.deserialize(db.serialize())
    sc = R.alloc_scratch(rdb)
    rv, got = R.scan(rdb, sc, text)
    R.L.hs_free_scratch(sc)
    R.L.hs_free_database(rdb)
    assert rv == 0
    return got


@pytest.mark.parametrize("seed", [1, 2])
def test_reference_hs_scan_vector_on_written_databases(ref, seed):
    rng = random.Random(100 + seed)
    total = 0
    for lits, flags, ids, text, pieces in cases(seed):
        vdb = litdb.compile_lit_multi(lits, flags, ids, mode=HS_MODE_VECTORED)
        bufs = cut(rng, text, pieces)
        rv, got = ref_vectored(ref, vdb, bufs)
        assert rv == 0
        want = litgen.brute_force(lits, flags, ids, text)
        assert sorted(set(got)) == want
        # the same literals as a block-mode database on the concatenation: same callbacks in the same order
        assert got == ref_block(ref, litdb.compile_lit_multi(lits, flags, ids), text)
        # the port, block by block with history
        prv, pgot = oracle_port.scan_vector(vdb, bufs)
        assert prv == 0 and pgot == got
        total += len(got)
    assert total > 2000


def test_block_and_vectored_entry_points_check_the_mode(ref):
    lits = [b"needle", b"hay"]
    vdb = litdb.compile_lit_multi(lits, [0, 0], [1, 2], mode=HS_MODE_VECTORED)
    bdb = litdb.compile_lit_multi(lits, [0, 0], [1, 2])
    # hs_scan on a vectored database and hs_scan_vector on a block database: HS_DB_MODE_ERROR (-7) in the reference
    rdb = ref.deserialize(vdb.serialize())
    sc = ref.alloc_scratch(rdb)
    assert ref.scan(rdb, sc, b"a needle in the hay")[0] == -7
    rv, got = ref.scan_vector(rdb, sc, [b"a nee", b"dle in the h", b"", b"ay"])
    assert rv == 0 and got == [(1, 8), (2, 19)]
    rdb2 = ref.deserialize(bdb.serialize())
    sc2 = ref.alloc_scratch(rdb2)
    assert ref.scan_vector(rdb2, sc2, [b"needle"])[0] == -7
    assert oracle_port.scan_vector(bdb, [b"needle"])[0] == -7
    assert oracle_port.scan(vdb, b"needle")[0] == -7
    assert oracle_port.scan_vector(vdb, [b"a nee", b"dle in the h", b"", b"ay"]) == (0, [(1, 8), (2, 19)])


def test_termination_and_single_match_across_blocks(ref):
    lits = [b"abc", b"bcd", b"zz"]
    flags = [8, 0, 8]   # single-match literals stay exhausted from block to block
    vdb = litdb.compile_lit_multi(lits, flags, [5, 6, 7], mode=HS_MODE_VECTORED)
    bufs = [b"xabcd", b"abcd", b"ab", b"cdzzz", b"zzabc"]
    rv, got = ref_vectored(ref, vdb, bufs)
    assert rv == 0
    assert sorted(got) == litgen.brute_force(lits, flags, [5, 6, 7], b"".join(bufs))
    assert [g for g in got if g[0] == 5] == [(5, 4)] and [g for g in got if g[0] == 7] == [(7, 15)]
    assert oracle_port.scan_vector(vdb, bufs) == (0, got)
    # callback asks to stop after the second match: HS_SCAN_TERMINATED (-3), nothing from later blocks
    rv, part = ref_vectored(ref, vdb, bufs, stop_after=2)
    assert rv == -3 and part == got[:2]
    assert oracle_port.scan_vector(vdb, bufs, stop_after=2) == (-3, part)


def test_lookaround_reaches_into_the_previous_block(ref):
    rng = random.Random(5)
    checked = 0
    for kind in ("byte", "mask", "mask32"):
        for n in (1, 6, 40):
            lits, flags, ids, ext, text = litgen.random_ext_case(rng, n, b"abcdefgh", 1500, kind)
            try:
                vdb = litdb.compile_lit_multi(lits, flags, ids, ext=ext, mode=HS_MODE_VECTORED)
            except litdb.CompileError as e:
                assert "bucket" in str(e) or "spans" in str(e), str(e)
                continue
            bufs = cut(rng, text, 40)
            rv, got = ref_vectored(ref, vdb, bufs)
            assert rv == 0
            assert got == ref_block(ref, litdb.compile_lit_multi(lits, flags, ids, ext=ext), text)
            assert oracle_port.scan_vector(vdb, bufs) == (0, got)
            checked += len(got)
    assert checked > 50
