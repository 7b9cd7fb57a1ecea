"""CPU tests of the literal-set database writer: the UNMODIFIED reference runtime (oracle/_ref) accepts the
emitted blob and reports exactly the brute-force matches; serialization is byte-compatible."""
import random

import pytest

import litgen
from vectorscan_b200 import litdb


def test_serialize_roundtrip_and_crc(ref):
    db = litdb.compile_lit_multi([b"hello", b"worldwide web", b"x"], [1, 0, 8], [7, 8, 9])
    ser = db.serialize()
    assert len(ser) == db.size
    again = litdb.Database.deserialize(ser)
    assert again.bytecode()[:db.size - 104] == db.bytecode()[:db.size - 104]
    assert litdb.lib().hsb_crc32c(b"123456789", 9) == ref.L.ref_shim_crc32c(b"123456789", 9) == 0x58e3fa20
    rdb = ref.deserialize(ser)
    assert "BLOCK" in ref.info(rdb) and "5.4.11" in ref.info(rdb)
    assert ref.L.ref_shim_db_is_valid(rdb) == 0  # dbIsValid: magic, version, platform, alignment, CRC


def test_engine_selection_follows_reference_rules():
    def engine(db):
        bc = db.bytecode()
        f = int.from_bytes(bc[96:100], "little")
        return bc[f], int.from_bytes(bc[f + 192:f + 196], "little")
    assert engine(litdb.compile_lit_multi([b"single"]))[0] == 16                     # noodle
    t, eid = engine(litdb.compile_lit_multi([b"abc", b"defg", b"hijkl"]))
    assert t == 12 and 11 <= eid <= 18                                                # Teddy
    rng = random.Random(1)
    many = sorted({bytes(rng.choice(b"abcdefghijklmnop") for _ in range(rng.randint(4, 9))) for _ in range(400)})
    t, eid = engine(litdb.compile_lit_multi(many))
    assert t == 12 and eid == 0                                                       # FDR


def test_rejects_bad_input():
    with pytest.raises(litdb.CompileError):
        litdb.compile_lit_multi([b""])
    with pytest.raises(litdb.CompileError):
        litdb.compile_lit_multi([b"ab"], mode=3)
    with pytest.raises(litdb.CompileError):
        litdb.compile_lit_multi([b"ab", b"cd"], engine=litdb.ENGINE_NOODLE)
    with pytest.raises(litdb.CompileError):
        litdb.compile_lit_multi([b"ab"], flags=[256])  # HS_FLAG_SOM_LEFTMOST


@pytest.mark.parametrize("seed", [1, 2])
def test_reference_runtime_on_emitted_databases_equals_brute_force(ref, seed):
    rng = random.Random(seed)
    total = 0
    for n in (1, 2, 8, 48, 49, 300, 1200):
        for alpha in (b"ab", b"abcdefgh", bytes(range(32, 127)), bytes(range(256))):
            tl = rng.choice([0, 1, 15, 16, 17, 33, 100, 1000, 5000])
            lits, flags, ids, text = litgen.random_case(rng, None, n, alpha, tl)
            exp = litgen.brute_force(lits, flags, ids, text)
            engines = [dict()]
            if len(lits) > 1:
                ml = min(min(len(l) for l in lits), 8)
                d, s = rng.choice([9, 11, 13, 15]), rng.choice([1, 2, 4])
                if s <= ml:
                    engines.append(dict(engine=litdb.ENGINE_FDR, fdr_domain=d, fdr_stride=s))
                if len(lits) <= 48:
                    engines.append(dict(engine=litdb.ENGINE_TEDDY, teddy_masks=rng.randint(1, 4)))
            for opts in engines:
                try:
                    db = litdb.compile_lit_multi(lits, flags, ids, **opts)
                except litdb.CompileError:
                    assert opts.get("engine") == litdb.ENGINE_TEDDY
                    continue
                rdb = ref.deserialize(db.serialize())
                sc = ref.alloc_scratch(rdb)
                rv, got = ref.scan(rdb, sc, text)
                ref.L.hs_free_scratch(sc)
                ref.L.hs_free_database(rdb)
                assert rv == 0 and sorted(got) == exp, (n, len(alpha), tl, opts)
                total += 1
    assert total > 40


@pytest.mark.parametrize("seed", [3, 4])
def test_fat_teddy_databases_on_reference_runtime(ref, seed):
    """16-bucket "fat" Teddy (engines 3-10: what an AVX2 target gets for up to 96 fragments, teddy_engine_description.cpp:
    52-62; masks teddy_compile.cpp:436-500, dup masks :355-435): accepted and scanned by the AVX2 / AVX512 builds of the
    unmodified runtime (fdr_exec_fat_teddy_*, src/fdr/teddy_avx2.c / teddy.c) == brute force."""
    rng = random.Random(seed)
    ran, eids = 0, set()
    for trial in range(60):
        nl = rng.choice([2, 9, 17, 30, 60, 96])
        alpha = rng.choice([b"abcdefgh", bytes(range(32, 127)), bytes(range(256))])
        lits, flags, ids, text = litgen.random_case(rng, None, nl, alpha, rng.choice([0, 5, 40, 600, 5000]))
        try:
            db = litdb.compile_lit_multi(lits, flags, ids, engine=litdb.ENGINE_TEDDY, teddy_masks=rng.randint(1, 4), teddy_fat=1)
        except litdb.CompileError:
            continue
        bc = db.bytecode()
        f = int.from_bytes(bc[96:100], "little")
        eid = int.from_bytes(bc[f + 192:f + 196], "little")
        assert 3 <= eid <= 10
        eids.add(eid)
        rdb = ref.deserialize(db.serialize())
        sc = ref.alloc_scratch(rdb)
        rv, got = ref.scan(rdb, sc, text)
        ref.L.hs_free_scratch(sc)
        ref.L.hs_free_database(rdb)
        assert rv == 0 and sorted(got) == litgen.brute_force(lits, flags, ids, text), (nl, len(text))
        import oracle_port
        assert oracle_port.scan(db, text) == (0, got)   # the C restatement: same matches, same callback order
        ran += 1
    assert ran > 30 and len(eids) >= 6
