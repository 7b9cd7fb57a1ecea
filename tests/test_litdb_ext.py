"""CPU tests of the writer's extended literal form (include/hs_litdb.h hsb_lit_ext): databases whose literal programs
carry the instructions the reference compiler attaches to masked / mixed-case / grouped / quiet / suffix-included
literals -- CHECK_BYTE, CHECK_MASK, CHECK_MASK_32, CHECK_MASK_64, CHECK_SHUFTI_*, CHECK_GROUPS, SQUASH_GROUPS,
SET_EXHAUST, INCLUDED_JUMP -- are accepted by the UNMODIFIED reference runtime (oracle/_ref) and reported exactly as the
brute-force ground truth says."""
import random

import pytest

import litgen
import oracle_port
from vectorscan_b200 import gpu, litdb

OP = dict(CHECK_GROUPS=3, CHECK_MASK=9, CHECK_MASK_32=10, CHECK_BYTE=11, CHECK_SHUFTI_16x8=12, CHECK_SHUFTI_32x8=13,
          CHECK_SHUFTI_16x16=14, CHECK_SHUFTI_32x16=15, SQUASH_GROUPS=43, INCLUDED_JUMP=61, SET_EXHAUST=65,
          CHECK_SHUFTI_64x8=67, CHECK_SHUFTI_64x16=68, CHECK_MASK_64=69)
KINDS = ["byte", "mask", "mask32", "groups", "quiet"]
KINDS_512 = ["mask64"]
KINDS_SHUFTI = ["shufti16", "shufti32"]  # full roseRunProgram only: roseRunProgram_l has no CHECK_SHUFTI_* case


def ext_cases(seed, kinds):
    rng = random.Random(seed)
    for kind in kinds:
        for n in (1, 6, 40, 300):
            for alpha in (b"abcdefgh", bytes(range(32, 127)), bytes(range(256))):
                tl = rng.choice([0, 9, 33, 100, 1000, 4000])
                lits, flags, ids, ext, text = litgen.random_ext_case(rng, n, alpha, tl, kind)
                goff = 0
                if kind == "groups" and rng.random() < 0.5:
                    goff = 1 << rng.randrange(6)
                yield kind, lits, flags, ids, ext, text, goff


def compile_case(kind, lits, flags, ids, ext, goff):
    return litdb.compile_lit_multi(lits, flags, ids, ext=ext, groups_off=goff,
                                   target_avx512=1 if kind in KINDS_512 else 0)


def ref_scan(R, db, text):
    rdb = R.deserialize(db.serialize())
    assert R.L.ref_shim_db_is_valid(rdb) == 0
    sc = R.alloc_scratch(rdb)
    rv, got = R.scan(rdb, sc, text)
    R.L.hs_free_scratch(sc)
    R.L.hs_free_database(rdb)
    assert rv == 0
    return got


@pytest.mark.parametrize("seed", [1, 2])
def test_reference_runtime_on_ext_databases_equals_brute_force(ref, seed):
    seen = [0] * 128
    total = 0
    for kind, lits, flags, ids, ext, text, goff in ext_cases(seed, KINDS):
        try:
            db = compile_case(kind, lits, flags, ids, ext, goff)
        except litdb.CompileError as e:
            assert "bucket" in str(e) or "spans" in str(e), str(e)
            continue
        for k, v in enumerate(gpu.database_opcodes(db)):
            seen[k] += v
        exp = litgen.brute_force_ext(lits, flags, ids, ext, text, goff)
        got = ref_scan(ref, db, text)
        assert sorted(got) == exp, (kind, len(lits), len(text))
        prv, pgot = oracle_port.scan(db, text)  # the C restatement, callback order included
        assert prv == 0 and pgot == got, (kind, len(lits), len(text))
        total += 1
    assert total >= 50
    for name in ("CHECK_GROUPS", "CHECK_MASK", "CHECK_MASK_32", "CHECK_BYTE", "SQUASH_GROUPS", "SET_EXHAUST"):
        assert seen[OP[name]] > 0, name


def test_avx512_only_instructions_on_the_avx512_reference_build(ref):
    """CHECK_MASK_64 / CHECK_SHUFTI_64x8 / 64x16 exist only in AVX512 builds of the runtime (program_runtime.c:2233-2238)."""
    if ref.isa != "avx512":
        pytest.skip("host CPU cannot run the AVX512 build of the reference runtime")
    seen = [0] * 128
    for kind, lits, flags, ids, ext, text, goff in ext_cases(3, KINDS_512):
        try:
            db = compile_case(kind, lits, flags, ids, ext, goff)
        except litdb.CompileError:
            continue
        for k, v in enumerate(gpu.database_opcodes(db)):
            seen[k] += v
        assert sorted(ref_scan(ref, db, text)) == litgen.brute_force_ext(lits, flags, ids, ext, text, goff), kind
        assert sorted(oracle_port.scan(db, text)[1]) == litgen.brute_force_ext(lits, flags, ids, ext, text, goff), kind
    assert seen[OP["CHECK_MASK_64"]] > 0


def included_case(rng, n, alpha, tl):
    """Literal sets rich in suffix pairs (x + lit, lit), all plain reports."""
    base = sorted({bytes(rng.choice(alpha) for _ in range(rng.randint(2, 5))) for _ in range(n)})
    lits = set(base)
    for b in base:
        if rng.random() < 0.6:
            lits.add(bytes(rng.choice(alpha) for _ in range(rng.randint(1, 3))) + b)
    lits = sorted(lits)
    nocase = rng.random() < 0.3
    flags = [1 if nocase else 0] * len(lits)
    ids = list(range(10, 10 + len(lits)))
    text = bytearray(rng.choice(alpha) for _ in range(tl))
    for _ in range(tl // 20 + 1):
        l = rng.choice(lits)
        if len(l) <= tl:
            p = rng.randrange(0, tl - len(l) + 1)
            text[p:p + len(l)] = l
    return lits, flags, ids, bytes(text)


@pytest.mark.parametrize("seed", [5, 6])
def test_included_jump_databases_on_reference_runtime(ref, seed):
    rng = random.Random(seed)
    jumps = 0
    for n in (20, 120, 600):
        for alpha in (b"abcd", b"abcdefghijklmnop", bytes(range(97, 123)) + bytes(range(65, 91))):
            lits, flags, ids, text = included_case(rng, n, alpha, rng.choice([50, 500, 3000]))
            db = litdb.compile_lit_multi(lits, flags, ids, engine=litdb.ENGINE_FDR, fdr_domain=rng.choice([9, 11, 13]),
                                         fdr_stride=1, included_jump=1)
            jumps += gpu.database_opcodes(db)[OP["INCLUDED_JUMP"]]
            got = ref_scan(ref, db, text)
            exp = litgen.brute_force(lits, flags, ids, text)
            # the jump must neither lose nor duplicate a match: compare as multisets
            assert sorted(got) == exp, (n, len(alpha))
            assert sorted(oracle_port.scan(db, text)[1]) == exp
    assert jumps > 20


def test_max_bi_anchored_width_gate(ref):
    """hs_scan drops buffers longer than RoseEngine.maxBiAnchoredWidth (src/runtime.c:356-361)."""
    db = litdb.compile_lit_multi([b"needle", b"pin"], [0, 0], [1, 2], max_bi_anchored_width=40)
    short, long_ = b"x" * 10 + b"needle" + b"pin", b"x" * 60 + b"needle" + b"pin"
    assert sorted(ref_scan(ref, db, short)) == [(1, 16), (2, 19)]
    assert ref_scan(ref, db, long_) == []


# ---- full-Rose databases: literal programs run through the full roseRunProgram; regexes ride along as outfix engines ----
@pytest.mark.parametrize("seed", [21, 22])
def test_full_rose_literal_databases_with_class_lookarounds(ref, seed):
    """CHECK_SHUFTI_16x8 / 32x8 / 16x16 / 32x16 exist only in the full interpreter (program_runtime.c:2379-2470): a
    literal-only FULL_ROSE database reaches it through roseFloatingCallback (src/rose/match.c:479-523)."""
    seen = [0] * 128
    total = 0
    for kind, lits, flags, ids, ext, text, goff in ext_cases(seed, KINDS_SHUFTI + ["mask", "groups", "quiet"]):
        try:
            db = litdb.compile_lit_multi(lits, flags, ids, ext=ext, groups_off=goff, full_rose=True)
        except litdb.CompileError as e:
            assert "bucket" in str(e) or "spans" in str(e), str(e)
            continue
        assert db.bytecode()[4] == 0   # ROSE_RUNTIME_FULL_ROSE
        for k, v in enumerate(gpu.database_opcodes(db)):
            seen[k] += v
        exp = litgen.brute_force_ext(lits, flags, ids, ext, text, goff)
        got = ref_scan(ref, db, text)
        assert sorted(got) == exp, (kind, len(lits), len(text))
        assert oracle_port.scan(db, text) == (0, got)   # the C restatement incl. CHECK_SHUFTI_*: same callback sequence
        total += 1
    assert total >= 40
    assert seen[OP["CHECK_SHUFTI_16x8"]] > 0 and seen[OP["CHECK_SHUFTI_32x8"]] > 0
    assert seen[OP["CHECK_SHUFTI_16x16"]] + seen[OP["CHECK_SHUFTI_32x16"]] > 0


def test_avx512_shufti_shapes_in_full_rose(ref):
    if ref.isa != "avx512":
        pytest.skip("host CPU cannot run the AVX512 build of the reference runtime")
    seen = [0] * 128
    for kind, lits, flags, ids, ext, text, goff in ext_cases(23, ["shufti64"]):
        try:
            db = litdb.compile_lit_multi(lits, flags, ids, ext=ext, full_rose=True, target_avx512=1)
        except litdb.CompileError:
            continue
        for k, v in enumerate(gpu.database_opcodes(db)):
            seen[k] += v
        assert sorted(ref_scan(ref, db, text)) == litgen.brute_force_ext(lits, flags, ids, ext, text, goff)
    assert seen[OP["CHECK_SHUFTI_64x8"]] + seen[OP["CHECK_SHUFTI_64x16"]] > 0


def mixed_case(rng, n_lits, n_rx, tl):
    import regen
    alpha = bytes(range(97, 123)) + b"0123456789 =&/."
    lits, flags, ids, text = litgen.random_case(rng, None, n_lits, alpha, tl)
    ids = [k + 1 for k in range(len(lits))]
    ex, fl, rids, smp = regen.random_set(rng, n_rx, True)
    rids = [100000 + k for k in range(len(ex))]
    t = bytearray(text)
    for s_ in smp[:8]:
        if len(s_) < len(t):
            p = rng.randrange(0, len(t) - len(s_))
            t[p:p + len(s_)] = s_
    return lits, flags, ids, ex, fl, rids, bytes(t)


def mixed_expected(lits, flags, ids, ex, fl, rids, text):
    import regen
    return sorted(litgen.brute_force(lits, flags, ids, text) + [tuple(x) for x in regen.brute_force(ex, fl, rids, text)])


@pytest.mark.parametrize("seed", range(8))
def test_mixed_literal_and_regex_databases(ref, seed):
    """Floating literal matcher + outfix engines in one FULL_ROSE database (C4's shape): reference hs_scan == literal brute
    force + `re` brute force, delivered in offset order."""
    rng = random.Random(700 + seed)
    lits, flags, ids, ex, fl, rids, text = mixed_case(rng, rng.choice([1, 5, 40, 300]), rng.randint(1, 8), rng.choice([60, 400, 3000]))
    db = litdb.compile_lit_multi(lits, flags, ids, regexes=ex, regex_flags=fl, regex_ids=rids, dfa_states=rng.choice([30, 300]))
    assert db.bytecode()[4] == 0
    for t in (text, text[:17], b""):
        got = ref_scan(ref, db, t)
        assert sorted(set(got)) == mixed_expected(lits, flags, ids, ex, fl, rids, t)
        assert [x[1] for x in got] == sorted(x[1] for x in got)   # CATCH_UP keeps the callbacks in offset order
        prv, pgot = oracle_port.scan(db, t)
        assert prv == 0 and pgot == got   # the C restatement: same matches, same callback order
