"""GPU tests (-m gpu) of the literal-program instructions beyond the plain report shapes: databases written with the
extended literal form (tests/test_litdb_ext.py validates the same databases on the unmodified reference runtime) are
scanned through the C ABI by every literal front end and compared with the reference runtime and the brute-force ground
truth; device-side counters prove that every case of the program interpreter was executed."""
import random

import pytest

import litgen
from test_litdb_ext import KINDS, KINDS_512, OP, compile_case, ext_cases, included_case
from vectorscan_b200 import gpu, litdb

pytestmark = pytest.mark.gpu
FRONTENDS = [0, 1, 2, 3]


def split_buffers(rng, text):
    """The text as one buffer plus the same text cut into ragged pieces (independent block scans)."""
    bufs = [text]
    p = 0
    while p < len(text):
        n = rng.choice([1, 7, 64, 300, 1500])
        bufs.append(text[p:p + n])
        p += n
    bufs.append(b"")
    return bufs


def gpu_batch(db, bufs, frontend, counts):
    sc = gpu.alloc_scratch(db)
    sc.set_option("frontend", frontend)
    sc.set_option("count_ops", 1)
    got = [(int(b), int(i), int(t)) for b, i, t in gpu.scan_vector_ex(db, bufs, sc)]
    for k, v in enumerate(sc.opcode_counts()):
        counts[k] += v
    sc.close()
    return got


def expected(lits, flags, ids, ext, bufs, goff):
    out = []
    for bi, b in enumerate(bufs):
        out += [(bi, i, t) for (i, t) in litgen.brute_force_ext(lits, flags, ids, ext, b, goff)]
    return sorted(out, key=lambda x: (x[0], x[2], x[1]))


@pytest.mark.parametrize("frontend", FRONTENDS)
def test_ext_programs_equal_reference_and_brute_force(ref, frontend):
    rng = random.Random(77 + frontend)
    counts = [0] * 128
    kinds = KINDS + KINDS_512  # the device implements the 64-byte instructions whatever ISA the host has
    ran = 0
    for kind, lits, flags, ids, ext, text, goff in ext_cases(11 + frontend, kinds):
        try:
            db = compile_case(kind, lits, flags, ids, ext, goff)
        except litdb.CompileError:
            continue
        bufs = split_buffers(rng, text)
        got = gpu_batch(db, bufs, frontend, counts)
        assert got == expected(lits, flags, ids, ext, bufs, goff), (kind, len(lits), len(text))
        if kind not in KINDS_512 or ref.isa == "avx512":
            rdb = ref.deserialize(db.serialize())
            rs = ref.alloc_scratch(rdb)
            rexp = []
            for bi, b in enumerate(bufs):
                rv, m = ref.scan(rdb, rs, b)
                assert rv == 0
                rexp += [(bi, i, t) for (i, t) in m]
            ref.L.hs_free_scratch(rs)
            ref.L.hs_free_database(rdb)
            assert got == sorted(rexp, key=lambda x: (x[0], x[2], x[1])), kind
        ran += 1
    assert ran >= 60
    for name in ("CHECK_GROUPS", "CHECK_MASK", "CHECK_MASK_32", "CHECK_MASK_64", "CHECK_BYTE", "SQUASH_GROUPS", "SET_EXHAUST"):
        assert counts[OP[name]] > 0, f"{name} never executed on the device"
    # and the everyday ones (the turbo front ends pre-decode plain report programs and never interpret them):
    # END 0, REPORT 33 / FINAL_REPORT 38, REPORT_EXHAUST 34, CHECK_EXHAUSTED 39, CHECK_MED_LIT 53
    assert counts[0] > 0 and counts[34] > 0 and counts[39] > 0 and counts[53] > 0 and counts[33] + counts[38] > 0


@pytest.mark.parametrize("frontend", FRONTENDS)
def test_included_jump_programs(ref, frontend):
    rng = random.Random(5)
    counts = [0] * 128
    for n in (20, 120, 600):
        for alpha in (b"abcd", b"abcdefghijklmnop"):
            lits, flags, ids, text = included_case(rng, n, alpha, rng.choice([500, 3000]))
            db = litdb.compile_lit_multi(lits, flags, ids, engine=litdb.ENGINE_FDR, fdr_domain=rng.choice([9, 11, 13]),
                                         fdr_stride=1, included_jump=1)
            bufs = split_buffers(rng, text)
            got = gpu_batch(db, bufs, frontend, counts)
            exp = []
            for bi, b in enumerate(bufs):
                exp += [(bi, i, t) for (i, t) in litgen.brute_force(lits, flags, ids, b)]
            assert got == sorted(exp, key=lambda x: (x[0], x[2], x[1])), (n, len(alpha))
    assert counts[OP["INCLUDED_JUMP"]] > 0


def test_every_interpreter_case_is_reached():
    """One database per remaining case of run_program: DEDUPE (+REPORT_EXHAUST / +SET_EXHAUST), DEDUPE_AND_REPORT,
    FINAL_REPORT, so that together with the tests above no case of the device interpreter is left unexecuted."""
    counts = [0] * 128
    # shared report id between two literals -> dkey; SINGLEMATCH -> DEDUPE + REPORT_EXHAUST; quiet -> DEDUPE + SET_EXHAUST
    lits = [b"alpha", b"alphabet", b"beta", b"betamax", b"gamma", b"gammaray", b"delta"]
    flags = [0, 0, 8, 8, 8, 8, 0]
    ids = [1, 1, 2, 2, 3, 3, 4]
    ext = [None, None, None, None, {"flags": 4}, {"flags": 4}, None]
    db = litdb.compile_lit_multi(lits, flags, ids, ext=ext)
    text = b"alphabet betamax gammaray delta alphabet betamax delta"
    got = gpu_batch(db, [text, text[:20]], 1, counts)
    exp = []
    for bi, b in enumerate([text, text[:20]]):
        exp += [(bi, i, t) for (i, t) in litgen.brute_force_ext(lits, flags, ids, ext, b)]
    assert got == sorted(exp, key=lambda x: (x[0], x[2], x[1]))
    # DEDUPE (its REPORT_EXHAUST / SET_EXHAUST successor is executed inside the same case), DEDUPE_AND_REPORT, FINAL_REPORT
    for code in (28, 37, 38):
        assert counts[code] > 0, code


def test_max_bi_anchored_width_gate_on_device():
    db = litdb.compile_lit_multi([b"needle", b"pin"], [0, 0], [1, 2], max_bi_anchored_width=40)
    short, long_ = b"x" * 10 + b"needle" + b"pin", b"x" * 60 + b"needle" + b"pin"
    sc = gpu.alloc_scratch(db)
    got = [(int(b), int(i), int(t)) for b, i, t in gpu.scan_vector_ex(db, [short, long_, short], sc)]
    assert got == [(0, 1, 16), (0, 2, 19), (2, 1, 16), (2, 2, 19)]


# ---- full-Rose databases: literal matcher (+ outfix engines) -------------------------------------------------------
from test_litdb_ext import KINDS_SHUFTI, mixed_case, mixed_expected  # noqa: E402


@pytest.mark.parametrize("frontend", FRONTENDS)
def test_full_rose_literal_databases_with_class_lookarounds(ref, frontend):
    """CHECK_SHUFTI_* (all six shapes) on the device against the reference's full interpreter and the ground truth."""
    rng = random.Random(170 + frontend)
    counts = [0] * 128
    ran = 0
    for kind, lits, flags, ids, ext, text, goff in ext_cases(31 + frontend, KINDS_SHUFTI + ["shufti64", "mask"]):
        try:
            db = litdb.compile_lit_multi(lits, flags, ids, ext=ext, groups_off=goff, full_rose=True,
                                         target_avx512=1 if kind == "shufti64" else 0)
        except litdb.CompileError:
            continue
        bufs = split_buffers(rng, text)
        got = gpu_batch(db, bufs, frontend, counts)
        assert got == expected(lits, flags, ids, ext, bufs, goff), (kind, len(lits), len(text))
        if kind != "shufti64" or ref.isa == "avx512":
            rdb = ref.deserialize(db.serialize())
            rs = ref.alloc_scratch(rdb)
            rexp = []
            for bi, b in enumerate(bufs):
                rv, m = ref.scan(rdb, rs, b)
                assert rv == 0
                rexp += [(bi, i, t) for (i, t) in m]
            ref.L.hs_free_scratch(rs)
            ref.L.hs_free_database(rdb)
            assert got == sorted(rexp, key=lambda x: (x[0], x[2], x[1])), kind
        ran += 1
    assert ran >= 30
    for name in ("CHECK_SHUFTI_16x8", "CHECK_SHUFTI_32x8", "CHECK_SHUFTI_64x8"):
        assert counts[OP[name]] > 0, f"{name} never executed on the device"
    assert counts[OP["CHECK_SHUFTI_16x16"]] + counts[OP["CHECK_SHUFTI_32x16"]] + counts[OP["CHECK_SHUFTI_64x16"]] > 0


@pytest.mark.parametrize("seed", range(8))
def test_mixed_literal_and_regex_databases(ref, seed):
    """C4's shape: floating literal matcher + outfix engines in one full-Rose database; the literal kernel and every
    engine add to one arena, the finalize merges by (buffer, offset, id)."""
    rng = random.Random(800 + seed)
    lits, flags, ids, ex, fl, rids, text = mixed_case(rng, rng.choice([1, 5, 40, 300]), rng.randint(1, 8), 4000)
    db = litdb.compile_lit_multi(lits, flags, ids, regexes=ex, regex_flags=fl, regex_ids=rids, dfa_states=rng.choice([30, 300]))
    bufs = split_buffers(rng, text)
    counts = [0] * 128
    got = gpu_batch(db, bufs, seed % 4, counts)
    exp = []
    for bi, b in enumerate(bufs):
        exp += [(bi, i, t) for (i, t) in mixed_expected(lits, flags, ids, ex, fl, rids, b)]
    assert got == sorted(exp, key=lambda x: (x[0], x[2], x[1]))
    rdb = ref.deserialize(db.serialize())
    rs = ref.alloc_scratch(rdb)
    rexp = []
    for bi, b in enumerate(bufs):
        rv, m = ref.scan(rdb, rs, b)
        assert rv == 0
        rexp += [(bi, i, t) for (i, t) in set(m)]
    assert got == sorted(rexp, key=lambda x: (x[0], x[2], x[1]))
    assert counts[20] > 0 or seed % 4 != 1   # CATCH_UP is interpreted (native front end runs every program)
