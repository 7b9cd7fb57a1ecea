"""CPU tests (not gpu): the plain-C oracle restatement (oracle/oracle_port.c) is pinned against
(a) known-answer vectors transcribed from the reference's unit tests, (b) matches recorded from the unmodified
reference runtime, (c) the live reference runtime where oracle/_ref is present, (d) brute force."""
import json
import os
import random

import pytest

import litgen
import oracle_port
from vectorscan_b200 import litdb

GOLD = os.path.join(os.path.dirname(__file__), "golden")
ENGINES = [dict(engine=litdb.ENGINE_AUTO)] + \
          [dict(engine=litdb.ENGINE_FDR, fdr_domain=d, fdr_stride=s) for d in (9, 12, 15) for s in (1, 2, 4)] + \
          [dict(engine=litdb.ENGINE_TEDDY, teddy_masks=m) for m in (1, 2, 3, 4)]


def _build(lits, flags, ids, opts):
    if opts.get("fdr_stride", 1) > min(min(len(l) for l in lits), 8):
        return None
    if opts.get("engine") == litdb.ENGINE_TEDDY and len(lits) < 2:
        return None
    try:
        return litdb.compile_lit_multi(lits, flags, ids, **opts)
    except litdb.CompileError:
        return None


def test_unit_vectors_every_engine():
    """unit/internal/fdr.cpp runs each vector for every valid FDR/Teddy engine id; so do we."""
    with open(os.path.join(GOLD, "ref_unit_vectors.json")) as f:
        cases = json.load(f)
    ran = 0
    dbs = {}
    for c in cases:
        lits = [x.encode("latin1") for x in c["lits"]]
        text = bytes.fromhex(c["text"])
        exp = sorted(tuple(x) for x in c["expect"])
        for opts in ENGINES:
            key = (tuple(lits), tuple(c["ids"]), tuple(sorted(opts.items())))
            if key not in dbs:
                dbs[key] = _build(lits, c["flags"], c["ids"], opts)
            db = dbs[key]
            if db is None:
                continue
            rv, got = oracle_port.scan(db, text)
            assert rv == 0
            assert sorted(got) == exp, (c["name"], opts)
            ran += 1
    assert ran > 5000


def test_recorded_reference_runtime_cases():
    with open(os.path.join(GOLD, "ref_runtime_cases.json")) as f:
        cases = json.load(f)
    assert len(cases) > 50
    for c in cases:
        lits = [bytes.fromhex(x) for x in c["lits"]]
        db = litdb.compile_lit_multi(lits, c["flags"], c["ids"], **c["opts"])
        rv, got = oracle_port.scan(db, bytes.fromhex(c["text"]))
        assert rv == 0
        assert sorted(got, key=lambda x: (x[1], x[0])) == [tuple(x) for x in c["expect"]]


def test_short_writings():
    """unit/internal/fdr.cpp:594-685 ShortWritings: exhaustive short buffers over a 3-symbol alphabet against
    exhaustive literals of length 1..8 over its first two symbols, 32 literals per table."""
    def fib(n):
        a = b = c = 1
        for _ in range(n):
            c = a + b
            a, b = b, c
        return c
    for alphabet in (b"abx", b"\x00A\x20"):
        bufs = []
        for ln in range(1, 7):
            for j in range(3 ** ln):
                bufs.append(bytes(alphabet[(j // 3 ** k) % 3] for k in range(ln)))
        base = len(bufs)
        for ln in range(7, 64, 8):
            for i in range(3):
                s, j = b"", 0
                while len(s) < ln:
                    s += bufs[fib(i * 5 + j + (ln - 6) * 10) % base]
                    j += 1
                bufs.append(s)
        pats = [bytes(alphabet[(j >> k) & 1] for k in range(ln)) for ln in range(1, 9) for j in range(2 ** ln)]
        sample = bufs[::7]
        for g in range(0, len(pats), 32):
            sigs = pats[g:g + 32]
            ids = list(range(g, g + len(sigs)))
            db = litdb.compile_lit_multi(sigs, [0] * len(sigs), ids, engine=litdb.ENGINE_FDR, fdr_domain=9, fdr_stride=1)
            for b in sample:
                rv, got = oracle_port.scan(db, b)
                assert sorted(got) == litgen.brute_force(sigs, [0] * len(sigs), ids, b)


def test_port_equals_live_reference_and_brute_force(ref):
    rng = random.Random(99)
    for n in (1, 4, 30, 48, 120, 700):
        for alpha in (b"ab", b"abcdefgh", bytes(range(32, 127)), bytes(range(256))):
            lits, flags, ids, text = litgen.random_case(rng, None, n, alpha, rng.choice([0, 3, 16, 40, 900, 4000]))
            for opts in (ENGINES[0], rng.choice(ENGINES[1:10]), rng.choice(ENGINES[10:])):
                db = _build(lits, flags, ids, opts)
                if db is None:
                    continue
                rv, got = oracle_port.scan(db, text)
                assert rv == 0
                rdb = ref.deserialize(db.serialize())
                rs = ref.alloc_scratch(rdb)
                rrv, rgot = ref.scan(rdb, rs, text)
                ref.L.hs_free_scratch(rs)
                ref.L.hs_free_database(rdb)
                assert sorted(got) == sorted(rgot) == litgen.brute_force(lits, flags, ids, text)


def test_callback_termination(ref):
    """hs_scan returns HS_SCAN_TERMINATED (-3) when the callback asks to stop (unit/hyperscan/single.cpp:517-632)."""
    db = litdb.compile_lit_multi([b"abc", b"bcd"], [0, 0], [1, 2])
    rv, got = oracle_port.scan(db, b"xxabcdxxabcd", stop_after=1)
    assert rv == -3 and len(got) == 1
    rdb = ref.deserialize(db.serialize())
    rs = ref.alloc_scratch(rdb)
    rrv, rgot = ref.scan(rdb, rs, b"xxabcdxxabcd", stop_after=1)
    assert rrv == -3 and rgot == got


def test_c_brute_force_agrees_with_python():
    rng = random.Random(5)
    lits, flags, ids, text = litgen.random_case(rng, None, 40, b"abcd", 500)
    flags = [f & 1 for f in flags]
    got = sorted(set(oracle_port.brute_force(lits, flags, ids, text)))
    assert got == litgen.brute_force(lits, flags, ids, text)


# ---- LimEx (single-outfix databases): the C restatement's limex_scan ---------------------------------------------
def test_limex_port_vs_recorded_reference_cases():
    import json
    import os
    from vectorscan_b200 import litdb
    with open(os.path.join(os.path.dirname(__file__), "golden", "ref_limex_cases.json")) as f:
        cases = json.load(f)
    n = 0
    for c in cases:
        db = litdb.compile_regex_multi([e.encode("latin-1") for e in c["exprs"]], c["flags"], c["ids"])
        for t, exp in zip(c["texts"], c["expect"]):
            rv, got = oracle_port.scan(db, bytes.fromhex(t))
            assert rv == 0 and sorted(got, key=lambda x: (x[1], x[0])) == [tuple(x) for x in exp], c["seed"]
            n += len(exp)
    assert n > 500


def test_limex_port_vs_live_reference_runtime(ref):
    import random
    import regen
    from vectorscan_b200 import litdb
    done = 0
    for seed in range(60):
        rng = random.Random(9000 + seed)
        ex, fl, ids, smp = regen.random_set(rng, rng.randint(1, 16), True)
        try:
            db = litdb.compile_regex_multi(ex, fl, ids)
        except litdb.CompileError:
            continue
        rdb = ref.deserialize(db.serialize())
        rs = ref.alloc_scratch(rdb)
        for n in (0, 1, 33, 500, 3000):
            t = regen.random_text(rng, ex, n, smp)
            rv, exp = ref.scan(rdb, rs, t)
            rv2, got = oracle_port.scan(db, t)
            assert rv == rv2 == 0 and got == exp, (seed, n)   # same callback ORDER too (ascending state index)
        done += 1
    assert done > 20


# ---- McClellan DFA (single-outfix databases): the C restatement's mcclellan_scan ---------------------------------
def test_mcclellan_port_vs_recorded_reference_cases():
    import json
    import os
    from vectorscan_b200 import litdb
    with open(os.path.join(os.path.dirname(__file__), "golden", "ref_mcclellan_cases.json")) as f:
        cases = json.load(f)
    n, kinds = 0, set()
    for c in cases:
        db = litdb.compile_regex_multi([e.encode("latin-1") for e in c["exprs"]], c["flags"], c["ids"], engine=c["engine"])
        assert db.outfix_engine_type() == c["type"]
        kinds.add((c["type"], c["engine"]))
        for t, exp in zip(c["texts"], c["expect"]):
            rv, got = oracle_port.scan(db, bytes.fromhex(t))
            assert rv == 0 and sorted(got, key=lambda x: (x[1], x[0])) == [tuple(x) for x in exp], c["seed"]
            n += len(exp)
    assert n > 300 and {(6, 1), (7, 1), (7, 2)} <= kinds


def test_mcclellan_port_vs_live_reference_runtime(ref):
    import random
    import regen
    from vectorscan_b200 import litdb
    done = 0
    for seed in range(60):
        rng = random.Random(9500 + seed)
        ex, fl, ids, smp = regen.random_set(rng, rng.randint(1, 12), True)
        try:
            db = litdb.compile_regex_multi(ex, fl, ids, engine=1 + seed % 2)
        except litdb.CompileError:
            continue
        rdb = ref.deserialize(db.serialize())
        rs = ref.alloc_scratch(rdb)
        for n in (0, 1, 33, 500, 3000):
            t = regen.random_text(rng, ex, n, smp)
            rv, exp = ref.scan(rdb, rs, t)
            rv2, got = oracle_port.scan(db, t)
            assert rv == rv2 == 0 and got == exp, (seed, n)   # same callback ORDER too (report-list order)
        done += 1
    assert done > 20


# ---- end-of-data accepts: LimEx acceptAtEOD and McClellan accept_eod through the C restatement --------------------
def test_port_vs_recorded_eod_cases():
    import json
    import os
    from vectorscan_b200 import litdb
    with open(os.path.join(os.path.dirname(__file__), "golden", "ref_eod_cases.json")) as f:
        cases = json.load(f)
    n_eod, engines = 0, set()
    for c in cases:
        db = litdb.compile_regex_multi([e.encode("latin-1") for e in c["exprs"]], c["flags"], c["ids"], engine=c["engine"])
        assert db.outfix_engine_type() == c["type"]
        engines.add(c["engine"])
        for t, exp in zip(c["texts"], c["expect"]):
            rv, got = oracle_port.scan(db, bytes.fromhex(t))
            assert rv == 0 and sorted(set(got), key=lambda x: (x[1], x[0])) == [tuple(x) for x in exp], c["seed"]
        n_eod += c["eod_matches"]
    assert engines == {0, 1, 2} and n_eod >= 20


# ---- Sheng DFA through the C restatement (sheng_scan) ----------------------------------------------------------
def test_sheng_port_vs_recorded_cases_and_live_reference(ref):
    import json
    import os
    import random
    import regen
    from vectorscan_b200 import litdb
    with open(os.path.join(os.path.dirname(__file__), "golden", "ref_sheng_cases.json")) as f:
        cases = json.load(f)
    n = 0
    for c in cases:
        db = litdb.compile_regex_multi([e.encode("latin-1") for e in c["exprs"]], c["flags"], c["ids"], engine=c["engine"])
        assert db.outfix_engine_type() == c["type"] == 17
        for t, exp in zip(c["texts"], c["expect"]):
            rv, got = oracle_port.scan(db, bytes.fromhex(t))
            assert rv == 0 and sorted(set(got), key=lambda x: (x[1], x[0])) == [tuple(x) for x in exp], c["seed"]
            n += len(exp)
    assert n > 300
    done = 0
    for seed in range(300):
        rng = random.Random(52000 + seed)
        ex, fl, ids, smp = regen.random_set(rng, rng.randint(1, 3), True)
        ex = regen.with_eod(rng, ex, 0.3)
        try:
            db = litdb.compile_regex_multi(ex, fl, ids, engine=litdb.ENGINE_SHENG)
        except litdb.CompileError:
            continue
        rdb = ref.deserialize(db.serialize())
        rs = ref.alloc_scratch(rdb)
        for ln in (0, 1, 33, 500):
            t = regen.random_text(rng, ex, ln, smp)
            assert ref.scan(rdb, rs, t) == oracle_port.scan(db, t), (seed, ln)   # callback order included
        done += 1
    assert done > 20
