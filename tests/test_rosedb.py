"""Full-Rose databases of outfix engines only (csrc/nfadb_writer.cpp build_rose): the UNMODIFIED
reference runtime must accept the blob and scan it through roseBlockExec -> blockInitSufPQ -> roseCatchUpTo
(src/rose/block.c:345-422, src/rose/catchup.c:560-672); its matches must equal the Python-`re` ground truth.  The C
restatement (oracle/oracle_port.c) is pinned on the same databases, callback order included."""
import random

import pytest

import oracle_port
import regen
from vectorscan_b200 import litdb


def rose_header(db):
    bc = db.bytecode()
    return dict(impl=bc[4], queues=int.from_bytes(bc[156:160], "little"))


def scan_ref(ref, db, texts, ordered=False):
    rdb = ref.deserialize(db.serialize())
    assert ref.L.ref_shim_db_is_valid(rdb) == 0
    rs = ref.alloc_scratch(rdb)
    out = []
    for t in texts:
        rv, got = ref.scan(rdb, rs, t)
        assert rv == 0
        out.append(got if ordered else sorted(set(got), key=lambda x: (x[1], x[0])))
    ref.L.hs_free_scratch(rs)
    ref.L.hs_free_database(rdb)
    return out


@pytest.mark.parametrize("engine,dfa_states", [(litdb.ENGINE_MCCLELLAN, 24), (litdb.ENGINE_MCCLELLAN, 200),
                                               (litdb.ENGINE_MCCLELLAN_SHERMAN, 600), (litdb.ENGINE_LIMEX, 0)])
@pytest.mark.parametrize("seed", range(10))
def test_multi_outfix_sets_match_brute_force(ref, seed, engine, dfa_states):
    rng = random.Random(9000 + seed)
    ex, fl, ids, smp = regen.random_set(rng, rng.randint(2, 14), True)
    db = litdb.compile_regex_rose(ex, fl, ids, engine=engine, dfa_states=dfa_states)
    h = rose_header(db)
    assert h["impl"] == 0 and h["queues"] >= 1          # ROSE_RUNTIME_FULL_ROSE
    texts = [regen.random_text(rng, ex, 200, smp) for _ in range(3)] + [b"", b"x"]
    got = scan_ref(ref, db, texts)
    for t, g in zip(texts, got):
        assert g == regen.brute_force(ex, fl, ids, t), (ex, fl, t)


def test_many_engines_and_shared_ids(ref):
    """A set that splits into dozens of engines; two patterns share a report id across engines (one dedupe key)."""
    rng = random.Random(4)
    words = [b"admin", b"login", b"passwd", b"select", b"union", b"script", b"eval", b"cmd", b"shell", b"upload"]
    ex, fl, ids = [], [], []
    for k in range(60):
        w, w2 = rng.choice(words), rng.choice(words)
        ex.append(rng.choice([w + b"[0-9a-f]{1,4}" + w2, w + b"\\s+[a-z]{2,5}=", b"/" + w + b"/[^ ?]*\\.\\./", b"(" + w + b"|" + w2 + b")[^&]*&x="]))
        fl.append(1 if k % 3 == 0 else 0)
        ids.append(k if k % 7 else 1000)
    db = litdb.compile_regex_rose(ex, fl, ids, dfa_states=120)
    assert rose_header(db)["queues"] > 8
    texts = [regen.random_text(rng, ex, 400) for _ in range(3)]
    texts.append(b"admin12login /cmd/a/../ LOGIN  ab= (eval x&x= adminffeval shell1cmd")
    got = scan_ref(ref, db, texts)
    total = 0
    for t, g in zip(texts, got):
        assert g == regen.brute_force(ex, fl, ids, t)
        total += len(g)
    assert total > 5


def test_refuses_eod_patterns_and_bad_engine():
    with pytest.raises(litdb.CompileError):
        litdb.compile_regex_rose([b"abc\\z", b"def"])
    with pytest.raises(litdb.CompileError):
        litdb.compile_regex_rose([b"abc"], engine=litdb.ENGINE_SHENG)


@pytest.mark.parametrize("seed", range(6))
def test_oracle_port_on_multi_outfix_databases(ref, seed):
    rng = random.Random(9500 + seed)
    ex, fl, ids, smp = regen.random_set(rng, rng.randint(3, 10), True)
    db = litdb.compile_regex_rose(ex, fl, ids, engine=rng.choice([litdb.ENGINE_MCCLELLAN, litdb.ENGINE_LIMEX]),
                                  dfa_states=rng.choice([30, 300]))
    texts = [regen.random_text(rng, ex, 160, smp) for _ in range(3)] + [b""]
    for t, g in zip(texts, scan_ref(ref, db, texts, ordered=True)):
        rv, pg = oracle_port.scan(db, t)
        assert rv == 0 and pg == g, (ex, t)   # same matches in the same callback order
