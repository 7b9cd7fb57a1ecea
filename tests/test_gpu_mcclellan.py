"""GPU tests (-m gpu) of McClellan DFA stepping (csrc/mcclellan.cuh) through the C ABI:
single-outfix databases holding one McClellan8 / McClellan16 (with and without Sherman states), against (a) matches
recorded from the unmodified reference runtime (tests/golden/ref_mcclellan_cases.json), (b) the live reference
runtime (oracle/_ref) and the C restatement (oracle/oracle_port.c mcclellan_scan) on seeded batches, (c) the `re`
brute force.  Bit-exact: sorted (buffer, to, id) lists."""
import json
import os
import random

import pytest

import oracle_port
import regen
from vectorscan_b200 import gpu, litdb
from test_gpu_limex import gpu_batch, ref_batch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def port_batch(db, bufs):
    out = []
    for bi, b in enumerate(bufs):
        rv, m = oracle_port.scan(db, b)
        assert rv == 0
        out += [(bi, i, t) for (i, t) in m]
    return sorted(set(out), key=lambda x: (x[0], x[2], x[1]))


def test_recorded_reference_cases():
    with open(os.path.join(GOLD, "ref_mcclellan_cases.json")) as f:
        cases = json.load(f)
    kinds, n_matches = set(), 0
    for c in cases:
        db = litdb.compile_regex_multi([e.encode("latin-1") for e in c["exprs"]], c["flags"], c["ids"], engine=c["engine"])
        assert db.outfix_engine_type() == c["type"]
        kinds.add((c["type"], c["engine"]))
        bufs = [bytes.fromhex(t) for t in c["texts"]]
        exp = [(bi, i, t) for bi, m in enumerate(c["expect"]) for (i, t) in m]
        assert gpu_batch(db, bufs) == exp, c["seed"]
        sc = gpu.alloc_scratch(db)   # one buffer at a time through hs_gpu_scan (callback order = (to, id))
        for b, m in zip(bufs, c["expect"]):
            one = []
            rv = gpu.scan(db, b, sc, lambda i, f, t, fl, cx: one.append([i, t]) or 0)
            assert rv == 0 and one == m
        sc.close()
        n_matches += len(exp)
    assert {(6, 1), (7, 1), (7, 2)} <= kinds and n_matches > 300


@pytest.mark.parametrize("seed", range(12))
def test_random_batches_vs_oracle(seed):
    rng = random.Random(8000 + seed)
    for _ in range(200):
        ex, fl, ids, smp = regen.random_set(rng, rng.randint(1, 12), True)
        try:
            db = litdb.compile_regex_multi(ex, fl, ids, engine=1 + seed % 2)
            break
        except litdb.CompileError:
            continue
    lens = [0, 1, 2, 15, 16, 17, 31, 33, 64, 200, 577, 1500, 4096] + [rng.randint(1, 1600) for _ in range(1500)]
    bufs = [regen.random_text(rng, ex, n, smp) for n in lens]
    got = gpu_batch(db, bufs)
    exp = port_batch(db, bufs)
    assert len(exp) > 50
    assert got == exp, (seed, db.outfix_engine_type())


def test_vs_live_reference_runtime_and_brute_force(ref):
    rng = random.Random(77)
    ex = [b"passwd\\s+[A-Za-z0-9_]{4,9}=", b"(login|admin)[^&]*&id=", b"User-Agent: [^\\r\\n]{8}", b"select.{1,12}from", b"^GET /[^ ]*\\.php"]
    fl, ids = [1, 0, 0, 1, 0], [3, 1, 4, 1, 5]   # two patterns share id 1 (dedupe key)
    for engine in (litdb.ENGINE_MCCLELLAN, litdb.ENGINE_MCCLELLAN_SHERMAN):
        db = litdb.compile_regex_multi(ex, fl, ids, engine=engine)
        assert db.outfix_engine_type() == 7
        bufs = [b"GET /a/b.php x PASSWD  user_1= login=1&id= admin&id= select a,b from User-Agent: abcdefghij", b""] + \
               [regen.random_text(rng, ex, rng.randint(1, 900)) for _ in range(120)]
        got = gpu_batch(db, bufs)
        assert got == ref_batch(ref, db, bufs)
        bf = []
        for bi, b in enumerate(bufs):
            bf += [(bi, i, t) for (i, t) in regen.brute_force(ex, fl, ids, b)]
        assert got == bf and len(got) >= 6


def test_long_buffer_anchored_only_and_table_outside_shared_memory(ref):
    """One 1 MiB buffer (a single thread steps it); anchored-only patterns (the DFA dies: early exit must not change
    the result); a DFA whose table does not fit shared memory (read through L1/L2)."""
    rng = random.Random(5)
    ex, fl, ids, smp = regen.random_set(rng, 6, True)
    db = litdb.compile_regex_multi(ex, fl, ids, engine=litdb.ENGINE_MCCLELLAN)
    big = regen.random_text(rng, ex, 1 << 20, smp)
    assert gpu_batch(db, [big]) == ref_batch(ref, db, [big])
    ex2 = [b"^GET /[a-z]+\\.php", b"^POST /[^ ]* HTTP"]
    db2 = litdb.compile_regex_multi(ex2, [0, 0], [1, 2], engine=litdb.ENGINE_MCCLELLAN)
    bufs = [b"GET /index.php HTTP/1.1", b"POST /a/b HTTP/1.0\r\n", b"xGET /index.php", b"", b"GET /x.php" * 50]
    got = gpu_batch(db2, bufs)
    assert got == ref_batch(ref, db2, bufs) and len(got) == 2 + 1
    # > 200 KiB of table: long literals over a 180-symbol alphabet (alphaShift 8)
    words = [bytes(rng.randrange(40, 220) for _ in range(24)) for _ in range(18)]
    ex3 = [b"".join(b"\\x%02x" % c for c in w[:12]) + b"[0-9a-f]{2}" + b"".join(b"\\x%02x" % c for c in w[12:]) for w in words]
    db3 = litdb.compile_regex_multi(ex3, [0] * len(ex3), list(range(len(ex3))), engine=litdb.ENGINE_MCCLELLAN)
    assert db3.outfix_engine_type() == 7 and len(db3.bytecode()) > 256 * 1024
    bufs3 = []
    for k in range(200):
        t = bytearray(rng.randrange(40, 220) for _ in range(rng.randint(100, 2000)))
        w = words[k % len(words)]
        if k % 3 == 0:
            t[50:76] = w[:12] + b"7f" + w[12:]
        bufs3.append(bytes(t))
    got3 = gpu_batch(db3, bufs3)
    assert got3 == ref_batch(ref, db3, bufs3) and len(got3) > 10


def test_sherman_states_expanded_and_exact_forms_agree(ref, monkeypatch):
    """16-bit engines with Sherman states: by default the device gives every Sherman state its full row at upload and runs
    the branch-free loop; HS_GPU_SHERMAN_EXACT=1 keeps the reference's compressed form and steps it with doSherman16 per
    byte.  Same matches both ways, equal to the reference runtime's."""
    rng = random.Random(41)
    ex = [b"passwd\\s+[A-Za-z0-9_]{4,9}=", b"(login|admin)[^&]*&id=", b"User-Agent: [^\\r\\n]{8}", b"select.{1,12}from", b"^GET /[^ ]*\\.php"]
    fl, ids = [1, 0, 0, 1, 0], [3, 1, 4, 1, 5]
    db = litdb.compile_regex_multi(ex, fl, ids, engine=litdb.ENGINE_MCCLELLAN_SHERMAN)
    bc = db.bytecode()
    info = int.from_bytes(bc[236:240], "little")
    nfa = int.from_bytes(bc[info:info + 4], "little")
    m = nfa + 64
    assert int.from_bytes(bc[m + 28:m + 30], "little") < int.from_bytes(bc[m:m + 2], "little")    # sherman_limit < state_count
    bufs = [b"GET /a/b.php x PASSWD  user_1= login=1&id= admin&id= select a,b from User-Agent: abcdefghij", b""] + \
           [regen.random_text(rng, ex, rng.randint(1, 900)) for _ in range(150)]
    want = ref_batch(ref, db, bufs)
    assert gpu_batch(db, bufs) == want and len(want) >= 6
    monkeypatch.setenv("HS_GPU_SHERMAN_EXACT", "1")
    assert gpu_batch(db, bufs) == want
