"""GPU tests (-m gpu) of LimEx stepping (csrc/limex.cuh) through the C ABI: single-outfix
databases holding one LimExNFA{32..512}, against (a) matches recorded from the unmodified reference runtime
(tests/golden/ref_limex_cases.json), (b) the live reference runtime (oracle/_ref) on seeded batches, (c) the
`re` brute force.  Bit-exact: sorted (buffer, to, id) lists."""
import json
import os
import random

import pytest

import regen
from vectorscan_b200 import gpu, litdb

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def gpu_batch(db, bufs, sc=None):
    own = sc is None
    if own:
        sc = gpu.alloc_scratch(db)
    got = [(int(b), int(i), int(t)) for b, i, t in gpu.scan_vector_ex(db, bufs, sc)]
    if own:
        sc.close()
    return got


def ref_batch(ref, db, bufs):
    rdb = ref.deserialize(db.serialize())
    rs = ref.alloc_scratch(rdb)
    out = []
    for bi, b in enumerate(bufs):
        rv, m = ref.scan(rdb, rs, b)
        assert rv == 0
        out += [(bi, i, t) for (i, t) in m]
    ref.L.hs_free_scratch(rs)
    ref.L.hs_free_database(rdb)
    return sorted(out, key=lambda x: (x[0], x[2], x[1]))


def test_recorded_reference_cases():
    with open(os.path.join(GOLD, "ref_limex_cases.json")) as f:
        cases = json.load(f)
    seen_bits, n_matches = set(), 0
    for c in cases:
        db = litdb.compile_regex_multi([e.encode("latin-1") for e in c["exprs"]], c["flags"], c["ids"])
        assert db.limex_bits() == c["bits"]
        seen_bits.add(c["bits"])
        bufs = [bytes.fromhex(t) for t in c["texts"]]
        exp = [(bi, i, t) for bi, m in enumerate(c["expect"]) for (i, t) in m]
        got = gpu_batch(db, bufs)
        assert got == exp, c["seed"]
        # one buffer at a time through hs_gpu_scan (callback order = (to, id))
        sc = gpu.alloc_scratch(db)
        for b, m in zip(bufs, c["expect"]):
            one = []
            rv = gpu.scan(db, b, sc, lambda i, f, t, fl, cx: one.append([i, t]) or 0)
            assert rv == 0 and one == m
        sc.close()
        n_matches += len(exp)
    assert seen_bits == {32, 64, 128, 256, 384, 512} and n_matches > 500


@pytest.mark.parametrize("seed", range(12))
def test_random_batches_vs_reference_runtime(ref, seed):
    rng = random.Random(7000 + seed)
    for _ in range(200):
        ex, fl, ids, smp = regen.random_set(rng, rng.randint(1, 20), True)
        try:
            db = litdb.compile_regex_multi(ex, fl, ids)
            break
        except litdb.CompileError:
            continue
    lens = [0, 1, 2, 15, 16, 17, 31, 33, 64, 200, 577, 1500, 4096] + [rng.randint(1, 1600) for _ in range(400)]
    bufs = [regen.random_text(rng, ex, n, smp) for n in lens]
    got = gpu_batch(db, bufs)
    exp = ref_batch(ref, db, bufs)
    assert len(exp) > 50
    assert got == exp, (seed, db.limex_bits())


def test_brute_force_small():
    rng = random.Random(99)
    ex, fl, ids, smp = regen.random_set(rng, 5, True)
    db = litdb.compile_regex_multi(ex, fl, ids)
    bufs = [regen.random_text(rng, ex, 120, smp) for _ in range(40)]
    got = gpu_batch(db, bufs)
    exp = []
    for bi, b in enumerate(bufs):
        exp += [(bi, i, t) for (i, t) in regen.brute_force(ex, fl, ids, b)]
    assert got == exp and len(exp) > 20


def test_one_long_buffer_and_anchored_only(ref):
    """A single 1 MiB buffer (one lane group steps it alone) and a set with only anchored patterns (the automaton
    dies after a few bytes: the early exit must not change the result)."""
    rng = random.Random(5)
    ex, fl, ids, smp = regen.random_set(rng, 6, True)
    db = litdb.compile_regex_multi(ex, fl, ids)
    big = regen.random_text(rng, ex, 1 << 20, smp)
    assert gpu_batch(db, [big]) == ref_batch(ref, db, [big])
    ex2 = [b"^GET /[a-z]+\\.php", b"^POST /[^ ]* HTTP"]
    db2 = litdb.compile_regex_multi(ex2, [0, 0], [1, 2])
    bufs = [b"GET /index.php HTTP/1.1", b"POST /a/b HTTP/1.0\r\n", b"xGET /index.php", b"", b"GET /x.php" * 50]
    got = gpu_batch(db2, bufs)
    assert got == ref_batch(ref, db2, bufs) and len(got) == 2 + 1
