"""GPU tests (-m gpu): Sheng DFA engines (<= 16 states, sheng_internal.h) through the C ABI.  The device path re-lays
the 256 shuffle masks as a state-major table at upload and steps it with k_mcclellan (csrc/hs_gpu.cu
build_sheng_tables).  Against matches recorded from the unmodified reference runtime
(tests/golden/ref_sheng_cases.json), the live reference runtime and the `re` brute force."""
import json
import os
import random

import pytest

import regen
from vectorscan_b200 import gpu, litdb
from test_gpu_limex import gpu_batch, ref_batch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_recorded_sheng_cases():
    with open(os.path.join(GOLD, "ref_sheng_cases.json")) as f:
        cases = json.load(f)
    n = 0
    for c in cases:
        db = litdb.compile_regex_multi([e.encode("latin-1") for e in c["exprs"]], c["flags"], c["ids"], engine=c["engine"])
        assert db.outfix_engine_type() == 17
        bufs = [bytes.fromhex(t) for t in c["texts"]]
        exp = [(bi, i, t) for bi, m in enumerate(c["expect"]) for (i, t) in m]
        assert gpu_batch(db, bufs) == exp, c["seed"]
        n += len(exp)
    assert n > 300


@pytest.mark.parametrize("seed", range(6))
def test_sheng_batches_vs_live_reference(ref, seed):
    rng = random.Random(7300 + seed)
    for _ in range(400):
        ex, fl, ids, smp = regen.random_set(rng, rng.randint(1, 3), True)
        ex = regen.with_eod(rng, ex, 0.3)
        try:
            db = litdb.compile_regex_multi(ex, fl, ids, engine=litdb.ENGINE_SHENG)
            break
        except litdb.CompileError:
            continue
    bufs = []
    for k in range(800):
        t = regen.random_text(rng, ex, rng.choice([0, 1, 15, 16, 17, 64, 300, 1400]), smp)
        if smp and k % 3 == 0:
            x = smp[k % len(smp)]
            t = t[:max(0, len(t) - len(x))] + x
        bufs.append(t)
    got = gpu_batch(db, bufs)
    assert got == ref_batch(ref, db, bufs) and len(got) > 20


def test_sheng_anchored_only_dies(ref):
    ex = [b"^GET /a", b"^POST"]
    db = litdb.compile_regex_multi(ex, [0, 0], [1, 2], engine=litdb.ENGINE_SHENG)
    bufs = [b"GET /a/b", b"POST /x", b"xGET /a", b"", b"GET /" + b"a" * 5000]
    got = gpu_batch(db, bufs)
    assert got == ref_batch(ref, db, bufs) and len(got) == 3


@pytest.mark.parametrize("seed", range(6))
def test_sheng32_and_sheng64_batches(seed):
    """The AVX512VBMI forms (struct sheng32 / sheng64, 64-byte masks, NFA types 21 / 22): the device re-lays them like the
    16-state form.  Checked against the `re` brute force and the C restatement (the reference runtime only executes them on
    an AVX512VBMI host: tests/test_nfadb.py pins the writer and the port there)."""
    import oracle_port
    rng = random.Random(7600 + seed)
    for _ in range(400):
        ex, fl, ids, smp = regen.random_set(rng, rng.randint(2, 5), True)
        ex = regen.with_eod(rng, ex, 0.3)
        try:
            db = litdb.compile_regex_multi(ex, fl, ids, engine=litdb.ENGINE_SHENG_WIDE)
            if db.outfix_engine_type() == (21 if seed % 2 else 22):
                break
        except litdb.CompileError:
            continue
    assert db.outfix_engine_type() in (21, 22)
    bufs = []
    for k in range(300):
        t = regen.random_text(rng, ex, rng.choice([0, 1, 15, 16, 17, 64, 300, 1400]), smp)
        if smp and k % 3 == 0:
            x = smp[k % len(smp)]
            t = t[:max(0, len(t) - len(x))] + x
        bufs.append(t)
    got = gpu_batch(db, bufs)
    exp = []
    for bi, b in enumerate(bufs):
        exp += [(bi, i, t) for (i, t) in regen.brute_force(ex, fl, ids, b)]
    assert got == sorted(exp, key=lambda x: (x[0], x[2], x[1])) and len(got) > 20
    port = []
    for bi, b in enumerate(bufs[:60]):
        rv, m = oracle_port.scan(db, b)
        assert rv == 0
        port += [(bi, i, t) for (i, t) in sorted(set(m), key=lambda x: (x[1], x[0]))]
    assert [g for g in got if g[0] < 60] == port
