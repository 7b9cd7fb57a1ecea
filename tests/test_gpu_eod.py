"""GPU tests (-m gpu): end-of-data accepts (patterns ending in \\z of the writer's dialect) through the C ABI for the
three single-outfix engines (LimEx acceptAtEOD, McClellan accept_eod, McClellan with Sherman states), against matches
recorded from the unmodified reference runtime (tests/golden/ref_eod_cases.json), the live reference runtime and the
C restatement.  Reference: src/runtime.c:273-278 (nfaCheckFinalState after the block run)."""
import json
import os
import random

import pytest

import regen
from vectorscan_b200 import gpu, litdb
from test_gpu_limex import gpu_batch, ref_batch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_recorded_eod_cases():
    with open(os.path.join(GOLD, "ref_eod_cases.json")) as f:
        cases = json.load(f)
    engines, n_eod = set(), 0
    for c in cases:
        db = litdb.compile_regex_multi([e.encode("latin-1") for e in c["exprs"]], c["flags"], c["ids"], engine=c["engine"])
        bufs = [bytes.fromhex(t) for t in c["texts"]]
        exp = [(bi, i, t) for bi, m in enumerate(c["expect"]) for (i, t) in m]
        assert gpu_batch(db, bufs) == exp, (c["seed"], c["engine"])
        engines.add(c["engine"])
        n_eod += c["eod_matches"]
    assert engines == {0, 1, 2} and n_eod >= 20


@pytest.mark.parametrize("engine", [litdb.ENGINE_LIMEX, litdb.ENGINE_MCCLELLAN, litdb.ENGINE_MCCLELLAN_SHERMAN])
def test_eod_batches_vs_live_reference(ref, engine):
    rng = random.Random(4100 + engine)
    for _ in range(200):
        ex, fl, ids, smp = regen.random_set(rng, rng.randint(3, 10), True)
        ex = regen.with_eod(rng, ex)
        try:
            db = litdb.compile_regex_multi(ex, fl, ids, engine=engine)
            break
        except litdb.CompileError:
            continue
    bufs = []
    for k in range(600):
        t = regen.random_text(rng, ex, rng.choice([0, 1, 15, 16, 17, 64, 300, 1400]), smp)
        if smp and k % 2 == 0:
            x = smp[k % len(smp)]
            t = t[:max(0, len(t) - len(x))] + x
        bufs.append(t)
    got = gpu_batch(db, bufs)
    exp = ref_batch(ref, db, bufs)
    eod_ids = {i for e, i in zip(ex, ids) if e.endswith(b"\\z")}
    assert got == exp
    assert sum(1 for (b, i, t) in got if i in eod_ids and t == len(bufs[b])) > 5
