"""Generates tests/golden/ref_mcclellan_cases.json: seeded regex sets (tests/regen.py), compiled by the McClellan
database writer (8-bit, 16-bit, 16-bit with Sherman states), scanned by the UNMODIFIED reference runtime
(oracle/_ref) in this container; the recorded (id, to) lists are what the GPU path must reproduce on the GPU box,
where the reference sources do not exist.   Run from the repo root:  python tests/golden/make_golden_mcclellan.py"""
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refrt  # noqa: E402
import regen  # noqa: E402
from vectorscan_b200 import litdb  # noqa: E402


def main():
    R = refrt.RefRuntime()
    cases = []
    seed = 20000
    want = {(6, 1): 5, (7, 1): 5, (7, 2): 5}   # (NFA.type, writer engine) -> cases
    while any(want.values()) and seed < 24000:
        seed += 1
        rng = random.Random(seed)
        engine = rng.choice([1, 2])
        ex, fl, ids, smp = regen.random_set(rng, rng.randint(1, 10), True)
        try:
            db = litdb.compile_regex_multi(ex, fl, ids, engine=engine)
        except litdb.CompileError:
            continue
        key = (db.outfix_engine_type(), engine if db.outfix_engine_type() == 7 else 1)
        if not want.get(key):
            continue
        want[key] -= 1
        rdb = R.deserialize(db.serialize())
        rs = R.alloc_scratch(rdb)
        texts, expect = [], []
        for k in range(6):
            t = regen.random_text(rng, ex, rng.choice([0, 1, 7, 64, 300, 1500]), smp)
            rv, got = R.scan(rdb, rs, t)
            assert rv == 0
            texts.append(t.hex())
            expect.append(sorted(set(got), key=lambda x: (x[1], x[0])))
        cases.append({"seed": seed, "type": key[0], "engine": engine, "exprs": [e.decode("latin-1") for e in ex],
                      "flags": fl, "ids": ids, "texts": texts, "expect": expect})
    out = os.path.join(ROOT, "tests", "golden", "ref_mcclellan_cases.json")
    with open(out, "w") as f:
        json.dump(cases, f)
    print(len(cases), "cases,", sum(len(e) for c in cases for e in c["expect"]), "matches ->", out)


if __name__ == "__main__":
    main()
