"""Generates tests/golden/ref_sheng_cases.json: small seeded regex sets (some with end-of-data anchored patterns) that
determinise to at most 16 states, compiled as a Sheng engine (sheng_internal.h), scanned by the UNMODIFIED reference
runtime (oracle/_ref) in this container.   python tests/golden/make_golden_sheng.py"""
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
    for seed in range(31000, 33000):
        if len(cases) >= 16:
            break
        rng = random.Random(seed)
        engine = 3
        ex, fl, ids, smp = regen.random_set(rng, rng.randint(1, 3), True)
        ex = regen.with_eod(rng, ex, 0.3)
        try:
            db = litdb.compile_regex_multi(ex, fl, ids, engine=engine)
        except litdb.CompileError:
            continue
        rdb = R.deserialize(db.serialize())
        rs = R.alloc_scratch(rdb)
        texts, expect, n_eod = [], [], 0
        eod_ids = {i for e, i in zip(ex, ids) if e.endswith(b"\\z")}
        for k in range(8):
            t = regen.random_text(rng, ex, rng.choice([1, 7, 64, 300, 1500]), smp)
            if smp and k % 2 == 0:
                x = smp[k % len(smp)]
                t = t[:max(0, len(t) - len(x))] + x
            rv, got = R.scan(rdb, rs, t)
            assert rv == 0
            texts.append(t.hex())
            expect.append(sorted(set(got), key=lambda x: (x[1], x[0])))
            n_eod += sum(1 for (i, to) in got if i in eod_ids and to == len(t))
        if not sum(len(e) for e in expect):
            continue
        cases.append({"seed": seed, "engine": engine, "type": db.outfix_engine_type(),
                      "exprs": [e.decode("latin-1") for e in ex], "flags": fl, "ids": ids, "texts": texts,
                      "expect": expect, "eod_matches": n_eod})
    out = os.path.join(ROOT, "tests", "golden", "ref_sheng_cases.json")
    with open(out, "w") as f:
        json.dump(cases, f)
    print(len(cases), "cases,", sum(len(e) for c in cases for e in c["expect"]), "matches,",
          sum(c["eod_matches"] for c in cases), "at end of data ->", out)


if __name__ == "__main__":
    main()
