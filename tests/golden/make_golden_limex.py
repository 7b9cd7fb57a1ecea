"""Generates tests/golden/ref_limex_cases.json: seeded regex sets (tests/regen.py), compiled by the LimEx database
writer, scanned by the UNMODIFIED reference runtime (oracle/_ref) in this container; the recorded (id, to) lists
are what the GPU path must reproduce on the GPU box, where the reference sources do not exist.
Run from the repo root:  python tests/golden/make_golden_limex.py"""
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
    seed = 0
    want_sizes = {32: 4, 64: 4, 128: 4, 256: 4, 384: 3, 512: 3}
    while any(want_sizes.values()) and seed < 4000:
        seed += 1
        rng = random.Random(seed)
        ex, fl, ids, smp = regen.random_set(rng, rng.randint(1, 14), True)
        try:
            db = litdb.compile_regex_multi(ex, fl, ids)
        except litdb.CompileError:
            continue
        bits = db.limex_bits()
        if not want_sizes.get(bits):
            continue
        want_sizes[bits] -= 1
        rdb = R.deserialize(db.serialize())
        rs = R.alloc_scratch(rdb)
        texts, expect = [], []
        for k in range(6):
            t = regen.random_text(rng, ex, rng.choice([0, 1, 7, 64, 300, 1500]), smp)
            rv, got = R.scan(rdb, rs, t)
            assert rv == 0
            texts.append(t.hex())
            expect.append(sorted(got, key=lambda x: (x[1], x[0])))
        cases.append({"seed": seed, "bits": bits, "exprs": [e.decode("latin-1") for e in ex], "flags": fl, "ids": ids,
                      "texts": texts, "expect": expect})
    out = os.path.join(ROOT, "tests", "golden", "ref_limex_cases.json")
    with open(out, "w") as f:
        json.dump(cases, f)
    print(len(cases), "cases,", sum(len(e) for c in cases for e in c["expect"]), "matches ->", out)


if __name__ == "__main__":
    main()
