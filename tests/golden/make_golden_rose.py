"""Generates tests/golden/ref_rose_cases.json: databases of the full-Rose shapes and of the extended literal form --
(a) outfix engines only (hsb_compile_regex_rose), (b) literals + regexes in one database (hsb_compile_mixed), (c) literal
sets whose programs carry lookaround / group / quiet / inclusion instructions -- scanned by the UNMODIFIED reference
runtime (oracle/_ref) in this container; the recorded callback sequences are what the oracle port (CPU tests) and the GPU
path (-m gpu, where /root/reference does not exist) must reproduce.   Run:  python tests/golden/make_golden_rose.py"""
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import litgen  # noqa: E402
import refrt  # noqa: E402
import regen  # noqa: E402
from test_litdb_ext import ext_cases, included_case, mixed_case  # noqa: E402
from vectorscan_b200 import litdb  # noqa: E402


def record(R, db, texts):
    rdb = R.deserialize(db.serialize())
    assert R.L.ref_shim_db_is_valid(rdb) == 0
    rs = R.alloc_scratch(rdb)
    out = []
    for t in texts:
        rv, got = R.scan(rdb, rs, t)
        assert rv == 0
        out.append([[int(i), int(to)] for i, to in got])   # callback order
    R.L.hs_free_scratch(rs)
    R.L.hs_free_database(rdb)
    return out


def ext_json(ext):
    out = []
    for e in ext:
        e = e or {}
        look = [[le[0], le[1].hex()] if len(le) == 2 else [le[0], le[1], le[2], bool(le[3])] for le in e.get("look", [])]
        out.append({"look": look, "group": e.get("group", 0), "flags": e.get("flags", 0)})
    return out


def main():
    R = refrt.RefRuntime("avx2")   # nothing here needs the 64-byte instructions
    cases = []
    for seed in range(6):          # (a) outfix engines only
        rng = random.Random(41000 + seed)
        ex, fl, ids, smp = regen.random_set(rng, rng.randint(3, 12), True)
        engine, ds = rng.choice([(1, 24), (1, 300), (2, 600), (0, 0)])
        db = litdb.compile_regex_rose(ex, fl, ids, engine=engine, dfa_states=ds)
        texts = [regen.random_text(rng, ex, rng.choice([0, 1, 40, 300, 1400]), smp) for _ in range(6)]
        cases.append({"kind": "outfixes", "exprs": [e.decode("latin-1") for e in ex], "flags": fl, "ids": ids, "engine": engine,
                      "dfa_states": ds, "texts": [t.hex() for t in texts], "expect": record(R, db, texts)})
    for seed in range(4):          # (a') outfix engines with end-of-data patterns: EOD program (ENGINES_EOD)
        rng = random.Random(41500 + seed)
        ex, fl, ids, smp = regen.random_set(rng, rng.randint(3, 10), True)
        ex = regen.with_eod(rng, ex, 0.5)
        engine, ds = [(1, 24), (1, 300), (0, 0), (2, 600)][seed]
        db = litdb.compile_regex_rose(ex, fl, ids, engine=engine, dfa_states=ds)
        texts = []
        for k in range(6):
            t = regen.random_text(rng, ex, rng.choice([20, 60, 140]), smp)
            if smp and k % 2 == 0:
                t = t[:max(0, len(t) - len(smp[k % len(smp)]))] + smp[k % len(smp)]
            texts.append(t)
        cases.append({"kind": "outfixes", "exprs": [e.decode("latin-1") for e in ex], "flags": fl, "ids": ids, "engine": engine,
                      "dfa_states": ds, "texts": [t.hex() for t in texts], "expect": record(R, db, texts)})
    for seed in range(6):          # (b) literals + regexes
        rng = random.Random(42000 + seed)
        lits, flags, ids, ex, fl, rids, text = mixed_case(rng, rng.choice([1, 5, 40, 300]), rng.randint(1, 8), 3000)
        ds = rng.choice([30, 300])
        db = litdb.compile_lit_multi(lits, flags, ids, regexes=ex, regex_flags=fl, regex_ids=rids, dfa_states=ds)
        texts = [text, text[:17], text[100:900], b""]
        cases.append({"kind": "mixed", "lits": [l.hex() for l in lits], "flags": flags, "ids": ids,
                      "exprs": [e.decode("latin-1") for e in ex], "regex_flags": fl, "regex_ids": rids, "dfa_states": ds,
                      "texts": [t.hex() for t in texts], "expect": record(R, db, texts)})
    n_ext = 0                      # (c) extended literal programs, pure-literal and full-Rose
    for full_rose, kinds in ((False, ["byte", "mask", "mask32", "groups", "quiet"]), (True, ["shufti16", "shufti32", "mask"])):
        for kind, lits, flags, ids, ext, text, goff in ext_cases(43000 + int(full_rose), kinds):
            if len(lits) > 60 or len(text) > 1200 or not text:
                continue
            try:
                db = litdb.compile_lit_multi(lits, flags, ids, ext=ext, groups_off=goff, full_rose=full_rose)
            except litdb.CompileError:
                continue
            texts = [text, text[: len(text) // 2]]
            cases.append({"kind": "ext", "ext_kind": kind, "full_rose": full_rose, "lits": [l.hex() for l in lits], "flags": flags,
                          "ids": ids, "ext": ext_json(ext), "groups_off": goff, "texts": [t.hex() for t in texts],
                          "expect": record(R, db, texts)})
            n_ext += 1
    for seed in range(3):          # suffix inclusion
        rng = random.Random(44000 + seed)
        lits, flags, ids, text = included_case(rng, 60, b"abcdefghijklmnop", 1500)
        db = litdb.compile_lit_multi(lits, flags, ids, engine=litdb.ENGINE_FDR, fdr_domain=11, fdr_stride=1, included_jump=1)
        cases.append({"kind": "included", "lits": [l.hex() for l in lits], "flags": flags, "ids": ids,
                      "texts": [text.hex()], "expect": record(R, db, [text])})
    out = os.path.join(ROOT, "tests", "golden", "ref_rose_cases.json")
    with open(out, "w") as f:
        json.dump(cases, f)
    print(len(cases), "cases (", n_ext, "ext ),", sum(len(e) for c in cases for e in c["expect"]), "matches ->", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
