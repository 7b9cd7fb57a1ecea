"""Generates tests/golden/ref_mode_cases.json: what the UNMODIFIED reference runtime (oracle/_ref) reports, in this
container, for the database modes and engines added late in round 2 --
  vectored   pure-literal HS_MODE_VECTORED databases through hs_scan_vector (blocks of one stream),
  stream     pure-literal HS_MODE_STREAM databases through hs_open_stream / hs_scan_stream / hs_close_stream,
  stream_dfa single-outfix HS_MODE_STREAM databases holding one DFA engine, with end-of-data patterns (reported at close),
  engine     block-mode single-outfix databases: McSheng (types 19/20), McClellan with wide states.
The recorded callback sequences are what the oracle port (CPU tests) and the device path (-m gpu, where neither
/root/reference nor, possibly, oracle/_ref exists) must reproduce.   Run:  python tests/golden/make_golden_modes.py"""
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
from test_nfadb import wide_case  # noqa: E402
from test_vectored import cut  # noqa: E402
from vectorscan_b200 import litdb  # noqa: E402


def pairs(got):
    return [[int(i), int(t)] for i, t in got]


def main():
    R = refrt.RefRuntime("avx2")
    cases = []
    for seed in range(8):          # vectored and stream, pure literal
        rng = random.Random(52000 + seed)
        n = [1, 3, 20, 70, 300, 1500, 40, 8][seed]
        alpha = [b"abcd", bytes(range(97, 123)), bytes(range(256))][seed % 3]
        lits, flags, ids, text = litgen.random_case(rng, None, n, alpha, [7, 200, 3000][seed % 3])
        bufs = cut(rng, text, [1, 2, 4, 16, 64][seed % 5])
        for kind, mode in (("vectored", litdb.HS_MODE_VECTORED), ("stream", litdb.HS_MODE_STREAM)):
            db = litdb.compile_lit_multi(lits, flags, ids, mode=mode)
            rdb = R.deserialize(db.serialize())
            sc = R.alloc_scratch(rdb)
            if kind == "vectored":
                rv, got = R.scan_vector(rdb, sc, bufs)
                assert rv == 0
            else:
                rvs, got = R.scan_stream_blocks(rdb, sc, bufs)
                assert not any(rvs)
            assert sorted(set(got)) == litgen.brute_force(lits, flags, ids, text)
            cases.append({"kind": kind, "lits": [l.hex() for l in lits], "flags": flags, "ids": ids,
                          "blocks": [b.hex() for b in bufs], "expect": pairs(got)})
    engines = [litdb.ENGINE_MCCLELLAN, litdb.ENGINE_MCCLELLAN_SHERMAN, litdb.ENGINE_MCSHENG, litdb.ENGINE_SHENG, litdb.ENGINE_MCCLELLAN_WIDE]
    made = 0
    for seed in range(40):         # stream mode, one DFA engine, half of the sets with \z patterns
        if made == 10:
            break
        rng = random.Random(53000 + seed)
        ex, fl, ids, smp = regen.random_set(rng, rng.randint(1, 6), True)
        if seed % 2:
            ex = regen.with_eod(rng, ex, 0.4)
        engine = engines[seed % len(engines)]
        try:
            db = litdb.compile_regex_multi(ex, fl, ids, engine=engine, mode=litdb.HS_MODE_STREAM)
        except litdb.CompileError:
            continue
        rdb = R.deserialize(db.serialize())
        sc = R.alloc_scratch(rdb)
        runs = []
        for k in range(4):
            text = regen.random_text(rng, ex, rng.choice([5, 60, 300]), smp)
            if smp and k % 2 == 0:
                x = smp[k % len(smp)]
                text = text[:max(0, len(text) - len(x))] + x
            bufs = cut(rng, text, rng.choice([1, 2, 5, 12]))
            rvs, got = R.scan_stream_blocks(rdb, sc, bufs)
            assert not any(rvs)
            assert sorted(set(got), key=lambda m: (m[1], m[0])) == regen.brute_force(ex, fl, ids, text)
            runs.append({"blocks": [b.hex() for b in bufs], "expect": pairs(got)})
        cases.append({"kind": "stream_dfa", "exprs": [e.decode("latin-1") for e in ex], "flags": fl, "ids": ids, "engine": engine,
                      "type": db.outfix_engine_type(), "runs": runs})
        made += 1
    for seed in range(6):          # block mode: McSheng and wide-state McClellan
        rng = random.Random(54000 + seed)
        if seed < 3:
            ex, fl, ids, texts = wide_case(rng)
            engine, texts = litdb.ENGINE_MCCLELLAN_WIDE, texts[:40]
        else:
            ex, fl, ids, smp = regen.random_set(rng, rng.randint(2, 8), True)
            engine = litdb.ENGINE_MCSHENG
            texts = [regen.random_text(rng, ex, rng.choice([0, 9, 80, 500]), smp) for _ in range(12)]
        try:
            db = litdb.compile_regex_multi(ex, fl, ids, engine=engine)
        except litdb.CompileError:
            continue
        rdb = R.deserialize(db.serialize())
        sc = R.alloc_scratch(rdb)
        expect = []
        for t in texts:
            rv, got = R.scan(rdb, sc, t)
            assert rv == 0
            expect.append(pairs(got))
        cases.append({"kind": "engine", "exprs": [e.decode("latin-1") for e in ex], "flags": fl, "ids": ids, "engine": engine,
                      "type": db.outfix_engine_type(), "texts": [t.hex() for t in texts], "expect": expect})
    out = os.path.join(ROOT, "tests", "golden", "ref_mode_cases.json")
    with open(out, "w") as f:
        json.dump(cases, f, separators=(",", ":"))
    n = sum(len(c["expect"]) if c["kind"] in ("vectored", "stream") else sum(len(r["expect"]) for r in c["runs"])
            if c["kind"] == "stream_dfa" else sum(len(e) for e in c["expect"]) for c in cases)
    print(f"{len(cases)} cases, {n} recorded matches -> {out} ({os.path.getsize(out)} bytes)")


if __name__ == "__main__":
    main()
