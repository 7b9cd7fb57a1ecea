#!/usr/bin/env python
"""Generates the committed golden fixtures.  Run in the build container (needs oracle/_ref for part b):

    python tests/golden/make_golden.py

(a) ref_unit_vectors.json -- known-answer vectors TRANSCRIBED from the reference's own unit tests
    (/root/reference/unit/internal/fdr.cpp: Simple :167-190, SimpleSingle :192-216, MultiLocation :218-244,
    AlignAndTooEarly :496-561 with its three pattern/alien pairs :563-567).  Inputs and expected (end, id)
    lists are literals in that source; `end` there is the index OF the last byte, so `to` = end + 1.
(b) ref_runtime_cases.json -- seeded random literal sets / inputs with the matches the UNMODIFIED reference
    runtime (oracle/_ref, compiled from /root/reference) reports for the database our writer emits.
"""
import json
import os
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def unit_vectors():
    cases = []
    # FDRp.Simple: scanned with sizeof(data), i.e. including the terminating NUL
    data = b"mnopqrabcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ12345678901234567890mnopqr\x00"
    cases.append({"name": "fdr.Simple", "lits": ["mnopqr"], "ids": [0], "flags": [0], "text": data.hex(),
                  "expect": [[0, 5 + 1], [0, 23 + 1], [0, 83 + 1]]})
    data = b"mnopqrabcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ12345678901234567890m0m"
    cases.append({"name": "fdr.SimpleSingle", "lits": ["m"], "ids": [0], "flags": [0], "text": data.hex(),
                  "expect": [[0, 0 + 1], [0, 18 + 1], [0, 78 + 1], [0, 80 + 1]]})
    for i in range(128 - 3):
        buf = bytearray(128)
        buf[i:i + 3] = b"abc"
        cases.append({"name": f"fdr.MultiLocation[{i}]", "lits": ["abc"], "ids": [1], "flags": [0],
                      "text": bytes(buf).hex(), "expect": [[1, i + 2 + 1]]})
    for pattern, alien in ((b"abaabaaa", ord("x")), (b"zzzyyzyz", 0x99), (b"abcdef l", 0)):
        for lit_len in range(1, len(pattern) + 1):
            for i in (0, 1, 7, 15, 16, 17, 31):  # a subset of the 32 alignments, all j
                base = bytearray([alien]) * 160
                base[i:i + lit_len] = pattern[:lit_len]
                base[i + 128 - lit_len:i + 128] = pattern[:lit_len]
                for j in range(lit_len + 1):
                    seg = bytes(base[i + j:i + j + 128 - 2 * j])
                    exp = [[0, lit_len], [0, 128]] if j == 0 else []
                    cases.append({"name": f"fdr.AlignAndTooEarly[{pattern.decode('latin1')},{lit_len},{i},{j}]",
                                  "lits": [pattern[:lit_len].decode("latin1")], "ids": [0], "flags": [0],
                                  "text": seg.hex(), "expect": exp})
    return cases


def runtime_cases():
    import litgen
    import refrt
    from vectorscan_b200 import litdb
    R = refrt.RefRuntime()
    rng = random.Random(20260104)
    out = []
    plan = [(1, {}), (3, {}), (12, {"engine": 2, "teddy_masks": 2}), (30, {"engine": 2, "teddy_masks": 4}),
            (40, {"engine": 1, "fdr_domain": 9, "fdr_stride": 1}), (200, {"engine": 1, "fdr_domain": 13, "fdr_stride": 2}),
            (300, {"engine": 1, "fdr_domain": 15, "fdr_stride": 1})]
    for n, opts in plan:
        for alpha in (b"abcdefgh", bytes(range(32, 127)), bytes(range(256))):
            for tl in (0, 1, 9, 33, 700):
                lits, flags, ids, text = litgen.random_case(rng, None, n, alpha, tl)
                if opts.get("fdr_stride", 1) > min(len(l) for l in lits):
                    continue
                try:
                    db = litdb.compile_lit_multi(lits, flags, ids, **opts)
                except litdb.CompileError:
                    continue
                rdb = R.deserialize(db.serialize())
                sc = R.alloc_scratch(rdb)
                rv, got = R.scan(rdb, sc, text)
                assert rv == 0
                out.append({"lits": [l.hex() for l in lits], "flags": flags, "ids": ids, "opts": opts,
                            "text": text.hex(), "expect": sorted([list(x) for x in got], key=lambda x: (x[1], x[0]))})
                R.L.hs_free_scratch(sc)
                R.L.hs_free_database(rdb)
    return out


if __name__ == "__main__":
    uv = unit_vectors()
    with open(os.path.join(HERE, "ref_unit_vectors.json"), "w") as f:
        json.dump(uv, f, separators=(",", ":"))
    print("ref_unit_vectors.json:", len(uv), "cases")
    rc = runtime_cases()
    with open(os.path.join(HERE, "ref_runtime_cases.json"), "w") as f:
        json.dump(rc, f, separators=(",", ":"))
    print("ref_runtime_cases.json:", len(rc), "cases")
