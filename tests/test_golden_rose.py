"""tests/golden/ref_rose_cases.json (recorded from the unmodified reference runtime by make_golden_rose.py): the full-Rose
shapes -- outfix engines only, literals + regexes -- and the extended literal programs.  CPU: the oracle port reproduces
the recorded callback sequences.  GPU: the device path reproduces the recorded matches; needs neither /root/reference nor
oracle/_ref at run time."""
import json
import os

import pytest

import oracle_port
from vectorscan_b200 import gpu, litdb

GOLD = os.path.join(os.path.dirname(__file__), "golden", "ref_rose_cases.json")


def load():
    with open(GOLD) as f:
        return json.load(f)


def build(c):
    if c["kind"] == "outfixes":
        return litdb.compile_regex_rose([e.encode("latin-1") for e in c["exprs"]], c["flags"], c["ids"], engine=c["engine"],
                                        dfa_states=c["dfa_states"])
    lits = [bytes.fromhex(x) for x in c["lits"]]
    if c["kind"] == "mixed":
        return litdb.compile_lit_multi(lits, c["flags"], c["ids"], regexes=[e.encode("latin-1") for e in c["exprs"]],
                                       regex_flags=c["regex_flags"], regex_ids=c["regex_ids"], dfa_states=c["dfa_states"])
    if c["kind"] == "included":
        return litdb.compile_lit_multi(lits, c["flags"], c["ids"], engine=litdb.ENGINE_FDR, fdr_domain=11, fdr_stride=1, included_jump=1)
    ext = []
    for e in c["ext"]:
        look = [(le[0], bytes.fromhex(le[1])) if len(le) == 2 else (le[0], le[1], le[2], le[3]) for le in e["look"]]
        ext.append({"look": look, "group": e["group"], "flags": e["flags"]})
    return litdb.compile_lit_multi(lits, c["flags"], c["ids"], ext=ext, groups_off=c["groups_off"], full_rose=c["full_rose"])


def test_fixture_covers_every_shape():
    kinds = {}
    for c in load():
        k = c["kind"] + (":" + c["ext_kind"] + (":full" if c["full_rose"] else "") if c["kind"] == "ext" else "")
        kinds[k] = kinds.get(k, 0) + sum(len(e) for e in c["expect"])
    for k in ("outfixes", "mixed", "included", "ext:byte", "ext:mask", "ext:mask32", "ext:groups", "ext:quiet",
              "ext:shufti16:full", "ext:shufti32:full"):
        assert kinds.get(k, 0) > 0, k


def test_oracle_port_reproduces_the_recorded_callback_sequences():
    n = 0
    for c in load():
        db = build(c)
        for t, exp in zip(c["texts"], c["expect"]):
            rv, got = oracle_port.scan(db, bytes.fromhex(t))
            assert rv == 0 and [list(x) for x in got] == exp, (c["kind"], c.get("ext_kind"))
            n += len(exp)
    assert n > 1500


@pytest.mark.gpu
@pytest.mark.parametrize("frontend", [0, 1])
def test_device_reproduces_the_recorded_matches(frontend):
    n = 0
    for c in load():
        db = build(c)
        bufs = [bytes.fromhex(t) for t in c["texts"]]
        sc = gpu.alloc_scratch(db)
        sc.set_option("frontend", frontend)
        got = [(int(b), int(i), int(t)) for b, i, t in gpu.scan_vector_ex(db, bufs, sc)]
        sc.close()
        exp = sorted({(bi, i, t) for bi, m in enumerate(c["expect"]) for (i, t) in m}, key=lambda x: (x[0], x[2], x[1]))
        assert got == exp, (c["kind"], c.get("ext_kind"))
        n += len(exp)
    assert n > 1500
