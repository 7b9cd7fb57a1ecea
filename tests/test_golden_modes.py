"""tests/golden/ref_mode_cases.json (recorded from the unmodified reference runtime by make_golden_modes.py): vectored and
stream mode of pure-literal databases, stream mode of one DFA engine (end-of-data patterns reported at close), McSheng
and wide-state McClellan engines.  CPU: the oracle port reproduces the recorded callback sequences.  GPU: the device path
reproduces the recorded matches; needs neither /root/reference nor oracle/_ref at run time."""
import json
import os

import pytest

import oracle_port
from vectorscan_b200 import gpu, litdb

GOLD = os.path.join(os.path.dirname(__file__), "golden", "ref_mode_cases.json")


def load():
    with open(GOLD) as f:
        return json.load(f)


def build(c):
    if c["kind"] in ("vectored", "stream"):
        mode = litdb.HS_MODE_VECTORED if c["kind"] == "vectored" else litdb.HS_MODE_STREAM
        return litdb.compile_lit_multi([bytes.fromhex(x) for x in c["lits"]], c["flags"], c["ids"], mode=mode)
    mode = litdb.HS_MODE_STREAM if c["kind"] == "stream_dfa" else litdb.HS_MODE_BLOCK
    db = litdb.compile_regex_multi([e.encode("latin-1") for e in c["exprs"]], c["flags"], c["ids"], engine=c["engine"], mode=mode)
    assert db.outfix_engine_type() == c["type"]
    return db


def by_offset(pairs):
    """Callback order is (offset, then the engine's own order); the device delivers (offset, id): compare per offset."""
    return sorted((t, i) for i, t in pairs)


def test_fixture_covers_every_kind():
    kinds, types = {}, set()
    for c in load():
        n = len(c["expect"]) if c["kind"] in ("vectored", "stream") else \
            sum(len(r["expect"]) for r in c["runs"]) if c["kind"] == "stream_dfa" else sum(len(e) for e in c["expect"])
        kinds[c["kind"]] = kinds.get(c["kind"], 0) + n
        types.add(c.get("type"))
    assert all(kinds.get(k, 0) > 50 for k in ("vectored", "stream", "stream_dfa", "engine")), kinds
    assert {6, 7} <= types and types & {19, 20}      # McClellan 8 / 16 (Sherman, wide states), McSheng


def test_oracle_port_reproduces_the_recorded_callback_sequences():
    n = 0
    for c in load():
        db = build(c)
        if c["kind"] in ("vectored", "stream"):
            rv, got = oracle_port.scan_vector(db, [bytes.fromhex(b) for b in c["blocks"]])
            assert rv == 0 and [list(x) for x in got] == c["expect"], c["kind"]
            n += len(got)
        elif c["kind"] == "stream_dfa":
            for r in c["runs"]:
                rv, got = oracle_port.scan_stream_blocks(db, [bytes.fromhex(b) for b in r["blocks"]])
                assert rv == 0 and [list(x) for x in got] == r["expect"], c["exprs"]
                n += len(got)
        else:
            for t, exp in zip(c["texts"], c["expect"]):
                rv, got = oracle_port.scan(db, bytes.fromhex(t))
                assert rv == 0 and [list(x) for x in got] == exp, c["exprs"]
                n += len(got)
    assert n > 2500


@pytest.mark.gpu
def test_device_reproduces_the_recorded_matches():
    n = 0
    for c in load():
        db = build(c)
        sc = gpu.alloc_scratch(db)
        if c["kind"] == "vectored":
            got = [(int(i), int(t)) for _, i, t in gpu.scan_vector_ex(db, [bytes.fromhex(b) for b in c["blocks"]], sc)]
            assert by_offset(got) == by_offset(c["expect"])
            n += len(got)
        elif c["kind"] in ("stream", "stream_dfa"):
            for r in ([c] if c["kind"] == "stream" else c["runs"]):
                st, got = gpu.Stream(db), []
                cb = lambda i, f, t, fl, ctx: got.append((int(i), int(t))) or 0   # noqa: E731
                for b in r["blocks"]:
                    assert st.scan(bytes.fromhex(b), sc, cb) == gpu.HS_SUCCESS
                assert st.close(sc, cb) == gpu.HS_SUCCESS
                assert by_offset(got) == by_offset(r["expect"]), c["kind"]
                assert [t for _, t in got] == [t for _, t in r["expect"]]       # same offsets in the same order
                n += len(got)
        else:
            bufs = [bytes.fromhex(t) for t in c["texts"]]
            got = [(int(b), int(i), int(t)) for b, i, t in gpu.scan_vector_ex(db, bufs, sc)]
            exp = sorted((bi, i, t) for bi, m in enumerate(c["expect"]) for (i, t) in set(map(tuple, m)))
            assert sorted(got) == exp, c["exprs"]
            n += len(got)
        sc.close()
    assert n > 2500
