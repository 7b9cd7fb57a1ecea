"""The LimEx database writer (csrc/nfadb_writer.cpp) against the UNMODIFIED reference runtime (oracle/_ref) and an
independent ground truth (Python `re`, all end offsets).  Acceptance is mechanical: hs_deserialize_database and
hs_alloc_scratch must accept the blob and hs_scan (runtime.c:246-279 single-outfix path ->
limex_runtime_impl.h nfaExecLimEx*_Q) must report exactly the brute-force matches."""
import random

import pytest

import regen
from vectorscan_b200 import litdb


def _scan_ref(ref, db, texts):
    rdb = ref.deserialize(db.serialize())
    rs = ref.alloc_scratch(rdb)
    out = []
    for t in texts:
        rv, got = ref.scan(rdb, rs, t)
        assert rv == 0
        out.append(sorted(set(got), key=lambda x: (x[1], x[0])))
    return out


@pytest.mark.parametrize("seed", range(24))
def test_random_sets_match_brute_force(ref, seed):
    rng = random.Random(seed)
    ex, fl, ids, smp = regen.random_set(rng, rng.randint(1, 6), True)
    try:
        db = litdb.compile_regex_multi(ex, fl, ids)
    except litdb.CompileError as e:
        assert "too large" in str(e)
        return
    texts = [regen.random_text(rng, ex, 140, smp) for _ in range(3)]
    got = _scan_ref(ref, db, texts)
    for t, g in zip(texts, got):
        assert g == regen.brute_force(ex, fl, ids, t), (ex, fl, t)


@pytest.mark.parametrize("n_states_hint,reps", [(20, 1), (50, 1), (100, 2), (200, 4), (400, 8)])
def test_every_limex_size(ref, n_states_hint, reps):
    """LimEx32/64/128/256/384/512 all come out of the writer and are all accepted (nfa_api_dispatch.c)."""
    ex = [b"ab[c-f]{%d}xy%d" % (n_states_hint // (2 * reps), k) for k in range(reps)]
    ex.append(b"^GET /[^ ]*\\.php")
    fl = [0] * len(ex)
    ids = list(range(len(ex)))
    db = litdb.compile_regex_multi(ex, fl, ids)
    rng = random.Random(n_states_hint)
    text = b"GET /a/b.php ab" + b"cdef" * 60 + b"xy0 abcdefxy1" + regen.random_text(rng, ex, 64)
    got = _scan_ref(ref, db, [text])[0]
    assert got == regen.brute_force(ex, fl, ids, text)


def test_refusals():
    for bad in (b"a*", b"(a|)", b"", b"a{0,3}"):
        with pytest.raises(litdb.CompileError):
            litdb.compile_regex_multi([bad])
    with pytest.raises(litdb.CompileError):
        litdb.compile_regex_multi([b"a[b-z]{600}c"])
