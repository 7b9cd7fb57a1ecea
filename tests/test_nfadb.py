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


# ---- McClellan DFA engine (mcclellan_internal.h), same acceptance: the unmodified runtime must scan it ----------
@pytest.mark.parametrize("engine", [litdb.ENGINE_MCCLELLAN, litdb.ENGINE_MCCLELLAN_SHERMAN, litdb.ENGINE_MCSHENG])
@pytest.mark.parametrize("seed", range(16))
def test_mcclellan_random_sets_match_brute_force(ref, seed, engine):
    rng = random.Random(1000 + seed)
    ex, fl, ids, smp = regen.random_set(rng, rng.randint(1, 6), True)
    try:
        db = litdb.compile_regex_multi(ex, fl, ids, engine=engine)
    except litdb.CompileError as e:
        assert "too large" in str(e) or "no Sheng region" in str(e)
        return
    assert db.outfix_engine_type() in ((19, 20) if engine == litdb.ENGINE_MCSHENG else (6, 7))
    texts = [regen.random_text(rng, ex, 140, smp) for _ in range(3)] + [b""]
    got = _scan_ref(ref, db, texts)
    for t, g in zip(texts, got):
        assert g == regen.brute_force(ex, fl, ids, t), (ex, fl, t)


@pytest.mark.parametrize("engine", [litdb.ENGINE_MCCLELLAN, litdb.ENGINE_MCCLELLAN_SHERMAN, litdb.ENGINE_MCSHENG])
def test_mcclellan_8_and_16_bit(ref, engine):
    """<= 256 states -> McClellan8 (NFA.type 6), more -> McClellan16 (type 7); both accepted by nfa_api_dispatch.c.
    McSheng (the hybrid the reference compiler prefers for DFAs too large for Sheng): types 19 / 20."""
    small = [b"abc", b"^GET /[^ ]*\\.php", b"x[0-9]{3}y"]
    big = [b"passwd\\s+[A-Za-z0-9_]{4,9}=", b"(login|admin)[^&]*&id=", b"User-Agent: [^\\r\\n]{8}", b"select.{1,12}from"]
    for ex, want in ((small, 6), (big, 7)):
        fl = [1 if k % 2 else 0 for k in range(len(ex))]
        ids = [10 + k for k in range(len(ex))]
        db = litdb.compile_regex_multi(ex, fl, ids, engine=engine)
        assert db.outfix_engine_type() == (want + 13 if engine == litdb.ENGINE_MCSHENG else want)
        rng = random.Random(want)
        texts = [b"GET /a/b.php x123y abc PASSWD  user_1= login=1&id= select a,b from User-Agent: abcdefghij"] + \
                [regen.random_text(rng, ex, 200) for _ in range(4)]
        got = _scan_ref(ref, db, texts)
        for t, g in zip(texts, got):
            assert g == regen.brute_force(ex, fl, ids, t), (ex, t)


# ---- end-of-data accepts (\\z): LimEx acceptAtEOD / McClellan accept_eod, NFA_ACCEPTS_EOD (runtime.c:273-278) -----
@pytest.mark.parametrize("engine", [litdb.ENGINE_LIMEX, litdb.ENGINE_MCCLELLAN, litdb.ENGINE_MCCLELLAN_SHERMAN, litdb.ENGINE_MCSHENG])
@pytest.mark.parametrize("seed", range(10))
def test_eod_patterns_match_brute_force(ref, seed, engine):
    rng = random.Random(3000 + seed)
    ex, fl, ids, smp = regen.random_set(rng, rng.randint(2, 6), True)
    ex = regen.with_eod(rng, ex)
    try:
        db = litdb.compile_regex_multi(ex, fl, ids, engine=engine)
    except litdb.CompileError as e:
        assert "too large" in str(e) or "no Sheng region" in str(e)
        return
    texts = []
    for k in range(6):
        t = regen.random_text(rng, ex, rng.choice([20, 60, 140]), smp)
        if smp and k % 2 == 0:   # end the text with a sample of one of the patterns, so that \\z patterns fire
            x = smp[k % len(smp)]
            t = t[:max(0, len(t) - len(x))] + x
        texts.append(t)
    got = _scan_ref(ref, db, texts)
    n = 0
    for t, g in zip(texts, got):
        assert g == regen.brute_force(ex, fl, ids, t), (ex, fl, t)
        n += len(g)


# ---- Sheng engine (sheng_internal.h): DFAs of at most 16 states as byte-shuffle masks ------------------------------
@pytest.mark.parametrize("seed", range(40))
def test_sheng_sets_match_brute_force(ref, seed):
    rng = random.Random(6000 + seed)
    ex, fl, ids, smp = regen.random_set(rng, rng.randint(1, 3), True)
    ex = regen.with_eod(rng, ex, 0.3)
    try:
        db = litdb.compile_regex_multi(ex, fl, ids, engine=litdb.ENGINE_SHENG)
    except litdb.CompileError as e:
        assert "too large" in str(e)
        return
    assert db.outfix_engine_type() == 17
    texts = [regen.random_text(rng, ex, 120, smp) for _ in range(3)] + [b""]
    if smp:
        texts.append(regen.random_text(rng, ex, 40, smp)[:20] + smp[0])
    got = _scan_ref(ref, db, texts)
    for t, g in zip(texts, got):
        assert g == regen.brute_force(ex, fl, ids, t), (ex, fl, t)


# ---- Sheng32 / Sheng64 (sheng_internal.h:81-107): the AVX512VBMI forms, 64-byte succ_masks ----------------------------
@pytest.mark.parametrize("seed", range(30))
def test_sheng32_and_sheng64_match_brute_force(seed):
    import oracle_port
    import refrt
    if refrt.pick_isa() != "avx512":
        pytest.skip("host CPU cannot run the AVX512VBMI build of the reference runtime")
    ref = refrt.RefRuntime("avx512")
    rng = random.Random(7000 + seed)
    ex, fl, ids, smp = regen.random_set(rng, rng.randint(1, 5), True)
    ex = regen.with_eod(rng, ex, 0.3)
    try:
        db = litdb.compile_regex_multi(ex, fl, ids, engine=litdb.ENGINE_SHENG_WIDE)
    except litdb.CompileError as e:
        assert "too large" in str(e)
        return
    assert db.outfix_engine_type() in (21, 22)
    texts = [regen.random_text(rng, ex, 120, smp) for _ in range(3)] + [b""]
    rdb = ref.deserialize(db.serialize())
    rs = ref.alloc_scratch(rdb)
    for t in texts:
        rv, got = ref.scan(rdb, rs, t)
        assert rv == 0 and sorted(set(got), key=lambda x: (x[1], x[0])) == regen.brute_force(ex, fl, ids, t), (ex, fl, t)
        assert oracle_port.scan(db, t) == (0, got)


# ---- McSheng hybrid (mcsheng_internal.h): the engine the reference compiler prefers for DFAs too large for Sheng ----
@pytest.mark.parametrize("seed", range(12))
def test_mcsheng_port_equals_reference_callback_order(ref, seed):
    """McSheng8 / McSheng16 databases, single outfix and as the DFA engine of multi-engine databases: the C restatement
    (oracle_port.c mcsheng_scan) raises the reference runtime's callbacks in the reference runtime's order."""
    import oracle_port
    rng = random.Random(8800 + seed)
    ex, fl, ids, smp = regen.random_set(rng, rng.randint(2, 9), True)
    if seed % 3 == 2:
        ex = regen.with_eod(rng, ex, 0.4)
    if seed % 2:
        db = litdb.compile_regex_rose(ex, fl, ids, engine=litdb.ENGINE_MCSHENG, dfa_states=[40, 400, 2000][seed % 3])
    else:
        try:
            db = litdb.compile_regex_multi(ex, fl, ids, engine=litdb.ENGINE_MCSHENG)
        except litdb.CompileError as e:
            assert "too large" in str(e) or "no Sheng region" in str(e)
            return
    rdb = ref.deserialize(db.serialize())
    rs = ref.alloc_scratch(rdb)
    n = 0
    for k in range(12):
        t = regen.random_text(rng, ex, rng.choice([0, 1, 9, 80, 500]), smp)
        if smp and k % 3 == 0:
            x = smp[k % len(smp)]
            t = t[:max(0, len(t) - len(x))] + x
        rv, got = ref.scan(rdb, rs, t)
        assert rv == 0
        assert sorted(set(got), key=lambda x: (x[1], x[0])) == regen.brute_force(ex, fl, ids, t)
        assert oracle_port.scan(db, t) == (0, got)
        n += len(got)
    assert n > 0


def test_mcsheng_engine_types_in_multi_engine_databases():
    """With ENGINE_MCSHENG a multi-engine database holds Sheng engines for DFAs of at most 16 states and McSheng engines
    (NFA.type 19 / 20) for the larger ones -- the order of getDfa (rose_build_bytecode.cpp)."""
    ex = [b"abc", b"passwd\\s+[A-Za-z0-9_]{4,9}=", b"(login|admin)[^&]*&id=", b"User-Agent: [^\\r\\n]{8}", b"select.{1,12}from"]
    db = litdb.compile_regex_rose(ex, [0] * 5, list(range(5)), engine=litdb.ENGINE_MCSHENG, dfa_states=300)
    bc = db.bytecode()
    nq = int.from_bytes(bc[156:160], "little")
    info = int.from_bytes(bc[236:240], "little")
    types = set()
    for q in range(nq):
        nfa = int.from_bytes(bc[info + 20 * q:info + 20 * q + 4], "little")
        types.add(bc[nfa + 8])
    assert types & {19, 20} and types <= {17, 19, 20, 6, 7}


# ---- McClellan wide states (mcclellan_internal.h:53-61): runs of single-entry states stored as one state ------------
def wide_case(rng):
    """A pattern set whose DFA has long single-entry runs (anchored literals, so that no other pattern's states merge into
    the runs) and more than 256 states, plus texts that walk the runs fully, partly, and leave them at every point."""
    def word(n):
        return bytes(rng.choice(b"abcdefghijklmnopqrstuvwxyz") for _ in range(n))
    lits = [word(rng.randint(11, 20)) for _ in range(24)]
    pats = [b"^" + l for l in lits[:20]] + [b"^" + lits[0][:10] + b"[0-9]+" + lits[1][:12], lits[20] + b"\\z",
                                             b"^" + lits[2][:9] + b".*" + lits[21][:10]]
    texts = [b""]
    for _ in range(160):
        l = lits[rng.randrange(len(lits))]
        r = rng.random()
        if r < .3:
            t = l[:rng.randint(0, len(l))]
        elif r < .5:
            t = l + word(rng.randint(0, 5))
        elif r < .6:
            t = lits[0][:10] + b"123" + lits[1][:12]
        elif r < .7:
            t = lits[2][:9] + word(rng.randint(0, 30)) + lits[21][:10] + word(3)
        elif r < .8:
            t = word(5) + lits[20]
        elif r < .9:
            b = bytearray(l)
            b[rng.randrange(len(b))] = rng.choice(b"abcxyz")
            t = bytes(b)
        else:
            t = lits[2][:9] + lits[21][:rng.randint(0, 10)]
        texts.append(t)
    return pats, [0] * len(pats), list(range(len(pats))), texts


def wide_header(db):
    """(NFA.type, has_wide, wide_limit, number of wide head states) of a single-outfix McClellan database."""
    bc = db.bytecode()
    info = int.from_bytes(bc[236:240], "little")
    nfa = int.from_bytes(bc[info:info + 4], "little")
    m = nfa + 64
    wo = int.from_bytes(bc[m + 304:m + 308], "little")
    return bc[nfa + 8], bc[m + 35], int.from_bytes(bc[m + 30:m + 32], "little"), \
        int.from_bytes(bc[nfa + wo + 2:nfa + wo + 4], "little") if wo else 0


@pytest.mark.parametrize("seed", range(6))
def test_mcclellan_wide_states_reference_brute_force_and_port(ref, seed):
    import oracle_port
    rng = random.Random(9900 + seed)
    ex, fl, ids, texts = wide_case(rng)
    db = litdb.compile_regex_multi(ex, fl, ids, engine=litdb.ENGINE_MCCLELLAN_WIDE)
    t, has_wide, wide_limit, heads = wide_header(db)
    assert t == 7 and has_wide == 1 and heads >= 10 and wide_limit > 0
    rdb = ref.deserialize(db.serialize())
    rs = ref.alloc_scratch(rdb)
    n = 0
    for text in texts:
        rv, got = ref.scan(rdb, rs, text)
        assert rv == 0
        assert sorted(set(got), key=lambda x: (x[1], x[0])) == regen.brute_force(ex, fl, ids, text), text
        assert oracle_port.scan(db, text) == (0, got)
        n += len(got)
    assert n > 40
    # the same set without wide states reports the same
    plain = litdb.compile_regex_multi(ex, fl, ids, engine=litdb.ENGINE_MCCLELLAN)
    assert wide_header(plain)[1] == 0 and db.size < plain.size


# ---- McSheng64 (mcsheng_internal.h:94-119): the AVX512VBMI form of the hybrid, up to 64 region states ----------------
@pytest.mark.parametrize("seed", range(12))
def test_mcsheng64_reference_brute_force_and_port(seed):
    import oracle_port
    import refrt
    if refrt.pick_isa() != "avx512":
        pytest.skip("host CPU cannot run the AVX512VBMI build of the reference runtime")
    ref = refrt.RefRuntime("avx512")
    rng = random.Random(8900 + seed)
    ex, fl, ids, smp = regen.random_set(rng, rng.randint(2, 9), True)
    if seed % 3 == 2:
        ex = regen.with_eod(rng, ex, 0.4)
    if seed % 2:
        db = litdb.compile_regex_rose(ex, fl, ids, engine=litdb.ENGINE_MCSHENG64, dfa_states=[40, 400, 2000][seed % 3])
    else:
        try:
            db = litdb.compile_regex_multi(ex, fl, ids, engine=litdb.ENGINE_MCSHENG64)
        except litdb.CompileError as e:
            assert "too large" in str(e) or "no Sheng region" in str(e)
            return
        assert db.outfix_engine_type() in (23, 24)
    rdb = ref.deserialize(db.serialize())
    rs = ref.alloc_scratch(rdb)
    n = 0
    for k in range(12):
        t = regen.random_text(rng, ex, rng.choice([0, 1, 9, 80, 500]), smp)
        if smp and k % 3 == 0:
            x = smp[k % len(smp)]
            t = t[:max(0, len(t) - len(x))] + x
        rv, got = ref.scan(rdb, rs, t)
        assert rv == 0
        assert sorted(set(got), key=lambda x: (x[1], x[0])) == regen.brute_force(ex, fl, ids, t)
        assert oracle_port.scan(db, t) == (0, got)
        n += len(got)
    assert n > 0
