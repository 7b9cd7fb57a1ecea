"""GPU tests (-m gpu) of McSheng hybrid DFAs (NFA types 19 / 20, src/nfa/mcsheng_internal.h): the
engine the reference compiler builds for DFAs too large for Sheng.  The device completes the successor table at upload
(rows of the Sheng region made from the shuffle masks, hs_gpu.cu build_mcsheng_tables) and steps it with k_mcclellan.
Against the live reference runtime (oracle/_ref), the C restatement (tests/test_nfadb.py pins it on the reference,
callback order included) and the `re` brute force.  Bit-exact: sorted (buffer, to, id) lists."""
import random

import pytest

import regen
from test_gpu_limex import gpu_batch, ref_batch
from test_gpu_mcclellan import port_batch
from vectorscan_b200 import gpu, litdb

pytestmark = pytest.mark.gpu


def brute_batch(ex, fl, ids, bufs):
    out = []
    for bi, b in enumerate(bufs):
        out += [(bi, i, t) for (i, t) in regen.brute_force(ex, fl, ids, b)]
    return sorted(out, key=lambda x: (x[0], x[2], x[1]))


def vbmi_ref(ref, engine):
    """McSheng64 runs on the AVX512VBMI build of the reference only."""
    if engine != litdb.ENGINE_MCSHENG64:
        return ref
    import refrt
    if refrt.pick_isa() != "avx512":
        pytest.skip("host CPU cannot run the AVX512VBMI build of the reference runtime")
    return refrt.RefRuntime("avx512")


@pytest.mark.parametrize("engine", [litdb.ENGINE_MCSHENG, litdb.ENGINE_MCSHENG64])
@pytest.mark.parametrize("seed", range(3))
def test_single_outfix_mcsheng_batches(ref, seed, engine):
    ref = vbmi_ref(ref, engine)
    rng = random.Random(8600 + seed)
    for _ in range(200):
        ex, fl, ids, smp = regen.random_set(rng, rng.randint(1, 10), True)
        if seed % 4 == 3:
            ex = regen.with_eod(rng, ex, 0.4)
        try:
            db = litdb.compile_regex_multi(ex, fl, ids, engine=engine)
            break
        except litdb.CompileError:
            continue
    assert db.outfix_engine_type() in ((23, 24) if engine == litdb.ENGINE_MCSHENG64 else (19, 20))
    lens = [0, 1, 2, 15, 16, 17, 31, 33, 64, 200, 577, 1500, 4096] + [rng.randint(1, 1600) for _ in range(150)]
    bufs = [regen.random_text(rng, ex, n, smp) for n in lens]
    if smp:
        bufs += [b[:max(0, len(b) - len(smp[k % len(smp)]))] + smp[k % len(smp)] for k, b in enumerate(bufs[:40])]
    got = gpu_batch(db, bufs)
    assert got == port_batch(db, bufs)
    sub = bufs[:13] + bufs[-30:]
    assert gpu_batch(db, sub) == ref_batch(ref, db, sub) == brute_batch(ex, fl, ids, sub)
    assert len(got) > 20


def test_both_widths_are_exercised():
    small = [b"abc", b"^GET /[^ ]*\\.php", b"x[0-9]{3}y"]
    big = [b"passwd\\s+[A-Za-z0-9_]{4,9}=", b"(login|admin)[^&]*&id=", b"User-Agent: [^\\r\\n]{8}", b"select.{1,12}from"]
    text = b"GET /a/b.php x123y abc PASSWD  user_1= login=1&id= select a,b from User-Agent: abcdefghij"
    for ex, want, engine in ((small, 19, litdb.ENGINE_MCSHENG), (big, 20, litdb.ENGINE_MCSHENG),
                             (small, 23, litdb.ENGINE_MCSHENG64), (big, 24, litdb.ENGINE_MCSHENG64)):
        fl = [1 if k % 2 else 0 for k in range(len(ex))]
        ids = [10 + k for k in range(len(ex))]
        db = litdb.compile_regex_multi(ex, fl, ids, engine=engine)
        assert db.outfix_engine_type() == want
        got = gpu_batch(db, [text, b"", text[3:]])
        assert got == brute_batch(ex, fl, ids, [text, b"", text[3:]]) and len(got) >= 3
        sc = gpu.alloc_scratch(db)     # one buffer through hs_gpu_scan: callbacks in (to, id) order
        one = []
        assert gpu.scan(db, text, sc, lambda i, f, t, fl_, cx: one.append((i, t)) or 0) == 0
        assert one == sorted(regen.brute_force(ex, fl, ids, text), key=lambda m: (m[1], m[0]))
        sc.close()


@pytest.mark.parametrize("seed", range(4))
def test_mcsheng_outfixes_in_multi_engine_and_mixed_databases(ref, seed):
    engine = litdb.ENGINE_MCSHENG64 if seed == 3 else litdb.ENGINE_MCSHENG
    ref = vbmi_ref(ref, engine)
    rng = random.Random(8700 + seed)
    ex, fl, ids, smp = regen.random_set(rng, rng.randint(4, 14), True)
    if seed % 2:
        ex = regen.with_eod(rng, ex, 0.3)
    db = litdb.compile_regex_rose(ex, fl, ids, engine=engine, dfa_states=[40, 300, 1000, 24][seed])
    bufs = [regen.random_text(rng, ex, rng.choice([0, 1, 30, 200, 1400]), smp) for _ in range(50)]
    if smp:
        bufs += [b[:max(0, len(b) - len(smp[k % len(smp)]))] + smp[k % len(smp)] for k, b in enumerate(bufs[:30])]
    want = ref_batch(ref, db, bufs)
    for pack in (1, 0, 2):
        sc = gpu.alloc_scratch(db)
        sc.set_option("pack", pack)
        assert gpu_batch(db, bufs, sc) == want, pack
        sc.close()
    assert want == port_batch(db, bufs) == brute_batch(ex, fl, ids, bufs)
    # literals + McSheng regexes in one database
    lits = [b"needle", b"hay", b"GET /"]
    mdb = litdb.compile_lit_multi(lits, [0, 1, 0], [900, 901, 902], full_rose=True, regexes=ex, regex_flags=fl, regex_ids=ids,
                                  regex_engine=engine, dfa_states=300)
    mb = [b + b" a needle in the HAY GET /x" for b in bufs[:25]]
    assert gpu_batch(mdb, mb) == ref_batch(ref, mdb, mb)
