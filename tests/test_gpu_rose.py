"""GPU tests (-m gpu) of full-Rose databases of outfix engines (outfix slice): many McClellan / LimEx
engines on their own queues, every one stepping the whole batch on the device, one report pass, one finalize.  Against
the unmodified reference runtime (roseBlockExec -> blockInitSufPQ -> roseCatchUpTo), the C restatement and `re`."""
import random

import pytest

import regen
from test_gpu_limex import gpu_batch, ref_batch
from test_gpu_mcclellan import port_batch
from vectorscan_b200 import gpu, litdb, workloads as W

pytestmark = pytest.mark.gpu


def brute_batch(ex, fl, ids, bufs):
    out = []
    for bi, b in enumerate(bufs):
        out += [(bi, i, t) for (i, t) in regen.brute_force(ex, fl, ids, b)]
    return sorted(out, key=lambda x: (x[0], x[2], x[1]))


@pytest.mark.parametrize("engine,dfa_states", [(litdb.ENGINE_MCCLELLAN, 24), (litdb.ENGINE_MCCLELLAN, 300),
                                               (litdb.ENGINE_MCCLELLAN_SHERMAN, 600), (litdb.ENGINE_LIMEX, 0)])
@pytest.mark.parametrize("seed", [0, 1])
def test_multi_outfix_batches(ref, seed, engine, dfa_states):
    rng = random.Random(9100 + seed)
    ex, fl, ids, smp = regen.random_set(rng, rng.randint(2, 14), True)
    db = litdb.compile_regex_rose(ex, fl, ids, engine=engine, dfa_states=dfa_states)
    bufs = [regen.random_text(rng, ex, rng.choice([0, 1, 30, 200, 1400]), smp) for _ in range(60)]
    got = gpu_batch(db, bufs)
    assert got == ref_batch(ref, db, bufs)
    assert got == port_batch(db, bufs)
    assert got == brute_batch(ex, fl, ids, bufs)


def test_c3_shaped_set_of_several_hundred_patterns(ref):
    """The C3 templates (workloads.c3_patterns) at a few hundred patterns: dozens of engines, shared words, caseless."""
    ex, fl, ids = W.c3_patterns(max_patterns=300, state_budget=1 << 30)
    db = litdb.compile_regex_rose(ex, fl, ids, dfa_states=512)
    nq = int.from_bytes(db.bytecode()[156:160], "little")
    assert nq > 4
    flat, offs = W.c3_corpus(1 << 20)
    data = flat.tobytes()
    bufs = [data[int(offs[k]):int(offs[k + 1])] for k in range(len(offs) - 1)]
    got = gpu_batch(db, bufs)
    assert len(got) > 100
    assert got == ref_batch(ref, db, bufs)


def test_scan_one_buffer_callback_order(ref):
    ex = [b"abc[0-9]{2,5}x", b"foo.*bar", b"(login|admin)[^&]*&id=", b"^GET /[^ ]*\\.php"]
    db = litdb.compile_regex_rose(ex, [0, 2, 0, 0], [5, 6, 7, 8], dfa_states=40)
    text = b"GET /a.php login=1&id=2 abc123x foo abc12x bar admin&id="
    sc = gpu.alloc_scratch(db)
    got = []
    assert gpu.scan(db, text, sc, lambda i, f, t, fl, c: got.append((i, t)) or 0) == 0
    exp = sorted(regen.brute_force(ex, [0, 2, 0, 0], [5, 6, 7, 8], text), key=lambda x: (x[1], x[0]))
    assert got == exp


def test_packed_and_single_engine_launches_agree(ref):
    """Multi-engine databases, option "pack": 1 (default) = k_mcclellan_pack, work item = (buffer, engine); 2 = k_dfa_pack,
    several DFAs stepped together per buffer on a joint alphabet; 0 = one launch per engine.  Option "order": buffers
    handed out by descending length (default) or as they come.  Same matches every way, and as the reference runtime's."""
    ex, fl, ids = W.c3_patterns(max_patterns=400, state_budget=1 << 30)
    for dfa_states in (64, 256, 1024):      # McClellan8 packs of 8, mixed, McClellan16 pairs
        db = litdb.compile_regex_rose(ex, fl, ids, dfa_states=dfa_states)
        flat, offs = W.c3_corpus(1 << 19, seed=5 + dfa_states)
        data = flat.tobytes()
        bufs = [data[int(offs[k]):int(offs[k + 1])] for k in range(len(offs) - 1)] + [b"", b"x"]
        out, launches = {}, {}
        for pack in (2, 1, 0):
            sc = gpu.alloc_scratch(db)
            sc.set_option("pack", pack)
            sc.set_option("order", pack != 0)
            out[pack] = gpu_batch(db, bufs, sc)
            launches[pack] = sc.stats()["kernel_launches"]
            sc.close()
        assert out[2] == out[1] == out[0] and len(out[0]) > 50
        assert launches[2] < launches[0] and launches[1] < launches[0]      # fewer launches when packed
        if dfa_states == 256:
            assert out[2] == ref_batch(ref, db, bufs)


@pytest.mark.parametrize("seed", [1, 2, 3, 5])
def test_fused_packs_on_random_sets(ref, seed):
    """k_dfa_pack on seeded sets with small DFA groups (many engines per pack: D = 2, 4, 8), ragged buffers."""
    rng = random.Random(9300 + seed)
    ex, fl, ids, smp = regen.random_set(rng, rng.randint(6, 14), True)
    db = litdb.compile_regex_rose(ex, fl, ids, dfa_states=rng.choice([12, 24, 60]))
    bufs = [regen.random_text(rng, ex, rng.choice([0, 1, 15, 16, 17, 33, 200, 1400]), smp) for _ in range(80)]
    sc = gpu.alloc_scratch(db)
    sc.set_option("pack", 2)
    got = gpu_batch(db, bufs, sc)
    sc.close()
    assert got == ref_batch(ref, db, bufs)
    assert got == brute_batch(ex, fl, ids, bufs)


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_end_of_data_patterns_in_multi_engine_databases(ref, seed):
    """\\z patterns spread over several outfix engines (EOD program = ENGINES_EOD): every engine kernel raises its own
    end-of-data accepts; the finalize dedupes them against regular accepts at the same offset."""
    rng = random.Random(5100 + seed)
    ex, fl, ids, smp = regen.random_set(rng, rng.randint(3, 12), True)
    ex = regen.with_eod(rng, ex, 0.5)
    engine, ds = [(litdb.ENGINE_MCCLELLAN, 24), (litdb.ENGINE_MCCLELLAN, 300), (litdb.ENGINE_LIMEX, 0),
                  (litdb.ENGINE_MCCLELLAN_SHERMAN, 600)][seed % 4]
    db = litdb.compile_regex_rose(ex, fl, ids, engine=engine, dfa_states=ds)
    bufs = []
    for k in range(120):
        t = regen.random_text(rng, ex, rng.choice([0, 1, 20, 60, 140, 1400]), smp)
        if smp and k % 2 == 0:
            x = smp[k % len(smp)]
            t = t[:max(0, len(t) - len(x))] + x
        bufs.append(t)
    want = ref_batch(ref, db, bufs)
    for pack in (1, 0, 2):
        sc = gpu.alloc_scratch(db)
        sc.set_option("pack", pack)
        got = gpu_batch(db, bufs, sc)
        sc.close()
        assert got == want, pack
    assert got == brute_batch(ex, fl, ids, bufs) and len(got) > 10


def test_mixed_database_with_eod_regexes(ref):
    lits = [b"needle", b"hay", b"stack"]
    ex = [b"abc[0-9]+\\z", b"tail\\z", b"mid[a-z]{2}x"]
    db = litdb.compile_lit_multi(lits, [0, 1, 0], [1, 2, 3], regexes=ex, regex_flags=[0, 0, 0], regex_ids=[10, 11, 12], dfa_states=20)
    bufs = [b"needle abc123", b"HAY midqqx tail", b"stack tail x", b"", b"abc1", b"tailtail"]
    got = gpu_batch(db, bufs)
    assert got == ref_batch(ref, db, bufs) and len(got) >= 8
