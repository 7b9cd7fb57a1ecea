"""GPU tests (-m gpu) of 16-bit McClellan engines with WIDE states (mcclellan_internal.h:53-61, doWide16
mcclellan_common_impl.h:86-188): the device gives every chain position (and every Sherman state) its
table row back at upload (hs_gpu.cu build_wide_tables) and steps the engine with k_mcclellan.  Against the live reference
runtime, the C restatement (tests/test_nfadb.py pins it on the reference, callback order included) and the `re` brute
force.  Bit-exact: sorted (buffer, to, id) lists."""
import random

import pytest

import regen
from test_gpu_limex import gpu_batch, ref_batch
from test_gpu_mcclellan import port_batch
from test_nfadb import wide_case, wide_header
from vectorscan_b200 import gpu, litdb

pytestmark = pytest.mark.gpu


def brute_batch(ex, fl, ids, bufs):
    out = []
    for bi, b in enumerate(bufs):
        out += [(bi, i, t) for (i, t) in regen.brute_force(ex, fl, ids, b)]
    return sorted(out, key=lambda x: (x[0], x[2], x[1]))


@pytest.mark.parametrize("seed", range(5))
def test_wide_state_engines(ref, seed):
    rng = random.Random(9950 + seed)
    ex, fl, ids, texts = wide_case(rng)
    db = litdb.compile_regex_multi(ex, fl, ids, engine=litdb.ENGINE_MCCLELLAN_WIDE)
    assert wide_header(db)[:2] == (7, 1)
    # long buffers too: the texts joined (anchored patterns die early, the floating ones go on) and 16-byte edge lengths
    bufs = texts + [texts[3] + b" " * k + texts[5] for k in (0, 1, 15, 16, 17)] + [b"".join(texts)]
    got = gpu_batch(db, bufs)
    assert got == ref_batch(ref, db, bufs)
    assert got == port_batch(db, bufs)
    assert got == brute_batch(ex, fl, ids, bufs) and len(got) > 40
    sc = gpu.alloc_scratch(db)     # one buffer through hs_gpu_scan
    one = []
    assert gpu.scan(db, texts[7], sc, lambda i, f, t, fl_, cx: one.append((i, t)) or 0) == 0
    assert one == sorted(regen.brute_force(ex, fl, ids, texts[7]), key=lambda m: (m[1], m[0]))
    sc.close()


@pytest.mark.parametrize("seed", range(2))
def test_wide_state_engines_as_outfixes(ref, seed):
    """Multi-engine database whose DFA engines carry wide states, beside other engines."""
    rng = random.Random(9970 + seed)
    ex, fl, ids, texts = wide_case(rng)
    more, mfl, mids, smp = regen.random_set(rng, 6, True)
    ex2, fl2, ids2 = ex + more, fl + mfl, ids + [100 + i for i in mids]
    db = litdb.compile_regex_rose(ex2, fl2, ids2, engine=litdb.ENGINE_MCCLELLAN_WIDE, dfa_states=3000)
    bufs = texts[:80] + [regen.random_text(rng, more, rng.choice([1, 30, 200]), smp) for _ in range(40)]
    want = ref_batch(ref, db, bufs)
    for pack in (1, 0, 2):
        sc = gpu.alloc_scratch(db)
        sc.set_option("pack", pack)
        assert gpu_batch(db, bufs, sc) == want, pack
        sc.close()
    assert want == port_batch(db, bufs) == brute_batch(ex2, fl2, ids2, bufs)
