"""GPU tests (-m gpu): the CUDA path, called through the C ABI, against (a) the committed golden fixtures,
(b) the unmodified reference runtime (oracle/_ref) and the C oracle port on seeded inputs, (c) size-independent
properties at BASELINE.json's full sizes.  Bit-exact: sorted (id, to) multisets per buffer."""
import ctypes as C
import json
import os
import random

import numpy as np
import pytest

import litgen
import oracle_port
from vectorscan_b200 import gpu, litdb, workloads as W

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
FRONTENDS = [0, 1, 2, 3]  # chosen by the tables' estimate, native (database tables), k_turbo (8-byte rows), k_turbo2 (4-byte rows)


def gpu_scan(db, text, frontend=0):
    sc = gpu.alloc_scratch(db)
    if frontend:
        sc.set_option("frontend", frontend)
    got = []
    rv = gpu.scan(db, text, sc, lambda i, f, t, fl, c: got.append((i, t)) or 0)
    sc.close()
    assert rv == 0
    return got


def ref_scan(ref, db, bufs):
    rdb = ref.deserialize(db.serialize())
    rs = ref.alloc_scratch(rdb)
    out = []
    for bi, b in enumerate(bufs):
        rv, m = ref.scan(rdb, rs, b)
        assert rv == 0
        out += [(bi, i, t) for (i, t) in m]
    ref.L.hs_free_scratch(rs)
    ref.L.hs_free_database(rdb)
    return sorted(out, key=lambda x: (x[0], x[2], x[1]))


@pytest.mark.parametrize("frontend", FRONTENDS)
def test_reference_unit_vectors(frontend):
    with open(os.path.join(GOLD, "ref_unit_vectors.json")) as f:
        cases = json.load(f)
    engines = [dict(), dict(engine=litdb.ENGINE_FDR, fdr_domain=9, fdr_stride=1),
               dict(engine=litdb.ENGINE_FDR, fdr_domain=13, fdr_stride=2)]
    dbs, scs = {}, {}
    ran = 0
    for c in cases[::3]:
        lits = [x.encode("latin1") for x in c["lits"]]
        for k, opts in enumerate(engines):
            if opts.get("fdr_stride", 1) > len(lits[0]):
                continue
            key = (tuple(lits), tuple(c["ids"]), k)
            if key not in dbs:
                dbs[key] = litdb.compile_lit_multi(lits, c["flags"], c["ids"], **opts)
                scs[key] = gpu.alloc_scratch(dbs[key])
                scs[key].set_option("frontend", frontend)
            got = []
            rv = gpu.scan(dbs[key], bytes.fromhex(c["text"]), scs[key], lambda i, f, t, fl, cx: got.append((i, t)) or 0)
            assert rv == 0 and sorted(got) == sorted(tuple(x) for x in c["expect"]), (c["name"], opts)
            ran += 1
    assert ran > 600


@pytest.mark.parametrize("frontend", FRONTENDS)
def test_recorded_reference_runtime_cases(frontend):
    with open(os.path.join(GOLD, "ref_runtime_cases.json")) as f:
        cases = json.load(f)
    for c in cases:
        lits = [bytes.fromhex(x) for x in c["lits"]]
        db = litdb.compile_lit_multi(lits, c["flags"], c["ids"], **c["opts"])
        got = gpu_scan(db, bytes.fromhex(c["text"]), frontend)
        assert sorted(got, key=lambda x: (x[1], x[0])) == [tuple(x) for x in c["expect"]], c["opts"]


@pytest.mark.parametrize("frontend", FRONTENDS)
def test_random_databases_vs_reference_and_port(ref, frontend):
    rng = random.Random(31 + frontend)
    for n in (1, 3, 20, 48, 100, 1200):
        for alpha in (b"ab", b"abcdefgh", bytes(range(32, 127)), bytes(range(256))):
            for tl in (0, 5, 17, 600, 40000):
                lits, flags, ids, text = litgen.random_case(rng, None, n, alpha, tl)
                opts_list = [dict()]
                ml = min(min(len(l) for l in lits), 8)
                if len(lits) > 1:
                    d, s = rng.choice([9, 12, 14, 15]), rng.choice([1, 2, 4])
                    if s <= ml:
                        opts_list.append(dict(engine=litdb.ENGINE_FDR, fdr_domain=d, fdr_stride=s))
                    if len(lits) <= 48:
                        opts_list.append(dict(engine=litdb.ENGINE_TEDDY, teddy_masks=rng.randint(1, 4)))
                for opts in opts_list:
                    try:
                        db = litdb.compile_lit_multi(lits, flags, ids, **opts)
                    except litdb.CompileError:
                        continue
                    got = sorted(gpu_scan(db, text, frontend))
                    exp = [(i, t) for (_, i, t) in ref_scan(ref, db, [text])]
                    assert got == sorted(exp), (n, len(alpha), tl, opts)
                    if tl <= 600:
                        assert got == sorted(oracle_port.scan(db, text)[1])


@pytest.mark.parametrize("frontend", FRONTENDS)
def test_batch_of_ragged_buffers(ref, frontend):
    """hs_scan_vector-shaped batch: empty, 1-byte, ragged and maximum-size packets; per-buffer offsets."""
    rng = random.Random(77)
    lits, flags, ids, _ = litgen.random_case(rng, None, 500, bytes(range(32, 127)), 10)
    db = litdb.compile_lit_multi(lits, flags, ids)
    bufs = []
    for k in range(4000):
        ln = rng.choice([0, 0, 1, 2, 7, 8, 15, 16, 17, 64, 300, 1500, 1500, 9000])
        t = bytearray(rng.choice(bytes(range(32, 127))) for _ in range(ln))
        l = rng.choice(lits)
        if len(l) <= ln:
            p = rng.choice([0, ln - len(l), rng.randrange(0, ln - len(l) + 1)])
            t[p:p + len(l)] = l
        bufs.append(bytes(t))
    # literals straddling a packet boundary must not match
    l = max(lits, key=len)
    bufs.append(b"zz" + l[:len(l) // 2])
    bufs.append(l[len(l) // 2:] + b"zz")
    sc = gpu.alloc_scratch(db)
    sc.set_option("frontend", frontend)
    got = [(int(a), int(b), int(c)) for a, b, c in gpu.scan_vector_ex(db, bufs, sc)]
    assert got == ref_scan(ref, db, bufs)
    # callback delivery: buffer by buffer, and per-buffer contexts
    seen = []
    rv = gpu.scan_vector(db, bufs, sc, lambda i, f, t, fl, c: seen.append((i, t)) or 0)
    assert rv == 0 and seen == [(i, t) for (_, i, t) in got]
    # early termination: a non-zero return ends THAT buffer's deliveries (as it ends that packet's hs_scan in the loop the
    # batch call replaces); the other buffers are still served and the call returns HS_SCAN_TERMINATED
    first = []
    rv = gpu.scan_vector(db, bufs, sc, lambda i, f, t, fl, c: first.append((i, t)) or 1)
    heads, last_buf = [], None
    for (b, i, t) in got:
        if b != last_buf:
            heads.append((i, t))
            last_buf = b
    assert rv == gpu.HS_SCAN_TERMINATED and first == heads and len(heads) > 1
    st = sc.stats()
    assert st["buffers"] == len(bufs) and st["bytes"] == sum(map(len, bufs))


@pytest.mark.parametrize("frontend", FRONTENDS)
def test_singlematch_dedupe_and_shared_ids(ref, frontend):
    lits = [b"abc", b"bc", b"abcabc", b"XYZ", b"xyz", b"longer than eight bytes", b"than eight bytes"]
    flags = [8, 8, 0, 1, 0, 0, 9]
    ids = [5, 5, 6, 7, 7, 9, 9]
    flags[5] = 9  # ids sharing a report must agree on SINGLEMATCH
    db = litdb.compile_lit_multi(lits, flags, ids)
    text = b"..abcabc..xyz XYZ xYz...longer than eight bytes LONGER THAN EIGHT BYTES abc" * 3
    bufs = [text, text[5:], b"", text[:20]]
    sc = gpu.alloc_scratch(db)
    sc.set_option("frontend", frontend)
    got = [(int(a), int(b), int(c)) for a, b, c in gpu.scan_vector_ex(db, bufs, sc)]
    assert got == ref_scan(ref, db, bufs)


def test_flood_and_dense_matches_grow_the_arena(ref):
    db = litdb.compile_lit_multi([b"a", b"aa", b"aaaa", b"aaaaaaaaa", b"ab"], [0] * 5, [1, 2, 3, 4, 5])
    text = b"a" * 200000 + b"b" + b"a" * 1000
    sc = gpu.alloc_scratch(db)
    sc.set_option("arena", 1024)
    got = [(int(a), int(b), int(c)) for a, b, c in gpu.scan_vector_ex(db, [text], sc)]
    assert got == ref_scan(ref, db, [text])
    assert sc.stats()["arena_retries"] >= 1


def test_scratch_reuse_clone_and_two_databases(ref):
    db1 = litdb.compile_lit_multi([b"alpha", b"beta"], [0, 0], [1, 2])
    db2 = litdb.compile_lit_multi([b"gamma"], [1], [3])
    sc = gpu.alloc_scratch(db1)
    sc.grow(db2)  # one scratch may serve several databases
    t = b"alpha Gamma beta gamma"
    a, b = [], []
    assert gpu.scan(db1, t, sc, lambda i, f, to, fl, c: a.append((i, to)) or 0) == 0
    assert gpu.scan(db2, t, sc, lambda i, f, to, fl, c: b.append((i, to)) or 0) == 0
    assert a == [(1, 5), (2, 16)] and b == [(3, 11), (3, 22)]
    clone = C.c_void_p()
    assert gpu.lib().hs_gpu_clone_scratch(sc.ptr, C.byref(clone)) == 0
    hb, dbytes = C.c_size_t(), C.c_size_t()
    assert gpu.lib().hs_gpu_scratch_size(clone, C.byref(hb), C.byref(dbytes)) == 0 and dbytes.value > 0
    assert gpu.lib().hs_gpu_free_scratch(clone) == 0
    # NULL callback suppresses matches; NULL data is HS_INVALID
    assert gpu.lib().hs_gpu_scan(db1.ptr, t, len(t), 0, sc.ptr, C.cast(None, gpu.MATCH_CB), None) == 0
    assert gpu.lib().hs_gpu_scan(db1.ptr, None, 0, 0, sc.ptr, C.cast(None, gpu.MATCH_CB), None) == gpu.HS_INVALID


@pytest.mark.parametrize("config", ["c2", "c1"])
def test_full_size_properties(ref, config):
    """At BASELINE sizes the reference is too slow to run in full; check size-independent properties:
    (1) match count and per-id histogram equal the reference's on a 4 MiB prefix, (2) scanning the batch in two
    halves split at a buffer boundary gives the same multiset, (3) native and turbo front ends agree exactly,
    (4) output is sorted and free of duplicates, (5) every reported match re-verifies by direct comparison."""
    if config == "c2":
        lits, flags, ids = W.c2_patterns()
        flat, off = W.c2_batch(64 << 20, lits)
    else:
        lits, flags, ids = W.c1_patterns()
        flat = W.c1_text(256 << 20, lits)
        off = np.array([0, len(flat)], dtype=np.uint64)
    db = litdb.compile_lit_multi(lits, flags, ids)
    sc = gpu.alloc_scratch(db)
    res = {}
    for fe in FRONTENDS:
        sc.set_option("frontend", fe)
        corpus = gpu.Corpus(flat, off)
        gpu.scan_corpus(db, corpus, sc)
        res[fe] = gpu.fetch_matches(sc).copy()
        corpus.close()
    m = res[FRONTENDS[0]]
    assert len(m) > 1000
    for fe in FRONTENDS[1:]:
        assert np.array_equal(res[fe], m)
    key = np.stack([m["buffer"].astype(np.uint64), m["to"], m["id"].astype(np.uint64)], axis=1)
    order = np.lexsort((key[:, 2], key[:, 1], key[:, 0]))
    assert np.array_equal(order, np.arange(len(m)))           # sorted by (buffer, to, id)
    assert len(np.unique(key, axis=0)) == len(m)                # no duplicates (ids are distinct literals)
    # (5) re-verify a sample of matches byte for byte
    rng = np.random.default_rng(1)
    for k in rng.choice(len(m), size=2000, replace=False):
        b, i, to = int(m["buffer"][k]), int(m["id"][k]), int(m["to"][k])
        lit = lits[i]
        seg = bytes(flat[int(off[b]) + to - len(lit):int(off[b]) + to])
        assert to >= len(lit)
        assert (litgen.fold(seg) == litgen.fold(lit)) if flags[i] & 1 else seg == lit
    # (1) prefix against the reference
    if config == "c2":
        nb = int(np.searchsorted(off, 4 << 20))
        bufs = [bytes(flat[int(off[k]):int(off[k + 1])]) for k in range(nb)]
        exp = ref_scan(ref, db, bufs)
        got = [(int(a), int(b), int(c)) for a, b, c in zip(m["buffer"], m["id"], m["to"]) if a < nb]
        assert got == exp
        # (2) halves
        half = len(off) // 2
        c1 = gpu.Corpus(flat[:int(off[half])], off[:half + 1])
        c2 = gpu.Corpus(flat[int(off[half]):], off[half:] - off[half])
        n1 = gpu.scan_corpus(db, c1, sc)
        m1 = gpu.fetch_matches(sc).copy()
        n2 = gpu.scan_corpus(db, c2, sc)
        m2 = gpu.fetch_matches(sc).copy()
        m2["buffer"] += half
        assert n1 + n2 == len(m) and np.array_equal(np.concatenate([m1, m2]), m)
    else:
        exp = ref_scan(ref, db, [bytes(flat[:8 << 20])])
        got = [(int(a), int(b), int(c)) for a, b, c in zip(m["buffer"], m["id"], m["to"]) if c <= (8 << 20)]
        assert got == exp


@pytest.mark.parametrize("pipeline", [0, 1, 2])
def test_host_staging_paths_agree(ref, pipeline):
    """Host staging: one copy then one scan (0), or chunks copied on a second stream with the scans of earlier chunks
    running beside them (2; 1 = only for input that has to be gathered).  hs_gpu_scan_vector_ex takes buffers (a) scattered in pageable memory (gathered block by block into the
    scratch's pinned staging), (b) back to back in one page-locked region (copied straight from it), (c) a mix of
    both with empty buffers in between; more than one staging block (> 4096 buffers) in every case."""
    rng = random.Random(77)
    lits, flags, ids, _ = litgen.random_case(rng, None, 200, bytes(range(97, 107)), 8)
    db = litdb.compile_lit_multi(lits, flags, ids)
    nrng = np.random.default_rng(77)
    bufs = []
    for k in range(30000):  # ~9.4 MB: above the 8 MiB floor of the chunked ("pipeline") staging
        n = rng.choice([0, 1, 9, 60, 300, 1500])
        bufs.append(nrng.integers(97, 107, size=n, dtype=np.uint8).tobytes())
    exp = ref_scan(ref, db, bufs)
    sc = gpu.alloc_scratch(db)
    sc.set_option("pipeline", pipeline)
    as_list = lambda m: [(int(b), int(i), int(t)) for b, i, t in m]
    assert as_list(gpu.scan_vector_ex(db, bufs, sc)) == exp                         # (a)
    L = gpu.lib()
    total = sum(len(b) for b in bufs)
    hp = L.hs_gpu_host_alloc(total + 16)
    assert hp
    C.memmove(hp, b"".join(bufs), total)
    n = len(bufs)
    lens = (C.c_uint * n)(*[len(b) for b in bufs])
    offs = np.concatenate([[0], np.cumsum([len(b) for b in bufs])]).astype(np.uint64)
    mp, cnt = C.POINTER(gpu.Match)(), C.c_uint64()

    def run(ptr_values):
        ptrs = (C.c_char_p * n)()
        C.memmove(ptrs, np.asarray(ptr_values, dtype=np.uint64).ctypes.data, n * 8)
        assert L.hs_gpu_scan_vector_ex(db.ptr, ptrs, lens, n, 0, sc.ptr, C.byref(mp), C.byref(cnt)) == 0
        return [(int(mp[k].buffer), int(mp[k].id), int(mp[k].to)) for k in range(cnt.value)]

    pinned = np.uint64(hp) + offs[:-1]
    assert run(pinned) == exp                                                       # (b)
    keep = [C.create_string_buffer(b, len(b)) if len(b) else None for b in bufs]
    mixed = pinned.copy()
    for k in range(n):
        if (k // 1000) % 2 and keep[k] is not None:
            mixed[k] = C.addressof(keep[k])
    assert run(mixed) == exp                                                        # (c)
    L.hs_gpu_host_free(hp)
    sc.close()


@pytest.mark.parametrize("config", ["c3d", "c3", "c4"])
def test_full_size_properties_engines_and_c4(ref, config):
    """Bench-size batches of the other configs (bench.py make_workload): (1) the first 2 MiB of buffers equal the
    unmodified reference runtime, (2) the batch scanned in two halves split at a buffer boundary gives the same
    records, (3) a second scan of the same corpus gives the same records (no state leaks between scans), (4) output
    is sorted by (buffer, to, id) and free of duplicates."""
    import bench
    nbytes = {"c3d": 64 << 20, "c3": 16 << 20, "c4": 64 << 20}[config]
    db, n_ids, batches, name = bench.make_workload(config, 0, nbytes, 6)
    flat, off = batches[0]
    sc = gpu.alloc_scratch(db)
    corpus = gpu.Corpus(flat, off)
    n = gpu.scan_corpus(db, corpus, sc)
    m = gpu.fetch_matches(sc).copy()
    n_again = gpu.scan_corpus(db, corpus, sc)
    assert n_again == n and np.array_equal(gpu.fetch_matches(sc), m)
    corpus.close()
    assert n == len(m) and n > 100
    key = np.stack([m["buffer"].astype(np.uint64), m["to"], m["id"].astype(np.uint64)], axis=1)
    order = np.lexsort((key[:, 2], key[:, 1], key[:, 0]))
    assert np.array_equal(order, np.arange(len(m)))
    assert len(np.unique(key, axis=0)) == len(m)
    nb = int(np.searchsorted(off, 2 << 20))
    bufs = [bytes(flat[int(off[k]):int(off[k + 1])]) for k in range(nb)]
    exp = ref_scan(ref, db, bufs)
    got = [(int(a), int(b), int(c)) for a, b, c in zip(m["buffer"], m["id"], m["to"]) if a < nb]
    assert got == exp and len(exp) > 5
    half = len(off) // 2
    c1 = gpu.Corpus(flat[:int(off[half])], off[:half + 1])
    c2 = gpu.Corpus(flat[int(off[half]):], off[half:] - off[half])
    n1 = gpu.scan_corpus(db, c1, sc)
    m1 = gpu.fetch_matches(sc).copy()
    n2 = gpu.scan_corpus(db, c2, sc)
    m2 = gpu.fetch_matches(sc).copy()
    m2["buffer"] += half
    assert n1 + n2 == len(m) and np.array_equal(np.concatenate([m1, m2]), m)


@pytest.mark.parametrize("frontend", FRONTENDS)
def test_fat_teddy_databases(ref, frontend):
    """16-bucket (AVX2 "fat") Teddy databases: the native front end folds the 16 buckets onto 8 filter bits and confirms
    both buckets of a bit; the turbo front ends re-encode the LitInfo records as for any other matcher."""
    rng = random.Random(91 + frontend)
    ran = 0
    for trial in range(30):
        nl = rng.choice([2, 9, 17, 30, 60, 96])
        alpha = rng.choice([b"abcdefgh", bytes(range(32, 127)), bytes(range(256))])
        lits, flags, ids, text = litgen.random_case(rng, None, nl, alpha, rng.choice([0, 5, 40, 600, 20000]))
        try:
            db = litdb.compile_lit_multi(lits, flags, ids, engine=litdb.ENGINE_TEDDY, teddy_masks=rng.randint(1, 4), teddy_fat=1)
        except litdb.CompileError:
            continue
        got = sorted(gpu_scan(db, text, frontend))
        assert got == litgen.brute_force(lits, flags, ids, text), (nl, len(text))
        assert got == sorted((i, t) for (_, i, t) in ref_scan(ref, db, [text]))
        ran += 1
    assert ran > 15


def test_single_match_sets_are_resolved_on_the_device(ref):
    """HS_FLAG_SINGLEMATCH: the device array hs_gpu_scan_corpus leaves behind is already the resolved one (first record per
    (buffer, exhaustion key), quiet records dropped) -- count, fetched matches and hs_gpu_histogram agree with the
    reference runtime."""
    import ctypes as C
    import torch
    rng = random.Random(12)
    lits, flags, ids, _ = litgen.random_case(rng, None, 300, b"abcdefgh", 10)
    flags = [f | (8 if k % 2 else 0) for k, f in enumerate(flags)]
    by_id = {}
    for k, i in enumerate(ids):
        by_id.setdefault(i, flags[k] & 8)
        flags[k] = (flags[k] & 1) | by_id[i]
    db = litdb.compile_lit_multi(lits, flags, ids)
    bufs = [bytes(rng.choice(b"abcdefgh") for _ in range(rng.choice([0, 3, 200, 3000]))) for _ in range(300)]
    flat = np.frombuffer(b"".join(bufs), dtype=np.uint8)
    offs = np.concatenate([[0], np.cumsum([len(b) for b in bufs])]).astype(np.uint64)
    sc = gpu.alloc_scratch(db)
    n = gpu.scan_corpus(db, gpu.Corpus(flat, offs), sc)
    exp = ref_scan(ref, db, bufs)
    assert n == len(exp) and len(exp) > 1000
    hist = torch.zeros(max(ids) + 1, dtype=torch.int64, device="cuda")
    assert gpu.lib().hs_gpu_histogram(sc.ptr, C.c_void_p(hist.data_ptr()), max(ids) + 1) == 0
    got = [(int(b), int(i), int(t)) for b, i, t in gpu.fetch_matches(sc)]
    assert got == exp
    want = np.bincount([i for _, i, _ in exp], minlength=max(ids) + 1)
    assert (hist.cpu().numpy() == want).all()
