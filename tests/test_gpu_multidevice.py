"""GPU tests that need two devices in ONE process (skipped on a single-GPU box): a scratch per device, each with its own
kernels' shared-memory opt-in (the attribute is per device: ADVICE r1), and a sharded corpus whose merged matches equal
the single-GPU scan (vectorscan_b200/shard.py -- the partitioning bench.py uses across ranks)."""
import random

import numpy as np
import pytest

import litgen
from vectorscan_b200 import gpu, litdb, shard, workloads as W

pytestmark = pytest.mark.gpu


def need_two():
    if gpu.lib().hs_gpu_device_count() < 2:
        pytest.skip("needs two GPUs in one process")


def test_scratch_per_device_in_one_process():
    need_two()
    rng = random.Random(2)
    lits, flags, ids, text = litgen.random_case(rng, None, 500, bytes(range(32, 127)), 200000)
    db = litdb.compile_lit_multi(lits, flags, ids)
    ex, fl, rid = W.c3_patterns(max_patterns=40, state_budget=1 << 30)
    rdb = litdb.compile_regex_rose(ex, fl, rid, dfa_states=256)
    outs = []
    for dev in (0, 1, 0):
        assert gpu.lib().hs_gpu_set_device(dev) == 0
        sc = gpu.alloc_scratch(db)
        got = []
        assert gpu.scan(db, text, sc, lambda i, f, t, fl_, c: got.append((i, t)) or 0) == 0
        sc.grow(rdb)
        rgot = [(int(b), int(i), int(t)) for b, i, t in gpu.scan_vector_ex(rdb, [text[:5000], text[5000:9000]], sc)]
        outs.append((sorted(got), rgot))
        sc.close()
    assert outs[0] == outs[1] == outs[2]
    assert outs[0][0] == litgen.brute_force(lits, flags, ids, text)
    gpu.lib().hs_gpu_set_device(0)


def test_sharded_corpus_on_two_devices_equals_single():
    need_two()
    lits, flags, ids = W.c2_patterns(2000)
    flat, offs = W.c2_batch(8 << 20, lits)
    db = litdb.compile_lit_multi(lits, flags, ids)
    gpu.lib().hs_gpu_set_device(0)
    sc0 = gpu.alloc_scratch(db)
    bufs = [flat[int(offs[i]):int(offs[i + 1])].tobytes() for i in range(len(offs) - 1)]
    single = [tuple(int(v) for v in m) for m in gpu.scan_vector_ex(db, bufs, sc0)]
    bounds = shard.shard_bounds(offs, 2)
    parts = []
    for rank in range(2):
        gpu.lib().hs_gpu_set_device(rank)
        sc = gpu.alloc_scratch(db)
        sflat, soff, first = shard.take_shard(flat, offs, bounds, rank)
        sb = [sflat[int(soff[i]):int(soff[i + 1])].tobytes() for i in range(len(soff) - 1)]
        parts.append([tuple(int(v) for v in m) for m in gpu.scan_vector_ex(db, sb, sc)])
        sc.close()
    gpu.lib().hs_gpu_set_device(0)
    assert shard.merge_shards(parts, bounds) == single and len(single) > 1000
