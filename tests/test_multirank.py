"""World-size-2 run of the multi-GPU plumbing on the CPU (gloo): shard split, the one all-reduce of the path, and
the shard merge (vectorscan_b200/shard.py).  Each rank scans its shard with the reference runtime
standing in for the device scan (oracle = checker here, the GPU tests cover the device scan itself); rank 0
compares the merged result and the reduced counters with one unsharded scan."""
import os
import random
import socket

import numpy as np
import pytest

import litgen
from vectorscan_b200 import litdb, shard


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _corpus(seed=3):
    rng = random.Random(seed)
    lits, flags, ids, _ = litgen.random_case(rng, None, 60, b"abcdefgh", 6)
    bufs = []
    for k in range(97):
        _, _, _, t = litgen.random_case(random.Random(seed * 1000 + k), None, 1, b"abcdefgh", rng.choice([0, 5, 64, 700, 1500]))
        bufs.append(bytes(t))
    off = np.zeros(len(bufs) + 1, dtype=np.uint64)
    off[1:] = np.cumsum([len(b) for b in bufs])
    return lits, flags, ids, b"".join(bufs), off


def _scan_shard(ser, flat, off):
    import refrt
    R = refrt.RefRuntime()
    rdb = R.deserialize(ser)
    rs = R.alloc_scratch(rdb)
    out = []
    for b in range(len(off) - 1):
        rv, m = R.scan(rdb, rs, flat[int(off[b]):int(off[b + 1])])
        assert rv == 0
        out += [(b, i, t) for (i, t) in sorted(m, key=lambda x: (x[1], x[0]))]
    return out


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lits, flags, ids, flat, off = _corpus()
    ser = litdb.compile_lit_multi(lits, flags, ids).serialize()
    bounds = shard.shard_bounds(off, world)
    sflat, soff, first = shard.take_shard(flat, off, bounds, rank)
    mine = _scan_shard(ser, sflat, soff)
    hist = np.zeros(max(ids) + 1, dtype=np.int64)
    for _, i, _ in mine:
        hist[i] += 1
    tot_m, tot_b, tot_h = shard.reduce_counts(len(mine), int(soff[-1]), hist)
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    if rank == 0:
        q.put((tot_m, tot_b, tot_h.tolist(), shard.merge_shards(gathered, bounds), bounds.tolist()))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_bounds_properties():
    rng = np.random.default_rng(0)
    for world in (1, 2, 3, 8):
        for _ in range(20):
            lens = rng.integers(0, 3000, rng.integers(1, 200))
            off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
            b = shard.shard_bounds(off, world)
            assert b[0] == 0 and b[-1] == len(lens) and np.all(np.diff(b) >= 0)
            if len(lens) >= 8 * world and off[-1] > 0:
                share = np.diff(off[b].astype(np.int64))
                assert share.max() - share.min() <= 2 * 3000  # balanced to within a buffer or two
    # fewer buffers than ranks: some shards are empty, none is lost
    b = shard.shard_bounds(np.array([0, 10, 20], dtype=np.uint64), 8)
    assert b[0] == 0 and b[-1] == 2 and np.all(np.diff(b) >= 0)


def test_world_size_2_gloo(ref):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    tot_m, tot_b, tot_h, merged, bounds = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    lits, flags, ids, flat, off = _corpus()
    whole = _scan_shard(litdb.compile_lit_multi(lits, flags, ids).serialize(), flat, off)
    assert 0 < bounds[1] < len(off) - 1
    assert merged == whole and tot_m == len(whole) > 100 and tot_b == int(off[-1])
    hist = np.zeros(max(ids) + 1, dtype=np.int64)
    for _, i, _ in whole:
        hist[i] += 1
    assert tot_h == hist.tolist()


# ---- stream mode: flows shard by id; each rank keeps its flows' streams; one all-reduce of the counters ----------------
def _flows(seed=5):
    rng = random.Random(seed)
    lits, flags, ids, _ = litgen.random_case(rng, None, 40, b"abcdefgh", 6)
    flows = {}
    for f in range(37):
        text = bytes(rng.choice(b"abcdefgh") for _ in range(rng.choice([3, 90, 700])))
        n = rng.choice([1, 2, 5])
        cuts = sorted(rng.randrange(0, len(text) + 1) for _ in range(n - 1))
        flows[1000 + 7 * f] = [text[a:b] for a, b in zip([0] + cuts, cuts + [len(text)])]
    return lits, flags, ids, flows


def _stream_worker(rank, world, port, q):
    import oracle_port
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lits, flags, ids, flows = _flows()
    db = litdb.compile_lit_multi(lits, flags, ids, mode=litdb.HS_MODE_STREAM)
    seen = {}                                  # flow id -> blocks this rank was handed, in arrival order
    for rnd in range(5):                       # round = the next block of every flow that still has one
        fids = [f for f in sorted(flows) if rnd < len(flows[f])]
        _, mine, blocks = shard.take_stream_round(fids, [flows[f][rnd] for f in fids], world, rank)
        for f, b in zip(mine, blocks):
            seen.setdefault(f, []).append(b)
    got, n_bytes = {}, 0
    for f, blocks in seen.items():             # the port's block-by-block stream scan stands in for hs_gpu_scan_streams
        rv, m = oracle_port.scan_stream_blocks(db, blocks)
        assert rv == 0
        got[f] = m
        n_bytes += sum(len(b) for b in blocks)
    hist = np.zeros(max(ids) + 1, dtype=np.int64)
    for m in got.values():
        for i, _ in m:
            hist[i] += 1
    tot_m, tot_b, tot_h = shard.reduce_counts(sum(len(m) for m in got.values()), n_bytes, hist)
    gathered = [None] * world
    dist.all_gather_object(gathered, got)
    if rank == 0:
        q.put((tot_m, tot_b, tot_h.tolist(), gathered))
    dist.barrier()
    dist.destroy_process_group()


def test_stream_owner_is_total_and_balanced():
    for world in (1, 2, 3, 8):
        ids = np.arange(100000, dtype=np.uint64) * 4 + 17          # a skewed id space
        own = shard.stream_owner(ids, world)
        assert own.min() >= 0 and own.max() < world
        share = np.bincount(own, minlength=world) / len(ids)
        assert share.max() - share.min() < 0.02
        assert np.array_equal(own, shard.stream_owner(ids, world))  # a flow always lands on the same rank


def test_two_ranks_stream_flows_gloo():
    import multiprocessing as mp
    import oracle_port
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_stream_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    tot_m, tot_b, tot_h, gathered = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    lits, flags, ids, flows = _flows()
    db = litdb.compile_lit_multi(lits, flags, ids, mode=litdb.HS_MODE_STREAM)
    merged = {}
    for part in gathered:
        assert not (set(part) & set(merged))                        # every flow on exactly one rank
        merged.update(part)
    assert set(merged) == set(flows) and all(len(p) > 0 for p in gathered)
    n = 0
    for f, blocks in flows.items():
        want = litgen.brute_force(lits, flags, ids, b"".join(blocks))
        assert sorted(merged[f]) == want and oracle_port.scan_stream_blocks(db, blocks) == (0, merged[f])
        n += len(want)
    assert tot_m == n and tot_b == sum(len(b) for bl in flows.values() for b in bl)
    hist = [0] * (max(ids) + 1)
    for m in merged.values():
        for i, _ in m:
            hist[i] += 1
    assert tot_h == hist and n >= 20
