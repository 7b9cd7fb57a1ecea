"""Multi-GPU plumbing of the scan path: the buffer list is split into contiguous, byte-balanced
shards, one per rank (one process per GPU, full database replica each); nothing is exchanged during the scan.
The only collective is one all-reduce of [matches, bytes] + the per-pattern histogram at batch end (NCCL on the
GPU box, gloo in the CPU tests); match tuples stay on their rank and are concatenated by shard -- shards are
contiguous in buffer index, so the concatenation is already globally ordered.

Mirrors what pcapscan / hsbench do with threads: every hs_scan call is independent given (db, buffer)
(tools/hsbench/main.cpp:486-512)."""
import numpy as np


def shard_bounds(offsets, world):
    """offsets: uint64[count+1] of the flat corpus.  Returns int64[world+1] buffer indices: rank r owns buffers
    [b[r], b[r+1]).  Contiguous; cut points are the buffer boundaries nearest to equal byte shares."""
    offsets = np.asarray(offsets, dtype=np.uint64)
    count = len(offsets) - 1
    total = int(offsets[-1])
    b = np.zeros(world + 1, dtype=np.int64)
    b[world] = count
    for r in range(1, world):
        target = total * r // world
        k = int(np.searchsorted(offsets, target, side="left"))
        if k > 0 and k <= count and int(offsets[k]) - target > target - int(offsets[k - 1]):
            k -= 1
        b[r] = min(max(k, int(b[r - 1])), count)
    return b


def take_shard(flat, offsets, bounds, rank):
    """(flat bytes of the shard, its offsets rebased to 0, index of its first buffer)."""
    offsets = np.asarray(offsets, dtype=np.uint64)
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    start, end = int(offsets[lo]), int(offsets[hi])
    return flat[start:end], (offsets[lo:hi + 1] - offsets[lo]).astype(np.uint64), lo


def reduce_counts(n_matches, n_bytes, hist, device="cpu", group=None):
    """All-reduce (sum) of the batch-end counters.  hist: integer array/tensor per pattern id.  Returns
    (matches, bytes, hist tensor) summed over ranks; a no-op without an initialised process group."""
    import torch
    import torch.distributed as dist
    h = torch.as_tensor(hist, dtype=torch.int64, device=device).clone()
    t = torch.cat([torch.tensor([int(n_matches), int(n_bytes)], dtype=torch.int64, device=device), h])
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return int(t[0]), int(t[1]), t[2:]


def merge_shards(per_rank_matches, bounds):
    """per_rank_matches[r]: iterable of (buffer, id, to) with shard-local buffer indices, sorted by (buffer, to, id).
    Returns the global list, buffer indices rebased; globally sorted because shards are contiguous."""
    out = []
    for r, ms in enumerate(per_rank_matches):
        base = int(bounds[r])
        out.extend((int(b) + base, int(i), int(t)) for b, i, t in ms)
    return out


# ---- stream mode: flows, not buffers, are what shards ------------------------------------------------------------------
def stream_owner(flow_ids, world):
    """Rank that owns each flow.  A stream's blocks must all reach the same rank (its history / engine state lives with the
    hs_gpu_stream object there), so the split is by flow id, not by position in the batch: a multiplicative hash of the id,
    which spreads consecutive ids and skewed id spaces evenly.  flow_ids: integer array -> int64 array of ranks."""
    import numpy as np
    f = np.asarray(flow_ids, dtype=np.uint64)
    h = (f * np.uint64(0x9E3779B97F4A7C15)) >> np.uint64(33)
    return (h % np.uint64(world)).astype(np.int64)


def take_stream_round(flow_ids, blocks, world, rank):
    """The part of one round (flow_ids[k] sends blocks[k]) that belongs to `rank`: (indices into the round, flow ids, blocks),
    in the round's order."""
    own = stream_owner(flow_ids, world)
    idx = [k for k in range(len(blocks)) if own[k] == rank]
    return idx, [flow_ids[k] for k in idx], [blocks[k] for k in idx]
