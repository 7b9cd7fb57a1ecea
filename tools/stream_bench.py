"""Stream-mode throughput: the C2 literal set compiled HS_MODE_STREAM, the C2 packets dealt to F flows
(packet k belongs to flow k mod F), one hs_gpu_scan_streams call per round = the next packet of every flow.  End to end
from host packets (wall clock; the library gathers history + block per flow, copies, scans, delivers), with the
reference's own hs_scan_stream on one host core over a bounded sample of the same flows beside it, which is also the
parity check (same callbacks per flow).  One JSON line.

usage: python tools/stream_bench.py [MiB=64] [flows=4096] [literal|dfa]

dfa: the c3d set (6 Snort-like regexes as one McClellan16 DFA, bench.py --config c3d) in stream mode over the C3 HTTP
payloads, each payload cut into 256-byte blocks dealt round by round; the engine's state id per stream crosses the
blocks."""
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

from vectorscan_b200 import gpu, litdb, workloads  # noqa: E402


def main():
    mib = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    flows = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    nbytes = mib << 20
    kind = sys.argv[3] if len(sys.argv) > 3 else "literal"
    if kind == "dfa":
        ex, fl, ids = workloads.c3_patterns(max_patterns=6, templates=(2, 4, 5))
        sdb = litdb.compile_regex_multi(ex, fl, ids, engine=litdb.ENGINE_MCCLELLAN, mode=litdb.HS_MODE_STREAM)
        pflat, poff = workloads.c3_corpus(nbytes)
        # flow k = the k-th payload of at least 1 KiB; round r = its bytes [256 r, 256 r + 256)
        plen = np.diff(poff).astype(np.int64)
        pick = np.nonzero(plen >= 1024)[0][:flows]      # payloads of at least 1 KiB: 4 rounds of 256 bytes
        assert len(pick) == flows, "corpus too small for that many flows"
        rounds = 4
        starts = poff[:-1][pick].astype(np.int64)
        off = np.zeros(rounds * flows + 1, dtype=np.uint64)
        flat = np.empty(rounds * flows * 256, dtype=np.uint8)
        for r in range(rounds):
            idx = (starts + 256 * r)[:, None] + np.arange(256)[None, :]
            flat[r * flows * 256:(r + 1) * flows * 256] = pflat[idx].reshape(-1)
        off[:] = np.arange(rounds * flows + 1, dtype=np.uint64) * 256
        what = f"c3d set (6 regexes, one McClellan16 DFA) in HS_MODE_STREAM, {flows} HTTP payloads as flows, 256-byte blocks"
    else:
        lits, flags, ids = workloads.c2_patterns()
        sdb = litdb.compile_lit_multi(lits, flags, ids, mode=litdb.HS_MODE_STREAM)
        flat, off = workloads.c2_batch(nbytes, lits)
        rounds = (len(off) - 1) // flows
        what = f"C2 literal set (10000 literals) in HS_MODE_STREAM, {mib} MiB of 1500-byte packets dealt to {flows} flows"
    count = len(off) - 1
    L = gpu.lib()
    sc = gpu.alloc_scratch(sdb)
    streams = [gpu.Stream(sdb) for _ in range(flows)]
    sid = (C.c_void_p * flows)(*[s.ptr.value for s in streams])
    lens_all = np.diff(off).astype(np.uint32)
    ptr_all = (np.uint64(flat.ctypes.data) + off[:-1]).astype(np.uint64)
    hits = [0]
    per_flow = {}
    sample_flows = 8

    def on_event(i, f, t, fl, ctx):
        hits[0] += 1
        k = int(ctx or 0)
        if k < sample_flows:
            per_flow.setdefault(k, []).append((int(i), int(t)))
        return 0
    cb = gpu.MATCH_CB(on_event)
    null_cb = C.cast(None, gpu.MATCH_CB)
    ctxs = (C.c_void_p * flows)(*range(flows))

    def run(fn):
        scanned = 0
        for r in range(rounds):
            ptrs = (C.c_char_p * flows)()
            C.memmove(ptrs, ptr_all[r * flows:(r + 1) * flows].ctypes.data, flows * 8)
            lens = (C.c_uint * flows)()
            C.memmove(lens, lens_all[r * flows:(r + 1) * flows].ctypes.data, flows * 4)
            rv = L.hs_gpu_scan_streams(sid, ptrs, lens, flows, 0, sc.ptr, fn, ctxs)
            assert rv == 0, rv
            scanned += int(lens_all[r * flows:(r + 1) * flows].sum())
        return scanned
    run(cb)                      # recording pass (Python callbacks): parity
    secs = 1e9
    for _ in range(3):           # timed passes: NULL handler, as the reference arm below
        for s_ in streams:
            s_.reset()
        t0 = time.perf_counter()
        scanned = run(null_cb)
        secs = min(secs, time.perf_counter() - t0)
    # the reference on one core: the first `sample_flows` flows, all their rounds
    import refrt
    R = refrt.RefRuntime()
    rdb = R.deserialize(sdb.serialize())
    rsc = R.alloc_scratch(rdb)
    ref_hits = {}
    cpu_s, cpu_bytes = 0.0, 0
    for k in range(sample_flows):
        out = []
        fn = refrt.MATCH_CB(lambda i, f, t, fl, ctx: out.append((int(i), int(t))) or 0)
        blocks = [flat[int(off[r * flows + k]):int(off[r * flows + k + 1])].tobytes() for r in range(rounds)]
        for timed in (False, True):
            st = C.c_void_p()
            assert R.L.hs_open_stream(rdb, 0, C.byref(st)) == 0
            h = C.cast(None, refrt.MATCH_CB) if timed else fn
            t0 = time.perf_counter()
            for b in blocks:
                assert R.L.hs_scan_stream(st, b, len(b), 0, rsc, h, None) == 0
            if timed:
                cpu_s += time.perf_counter() - t0
                cpu_bytes += sum(len(b) for b in blocks)
            R.L.hs_close_stream(st, rsc, h, None)
        ref_hits[k] = out
    parity = all(per_flow.get(k, []) == ref_hits[k] for k in range(sample_flows))
    print(json.dumps({
        "workload": what + f", {rounds} rounds of one block per flow",
        "e2e_gbs": round(scanned / secs / 1e9, 2), "ms_per_round": round(1e3 * secs / rounds, 3), "matches": hits[0],
        "ref_hs_scan_stream_1core_gbs": round(cpu_bytes / cpu_s / 1e9, 3), "ref_isa": R.isa,
        "ref_sample": f"{sample_flows} flows x {rounds} blocks", "parity_on_sample": bool(parity),
        "note": "timed passes run with a NULL match handler on both sides; parity from a recording pass"}))


if __name__ == "__main__":
    main()
