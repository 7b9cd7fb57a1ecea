"""Vectored-mode throughput: the C2 literal set compiled HS_MODE_VECTORED and the C2 packets scanned
as ONE stream by hs_gpu_scan_vector -- resident (device time, CUDA events) and end to end from page-locked host packets
(wall time, copies inside) -- with the reference's own hs_scan_vector on one host core beside it (one stream is one
sequential scan on the CPU: src/runtime.c:1143-1153 walks the blocks in order with history carried over).  Parity: the
device's match list == the reference's on a bounded prefix of the stream.  One JSON line.

usage: python tools/vectored_bench.py [MiB=64] [steps=20]"""
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
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    nbytes = mib << 20
    lits, flags, ids = workloads.c2_patterns()
    vdb = litdb.compile_lit_multi(lits, flags, ids, mode=4)
    sc = gpu.alloc_scratch(vdb)
    L = gpu.lib()
    # four different streams in rotation (256 MiB > the 126 MB L2 between two scans of the same one)
    streams = [workloads.c2_batch(nbytes, lits, seed=20260102 + 31 * k) for k in range(4)]
    corpora = [gpu.Corpus(f, o) for f, o in streams]
    for c in corpora:
        gpu.scan_corpus(vdb, c, sc)
    dev_ms, n_matches = 0.0, 0
    for k in range(steps):
        n_matches = gpu.scan_corpus(vdb, corpora[k % 4], sc)
        st = sc.stats()
        dev_ms += st["scan_ms"] + st["sort_ms"]
    dev_gbs = steps * nbytes / (dev_ms * 1e-3) / 1e9
    # end to end: packets back to back in page-locked memory, one hs_gpu_scan_vector_ex per step
    flat0, off0 = streams[0]
    count0 = len(off0) - 1
    lens = (C.c_uint * count0)(*np.diff(off0).astype(np.uint32).tolist())
    hp = L.hs_gpu_host_alloc(int(flat0.nbytes))
    C.memmove(hp, flat0.ctypes.data, flat0.nbytes)
    ptrs = (C.c_char_p * count0)()
    pa = (np.uint64(hp) + off0[:-1]).astype(np.uint64)
    C.memmove(ptrs, pa.ctypes.data, count0 * 8)
    mp, cnt = C.POINTER(gpu.Match)(), C.c_uint64()

    def once():
        rv = L.hs_gpu_scan_vector_ex(vdb.ptr, ptrs, lens, count0, 0, sc.ptr, C.byref(mp), C.byref(cnt))
        assert rv == 0, rv
        return cnt.value
    for _ in range(3):
        once()
    t0 = time.perf_counter()
    for _ in range(steps):
        e2e_matches = once()
    e2e_gbs = steps * nbytes / (time.perf_counter() - t0) / 1e9
    got = gpu._matches_to_list(mp, cnt.value)
    # the reference on one core: a bounded prefix of the same stream (about 8 MiB), also the parity check
    import refrt
    R = refrt.RefRuntime()
    rdb = R.deserialize(vdb.serialize())
    rsc = R.alloc_scratch(rdb)
    npk = min(count0, (8 << 20) // 1500)
    sample = int(off0[npk])
    hits = []
    cb = refrt.MATCH_CB(lambda i, f, t, fl, ctx: hits.append((i, t)) or 0)
    ptrs_r = (C.c_char_p * npk)()
    C.memmove(ptrs_r, pa.ctypes.data, npk * 8)
    assert R.L.hs_scan_vector(rdb, ptrs_r, lens, npk, 0, rsc, cb, None) == 0        # recording pass: parity
    cpu_s = 1e9
    for _ in range(3):      # timed passes: NULL handler (the reference substitutes null_onEvent, src/runtime.c:127)
        t0 = time.perf_counter()
        rv = R.L.hs_scan_vector(rdb, ptrs_r, lens, npk, 0, rsc, C.cast(None, refrt.MATCH_CB), None)
        cpu_s = min(cpu_s, time.perf_counter() - t0)
        assert rv == 0
    mine = sorted((int(i), int(t)) for _, i, t in got if int(t) <= sample)
    parity = sorted(hits) == mine
    L.hs_gpu_host_free(hp)
    print(json.dumps({
        "workload": f"C2 literal set (10000 literals) in HS_MODE_VECTORED, {mib} MiB of 1500-byte packets as one stream",
        "device_gbs": round(dev_gbs, 1), "e2e_gbs": round(e2e_gbs, 1), "matches": int(n_matches), "e2e_matches": int(e2e_matches),
        "ref_hs_scan_vector_1core_gbs": round(sample / cpu_s / 1e9, 3), "ref_isa": R.isa, "ref_sample_bytes": sample,
        "ref_callbacks": len(hits), "parity_on_sample": bool(parity), "steps": steps}))


if __name__ == "__main__":
    main()
