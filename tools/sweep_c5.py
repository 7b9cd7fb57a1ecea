"""C5 of BASELINE.json: hsbench-style sweep of pattern count x buffer size on one GPU (SURVEY.md 8d "C5 concrete
input"), literal sets (default) or regex sets (--regex: the C3 templates as a full-Rose database of outfix engines,
pattern counts 100 ... 2000).  Every cell: database from the literal writer, a corpus of `--bytes` bytes cut into equal
buffers, `--steps` device-timed scans through the C ABI (corpus resident in HBM, L2 defeated by rotating 3
corpora when they are smaller than L2), a parity check of the first `--check` buffers against the unmodified
reference runtime (oracle/_ref, when present: the checker, never the thing measured) and the reference's own
throughput on those buffers on one host core.  One JSON line per cell on stdout.

    python tools/sweep_c5.py --bytes $((128<<20)) --steps 10 > profiles/r01_c5_sweep.jsonl
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from vectorscan_b200 import gpu, litdb, workloads as W  # noqa: E402

PATTERNS = [100, 300, 1000, 3000, 10000, 30000, 50000]
BUFSIZES = [64, 256, 1500, 16 << 10, 256 << 10, 1 << 20]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bytes", type=int, default=128 << 20)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--check", type=int, default=64, help="buffers per cell compared with the reference runtime (at least)")
    ap.add_argument("--check-bytes", type=int, default=4 << 20, help="... and at least this many bytes per cell")
    ap.add_argument("--regex", action="store_true", help="regex rows: C3 templates as outfix engines of a full-Rose database")
    ap.add_argument("--dfa-states", type=int, default=1024)
    ap.add_argument("--patterns", type=int, nargs="*", default=PATTERNS)
    ap.add_argument("--bufsizes", type=int, nargs="*", default=BUFSIZES)
    args = ap.parse_args()
    gpu.lib().hs_gpu_set_device(0)
    ref = None
    try:
        import refrt
        ref = refrt.RefRuntime()
    except Exception:
        pass
    for ci, npat in enumerate(args.patterns):
        t0 = time.perf_counter()
        if args.regex:
            ex, fl, ids = W.c3_patterns(seed=20260105 + ci, max_patterns=npat, state_budget=1 << 30)
            db = litdb.compile_regex_rose(ex, fl, ids, dfa_states=args.dfa_states)
            lits = [w + b"=" for w in W._C3_WORDS]   # planted into the sweep text so that the engines have work to do
        else:
            lits, flags, ids = W.sweep_patterns(npat, 20260105 + ci)
            db = litdb.compile_lit_multi(lits, flags, ids)
        compile_s = time.perf_counter() - t0
        sc = gpu.alloc_scratch(db)
        rdb = rs = None
        if ref is not None:
            rdb = ref.deserialize(db.serialize())
            rs = ref.alloc_scratch(rdb)
        for bi, bufsize in enumerate(args.bufsizes):
            rot = 3 if args.bytes < (128 << 20) else 1   # 128 MiB > the 126 MB L2
            corpora, host = [], None
            for r in range(rot):
                if args.regex:
                    # the C3 corpus (synthetic HTTP payloads with the templates' words) cut into equal buffers
                    flat, _ = W.c3_corpus(args.bytes, seed=20260105 + 1000 * ci + 10 * bi + r)
                    off = np.minimum(np.arange(0, args.bytes + bufsize, bufsize, dtype=np.uint64), np.uint64(args.bytes))
                    off = np.unique(off)
                else:
                    flat, off = W.sweep_batch(args.bytes, bufsize, lits, 20260105 + 1000 * ci + 10 * bi + r)
                if r == 0:
                    host = (flat, off)
                corpora.append(gpu.Corpus(flat, off))
            for w in range(3):
                gpu.scan_corpus(db, corpora[w % rot], sc)
            dev_ms = kern_ms = 0.0
            nmatch = 0
            for k in range(args.steps):
                nmatch += gpu.scan_corpus(db, corpora[k % rot], sc)
                st = sc.stats()
                dev_ms += st["scan_ms"] + st["sort_ms"]
                kern_ms += st["kernel_ms"]
            line = {"patterns": npat, "buffer_bytes": bufsize, "buffers": len(host[1]) - 1, "corpus_bytes": args.bytes,
                    "steps": args.steps, "gbps_device": round(args.bytes * args.steps / (dev_ms * 1e-3) / 1e9, 2),
                    "gbps_kernel": round(args.bytes * args.steps / (kern_ms * 1e-3) / 1e9, 2),
                    "matches_per_step": nmatch // args.steps, "db_bytes": len(db.bytecode()),
                    "kind": "regex (outfix engines)" if args.regex else "literal",
                    "engines": int.from_bytes(db.bytecode()[156:160], "little"),
                    "compile_s": round(compile_s, 2)}
            # parity: the first buffers of corpus 0, GPU (through hs_gpu_scan_vector_ex, host buffers) vs reference
            if ref is not None and args.check:
                flat, off = host
                n = min(args.check, len(off) - 1)
                while n < len(off) - 1 and int(off[n]) < args.check_bytes // (40 if args.regex else 1):
                    n += 1   # (the regex databases run ~100 engines per byte on the CPU: a smaller byte budget)
                bufs = [bytes(flat[int(off[i]):int(off[i + 1])]) for i in range(n)]
                got = [(int(b), int(i), int(t)) for b, i, t in gpu.scan_vector_ex(db, bufs, sc)]
                exp = []
                t0 = time.perf_counter()
                for bidx, b in enumerate(bufs):
                    rv, m = ref.scan(rdb, rs, b)
                    exp += [(bidx, i, t) for (i, t) in m]
                ref_s = time.perf_counter() - t0
                exp.sort(key=lambda x: (x[0], x[2], x[1]))
                line["parity_buffers"] = n
                line["parity_matches"] = len(exp)
                line["parity"] = "ok" if got == exp else "MISMATCH"
                line["ref_1core_gbps_incl_ctypes"] = round(sum(len(b) for b in bufs) / ref_s / 1e9, 3)
            print(json.dumps(line), flush=True)
            for c in corpora:
                c.close()
        sc.close()


if __name__ == "__main__":
    main()
