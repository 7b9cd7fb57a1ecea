#!/usr/bin/env python3
"""hsbench-shaped command line over the GPU runtime (the block-mode subset of /root/reference/tools/hsbench/main.cpp):

    python tools/hsbench_gpu.py -e PATTERNS -c CORPUS [-n 20] [-N] [--echo-matches] [--frontend K] [--write-corpus ...]

-e PATTERNS   pattern file in the reference's format, one `ID:/EXPRESSION/FLAGS` per line, `#` comments
              (util/ExpressionParser.rl, tools/hsbench/README / doc/dev-reference/tools.rst).  Flags understood: i
              (HS_FLAG_CASELESS), s (DOTALL), H (SINGLEMATCH); anything else is refused.  Expressions without regex
              syntax go to the literal matcher; the rest become outfix engines of the same database (hsb_compile_mixed),
              as far as the stand-in compiler's dialect goes (csrc/nfadb_writer.cpp).
-c CORPUS     hsbench's SQLite corpus (table chunk(id, stream_id, data), tools/hsbench/data_corpus.cpp:86-87; every row is
              one block) or any other file, cut into --block-size byte blocks.
-n NUMBER     repeat the scan NUMBER times (default 20, main.cpp:185); -N block mode (the default here); -V vectored mode
              (main.cpp:673-700: the blocks of one stream_id are ONE hs_scan_vector call; literal pattern files only, the
              database is compiled HS_MODE_VECTORED; with a plain file the whole file is one stream); -T is accepted and
              ignored (one GPU stream replaces the CPU thread list).
The corpus is made resident once (hs_gpu_corpus_create) and each repeat is one hs_gpu_scan_corpus -- what hsbench's
per-thread loop over blocks (main.cpp:486-512) is to the CPU runtime.  The summary lines are hsbench's (main.cpp:769-799)."""
import argparse
import os
import re
import sqlite3
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from vectorscan_b200 import gpu, litdb  # noqa: E402

FLAG_OF = {"i": litdb.HS_FLAG_CASELESS, "s": litdb.HS_FLAG_DOTALL, "H": litdb.HS_FLAG_SINGLEMATCH}
META = re.compile(rb"[\\.\[\](){}|*+?^$]")


def read_patterns(path):
    lits, regs = [], []
    with open(path, "rb") as f:
        for ln, line in enumerate(f, 1):
            line = line.strip()
            if not line or line.startswith(b"#"):
                continue
            m = re.match(rb"(\d+):/(.*)/([A-Za-z0-9]*)$", line)
            if not m:
                raise SystemExit(f"{path}:{ln}: expected ID:/EXPRESSION/FLAGS")
            pid, expr, fl = int(m.group(1)), m.group(2), m.group(3).decode()
            flags = 0
            for ch in fl:
                if ch not in FLAG_OF:
                    raise SystemExit(f"{path}:{ln}: flag '{ch}' is not supported by the GPU runtime")
                flags |= FLAG_OF[ch]
            (regs if META.search(expr) else lits).append((pid, expr, flags))
    return lits, regs


def read_corpus(path, block_size, with_streams=False):
    """-> blocks, or (with_streams) [(stream_id, block)] in id order."""
    with open(path, "rb") as f:
        head = f.read(16)
    if head.startswith(b"SQLite format 3"):
        con = sqlite3.connect(f"file:{path}?mode=ro", uri=True)
        rows = [(int(r[1]), bytes(r[2])) for r in con.execute("SELECT id, stream_id, data FROM chunk ORDER BY id")]
        con.close()
    else:
        data = open(path, "rb").read()
        rows = [(0, data[i:i + block_size]) for i in range(0, len(data), block_size)]
    return rows if with_streams else [b for _, b in rows]


def write_corpus(path, blocks, stream_ids=None):
    """The inverse of read_corpus for the SQLite shape (what tools/hsbench/scripts/*.py produce)."""
    con = sqlite3.connect(path)
    con.execute("CREATE TABLE chunk (id integer primary key, stream_id integer not null, data blob)")
    con.executemany("INSERT INTO chunk VALUES (?, ?, ?)", [(i, stream_ids[i] if stream_ids else i, b) for i, b in enumerate(blocks)])
    con.commit()
    con.close()


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("-e", dest="patterns", required=True)
    ap.add_argument("-c", dest="corpus", required=True)
    ap.add_argument("-n", dest="repeats", type=int, default=20)
    ap.add_argument("-N", dest="block", action="store_true")
    ap.add_argument("-V", dest="vectored", action="store_true")
    ap.add_argument("-T", dest="threads", default=None)
    ap.add_argument("--block-size", type=int, default=1500)
    ap.add_argument("--echo-matches", action="store_true")
    ap.add_argument("--frontend", type=int, default=0)
    ap.add_argument("--dfa-states", type=int, default=1024)
    args = ap.parse_args()

    lits, regs = read_patterns(args.patterns)
    t0 = time.perf_counter()
    if args.vectored:
        return vectored_main(args, lits, regs, t0)
    if lits:
        db = litdb.compile_lit_multi([e for _, e, _ in lits], [f & ~litdb.HS_FLAG_DOTALL for _, _, f in lits], [p for p, _, _ in lits],
                                     regexes=[e for _, e, _ in regs] or None, regex_flags=[f for _, _, f in regs],
                                     regex_ids=[p for p, _, _ in regs], dfa_states=args.dfa_states)
    elif regs:
        db = litdb.compile_regex_rose([e for _, e, _ in regs], [f for _, _, f in regs], [p for p, _, _ in regs], dfa_states=args.dfa_states)
    else:
        raise SystemExit("no patterns")
    print(f"Signatures:        {args.patterns}")
    print(f"Compile:           {len(lits)} literals + {len(regs)} regular expressions, {db.size} bytes, {time.perf_counter() - t0:.3f} seconds")
    print("Scan mode:         block (one GPU stream)")
    blocks = read_corpus(args.corpus, args.block_size)
    flat = np.frombuffer(b"".join(blocks), dtype=np.uint8)
    offs = np.concatenate([[0], np.cumsum([len(b) for b in blocks])]).astype(np.uint64)
    sc = gpu.alloc_scratch(db)
    if args.frontend:
        sc.set_option("frontend", args.frontend)
    corpus = gpu.Corpus(flat, offs)
    gpu.scan_corpus(db, corpus, sc)   # warm-up (also sizes the arenas)
    dev_ms, matches = 0.0, None
    t0 = time.perf_counter()
    for r in range(args.repeats):
        m = gpu.scan_corpus(db, corpus, sc)
        st = sc.stats()
        dev_ms += st["scan_ms"] + st["sort_ms"]
        if matches is not None and m != matches:
            print("\nWARNING: PER-SCAN MATCH COUNTS ARE INCONSISTENT!\n")
        matches = m
    secs = time.perf_counter() - t0
    if args.echo_matches:
        for rec in gpu.fetch_matches(sc):
            print(f"Match @ {int(rec['buffer'])}:{int(rec['to'])} for {int(rec['id'])}")
    nbytes = int(flat.nbytes)
    print(f"Time spent scanning:       {secs:,.3f} seconds ({dev_ms / 1e3:,.3f} on the device)")
    print(f"Corpus size:               {nbytes:,} bytes ({len(blocks):,} blocks)")
    print(f"Matches per iteration:     {matches:,} ({matches * 1024.0 / max(1, nbytes):,.3f} matches/kilobyte)")
    print(f"Overall block rate:        {len(blocks) * args.repeats / secs:,.2f} blocks/sec")
    print(f"Mean throughput (overall): {nbytes * args.repeats * 8 / secs / 1e6:,.2f} Mbit/sec")
    print(f"Device throughput:         {nbytes * args.repeats * 8 / (dev_ms * 1e-3) / 1e6:,.2f} Mbit/sec")


def vectored_main(args, lits, regs, t0):
    if regs or not lits:
        raise SystemExit("-V: vectored mode is served for literal pattern files only")
    db = litdb.compile_lit_multi([e for _, e, _ in lits], [f & ~litdb.HS_FLAG_DOTALL for _, _, f in lits], [p for p, _, _ in lits],
                                 mode=litdb.HS_MODE_VECTORED)
    print(f"Signatures:        {args.patterns}")
    print(f"Compile:           {len(lits)} literals, {db.size} bytes, {time.perf_counter() - t0:.3f} seconds")
    print("Scan mode:         vectored (one hs_scan_vector per stream, one GPU stream)")
    rows = read_corpus(args.corpus, args.block_size, with_streams=True)
    streams = {}
    for sid, b in rows:
        streams.setdefault(sid, []).append(b)
    sc = gpu.alloc_scratch(db)
    if args.frontend:
        sc.set_option("frontend", args.frontend)
    corpora = {}
    for sid, blocks in streams.items():
        offs = np.concatenate([[0], np.cumsum([len(b) for b in blocks])]).astype(np.uint64)
        if offs[-1]:
            corpora[sid] = gpu.Corpus(np.frombuffer(b"".join(blocks), dtype=np.uint8), offs)
    for c in corpora.values():
        gpu.scan_corpus(db, c, sc)        # warm-up
    dev_ms, matches = 0.0, 0
    t0 = time.perf_counter()
    for r in range(args.repeats):
        matches = 0
        for sid, c in corpora.items():
            matches += gpu.scan_corpus(db, c, sc)
            st = sc.stats()
            dev_ms += st["scan_ms"] + st["sort_ms"]
            if args.echo_matches and r == 0:
                for rec in gpu.fetch_matches(sc):
                    print(f"Match @ {sid}:{int(rec['to'])} for {int(rec['id'])}")
    secs = time.perf_counter() - t0
    nbytes = sum(len(b) for _, b in rows)
    print(f"Time spent scanning:       {secs:,.3f} seconds ({dev_ms / 1e3:,.3f} on the device)")
    print(f"Corpus size:               {nbytes:,} bytes ({len(rows):,} blocks in {len(streams):,} streams)")
    print(f"Matches per iteration:     {matches:,} ({matches * 1024.0 / max(1, nbytes):,.3f} matches/kilobyte)")
    print(f"Overall block rate:        {len(rows) * args.repeats / secs:,.2f} blocks/sec")
    print(f"Mean throughput (overall): {nbytes * args.repeats * 8 / secs / 1e6:,.2f} Mbit/sec")
    print(f"Device throughput:         {nbytes * args.repeats * 8 / max(dev_ms, 1e-6) / 1e3:,.2f} Mbit/sec")


if __name__ == "__main__":
    main()
