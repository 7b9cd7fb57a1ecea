#!/usr/bin/env python
"""bench.py -- headline benchmark: block-mode literal-set scanning throughput (GB/s, Mmatches/s).

  python bench.py --gpus N --steps K --warmup W [--config c2|c1] [--impl reference]

A "step" is one pass of the hot path over one batch of synthetic input.  Default workload (BASELINE.json
configs[1]): 10 000 literals (FDR) over a 64 MiB batch of 1500-byte packets, scanned as independent block
scans (hs_scan_vector shape).  Per rank, ROT batches (> L2 in total) are kept resident in HBM and scanned
in rotation so no step finds its input in L2.

  value        bytes scanned by all ranks / max-over-ranks device time, CUDA events around every step's
               device work (front end + confirm + Rose programs + arena + sort/dedupe), inputs resident in HBM
  e2e          the same metric through the reference-facing call hs_gpu_scan_vector_ex() with HOST buffers
               (pinned), wall clock around the call: H2D of the batch, scan, D2H of the sorted matches
  roofline     dominant kernel k_turbo: algorithmic bytes per launch / its CUDA-event time, vs MEASURED_PEAKS hbm
  cpu_baseline the unmodified reference runtime (oracle/_ref) on host cores, bounded sample
  --impl reference   the same workload on the reference's CPU hs_scan with all host threads
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

from vectorscan_b200 import litdb, shard, workloads as W  # noqa: E402

ROT = 4  # resident batches per rank: 4 x 64 MiB = 256 MiB > 126 MB of L2


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(config, nbytes, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, per launch, from the committed ncu --set full
    capture of this workload (profiles/ncu_traffic.json, which names the kernel and the capture it came from); None when
    no capture matches config, batch size AND kernel name -- a capture of another kernel says nothing about this one."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            for e in json.load(f):
                if e["config"] == config and e["bytes"] == nbytes and e.get("kernel") == kernel:
                    return e["dram_bytes_per_launch"]
    except Exception:
        pass
    return None


def make_workload(config, rank, nbytes, npat=6):
    """-> (database, number of pattern ids, [(flat, offsets)] x ROT, description)"""
    if config == "c3d":
        # the C3 templates that determinise (no counted gaps), as ONE McClellan DFA: the engine the reference compiler
        # prefers for outfixes; 6 patterns keep the 16-bit table in shared memory, 8 push it out to L2
        ex, fl, ids = W.c3_patterns(max_patterns=npat, templates=(2, 4, 5))
        batches = [W.c3_corpus(nbytes, seed=20260103 + 100 * rank + r) for r in range(ROT)]
        db = litdb.compile_regex_multi(ex, fl, ids, engine=int(os.environ.get("HS_BENCH_C3D_ENGINE", litdb.ENGINE_MCCLELLAN)))
        name = (f"c3d: {len(ex)} Snort-like regexes (C3 templates without counted gaps) as one McClellan"
                f"{8 if db.outfix_engine_type() == 6 else 16} DFA (single-outfix database) over a "
                f"{nbytes >> 20} MiB batch of 200-1400-byte synthetic HTTP payloads per step")
        return db, len(ex), batches, name
    if config == "c3":
        # C3 at its stated size: 2000 regexes.  Without the reference's host compiler the set is laid
        # out as a full-Rose database of outfix engines (hsb_compile_regex_rose): McClellan DFAs of <= `dfa_states`
        # states, LimEx NFAs for what does not determinise -- the shape hs_compile_multi gives patterns without literals.
        ex, fl, ids = W.c3_patterns(max_patterns=2000, state_budget=1 << 30)
        batches = [W.c3_corpus(nbytes, seed=20260103 + 100 * rank + r) for r in range(ROT)]
        db = litdb.compile_regex_rose(ex, fl, ids, dfa_states=int(os.environ.get("HS_BENCH_DFA_STATES", "1024")))
        nq = int.from_bytes(db.bytecode()[156:160], "little")
        name = (f"c3: {len(ex)} Snort-like regexes with character classes as {nq} outfix engines (McClellan / LimEx) of one "
                f"full-Rose database over a {nbytes >> 20} MiB batch of 200-1400-byte synthetic HTTP payloads per step")
        return db, len(ex), batches, name
    if config == "c3n":
        ex, fl, ids = W.c3_patterns()
        batches = [W.c3_corpus(nbytes, seed=20260103 + 100 * rank + r) for r in range(ROT)]
        name = (f"c3n: {len(ex)} Snort-like regexes with character classes in one LimEx512 NFA (single-outfix database; "
                f"as many of the C3 templates as 512 states hold) over a "
                f"{nbytes >> 20} MiB batch of 200-1400-byte synthetic HTTP payloads per step")
        return litdb.compile_regex_multi(ex, fl, ids), len(ex), batches, name
    if config == "c4":
        # C4 at its stated mix: 20000 patterns = 16000 literals + 4000 regexes in ONE full-Rose database (floating FDR
        # matcher + outfix engines, hsb_compile_mixed) over pcap-shaped packets
        lits, flags, ids = W.c4_patterns()
        ex, fl, _ = W.c3_patterns(seed=20260104, max_patterns=4000, state_budget=1 << 30)
        rids = [100000 + k for k in range(len(ex))]
        batches = [W.c4_corpus(nbytes, lits, seed=20260104 + 100 * rank + r) for r in range(ROT)]
        db = litdb.compile_lit_multi(lits, flags, ids, regexes=ex, regex_flags=fl, regex_ids=rids,
                                     dfa_states=int(os.environ.get("HS_BENCH_DFA_STATES", "1024")))
        nq = int.from_bytes(db.bytecode()[156:160], "little")
        name = (f"c4: 20000 mixed patterns = 16000 literals (floating FDR matcher) + {len(ex)} regexes ({nq} outfix engines) in "
                f"one full-Rose database over a {nbytes >> 20} MiB shard of pcap-shaped packets (64/576/1500 B = 40/20/40 %), "
                f"one block scan per packet")
        return db, 100000 + len(ex), batches, name
    if config == "c4l":
        lits, flags, ids = W.c4_patterns()
        batches = [W.c4_corpus(nbytes, lits, seed=20260104 + 100 * rank + r) for r in range(ROT)]
        name = (f"c4l (literal part of c4): 16000 literals over a {nbytes >> 20} MiB shard of "
                f"pcap-shaped packets (64/576/1500 B = 40/20/40 %), one block scan per packet")
        return litdb.compile_lit_multi(lits, flags, ids), len(lits), batches, name
    if config == "c1":
        lits, flags, ids = W.c1_patterns()
        batches = []
        for r in range(ROT):
            t = W.c1_text(nbytes, lits, seed=20260101 + 100 * rank + r)
            batches.append((t, np.array([0, len(t)], dtype=np.uint64)))
        name = f"c1: 1000 literals (len 4-16, a-z, 20% caseless) over one {nbytes >> 20} MiB text buffer per step"
    else:
        lits, flags, ids = W.c2_patterns()
        batches = [W.c2_batch(nbytes, lits, seed=20260102 + 100 * rank + r) for r in range(ROT)]
        name = f"c2: 10000 literals (len 3-12, printable) over a {nbytes >> 20} MiB batch of 1500-byte packets per step"
    return litdb.compile_lit_multi(lits, flags, ids), len(lits), batches, name


class ClockSampler:
    def __init__(self, dev):
        self.dev, self.samples, self.reasons, self.maxmhz = dev, [], set(), None
        self._stop = threading.Event()
        self._ready = threading.Event()
        self._go = threading.Event()
        self._t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        # NVML in-process (one init, then ~20 us per sample): spawning nvidia-smi inside the timed region, once per rank,
        # initialises the driver for every GPU and delays the launches being timed
        try:
            import pynvml as N
            N.nvmlInit()
            h = N.nvmlDeviceGetHandleByIndex(self.dev)
            self.maxmhz = float(N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM))
            get_reasons = getattr(N, "nvmlDeviceGetCurrentClocksEventReasons", None) or N.nvmlDeviceGetCurrentClocksThrottleReasons
            bits = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40, "sw_power_cap": 0x4}
            self._ready.set()
            self._go.wait()
            while not self._stop.is_set():
                self.samples.append(float(N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM)))
                r = int(get_reasons(h))
                for n, b in bits.items():
                    if r & b:
                        self.reasons.add(n)
                self._stop.wait(0.001)
            return
        except Exception:
            self._ready.set()
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.dev), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                self.samples.append(float(out[0]))
                self.maxmhz = float(out[1])
                for n, v in zip(names, out[2:]):
                    if v.strip().lower().startswith("active"):
                        self.reasons.add(n)
            except Exception:
                pass
            self._stop.wait(0.2)

    def start(self):
        self._t.start()
        self._ready.wait(timeout=10)   # NVML initialised before the timed region begins

    def go(self):
        self._go.set()

    def stop(self):
        self._stop.set()
        self._go.set()
        self._t.join(timeout=6)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.maxmhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def cpu_reference_run(ser, flat, off, threads, repeats, budget_bytes=None):
    """Times the unmodified reference runtime over (a prefix of) the batch the way hsbench does (oracle/cpu_bench.c:
    persistent pinned workers, each scanning the whole sample `repeats` times, clock around the passes only).
    Returns (GB/s over all threads, Mmatch/s, seconds, sample bytes, isa)."""
    import refrt
    isa = refrt.pick_isa()
    lib = os.path.join(ROOT, "oracle", "_ref", f"libhs_ref_{isa}.so")
    bl = C.CDLL(os.path.join(ROOT, "oracle", "liboracle_bench.so"))
    bl.cpu_bench_run.restype = C.c_double
    bl.cpu_bench_run.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int,
                                 C.c_int, C.POINTER(C.c_uint64)]
    count = len(off) - 1
    if budget_bytes is not None and int(off[-1]) > budget_bytes:
        count = max(1, int(np.searchsorted(off, budget_bytes, side="right")) - 1)
    nbytes = int(off[count])
    m = C.c_uint64()
    secs = bl.cpu_bench_run(lib.encode(), ser, len(ser), flat.ctypes.data, off.ctypes.data, count, threads, repeats,
                            C.byref(m))
    if secs <= 0:
        raise RuntimeError(f"cpu_bench_run failed: {secs}")
    gbs = nbytes * repeats * threads / secs / 1e9
    return gbs, m.value / secs / 1e6, secs, nbytes, isa


def workload_config(name):
    """The `config` object both arms print (identical, so the driver can tell they ran the same workload)."""
    return {"workload": name, "l2": f"inputs larger than L2: {ROT} resident batches scanned in rotation"}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU hs_scan loop on all host threads (rank 0 only)."""
    if rank != 0:
        return
    nbytes = args.bytes
    db, n_ids, batches, name = make_workload(args.config, 0, nbytes, args.patterns)
    ser = db.serialize()
    threads = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    # bounded sample per step: every worker scans `sample` bytes `repeats` times; repeats sized from one calibration
    # pass so that a step takes about half a second (hsbench's default is 20 passes)
    sample = min(nbytes, args.ref_sample_bytes)
    g0, _, s0, nb0, isa = cpu_reference_run(ser, batches[0][0], batches[0][1], threads, 1, sample)
    repeats = int(max(2, min(40, round(0.5 / max(s0, 1e-4)))))
    for w in range(args.warmup):
        cpu_reference_run(ser, batches[w % ROT][0], batches[w % ROT][1], threads, 1, sample)
    tot_s, tot_b, tot_m = 0.0, 0, 0
    for k in range(args.steps):
        flat, off = batches[k % ROT]
        gbs, mms, secs, nb, isa = cpu_reference_run(ser, flat, off, threads, repeats, sample)
        tot_s += secs
        tot_b += nb * repeats * threads
        tot_m += mms * secs
    value = tot_b / tot_s / 1e9
    g1, _, _, _, _ = cpu_reference_run(ser, batches[0][0], batches[0][1], 1, max(2, repeats // 2), sample)
    line = {
        "impl": "reference", "metric": "scan throughput", "value": round(value, 4), "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1e3 * tot_s / args.steps, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "mmatches_per_s": round(tot_m / tot_s, 4),
        "config": workload_config(name),
        "cpu_baseline": {"value": round(value, 4), "unit": "GB/s", "cores": threads, "kind": "reference",
                         "single_core_value": round(g1, 4), "parallel_efficiency": round(value / (threads * g1), 3),
                         "isa": isa,
                         "sample": f"hsbench-shaped: {threads} pinned workers, each scanning the first {sample >> 20} MiB of "
                                   f"the step's batch {repeats} times (hs_scan per buffer), clock around the passes only; "
                                   f"unmodified reference runtime, {isa} build"},
        "e2e": {"value": round(value, 4), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--config", default="c2", choices=["c1", "c2", "c3", "c3n", "c3d", "c4", "c4l"])
    ap.add_argument("--patterns", type=int, default=6, help="c3d: number of patterns in the DFA")
    ap.add_argument("--bytes", type=int, default=0, help="batch size (default: 64 MiB; c3: the 256 MiB corpus SURVEY 8d states)")
    ap.add_argument("--ref-sample-bytes", type=int, default=0)
    ap.add_argument("--cpu-sample-bytes", type=int, default=0)
    ap.add_argument("--frontend", type=int, default=0)
    ap.add_argument("--batches-per-step", type=int, default=0, help="batches queued back to back per step (0: about 1 GiB)")
    ap.add_argument("--pipeline", type=int, default=-1, help="host path: 0 one copy then scan, 1 chunked for gathered input, 2 always chunked")
    ap.add_argument("--no-verify", action="store_true", help="skip the sharded-equals-single check")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if not args.bytes:
        args.bytes = (256 << 20) if args.config == "c3" else (64 << 20)
    # bounded CPU samples (the NFA path is ~100x slower per byte on the CPU than the literal path)
    if not args.ref_sample_bytes:
        args.ref_sample_bytes = (1 << 20) if args.config in ("c3", "c4") else (8 << 20) if args.config in ("c3n", "c3d") else (64 << 20)
    if not args.cpu_sample_bytes:
        args.cpu_sample_bytes = (1 << 20) if args.config in ("c3", "c4") else (4 << 20) if args.config in ("c3n", "c3d") else (64 << 20)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from vectorscan_b200 import gpu

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the scanning runtime has no CPU fallback")
    torch.cuda.set_device(local_rank)
    gpu.lib().hs_gpu_set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    nbytes = args.bytes
    db, n_ids, batches, name = make_workload(args.config, rank, nbytes, args.patterns)
    sc = gpu.alloc_scratch(db)
    if args.frontend:
        sc.set_option("frontend", args.frontend)
    if os.environ.get("HS_GPU_ORDER"):
        sc.set_option("order", int(os.environ["HS_GPU_ORDER"]))  # engine kernels: buffers handed out by descending length (1) or as they come (0)
    if os.environ.get("HS_GPU_PACK"):
        sc.set_option("pack", int(os.environ["HS_GPU_PACK"]))   # multi-engine databases: 0 / 1 / 2 (csrc/hs_gpu.cu launch_outfixes)
    if args.pipeline >= 0:
        sc.set_option("pipeline", args.pipeline)
    corpora = [gpu.Corpus(flat, off) for flat, off in batches]

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput ("value") -------------------------------------------------------
    # A step = `per_step` batches queued back to back (hs_gpu_scan_corpora: one host round trip per step), so that a
    # step is >= 1 ms of device work and a launch gap on one rank does not decide the max over ranks; consecutive
    # members are different resident batches (ROT x bytes > L2).  Databases outside the sort-free path (regex engines)
    # scan one batch per step.
    per_step = args.batches_per_step or max(1, min(16, (1 << 30) // max(1, nbytes)))
    try:
        gpu.scan_corpora(db, [corpora[0], corpora[1 % ROT]], sc)
        train_ok = True
    except gpu.HsGpuError:
        train_ok, per_step = False, 1

    def step(k):
        if train_ok:
            members = [corpora[(k * per_step + j) % ROT] for j in range(per_step)]
            n = gpu.scan_corpora(db, members, sc)
            st = sc.stats()
            return n, st, st["train_ms"], sum(c.total for c in members), sum(c.count for c in members)
        c = corpora[k % ROT]
        n = gpu.scan_corpus(db, c, sc)
        st = sc.stats()
        return n, st, st["scan_ms"] + st["sort_ms"], c.total, c.count

    for w in range(max(args.warmup, 3)):
        step(w)
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    sampler.go()
    dev_ms = kern_ms = 0.0
    nmatch = raw = launches = nbytes_scanned = nbuffers = cands = probes = scans = 0
    t0 = time.perf_counter()
    for k in range(args.steps):
        n, st, ms, nb, nbuf = step(k)
        nmatch += n
        dev_ms += ms
        kern_ms += st["kernel_ms"]
        raw += st["raw_matches"]
        cands += st["candidates"]
        probes += st["confirms"]
        launches += st["kernel_launches"]
        scans += st["scan_launches"]
        nbytes_scanned += nb
        nbuffers += nbuf
    barrier()
    wall_s = time.perf_counter() - t0
    clocks = sampler.stop()

    # the only collective of the path: match counts (+ per-pattern histogram of the last step) over NCCL
    hist = torch.zeros(n_ids, dtype=torch.int64, device="cuda")
    gpu.lib().hs_gpu_histogram(sc.ptr, C.c_void_p(hist.data_ptr()), n_ids)
    tot_matches, tot_bytes, hist = shard.reduce_counts(nmatch, nbytes_scanned, hist, device="cuda")
    rank_ms = torch.tensor([dev_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        allms = [torch.zeros_like(rank_ms) for _ in range(world)]
        dist.all_gather(allms, rank_ms)
        allms = sorted(float(x[0]) for x in allms)
        mx = torch.tensor([wall_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        max_wall = float(mx[0])
    else:
        allms, max_wall = [dev_ms], wall_s
    max_dev_ms = allms[-1]
    tot_bytes, tot_matches = float(tot_bytes), float(tot_matches)
    value = tot_bytes / (max_dev_ms * 1e-3) / 1e9

    # ---- end to end through the C ABI with host buffers -----------------------------------------------
    # (a) the friendly shape: the batch back to back in page-locked memory (hs_gpu_host_alloc);
    # (b) the pcapscan shape (examples/pcapscan.cc:305-316): every packet its own pageable allocation slot.
    L = gpu.lib()
    flat0, off0 = batches[0]
    count0 = len(off0) - 1
    lens_np = np.diff(off0).astype(np.uint32)
    lens = (C.c_uint * count0)(*lens_np.tolist())
    mp, cnt = C.POINTER(gpu.Match)(), C.c_uint64()

    def e2e_measure(ptrs, steps):
        def once():
            rv = L.hs_gpu_scan_vector_ex(db.ptr, ptrs, lens, count0, 0, sc.ptr, C.byref(mp), C.byref(cnt))
            if rv != 0:
                raise RuntimeError(f"hs_gpu_scan_vector_ex -> {rv}")
            return cnt.value
        for _ in range(3):
            once()
        barrier()
        t0 = time.perf_counter()
        matches = 0
        brk = {"h2d_ms": 0.0, "scan_ms": 0.0, "sort_ms": 0.0, "d2h_ms": 0.0}
        for _ in range(steps):
            matches += once()
            st = sc.stats()
            for k in brk:
                brk[k] += st[k]
        barrier()
        secs = time.perf_counter() - t0
        brk = {k: round(v / steps, 4) for k, v in brk.items()}
        brk["wall_ms"] = round(1e3 * secs / steps, 4)
        t = torch.tensor([secs], dtype=torch.float64, device="cuda")
        if world > 1:
            # per-rank H2D time: at N = 8 the ranks share the host's memory bandwidth and PCIe switches -- the spread names
            # the limiting resource
            h = torch.tensor([brk["h2d_ms"]], dtype=torch.float64, device="cuda")
            hs = [torch.zeros_like(h) for _ in range(world)]
            dist.all_gather(hs, h)
            brk["rank_h2d_ms"] = [round(float(x[0]), 4) for x in hs]
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return world * steps * float(flat0.nbytes) / float(t[0]) / 1e9, matches, brk

    e2e_steps = max(3, min(args.steps, 20))
    hp = L.hs_gpu_host_alloc(int(flat0.nbytes))
    C.memmove(hp, flat0.ctypes.data, flat0.nbytes)
    ptrs = (C.c_char_p * count0)()
    ptr_arr = (np.uint64(hp) + off0[:-1]).astype(np.uint64)
    C.memmove(ptrs, ptr_arr.ctypes.data, count0 * 8)
    e2e_value, e2e_matches, brk = e2e_measure(ptrs, e2e_steps)
    d2h_per_step = int(e2e_matches / e2e_steps) * 16
    L.hs_gpu_host_free(hp)
    # (b): packets at a stride of their length rounded up to 64 + 64 in ordinary (pageable) memory
    slot = ((lens_np.astype(np.int64) + 63) // 64 + 1) * 64
    starts = np.concatenate([[0], np.cumsum(slot)[:-1]]).astype(np.int64)
    pageable = np.zeros(int(slot.sum()) + 64, dtype=np.uint8)
    idx = np.repeat(starts - off0[:-1].astype(np.int64), lens_np) + np.arange(int(off0[-1]), dtype=np.int64)
    pageable[idx] = flat0
    ptrs2 = (C.c_char_p * count0)()
    ptr_arr2 = (np.uint64(pageable.ctypes.data) + starts.astype(np.uint64)).astype(np.uint64)
    C.memmove(ptrs2, ptr_arr2.ctypes.data, count0 * 8)
    e2e2_value, e2e2_matches, brk2 = e2e_measure(ptrs2, max(3, e2e_steps // 2))
    assert e2e2_matches // max(3, e2e_steps // 2) == e2e_matches // e2e_steps, "pageable and pinned inputs disagree"

    # ---- N GPUs scanning a sharded corpus reproduce the 1-GPU match list and histogram -----------------
    verify = None
    if not args.no_verify:
        # every rank shards the SAME corpus: rank 0's first batch when it is at most 64 MiB, else a 64 MiB one built from
        # rank 0's seed on every rank (r2u: ranks > 0 used to rebuild a capped corpus while rank 0 kept its 256 MiB one)
        vbytes = min(nbytes, 64 << 20)
        vflat, voff = batches[0] if (rank == 0 and vbytes == nbytes) else make_workload(args.config, 0, vbytes, args.patterns)[2][0]
        if len(voff) - 1 >= world:
            bounds = shard.shard_bounds(voff, world)
            sflat, soff, first = shard.take_shard(vflat, voff, bounds, rank)
            sbufs = [sflat[int(soff[i]):int(soff[i + 1])].tobytes() for i in range(len(soff) - 1)]
            mine = gpu.scan_vector_ex(db, sbufs, sc)
            myhist = torch.zeros(n_ids, dtype=torch.int64, device="cuda")
            if len(mine):
                myhist.index_add_(0, torch.as_tensor(mine["id"].astype(np.int64), device="cuda"),
                                  torch.ones(len(mine), dtype=torch.int64, device="cuda"))
            if world > 1:
                dist.all_reduce(myhist, op=dist.ReduceOp.SUM)
                gathered = [None] * world
                dist.all_gather_object(gathered, [tuple(int(v) for v in m) for m in mine])
            else:
                gathered = [[tuple(int(v) for v in m) for m in mine]]
            if rank == 0:
                merged = shard.merge_shards(gathered, bounds)
                vb = [vflat[int(voff[i]):int(voff[i + 1])].tobytes() for i in range(len(voff) - 1)]
                single = gpu.scan_vector_ex(db, vb, sc)
                single_l = [tuple(int(v) for v in m) for m in single]
                shist = np.bincount(single["id"].astype(np.int64), minlength=n_ids)[:n_ids] if len(single) else np.zeros(n_ids, dtype=np.int64)
                verify = {"sharded_equals_single": merged == single_l,
                          "histogram_equals_single": bool((myhist.cpu().numpy() == shist).all()),
                          "matches": len(single_l), "shards": world, "bytes": int(voff[-1])}

    if rank == 0:
        peak, peak_src = peaks()
        n_launch = max(1, scans)   # launches of the dominant kernel inside the timed region
        kernel = gpu.KERNEL_NAMES.get(sc.stats()["front_end"], "?")
        alg_bytes = (nbytes_scanned + 16 * raw + 12 * nbuffers) / n_launch
        achieved = alg_bytes / (kern_ms / n_launch * 1e-3) / 1e9
        line = {
            "metric": "scan throughput", "value": round(value, 2), "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": round(max_dev_ms / args.steps, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "mmatches_per_s": round(tot_matches / (max_dev_ms * 1e-3) / 1e6, 2),
            "matches_per_step": int(tot_matches / args.steps / world),
            "wall_ms_per_step": round(1e3 * max_wall / args.steps, 4),
            "config": workload_config(name),
            "step": (f"{per_step} batches of {nbytes >> 20} MiB queued back to back (hs_gpu_scan_corpora), CUDA events around the "
                     f"whole train" if train_ok else f"one {nbytes >> 20} MiB batch (hs_gpu_scan_corpus)"),
            "rank_ms_per_step": {"min": round(allms[0] / args.steps, 4), "median": round(allms[len(allms) // 2] / args.steps, 4),
                                 "max": round(allms[-1] / args.steps, 4)},
            "diag": {"frontend": args.frontend, "candidates_per_kb": round(1024.0 * cands / max(1, nbytes_scanned), 3),
                     "confirm_probes_per_kb": round(1024.0 * probes / max(1, nbytes_scanned), 3), "hist_nonzero_ids": int((hist > 0).sum())},
            "clocks": clocks,
            "e2e": {"value": round(e2e_value, 3), "unit": "GB/s", "h2d_bytes_per_step": int(flat0.nbytes + off0.nbytes),
                    "d2h_bytes_per_step": d2h_per_step, "steps": e2e_steps,
                    "call": "hs_gpu_scan_vector_ex(host pinned buffers)", "per_step_ms": brk,
                    "pageable_scattered": {"value": round(e2e2_value, 3), "unit": "GB/s", "per_step_ms": brk2,
                                           "shape": "every packet in its own slot of ordinary (pageable) memory, as pcapscan holds them"}},
            "verify": verify,
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": kernel, "achieved": round(achieved, 2),
                         "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
                         "traffic": ncu_traffic(args.config, nbytes, kernel), "peak_source": peak_src, "launches": n_launch,
                         "alg_bytes_per_launch": int(alg_bytes), "kernel_ms": round(kern_ms / n_launch, 4)},
        }
        if not args.no_cpu_baseline:
            try:
                ser = db.serialize()
                threads = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
                sample = args.cpu_sample_bytes  # same sizing as the --impl reference arm
                g0, _, s0, nb1, isa = cpu_reference_run(ser, flat0, off0, threads, 1, sample)
                reps = int(max(2, min(40, round(1.0 / max(s0, 1e-4)))))
                gN, mN, sN, nbN, _ = cpu_reference_run(ser, flat0, off0, threads, reps, sample)
                g1, m1, s1, _, _ = cpu_reference_run(ser, flat0, off0, 1, max(2, reps // 2), sample)
                line["cpu_baseline"] = {"value": round(gN, 4), "unit": "GB/s", "cores": threads, "kind": "reference",
                                        "single_core_value": round(g1, 4), "parallel_efficiency": round(gN / (threads * g1), 3),
                                        "isa": isa,
                                        "sample": f"hsbench-shaped: {threads} pinned workers, each scanning the first {nb1 >> 20} MiB "
                                                  f"of the step's batch {reps} times (hs_scan per buffer), clock around the passes "
                                                  f"only; unmodified reference runtime, {isa} build"}
            except Exception as e:  # the oracle is test infrastructure: report, do not fail the bench
                line["cpu_baseline"] = {"value": None, "unit": "GB/s", "cores": 0, "kind": "reference", "sample": f"failed: {e}"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
