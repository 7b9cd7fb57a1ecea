"""The link-time drop-in of INTEGRATION.md section 1 as real code: integration/hs_gpu_shim.c defines the reference's own
entry points (hs_alloc_scratch, hs_scan, ...) over libhs_gpu.so.  CPU: it compiles, warnings as errors, with the
reference's own src/hs_runtime.h included -- a prototype that differed from the reference's declaration would not.  GPU:
a caller written against the reference API only (tests/shim_caller.c, the scan half of examples/simplegrep.c:150-219)
is linked once with the unmodified reference runtime and once with the shim; both print the same matches."""
import os
import random
import re
import subprocess

import pytest

import litgen
import refrt
from vectorscan_b200 import litdb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "bin")


def test_shim_compiles_against_the_reference_header(tmp_path):
    hdr = "/root/reference/src/hs_runtime.h"
    if not os.path.exists(hdr):
        pytest.skip("reference sources not present on this machine")
    out = tmp_path / "libshim.so"
    cmd = ["gcc", "-std=c11", "-Wall", "-Wextra", "-Werror", "-shared", "-fPIC", os.path.join(ROOT, "integration", "hs_gpu_shim.c"),
           "-I" + os.path.join(ROOT, "include"), "-I/root/reference/src", "-DHS_SHIM_WITH_REFERENCE_HEADER",
           "-L" + os.path.join(ROOT, "vectorscan_b200"), "-lhs_gpu", "-o", str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_shim_exports_the_reference_entry_points():
    so = os.path.join(ROOT, "vectorscan_b200", "libhs_gpushim.so")
    syms = subprocess.run(["nm", "-D", "--defined-only", so], capture_output=True, text=True, check=True).stdout
    for name in ("hs_alloc_scratch", "hs_clone_scratch", "hs_scratch_size", "hs_free_scratch", "hs_scan", "hs_scan_vector",
                 "hs_open_stream", "hs_scan_stream", "hs_close_stream", "hs_reset_stream", "hs_copy_stream"):
        assert re.search(rf" T {name}$", syms, re.M), name


def run_caller(exe, db_path, text_path, budget=None):
    cmd = [exe, str(db_path), str(text_path)] + ([str(budget)] if budget else [])
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.strip().splitlines()
    rv = int(lines[-1].split()[-1])
    matches = [(int(m.group(1)), int(m.group(2))) for m in (re.match(r"Match for pattern (\d+) at offset (\d+)", l) for l in lines[:-1])]
    return rv, matches


def make_case(tmp_path, seed):
    rng = random.Random(seed)
    lits, flags, ids, text = litgen.random_case(rng, None, rng.choice([1, 30, 400]), bytes(range(32, 127)), 20000)
    db = litdb.compile_lit_multi(lits, flags, ids)
    dbp, tp = tmp_path / "db.bin", tmp_path / "in.txt"
    dbp.write_bytes(db.serialize())
    tp.write_bytes(text)
    return lits, flags, ids, text, dbp, tp


def test_caller_on_the_reference_runtime(tmp_path):
    exe = os.path.join(BIN, f"shim_caller_ref_{refrt.pick_isa()}")
    if not os.path.exists(exe):
        pytest.skip("tests/bin not built (build() with oracle/_ref present)")
    lits, flags, ids, text, dbp, tp = make_case(tmp_path, 1)
    rv, got = run_caller(exe, dbp, tp)
    assert rv == 0 and sorted(got) == litgen.brute_force(lits, flags, ids, text)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_same_caller_through_the_shim_on_the_gpu(tmp_path, seed):
    ref_exe, gpu_exe = os.path.join(BIN, f"shim_caller_ref_{refrt.pick_isa()}"), os.path.join(BIN, "shim_caller_gpu")
    if not os.path.exists(ref_exe) or not os.path.exists(gpu_exe):
        pytest.skip("tests/bin not built")
    lits, flags, ids, text, dbp, tp = make_case(tmp_path, seed)
    rv_r, ref = run_caller(ref_exe, dbp, tp)
    rv_g, got = run_caller(gpu_exe, dbp, tp)
    assert rv_r == rv_g == 0
    assert sorted(got) == sorted(ref) == litgen.brute_force(lits, flags, ids, text)
    assert [t for _, t in got] == sorted(t for _, t in got)          # delivered in offset order, like the reference
    # a handler that asks to stop: both return HS_SCAN_TERMINATED after the same number of callbacks
    if len(ref) > 3:
        rv_r, ref3 = run_caller(ref_exe, dbp, tp, 3)
        rv_g, got3 = run_caller(gpu_exe, dbp, tp, 3)
        assert rv_r == rv_g == -3 and len(ref3) == len(got3) == 3
