"""CPU test of the literal front end's upload-time tables (csrc/turbo_tables.h): evaluated on the host the way
k_turbo2 reads them (tests/turbo_emul.cpp), neither filter tier may drop a position where a literal really ends --
for the 4-byte union rows of k_turbo2, caseless and shortest-possible literals included."""
import ctypes as C
import os
import random
import subprocess

import numpy as np
import pytest

import litgen
from vectorscan_b200 import litdb, workloads as W

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def emul(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("emul") / "libturbo_emul.so")
    subprocess.run(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-I" + os.path.join(ROOT, "vectorscan_b200", "csrc"),
                    "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "turbo_emul.cpp"), "-o", so], check=True)
    L = C.CDLL(so)
    L.turbo_emul_check.restype = C.c_long
    L.turbo_emul_check.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.c_void_p, C.c_uint32, C.c_void_p]
    return L


def check(emul, lits, flags, ids, text, wide):
    os.environ["HS_GPU_TURBO_WIDE"] = str(wide)
    try:
        db = litdb.compile_lit_multi(lits, [f & 1 for f in flags], ids)
        ends = np.array(sorted({to - 1 for (_, to) in litgen.brute_force(lits, [f & 1 for f in flags], ids, text)}), dtype=np.uint64)
        stats = np.zeros(8, dtype=np.uint64)
        dropped = emul.turbo_emul_check(db.ptr, text, len(text), ends.ctypes.data, len(ends), stats.ctypes.data)
    finally:
        del os.environ["HS_GPU_TURBO_WIDE"]
    assert dropped == 0, (dropped, len(ends), stats[:5])
    return len(ends), [int(x) for x in stats[:5]]


@pytest.mark.parametrize("wide", [0])
def test_no_literal_end_is_filtered_out(emul, wide):
    total_ends, forms = 0, set()
    for seed in range(60):
        rng = random.Random(7000 + seed)
        alpha = rng.choice([b"ab", b"abcXYZ", bytes(range(97, 123)), bytes(range(32, 127)), bytes(range(256))])
        lits, flags, ids, text = litgen.random_case(rng, None, rng.choice([1, 3, 20, 200]), alpha, rng.choice([64, 1500, 5000]))
        n, st = check(emul, lits, flags, ids, text, wide)
        total_ends += n
        forms.add(st[1])
    assert total_ends > 2000


def test_workload_tables(emul):
    """C1 (K = 4) is estimated fit for the 4-byte union rows, C2 (10 000 literals behind two shared bytes) for the pair
    table's 8-byte rows; the union rows keep every planted literal of both."""
    lits, flags, ids = W.c1_patterns()
    text = W.c1_text(1 << 18, lits).tobytes()
    n, st = check(emul, lits, flags, ids, text, -1)
    assert n > 50 and st[0] == 4 and st[1] == 0, st
    assert st[3] < 0.05 * len(text)
    lits, flags, ids = W.c2_patterns()
    flat, _ = W.c2_batch(1 << 18, lits)
    n, st = check(emul, lits, flags, ids, flat.tobytes(), -1)
    assert n > 100 and st[0] == 3 and st[1] == 1, st
    assert st[3] < 0.08 * len(flat) and st[4] < 0.01 * len(flat)
