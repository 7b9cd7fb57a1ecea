"""Byte-class search microbenchmark (north star item 2; shape of /root/reference/benchmarks/benchmarks.cpp:112-250):
shufti / truffle / vermicelli / double vermicelli / double shufti over 16 KiB ... 256 MiB, no match (full scan) or a
match 5 bytes before the end; device kernel time by CUDA events with the data resident, the reference's own routine on
one host core beside it (oracle/_ref through ref_shim.c).  One JSON line per (routine, size, shape)."""
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

from vectorscan_b200 import gpu  # noqa: E402


def main():
    import refrt
    R = refrt.RefRuntime()
    L = gpu.lib()
    L.hs_gpu_accel_search_staged.restype = C.c_longlong
    L.hs_gpu_accel_search_staged.argtypes = [C.c_int, C.c_char_p, C.c_ubyte, C.c_ubyte, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float)]
    L.hs_gpu_accel_stage.argtypes = [C.c_void_p, C.c_size_t]
    for f, at in (("ref_shim_shufti", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]),
                  ("ref_shim_truffle", [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]),
                  ("ref_shim_verm", [C.c_ubyte, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int]),
                  ("ref_shim_dverm", [C.c_ubyte, C.c_ubyte, C.c_int, C.c_void_p, C.c_size_t]),
                  ("ref_shim_dshufti", [C.c_void_p] * 5 + [C.c_size_t])):
        getattr(R.L, f).restype = C.c_longlong
        getattr(R.L, f).argtypes = at
    # class {'X','Y'}: one shufti bucket; truffle masks; pair "XY" for the double forms
    lo, hi = bytearray(16), bytearray(16)
    for c in b"XY":
        lo[c & 15] |= 1
        hi[c >> 4] |= 1
    t1 = bytearray(16)
    for c in b"XY":
        t1[c & 15] |= 1 << ((c >> 4) & 7)
    d = [bytearray([0xff] * 16) for _ in range(4)]
    d[0][ord("X") & 15] &= 0xfe; d[1][ord("X") >> 4] &= 0xfe; d[2][ord("Y") & 15] &= 0xfe; d[3][ord("Y") >> 4] &= 0xfe
    routines = [
        ("shufti", 0, bytes(lo) + bytes(hi) + bytes(32), 0, 0, lambda p, n: R.L.ref_shim_shufti(bytes(lo), bytes(hi), p, n, 0)),
        ("truffle", 1, bytes(t1) + bytes(16) + bytes(32), 0, 0, lambda p, n: R.L.ref_shim_truffle(bytes(t1), bytes(16), p, n, 0)),
        ("vermicelli", 2, bytes(64), ord("X"), 0, lambda p, n: R.L.ref_shim_verm(ord("X"), 0, p, n, 0, 0)),
        ("vermicelli_double", 4, bytes(64), ord("X"), ord("Y"), lambda p, n: R.L.ref_shim_dverm(ord("X"), ord("Y"), 0, p, n)),
        ("shufti_double", 5, b"".join(bytes(x) for x in d), 0, 0,
         lambda p, n: R.L.ref_shim_dshufti(bytes(d[0]), bytes(d[1]), bytes(d[2]), bytes(d[3]), p, n)),
    ]
    sizes = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [16 << 10, 256 << 10, 4 << 20, 64 << 20, 256 << 20]
    rng = np.random.default_rng(5)
    for n in sizes:
        base = rng.integers(97, 123, n, dtype=np.uint8)       # a-z: no member of any class
        for shape in ("no match", "match 5 bytes before the end"):
            data = base.copy()
            if shape != "no match":
                data[n - 6] = ord("X"); data[n - 5] = ord("Y")
            assert L.hs_gpu_accel_stage(data.ctypes.data, n) == 0
            for name, kind, masks, c1, c2, cpu in routines:
                ms = C.c_float()
                reps = 20 if n <= (4 << 20) else 5
                L.hs_gpu_accel_search_staged(kind, masks, c1, c2, 0, 0, 3, C.byref(ms))          # warm-up
                got = L.hs_gpu_accel_search_staged(kind, masks, c1, c2, 0, 0, reps, C.byref(ms))
                creps = 3 if n >= (64 << 20) else 20
                cpu(data.ctypes.data, n)
                t0 = time.perf_counter()
                for _ in range(creps):
                    exp = cpu(data.ctypes.data, n)
                cpu_s = (time.perf_counter() - t0) / creps
                want = n if shape == "no match" else n - 6
                assert got == want, (name, n, shape, got)
                # (the reference's 512-bit vermicelli answers "not found" for some matches close to the end of the buffer --
                # tests/test_gpu_accel.py notes the defect; its answer is recorded, not asserted)
                print(json.dumps({"routine": name, "bytes": n, "shape": shape, "gpu_kernel_ms": round(ms.value, 4),
                                  "gpu_GBps": round(n / (ms.value * 1e-3) / 1e9, 1), "cpu_ms": round(cpu_s * 1e3, 4),
                                  "cpu_GBps": round(n / cpu_s / 1e9, 2), "cpu": f"reference routine, 1 core, {R.isa} build",
                                  "result": int(got), "cpu_result": int(exp)}), flush=True)


if __name__ == "__main__":
    main()
