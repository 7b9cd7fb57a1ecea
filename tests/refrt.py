"""TEST INFRASTRUCTURE: ctypes driver for oracle/_ref/libhs_ref_*.so -- the UNMODIFIED reference
runtime (Vectorscan 5.4.11) compiled from /root/reference by oracle/Makefile.  Only tests, smoke()
and bench.py's cpu_baseline / --impl reference legs may import this."""
import ctypes as C
import os

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MATCH_CB = C.CFUNCTYPE(C.c_int, C.c_uint, C.c_ulonglong, C.c_ulonglong, C.c_uint, C.c_void_p)


def _cpu_flags():
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def pick_isa():
    fl = _cpu_flags()
    need512 = {"avx512f", "avx512bw", "avx512vl", "avx512dq", "avx512cd", "avx512vbmi", "avx512_vbmi2",
               "avx512_vnni", "avx512_bitalg", "avx512_vpopcntdq", "gfni", "vaes", "vpclmulqdq"}
    if need512 <= fl:
        return "avx512"
    return "avx2"


class RefRuntime:
    def __init__(self, isa=None):
        self.isa = isa or os.environ.get("HSB_REF_ISA") or pick_isa()
        path = os.path.join(_ROOT, "oracle", "_ref", f"libhs_ref_{self.isa}.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        L = C.CDLL(path)
        L.hs_deserialize_database.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_void_p)]
        L.hs_alloc_scratch.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.hs_clone_scratch.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.hs_free_scratch.argtypes = [C.c_void_p]
        L.hs_free_database.argtypes = [C.c_void_p]
        L.hs_scan.argtypes = [C.c_void_p, C.c_char_p, C.c_uint, C.c_uint, C.c_void_p, MATCH_CB, C.c_void_p]
        L.hs_scan.restype = C.c_int
        L.hs_scan_vector.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.POINTER(C.c_uint), C.c_uint, C.c_uint, C.c_void_p,
                                     MATCH_CB, C.c_void_p]
        L.hs_scan_vector.restype = C.c_int
        L.hs_open_stream.argtypes = [C.c_void_p, C.c_uint, C.POINTER(C.c_void_p)]
        L.hs_scan_stream.argtypes = [C.c_void_p, C.c_char_p, C.c_uint, C.c_uint, C.c_void_p, MATCH_CB, C.c_void_p]
        L.hs_close_stream.argtypes = [C.c_void_p, C.c_void_p, MATCH_CB, C.c_void_p]
        L.hs_reset_stream.argtypes = [C.c_void_p, C.c_uint, C.c_void_p, MATCH_CB, C.c_void_p]
        L.hs_copy_stream.argtypes = [C.POINTER(C.c_void_p), C.c_void_p]
        L.hs_database_info.argtypes = [C.c_void_p, C.POINTER(C.c_char_p)]
        L.ref_shim_crc32c.argtypes = [C.c_char_p, C.c_size_t]
        L.ref_shim_crc32c.restype = C.c_uint
        L.ref_shim_db_is_valid.argtypes = [C.c_void_p]
        self.L = L

    def deserialize(self, data):
        db = C.c_void_p()
        rv = self.L.hs_deserialize_database(data, len(data), C.byref(db))
        if rv != 0:
            raise RuntimeError(f"hs_deserialize_database -> {rv}")
        return db

    def alloc_scratch(self, db):
        s = C.c_void_p()
        rv = self.L.hs_alloc_scratch(db, C.byref(s))
        if rv != 0:
            raise RuntimeError(f"hs_alloc_scratch -> {rv}")
        return s

    def scan(self, db, scratch, data, stop_after=0):
        """hs_scan with a recording callback; returns (rv, [(id, to), ...]) in callback order."""
        out = []

        def cb(id_, from_, to, flags, ctx):
            out.append((id_, to))
            return 1 if stop_after and len(out) >= stop_after else 0

        rv = self.L.hs_scan(db, data, len(data), 0, scratch, MATCH_CB(cb), None)
        return rv, out

    def scan_vector(self, db, scratch, bufs, stop_after=0):
        """hs_scan_vector (HS_MODE_VECTORED databases): (rv, [(id, to), ...]) in callback order."""
        out = []

        def cb(id_, from_, to, flags, ctx):
            out.append((id_, to))
            return 1 if stop_after and len(out) >= stop_after else 0

        n = len(bufs)
        arr = (C.c_char_p * n)(*bufs)
        lens = (C.c_uint * n)(*[len(b) for b in bufs])
        rv = self.L.hs_scan_vector(db, arr, lens, n, 0, scratch, MATCH_CB(cb), None)
        return rv, out

    def scan_stream_blocks(self, db, scratch, bufs, stop_after=0):
        """hs_open_stream, hs_scan_stream per block, hs_close_stream (HS_MODE_STREAM databases): ([rv per call], matches in
        callback order with stream offsets)."""
        out, rvs = [], []

        def cb(id_, from_, to, flags, ctx):
            out.append((id_, to))
            return 1 if stop_after and len(out) >= stop_after else 0

        fn = MATCH_CB(cb)
        st = C.c_void_p()
        rv = self.L.hs_open_stream(db, 0, C.byref(st))
        if rv != 0:
            return [rv], out
        for b in bufs:
            rvs.append(self.L.hs_scan_stream(st, b, len(b), 0, scratch, fn, None))
        rvs.append(self.L.hs_close_stream(st, scratch, fn, None))
        return rvs, out

    def info(self, db):
        s = C.c_char_p()
        self.L.hs_database_info(db, C.byref(s))
        return s.value.decode()
