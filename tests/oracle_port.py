"""TEST INFRASTRUCTURE: ctypes driver of oracle/liboracle_port.so (the plain-C restatement)."""
import ctypes as C
import os

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(os.path.join(_ROOT, "oracle", "liboracle_port.so"))
        L.op_scan_collect.argtypes = [C.c_void_p, C.c_char_p, C.c_uint, C.POINTER(C.c_uint32), C.POINTER(C.c_uint64),
                                      C.c_size_t, C.c_size_t, C.POINTER(C.c_size_t)]
        L.op_scan_vector_collect.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.POINTER(C.c_uint), C.c_uint,
                                             C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.c_size_t, C.c_size_t,
                                             C.POINTER(C.c_size_t)]
        L.op_brute_force.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.POINTER(C.c_uint),
                                     C.POINTER(C.c_uint), C.c_uint, C.c_char_p, C.c_size_t, C.POINTER(C.c_uint32),
                                     C.POINTER(C.c_uint64), C.c_size_t]
        L.op_brute_force.restype = C.c_size_t
        _lib = L
    return _lib


def scan(db, data, stop_after=0, cap=1 << 16):
    """db: litdb.Database (in-memory image). Returns (rv, [(id, to), ...]) in callback order."""
    while True:
        ids = (C.c_uint32 * cap)()
        tos = (C.c_uint64 * cap)()
        n = C.c_size_t()
        rv = lib().op_scan_collect(db.ptr, data, len(data), ids, tos, cap, stop_after, C.byref(n))
        if n.value <= cap:
            return rv, [(ids[i], tos[i]) for i in range(n.value)]
        cap = n.value + 16


def scan_vector(db, bufs, stop_after=0, cap=1 << 16):
    """hs_scan_vector on a vectored-mode database: (rv, [(id, to), ...]) in callback order, offsets running on."""
    n = len(bufs)
    arr = (C.c_char_p * n)(*bufs)
    lens = (C.c_uint * n)(*[len(b) for b in bufs])
    while True:
        ids = (C.c_uint32 * cap)()
        tos = (C.c_uint64 * cap)()
        m = C.c_size_t()
        rv = lib().op_scan_vector_collect(db.ptr, arr, lens, n, ids, tos, cap, stop_after, C.byref(m))
        if m.value <= cap:
            return rv, [(ids[i], tos[i]) for i in range(m.value)]
        cap = m.value + 16


scan_stream_blocks = scan_vector   # HS_MODE_STREAM databases: the same block-by-block loop (open, scan per block, close)


def brute_force(lits, flags, ids, data, cap=1 << 16):
    n = len(lits)
    arr = (C.c_char_p * n)(*lits)
    lens = (C.c_size_t * n)(*[len(x) for x in lits])
    fl = (C.c_uint * n)(*flags)
    idv = (C.c_uint * n)(*ids)
    while True:
        oi = (C.c_uint32 * cap)()
        ot = (C.c_uint64 * cap)()
        m = lib().op_brute_force(arr, lens, fl, idv, n, data, len(data), oi, ot, cap)
        if m <= cap:
            return [(oi[i], ot[i]) for i in range(m)]
        cap = m + 16
