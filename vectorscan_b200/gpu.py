"""ctypes binding of libhs_gpu.so (include/hs_gpu.h) -- the B200 block-mode scanning runtime.

The Python names mirror the reference's runtime API (/root/reference/src/hs_runtime.h): alloc_scratch,
scan, scan_vector, free_scratch; a match callback receives (id, from, to, flags, context).  There is
no CPU fallback: loading fails loudly when the CUDA library has not been built, and every scanning
call returns HS_ARCH_ERROR without a CUDA device.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))

HS_SUCCESS, HS_INVALID, HS_NOMEM, HS_SCAN_TERMINATED = 0, -1, -2, -3
HS_DB_VERSION_ERROR, HS_DB_PLATFORM_ERROR, HS_DB_MODE_ERROR = -5, -6, -7
HS_SCRATCH_IN_USE, HS_ARCH_ERROR, HS_UNKNOWN_ERROR = -10, -11, -13

MATCH_CB = C.CFUNCTYPE(C.c_int, C.c_uint, C.c_ulonglong, C.c_ulonglong, C.c_uint, C.c_void_p)


class Match(C.Structure):
    _fields_ = [("buffer", C.c_uint32), ("id", C.c_uint32), ("to", C.c_uint64)]


class Stats(C.Structure):
    _fields_ = [("bytes", C.c_uint64), ("buffers", C.c_uint64), ("matches", C.c_uint64),
                ("raw_matches", C.c_uint64), ("candidates", C.c_uint64), ("confirms", C.c_uint64),
                ("h2d_ms", C.c_float), ("scan_ms", C.c_float), ("sort_ms", C.c_float), ("d2h_ms", C.c_float),
                ("scan_launches", C.c_uint32), ("arena_retries", C.c_uint32), ("kernel_ms", C.c_float),
                ("kernel_launches", C.c_uint32), ("train_ms", C.c_float), ("front_end", C.c_uint32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class HsGpuError(RuntimeError):
    def __init__(self, code, what):
        super().__init__(f"{what} -> {code}")
        self.code = code


KERNEL_NAMES = {1: "k_scan", 2: "k_turbo", 3: "k_turbo2", 4: "k_limex", 5: "k_mcclellan", 6: "k_mcclellan/k_limex (several engines)"}

_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "libhs_gpu.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} missing: the CUDA extension is not built "
                               "(python -c 'import __graft_entry__ as g; g.build()'); there is no CPU fallback")
        L = C.CDLL(path)
        vp, u32p = C.c_void_p, C.POINTER(C.c_uint)
        L.hs_gpu_alloc_scratch.argtypes = [vp, C.POINTER(vp)]
        L.hs_gpu_clone_scratch.argtypes = [vp, C.POINTER(vp)]
        L.hs_gpu_scratch_size.argtypes = [vp, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]
        L.hs_gpu_free_scratch.argtypes = [vp]
        L.hs_gpu_scan.argtypes = [vp, C.c_char_p, C.c_uint, C.c_uint, vp, MATCH_CB, vp]
        L.hs_gpu_scan_vector.argtypes = [vp, C.POINTER(C.c_char_p), u32p, C.c_uint, C.c_uint, vp, MATCH_CB, vp]
        L.hs_gpu_scan_batch.argtypes = [vp, C.POINTER(C.c_char_p), u32p, C.c_uint, C.c_uint, vp, MATCH_CB,
                                        C.POINTER(vp)]
        L.hs_gpu_scan_vector_ex.argtypes = [vp, C.POINTER(C.c_char_p), u32p, C.c_uint, C.c_uint, vp,
                                            C.POINTER(C.POINTER(Match)), C.POINTER(C.c_uint64)]
        L.hs_gpu_corpus_create.argtypes = [vp, C.POINTER(C.c_uint64), C.c_uint, C.POINTER(vp)]
        L.hs_gpu_corpus_free.argtypes = [vp]
        L.hs_gpu_scan_corpus.argtypes = [vp, vp, vp, C.POINTER(C.c_uint64)]
        L.hs_gpu_scan_corpora.argtypes = [vp, C.POINTER(vp), C.c_uint, vp, C.POINTER(C.c_uint64)]
        L.hs_gpu_fetch_matches.argtypes = [vp, C.POINTER(C.POINTER(Match)), C.POINTER(C.c_uint64)]
        L.hs_gpu_last_stats.argtypes = [vp, C.POINTER(Stats)]
        L.hs_gpu_histogram.argtypes = [vp, vp, C.c_uint32]
        L.hs_gpu_host_alloc.argtypes = [C.c_size_t]
        L.hs_gpu_host_alloc.restype = vp
        L.hs_gpu_host_free.argtypes = [vp]
        L.hs_gpu_set_option.argtypes = [vp, C.c_char_p, C.c_long]
        L.hs_gpu_opcode_counts.argtypes = [vp, C.POINTER(C.c_ulonglong), C.c_uint]
        L.hs_gpu_database_opcodes.argtypes = [vp, C.POINTER(C.c_ulonglong), C.c_uint]
        L.hs_gpu_set_device.argtypes = [C.c_int]
        L.hs_gpu_open_stream.argtypes = [vp, C.c_uint, C.POINTER(vp)]
        L.hs_gpu_scan_stream.argtypes = [vp, C.c_char_p, C.c_uint, C.c_uint, vp, MATCH_CB, vp]
        L.hs_gpu_close_stream.argtypes = [vp, vp, MATCH_CB, vp]
        L.hs_gpu_reset_stream.argtypes = [vp, C.c_uint, vp, MATCH_CB, vp]
        L.hs_gpu_copy_stream.argtypes = [C.POINTER(vp), vp]
        L.hs_gpu_scan_streams.argtypes = [C.POINTER(vp), C.POINTER(C.c_char_p), u32p, C.c_uint, C.c_uint, vp, MATCH_CB,
                                          C.POINTER(vp)]
        L.hs_gpu_version.restype = C.c_char_p
        _lib = L
    return _lib


def _dbptr(db):
    return db.ptr if hasattr(db, "ptr") else db


def _matches_to_list(mp, n):
    if not n:
        return []
    import numpy as np
    arr = np.ctypeslib.as_array(C.cast(mp, C.POINTER(C.c_uint8)), shape=(n * 16,)).view(
        np.dtype([("buffer", "<u4"), ("id", "<u4"), ("to", "<u8")]))
    return arr.copy()


class Scratch:
    """hs_scratch_t stand-in: owns the device copy of each database it was allocated for, staging,
    the match arena and a CUDA stream."""

    def __init__(self, db):
        self.ptr = C.c_void_p()
        self.grow(db)

    def grow(self, db):
        rv = lib().hs_gpu_alloc_scratch(_dbptr(db), C.byref(self.ptr))
        if rv != HS_SUCCESS:
            raise HsGpuError(rv, "hs_gpu_alloc_scratch")

    def close(self):
        if self.ptr:
            lib().hs_gpu_free_scratch(self.ptr)
            self.ptr = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_option(self, name, value):
        rv = lib().hs_gpu_set_option(self.ptr, name.encode(), int(value))
        if rv != HS_SUCCESS:
            raise HsGpuError(rv, "hs_gpu_set_option")

    def opcode_counts(self):
        """Executions per Rose opcode since set_option("count_ops", 1) (test instrumentation)."""
        out = (C.c_ulonglong * 128)()
        rv = lib().hs_gpu_opcode_counts(self.ptr, out, 128)
        if rv != HS_SUCCESS:
            raise HsGpuError(rv, "hs_gpu_opcode_counts")
        return list(out)

    def stats(self):
        st = Stats()
        lib().hs_gpu_last_stats(self.ptr, C.byref(st))
        return st.as_dict()


def database_opcodes(db):
    """Static histogram of the Rose opcodes in the database's literal/report programs (host only)."""
    out = (C.c_ulonglong * 128)()
    rv = lib().hs_gpu_database_opcodes(db.ptr, out, 128)
    if rv != HS_SUCCESS:
        raise HsGpuError(rv, "hs_gpu_database_opcodes")
    return list(out)


def alloc_scratch(db):
    return Scratch(db)


def scan(db, data, scratch, on_event=None, context=None):
    """hs_scan: returns the hs_error_t code; on_event(id, from, to, flags, ctx) -> int."""
    cb = MATCH_CB(on_event) if on_event else C.cast(None, MATCH_CB)
    return lib().hs_gpu_scan(_dbptr(db), data, len(data), 0, scratch.ptr, cb, context)


def _arrays(buffers):
    n = len(buffers)
    arr = (C.c_char_p * n)(*buffers)
    lens = (C.c_uint * n)(*[len(b) for b in buffers])
    return arr, lens, n


def scan_vector(db, buffers, scratch, on_event=None, context=None):
    arr, lens, n = _arrays(buffers)
    cb = MATCH_CB(on_event) if on_event else C.cast(None, MATCH_CB)
    return lib().hs_gpu_scan_vector(_dbptr(db), arr, lens, n, 0, scratch.ptr, cb, context)


def scan_vector_ex(db, buffers, scratch):
    """Returns a numpy structured array (buffer, id, to) sorted by (buffer, to, id)."""
    arr, lens, n = _arrays(buffers)
    mp, cnt = C.POINTER(Match)(), C.c_uint64()
    rv = lib().hs_gpu_scan_vector_ex(_dbptr(db), arr, lens, n, 0, scratch.ptr, C.byref(mp), C.byref(cnt))
    if rv != HS_SUCCESS:
        raise HsGpuError(rv, "hs_gpu_scan_vector_ex")
    return _matches_to_list(mp, cnt.value)


class Corpus:
    """Device-resident batch of buffers (flat bytes + offsets)."""

    def __init__(self, flat, offsets):
        import numpy as np
        self.offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        self.count = len(self.offsets) - 1
        self.total = int(self.offsets[-1])
        flat = np.ascontiguousarray(flat, dtype=np.uint8) if not isinstance(flat, (bytes, bytearray)) else flat
        fp = flat.ctypes.data_as(C.c_void_p) if hasattr(flat, "ctypes") else C.cast(C.c_char_p(bytes(flat)), C.c_void_p)
        self.ptr = C.c_void_p()
        rv = lib().hs_gpu_corpus_create(fp, self.offsets.ctypes.data_as(C.POINTER(C.c_uint64)), self.count,
                                        C.byref(self.ptr))
        if rv != HS_SUCCESS:
            raise HsGpuError(rv, "hs_gpu_corpus_create")

    def close(self):
        if self.ptr:
            lib().hs_gpu_corpus_free(self.ptr)
            self.ptr = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def scan_corpus(db, corpus, scratch):
    n = C.c_uint64()
    rv = lib().hs_gpu_scan_corpus(_dbptr(db), corpus.ptr, scratch.ptr, C.byref(n))
    if rv != HS_SUCCESS:
        raise HsGpuError(rv, "hs_gpu_scan_corpus")
    return n.value


def scan_corpora(db, corpora, scratch):
    """hs_gpu_scan_corpora: the resident corpora scanned back to back as one logical batch (one host round trip);
    buffer indices of the matches continue from corpus to corpus.  Returns the match count."""
    arr = (C.c_void_p * len(corpora))(*[c.ptr for c in corpora])
    n = C.c_uint64()
    rv = lib().hs_gpu_scan_corpora(_dbptr(db), arr, len(corpora), scratch.ptr, C.byref(n))
    if rv != HS_SUCCESS:
        raise HsGpuError(rv, "hs_gpu_scan_corpora")
    return n.value


def fetch_matches(scratch):
    mp, cnt = C.POINTER(Match)(), C.c_uint64()
    rv = lib().hs_gpu_fetch_matches(scratch.ptr, C.byref(mp), C.byref(cnt))
    if rv != HS_SUCCESS:
        raise HsGpuError(rv, "hs_gpu_fetch_matches")
    return _matches_to_list(mp, cnt.value)


# ---- stream mode (HS_MODE_STREAM pure-literal databases) ----------------------------------------------------------------
class Stream:
    """hs_stream_t stand-in: hs_gpu_open_stream ... hs_gpu_close_stream."""

    def __init__(self, db=None, ptr=None):
        self.ptr = C.c_void_p() if ptr is None else ptr
        if db is not None:
            rv = lib().hs_gpu_open_stream(_dbptr(db), 0, C.byref(self.ptr))
            if rv != HS_SUCCESS:
                raise HsGpuError(rv, "hs_gpu_open_stream")

    def scan(self, data, scratch, on_event=None, context=None):
        cb = MATCH_CB(on_event) if on_event else C.cast(None, MATCH_CB)
        return lib().hs_gpu_scan_stream(self.ptr, data, len(data), 0, scratch.ptr, cb, context)

    def reset(self):
        return lib().hs_gpu_reset_stream(self.ptr, 0, None, C.cast(None, MATCH_CB), None)

    def copy(self):
        out = C.c_void_p()
        rv = lib().hs_gpu_copy_stream(C.byref(out), self.ptr)
        if rv != HS_SUCCESS:
            raise HsGpuError(rv, "hs_gpu_copy_stream")
        return Stream(ptr=out)

    def close(self, scratch=None, on_event=None, context=None):
        """hs_close_stream: with a scratch and a handler the end-of-data matches are reported; returns the hs_error_t."""
        rv = HS_SUCCESS
        if self.ptr:
            cb = MATCH_CB(on_event) if on_event else C.cast(None, MATCH_CB)
            rv = lib().hs_gpu_close_stream(self.ptr, scratch.ptr if scratch is not None else None, cb, context)
            self.ptr = C.c_void_p()
        return rv

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def scan_streams(streams, blocks, scratch, on_event=None, contexts=None):
    """hs_gpu_scan_streams: one block for each of several different streams in one device pass.  on_event gets
    contexts[i] (an integer or None) as its last argument for the callbacks of streams[i]."""
    n = len(streams)
    ids = (C.c_void_p * n)(*[s.ptr.value for s in streams])
    arr, lens, _ = _arrays(blocks)
    ctxs = (C.c_void_p * n)(*(contexts if contexts is not None else [None] * n))
    cb = MATCH_CB(on_event) if on_event else C.cast(None, MATCH_CB)
    return lib().hs_gpu_scan_streams(ids, arr, lens, n, 0, scratch.ptr, cb, ctxs)
