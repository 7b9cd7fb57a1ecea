"""ctypes binding of libhs_litdb.so -- the literal-set database writer (include/hs_litdb.h).

Mirrors the reference's hs_compile_lit_multi (/root/reference/src/hs_compile.h:640-648): literals,
per-literal flags, ids -> an hs_database_t in the Vectorscan 5.4.11 bytecode format.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))

HS_FLAG_CASELESS = 1
HS_FLAG_DOTALL = 2
HS_FLAG_SINGLEMATCH = 8
HS_MODE_BLOCK = 1
HS_MODE_STREAM = 2
HS_MODE_VECTORED = 4

ENGINE_AUTO, ENGINE_FDR, ENGINE_TEDDY, ENGINE_NOODLE = 0, 1, 2, 3


class LitOptions(C.Structure):
    _fields_ = [("engine", C.c_int), ("fdr_domain", C.c_int), ("fdr_stride", C.c_int),
                ("teddy_masks", C.c_int), ("no_flood", C.c_int), ("teddy_fat", C.c_int)]


class LitLook(C.Structure):
    _fields_ = [("offset", C.c_int), ("and_mask", C.c_ubyte), ("cmp_mask", C.c_ubyte), ("negate", C.c_ubyte),
                ("reserved", C.c_ubyte), ("cls", C.c_void_p)]


class LitExt(C.Structure):
    _fields_ = [("look", C.POINTER(LitLook)), ("look_count", C.c_uint), ("group", C.c_uint), ("flags", C.c_uint)]


class LitBuildExt(C.Structure):
    _fields_ = [("included_jump", C.c_int), ("target_avx512", C.c_int), ("groups_off", C.c_ulonglong),
                ("max_bi_anchored_width", C.c_uint), ("reserved_pure_literal", C.c_uint)]


EXT_CHECK_GROUPS, EXT_SQUASH_GROUP, EXT_QUIET = 1, 2, 4

_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "libhs_litdb.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
        L = C.CDLL(path)
        L.hsb_compile_lit_multi.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_uint), C.POINTER(C.c_uint),
                                            C.POINTER(C.c_size_t), C.c_uint, C.c_uint, C.POINTER(LitOptions),
                                            C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.hsb_compile_lit_multi.restype = C.c_int
        L.hsb_compile_lit_multi_ext.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_uint), C.POINTER(C.c_uint),
                                                C.POINTER(C.c_size_t), C.POINTER(LitExt), C.c_uint, C.c_uint,
                                                C.POINTER(LitOptions), C.POINTER(LitBuildExt),
                                                C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.hsb_compile_lit_multi_ext.restype = C.c_int
        L.hsb_compile_mixed.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_uint), C.POINTER(C.c_uint),
                                        C.POINTER(C.c_size_t), C.POINTER(LitExt), C.c_uint,
                                        C.POINTER(C.c_char_p), C.POINTER(C.c_uint), C.POINTER(C.c_uint), C.c_uint, C.c_uint,
                                        C.POINTER(LitOptions), C.POINTER(LitBuildExt), C.c_uint, C.c_uint,
                                        C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.hsb_compile_mixed.restype = C.c_int
        L.hsb_compile_regex_multi.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_uint), C.POINTER(C.c_uint), C.c_uint,
                                              C.c_uint, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.hsb_compile_regex_multi.restype = C.c_int
        L.hsb_compile_regex_multi_ex.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_uint), C.POINTER(C.c_uint), C.c_uint,
                                                 C.c_uint, C.c_uint, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.hsb_compile_regex_multi_ex.restype = C.c_int
        L.hsb_compile_regex_rose.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_uint), C.POINTER(C.c_uint), C.c_uint,
                                             C.c_uint, C.c_uint, C.c_uint, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.hsb_compile_regex_rose.restype = C.c_int
        L.hsb_free_database.argtypes = [C.c_void_p]
        L.hsb_free_error.argtypes = [C.c_void_p]
        L.hsb_database_size.argtypes = [C.c_void_p, C.POINTER(C.c_size_t)]
        L.hsb_serialize_database.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
        L.hsb_deserialize_database.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_void_p)]
        L.hsb_crc32c.argtypes = [C.c_char_p, C.c_size_t]
        L.hsb_crc32c.restype = C.c_uint
        _lib = L
    return _lib


class CompileError(Exception):
    pass


class Database:
    """Owns an hs_database_t produced by the writer (or deserialized)."""

    def __init__(self, ptr):
        self.ptr = C.c_void_p(ptr)

    def __del__(self):
        try:
            if self.ptr:
                lib().hsb_free_database(self.ptr)
                self.ptr = None
        except Exception:
            pass

    @property
    def size(self):
        n = C.c_size_t()
        lib().hsb_database_size(self.ptr, C.byref(n))
        return n.value

    def raw(self):
        """The in-memory image (header + padding + bytecode)."""
        return C.string_at(self.ptr, self.size)

    def bytecode(self):
        raw = self.raw()
        off = int.from_bytes(raw[36:40], "little")
        return raw[off:]

    def limex_bits(self):
        """State width of the LimEx NFA of a single-outfix database (0 for any other database):
        RoseEngine.nfaInfoOffset -> NfaInfo.nfaOffset -> NFA.type (nfa_internal.h:89-128)."""
        bc = self.bytecode()
        if bc[4] != 2:  # RoseEngine.runtimeImpl == ROSE_RUNTIME_SINGLE_OUTFIX
            return 0
        info = int.from_bytes(bc[236:240], "little")
        nfa = int.from_bytes(bc[info:info + 4], "little")
        t = bc[nfa + 8]
        return [32, 64, 128, 256, 384, 512][t] if t < 6 else 0

    def outfix_engine_type(self):
        """NFA.type of the sole outfix engine (nfa_internal.h:52-80: 0-5 LimEx32..512, 6/7 McClellan 8/16-bit), or
        -1 when the database is not a single-outfix database."""
        bc = self.bytecode()
        if bc[4] != 2:
            return -1
        info = int.from_bytes(bc[236:240], "little")
        nfa = int.from_bytes(bc[info:info + 4], "little")
        return bc[nfa + 8]

    def serialize(self):
        out, n = C.c_void_p(), C.c_size_t()
        if lib().hsb_serialize_database(self.ptr, C.byref(out), C.byref(n)) != 0:
            raise RuntimeError("serialize failed")
        data = C.string_at(out, n.value)
        C.CDLL(None).free(out)
        return data

    @staticmethod
    def deserialize(data):
        db = C.c_void_p()
        if lib().hsb_deserialize_database(data, len(data), C.byref(db)) != 0:
            raise RuntimeError("deserialize failed")
        return Database(db.value)


def compile_lit_multi(literals, flags=None, ids=None, mode=HS_MODE_BLOCK, engine=ENGINE_AUTO, fdr_domain=0,
                      fdr_stride=0, teddy_masks=0, no_flood=0, ext=None, included_jump=0, target_avx512=0,
                      groups_off=0, max_bi_anchored_width=0, full_rose=False, regexes=None, regex_flags=None,
                      regex_ids=None, regex_engine=1, dfa_states=0, teddy_fat=0):
    """literals: list of bytes. Returns Database. Same argument meaning as hs_compile_lit_multi.
    ext: optional per-literal dicts {"look": [(offset, and, cmp, negate) | (offset, class_bitmap32)], "group": g,
    "flags": EXT_*} (include/hs_litdb.h, hsb_lit_ext); the remaining keywords are hsb_lit_build_ext."""
    n = len(literals)
    lits = [bytes(x) for x in literals]
    arr = (C.c_char_p * n)(*lits)
    lens = (C.c_size_t * n)(*[len(x) for x in lits])
    fl = (C.c_uint * n)(*(flags if flags is not None else [0] * n))
    idv = (C.c_uint * n)(*(ids if ids is not None else list(range(n))))
    opt = LitOptions(engine, fdr_domain, fdr_stride, teddy_masks, no_flood, teddy_fat)
    keep = []
    extv = None
    if ext is not None:
        extv = (LitExt * n)()
        for i, e in enumerate(ext):
            e = e or {}
            look = e.get("look", [])
            if look:
                lv = (LitLook * len(look))()
                for k, le in enumerate(look):
                    lv[k].offset = le[0]
                    if len(le) == 2:
                        cb = C.create_string_buffer(bytes(le[1]), 32)
                        keep.append(cb)
                        lv[k].cls = C.cast(cb, C.c_void_p)
                    else:
                        lv[k].and_mask, lv[k].cmp_mask, lv[k].negate = le[1], le[2], 1 if le[3] else 0
                keep.append(lv)
                extv[i].look = lv
                extv[i].look_count = len(look)
            extv[i].group = e.get("group", 0)
            extv[i].flags = e.get("flags", 0)
    bx = LitBuildExt(included_jump, target_avx512, groups_off, max_bi_anchored_width, 0)
    db, err = C.c_void_p(), C.c_void_p()
    if full_rose or regexes:
        # hsb_compile_mixed: a full-Rose database, the regexes (if any) as outfix engines beside the literal matcher
        rx = [e.encode("latin-1") if isinstance(e, str) else bytes(e) for e in (regexes or [])]
        m = len(rx)
        rarr = (C.c_char_p * max(1, m))(*rx)
        rfl = (C.c_uint * max(1, m))(*(regex_flags if regex_flags is not None else [0] * m))
        rid = (C.c_uint * max(1, m))(*(regex_ids if regex_ids is not None else list(range(m))))
        rv = lib().hsb_compile_mixed(arr, fl, idv, lens, extv, n, rarr, rfl, rid, m, mode, C.byref(opt), C.byref(bx),
                                     regex_engine, dfa_states, C.byref(db), C.byref(err))
    else:
        rv = lib().hsb_compile_lit_multi_ext(arr, fl, idv, lens, extv, n, mode, C.byref(opt), C.byref(bx), C.byref(db),
                                             C.byref(err))
    if rv != 0:
        msg = C.string_at(err).decode() if err else "error"
        if err:
            lib().hsb_free_error(err)
        raise CompileError(msg)
    return Database(db.value)


ENGINE_LIMEX, ENGINE_MCCLELLAN, ENGINE_MCCLELLAN_SHERMAN, ENGINE_SHENG, ENGINE_SHENG_WIDE, ENGINE_MCSHENG = 0, 1, 2, 3, 4, 5
ENGINE_MCCLELLAN_WIDE = 6
ENGINE_MCSHENG64 = 7


def compile_regex_multi(expressions, flags=None, ids=None, mode=HS_MODE_BLOCK, engine=ENGINE_LIMEX):
    """expressions: list of bytes/str in the writer's dialect (csrc/nfadb_writer.cpp).  Returns a single-outfix
    Database holding one LimEx NFA (default) or one McClellan DFA.  Same argument meaning as hs_compile_multi."""
    n = len(expressions)
    ex = [e.encode("latin-1") if isinstance(e, str) else bytes(e) for e in expressions]
    if any(b"\0" in e for e in ex):
        raise CompileError("NUL byte in expression: write it as \\x00")
    arr = (C.c_char_p * n)(*ex)
    fl = (C.c_uint * n)(*(flags if flags is not None else [0] * n))
    idv = (C.c_uint * n)(*(ids if ids is not None else list(range(n))))
    db, err = C.c_void_p(), C.c_void_p()
    rv = lib().hsb_compile_regex_multi_ex(arr, fl, idv, n, mode, engine, C.byref(db), C.byref(err))
    if rv != 0:
        msg = C.string_at(err).decode() if err else "error"
        if err:
            lib().hsb_free_error(err)
        raise CompileError(msg)
    return Database(db.value)


def compile_regex_rose(expressions, flags=None, ids=None, mode=HS_MODE_BLOCK, engine=ENGINE_MCCLELLAN, dfa_states=0):
    """Large regex sets: a full-Rose Database whose patterns are spread over many outfix engines (McClellan DFAs of at
    most dfa_states states, LimEx NFAs for what does not determinise).  include/hs_litdb.h hsb_compile_regex_rose."""
    n = len(expressions)
    ex = [e.encode("latin-1") if isinstance(e, str) else bytes(e) for e in expressions]
    if any(b"\0" in e for e in ex):
        raise CompileError("NUL byte in expression: write it as \\x00")
    arr = (C.c_char_p * n)(*ex)
    fl = (C.c_uint * n)(*(flags if flags is not None else [0] * n))
    idv = (C.c_uint * n)(*(ids if ids is not None else list(range(n))))
    db, err = C.c_void_p(), C.c_void_p()
    rv = lib().hsb_compile_regex_rose(arr, fl, idv, n, mode, engine, dfa_states, C.byref(db), C.byref(err))
    if rv != 0:
        msg = C.string_at(err).decode() if err else "error"
        if err:
            lib().hsb_free_error(err)
        raise CompileError(msg)
    return Database(db.value)
