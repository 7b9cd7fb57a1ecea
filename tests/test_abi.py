"""CPU tests: the C-ABI libraries load and export every symbol include/*.h declares; argument/error
conventions follow the reference (unit/hyperscan/arg_checks.cpp).  No compute calls without a GPU."""
import ctypes as C
import os
import re

from vectorscan_b200 import gpu, litdb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hs_gpu_\w+|hsb_\w+)\s*\(", text)))


def test_gpu_library_exports_every_declared_symbol():
    L = C.CDLL(os.path.join(ROOT, "vectorscan_b200", "libhs_gpu.so"))
    names = [n for n in _declared("hs_gpu.h") if not n.endswith("_t")]
    assert len(names) >= 24
    for n in names:
        assert hasattr(L, n), n


def test_litdb_library_exports_every_declared_symbol():
    L = C.CDLL(os.path.join(ROOT, "vectorscan_b200", "libhs_litdb.so"))
    names = [n for n in _declared("hs_litdb.h") if n not in ("hsb_lit_options",)]
    assert len(names) >= 7
    for n in names:
        assert hasattr(L, n), n


def test_no_torch_types_in_the_abi():
    text = open(os.path.join(ROOT, "include", "hs_gpu.h")).read()
    assert "torch" not in text.replace("no torch", "") and "at::" not in text and "cudaStream_t" not in text


def test_argument_errors_without_device():
    L = gpu.lib()
    s = C.c_void_p()
    assert L.hs_gpu_alloc_scratch(None, C.byref(s)) == gpu.HS_INVALID
    junk = C.create_string_buffer(256)
    assert L.hs_gpu_alloc_scratch(junk, C.byref(s)) == gpu.HS_INVALID  # bad magic
    assert L.hs_gpu_free_scratch(None) == gpu.HS_SUCCESS  # NULL ok, like hs_free_scratch
    db = litdb.compile_lit_multi([b"abc", b"defg"])
    # version mismatch -> HS_DB_VERSION_ERROR
    raw = bytearray(db.raw())
    raw[4:8] = (0x05040a00).to_bytes(4, "little")
    buf = (C.c_char * (len(raw) + 64))()
    base = (C.addressof(buf) + 63) & ~63
    C.memmove(base, bytes(raw), len(raw))
    assert L.hs_gpu_alloc_scratch(C.c_void_p(base), C.byref(s)) == gpu.HS_DB_VERSION_ERROR
    # a mode that is none of block / stream / vectored -> HS_DB_MODE_ERROR
    raw = bytearray(db.raw())
    raw[64 + 12:64 + 16] = (8).to_bytes(4, "little")
    C.memmove(base, bytes(raw), len(raw))
    assert L.hs_gpu_alloc_scratch(C.c_void_p(base), C.byref(s)) == gpu.HS_DB_MODE_ERROR
    # SOM database -> refused, never approximated
    raw = bytearray(db.raw())
    raw[64 + 7] = 1
    C.memmove(base, bytes(raw), len(raw))
    assert L.hs_gpu_alloc_scratch(C.c_void_p(base), C.byref(s)) == gpu.HS_ARCH_ERROR
    # unknown Rose opcode in a literal program -> HS_ARCH_ERROR from the census
    raw = bytearray(db.raw())
    bc = raw[64:]
    prog_candidates = [i for i in range(472, len(bc) - 8, 8) if bc[i] in (33, 38)]  # REPORT / FINAL_REPORT
    assert prog_candidates
    raw[64 + prog_candidates[0]] = 16  # CHECK_INFIX
    C.memmove(base, bytes(raw), len(raw))
    assert L.hs_gpu_alloc_scratch(C.c_void_p(base), C.byref(s)) == gpu.HS_ARCH_ERROR
    if L.hs_gpu_device_count() == 0:
        # a valid database but no device: loud failure, no CPU fallback
        assert L.hs_gpu_alloc_scratch(db.ptr, C.byref(s)) == gpu.HS_ARCH_ERROR
        assert L.hs_gpu_shufti(bytes(16), bytes(16), b"abc", 3, 0) == -2


def test_headers_are_plain_c_and_a_c_caller_links(tmp_path):
    """include/*.h must be usable from the reference's own language (C): compile a C caller with gcc -std=c99
    -pedantic that references the entry points of the drop-in path with the reference's callback typedef, link it
    against the two shared libraries, and run it far enough to get HS_INVALID for a NULL database (no device work)."""
    import subprocess
    src = tmp_path / "caller.c"
    src.write_text(r'''
#include "hs_gpu.h"
#include "hs_litdb.h"
#include <stdio.h>
static int on_match(unsigned int id, unsigned long long from, unsigned long long to, unsigned int flags, void *ctx) {
    (void)id; (void)from; (void)to; (void)flags; (void)ctx; return 0;
}
int main(void) {
    hs_gpu_scratch_t *scratch = NULL;
    int rv = hs_gpu_alloc_scratch(NULL, &scratch);               /* hs_alloc_scratch(NULL, ..) -> HS_INVALID */
    int rv2 = hs_gpu_scan(NULL, "x", 1, 0, NULL, on_match, NULL); /* hs_scan(NULL, ..)         -> HS_INVALID */
    const char *ex[1] = {"abc"};
    hs_database_t *db = NULL;
    char *err = NULL;
    int rv3 = hsb_compile_regex_multi_ex(ex, NULL, NULL, 1, 1 /* HS_MODE_BLOCK */, HSB_REGEX_ENGINE_SHENG, &db, &err);
    printf("%d %d %d %d\n", rv, rv2, rv3, db != NULL);
    hsb_free_database(db);
    return 0;
}
''')
    exe = tmp_path / "caller"
    pkg = os.path.join(ROOT, "vectorscan_b200")
    subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), str(src),
                    "-L" + pkg, "-lhs_gpu", "-lhs_litdb", "-Wl,-rpath," + pkg, "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()
    assert out == ["-1", "-1", "0", "1"]
