/* oracle/ref_shim.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Compiled INTO oracle/_ref/libhs_ref_*.so next to the unmodified reference runtime so that
 * tests can drive the reference's engine-level entry points the same way its own unit tests
 * do (unit/internal/fdr.cpp:94-108,180-183: a zeroed hs_scratch and a recording HWLMCallback).
 * It contains no matching logic of its own.
 */
#include "config.h"
#include "ue2common.h"
#include "scratch.h"
#include "hwlm/hwlm.h"
#include "fdr/fdr.h"
#include "database.h"
#include "crc32.h"
#include <string.h>
#include <stdlib.h>

struct ref_hit { unsigned long long end; unsigned int id; };
struct ref_rec { struct ref_hit *out; size_t cap; size_t n; hwlm_group_t ret; size_t stop_after; };

static __thread struct ref_rec *g_rec;

static hwlmcb_rv_t rec_cb(size_t end, u32 id, struct hs_scratch *scratch) {
    (void)scratch;
    struct ref_rec *r = g_rec;
    if (r->n < r->cap) { r->out[r->n].end = end; r->out[r->n].id = id; }
    r->n++;
    if (r->stop_after && r->n >= r->stop_after) return HWLM_TERMINATE_MATCHING;
    return r->ret;
}

/* hwlmExec over [start,len) with every group on; returns number of (end,id) hits recorded.
 * `hwlm` points at a struct HWLM (src/hwlm/hwlm_internal.h:48). */
size_t ref_shim_hwlm_exec(const void *hwlm, const unsigned char *buf, size_t len, size_t start,
                          struct ref_hit *out, size_t cap, size_t stop_after) {
    struct hs_scratch *s;
    if (posix_memalign((void **)&s, 64, sizeof(*s))) return (size_t)-1;
    memset(s, 0, sizeof(*s));
    struct ref_rec r = { out, cap, 0, HWLM_CONTINUE_MATCHING, stop_after };
    g_rec = &r;
    hwlmExec((const struct HWLM *)hwlm, buf, len, start, rec_cb, s, ~0ULL);
    g_rec = NULL;
    free(s);
    return r.n;
}

/* fdrExec on a bare FDR/Teddy table (src/fdr/fdr.c:801). */
size_t ref_shim_fdr_exec(const void *fdr, const unsigned char *buf, size_t len, size_t start,
                         struct ref_hit *out, size_t cap, size_t stop_after) {
    struct hs_scratch *s;
    if (posix_memalign((void **)&s, 64, sizeof(*s))) return (size_t)-1;
    memset(s, 0, sizeof(*s));
    struct ref_rec r = { out, cap, 0, HWLM_CONTINUE_MATCHING, stop_after };
    g_rec = &r;
    fdrExec((const struct FDR *)fdr, buf, len, start, rec_cb, s, ~0ULL);
    g_rec = NULL;
    free(s);
    return r.n;
}

/* The reference's own CRC (src/crc32.c) and DB validity check (src/database.c dbIsValid). */
unsigned int ref_shim_crc32c(const void *p, size_t n) { return Crc32c_ComputeBuf(0, p, n); }
int ref_shim_db_is_valid(const void *db) { return (int)dbIsValid((const struct hs_database *)db); }
unsigned long long ref_shim_current_platform(void) { return hs_current_platform; }

/* Byte-class primitives: thin offset-returning wrappers over the reference's own routines
 * (src/nfa/shufti.h, truffle.h, vermicelli.hpp).  Forward: offset of the hit or len; reverse: offset or -1. */
#include "nfa/shufti.h"
#include "nfa/truffle.h"
#include "util/simd_utils.h"
const u8 *vermicelliExec(char c, char noCase, const u8 *buf, const u8 *buf_end);
const u8 *nvermicelliExec(char c, char noCase, const u8 *buf, const u8 *buf_end);
const u8 *rvermicelliExec(char c, char nocase, const u8 *buf, const u8 *buf_end);
const u8 *rnvermicelliExec(char c, char nocase, const u8 *buf, const u8 *buf_end);
const u8 *vermicelliDoubleExec(char c1, char c2, char nocase, const u8 *buf, const u8 *buf_end);

long long ref_shim_shufti(const u8 *lo, const u8 *hi, const u8 *buf, size_t len, int reverse) {
    m128 l = loadu128(lo), h = loadu128(hi);
    const u8 *r = reverse ? rshuftiExec(l, h, buf, buf + len) : shuftiExec(l, h, buf, buf + len);
    return (long long)(r - buf);
}
long long ref_shim_truffle(const u8 *m1, const u8 *m2, const u8 *buf, size_t len, int reverse) {
    m128 a = loadu128(m1), b = loadu128(m2);
    const u8 *r = reverse ? rtruffleExec(a, b, buf, buf + len) : truffleExec(a, b, buf, buf + len);
    return (long long)(r - buf);
}
long long ref_shim_verm(u8 c, int nocase, const u8 *buf, size_t len, int reverse, int negate) {
    const u8 *r;
    if (!reverse) r = negate ? nvermicelliExec((char)c, (char)nocase, buf, buf + len) : vermicelliExec((char)c, (char)nocase, buf, buf + len);
    else r = negate ? rnvermicelliExec((char)c, (char)nocase, buf, buf + len) : rvermicelliExec((char)c, (char)nocase, buf, buf + len);
    return (long long)(r - buf);
}
long long ref_shim_dverm(u8 c1, u8 c2, int nocase, const u8 *buf, size_t len) {
    return (long long)(vermicelliDoubleExec((char)c1, (char)c2, (char)nocase, buf, buf + len) - buf);
}
long long ref_shim_dshufti(const u8 *lo1, const u8 *hi1, const u8 *lo2, const u8 *hi2, const u8 *buf, size_t len) {
    return (long long)(shuftiDoubleExec(loadu128(lo1), loadu128(hi1), loadu128(lo2), loadu128(hi2), buf, buf + len) - buf);
}
