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
