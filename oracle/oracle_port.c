/* oracle/oracle_port.c -- TEST INFRASTRUCTURE ONLY.  Never imported, linked or executed by the product
 * path (vectorscan_b200/): only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use it.
 *
 * Plain-C, sequential restatement of the reference's block-mode scan path -- pure-literal databases and
 * single-outfix databases holding a LimEx NFA (see limex_scan below) -- over the reference's own bytecode.  PARITY PINNED: tests/test_oracle.py checks it against (a) the known-answer vectors of the
 * reference's unit tests (tests/golden/ref_unit_vectors.json, transcribed from unit/internal/fdr.cpp) and
 * (b) outputs of the unmodified reference runtime compiled from /root/reference (oracle/_ref) on seeded
 * random databases/inputs (tests/golden/ref_runtime_cases.json, tests/golden/ref_limex_cases.json and live
 * comparison where _ref exists).
 *
 * Each function cites the reference code it follows (paths under /root/reference/src):
 *   op_scan            hs_scan                      runtime.c:316-475  (gates, engine switch)
 *   hwlm_exec          hwlmExec / fdrExec           hwlm/hwlm.c:178-205, fdr/fdr.c:801-825
 *   fdr_candidates     get_conf_stride_{1,2,4}      fdr/fdr.c:144-296 (+ getInitState :128-142, zones :624-659)
 *   teddy_candidates   fdr_exec_teddy_msksN         fdr/teddy.c:920-1066
 *   nood_match         noodExec / final             hwlm/noodle_engine.cpp:75-133
 *   confirm            do_confirm_fdr / confWithBit fdr/fdr.c:298-333, fdr/fdr_confirm_runtime.h:43-102
 *   run_program        roseCallback_i + roseRunProgram_l   rose/match.c:479-523, rose/program_runtime.c:3083-3504
 *   report             dedupeCatchup / roseDeliverReport   report.h:54-117,300-340
 * The walk is in increasing end offset, buckets low to high, chain order -- the order in which the reference
 * invokes its callbacks when no flood/acceleration shortcut is taken (those are result-neutral).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef uint8_t u8;
typedef uint16_t u16;
typedef uint32_t u32;
typedef uint64_t u64;

/* ---- byte layout (same facts as vectorscan_b200/csrc/hsb_layout.h; pinned by tests/test_layout.py) ---- */
enum {
    DB_MAGIC_OFF = 0, DB_VERSION_OFF = 4, DB_LENGTH_OFF = 8, DB_BYTECODE_OFF = 36,
    R_RUNTIME_IMPL = 4, R_HAS_SOM = 7, R_MODE = 12, R_EKEY_COUNT = 20, R_CKEY_COUNT = 32, R_DKEY_COUNT = 44,
    R_FMATCHER = 96, R_MIN_WIDTH = 200, R_MIN_WIDTH_EXCL = 204, R_MAX_BI = 208, R_INIT_GROUPS = 240,
    R_FLOAT_MASK = 248, R_SIZE = 256, R_BOUNDARY = 376,
    HWLM_BODY = 192, F_ENGINE_ID = 0, F_CONF_OFF = 16, F_STRIDE = 24, F_DOMAIN = 25, F_DOMAIN_MASK = 26, F_START = 32,
    F_TABLE = 64, C_ANDMSK = 0, C_MULT = 8, C_NBITS = 16, C_GROUPS = 24, C_SIZE = 32,
    L_V = 0, L_MSK = 8, L_GROUPS = 16, L_ID = 24, L_SIZE = 28, L_FLAGS = 29, L_NEXT = 30, L_STRUCT = 32,
    N_ID = 0, N_MSK = 8, N_CMP = 16, N_MSK_LEN = 24
};
enum { /* Rose opcodes, rose/rose_program.h:44-216 */
    OP_END = 0, OP_CHECK_GROUPS = 3, OP_CHECK_MASK = 9, OP_CHECK_MASK_32 = 10, OP_CHECK_MASK_64 = 69, OP_CHECK_BYTE = 11, OP_CATCH_UP = 20, OP_DEDUPE = 28,
    OP_REPORT = 33, OP_REPORT_EXHAUST = 34, OP_DEDUPE_AND_REPORT = 37, OP_FINAL_REPORT = 38, OP_CHECK_EXHAUSTED = 39,
    OP_SQUASH_GROUPS = 43, OP_CHECK_LONG_LIT = 51, OP_CHECK_LONG_LIT_NOCASE = 52, OP_CHECK_MED_LIT = 53,
    OP_CHECK_MED_LIT_NOCASE = 54, OP_CLEAR_WORK_DONE = 55, OP_INCLUDED_JUMP = 61, OP_SET_EXHAUST = 65
};
#define HS_SUCCESS 0
#define HS_INVALID (-1)
#define HS_SCAN_TERMINATED (-3)
#define HS_DB_VERSION_ERROR (-5)
#define HS_DB_MODE_ERROR (-7)
#define HS_ARCH_ERROR (-11)
#define HS_UNKNOWN_ERROR (-13)

static u32 rd32(const u8 *p) { u32 v; memcpy(&v, p, 4); return v; }
static u64 rd64(const u8 *p) { u64 v; memcpy(&v, p, 8); return v; }
static u16 rd16(const u8 *p) { u16 v; memcpy(&v, p, 2); return v; }

typedef int (*op_match_cb)(unsigned id, unsigned long long from, unsigned long long to, unsigned flags, void *ctx);

struct scan_state {
    const u8 *rose; const u8 *buf; u32 len;
    u32 start; u64 base;               /* vectored mode: buf = history + block; matches end at index >= start, stream offset = base + index */
    u64 groups;
    u8 *exhausted; u32 ekeys;          /* exhaustion vector (report.h:121-153) */
    u8 *dedupe; u32 dkeys; u64 dedupe_offset; /* dedupe log for the current report offset */
    op_match_cb cb; void *ctx;
    int terminated, error;
    /* multi-outfix catch-up (rose/catchup.c): engine reports are gathered, then replayed in offset order */
    int defer; u32 cur_queue; int cur_eod;   /* cur_eod: the engine is in its end-of-data check (ENGINES_EOD) */
    struct deferred { u64 to; u32 queue, prog, seq, eod; } *ev; size_t n_ev, cap_ev, ev_next;
};

/* roseDeliverReport + the dedupe test of dedupeCatchup for external reports without SOM */
static int report(struct scan_state *s, u32 id, u64 to, u32 ekey, u32 dkey) {
    to += s->base;
    if (dkey != 0xffffffffu) {
        if (s->dedupe_offset != to) { memset(s->dedupe, 0, (s->dkeys + 7) / 8 + 1); s->dedupe_offset = to; }
        if (s->dedupe[dkey >> 3] & (1u << (dkey & 7))) return 0; /* DEDUPE_SKIP */
        s->dedupe[dkey >> 3] |= (u8)(1u << (dkey & 7));
    }
    if (s->cb && s->cb(id, 0, to, 0, s->ctx)) { s->terminated = 1; return -1; }
    if (ekey != 0xffffffffu) s->exhausted[ekey >> 3] |= (u8)(1u << (ekey & 7));
    return 1;
}

static u8 upc(u8 c) { return (c >= 'a' && c <= 'z') ? (u8)(c - 32) : c; }

/* roseCheckMediumLiteral, rose/program_runtime.c:1954-2013 (block mode: no history) */
static int check_lit(const struct scan_state *s, u32 lit_off, u32 lit_len, int nocase, u64 end) {
    if (end < lit_len) return 0;
    const u8 *t = s->buf + (end - lit_len), *l = s->rose + lit_off;
    for (u32 i = 0; i < lit_len; i++)
        if ((nocase ? upc(t[i]) : t[i]) != l[i]) return 0;
    return 1;
}

/* roseCheckMask, rose/program_runtime.c:644-727 + validateMask, rose/validate_mask.h:80-104 */
static int check_mask(const struct scan_state *s, u64 and_m, u64 cmp_m, u64 neg_m, int32_t off, u64 end) {
    if (off < 0 && (u64)(-(int64_t)off) > end) return 0;
    int64_t base = (int64_t)end + off;
    for (int k = 0; k < 8; k++) {
        int64_t o = base + k;
        if (o < 0 || o >= (int64_t)s->len) continue;
        u8 r = (u8)((s->buf[o] & (u8)(and_m >> (8 * k))) ^ (u8)(cmp_m >> (8 * k)));
        if ((u8)(neg_m >> (8 * k)) ? r == 0 : r != 0) return 0;
    }
    return 1;
}

/* roseCheckByte, rose/program_runtime.c:600-641 */
static int check_byte(const struct scan_state *s, u8 and_m, u8 cmp_m, u8 neg, int32_t off, u64 end) {
    if (off < 0 && (u64)(-(int64_t)off) > end) return 0;
    int64_t o = (int64_t)end + off;
    if (o < 0 || o >= (int64_t)s->len) return 1;
    int differ = (s->buf[o] & and_m) != cmp_m;
    return !(differ ^ (neg != 0));
}

/* roseCheckMask32 / roseCheckMask64, rose/program_runtime.c:729-935 + validateMask32, rose/validate_mask.h:106-127:
 * n window bytes from end + off; a byte outside the buffer is don't-care, a byte inside passes when
 * ((c & and) != cmp) equals its neg_mask bit. */
static int check_mask_n(const struct scan_state *s, const u8 *and_m, const u8 *cmp_m, u64 neg, u32 n, int32_t off, u64 end) {
    if (off < 0 && (u64)(-(int64_t)off) > end) return 0;
    int64_t base = (int64_t)end + off;
    for (u32 k = 0; k < n; k++) {
        int64_t o = base + k;
        if (o < 0 || o >= (int64_t)s->len) continue;
        int differ = (u8)(s->buf[o] & and_m[k]) != cmp_m[k];
        if (differ != (int)((neg >> k) & 1)) return 0;
    }
    return 1;
}

/* roseCheckShufti16x8 ... 64x16 (rose/program_runtime.c:992-1250, validateShuftiMask* rose/validate_shufti.h:46-372):
 * window byte k is in its class when (lo[c & 15] & hi[c >> 4] & bucket_select[k]) != 0 for one of the two bucket
 * halves; every byte inside the buffer must have (not in class) == its neg_mask bit.  Layouts: rose/rose_program.h. */
static int check_shufti(const struct scan_state *s, const u8 *pc, u64 end) {
    u32 n, wide = 0, lane = 0;
    const u8 *hi1, *lo1, *hi2 = NULL, *lo2 = NULL, *sel_lo, *sel_hi = NULL;
    u64 neg;
    int32_t off;
    switch (pc[0]) {
    case 12: n = 16; lo1 = pc + 1; hi1 = pc + 17; sel_lo = pc + 33; neg = rd32(pc + 52); off = (int32_t)rd32(pc + 56); break;
    case 13: n = 32; hi1 = pc + 1; lo1 = pc + 17; sel_lo = pc + 33; neg = rd32(pc + 68); off = (int32_t)rd32(pc + 72); break;
    case 14: n = 16; wide = 1; hi1 = pc + 1; hi2 = pc + 17; lo1 = pc + 33; lo2 = pc + 49; sel_lo = pc + 65; sel_hi = pc + 81;
             neg = rd32(pc + 100); off = (int32_t)rd32(pc + 104); break;
    case 15: n = 32; wide = 1; hi1 = pc + 1; hi2 = pc + 17; lo1 = pc + 33; lo2 = pc + 49; sel_hi = pc + 65; sel_lo = pc + 97;
             neg = rd32(pc + 132); off = (int32_t)rd32(pc + 136); break;
    case 67: n = 64; lane = 16; hi1 = pc + 1; lo1 = pc + 65; sel_lo = pc + 129; neg = rd64(pc + 200); off = (int32_t)rd32(pc + 208); break;
    default: n = 64; wide = 1; lane = 16; hi1 = pc + 1; hi2 = pc + 65; lo1 = pc + 129; lo2 = pc + 193; sel_hi = pc + 257; sel_lo = pc + 321;
             neg = rd64(pc + 392); off = (int32_t)rd32(pc + 400); break;
    }
    if (off < 0 && (u64)(-(int64_t)off) > end) return 0;
    int64_t base = (int64_t)end + off;
    for (u32 k = 0; k < n; k++) {
        int64_t o = base + k;
        if (o < 0 || o >= (int64_t)s->len) continue;
        u8 c = s->buf[o];
        u32 ln = lane ? (k & ~15u) : 0;
        u8 hit = (u8)(lo1[ln + (c & 15)] & hi1[ln + (c >> 4)] & sel_lo[k]);
        if (wide) hit |= (u8)(lo2[ln + (c & 15)] & hi2[ln + (c >> 4)] & sel_hi[k]);
        if ((hit == 0) != (int)((neg >> k) & 1)) return 0;
    }
    return 1;
}
static u32 shufti_fail_at(u8 code) { return code == 12 ? 60 : code == 13 ? 76 : code == 14 ? 108 : code == 15 ? 140 : code == 67 ? 212 : 404; }

static int catch_up(struct scan_state *s, u64 end);

/* roseRunProgram_l.  `end` = offset one past the literal's last byte (lit_offset_adjust = 1).
 * Returns 0 to continue, -1 to stop the scan. */
static int run_program(struct scan_state *s, u32 prog, u64 end) {
    if (s->defer) {
        if (s->n_ev == s->cap_ev) {
            s->cap_ev = s->cap_ev ? 2 * s->cap_ev : 256;
            s->ev = realloc(s->ev, s->cap_ev * sizeof(*s->ev));
        }
        s->ev[s->n_ev] = (struct deferred){end, s->cur_queue, prog, (u32)s->n_ev, (u32)s->cur_eod};
        s->n_ev++;
        return 0;
    }
    const u8 *pc = s->rose + prog;
    int work_done = 0;
    for (;;) {
        switch (pc[0]) {
        case OP_END: return 0;
        case OP_CHECK_GROUPS: if (!(rd64(pc + 8) & s->groups)) return 0; pc += 16; break;
        case OP_CHECK_MASK:
            if (!check_mask(s, rd64(pc + 8), rd64(pc + 16), rd64(pc + 24), (int32_t)rd32(pc + 32), end)) pc += rd32(pc + 36);
            else pc += 40;
            break;
        case OP_CHECK_BYTE:
            if (!check_byte(s, pc[1], pc[2], pc[3], (int32_t)rd32(pc + 4), end)) pc += rd32(pc + 8); else pc += 16;
            break;
        case OP_CHECK_MASK_32: /* struct ROSE_STRUCT_CHECK_MASK_32, rose/rose_program.h: and 1, cmp 33, neg 68, offset 72, fail 76 */
            if (!check_mask_n(s, pc + 1, pc + 33, rd32(pc + 68), 32, (int32_t)rd32(pc + 72), end)) pc += rd32(pc + 76);
            else pc += 80;
            break;
        case OP_CHECK_MASK_64: /* and 1, cmp 65, neg 136, offset 144, fail 148 */
            if (!check_mask_n(s, pc + 1, pc + 65, rd64(pc + 136), 64, (int32_t)rd32(pc + 144), end)) pc += rd32(pc + 148);
            else pc += 152;
            break;
        case 12: case 13: case 14: case 15: case 67: case 68: { /* CHECK_SHUFTI_16x8, 32x8, 16x16, 32x16, 64x8, 64x16 */
            const u32 fj = shufti_fail_at(pc[0]);
            if (!check_shufti(s, pc, end)) pc += rd32(pc + fj); else pc += (fj + 4 + 7) & ~7u;
            break;
        }
        case OP_CATCH_UP: /* roseCatchUpTo(end), program_runtime.c:2499-2504: engine matches up to `end` come first */
            if (catch_up(s, end) < 0) return -1;
            pc += 8;
            break;
        case OP_CLEAR_WORK_DONE: work_done = 0; pc += 8; break;
        case OP_CHECK_LONG_LIT: case OP_CHECK_MED_LIT: case OP_CHECK_LONG_LIT_NOCASE: case OP_CHECK_MED_LIT_NOCASE: {
            int nocase = pc[0] == OP_CHECK_LONG_LIT_NOCASE || pc[0] == OP_CHECK_MED_LIT_NOCASE;
            if (!check_lit(s, rd32(pc + 4), rd32(pc + 8), nocase, end)) pc += rd32(pc + 12); else pc += 16;
            break;
        }
        case OP_CHECK_EXHAUSTED: {
            u32 ek = rd32(pc + 4);
            if (s->exhausted[ek >> 3] & (1u << (ek & 7))) pc += rd32(pc + 8); else pc += 16;
            break;
        }
        case OP_DEDUPE: {
            u32 dk = rd32(pc + 4);
            u64 to = end + (int32_t)rd32(pc + 8);
            if (s->dedupe_offset != to) { memset(s->dedupe, 0, (s->dkeys + 7) / 8 + 1); s->dedupe_offset = to; }
            if (s->dedupe[dk >> 3] & (1u << (dk & 7))) pc += rd32(pc + 12);
            else { s->dedupe[dk >> 3] |= (u8)(1u << (dk & 7)); pc += 16; }
            break;
        }
        case OP_REPORT:
            if (report(s, rd32(pc + 4), end + (int32_t)rd32(pc + 8), 0xffffffffu, 0xffffffffu) < 0) return -1;
            work_done = 1; pc += 16; break;
        case OP_FINAL_REPORT:
            return report(s, rd32(pc + 4), end + (int32_t)rd32(pc + 8), 0xffffffffu, 0xffffffffu) < 0 ? -1 : 0;
        case OP_REPORT_EXHAUST:
            if (report(s, rd32(pc + 4), end + (int32_t)rd32(pc + 8), rd32(pc + 12), 0xffffffffu) < 0) return -1;
            work_done = 1; pc += 16; break;
        case OP_DEDUPE_AND_REPORT: {
            int rv = report(s, rd32(pc + 8), end + (int32_t)rd32(pc + 12), 0xffffffffu, rd32(pc + 4));
            if (rv < 0) return -1;
            if (rv == 0) { pc += rd32(pc + 16); break; }
            work_done = 1; pc += 24; break;
        }
        case OP_SET_EXHAUST: { u32 ek = rd32(pc + 4); s->exhausted[ek >> 3] |= (u8)(1u << (ek & 7)); work_done = 1; pc += 8; break; }
        case OP_SQUASH_GROUPS: if (work_done) s->groups &= rd64(pc + 8); pc += 16; break;
        case OP_INCLUDED_JUMP: pc += 8; break; /* scratch->fdr_conf == NULL fall-through: every bit is confirmed on its own */
        default: s->error = 1; return -1;
        }
    }
}

/* do_confirm_fdr + confWithBit for one (end index i, bucket) pair */
static int confirm(struct scan_state *s, const u8 *conf_base, u32 bucket, u32 i) {
    u32 cf = rd32(conf_base + 4 * bucket);
    if (!cf) return 0;
    const u8 *fdrc = conf_base + cf;
    if (!(rd64(fdrc + C_GROUPS) & s->groups)) return 0;
    u64 v = 0; /* 8 bytes ending at i, zero before the buffer (block mode: fake zero history) */
    for (int k = 0; k < 8; k++) { int64_t o = (int64_t)i - 7 + k; if (o >= 0) v |= (u64)s->buf[o] << (8 * k); }
    u32 nbits = rd32(fdrc + C_NBITS);
    u32 h = (u32)(((v & rd64(fdrc + C_ANDMSK)) * rd64(fdrc + C_MULT)) >> (64 - nbits));
    u32 start = rd32(fdrc + C_SIZE + 4 * (size_t)h);
    if (!start) return 0;
    const u8 *li = fdrc + start;
    for (;;) {
        if ((v & rd64(li + L_MSK)) == rd64(li + L_V) && li[L_SIZE] <= i + 1 && (rd64(li + L_GROUPS) & s->groups)) {
            if (run_program(s, rd32(li + L_ID), (u64)i + 1) < 0) return -1;
        }
        if (!li[L_NEXT]) break;
        li += L_STRUCT;
    }
    return 0;
}

/* FDR: bucket mask of candidates at end index e (position-local form of the shift-or recurrence) */
static u32 fdr_candidates(const u8 *fdr, const struct scan_state *s, u32 e) {
    const u64 *ft = (const u64 *)(fdr + F_TABLE);
    u32 stride = fdr[F_STRIDE], dmask = rd16(fdr + F_DOMAIN_MASK);
    u8 st = e < 8 ? fdr[F_START + e] : 0; /* initial state: bit b set = bucket b cannot end here yet */
    for (u32 p = 0; p < 8 && p <= e; p++) {
        u32 x = e - p;
        if (x % stride) continue;
        u32 key = s->buf[x] | ((x + 1 < s->len ? s->buf[x + 1] : 0u) << 8);
        st |= (u8)(ft[key & dmask] >> (8 * p));
    }
    return (u8)~st;
}

/* Teddy: r_j[i] = lo_j[c & 15] | hi_j[c >> 4]; candidate buckets at e = ~(OR_j r_j[e - j]).  width = 1: 8 buckets;
 * width = 2 (fat Teddy, fdr/teddy_avx2.c): every mask twice, buckets 0-7 then 8-15 (fdr/teddy_compile.cpp:436-500) */
static u32 teddy_candidates(const u8 *teddy, u32 masks, u32 width, const struct scan_state *s, u32 e) {
    const u8 *mb = teddy + 64;
    u32 cand = 0;
    for (u32 g = 0; g < width; g++) {
        u8 r = 0;
        for (u32 j = 0; j < masks && j <= e; j++) {
            u8 c = s->buf[e - j];
            r |= (u8)(mb[((2 * j) * width + g) * 16 + (c & 15)] | mb[((2 * j + 1) * width + g) * 16 + (c >> 4)]);
        }
        cand |= (u32)(u8)~r << (8 * g);
    }
    return cand;
}

static int hwlm_exec(struct scan_state *s, const u8 *hwlm) {
    if (!s->groups) return 0;
    const u8 *eng = hwlm + HWLM_BODY;
    if (hwlm[0] == 16) { /* noodle */
        u32 n = eng[N_MSK_LEN];
        u64 msk = rd64(eng + N_MSK), cmp = rd64(eng + N_CMP);
        if (s->len < n) return 0;
        for (u32 e = n - 1 > s->start ? n - 1 : s->start; e < s->len; e++) {
            u64 v = 0;
            for (u32 k = 0; k < n; k++) v |= (u64)s->buf[e + 1 - n + k] << (8 * k);
            if ((v & msk) == cmp && run_program(s, rd32(eng + N_ID), (u64)e + 1) < 0) return -1;
        }
        return 0;
    }
    u32 id = rd32(eng + F_ENGINE_ID);
    const u8 *conf_base = eng + rd32(eng + F_CONF_OFF);
    for (u32 e = s->start; e < s->len; e++) {
        u32 cand;
        if (id == 0) cand = fdr_candidates(eng, s, e);
        else if (id >= 11 && id <= 18) cand = teddy_candidates(eng, (id - 11) / 2 + 1, 1, s, e);
        else if (id >= 3 && id <= 10) cand = teddy_candidates(eng, (id - 3) / 2 + 1, 2, s, e);
        else { s->error = 1; return -1; }
        for (u32 b = 0; cand; b++, cand >>= 1)
            if ((cand & 1) && confirm(s, conf_base, b, e) < 0) return -1;
    }
    return 0;
}


/* ------------------------------------------------------------------------------------------------------------
 * Single-outfix databases whose engine is a LimEx NFA.
 *   limex_scan         soleOutfixBlockExec          runtime.c:246-279   (START, TOP at 0, END at len; EOD accepts)
 *                      nfaExecLimEx*_Q / STREAM_FN  nfa/limex_runtime_impl.h:620-714, 246-367
 *   limex_step         NFA_EXEC_GET_LIM_SUCC        nfa/limex_runtime_impl.h:164-199 (shifts, per 64-bit lane)
 *                      processExceptional           nfa/limex_exceptional.h:92-204,212-368 (reports at offset i,
 *                                                   squash on succ, successors OR-ed in afterwards)
 *   limex_accepts      moProcessAccepts / TESTEOD   nfa/limex_common_impl.h:116-213
 *   reports            limexRunReports -> roseReportAdaptor -> roseRunProgram    nfa/limex_runtime.h:90-117,
 *                                                   rose/match.c:611-633
 * State vectors are handled as arrays of 64-bit words (LimEx32: one word whose upper half stays clear because
 * every reach row has it clear).  Acceleration and the exception cache are shortcuts without effect on results. */
enum {
    R_NFA_INFO = 236, NI_NFA_OFF = 0, NFA_FLAGS = 0, NFA_TYPE = 8, NFA_RACCEL = 9, NFA_SIZE = 64, NFA_ACCEPTS_EOD = 1,
    LX_REACH_SIZE = 256, LX_ACCEPT_COUNT = 276, LX_ACCEPT_OFF = 280, LX_ACCEPT_EOD_COUNT = 284, LX_ACCEPT_EOD_OFF = 288,
    LX_EXC_COUNT = 292, LX_EXC_OFF = 296, LX_REPEAT_COUNT = 300, LX_HEAD = 332
};
struct lx_layout { u32 bytes, words, init, accept, accept_eod, exc_mask, shift, shift_count, shift_amount, size, e_size; };
static struct lx_layout lx_layout_of(u32 type) {
    static const u32 bits_of[6] = {32, 64, 128, 256, 384, 512}, align_of[6] = {4, 8, 16, 32, 16, 64};
    struct lx_layout l;
    u32 S = bits_of[type] / 8, A = align_of[type];
    l.bytes = S; l.words = (S + 7) / 8;
    l.init = (LX_HEAD + A - 1) / A * A;            /* twelve state-sized masks follow the head, then shift[8] */
    l.accept = l.init + 2 * S; l.accept_eod = l.init + 3 * S; l.exc_mask = l.init + 9 * S;
    l.shift = l.init + 12 * S; l.shift_count = l.shift + 8 * S; l.shift_amount = l.shift_count + 4;
    l.size = (l.shift_amount + 8 + 63) / 64 * 64 + 192;  /* three m512 exception-extraction masks close the struct */
    l.e_size = (2 * S + 10 + A - 1) / A * A;       /* NFAException: squash, successors, reports, repeatOffset, 2 x u8 */
    return l;
}
static void lx_load(u64 *dst, const u8 *src, const struct lx_layout *l) {
    memset(dst, 0, 64);
    memcpy(dst, src, l->bytes);
}
static u32 lx_rank(const u64 *mask, u32 bit) { /* number of mask bits below `bit` */
    u32 n = 0;
    for (u32 w = 0; w < bit / 64; w++) n += (u32)__builtin_popcountll(mask[w]);
    return n + (u32)__builtin_popcountll(mask[bit / 64] & ((1ull << (bit % 64)) - 1));
}
static int lx_reports(struct scan_state *s, const u8 *limex, u32 list, u64 offset) {
    for (const u8 *p = limex + list; rd32(p) != 0xffffffffu; p += 4)
        if (run_program(s, rd32(p), offset) < 0) return -1;
    return 0;
}
static int lx_accepts(struct scan_state *s, const u8 *limex, const struct lx_layout *l, const u64 *state, u32 mask_at,
                      u32 table_off, u64 offset) {
    u64 mask[8];
    lx_load(mask, limex + mask_at, l);
    for (u32 b = 0; b < l->bytes * 8; b++) {
        if (!((state[b / 64] & mask[b / 64]) >> (b % 64) & 1)) continue;
        const u8 *a = limex + table_off + 12 * (size_t)lx_rank(mask, b); /* struct NFAAccept {u8 single; u32 reports; u32 squash} */
        if (a[0]) { if (run_program(s, rd32(a + 4), offset) < 0) return -1; }
        else if (lx_reports(s, limex, rd32(a + 4), offset) < 0) return -1;
    }
    return 0;
}

static int limex_scan(struct scan_state *s, const u8 *nfa) {
    u32 type = nfa[NFA_TYPE];
    if (type > 5 || nfa[NFA_RACCEL]) { s->error = 1; return -1; }
    const u8 *limex = nfa + NFA_SIZE;
    if (rd32(limex + LX_REPEAT_COUNT)) { s->error = 1; return -1; } /* bounded repeats: outside the restated path */
    const struct lx_layout l = lx_layout_of(type);
    const u8 *reach = limex + l.size, *exc = limex + rd32(limex + LX_EXC_OFF);
    u32 nshift = rd32(limex + l.shift_count);
    u64 st[8], excm[8], shm[8][8];
    lx_load(st, limex + l.init, &l);                       /* MQE_TOP at offset 0: s = init */
    lx_load(excm, limex + l.exc_mask, &l);
    for (u32 k = 0; k < nshift; k++) lx_load(shm[k], limex + l.shift + (size_t)k * l.bytes, &l);
    const u64 lane_mask = l.bytes == 4 ? 0xffffffffull : ~0ull;
    for (u32 i = 0; i < s->len; i++) {
        u64 succ[8] = {0}, local[8] = {0}, any = 0;
        for (u32 w = 0; w < l.words; w++) any |= st[w];
        if (!any) return 0;                                  /* nothing on: stays that way */
        for (u32 k = 0; k < nshift; k++)
            for (u32 w = 0; w < l.words; w++)
                succ[w] |= ((st[w] & shm[k][w]) << limex[l.shift_amount + k]) & lane_mask;
        for (u32 b = 0; b < l.bytes * 8; b++) {              /* exceptions, lowest state first */
            if (!((st[b / 64] & excm[b / 64]) >> (b % 64) & 1)) continue;
            const u8 *e = exc + (size_t)lx_rank(excm, b) * l.e_size;
            if (e[2 * l.bytes + 9]) { s->error = 1; return -1; }       /* trigger: repeats only */
            u32 reports = rd32(e + 2 * l.bytes);
            if (reports != 0xffffffffu && lx_reports(s, limex, reports, i) < 0) return -1;
            u64 t[8];
            lx_load(t, e + l.bytes, &l);
            for (u32 w = 0; w < l.words; w++) local[w] |= t[w];
            u8 sq = e[2 * l.bytes + 8];
            if (sq == 1 || sq == 3) {                        /* LIMEX_SQUASH_CYCLIC / _REPORT: on succ, not local */
                lx_load(t, e, &l);
                for (u32 w = 0; w < l.words; w++) succ[w] &= t[w];
            }
        }
        u64 r[8];
        lx_load(r, reach + (size_t)limex[s->buf[i]] * l.bytes, &l);  /* reach[reachMap[c]] */
        for (u32 w = 0; w < l.words; w++) st[w] = (succ[w] | local[w]) & r[w];
    }
    if (rd32(limex + LX_ACCEPT_COUNT) &&
        lx_accepts(s, limex, &l, st, l.accept, rd32(limex + LX_ACCEPT_OFF), s->len) < 0) return -1;
    u64 any = 0;
    for (u32 w = 0; w < l.words; w++) any |= st[w];
    s->cur_eod = 1;
    if (any && (rd32(nfa + NFA_FLAGS) & NFA_ACCEPTS_EOD) && rd32(limex + LX_ACCEPT_EOD_COUNT) &&
        lx_accepts(s, limex, &l, st, l.accept_eod, rd32(limex + LX_ACCEPT_EOD_OFF), s->len) < 0) { s->cur_eod = 0; return -1; }
    s->cur_eod = 0;
    return 0;
}


/* ------------------------------------------------------------------------------------------------------------
 * McClellan DFA, sole-outfix block run: src/runtime.c:246-279 -> nfaExecMcClellan{8,16}_Q (mcclellan.c:961-999,
 * Q2i :626-759 / :789-925: MQE_TOP at offset 0 sets s = start_anchored) -> mcclellanExec{16,8}_i (:227-366 /
 * :445-565) with doNormal16 (:122-167), doNormal8 (:413-443), doSherman16 (mcclellan_common_impl.h:54-84),
 * doComplexReport (mcclellan.c:43-89), then nfaCheckFinalState -> mcclellanCheckEOD (:609-623).
 * struct mcclellan / mstate_aux offsets: mcclellan_internal.h:63-101 (pinned by tests/test_layout.py).
 * Acceleration is a skip over bytes that cannot change the state; wide states are outside the restated path. */
enum { MC_STATE_COUNT = 0, MC_START_ANCHORED = 8, MC_AUX_OFFSET = 12, MC_SHERMAN_OFFSET = 16, MC_ACCEPT_LIMIT_8 = 26,
       MC_SHERMAN_LIMIT = 28, MC_WIDE_LIMIT = 30, MC_ALPHA_SHIFT = 32, MC_FLAGS = 33, MC_HAS_WIDE = 35, MC_WIDE_OFFSET = 304, MC_REMAP = 36, MC_ARB_REPORT = 292,
       MC_SIZE = 308, MC_AUX_SIZE = 16, MC_AUX_ACCEPT = 0, MC_AUX_ACCEPT_EOD = 4 };

static int mc_report(struct scan_state *s, const u8 *nfa, const u8 *m, u32 st, u64 loc, int eod) {
    if (!eod && (m[MC_FLAGS] & 1)) return run_program(s, rd32(m + MC_ARB_REPORT), loc);   /* MCCLELLAN_FLAG_SINGLE */
    const u8 *aux = nfa + rd32(m + MC_AUX_OFFSET) + (size_t)st * MC_AUX_SIZE;
    u32 off = rd32(aux + (eod ? MC_AUX_ACCEPT_EOD : MC_AUX_ACCEPT));
    if (!off) return 0;
    const u8 *rl = nfa + off;                                  /* struct report_list {u32 count; ReportID report[]} */
    for (u32 i = 0, n = rd32(rl); i < n; i++)
        if (run_program(s, rd32(rl + 4 + 4 * (size_t)i), loc) < 0) return -1;
    return 0;
}

static int mcclellan_scan(struct scan_state *s, const u8 *nfa) {
    const int wide16 = nfa[NFA_TYPE] == 7;
    const u8 *m = nfa + NFA_SIZE, *succ = m + MC_SIZE;
    if (nfa[NFA_RACCEL] || (m[MC_HAS_WIDE] && !wide16)) { s->error = 1; return -1; }
    /* wide states (16-bit engines): ids >= wide_limit name entries {u16 width; u8 symbols[width]; (pad to 2); u16 next;
     * u16 failure[alphabet]} found through the u32 offset list at wide_base + 4 (mcclellan_internal.h:53-61, :138-148) */
    const u32 wide_limit = m[MC_HAS_WIDE] ? rd16(m + MC_WIDE_LIMIT) : 0xffffffffu;
    const u8 *wide_base = nfa + rd32(m + MC_WIDE_OFFSET);
    const u32 as = m[MC_ALPHA_SHIFT], accept_limit = rd16(m + MC_ACCEPT_LIMIT_8), sherman_limit = rd16(m + MC_SHERMAN_LIMIT);
    u32 st = rd16(m + MC_START_ANCHORED);
    int in_wide = 0;
    for (u32 i = 0; i < s->len && st; i++) {
        const u32 cp = m[MC_REMAP + s->buf[i]];
        int accept;
        if (wide16) {
            u32 e;
            if (st >= wide_limit) {
                /* doWide16 (mcclellan_common_impl.h:86-188), one call of the block scan (offset 0): the data is compared
                 * with the state's symbols; all of them -> `next`, data ends first -> the scan ends inside the wide state
                 * (no accept, no end-of-data accept: :615), a differing byte -> failure[that byte], which consumes it */
                const u8 *we = wide_base + rd32(wide_base + 4 + 4 * (size_t)(st - wide_limit));
                const u32 width = rd16(we);
                const u8 *sym = we + 2, *tr = we + 2 + ((width + 1) & ~1u);
                u32 j = 0;
                while (j < width && i + j < s->len && m[MC_REMAP + s->buf[i + j]] == sym[j]) j++;
                if (j == width) { e = rd16(tr); i += width - 1; }
                else if (i + j == s->len) { in_wide = 1; break; }
                else { e = rd16(tr + 2 + 2 * (size_t)m[MC_REMAP + s->buf[i + j]]); i += j; }
            } else
            if (st < sherman_limit) e = rd16(succ + 2 * (((size_t)st << as) + cp));
            else {
                const u8 *sh = nfa + rd32(m + MC_SHERMAN_OFFSET) + 32 * (size_t)(st - sherman_limit);
                u32 len = sh[1], k;
                for (k = 0; k < len && sh[4 + k] != cp; k++) {}
                if (k < len) e = rd16(sh + 4 + len + 2 * k);
                else e = rd16(succ + 2 * (((size_t)rd16(sh + 2) << as) + cp));
            }
            accept = (e & 0x8000u) != 0;
            st = e & 0x3fffu;
        } else {
            st = succ[((size_t)st << as) + cp];
            accept = st >= accept_limit;
        }
        if (accept && mc_report(s, nfa, m, st, (u64)i + 1, 0) < 0) return -1;
    }
    s->cur_eod = 1;
    if (st && !in_wide && st < wide_limit && (rd32(nfa + NFA_FLAGS) & NFA_ACCEPTS_EOD) && mc_report(s, nfa, m, st, s->len, 1) < 0) { s->cur_eod = 0; return -1; }
    s->cur_eod = 0;
    return 0;
}

/* ------------------------------------------------------------------------------------------------------------
 * Sheng DFA (<= 16 states), sole-outfix block run: src/runtime.c:246-279 -> nfaExecSheng_Q -> runSheng
 * (sheng.c:510-700: MQE_TOP at offset 0 sets state = sh->anchored) -> runShengCb (:359-408) -> SHENG_IMPL
 * (sheng_impl.h:40-97: state = shuffle_masks[c][state]; ACCEPT flag -> fireSingleReport / fireReports, sheng.c:106-157),
 * MQE_END returns MO_DEAD for a dead state when SHENG_FLAG_CAN_DIE (:672-677), which skips the EOD check of
 * runtime.c:273; otherwise nfaExecSheng_testEOD (:796-812).  Layout: sheng_internal.h:56-79. */
/* struct sheng / sheng32 / sheng64 (nfa/sheng_internal.h:68-107): masks of `stride` bytes per byte value, then the trailer */
struct sh_shape { u32 stride, accept, dead, mask, trailer; };
static struct sh_shape sh_shape_of(u8 nfa_type) {
    if (nfa_type == 21) return (struct sh_shape){64, 0x20, 0x40, 0x1f, 16384};
    if (nfa_type == 22) return (struct sh_shape){64, 0x40, 0x80, 0x3f, 16384};
    return (struct sh_shape){16, 0x10, 0x20, 0x0f, 4096};
}
/* trailer: u32 length, aux_offset, report_offset, accel_offset; u8 n_states, anchored, floating, flags; ReportID report */

static int sh_report(struct scan_state *s, const u8 *nfa, const u8 *sh, const struct sh_shape *f, u8 st, u64 loc, int eod) {
    const u8 *tr = sh + f->trailer;
    if (!eod && (tr[19] & 1)) return run_program(s, rd32(tr + 20), loc);                    /* SHENG_FLAG_SINGLE_REPORT */
    const u8 *aux = nfa + rd32(tr + 4) + (size_t)(st & f->mask) * 16;                       /* struct sstate_aux */
    u32 off = rd32(aux + (eod ? 4 : 0));
    if (!off) return 0;
    const u8 *rl = nfa + off;
    for (u32 i = 0, n = rd32(rl); i < n; i++)
        if (run_program(s, rd32(rl + 4 + 4 * (size_t)i), loc) < 0) return -1;
    return 0;
}

static int sheng_scan(struct scan_state *s, const u8 *nfa) {
    const u8 *sh = nfa + NFA_SIZE;
    if (nfa[NFA_RACCEL]) { s->error = 1; return -1; }
    const struct sh_shape f = sh_shape_of(nfa[NFA_TYPE]);
    const u8 *tr = sh + f.trailer;
    const int can_die = (tr[19] & 2) != 0;
    u8 st = tr[17];                                                /* anchored */
    for (u32 i = 0; i < s->len; i++) {
        if (st & f.dead) break;                                    /* dead: every successor is the same state */
        st = sh[(size_t)f.stride * s->buf[i] + (st & f.mask)];
        if ((st & f.accept) && sh_report(s, nfa, sh, &f, st, (u64)i + 1, 0) < 0) return -1;
    }
    if (can_die && (st & f.dead)) return 0;                        /* MO_DEAD: no final-state check */
    if (rd32(nfa + NFA_FLAGS) & NFA_ACCEPTS_EOD) {
        const u8 *aux = nfa + rd32(tr + 4) + (size_t)(st & f.mask) * 16;
        s->cur_eod = 1;
        const int rv = rd32(aux + 4) ? sh_report(s, nfa, sh, &f, st, s->len, 1) : 0;
        s->cur_eod = 0;
        if (rv < 0) return -1;
    }
    return 0;
}

static int deferred_cmp(const void *a, const void *b) {
    const struct deferred *x = a, *y = b;
    if (x->eod != y->eod) return x->eod < y->eod ? -1 : 1;   /* the EOD program runs after every engine was caught up */
    if (x->to != y->to) return x->to < y->to ? -1 : 1;
    if (x->queue != y->queue) return x->queue < y->queue ? -1 : 1;
    return x->seq < y->seq ? -1 : x->seq > y->seq;
}

/* McSheng hybrid DFA (nfa types 19 / 20): nfaExecMcSheng{8,16}_Q -> mcshengExec{8,16}_i (nfa/mcsheng.c:423-590, :654-810).
 * A state below sheng_end steps through the byte-shuffle masks (doSheng :150-322: sheng_masks[c][s - 1]; the value
 * sheng_end - 1 is the dead state, a smaller one the region state value + 1, a larger one the id of a state outside the
 * region); any other state through the successor table, whose rows start at state sheng_end (doNormal16 :368-420,
 * doNormal8 :612-650; Sherman states as McClellan's).  Leaving the region, the 16-bit engine reads the accept from the
 * new state's aux record (:466), the 8-bit one from accept_limit_8 (:698).  EOD: mcshengCheckEOD (:830-840).
 * struct mcsheng: nfa/mcsheng_internal.h:63-88. */
enum { MS_SHENG_END = 24, MS_ACCEPT_LIMIT_8 = 30, MS_SHERMAN_LIMIT = 32, MS_ALPHA_SHIFT = 34, MS_FLAGS = 35, MS_REMAP = 37,
       MS_ARB_REPORT = 296, MS_MASKS = 304, MS_SIZE = 4400 };

static int ms_report(struct scan_state *s, const u8 *nfa, const u8 *m, u32 st, u64 loc, int eod) {
    if (!eod && (m[MS_FLAGS] & 1)) return run_program(s, rd32(m + MS_ARB_REPORT), loc);
    const u8 *aux = nfa + rd32(m + MC_AUX_OFFSET) + (size_t)st * MC_AUX_SIZE;
    u32 off = rd32(aux + (eod ? MC_AUX_ACCEPT_EOD : MC_AUX_ACCEPT));
    if (!off) return 0;
    const u8 *rl = nfa + off;
    for (u32 i = 0, n = rd32(rl); i < n; i++)
        if (run_program(s, rd32(rl + 4 + 4 * (size_t)i), loc) < 0) return -1;
    return 0;
}

static int mcsheng_scan(struct scan_state *s, const u8 *nfa) {
    /* types 23 / 24: struct mcsheng64 (mcsheng_internal.h:94-119, doSheng64 mcsheng.c:1489-1660): the same fields, up to 64
     * region states, 64-byte masks at offset 320 */
    const int wide64 = nfa[NFA_TYPE] == 23 || nfa[NFA_TYPE] == 24;
    const int wide16 = nfa[NFA_TYPE] == 20 || nfa[NFA_TYPE] == 24;
    const u32 masks_at = wide64 ? 320 : MS_MASKS, stride = wide64 ? 64 : 16;
    const u8 *m = nfa + NFA_SIZE, *succ = m + (wide64 ? 320 + 256 * 64 : MS_SIZE);
    if (nfa[NFA_RACCEL]) { s->error = 1; return -1; }
    const u32 as = m[MS_ALPHA_SHIFT], accept_limit = rd16(m + MS_ACCEPT_LIMIT_8), sheng_end = rd16(m + MS_SHENG_END);
    const u32 sherman_limit = wide16 ? rd16(m + MS_SHERMAN_LIMIT) : 0xffffffffu;
    const u8 *aux0 = nfa + rd32(m + MC_AUX_OFFSET);
    u32 st = rd16(m + MC_START_ANCHORED);
    for (u32 i = 0; i < s->len && st; i++) {
        const u8 c = s->buf[i];
        int accept;
        if (st < sheng_end) {
            const u32 v = m[masks_at + stride * (size_t)c + st - 1];
            st = v == sheng_end - 1 ? 0 : v < sheng_end - 1 ? v + 1 : v;
            accept = wide16 ? rd32(aux0 + (size_t)st * MC_AUX_SIZE + MC_AUX_ACCEPT) != 0 : st >= accept_limit;
        } else {
            const u32 cp = m[MS_REMAP + c];
            if (wide16) {
                u32 e;
                if (st < sherman_limit) e = rd16(succ + 2 * (((size_t)(st - sheng_end) << as) + cp));
                else {
                    const u8 *sh = nfa + rd32(m + MC_SHERMAN_OFFSET) + 32 * (size_t)(st - sherman_limit);
                    u32 len = sh[1], k;
                    for (k = 0; k < len && sh[4 + k] != cp; k++) {}
                    if (k < len) e = rd16(sh + 4 + len + 2 * k);
                    else e = rd16(succ + 2 * (((size_t)(rd16(sh + 2) - sheng_end) << as) + cp));
                }
                accept = (e & 0x8000u) != 0;
                st = e & 0x3fffu;
            } else {
                st = succ[((size_t)(st - sheng_end) << as) + cp];
                accept = st >= accept_limit;
            }
        }
        if (accept && ms_report(s, nfa, m, st, (u64)i + 1, 0) < 0) return -1;
    }
    s->cur_eod = 1;
    if (st && (rd32(nfa + NFA_FLAGS) & NFA_ACCEPTS_EOD) && ms_report(s, nfa, m, st, s->len, 1) < 0) { s->cur_eod = 0; return -1; }
    s->cur_eod = 0;
    return 0;
}

static void engine_scan(struct scan_state *s, const u8 *nfa) {
    if (nfa[NFA_TYPE] == 6 || nfa[NFA_TYPE] == 7) mcclellan_scan(s, nfa);
    else if (nfa[NFA_TYPE] == 19 || nfa[NFA_TYPE] == 20 || nfa[NFA_TYPE] == 23 || nfa[NFA_TYPE] == 24) mcsheng_scan(s, nfa);
    else if (nfa[NFA_TYPE] == 17 || nfa[NFA_TYPE] == 21 || nfa[NFA_TYPE] == 22) sheng_scan(s, nfa);
    else limex_scan(s, nfa);
}

/* roseCatchUpTo (rose/catchup.h:128-167): deliver the outfix engines' pending reports with offset <= end, in offset
 * order (queue order inside one offset), through their report programs. */
static int catch_up(struct scan_state *s, u64 end) {
    while (s->ev_next < s->n_ev && s->ev[s->ev_next].to <= end && !s->ev[s->ev_next].eod) {
        const struct deferred e = s->ev[s->ev_next++];
        if (run_program(s, e.prog, e.to) < 0) return -1;
    }
    return 0;
}

/* roseBlockExec (rose/block.c:345-422) for the restated full-Rose shapes: blockInitSufPQ (rose/catchup.c:612-672) starts
 * every outfix whose NFA.minWidth the block satisfies (nfaRevAccelCheck, nfa/nfa_rev_api.h:135-155); the floating literal
 * matcher runs with roseFloatingCallback, whose programs begin with CATCH_UP; roseCatchUpTo(length) delivers what is
 * left.  The engines are stepped up front and their reports replayed at the catch-up points, which gives the same
 * callback sequence as the reference's priority queue of engines. */
static void full_rose_exec(struct scan_state *s, const u8 *rose) {
    const u32 q0 = rd32(rose + 396), q1 = rd32(rose + 400); /* outfixBeginQueue, outfixEndQueue */
    s->defer = 1;
    for (u32 qi = q0; qi < q1 && !s->error; qi++) {
        const u8 *nfa = rose + rd32(rose + rd32(rose + R_NFA_INFO) + 20 * qi + NI_NFA_OFF);
        if (rd32(nfa + 36) > s->len) continue; /* NFA.minWidth */
        s->cur_queue = qi;
        engine_scan(s, nfa);
    }
    s->defer = 0;
    qsort(s->ev, s->n_ev, sizeof(*s->ev), deferred_cmp);
    s->ev_next = 0;
    if (rd32(rose + R_FMATCHER) && s->len >= rd32(rose + 116) /* fmatcherMinWidth */ && !s->error)
        hwlm_exec(s, rose + rd32(rose + R_FMATCHER));
    if (!s->terminated && !s->error) catch_up(s, s->len);
    /* roseBlockEodExec (rose/block.c:163-190): ENGINES_EOD -> nfaCheckFinalState of every engine still alive, queue order */
    while (!s->terminated && !s->error && s->ev_next < s->n_ev) {
        const struct deferred e = s->ev[s->ev_next++];
        if (run_program(s, e.prog, e.to) < 0) break;
    }
    free(s->ev);
    s->ev = NULL;
}

/* hs_scan for block-mode pure-literal, single-outfix and outfix-only full-Rose (LimEx, McClellan or Sheng) databases.  `db` = in-memory hs_database_t. */
int op_scan(const void *db, const char *data, unsigned length, op_match_cb cb, void *ctx) {
    const u8 *d = (const u8 *)db;
    if (!d || !data || rd32(d + DB_MAGIC_OFF) != 0xdbdbdbdbu) return HS_INVALID;
    if (rd32(d + DB_VERSION_OFF) != ((5u << 24) | (4u << 16) | (11u << 8))) return HS_DB_VERSION_ERROR;
    const u8 *rose = d + rd32(d + DB_BYTECODE_OFF);
    if (rd32(rose + R_MODE) != 1) return HS_DB_MODE_ERROR;
    if (rose[R_RUNTIME_IMPL] > 2 || rose[R_HAS_SOM] || rd32(rose + R_CKEY_COUNT) || rd32(rose + R_BOUNDARY) ||
        rd32(rose + R_BOUNDARY + 4) || rd32(rose + R_BOUNDARY + 8))
        return HS_ARCH_ERROR; /* outside the restated path */
    if (rose[R_RUNTIME_IMPL] == 0) {
        /* full Rose: a floating literal matcher and / or outfix engines; no roles, leftfixes, other matchers, EOD program */
        if (rd32(rose + 88) || rd32(rose + 92) || rd32(rose + 100) || rd32(rose + 104) || rd32(rose + 108) ||
            (rd32(rose + R_FMATCHER) && (rose[1] || rd32(rose + 224) || rd32(rose + 220) != 0xffffffffu || rd32(rose + 128) != 0xffffffffu)) ||
            rd32(rose + 84) || rd32(rose + 64) || rd32(rose + 152) ||
            (rd32(rose + 184) && (rose[rd32(rose + 184)] != 48 || rose[rd32(rose + 184) + 8] != 0)) || /* eodProgram: ENGINES_EOD; END */
            rd32(rose + 404) != rd32(rose + 156) || rd32(rose + 400) != rd32(rose + 156) || rd32(rose + 408) != 0xffffffffu)
            return HS_ARCH_ERROR;
    }
    if (rd32(rose + R_MIN_WIDTH) > length) return HS_SUCCESS;
    if (!length) return HS_SUCCESS;
    if (rd32(rose + R_MIN_WIDTH_EXCL) > length) return HS_SUCCESS;
    if (rd32(rose + R_MAX_BI) != 0xffffffffu && length > rd32(rose + R_MAX_BI)) return HS_SUCCESS;
    struct scan_state s;
    memset(&s, 0, sizeof(s));
    s.rose = rose; s.buf = (const u8 *)data; s.len = length; s.cb = cb; s.ctx = ctx;
    s.groups = rd64(rose + R_INIT_GROUPS) & rd64(rose + R_FLOAT_MASK);
    s.ekeys = rd32(rose + R_EKEY_COUNT); s.dkeys = rd32(rose + R_DKEY_COUNT);
    s.exhausted = calloc(s.ekeys / 8 + 2, 1);
    s.dedupe = calloc(s.dkeys / 8 + 2, 1);
    s.dedupe_offset = ~0ull;
    if (rose[R_RUNTIME_IMPL] == 0) full_rose_exec(&s, rose);
    else if (rose[R_RUNTIME_IMPL] == 2) engine_scan(&s, rose + rd32(rose + rd32(rose + R_NFA_INFO) + NI_NFA_OFF));
    else hwlm_exec(&s, rose + rd32(rose + R_FMATCHER));
    free(s.exhausted); free(s.dedupe);
    if (s.error) return HS_UNKNOWN_ERROR;
    return s.terminated ? HS_SCAN_TERMINATED : HS_SUCCESS;
}

/* hs_scan_vector (runtime.c:1106-1174) for HS_MODE_VECTORED pure-literal databases: the buffers are consecutive blocks
 * of one stream.  Block by block as hs_scan_stream_internal (runtime.c:880-975) does it: the last historyRequired bytes
 * of the stream stand in front of the block (getHistory :77-85, maintainHistoryBuffer :483-508); the literal matcher
 * reports only matches ending inside the block (hwlmExecStreaming) but confirms them, and runs their programs' literal
 * and lookaround checks, against history + block (fdr_confirm_runtime.h:77-88, program_runtime.c:1988-2011); offsets
 * run on from block to block; exhaustion lasts to the end of the call. */
int op_scan_vector(const void *db, const char *const *data, const unsigned *length, unsigned count, op_match_cb cb, void *ctx) {
    const u8 *d = (const u8 *)db;
    if (!d || !data || !length || rd32(d + DB_MAGIC_OFF) != 0xdbdbdbdbu) return HS_INVALID;
    if (rd32(d + DB_VERSION_OFF) != ((5u << 24) | (4u << 16) | (11u << 8))) return HS_DB_VERSION_ERROR;
    const u8 *rose = d + rd32(d + DB_BYTECODE_OFF);
    if (rd32(rose + R_MODE) != 4 && rd32(rose + R_MODE) != 2) return HS_DB_MODE_ERROR; /* mode 2 (HS_MODE_STREAM): the same loop is
                                                                                         * hs_open_stream, hs_scan_stream per block, hs_close_stream */
    if (rose[R_RUNTIME_IMPL] == 2 && rd32(rose + R_MODE) == 2 && !rose[R_HAS_SOM] && !rd32(rose + R_CKEY_COUNT)) {
        /* stream mode, one DFA engine (soleOutfixStreamExec, runtime.c:833-870): the engine's state is compressed into the
         * stream state after a block and expanded before the next (nfaQueueCompressState / nfaExpandState) -- for a DFA
         * that is its state id, so the blocks are one run of the automaton over their concatenation; restated as that.
         * (Engines with end-of-data accepts report at hs_close_stream: engine_scan's end-of-data check is that report.) */
        size_t total = 0;
        for (unsigned b = 0; b < count; b++) total += length[b];
        u8 *all = malloc(total + 1);
        size_t at = 0;
        for (unsigned b = 0; b < count; b++) { if (length[b]) memcpy(all + at, data[b], length[b]); at += length[b]; }
        struct scan_state e;
        memset(&e, 0, sizeof(e));
        e.rose = rose; e.buf = all; e.len = (u32)total; e.cb = cb; e.ctx = ctx;
        e.ekeys = rd32(rose + R_EKEY_COUNT); e.dkeys = rd32(rose + R_DKEY_COUNT);
        e.exhausted = calloc(e.ekeys / 8 + 2, 1);
        e.dedupe = calloc(e.dkeys / 8 + 2, 1);
        e.dedupe_offset = ~0ull;
        if (total) engine_scan(&e, rose + rd32(rose + rd32(rose + R_NFA_INFO) + NI_NFA_OFF));
        free(all); free(e.exhausted); free(e.dedupe);
        if (e.error) return HS_UNKNOWN_ERROR;
        return e.terminated ? HS_SCAN_TERMINATED : HS_SUCCESS;
    }
    if (rose[R_RUNTIME_IMPL] != 1 || rose[R_HAS_SOM] || rd32(rose + R_CKEY_COUNT) || rd32(rose + R_BOUNDARY) ||
        rd32(rose + R_BOUNDARY + 4) || rd32(rose + R_BOUNDARY + 8))
        return HS_ARCH_ERROR; /* outside the restated path */
    const u32 hreq = rd32(rose + 16); /* historyRequired */
    struct scan_state s;
    memset(&s, 0, sizeof(s));
    s.rose = rose; s.cb = cb; s.ctx = ctx;
    s.groups = rd64(rose + R_INIT_GROUPS) & rd64(rose + R_FLOAT_MASK);
    s.ekeys = rd32(rose + R_EKEY_COUNT); s.dkeys = rd32(rose + R_DKEY_COUNT);
    s.exhausted = calloc(s.ekeys / 8 + 2, 1);
    s.dedupe = calloc(s.dkeys / 8 + 2, 1);
    s.dedupe_offset = ~0ull;
    u8 *hist = malloc(hreq + 1), *local = NULL;
    u32 hlen = 0;
    u64 offset = 0;
    for (unsigned b = 0; b < count && !s.terminated && !s.error; b++) {
        if (!length[b]) continue;              /* zero-length block: no work (runtime.c:895-900) */
        if (!data[b]) { s.error = 1; break; }
        local = realloc(local, (size_t)hlen + length[b]);
        memcpy(local, hist, hlen);
        memcpy(local + hlen, data[b], length[b]);
        s.buf = local; s.len = hlen + length[b]; s.start = hlen; s.base = offset - hlen;
        hwlm_exec(&s, rose + rd32(rose + R_FMATCHER));
        offset += length[b];
        const u32 keep = offset < hreq ? (u32)offset : hreq;
        memmove(hist, local + s.len - keep, keep);
        hlen = keep;
    }
    free(hist); free(local); free(s.exhausted); free(s.dedupe);
    if (s.error) return HS_UNKNOWN_ERROR;
    return s.terminated ? HS_SCAN_TERMINATED : HS_SUCCESS;
}

/* Convenience for tests: collect (id, to) pairs; returns the hs_error_t, *n = pairs wanted (may exceed cap). */
struct collect { u32 *ids; u64 *tos; size_t cap, n, stop_after; };
static int collect_cb(unsigned id, unsigned long long from, unsigned long long to, unsigned flags, void *ctx) {
    (void)from; (void)flags;
    struct collect *c = (struct collect *)ctx;
    if (c->n < c->cap) { c->ids[c->n] = id; c->tos[c->n] = to; }
    c->n++;
    return c->stop_after && c->n >= c->stop_after;
}
int op_scan_collect(const void *db, const char *data, unsigned length, u32 *ids, u64 *tos, size_t cap, size_t stop_after,
                    size_t *n) {
    struct collect c = {ids, tos, cap, 0, stop_after};
    int rv = op_scan(db, data, length, collect_cb, &c);
    if (n) *n = c.n;
    return rv;
}

int op_scan_vector_collect(const void *db, const char *const *data, const unsigned *length, unsigned count, u32 *ids, u64 *tos,
                           size_t cap, size_t stop_after, size_t *n) {
    struct collect c = {ids, tos, cap, 0, stop_after};
    int rv = op_scan_vector(db, data, length, count, collect_cb, &c);
    if (n) *n = c.n;
    return rv;
}

/* Independent ground truth (no bytecode involved): every occurrence of every literal, `to` one past the
 * last byte, caseless = ASCII fold.  out_id/out_to receive up to cap pairs in (to, literal index) order. */
size_t op_brute_force(const char *const *lits, const size_t *lens, const unsigned *flags, const unsigned *ids, unsigned n,
                      const char *data, size_t length, u32 *out_id, u64 *out_to, size_t cap) {
    size_t m = 0;
    for (size_t e = 0; e < length; e++)
        for (unsigned k = 0; k < n; k++) {
            size_t L = lens[k];
            if (L > e + 1) continue;
            const u8 *t = (const u8 *)data + (e + 1 - L), *l = (const u8 *)lits[k];
            int nocase = flags && (flags[k] & 1), ok = 1;
            for (size_t i = 0; i < L && ok; i++) ok = nocase ? upc(t[i]) == upc(l[i]) : t[i] == l[i];
            if (ok) { if (m < cap) { out_id[m] = ids ? ids[k] : k; out_to[m] = e + 1; } m++; }
        }
    return m;
}
