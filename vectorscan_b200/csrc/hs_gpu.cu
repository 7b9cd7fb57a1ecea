// hs_gpu.cu -- libhs_gpu.so: B200 (sm_100a) block-mode scanning runtime for Vectorscan databases.
//
// Path restated (see DESIGN.md for the full map; citations are into /root/reference):
//   hs_scan gates / engine switch        src/runtime.c:316-475         -> host side of hs_gpu_scan*
//   hwlmExec -> fdrExec                  src/hwlm/hwlm.c:178-205, src/fdr/fdr.c:801-825
//   FDR shift-or filter                  src/fdr/fdr.c:144-296,663-764 -> k_scan<FdrNative>  (position-local form)
//   Teddy nibble filter                  src/fdr/teddy.c:920-1066      -> k_scan<TeddyNative>
//   noodle single literal                src/hwlm/noodle_engine.cpp:75-133 -> k_scan<Noodle>
//   bucket confirm                       src/fdr/fdr.c:298-333, src/fdr/fdr_confirm_runtime.h:43-102 -> confirm_entry()
//   roseCallback + roseRunProgram_l      src/rose/match.c:479-534, src/rose/program_runtime.c:3083-3504 -> run_program()
//   dedupe / report / exhaustion         src/report.h:54-153,300-340   -> match arena + finalize pass
//   scratch                              src/scratch.c:244-389         -> hs_gpu_scratch (device arenas, stream)
//
// Data-parallel formulation.  The reference walks one buffer sequentially carrying a 128-bit shift-or
// state.  Here the whole batch is one flat byte array in HBM (buffers back to back, offsets[] on the
// side) and every end position is decided independently: the FDR candidate bit for (end e, bucket b)
// depends only on bytes [e-7, e+1], so a warp takes a contiguous strip, each lane a 16-byte chunk per
// 512-byte row (one coalesced LDG.128 per lane), and the 7-byte state overhang travels between
// neighbouring lanes by shuffle instead of between loop iterations.  Buffer boundaries are ignored by
// the filter (it stays sound: it can only over-approximate) and enforced exactly by the confirm step.
// Candidates are compacted per warp into a shared-memory queue, confirmed one lane per candidate
// against the database's own FDRConfirm/LitInfo structures, the literal's Rose program is interpreted,
// and reports are appended to a device match arena (staged per warp in shared memory, one global
// atomic per flush).  A finalize pass sorts the arena by (buffer, to, id) and applies dedupe keys.
#include "hsb_layout.h"
#include "hsb_dbview.h"
#include "turbo_tables.h"
#include "limex_tables.h"
#include "hs_gpu.h"

#include <cuda_runtime.h>
#include <cub/cub.cuh>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

using namespace hsb;

#define HSG_VERSION "hs-gpu 0.1 (sm_100a) for Vectorscan 5.4.11 bytecode"

namespace {

// ------------------------------------------------------------------------------------------------
// geometry
// ------------------------------------------------------------------------------------------------
constexpr int WARPS = 8;
constexpr int THREADS = WARPS * 32;
constexpr int CHUNK = 16;                      // bytes per lane per row
constexpr int ROW_BYTES = 32 * CHUNK;          // 512: one coalesced warp load
constexpr int STRIP_ROWS = 32;
constexpr int STRIP_BYTES = STRIP_ROWS * ROW_BYTES;  // 16 KiB per warp strip
constexpr int QCAP = 1024;                     // candidate queue entries per warp
constexpr int STAGE_CAP = 96;                  // staged match records per warp
constexpr int CELL_BYTES = 1024;               // granularity of the position -> buffer index
constexpr size_t FRONT_PAD = 1024, BACK_PAD = 2048;  // zero bytes around the flat text in HBM

struct __align__(16) Rec {  // device match record
    u32 buf, to, id, aux;
};
constexpr u32 AUX_DEDUPE = 0x80000000u, AUX_EKEY = 0x40000000u, AUX_QUIET = 0x20000000u, AUX_KEYMASK = 0x1fffffffu;

struct ConfHdr {  // FDRConfirm header cached in shared memory
    u64 andmsk, mult, groups;
    u32 nbits, base;  // base: absolute bytecode offset of the FDRConfirm (0 = empty bucket)
};

struct DevDb {
    const u8 *bc;         // device copy of the bytecode (256-byte aligned)
    const u64 *fdr_tab;   // FDR reach table, folded to <= 2^14 rows (device global)
    const u32 *teddy_tab; // per byte value: byte j = lo_j[c&15] | hi_j[c>>4]
    u64 groups;
    u64 nood_msk, nood_cmp;
    u32 fdr_mask;         // rows - 1
    u32 fdr_stride;
    u32 conf_base;
    u32 n_buckets;
    u32 teddy_masks;
    u32 nood_len, nood_id;
    u32 min_width, min_width_excl, max_bi_width;
    u32 bc_len;
    // turbo front end (turbo_tables.h); tK == 0 when not built
    const u64 *pair_tab;
    const turbo::Slot *slots;
    const u32 *bloom;     // second filter tier (turbo_tables.h, 3.)
    u32 tK, tH1, tH2, tL, tBL;
    const u32 *gram_tab;  // first tier of k_turbo2 (turbo_tables.h, "gram table")
    u32 gG1, gmul, gwide;
    u32 lx_abs;           // absolute bytecode offset of the LimExNFA (single-outfix databases)
};

struct DevCorpus {
    const u8 *text;          // FRONT_PAD zero bytes precede, BACK_PAD follow
    const u64 *off;          // count + 1 offsets
    const u32 *cell_first;   // ncells + 1: buffer containing the first byte of each CELL_BYTES cell
    u64 total;
    u32 count;
    u32 nstrips;
    const u32 *order;        // engine kernels: work item k scans buffer order[k] (buffers by descending length, so that
                             // the lanes of a warp finish together); null = identity
    // stream mode of a DFA engine (k_mcclellan): buffer b is the next block of a stream that has come to DFA state
    // init_state[b] (device state ids); the state it leaves the block in goes to final_state[b]; no end-of-data check
    // (the stream is not over).  Both null outside stream scans.
    const u32 *init_state;
    u32 *final_state;
};

struct DevArena {
    Rec *recs;
    unsigned long long *count;  // records wanted (may exceed cap)
    unsigned long long *stats;  // [0] candidates, [1] confirm probes
    u32 *error;                 // sticky device-side error flag
    u64 cap;
    // ordered (segmented) mode: warp w of the grid owns recs[w * seg_cap, (w + 1) * seg_cap) and appends in
    // position order; seg_count[w] = records it wanted
    u32 *seg_count;
    u32 seg_cap;
    unsigned long long *op_counts;  // test instrumentation (option "count_ops"): executions per Rose opcode, else null
};

constexpr u32 ERR_BAD_OPCODE = 1, ERR_BAD_DEDUPE = 2, ERR_PROGRAM_LOOP = 4, ERR_COLLECTOR = 8, ERR_UNSORTED = 16;

struct Collector {  // per-lane, per-candidate match buffer of the ordered mode
    Rec r[4];
    u32 n;
};

struct WarpShared {
    u32 q[QCAP];
    Rec stage[STAGE_CAP];
    u32 stage_n;
    u32 pad[3];
};

// ------------------------------------------------------------------------------------------------
// small device helpers
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ u32 lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ u32 ld32(const u8 *bc, u32 off) { return *reinterpret_cast<const u32 *>(bc + off); }
__device__ __forceinline__ u64 ld64(const u8 *bc, u32 off) { return *reinterpret_cast<const u64 *>(bc + off); }

// 8 text bytes ending at global position g (last byte in the top byte), from 8-byte aligned loads
__device__ __forceinline__ u64 load_tail8(const u8 *text, u64 g) {
    const long long a = (long long)g - 7;           // may be negative: FRONT_PAD covers it
    const u8 *p = text + a;
    const uintptr_t ap = reinterpret_cast<uintptr_t>(p);
    const u64 *q = reinterpret_cast<const u64 *>(ap & ~(uintptr_t)7);
    const unsigned sh = (unsigned)(ap & 7) * 8;
    u64 lo = q[0];
    if (!sh) return lo;
    u64 hi = q[1];
    return (lo >> sh) | (hi << (64 - sh));
}

struct Ctx {
    const DevDb &db;
    const DevCorpus &cp;
    const DevArena &ar;
    u32 *q;         // per-warp candidate queue
    Rec *stage;     // per-warp staged match records
    u32 *stage_n;
    u32 stage_cap;
    const ConfHdr *conf;
    Collector *col;  // non-null: ordered mode, emit() collects instead of staging
};

__device__ __forceinline__ void emit(const Ctx &c, u32 buf, u32 to, u32 id, u32 aux) {
    Rec r{buf, to, id, aux};
    if (c.col) {
        if (c.col->n < 4) c.col->r[c.col->n] = r; else atomicOr(c.ar.error, ERR_COLLECTOR);
        c.col->n++;
        return;
    }
    u32 slot = atomicAdd(c.stage_n, 1u);
    if (slot < c.stage_cap) {
        c.stage[slot] = r;
    } else {  // staging full: slow path straight to the arena
        unsigned long long idx = atomicAdd(c.ar.count, 1ull);
        if (idx < c.ar.cap) c.ar.recs[idx] = r;
    }
}

// warp-synchronous: move staged records to the arena with one global atomic
__device__ __forceinline__ void flush_stage(const Ctx &c) {
    __syncwarp();
    u32 n = *(volatile u32 *)c.stage_n;
    if (n > c.stage_cap) n = c.stage_cap;
    if (n) {
        unsigned long long base = 0;
        if (lane_id() == 0) base = atomicAdd(c.ar.count, (unsigned long long)n);
        base = __shfl_sync(0xffffffffu, base, 0);
        for (u32 i = lane_id(); i < n; i += 32)
            if (base + i < c.ar.cap) c.ar.recs[base + i] = c.stage[i];
    }
    __syncwarp();
    if (lane_id() == 0) *c.stage_n = 0;
    __syncwarp();
}

// ------------------------------------------------------------------------------------------------
// Rose program interpreter (pure-literal subset of roseRunProgram_l).  `to` = end offset of the literal
// (one past its last byte) relative to the buffer [btext, btext + blen).
// Exhaustion cannot be evaluated in parallel; CHECK_EXHAUSTED always passes and reports carrying an
// ekey are resolved to "earliest offset per (buffer, ekey)" afterwards.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ u8 upcase(u8 c) { return (c >= 'a' && c <= 'z') ? (u8)(c - 0x20) : c; }

// 8 bytes starting at p (first byte lowest), from 8-byte aligned loads; may touch up to 7 bytes either side
__device__ __forceinline__ u64 load_u64_unaligned(const u8 *p) {
    const uintptr_t ap = reinterpret_cast<uintptr_t>(p);
    const u64 *q = reinterpret_cast<const u64 *>(ap & ~(uintptr_t)7);
    const unsigned sh = (unsigned)(ap & 7) * 8;
    const u64 lo = q[0];
    if (!sh) return lo;
    return (lo >> sh) | (q[1] << (64 - sh));
}

// clear bit 5 of every byte that is an ASCII lower-case letter (8 bytes at a time)
__device__ __forceinline__ u64 upcase8(u64 x) {
    const u64 k7f = 0x7f7f7f7f7f7f7f7full, low = x & k7f;
    // per byte, bit 7 set iff 'a' <= low byte <= 'z' and the original byte was below 0x80
    const u64 ge_a = low + (0x80 - 'a') * 0x0101010101010101ull;
    const u64 le_z = low + (0x7f - 'z') * 0x0101010101010101ull;
    const u64 is_lower = ge_a & ~le_z & ~x & 0x8080808080808080ull;
    return x ^ (is_lower >> 2);
}

// roseCheckMediumLiteral / roseCheckLongLiteral in block mode (src/rose/program_runtime.c:1954-2013): the
// lit_len bytes ending at `to` equal the literal (stored upper-cased when nocase).  Compared 8 bytes at a
// time; the bytecode copy is padded (upload_database) and the text has pads either side.
__device__ bool check_literal(const u8 *bc, u32 lit_off, u32 lit_len, bool nocase, const u8 *btext, u32 to) {
    if (to < lit_len) return false;  // "too short"
    const u8 *t = btext + (to - lit_len);
    const u8 *l = bc + lit_off;
    u64 diff = 0;
    for (u32 i = 0; i < lit_len; i += 8) {
        u64 a = load_u64_unaligned(t + i);
        const u64 b = load_u64_unaligned(l + i);
        if (nocase) a = upcase8(a);
        u64 d = a ^ b;
        const u32 left = lit_len - i;
        if (left < 8) d &= (1ull << (8 * left)) - 1;
        diff |= d;
    }
    return diff == 0;
}

// roseCheckMask (src/rose/program_runtime.c:644-727) in block mode: bytes outside the buffer are don't-care
__device__ bool check_mask8(u64 and_mask, u64 cmp_mask, u64 neg_mask, int check_off, const u8 *btext, u32 blen, u32 to) {
    if (check_off < 0 && (u32)(-check_off) > to) return false;
    const long long base = (long long)to + check_off;
    for (int k = 0; k < 8; k++) {
        long long o = base + k;
        if (o < 0 || o >= (long long)blen) continue;
        u8 a = (u8)(and_mask >> (8 * k)), cm = (u8)(cmp_mask >> (8 * k)), ng = (u8)(neg_mask >> (8 * k));
        u8 r = (u8)((btext[o] & a) ^ cm);
        if (ng ? (r == 0) : (r != 0)) return false;  // negated bytes must differ, plain bytes must be equal
    }
    return true;
}

__device__ bool check_byte(u8 and_mask, u8 cmp_mask, u8 negation, int check_off, const u8 *btext, u32 blen, u32 to) {
    if (check_off < 0 && (u32)(-check_off) > to) return false;
    const long long o = (long long)to + check_off;
    if (o < 0 || o >= (long long)blen) return true;
    const bool differ = (btext[o] & and_mask) != cmp_mask;
    return !(differ ^ (negation != 0));
}

// roseCheckMask32 / roseCheckMask64 (src/rose/program_runtime.c:729-935) in block mode: n window bytes starting at
// to + check_off; a byte outside the buffer is "don't care" (valid_data_mask), a byte inside must satisfy
// ((c & and) != cmp) == its neg_mask bit (validateMask32, src/rose/validate_mask.h:106-127).
__device__ bool check_mask_n(const u8 *andm, const u8 *cmpm, u64 neg, u32 n, int check_off, const u8 *btext, u32 blen, u32 to) {
    if (check_off < 0 && (u32)(-check_off) > to) return false;
    const long long base = (long long)to + check_off;
    for (u32 k = 0; k < n; k++) {
        const long long o = base + k;
        if (o < 0 || o >= (long long)blen) continue;
        const bool differ = (u8)(btext[o] & andm[k]) != cmpm[k];
        if (differ != (((neg >> k) & 1) != 0)) return false;
    }
    return true;
}

// roseCheckShufti16x8 ... 64x16 (src/rose/program_runtime.c:992-1250, validateShuftiMask* src/rose/validate_shufti.h:
// 46-372): window byte k is "in the class" when its nibble-mask product has a bit of bucket_select_mask[k]; every byte
// inside the buffer must have (not in class) == its neg_mask bit.  `ins` = the instruction; one routine for the six
// shapes (the 64-byte masks of the AVX512 shapes are four copies of a 16-byte mask, indexed per 16-byte lane as pshufb does).
__device__ bool check_shufti(const u8 *ins, u8 code, const u8 *btext, u32 blen, u32 to) {
    u32 n, wide;  // window bytes, 16 buckets?
    const u8 *hi1, *lo1, *hi2 = nullptr, *lo2 = nullptr, *sel_lo, *sel_hi = nullptr;
    u64 neg;
    int check_off;
    u32 lane_stride = 0;  // 16 for the 64-byte shapes: mask copy of the byte's 128-bit lane
    switch (code) {
    case op::CHECK_SHUFTI_16x8:
        n = 16; wide = 0; lo1 = ins + op::check_shufti_16x8::nib_mask; hi1 = lo1 + 16;
        sel_lo = ins + op::check_shufti_16x8::bucket_select_mask;
        neg = *reinterpret_cast<const u32 *>(ins + op::check_shufti_16x8::neg_mask);
        check_off = *reinterpret_cast<const int *>(ins + op::check_shufti_16x8::offset);
        break;
    case op::CHECK_SHUFTI_32x8:
        n = 32; wide = 0; hi1 = ins + op::check_shufti_32x8::hi_mask; lo1 = ins + op::check_shufti_32x8::lo_mask;
        sel_lo = ins + op::check_shufti_32x8::bucket_select_mask;
        neg = *reinterpret_cast<const u32 *>(ins + op::check_shufti_32x8::neg_mask);
        check_off = *reinterpret_cast<const int *>(ins + op::check_shufti_32x8::offset);
        break;
    case op::CHECK_SHUFTI_16x16:
        n = 16; wide = 1; hi1 = ins + op::check_shufti_16x16::hi_mask; hi2 = hi1 + 16;
        lo1 = ins + op::check_shufti_16x16::lo_mask; lo2 = lo1 + 16;
        sel_lo = ins + op::check_shufti_16x16::bucket_select_mask; sel_hi = sel_lo + 16;
        neg = *reinterpret_cast<const u32 *>(ins + op::check_shufti_16x16::neg_mask);
        check_off = *reinterpret_cast<const int *>(ins + op::check_shufti_16x16::offset);
        break;
    case op::CHECK_SHUFTI_32x16:
        n = 32; wide = 1; hi1 = ins + op::check_shufti_32x16::hi_mask; hi2 = hi1 + 16;
        lo1 = ins + op::check_shufti_32x16::lo_mask; lo2 = lo1 + 16;
        sel_hi = ins + op::check_shufti_32x16::bucket_select_mask_hi; sel_lo = ins + op::check_shufti_32x16::bucket_select_mask_lo;
        neg = *reinterpret_cast<const u32 *>(ins + op::check_shufti_32x16::neg_mask);
        check_off = *reinterpret_cast<const int *>(ins + op::check_shufti_32x16::offset);
        break;
    case op::CHECK_SHUFTI_64x8:
        n = 64; wide = 0; lane_stride = 16; hi1 = ins + op::check_shufti_64x8::hi_mask; lo1 = ins + op::check_shufti_64x8::lo_mask;
        sel_lo = ins + op::check_shufti_64x8::bucket_select_mask;
        neg = *reinterpret_cast<const u64 *>(ins + op::check_shufti_64x8::neg_mask);
        check_off = *reinterpret_cast<const int *>(ins + op::check_shufti_64x8::offset);
        break;
    default:  // CHECK_SHUFTI_64x16
        n = 64; wide = 1; lane_stride = 16; hi1 = ins + op::check_shufti_64x16::hi_mask_1; hi2 = ins + op::check_shufti_64x16::hi_mask_2;
        lo1 = ins + op::check_shufti_64x16::lo_mask_1; lo2 = ins + op::check_shufti_64x16::lo_mask_2;
        sel_hi = ins + op::check_shufti_64x16::bucket_select_mask_hi; sel_lo = ins + op::check_shufti_64x16::bucket_select_mask_lo;
        neg = *reinterpret_cast<const u64 *>(ins + op::check_shufti_64x16::neg_mask);
        check_off = *reinterpret_cast<const int *>(ins + op::check_shufti_64x16::offset);
        break;
    }
    if (check_off < 0 && (u32)(-check_off) > to) return false;
    const long long base = (long long)to + check_off;
    for (u32 k = 0; k < n; k++) {
        const long long o = base + k;
        if (o < 0 || o >= (long long)blen) continue;
        const u8 ch = btext[o];
        const u32 ln = lane_stride ? (k & ~15u) : 0u;
        u8 hit = (u8)(lo1[ln + (ch & 15)] & hi1[ln + (ch >> 4)] & sel_lo[k]);
        if (wide) hit |= (u8)(lo2[ln + (ch & 15)] & hi2[ln + (ch >> 4)] & sel_hi[k]);
        if ((hit == 0) != (((neg >> k) & 1) != 0)) return false;
    }
    return true;
}

__device__ void run_program(const Ctx &c, u32 prog, u32 buf, const u8 *btext, u32 blen, u32 to) {
    const u8 *bc = c.db.bc;
    u32 pc = prog;
    for (int guard = 0; guard < 4096; guard++) {
        const u8 code = bc[pc];
        if (c.ar.op_counts) atomicAdd(&c.ar.op_counts[code & 127], 1ull);
        switch (code) {
        case op::END:
            return;
        case op::CHECK_GROUPS:
            if (!(ld64(bc, pc + op::check_groups::groups) & c.db.groups)) return;
            pc += op::check_groups::SIZE;
            break;
        case op::CHECK_MASK:
            if (!check_mask8(ld64(bc, pc + op::check_mask::and_mask), ld64(bc, pc + op::check_mask::cmp_mask),
                             ld64(bc, pc + op::check_mask::neg_mask), (int)ld32(bc, pc + op::check_mask::offset),
                             btext, blen, to)) {
                pc += ld32(bc, pc + op::check_mask::fail_jump);
            } else {
                pc += op::check_mask::SIZE;
            }
            break;
        case op::CHECK_BYTE:
            if (!check_byte(bc[pc + op::check_byte::and_mask], bc[pc + op::check_byte::cmp_mask],
                            bc[pc + op::check_byte::negation], (int)ld32(bc, pc + op::check_byte::offset), btext,
                            blen, to)) {
                pc += ld32(bc, pc + op::check_byte::fail_jump);
            } else {
                pc += op::check_byte::SIZE;
            }
            break;
        case op::CHECK_MASK_32:
            if (!check_mask_n(bc + pc + op::check_mask32::and_mask, bc + pc + op::check_mask32::cmp_mask,
                              ld32(bc, pc + op::check_mask32::neg_mask), 32, (int)ld32(bc, pc + op::check_mask32::offset),
                              btext, blen, to)) {
                pc += ld32(bc, pc + op::check_mask32::fail_jump);
            } else {
                pc += op::check_mask32::SIZE;
            }
            break;
        case op::CHECK_MASK_64:
            if (!check_mask_n(bc + pc + op::check_mask64::and_mask, bc + pc + op::check_mask64::cmp_mask,
                              ld64(bc, pc + op::check_mask64::neg_mask), 64, (int)ld32(bc, pc + op::check_mask64::offset),
                              btext, blen, to)) {
                pc += ld32(bc, pc + op::check_mask64::fail_jump);
            } else {
                pc += op::check_mask64::SIZE;
            }
            break;
        case op::CHECK_SHUFTI_16x8: case op::CHECK_SHUFTI_32x8: case op::CHECK_SHUFTI_16x16:
        case op::CHECK_SHUFTI_32x16: case op::CHECK_SHUFTI_64x8: case op::CHECK_SHUFTI_64x16: {
            const u32 fj = code == op::CHECK_SHUFTI_16x8 ? op::check_shufti_16x8::fail_jump
                         : code == op::CHECK_SHUFTI_32x8 ? op::check_shufti_32x8::fail_jump
                         : code == op::CHECK_SHUFTI_16x16 ? op::check_shufti_16x16::fail_jump
                         : code == op::CHECK_SHUFTI_32x16 ? op::check_shufti_32x16::fail_jump
                         : code == op::CHECK_SHUFTI_64x8 ? op::check_shufti_64x8::fail_jump : op::check_shufti_64x16::fail_jump;
            // every shape's fail_jump is its last operand: the instruction ends 4 bytes later (rounded to 8)
            if (!check_shufti(bc + pc, code, btext, blen, to)) pc += ld32(bc, pc + fj);
            else pc += (fj + 4 + 7) & ~7u;
            break;
        }
        case op::CATCH_UP:
        case op::CLEAR_WORK_DONE:
            pc += op::SIZE_1;
            break;
        case op::CHECK_LONG_LIT:
        case op::CHECK_MED_LIT:
        case op::CHECK_LONG_LIT_NOCASE:
        case op::CHECK_MED_LIT_NOCASE: {
            const bool nocase = code == op::CHECK_LONG_LIT_NOCASE || code == op::CHECK_MED_LIT_NOCASE;
            if (!check_literal(bc, ld32(bc, pc + op::check_lit::lit_offset), ld32(bc, pc + op::check_lit::lit_length),
                               nocase, btext, to)) {
                pc += ld32(bc, pc + op::check_lit::fail_jump);
            } else {
                pc += op::check_lit::SIZE;
            }
            break;
        }
        case op::CHECK_EXHAUSTED:
            pc += op::check_exhausted::SIZE;  // resolved after the scan
            break;
        case op::SQUASH_GROUPS:
            pc += op::squash_groups::SIZE;  // result-neutral once exhaustion is exact
            break;
        case op::INCLUDED_JUMP:
            pc += op::included_jump::SIZE;  // every candidate bit is confirmed independently: no squash, no jump
            break;
        case op::DEDUPE: {
            // the dedupe key rides on the following report; a DEDUPE is always followed by REPORT_EXHAUST
            // or a quiet SET_EXHAUST in pure-literal programs (makeReport), handled by tagging below
            const u32 dkey = ld32(bc, pc + op::dedupe::dkey);
            pc += op::dedupe::SIZE;
            const u8 nxt = bc[pc];
            if (nxt == op::REPORT_EXHAUST) {
                const u32 id = ld32(bc, pc + op::report_exhaust::onmatch);
                const int adj = (int)ld32(bc, pc + op::report_exhaust::offset_adjust);
                const u32 ekey = ld32(bc, pc + op::report_exhaust::ekey);
                emit(c, buf, (u32)((int)to + adj), id, AUX_DEDUPE | AUX_EKEY | (ekey & AUX_KEYMASK));
                pc += op::report_exhaust::SIZE;
            } else if (nxt == op::REPORT) {
                emit(c, buf, (u32)((int)to + (int)ld32(bc, pc + op::report::offset_adjust)),
                     ld32(bc, pc + op::report::onmatch), AUX_DEDUPE);
                pc += op::report::SIZE;
            } else if (nxt == op::SET_EXHAUST) {
                emit(c, buf, to, 0, AUX_QUIET | AUX_EKEY | (ld32(bc, pc + op::set_exhaust::ekey) & AUX_KEYMASK));
                pc += op::set_exhaust::SIZE;
            } else {
                (void)dkey;
                atomicOr(c.ar.error, 2u);
                return;
            }
            break;
        }
        case op::REPORT:
            if (c.col && ld32(bc, pc + op::report::offset_adjust)) atomicOr(c.ar.error, ERR_UNSORTED);
            emit(c, buf, (u32)((int)to + (int)ld32(bc, pc + op::report::offset_adjust)),
                 ld32(bc, pc + op::report::onmatch), 0);
            pc += op::report::SIZE;
            break;
        case op::FINAL_REPORT:
            emit(c, buf, (u32)((int)to + (int)ld32(bc, pc + op::report::offset_adjust)),
                 ld32(bc, pc + op::report::onmatch), 0);
            return;
        case op::REPORT_EXHAUST:
            emit(c, buf, (u32)((int)to + (int)ld32(bc, pc + op::report_exhaust::offset_adjust)),
                 ld32(bc, pc + op::report_exhaust::onmatch),
                 AUX_EKEY | (ld32(bc, pc + op::report_exhaust::ekey) & AUX_KEYMASK));
            pc += op::report_exhaust::SIZE;
            break;
        case op::DEDUPE_AND_REPORT:
            emit(c, buf, (u32)((int)to + (int)ld32(bc, pc + op::dedupe_and_report::offset_adjust)),
                 ld32(bc, pc + op::dedupe_and_report::onmatch), AUX_DEDUPE);
            pc += op::dedupe_and_report::SIZE;
            break;
        case op::SET_EXHAUST:
            emit(c, buf, to, 0, AUX_QUIET | AUX_EKEY | (ld32(bc, pc + op::set_exhaust::ekey) & AUX_KEYMASK));
            pc += op::set_exhaust::SIZE;
            break;
        default:
            atomicOr(c.ar.error, 1u);  // cannot happen: the census rejected such databases
            return;
        }
    }
    atomicOr(c.ar.error, 4u);
}

// ------------------------------------------------------------------------------------------------
// buffer lookup + bucket confirm (do_confirm_fdr + confWithBit)
// ------------------------------------------------------------------------------------------------
struct BufRef {
    u32 idx;
    u64 start;
    u32 len;
};

__device__ __forceinline__ BufRef find_buffer(const DevCorpus &cp, u32 /*strip*/, u64 g) {
    const u32 cell = (u32)(g / CELL_BYTES);
    u32 lo = cp.cell_first[cell], hi = cp.cell_first[cell + 1];
    if (hi - lo <= 1) {  // common case (buffers of a cell's size or more): three independent loads, no search
        const u64 o0 = cp.off[lo], o1 = cp.off[lo + 1], o2 = cp.off[hi + 1];
        if (hi != lo && o1 <= g) return BufRef{hi, o1, (u32)(o2 - o1)};
        return BufRef{lo, o0, (u32)(o1 - o0)};
    }
    // largest b in [lo, hi] with off[b] <= g
    while (lo < hi) {
        u32 mid = (lo + hi + 1) >> 1;
        if (cp.off[mid] <= g) lo = mid; else hi = mid - 1;
    }
    const u64 s = cp.off[lo], e = cp.off[lo + 1];
    return BufRef{lo, s, (u32)(e - s)};
}

__device__ __forceinline__ bool passes_gates(const DevDb &db, u32 blen) {
    if (blen < db.min_width || blen < db.min_width_excl) return false;                 // src/runtime.c:346-350,385-389
    if (db.max_bi_width != rose::BOUND_INF && blen > db.max_bi_width) return false;     // :393-398
    return true;
}

// One candidate: global end position g, bitmask of buckets to confirm.
__device__ void confirm_entry(const Ctx &c, u32 strip, u64 g, u32 bucket_mask, u32 &n_probe) {
    if (g >= c.cp.total) return;
    const BufRef b = find_buffer(c.cp, strip, g);
    if (!passes_gates(c.db, b.len)) return;
    const u32 i = (u32)(g - b.start);  // index of the candidate's last byte inside its buffer
    u64 v = load_tail8(c.cp.text, g);
    if (i < 7) v &= ~0ull << (8 * (7 - i));  // bytes before the buffer start read as zero (block mode has no history)
    const u8 *btext = c.cp.text + b.start;
    const u8 *bc = c.db.bc;
    if (c.db.n_buckets == 16) bucket_mask |= bucket_mask << 8;  // fat Teddy: filter bit b stands for buckets b and b + 8
    while (bucket_mask) {
        const u32 bk = __ffs(bucket_mask) - 1;
        bucket_mask &= bucket_mask - 1;
        const ConfHdr &h = c.conf[bk];
        if (!h.base || !(h.groups & c.db.groups)) continue;
        n_probe++;
        const u32 hash = (u32)(((v & h.andmsk) * h.mult) >> (64 - h.nbits));
        const u32 start = ld32(bc, h.base + (u32)conf::SIZE + 4 * hash);
        if (!start) continue;
        u32 li = h.base + start;
        for (;;) {
            const u64 lv = ld64(bc, li + lit::v), lm = ld64(bc, li + lit::msk);
            const u64 tail = ld64(bc, li + lit::id);  // id | size << 32 | flags << 40 | next << 48
            const u32 size = (u32)(tail >> 32) & 0xff;
            if ((v & lm) == lv && size <= i + 1 && (ld64(bc, li + lit::groups) & c.db.groups))
                run_program(c, (u32)tail, b.idx, btext, b.len, i + 1);
            if (!((tail >> 48) & 0xff)) break;
            li += (u32)lit::SIZE;
        }
    }
}

// noodle has no confirm tables: its queue entries run the literal's program directly
__device__ void confirm_entry_noodle(const Ctx &c, u32 strip, u64 g) {
    if (g >= c.cp.total) return;
    const BufRef b = find_buffer(c.cp, strip, g);
    if (!passes_gates(c.db, b.len)) return;
    const u32 i = (u32)(g - b.start);
    if (i + 1 < c.db.nood_len) return;  // literal would start before the buffer
    run_program(c, c.db.nood_id, b.idx, c.cp.text + b.start, b.len, i + 1);
}

// warp-synchronous drain of the candidate queue: one lane per entry
template <bool NOODLE>
__device__ void drain_queue(const Ctx &c, u32 strip, u64 strip_base, u32 count, u32 &n_probe) {
    for (u32 base = 0; base < count; base += 32) {
        const u32 idx = base + lane_id();
        if (idx < count) {
            const u32 e = c.q[idx];
            if (NOODLE) confirm_entry_noodle(c, strip, strip_base + (e >> 8));
            else confirm_entry(c, strip, strip_base + (e >> 8), e & 0xff, n_probe);
        }
        __syncwarp();
        if (*(volatile u32 *)c.stage_n >= c.stage_cap / 2) flush_stage(c);
    }
}

// Push this row's candidate bytes (cw[k] covers chunk bytes 4k..4k+3, one bucket bit per bit) into
// the warp queue.  Warp-synchronous; returns the new queue length.
__device__ __forceinline__ u32 push_candidates(const Ctx &c, const u32 cw[4], u32 pos_in_strip, u32 qn) {
    u32 nz = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        u32 w = cw[k];
        nz |= ((w & 0xffu) ? 1u : 0u) << (4 * k);
        nz |= ((w & 0xff00u) ? 2u : 0u) << (4 * k);
        nz |= ((w & 0xff0000u) ? 4u : 0u) << (4 * k);
        nz |= ((w & 0xff000000u) ? 8u : 0u) << (4 * k);
    }
    const u32 n = __popc(nz);
    u32 incl = n;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        u32 t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane_id() >= (u32)d) incl += t;
    }
    u32 slot = qn + incl - n;
    while (nz) {
        const u32 k = __ffs(nz) - 1;
        nz &= nz - 1;
        const u32 byte = (cw[k >> 2] >> (8 * (k & 3))) & 0xff;
        c.q[slot++] = ((pos_in_strip + k) << 8) | byte;
    }
    return qn + __shfl_sync(0xffffffffu, incl, 31);
}

// ------------------------------------------------------------------------------------------------
// front ends.  Each provides:  smem_bytes(db), setup(smem, db) [CTA-wide], and
//   row(chunk words w[0..3], following word nxt, trailing words of the previous chunk, carry) -> cw[4]
// ------------------------------------------------------------------------------------------------

// ---- FDR, database's own reach table (position-local restatement of get_conf_stride_N) -----------
struct FdrNative {
    static __host__ size_t smem_bytes(const DevDb &db) { return ((size_t)db.fdr_mask + 1) * 8; }

    const u64 *tab;
    u32 mask;
    u32 stride;
    u32 carry4, carry5;  // state overhang of lane 31's previous chunk

    __device__ void setup(u8 *smem, const DevDb &db) {
        u64 *t = reinterpret_cast<u64 *>(smem);
        const uint4 *src = reinterpret_cast<const uint4 *>(db.fdr_tab);
        uint4 *dst = reinterpret_cast<uint4 *>(t);
        const u32 n16 = (db.fdr_mask + 1) / 2;
        for (u32 i = threadIdx.x; i < n16; i += blockDim.x) dst[i] = src[i];
        tab = t;
        mask = db.fdr_mask;
        stride = db.fdr_stride;
        carry4 = carry5 = 0;
    }
    __device__ void reset_carry() { carry4 = carry5 = 0; }

    // S = OR over probed x of row(x) << 8x for the 16 positions of this chunk (184 bits in 6 words)
    template <int STRIDE>
    __device__ __forceinline__ void probe(const u32 w[5], u32 S[6]) const {
        u32 A[4][5];
#pragma unroll
        for (int s = 0; s < 4; s++)
#pragma unroll
            for (int k = 0; k < 5; k++) A[s][k] = 0;
#pragma unroll
        for (int x = 0; x < 16; x += STRIDE) {
            const int wi = x >> 2, sh = x & 3;
            const u32 key = __funnelshift_r(w[wi], w[wi + 1], 8 * sh) & mask;
            const u64 r = tab[key];
            A[sh][wi] |= (u32)r;
            A[sh][wi + 1] |= (u32)(r >> 32);
        }
#pragma unroll
        for (int k = 0; k < 6; k++) {
            u32 v = k < 5 ? A[0][k] : 0;
#pragma unroll
            for (int s = STRIDE; s < 4; s += STRIDE) {
                const u32 lo = k >= 1 ? A[s][k - 1] : 0, hi = k < 5 ? A[s][k] : 0;
                v |= __funnelshift_l(lo, hi, 8 * s);
            }
            S[k] = v;
        }
    }

    __device__ __forceinline__ void row(const u32 w[5], u32 cw[4], bool warmup) {
        u32 S[6];
        if (stride == 1) probe<1>(w, S);
        else if (stride == 2) probe<2>(w, S);
        else probe<4>(w, S);
        u32 in4 = __shfl_up_sync(0xffffffffu, S[4], 1), in5 = __shfl_up_sync(0xffffffffu, S[5], 1);
        if (lane_id() == 0) { in4 = carry4; in5 = carry5; }
        carry4 = __shfl_sync(0xffffffffu, S[4], 31);
        carry5 = __shfl_sync(0xffffffffu, S[5], 31);
        if (warmup) return;
        cw[0] = ~(S[0] | in4);
        cw[1] = ~(S[1] | in5);
        cw[2] = ~S[2];
        cw[3] = ~S[3];
    }
};

// ---- Teddy, database's own nibble masks folded into one u32 per byte value -----------------------
// T[c] byte j = lo_j[c & 15] | hi_j[c >> 4]  (bit b set = bucket b cannot have c as its (j+1)-th byte from
// the end); candidate byte for end i = ~(T[c_i].b0 | T[c_{i-1}].b1 | T[c_{i-2}].b2 | T[c_{i-3}].b3).
// The table is replicated once per bank (word index c * 32 + lane) so lookups never conflict.
struct TeddyNative {
    static __host__ size_t smem_bytes(const DevDb &) { return 256 * 32 * 4; }
    const u32 *tab;
    u32 c1, c2, c3;  // T of the last three bytes of lane 31's previous chunk

    __device__ void setup(u8 *smem, const DevDb &db) {
        u32 *t = reinterpret_cast<u32 *>(smem);
        for (u32 i = threadIdx.x; i < 256 * 32; i += blockDim.x) t[i] = db.teddy_tab[i >> 5];
        tab = t;
        c1 = c2 = c3 = 0;
    }
    __device__ void reset_carry() { c1 = c2 = c3 = 0; }

    __device__ __forceinline__ void row(const u32 w[5], u32 cw[4], bool warmup) {
        u32 t[19];  // t[3 + i] = T[c_i]; t[0..2] = previous three bytes
        const u32 ln = lane_id();
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const u32 ch = (w[i >> 2] >> (8 * (i & 3))) & 0xff;
            t[3 + i] = tab[ch * 32 + ln];
        }
        u32 p1 = __shfl_up_sync(0xffffffffu, t[18], 1), p2 = __shfl_up_sync(0xffffffffu, t[17], 1),
            p3 = __shfl_up_sync(0xffffffffu, t[16], 1);
        if (ln == 0) { p1 = c1; p2 = c2; p3 = c3; }
        c1 = __shfl_sync(0xffffffffu, t[18], 31);
        c2 = __shfl_sync(0xffffffffu, t[17], 31);
        c3 = __shfl_sync(0xffffffffu, t[16], 31);
        if (warmup) return;
        t[2] = p1; t[1] = p2; t[0] = p3;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            u32 word = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int i = 4 * k + j;
                const u32 m = (t[3 + i] & 0xff) | ((t[2 + i] >> 8) & 0xff) | ((t[1 + i] >> 16) & 0xff) | (t[i] >> 24);
                word |= (~m & 0xff) << (8 * j);
            }
            cw[k] = word;
        }
    }
};

// ---- noodle: one literal of <= 8 bytes, compared directly (msk/cmp of the noodTable) ---------------
struct Noodle {
    static __host__ size_t smem_bytes(const DevDb &) { return 16; }
    u64 msk, cmp;
    u32 len;
    u32 cw2, cw3;  // last 8 bytes of lane 31's previous chunk

    __device__ void setup(u8 *, const DevDb &db) {
        // noodTable msk/cmp hold the literal's first byte in the low byte; align so its LAST byte is the top byte
        const u32 sh = 8 * (8 - db.nood_len);
        msk = db.nood_msk << sh;
        cmp = db.nood_cmp << sh;
        len = db.nood_len;
        cw2 = cw3 = 0;
    }
    __device__ void reset_carry() { cw2 = cw3 = 0; }

    __device__ __forceinline__ void row(const u32 w[5], u32 cw[4], bool warmup) {
        u32 p2 = __shfl_up_sync(0xffffffffu, w[2], 1), p3 = __shfl_up_sync(0xffffffffu, w[3], 1);
        if (lane_id() == 0) { p2 = cw2; p3 = cw3; }
        cw2 = __shfl_sync(0xffffffffu, w[2], 31);
        cw3 = __shfl_sync(0xffffffffu, w[3], 31);
        if (warmup) return;
        const u32 e[6] = {p2, p3, w[0], w[1], w[2], w[3]};  // 24 bytes: 8 before the chunk + the chunk
#pragma unroll
        for (int k = 0; k < 4; k++) {
            u32 word = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int i = 4 * k + j;            // end position inside the chunk
                const int first = i + 1;            // byte index in e[] of the window's first byte (8 + i - 7)
                const int wi = first >> 2, sh = (first & 3) * 8;
                const u32 lo = __funnelshift_r(e[wi], e[wi + 1], sh);
                const u32 hi = __funnelshift_r(e[wi + 1], wi + 2 < 6 ? e[wi + 2] : 0, sh);
                const u64 v = ((u64)hi << 32) | lo;
                word |= (((v & msk) == cmp) ? 1u : 0u) << (8 * j);  // single pseudo-bucket 0
            }
            cw[k] = word;
        }
    }
};

// ------------------------------------------------------------------------------------------------
// the scan kernel: persistent CTAs, one strip per warp at a time
// ------------------------------------------------------------------------------------------------
template <class FE, bool NOODLE>
__global__ void __launch_bounds__(THREADS, 1) k_scan(const DevDb db, const DevCorpus cp, const DevArena ar) {
    extern __shared__ __align__(16) u8 smem[];
    __shared__ WarpShared wsh[WARPS];
    __shared__ ConfHdr conf[16];

    FE fe;
    fe.setup(smem, db);
    if (threadIdx.x < 16) {
        ConfHdr h{0, 0, 0, 0, 0};
        if (!NOODLE && threadIdx.x < db.n_buckets) {
            const u32 cf = ld32(db.bc, db.conf_base + 4 * threadIdx.x);
            if (cf) {
                h.base = db.conf_base + cf;
                h.andmsk = ld64(db.bc, h.base + (u32)conf::andmsk);
                h.mult = ld64(db.bc, h.base + (u32)conf::mult);
                h.nbits = ld32(db.bc, h.base + (u32)conf::nBits);
                h.groups = ld64(db.bc, h.base + (u32)conf::groups);
            }
        }
        conf[threadIdx.x] = h;
    }
    const u32 warp = threadIdx.x >> 5, ln = lane_id();
    if (ln == 0) wsh[warp].stage_n = 0;
    __syncthreads();

    const Ctx c{db, cp, ar, wsh[warp].q, wsh[warp].stage, &wsh[warp].stage_n, STAGE_CAP, conf, nullptr};
    u32 n_cand = 0, n_probe = 0;
    const u32 warps_total = gridDim.x * WARPS;
    for (u32 strip = blockIdx.x * WARPS + warp; strip < cp.nstrips; strip += warps_total) {
        const u64 strip_base = (u64)strip * STRIP_BYTES;
        u32 qn = 0;
        fe.reset_carry();
        for (int row = strip_base ? -1 : 0; row < STRIP_ROWS; row++) {
            const long long g0 = (long long)strip_base + (long long)row * ROW_BYTES + ln * CHUNK;
            if (row >= 0 && (u64)((long long)strip_base + (long long)row * ROW_BYTES) >= cp.total) break;
            const uint4 v = *reinterpret_cast<const uint4 *>(cp.text + g0);
            u32 w[5] = {v.x, v.y, v.z, v.w, 0};
            w[4] = __shfl_down_sync(0xffffffffu, v.x, 1);
            if (ln == 31) w[4] = *reinterpret_cast<const u32 *>(cp.text + g0 + CHUNK);
            u32 cw[4] = {0, 0, 0, 0};
            fe.row(w, cw, row < 0);
            if (row < 0) continue;
            // clip end positions beyond the corpus
            const long long left = (long long)cp.total - g0;
            if (left < CHUNK) {
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const long long lw = left - 4 * k;
                    if (lw <= 0) cw[k] = 0;
                    else if (lw < 4) cw[k] &= (1u << (8 * (int)lw)) - 1;
                }
            }
            if (__any_sync(0xffffffffu, (cw[0] | cw[1] | cw[2] | cw[3]) != 0)) {
                const u32 before = qn;
                qn = push_candidates(c, cw, (u32)row * ROW_BYTES + ln * CHUNK, qn);
                n_cand += qn - before;
                __syncwarp();
                if (qn > QCAP - ROW_BYTES) {
                    drain_queue<NOODLE>(c, strip, strip_base, qn, n_probe);
                    qn = 0;
                }
            }
        }
        if (qn) drain_queue<NOODLE>(c, strip, strip_base, qn, n_probe);
    }
    flush_stage(c);
    // per-warp statistics: n_cand is warp-uniform, n_probe is per lane
#pragma unroll
    for (int d = 16; d; d >>= 1) n_probe += __shfl_down_sync(0xffffffffu, n_probe, d);
    if (ln == 0) {
        if (n_cand) atomicAdd(&ar.stats[0], (unsigned long long)n_cand);
        if (n_probe) atomicAdd(&ar.stats[1], (unsigned long long)n_probe);
    }
}

#include "turbo.cuh"
#include "turbo2.cuh"
#include "limex.cuh"
#include "mcclellan.cuh"
#include "dfapack.cuh"

// ------------------------------------------------------------------------------------------------
// finalize: Rec (sorted, deduped) -> hs_gpu_match_t
// ------------------------------------------------------------------------------------------------
struct RecLess {
    __host__ __device__ bool operator()(const Rec &a, const Rec &b) const {
        if (a.buf != b.buf) return a.buf < b.buf;
        if (a.to != b.to) return a.to < b.to;
        if (a.id != b.id) return a.id < b.id;
        return a.aux < b.aux;
    }
};
struct RecDup {  // equal = second is a dedupe-key duplicate of the first
    __host__ __device__ bool operator()(const Rec &a, const Rec &b) const {
        return (a.aux & b.aux & AUX_DEDUPE) && a.buf == b.buf && a.to == b.to && a.id == b.id &&
               ((a.aux ^ b.aux) & AUX_QUIET) == 0;
    }
};

// Small batches of records (the usual case for regex databases: a few matches per MiB): sort, dedupe, compact and
// convert in ONE single-CTA kernel -- a block merge sort with the records in registers -- instead of the device-wide
// merge sort + select chain (five launches and a host round trip for the count: ~55 us for a thousand records).
constexpr u32 SMALL_FIN_MAX = 4096, SMALL_FIN_ITEMS = 4;
template <int THREADS>
__global__ void __launch_bounds__(THREADS) k_small_finalize(const Rec *in, Rec *out_recs, hs_gpu_match_t *out,
                                                             unsigned long long *n_out,
                                                             const unsigned long long *n_raw, u64 arena_cap) {
    // launched right behind the scan, before the host knows the record count: do nothing when it is outside this
    // kernel's envelope (the host then runs the general finalize) or when the arena overflowed (the scan is repeated)
    const unsigned long long raw = *n_raw;
    if (raw == 0 || raw > (unsigned long long)(THREADS * SMALL_FIN_ITEMS) || raw > arena_cap) {
        if (threadIdx.x == 0) *n_out = raw ? ~0ull : 0ull;
        return;
    }
    const u32 n = (u32)raw;
    using Sort = cub::BlockMergeSort<Rec, THREADS, SMALL_FIN_ITEMS>;
    using Scan = cub::BlockScan<u32, THREADS>;
    extern __shared__ __align__(16) u8 fin_smem[];
    typename Sort::TempStorage &sort_tmp = *reinterpret_cast<typename Sort::TempStorage *>(fin_smem);
    __shared__ typename Scan::TempStorage scan_tmp;
    __shared__ Rec last_of[THREADS];
    const u32 tid = threadIdx.x;
    Rec it[SMALL_FIN_ITEMS];
#pragma unroll
    for (u32 q = 0; q < SMALL_FIN_ITEMS; q++) {
        const u32 i = tid * SMALL_FIN_ITEMS + q;
        it[q] = i < n ? in[i] : Rec{0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};  // sorts behind every record
    }
    Sort(sort_tmp).Sort(it, RecLess());   // blocked arrangement: thread t holds sorted positions 4t .. 4t+3
    last_of[tid] = it[SMALL_FIN_ITEMS - 1];
    __syncthreads();
    u32 keep = 0, cnt = 0;
#pragma unroll
    for (u32 q = 0; q < SMALL_FIN_ITEMS; q++) {
        const u32 i = tid * SMALL_FIN_ITEMS + q;
        bool dup = false;
        if (q) dup = RecDup()(it[q - 1], it[q]);
        else if (tid) dup = RecDup()(last_of[tid - 1], it[0]);
        if (i < n && !dup) { keep |= 1u << q; cnt++; }
    }
    u32 base = 0, total = 0;
    Scan(scan_tmp).ExclusiveSum(cnt, base, total);
#pragma unroll
    for (u32 q = 0; q < SMALL_FIN_ITEMS; q++)
        if ((keep >> q) & 1u) {
            out_recs[base] = it[q];
            out[base] = hs_gpu_match_t{it[q].buf, it[q].id, (uint64_t)it[q].to};
            base++;
        }
    if (tid == 0) *n_out = total;
}


// keep[i] = 0 iff record i repeats record i-1 under a dedupe key (dedupeCatchup, src/report.h:83-93)
__global__ void k_flag_dups(const Rec *in, u8 *keep, u64 n) {
    const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) keep[i] = (i && RecDup()(in[i - 1], in[i])) ? 0 : 1;
}

// ---- single-match (exhaustion key) resolution on the device ------------------------------------------------------
// The records are sorted by (buffer, to, id): the report an exhaustible pattern keeps is the FIRST record of its
// (buffer, ekey) pair; quiet records (SET_EXHAUST) exhaust the key and are dropped.  Pass 1 files every keyed record
// under its pair in an open-addressing table and keeps the smallest record index (atomicMin); pass 2 keeps a keyed
// record only when it is that smallest index.  What roseReportExhaust / CHECK_EXHAUSTED do one match at a time
// (src/report.h:121-153, program_runtime.c:2665-2680), restated for records that are all known at once.
__device__ __forceinline__ u32 ekey_slot(unsigned long long key, u32 mask) {
    return (u32)((key * 0x9E3779B97F4A7C15ull) >> 32) & mask;
}

__global__ void k_ekey_insert(const Rec *recs, u64 n, unsigned long long *keys, u32 *first, u32 mask) {
    const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Rec r = recs[i];
    if (!(r.aux & AUX_EKEY)) return;
    const unsigned long long key = ((unsigned long long)r.buf << 32) | (r.aux & AUX_KEYMASK);
    for (u32 h = ekey_slot(key, mask);; h = (h + 1) & mask) {
        const unsigned long long prev = atomicCAS(&keys[h], ~0ull, key);
        if (prev == ~0ull || prev == key) { atomicMin(&first[h], (u32)i); return; }
    }
}

__global__ void k_ekey_filter(const Rec *recs, u64 n, const unsigned long long *keys, const u32 *first, u32 mask, u8 *keep) {
    const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Rec r = recs[i];
    bool k = !(r.aux & AUX_QUIET);
    if (r.aux & AUX_EKEY) {
        const unsigned long long key = ((unsigned long long)r.buf << 32) | (r.aux & AUX_KEYMASK);
        u32 h = ekey_slot(key, mask);
        while (keys[h] != key) h = (h + 1) & mask;
        k = k && first[h] == (u32)i;
    }
    keep[i] = k ? 1 : 0;
}

// ---- buffers by descending length (16-byte classes), for the engine kernels: one thread steps one buffer, so a warp
// runs as long as its longest buffer; handing out buffers of similar length to neighbouring lanes removes the idle lanes
// (C3's 200-1400-byte payloads: ~1.75x fewer warp-steps).  Counting sort on the device: histogram, scan, scatter.
constexpr u32 ORDER_BUCKETS = 4096;
__global__ void k_order_hist(const u64 *off, u32 count, u32 *hist) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const u64 len = off[i + 1] - off[i];
    const u32 b = len >> 4 >= ORDER_BUCKETS ? ORDER_BUCKETS - 1 : (u32)(len >> 4);
    atomicAdd(&hist[ORDER_BUCKETS - 1 - b], 1u);   // bucket 0 = the longest buffers
}
__global__ void k_order_scan(u32 *hist) {  // one block of 1024: exclusive prefix sum over the 4096 buckets
    __shared__ u32 part[1024];
    const u32 t = threadIdx.x;
    u32 v[4], sum = 0;
    for (int k = 0; k < 4; k++) { v[k] = hist[4 * t + k]; sum += v[k]; }
    part[t] = sum;
    __syncthreads();
    for (u32 d = 1; d < 1024; d <<= 1) {
        const u32 x = t >= d ? part[t - d] : 0;
        __syncthreads();
        part[t] += x;
        __syncthreads();
    }
    u32 base = part[t] - sum;
    for (int k = 0; k < 4; k++) { hist[4 * t + k] = base; base += v[k]; }
}
__global__ void k_order_scatter(const u64 *off, u32 count, u32 *cursor, u32 *order) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const u64 len = off[i + 1] - off[i];
    const u32 b = len >> 4 >= ORDER_BUCKETS ? ORDER_BUCKETS - 1 : (u32)(len >> 4);
    order[atomicAdd(&cursor[ORDER_BUCKETS - 1 - b], 1u)] = i;
}

__global__ void k_histogram(const hs_gpu_match_t *m, u64 n, unsigned long long *hist, u32 n_ids) {
    const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && m[i].id < n_ids) atomicAdd(&hist[m[i].id], 1ull);
}

__global__ void k_convert(const Rec *in, hs_gpu_match_t *out, u64 n) {
    const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const Rec r = in[i];
        out[i] = hs_gpu_match_t{r.buf, r.id, (uint64_t)r.to};
    }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
#define CU_OK(expr)                                                                          \
    do {                                                                                     \
        cudaError_t e__ = (expr);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            if (getenv("HS_GPU_DEBUG"))                                                      \
                fprintf(stderr, "hs_gpu: %s -> %s (%s:%d)\n", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
            return e__ == cudaErrorMemoryAllocation ? HS_NOMEM : HS_UNKNOWN_ERROR;           \
        }                                                                                    \
    } while (0)

template <class T>
struct DevBuf {  // grow-only device allocation
    T *p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = n + n / 4 + 64;
        cudaError_t e = cudaMalloc((void **)&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    size_t bytes() const { return cap * sizeof(T); }
};

template <class T>
struct PinnedBuf {  // grow-only page-locked host array (so cudaMemcpyAsync from it really is asynchronous)
    T *p = nullptr;
    size_t cap = 0, n = 0;
    bool resize(size_t want) {
        if (want > cap) {
            if (p) cudaFreeHost(p);
            p = nullptr;
            cap = 0;
            const size_t c = want + want / 4 + 64;
            if (cudaHostAlloc((void **)&p, c * sizeof(T), cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); n = 0; return false; }
            cap = c;
        }
        n = want;
        return true;
    }
    T *data() { return p; }
    size_t size() const { return n; }
    T &operator[](size_t i) { return p[i]; }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = n = 0; }
};

struct DbImage {  // one uploaded database
    const void *host_db = nullptr;
    u32 crc = 0, bc_len = 0;
    DbView view;
    DevDb dev{};
    u8 *d_bc = nullptr;
    u64 *d_fdr = nullptr;
    u32 *d_teddy = nullptr;
    u64 *d_pair = nullptr;
    turbo::Slot *d_slots = nullptr;
    u32 *d_bloom = nullptr;
    u32 *d_gram = nullptr;
    turbo::Tables turbo_tabs;
    u64 *d_lx = nullptr;
    limex_dev::Tables lx_tabs;
    DevLimex lx{};
    DevMcc mcc{};
    bool ordered_ok = true;
    size_t device_bytes = 0;
    // full-Rose database of outfix engines (view.kind == ENG_MULTI): one image per queue, sharing this image's bytecode
    std::vector<DbImage *> subs;
    bool owns_bc = true;
    // McClellan engines of `subs` re-laid on a joint alphabet and stepped together (dfapack.cuh); packed[q] = engine q is
    // in one of them
    struct DfaPackImage { DevDfaPack dev; u8 *d_blob = nullptr; int D = 2; };
    std::vector<DfaPackImage> packs;
    std::vector<u8> packed;
    void release() {
        for (DbImage *sub : subs) { sub->release(); delete sub; }
        subs.clear();
        for (DfaPackImage &pk : packs) if (pk.d_blob) cudaFree(pk.d_blob);
        packs.clear();
        packed.clear();
        if (d_bc && owns_bc) cudaFree(d_bc);
        if (d_fdr) cudaFree(d_fdr);
        if (d_teddy) cudaFree(d_teddy);
        if (d_pair) cudaFree(d_pair);
        if (d_slots) cudaFree(d_slots);
        if (d_bloom) cudaFree(d_bloom);
        if (d_gram) cudaFree(d_gram);
        d_gram = nullptr;
        if (d_lx) cudaFree(d_lx);
        d_lx = nullptr;
        d_bloom = nullptr;
        d_bc = nullptr; d_fdr = nullptr; d_teddy = nullptr; d_pair = nullptr; d_slots = nullptr;
    }
};

}  // namespace

struct hs_gpu_corpus {
    u32 magic = 0x43525053;
    int device = 0;
    DevBuf<u8> text;          // FRONT_PAD + total + BACK_PAD
    DevBuf<u64> off;
    DevBuf<u32> cell_first;
    // buffers by descending length for the engine kernels (k_order_*): built by the first scan that wants it, kept while
    // the corpus keeps its contents (the staging corpus of host batches is refilled, and reordered, every call)
    mutable DevBuf<u32> order;
    mutable bool order_ready = false;
    // the same text seen as ONE buffer (vectored-mode databases: the buffers are consecutive blocks of one stream):
    // offsets {0, total} and an all-zero cell index, built by the first vectored scan of these contents
    mutable DevBuf<u64> stream_off;
    mutable DevBuf<u32> stream_cell;
    mutable bool stream_ready = false;
    u64 total = 0;
    u32 count = 0, nstrips = 0;
    DevCorpus dev() const {
        return DevCorpus{text.p + FRONT_PAD, off.p, cell_first.p, total, count, nstrips, nullptr, nullptr, nullptr};
    }
    DevCorpus dev_stream() const {
        return DevCorpus{text.p + FRONT_PAD, stream_off.p, stream_cell.p, total, 1, nstrips, nullptr, nullptr, nullptr};
    }
};

// hs_stream_t (src/state.h, runtime.c:543-600): what a pure-literal stream carries from block to block -- how far it has
// come, its last historyRequired bytes, the exhaustion keys of the single-match reports already delivered, its status.
struct hs_gpu_stream {
    u32 magic = 0x5354524d;
    const hs_database_t *db = nullptr;
    u32 crc = 0, hreq = 0;
    u64 offset = 0;
    std::vector<u8> hist;        // min(offset, hreq) bytes
    std::vector<u32> exhausted;  // sorted exhaustion keys
    bool terminated = false;     // a callback asked to stop (STATUS_TERMINATED): later blocks return HS_SCAN_TERMINATED
    bool in_call = false;
    // single-DFA databases: the engine's state (device state id) where the stream stands; 0 = dead (the reference marks
    // the stream exhausted, src/runtime.c:862-866); `started` false = no block scanned yet (the first one begins in the
    // engine's anchored start state, as MQE_TOP at stream offset 0 does, :850-854)
    bool started = false;
    u32 engine_state = 0;
};

struct hs_gpu_scratch {
    u32 magic = 0x47505553;  // 'GPUS'
    std::atomic<int> in_use{0};
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    std::vector<DbImage *> dbs;
    hs_gpu_corpus staging;       // device staging for host-buffer scans
    u8 *h_stage = nullptr;       // pinned host staging
    size_t h_stage_cap = 0;
    PinnedBuf<u64> h_off;
    PinnedBuf<u32> h_cell_first;
    PinnedBuf<u8> h_streams;     // hs_gpu_scan_streams: history + block of every stream of the call, back to back
    // stream mode of a DFA engine: per-buffer entry / exit states of the call being served (null pointers otherwise)
    PinnedBuf<u32> h_stream_state;   // [0, count): states the blocks start in; [count, 2 count): states they end in
    DevBuf<u32> d_stream_state;
    u32 stream_state_count = 0;      // > 0: the next scan_device carries DFA states in and out
    DevBuf<Rec> recs, recs2;
    DevBuf<hs_gpu_match_t> d_matches;
    DevBuf<u8> cub_tmp, keep;
    DevBuf<u32> order_hist;                 // engine kernels: buffers by descending length (k_order_*, histogram)
    long opt_order = 1;
    DevBuf<unsigned long long> ekey_keys;   // single-match resolution: open-addressing table (pair -> first record index)
    DevBuf<u32> ekey_first;
    DevBuf<u32> seg_count;
    u64 seg_cap_hint = 256;
    u64 out_hint = 1 << 16;                    // capacity of the ordered path's output arrays (grown on demand)
    DevBuf<unsigned long long> train_ctl;      // 2 x train_cap control blocks of 8 words (ordered path)
    PinnedBuf<unsigned long long> h_train_ctl;
    std::vector<cudaEvent_t> tev;              // 3 x train_cap events: (start, kernel end, compact end) per member
    u32 train_cap = 0;
    cudaStream_t copy_stream = nullptr;        // H2D of host batches, so that chunk k + 1 copies while chunk k is scanned
    std::vector<cudaEvent_t> chunk_ready;
    long opt_ordered = 1;
    DevBuf<unsigned long long> op_counts;      // option "count_ops": 128 counters, one per Rose opcode (tests)
    // control blocks of 8 words: [0] arena count, [1] candidates, [2] probes, [3] error(u32), [4] n_selected,
    // [5] largest segment.  Block 0 serves the general path; blocks 1 and 2 alternate between ordered scans
    // (k_compact clears the block of the next scan, so no memset sits in front of the kernel).
    unsigned long long *d_ctl = nullptr;
    u32 ctl_flip = 0;
    bool ctl_dirty = true;
    unsigned long long *h_ctl = nullptr;  // pinned mirror
    hs_gpu_match_t *h_matches = nullptr;  // pinned
    size_t h_matches_cap = 0;
    u64 n_matches = 0;           // valid entries in d_matches
    bool fetched = false;
    void *last_image = nullptr;  // DbImage that produced the current matches
    hs_gpu_stats_t stats{};
    int sm_count = 148;
    long opt_frontend = 0;   // 0: k_turbo2 or k_turbo by the tables' estimate, 1: the database's own FDR / Teddy / noodle
                             // tables (k_scan), 2: k_turbo, 3: k_turbo2
    long opt_grid = 0;
    long opt_warps = 0;      // k_turbo2: cap on warps per CTA (0 = as many as fit)
    long opt_pack = 1;       // multi-engine databases: 0 one launch per engine, 1 k_mcclellan_pack (work item = (buffer,
                             // engine); default), 2 k_dfa_pack (D engines stepped together per buffer on a joint alphabet;
                             // measured on C3: 18.9 ms against 18.5 ms for mode 1 -- kept as an option, not the default)
    long opt_pipeline = 0;   // host batches: 0 one copy then one scan (measured fastest), 1 chunk gathered input, 2 always chunk
    size_t smem_optin_max = 227 * 1024;
    u64 arena_hint = 1 << 16;
    u64 last_raw = ~0ull;  // record count of the previous scan (picks the finalize kernel size)
    u64 lx_event_hint = 1 << 16;
    // kernels whose > 48 KiB dynamic shared memory opt-in has been set ON THIS SCRATCH'S DEVICE (the attribute is per
    // device; a scratch has one device and one user at a time, so this needs no lock)
    std::vector<std::pair<const void *, size_t>> smem_optin;
};

namespace {

std::mutex g_init_mutex;

cudaError_t ensure_smem_optin(hs_gpu_scratch *s, const void *fn, size_t dyn) {
    for (auto &e : s->smem_optin)
        if (e.first == fn) {
            if (dyn <= e.second) return cudaSuccess;
            cudaError_t rv = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
            if (rv == cudaSuccess) e.second = dyn;
            return rv;
        }
    cudaError_t rv = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
    if (rv == cudaSuccess) s->smem_optin.emplace_back(fn, dyn);
    return rv;
}

// Sheng -> state-major table (see the upload).  Reference semantics kept: state byte = id | flags
// (shengcompile.cpp:343-357); a DEAD-flagged state is a trap without reports (census), so it may become state 0.
void build_sheng_tables(const DbView &v, u32 at, std::vector<u8> &out, DevMcc &D) {
    const DbView::Sheng &S = v.sheng;
    const sheng::Shape sh = sheng::shape(S.nfa_type);   // Sheng (16 states), Sheng32, Sheng64
    const u8 *m = v.bc + S.sh_off;
    const u32 n = S.n_states;
    bool dead[64] = {false}, acc[64] = {false};
    auto note = [&](u8 e) { if (e & sh.dead) dead[e & sh.mask] = true; };
    note((u8)S.anchored);
    for (u32 c = 0; c < 256; c++) for (u32 st = 0; st < n; st++) note(m[(size_t)sh.stride * c + st]);
    for (u32 st = 0; st < n; st++) acc[st] = rd32(v.bc, S.aux_off + st * sheng::aux::SIZE + sheng::aux::accept) != 0;
    u32 new_of[64], old_of[65], next = 1;
    for (u32 st = 0; st < n; st++) if (dead[st]) new_of[st] = 0;
    for (u32 st = 0; st < n; st++) if (!dead[st] && !acc[st]) { old_of[next] = st; new_of[st] = next++; }
    const u32 accept_limit = next;
    for (u32 st = 0; st < n; st++) if (!dead[st] && acc[st]) { old_of[next] = st; new_of[st] = next++; }
    const u32 rows = next, tab_bytes = rows * 256;
    out.assign(256 + tab_bytes + rows * mcclellan::aux::SIZE, 0);
    for (u32 c = 0; c < 256; c++) out[c] = (u8)c;  // remap: identity, alphaShift 8
    for (u32 q = 1; q < rows; q++) {
        for (u32 c = 0; c < 256; c++) out[256 + q * 256 + c] = (u8)new_of[m[(size_t)sh.stride * c + old_of[q]] & sh.mask];
        u8 *a = out.data() + 256 + tab_bytes + q * mcclellan::aux::SIZE;
        wr32(a, mcclellan::aux::accept, rd32(v.bc, S.aux_off + old_of[q] * sheng::aux::SIZE + sheng::aux::accept));
        wr32(a, mcclellan::aux::accept_eod, rd32(v.bc, S.aux_off + old_of[q] * sheng::aux::SIZE + sheng::aux::accept_eod));
    }
    D = DevMcc{};
    D.remap_abs = at; D.tab_abs = at + 256; D.tab_bytes = tab_bytes; D.aux_abs = at + 256 + tab_bytes;
    D.nfa_abs = v.nfa_off; D.sherman_abs = 0; D.wide16 = 0; D.as = 8; D.start = new_of[S.anchored & sh.mask];
    D.accept_limit_8 = accept_limit; D.sherman_limit = 0xffffffffu; D.single = (S.flags & sheng::FLAG_SINGLE_REPORT) ? 1u : 0u;
    D.arb_report = S.report; D.accepts_eod = (v.nfa_flags & nfa::ACCEPTS_EOD) ? 1u : 0u; D.in_smem = 1; D.has_sherman = 0;
}

// A McSheng engine (mcsheng_internal.h:63-88, mcsheng.c:150-322,368-420) is McClellan whose states below sheng_end step
// through byte-shuffle masks and have no row in the successor table.  At upload the rows are made: row s of the region =
// the mask bytes of one representative byte per symbol, translated to state ids (mask value sheng_end - 1 = dead,
// below it id - 1, above it the id itself), with the accept flag of 16-bit entries taken from the successor's aux record
// (the reference tests get_aux(s)->accept when the region is left, mcsheng.c:466); the rows of the other states are
// the engine's own.  The result is a complete McClellan table behind the bytecode, stepped by k_mcclellan.
void build_mcsheng_tables(const DbView &v, u32 at, std::vector<u8> &out, DevMcc &D) {
    const DbView::Mcc &M = v.mcc;
    const u32 rows = M.wide16 ? M.sherman_limit : M.states, ent = M.wide16 ? 2u : 1u, width = 1u << M.as;
    out.assign((size_t)rows * width * ent, 0);
    int rep[256];
    for (int &x : rep) x = -1;
    for (u32 c = 0; c < 256; c++) { const u32 cp = rd8(v.bc, M.remap_off + c); if (rep[cp] < 0) rep[cp] = (int)c; }
    for (u32 st = 1; st < M.sheng_end; st++)
        for (u32 cp = 0; cp < width; cp++) {
            if (rep[cp] < 0) continue;  // no byte maps to this symbol (TOP)
            const u32 mv = rd8(v.bc, M.masks_off + M.mask_stride * (u32)rep[cp] + st - 1);
            u32 nx = mv == M.sheng_end - 1 ? 0u : mv < M.sheng_end - 1 ? mv + 1 : mv;
            if (M.wide16 && rd32(v.bc, M.aux_off + nx * mcclellan::aux::SIZE + mcclellan::aux::accept)) nx |= mcclellan::ACCEPT_FLAG;
            const size_t i = ((size_t)st << M.as) + cp;
            if (M.wide16) { out[2 * i] = (u8)nx; out[2 * i + 1] = (u8)(nx >> 8); } else out[i] = (u8)nx;
        }
    std::memcpy(out.data() + ((size_t)M.sheng_end << M.as) * ent, v.bc + M.tab_off, M.tab_bytes);
    D = DevMcc{};
    D.tab_abs = at; D.tab_bytes = (u32)out.size(); D.remap_abs = M.remap_off;
    D.nfa_abs = v.nfa_off; D.aux_abs = M.aux_off; D.sherman_abs = M.sherman_off;
    D.wide16 = M.wide16; D.as = M.as; D.start = M.start; D.accept_limit_8 = M.accept_limit_8;
    D.sherman_limit = M.sherman_limit; D.single = M.single; D.arb_report = M.arb_report;
    D.accepts_eod = (v.nfa_flags & nfa::ACCEPTS_EOD) ? 1u : 0u;
    D.in_smem = MCC_RM_BYTES + (size_t)D.tab_bytes + 8 <= MCC_SMEM_MAX ? 1u : 0u;
    D.has_sherman = M.wide16 && M.sherman_limit < M.states ? 1u : 0u;
}

// A 16-bit McClellan engine with WIDE states (mcclellan_internal.h:53-61; runtime doWide16, mcclellan_common_impl.h:86-188):
// ids >= wide_limit stand for chains of states that the compiler stored as {symbols, successor after the last symbol,
// one shared failure row}.  At upload every chain position gets a table row again -- on its own symbol the next position
// (the last one: the stored successor), on any other symbol the failure row's entry -- and, while a complete table is
// being made anyway, every Sherman state its full row (daddy's row with the listed exceptions), so the engine is stepped
// by the branch-free k_mcclellan.  Chain position j of head h: j = 0 keeps id h, the others are numbered behind the
// heads.  The mstate_aux array is copied behind the table with empty records for the wide ids (they never accept, and
// the end-of-data check of a scan that stops inside a wide state reports nothing, mcclellan.c:615).
void build_wide_tables(const DbView &v, u32 at, std::vector<u8> &out, DevMcc &D) {
    namespace mc = mcclellan;
    const DbView::Mcc &M = v.mcc;
    const u32 rows = M.wide_limit + M.wide_rows, width = 1u << M.as;
    const size_t tab_bytes = (size_t)rows * width * 2, aux_at = round_up(tab_bytes, 16);
    out.assign(aux_at + (size_t)rows * mc::aux::SIZE, 0);
    auto put = [&](u32 row, u32 cp, u32 e) { const size_t i = ((size_t)row << M.as) + cp; out[2 * i] = (u8)e; out[2 * i + 1] = (u8)(e >> 8); };
    std::memcpy(out.data(), v.bc + M.tab_off, M.tab_bytes);                      // rows below sherman_limit
    for (u32 st = M.sherman_limit; st < M.wide_limit; st++) {                    // doSherman16, mcclellan_common_impl.h:54-84
        const u8 *sh = v.bc + M.sherman_off + (size_t)mc::SHERMAN_FIXED_SIZE * (st - M.sherman_limit);
        const u32 len = sh[mc::SHERMAN_LEN_OFFSET], daddy = rd16(sh, mc::SHERMAN_DADDY_OFFSET);
        std::memcpy(out.data() + ((size_t)st << M.as) * 2, out.data() + ((size_t)daddy << M.as) * 2, (size_t)width * 2);
        for (u32 k = 0; k < len; k++) put(st, sh[mc::SHERMAN_CHARS_OFFSET + k], rd16(sh, mc::SHERMAN_CHARS_OFFSET + len + 2 * k));
    }
    u32 inner = M.wide_limit + M.wide_number;                                    // next id of a chain position behind a head
    for (u32 k = 0; k < M.wide_number; k++) {
        const u8 *we = v.bc + M.wide_off + rd32(v.bc, M.wide_off + 4 + 4 * k);
        const u32 w = rd16(we, 0);
        const u8 *sym = we + 2, *tr = we + 2 + ((w + 1) & ~1u);
        for (u32 j = 0; j < w; j++) {
            const u32 row = j ? inner + j - 1 : M.wide_limit + k;
            for (u32 cp = 0; cp < M.n_alpha; cp++) put(row, cp, rd16(tr, 2 + 2 * (size_t)cp));
            put(row, sym[j], j + 1 < w ? inner + j : rd16(tr, 0));
        }
        inner += w - 1;
    }
    std::memcpy(out.data() + aux_at, v.bc + M.aux_off, (size_t)M.wide_limit * mc::aux::SIZE);
    D = DevMcc{};
    D.tab_abs = at; D.tab_bytes = (u32)tab_bytes; D.remap_abs = M.remap_off;
    D.nfa_abs = v.nfa_off; D.aux_abs = at + (u32)aux_at; D.sherman_abs = 0;
    D.wide16 = 1; D.as = M.as; D.start = M.start; D.accept_limit_8 = M.accept_limit_8;
    D.sherman_limit = rows; D.single = M.single; D.arb_report = M.arb_report;
    D.accepts_eod = (v.nfa_flags & nfa::ACCEPTS_EOD) ? 1u : 0u;
    D.in_smem = MCC_RM_BYTES + (size_t)D.tab_bytes + 8 <= MCC_SMEM_MAX ? 1u : 0u;
    D.has_sherman = 0;
}

// Device tables of ONE outfix engine (im->view describes it; im->dev.bc is the uploaded bytecode it lives in).
hs_error_t build_engine_image(hs_gpu_scratch *s, DbImage *im) {
    const DbView &v = im->view;
    DevDb &d = im->dev;
    if (v.kind == ENG_LIMEX) {
        limex_dev::Tables &t = im->lx_tabs;
        if (!limex_dev::build(v, t)) return HS_UNKNOWN_ERROR;
        if (cudaMalloc((void **)&im->d_lx, t.blob.size() * 8) != cudaSuccess) return HS_NOMEM;
        cudaMemcpyAsync(im->d_lx, t.blob.data(), t.blob.size() * 8, cudaMemcpyHostToDevice, s->stream);
        im->device_bytes += t.blob.size() * 8;
        d.lx_abs = t.limex_abs;
        DevLimex &L = im->lx;
        L = DevLimex{};
        L.tab = im->d_lx; L.G = t.G; L.n_exc = t.n_exc; L.shift_count = t.shift_count;
        L.any_squash = t.any_squash ? 1u : 0u; L.can_die = t.can_die ? 1u : 0u;
        for (u32 k = 0; k < 8; k++) L.shmul[k] = 1u << ((u32)(t.shift_amounts >> (8 * k)) & 31u);
        L.o_reach = t.o_reach; L.o_shift = t.o_shift; L.o_shamt = t.o_shamt; L.o_csucc = t.o_csucc; L.o_succ = t.o_succ;
        L.o_squash = t.o_squash; L.o_meta = t.o_meta; L.o_masks = t.o_masks; L.o_base = t.o_base;
        L.total_words = t.total_words; L.accept_abs = t.accept_abs; L.accept_eod_abs = t.accept_eod_abs;
        L.nfa_flags = t.nfa_flags;
    } else if (v.kind == ENG_SHENG || (v.kind == ENG_MCCLELLAN && im->mcc.tab_abs)) {
        // im->mcc was filled by build_sheng_tables / build_mcsheng_tables / build_wide_tables with the image
    } else if (v.kind == ENG_MCCLELLAN) {
        // the engine is stepped straight from the uploaded bytecode (mcclellan.cuh copies the table to shared memory)
        const DbView::Mcc &M = v.mcc;
        DevMcc &D = im->mcc;
        D = DevMcc{};
        D.tab_abs = M.tab_off; D.tab_bytes = M.tab_bytes; D.remap_abs = M.remap_off;
        D.nfa_abs = v.nfa_off; D.aux_abs = M.aux_off; D.sherman_abs = M.sherman_off;
        D.wide16 = M.wide16; D.as = M.as; D.start = M.start; D.accept_limit_8 = M.accept_limit_8;
        D.sherman_limit = M.sherman_limit; D.single = M.single; D.arb_report = M.arb_report;
        D.accepts_eod = (v.nfa_flags & nfa::ACCEPTS_EOD) ? 1u : 0u;
        D.in_smem = MCC_RM_BYTES + (size_t)M.tab_bytes + 8 <= MCC_SMEM_MAX ? 1u : 0u;
        D.has_sherman = M.wide16 && M.sherman_limit < M.states ? 1u : 0u;
    }
    return HS_SUCCESS;
}

// Packs of McClellan outfix engines for k_dfa_pack (dfapack.cuh): engines are taken in queue order while the pack's
// tables -- every engine's successor table re-laid on the pack's JOINT alphabet, 16-bit entries with the accept flag --
// fit shared memory beside the class table.  Two bytes share a joint class when every engine of the pack maps them to
// the same symbol.  A pack of one is not built (k_mcclellan serves it).
hs_error_t build_dfa_packs(hs_gpu_scratch *s, DbImage *im) {
    const size_t nq = im->subs.size();
    im->packed.assign(nq, 0);
    auto packable = [&](size_t q) {
        const DbImage *e = im->subs[q];
        const DbView::Mcc &M = e->view.mcc;
        return e->view.kind == ENG_MCCLELLAN && !M.sheng_end && !M.has_wide && !(M.wide16 && M.sherman_limit < M.states) && !e->mcc.has_sherman && !e->mcc.accepts_eod && M.states <= mcclellan::STATE_MASK;
    };
    auto joint = [&](const std::vector<size_t> &members, std::vector<u32> &cls_of) {  // -> number of joint classes
        std::map<std::vector<u8>, u32> seen;
        cls_of.assign(256, 0);
        for (u32 c = 0; c < 256; c++) {
            std::vector<u8> sig;
            for (size_t q : members) sig.push_back(rd8(im->view.bc, im->subs[q]->view.mcc.remap_off + c));
            auto it = seen.find(sig);
            if (it == seen.end()) it = seen.emplace(sig, (u32)seen.size()).first;
            cls_of[c] = it->second;
        }
        return (u32)seen.size();
    };
    auto fits = [&](const std::vector<size_t> &members, u32 &row_shift, size_t &bytes) {
        std::vector<u32> cls;
        const u32 J = joint(members, cls);
        row_shift = 1;
        while ((1u << (row_shift - 1)) < J) row_shift++;
        bytes = 0;
        for (size_t q : members) bytes += round_up((size_t)im->subs[q]->view.mcc.states << row_shift, 16);
        return MCC_RM_BYTES + bytes + 16 <= MCC_SMEM_MAX;
    };
    std::vector<size_t> cur;
    auto close = [&]() -> hs_error_t {
        if (cur.size() >= 2) {
            u32 row_shift = 1;
            size_t bytes = 0;
            fits(cur, row_shift, bytes);
            std::vector<u32> cls;
            joint(cur, cls);
            std::vector<u8> rep_of;  // a representative byte of every joint class
            {
                u32 J = 0;
                for (u32 c = 0; c < 256; c++) J = std::max(J, cls[c] + 1);
                rep_of.assign(J, 0);
                for (u32 c = 256; c-- > 0;) rep_of[cls[c]] = (u8)c;
            }
            std::vector<u8> blob(1024 + bytes, 0);
            for (u32 c = 0; c < 256; c++) wr32(blob.data(), 4 * c, cls[c] * 2);
            DbImage::DfaPackImage pk;
            pk.dev = DevDfaPack{};
            size_t at = 0;
            for (size_t k = 0; k < cur.size(); k++) {
                const DbImage *e = im->subs[cur[k]];
                const DbView::Mcc &M = e->view.mcc;
                pk.dev.tab_off[k] = (u32)at;
                pk.dev.start[k] = M.start;
                pk.dev.m[k] = e->mcc;
                for (u32 st = 0; st < M.states; st++)
                    for (u32 j = 0; j < rep_of.size(); j++) {
                        const u32 sym = rd8(im->view.bc, M.remap_off + rep_of[j]);
                        u32 out;
                        if (M.wide16) {
                            const u32 en = rd16(im->view.bc, M.tab_off + 2 * (((size_t)st << M.as) + sym));
                            out = (en & mcclellan::STATE_MASK) | ((en & mcclellan::ACCEPT_FLAG) ? 0x8000u : 0u);
                        } else {
                            const u32 nx = rd8(im->view.bc, M.tab_off + ((size_t)st << M.as) + sym);
                            out = nx | (nx >= M.accept_limit_8 ? 0x8000u : 0u);
                        }
                        wr16(blob.data(), 1024 + at + ((size_t)st << row_shift) + 2 * j, (u16)out);
                    }
                at += round_up((size_t)M.states << row_shift, 16);
                im->packed[cur[k]] = 1;
            }
            pk.dev.n = (u32)cur.size();
            pk.dev.row_shift = row_shift;
            pk.dev.tab_bytes = (u32)bytes;
            pk.D = cur.size() <= 2 ? 2 : cur.size() <= 4 ? 4 : 8;
            if (cudaMalloc((void **)&pk.d_blob, blob.size()) != cudaSuccess) return HS_NOMEM;
            cudaMemcpyAsync(pk.d_blob, blob.data(), blob.size(), cudaMemcpyHostToDevice, s->stream);
            cudaStreamSynchronize(s->stream);  // `blob` is a local
            pk.dev.blob = pk.d_blob;
            im->device_bytes += blob.size();
            im->packs.push_back(pk);
        }
        cur.clear();
        return HS_SUCCESS;
    };
    for (size_t q = 0; q < nq; q++) {
        if (!packable(q)) continue;
        cur.push_back(q);
        u32 rs_ = 0;
        size_t b_ = 0;
        if (cur.size() > (size_t)DP_MAX || !fits(cur, rs_, b_)) {
            cur.pop_back();
            const hs_error_t rv = close();
            if (rv != HS_SUCCESS) return rv;
            cur.push_back(q);
            if (!fits(cur, rs_, b_)) cur.clear();   // does not even fit alone: k_mcclellan reads its table through L2
        }
    }
    return close();
}

hs_error_t upload_database(hs_gpu_scratch *s, const hs_database_t *db, DbImage **out) {
    // cached by (pointer, crc)
    const u32 crc = rd32(db, db::crc32);
    for (DbImage *im : s->dbs)
        if (im->host_db == db && im->crc == crc) { *out = im; return HS_SUCCESS; }
    DbImage *im = new DbImage();
    std::string why;
    int rv = parse_database(db, im->view, why);
    if (rv != E_SUCCESS) {
        if (getenv("HS_GPU_DEBUG")) fprintf(stderr, "hs_gpu: database refused: %s\n", why.c_str());
        delete im;
        return rv;
    }
    const DbView &v = im->view;
    im->host_db = db;
    im->crc = crc;
    im->bc_len = v.bc_len;
    // A Sheng engine (<= 16 states as 256 shuffle masks) is re-laid at upload as a state-major byte table with the
    // McClellan8 conventions (dead = 0, accepting states numbered last) behind the bytecode, and stepped by k_mcclellan.
    // (Sheng outfixes of a multi-engine database get their tables one behind the other in the same area.)
    std::vector<u8> derived;
    const u32 derived_at = (u32)round_up(v.bc_len, 256) + 256;
    // McSheng and McClellan with wide states: their complete tables are made here too.  So are those of plain 16-bit
    // engines with Sherman states when the complete table stays small (Sherman rows are a size optimisation for a CPU
    // cache; here a full row per state buys the branch-free loop): HS_GPU_SHERMAN_EXACT=1 keeps the reference's form,
    // stepped by the exact loop (doSherman16 per byte).
    const bool expand_sherman = !getenv("HS_GPU_SHERMAN_EXACT");
    auto hybrid = [expand_sherman](const DbView &e) {
        if (e.kind != ENG_MCCLELLAN) return false;
        if (e.mcc.sheng_end != 0 || e.mcc.has_wide) return true;
        return expand_sherman && e.mcc.wide16 && e.mcc.sherman_limit < e.mcc.states && ((size_t)e.mcc.states << e.mcc.as) * 2 <= (4u << 20);
    };
    auto build_hybrid = [](const DbView &e, u32 at, std::vector<u8> &out, DevMcc &D) {
        if (e.mcc.sheng_end) build_mcsheng_tables(e, at, out, D); else build_wide_tables(e, at, out, D);
    };
    if (v.kind == ENG_SHENG) build_sheng_tables(v, derived_at, derived, im->mcc);
    if (hybrid(v)) build_hybrid(v, derived_at, derived, im->mcc);
    std::vector<DevMcc> sub_sheng(v.engines.size());
    for (size_t q = 0; q < v.engines.size(); q++) {
        if (v.engines[q].kind != ENG_SHENG && !hybrid(v.engines[q])) continue;
        std::vector<u8> one;
        const u32 at = derived_at + (u32)round_up(derived.size(), 256);
        if (hybrid(v.engines[q])) build_hybrid(v.engines[q], at, one, sub_sheng[q]);
        else build_sheng_tables(v.engines[q], at, one, sub_sheng[q]);
        derived.resize(at - derived_at, 0);
        derived.insert(derived.end(), one.begin(), one.end());
    }
    const size_t image_bytes = derived_at + (derived.empty() ? 0 : round_up(derived.size(), 256) + 256);
    cudaError_t e = cudaMalloc((void **)&im->d_bc, image_bytes);
    if (e != cudaSuccess) { delete im; return HS_NOMEM; }
    cudaMemcpyAsync(im->d_bc, v.bc, v.bc_len, cudaMemcpyHostToDevice, s->stream);
    if (!derived.empty()) {
        cudaMemcpyAsync(im->d_bc + derived_at, derived.data(), derived.size(), cudaMemcpyHostToDevice, s->stream);
        cudaStreamSynchronize(s->stream);  // `derived` is a local
    }
    im->device_bytes = image_bytes;
    DevDb &d = im->dev;
    d.bc = im->d_bc;
    d.groups = v.groups;
    d.conf_base = v.conf_base;
    d.n_buckets = v.n_buckets;
    d.min_width = v.min_width;
    d.min_width_excl = v.min_width_excl;
    d.max_bi_width = v.max_bi_width;
    d.bc_len = v.bc_len;
    std::vector<u64> folded;
    std::vector<u32> ttab;
    if (v.kind == ENG_FDR) {
        // Reach table for shared memory: domains up to 14 are used as they are; a domain-15 table (256 KiB)
        // is folded to 2^14 rows by AND-ing the two rows that differ in key bit 14.  A cleared bit means
        // "possible", so the fold only over-approximates: the filter stays sound, confirm stays exact.
        const u32 rows = 1u << v.fdr_domain, keep = std::min<u32>(rows, 1u << 14);
        folded.resize(keep);
        for (u32 i = 0; i < keep; i++) {
            u64 r = rd64(v.bc, v.fdr_tab + 8 * (size_t)i);
            for (u32 j = i + keep; j < rows; j += keep) r &= rd64(v.bc, v.fdr_tab + 8 * (size_t)j);
            folded[i] = r;
        }
        e = cudaMalloc((void **)&im->d_fdr, keep * 8);
        if (e != cudaSuccess) { im->release(); delete im; return HS_NOMEM; }
        cudaMemcpyAsync(im->d_fdr, folded.data(), keep * 8, cudaMemcpyHostToDevice, s->stream);
        im->device_bytes += keep * 8;
        d.fdr_tab = im->d_fdr;
        d.fdr_mask = keep - 1;
        d.fdr_stride = v.fdr_stride;
    } else if (v.kind == ENG_TEDDY) {
        ttab.assign(256, 0);
        for (u32 cch = 0; cch < 256; cch++) {
            u32 t = 0;
            for (u32 j = 0; j < v.teddy_masks; j++) {
                if (v.teddy_fat) {
                    // 16 buckets folded onto 8 filter bits: bit b stays clear ("possible") when bucket b OR bucket b + 8
                    // admits the byte; confirm_entry then walks both buckets' FDRConfirm (sound: over-approximation only)
                    const size_t lo0 = v.teddy_mask_off + (size_t)(4 * j) * 16, hi0 = v.teddy_mask_off + (size_t)(4 * j + 2) * 16;
                    const u8 a = rd8(v.bc, lo0 + (cch & 15)) | rd8(v.bc, hi0 + (cch >> 4));
                    const u8 b = rd8(v.bc, lo0 + 16 + (cch & 15)) | rd8(v.bc, hi0 + 16 + (cch >> 4));
                    t |= (u32)(a & b) << (8 * j);
                    continue;
                }
                const u8 lo = rd8(v.bc, v.teddy_mask_off + (size_t)(2 * j) * 16 + (cch & 15));
                const u8 hi = rd8(v.bc, v.teddy_mask_off + (size_t)(2 * j + 1) * 16 + (cch >> 4));
                t |= (u32)(lo | hi) << (8 * j);
            }
            ttab[cch] = t;
        }
        e = cudaMalloc((void **)&im->d_teddy, 256 * 4);
        if (e != cudaSuccess) { im->release(); delete im; return HS_NOMEM; }
        cudaMemcpyAsync(im->d_teddy, ttab.data(), 256 * 4, cudaMemcpyHostToDevice, s->stream);
        im->device_bytes += 1024;
        d.teddy_tab = im->d_teddy;
        d.teddy_masks = v.teddy_masks;
    } else if (v.kind == ENG_LIMEX || v.kind == ENG_SHENG || v.kind == ENG_MCCLELLAN) {
        const hs_error_t erv = build_engine_image(s, im);
        if (erv != HS_SUCCESS) { im->release(); delete im; return erv; }
    } else if (v.kind == ENG_MULTI) {
        // engines only: images built below
    } else {
        d.nood_msk = v.nood_msk;
        d.nood_cmp = v.nood_cmp;
        d.nood_len = v.nood_len;
        d.nood_id = v.nood_id;
    }
    // outfix engines of a full-Rose database (alone, or beside the literal matcher): one image per queue
    for (const DbView &ev : v.engines) {
        DbImage *sub = new DbImage();
        sub->view = ev;
        sub->host_db = db;
        sub->crc = crc;
        sub->bc_len = v.bc_len;
        sub->d_bc = im->d_bc;
        sub->owns_bc = false;
        sub->dev = im->dev;
        im->subs.push_back(sub);
        if (ev.kind == ENG_SHENG || hybrid(ev)) sub->mcc = sub_sheng[im->subs.size() - 1];   // tables built with the image (above)
        const hs_error_t erv = build_engine_image(s, sub);
        if (erv != HS_SUCCESS) { im->release(); delete im; return erv; }
        im->device_bytes += sub->device_bytes;
    }
    if (!im->subs.empty()) im->ordered_ok = false;  // engine matches merge with the literal ones in the sorting finalize
    im->packed.assign(im->subs.size(), 0);
    if (im->subs.size() >= 2) {
        const hs_error_t prv = build_dfa_packs(s, im);
        if (prv != HS_SUCCESS) { im->release(); delete im; return prv; }
    }
    d.tK = 0;
    if (v.kind != ENG_LIMEX && v.kind != ENG_MCCLELLAN && v.kind != ENG_SHENG && v.kind != ENG_MULTI && turbo::build(v, im->turbo_tabs)) {
        const turbo::Tables &tt = im->turbo_tabs;
        if (cudaMalloc((void **)&im->d_pair, tt.pair.size() * 8) != cudaSuccess ||
            cudaMalloc((void **)&im->d_slots, tt.slots.size() * sizeof(turbo::Slot)) != cudaSuccess ||
            cudaMalloc((void **)&im->d_bloom, tt.bloom.size() * 4) != cudaSuccess ||
            cudaMalloc((void **)&im->d_gram, tt.gram.size() * 4) != cudaSuccess) {
            im->release(); delete im; return HS_NOMEM;
        }
        cudaMemcpyAsync(im->d_gram, tt.gram.data(), tt.gram.size() * 4, cudaMemcpyHostToDevice, s->stream);
        d.gram_tab = im->d_gram;
        d.gG1 = tt.G1; d.gmul = tt.gmul; d.gwide = tt.wide;
        cudaMemcpyAsync(im->d_pair, tt.pair.data(), tt.pair.size() * 8, cudaMemcpyHostToDevice, s->stream);
        cudaMemcpyAsync(im->d_slots, tt.slots.data(), tt.slots.size() * sizeof(turbo::Slot), cudaMemcpyHostToDevice, s->stream);
        cudaMemcpyAsync(im->d_bloom, tt.bloom.data(), tt.bloom.size() * 4, cudaMemcpyHostToDevice, s->stream);
        im->device_bytes += tt.pair.size() * 8 + tt.slots.size() * sizeof(turbo::Slot) + tt.bloom.size() * 4 + tt.gram.size() * 4;
        d.pair_tab = im->d_pair;
        d.slots = im->d_slots;
        d.bloom = im->d_bloom;
        d.tK = tt.K; d.tH1 = tt.H1; d.tH2 = tt.H2; d.tL = tt.L; d.tBL = tt.BL;
    }
    if (cudaStreamSynchronize(s->stream) != cudaSuccess) { im->release(); delete im; return HS_UNKNOWN_ERROR; }
    s->dbs.push_back(im);
    *out = im;
    return HS_SUCCESS;
}

template <int THREADS>
cudaError_t launch_small_finalize(hs_gpu_scratch *s, const Rec *in, Rec *out_recs, hs_gpu_match_t *out, unsigned long long *n_out,
                                  const unsigned long long *n_raw, u64 arena_cap) {
    using Sort = cub::BlockMergeSort<Rec, THREADS, SMALL_FIN_ITEMS>;
    const size_t dyn = sizeof(typename Sort::TempStorage);
    cudaError_t e = ensure_smem_optin(s, (const void *)k_small_finalize<THREADS>, dyn);
    if (e != cudaSuccess) return e;
    k_small_finalize<THREADS><<<1, THREADS, dyn, s->stream>>>(in, out_recs, out, n_out, n_raw, arena_cap);
    return cudaGetLastError();
}

inline bool resize_ok(std::vector<u32> &v, size_t n) { v.resize(n); return true; }
inline bool resize_ok(PinnedBuf<u32> &v, size_t n) { return v.resize(n); }

__global__ void k_stream_offsets(u64 *off, u64 total) { off[0] = 0; off[1] = total; }

// out[c] = buffer holding the first byte of cell c (CELL_BYTES each), out[ncells] = buffer of the last byte
template <class Vec>
bool build_cell_first(const u64 *off, u32 count, u64 total, Vec &out, u32 &nstrips) {
    nstrips = (u32)((total + STRIP_BYTES - 1) / STRIP_BYTES);
    const u32 ncells = (u32)((total + CELL_BYTES - 1) / CELL_BYTES);
    if (!resize_ok(out, (size_t)ncells + 1)) return false;
    u32 b = 0;
    for (u32 s = 0; s <= ncells; s++) {
        u64 pos = std::min<u64>((u64)s * CELL_BYTES, total ? total - 1 : 0);
        while (b + 1 < count && off[b + 1] <= pos) b++;
        out[s] = b;
    }
    return true;
}

u32 turbo_grid(const hs_gpu_scratch *s, const DevCorpus &cp, u32 warps_per_cta) {
    u32 grid = s->opt_grid > 0 ? (u32)s->opt_grid : (u32)s->sm_count;
    if (grid > 8192 / warps_per_cta) grid = 8192 / warps_per_cta;  // k_compact scans at most 8192 warp segments
    // the kernels balance 1 KiB steps over all warps of the grid: shrink the grid only when a warp would see fewer than 4
    const u64 steps = (cp.total + 1023) / 1024;
    const u32 need = (u32)((steps + 4 * warps_per_cta - 1) / (4 * warps_per_cta));
    if (grid > need) grid = need ? need : 1;
    return grid;
}

template <int K, bool ORDERED>
hs_error_t launch_turbo(hs_gpu_scratch *s, const DevDb &db, const DevCorpus &cp, const DevArena &ar) {
    const size_t dyn = turbo_smem_bytes<ORDERED>(db);
    CU_OK(ensure_smem_optin(s, (const void *)k_turbo<K, ORDERED>, dyn));
    k_turbo<K, ORDERED><<<turbo_grid(s, cp, turbo_warps<ORDERED>()), turbo_warps<ORDERED>() * 32, dyn, s->stream>>>(db, cp, ar);
    CU_OK(cudaGetLastError());
    return HS_SUCCESS;
}

// k_turbo2: as many warps per CTA (one CTA per SM) as the shared memory left by the tables holds, T2MAXW at most
template <bool ORDERED>
u32 turbo2_warps(const hs_gpu_scratch *s, const DevDb &db) {
    const size_t budget = s->smem_optin_max > 2048 ? s->smem_optin_max - 1024 : 0, tables = turbo2_smem_bytes<ORDERED>(db, 0);
    u32 w = tables < budget ? (u32)((budget - tables) / sizeof(Turbo2WarpShared<ORDERED>)) : 0;
    if (w > (u32)T2MAXW) w = T2MAXW;
    if (s->opt_warps > 0 && (u32)s->opt_warps < w) w = (u32)s->opt_warps;
    return w;
}

template <int K, bool ORDERED>
hs_error_t launch_turbo2(hs_gpu_scratch *s, const DevDb &db, const DevCorpus &cp, const DevArena &ar, u32 grid, u32 warps) {
    const size_t dyn = turbo2_smem_bytes<ORDERED>(db, (int)warps);
    CU_OK(ensure_smem_optin(s, (const void *)k_turbo2<K, ORDERED>, dyn));
    k_turbo2<K, ORDERED><<<grid, warps * 32, dyn, s->stream>>>(db, cp, ar);
    CU_OK(cudaGetLastError());
    return HS_SUCCESS;
}

template <bool ORDERED>
hs_error_t launch_turbo2_any(hs_gpu_scratch *s, const DevDb &db, const DevCorpus &cp, const DevArena &ar, u32 grid, u32 warps) {
    switch (db.tK) {
    case 1: return launch_turbo2<1, ORDERED>(s, db, cp, ar, grid, warps);
    case 2: return launch_turbo2<2, ORDERED>(s, db, cp, ar, grid, warps);
    case 3: return launch_turbo2<3, ORDERED>(s, db, cp, ar, grid, warps);
    default: return launch_turbo2<4, ORDERED>(s, db, cp, ar, grid, warps);
    }
}

// which literal front end serves this database: 0 option = by the tables' own estimate (turbo_tables.h `wide`)
inline bool use_turbo2(const hs_gpu_scratch *s, const DevDb &db) {
    if (s->opt_frontend == 2) return false;
    if (s->opt_frontend == 3) return true;
    return !db.gwide;
}

template <class FE, bool NOODLE>
hs_error_t launch_scan(hs_gpu_scratch *s, const DevDb &db, const DevCorpus &cp, const DevArena &ar) {
    const size_t dyn = FE::smem_bytes(db);
    CU_OK(ensure_smem_optin(s, (const void *)k_scan<FE, NOODLE>, dyn));
    u32 grid = s->opt_grid > 0 ? (u32)s->opt_grid : (u32)s->sm_count;
    const u32 need = (cp.nstrips + WARPS - 1) / WARPS;
    if (grid > need) grid = need ? need : 1;
    k_scan<FE, NOODLE><<<grid, THREADS, dyn, s->stream>>>(db, cp, ar);
    CU_OK(cudaGetLastError());
    return HS_SUCCESS;
}

template <int G>
hs_error_t launch_limex_g(hs_gpu_scratch *s, const DbImage *im, const DevCorpus &cp, const DevArena &ar, bool report) {
    const size_t dyn = limex_smem_bytes(im->lx);
    CU_OK(ensure_smem_optin(s, (const void *)k_limex<G>, dyn));
    // persistent groups of G lanes; no more CTAs than there are buffers to hand out
    const u32 groups_per_cta = LX_THREADS / G;
    u32 grid = s->opt_grid > 0 ? (u32)s->opt_grid : 4u * (u32)s->sm_count;
    const u32 need = (cp.count + groups_per_cta - 1) / groups_per_cta;
    if (grid > need) grid = need ? need : 1;
    // events (what the NFA raised) go to recs2, then one thread per event runs the report programs into the arena
    CU_OK(s->recs2.reserve(s->lx_event_hint));
    const DevLxEvents evq{reinterpret_cast<LxEvent *>(s->recs2.p), s->d_ctl + 7, s->recs2.cap};
    k_limex<G><<<grid, LX_THREADS, dyn, s->stream>>>(im->dev, im->lx, cp, evq, s->d_ctl + 6);
    CU_OK(cudaGetLastError());
    s->stats.kernel_launches++;
    if (!report) return HS_SUCCESS;
    k_limex_report<<<(u32)s->sm_count, LX_THREADS, 0, s->stream>>>(im->dev, cp, ar, evq);
    CU_OK(cudaGetLastError());
    s->stats.kernel_launches++;
    return HS_SUCCESS;
}

hs_error_t launch_limex(hs_gpu_scratch *s, const DbImage *im, const DevCorpus &cp, const DevArena &ar, bool report = true) {
    switch (im->lx.G) {
    case 1: return launch_limex_g<1>(s, im, cp, ar, report);
    case 2: return launch_limex_g<2>(s, im, cp, ar, report);
    case 4: return launch_limex_g<4>(s, im, cp, ar, report);
    default: return launch_limex_g<8>(s, im, cp, ar, report);
    }
}

template <bool WIDE16, bool SMEM, bool SHERMAN>
hs_error_t launch_mcclellan_t(hs_gpu_scratch *s, const DbImage *im, const DevCorpus &cp, const DevArena &ar, bool report) {
    const size_t dyn = mcc_smem_bytes(im->mcc);
    CU_OK(ensure_smem_optin(s, (const void *)k_mcclellan<WIDE16, SMEM, SHERMAN>, dyn));
    // persistent threads, one buffer each at a time, fed from a counter: every SM gets a CTA as soon as there is a
    // warp's worth of buffers for it
    u32 grid = s->opt_grid > 0 ? (u32)s->opt_grid : (u32)s->sm_count;
    const u32 need = (cp.count + 31) / 32;
    if (grid > need) grid = need ? need : 1;
    // 1024 threads per SM even when that leaves a thread with one buffer or none: the step is a dependent
    // shared-memory read, and chains in flight are what hides it (measured: 128/256/512/1024 threads -> 0.36/0.20/
    // 0.12/0.115 ms on 64 MiB of 200-1400-byte buffers)
    u32 block = MCC_THREADS;
    if (const char *e = getenv("HS_GPU_MCC_BLOCK")) { const int b = atoi(e); if (b >= 32 && b <= MCC_THREADS && !(b & 31)) block = (u32)b; }
    CU_OK(s->recs2.reserve(s->lx_event_hint));
    const DevLxEvents evq{reinterpret_cast<LxEvent *>(s->recs2.p), s->d_ctl + 7, s->recs2.cap};
    k_mcclellan<WIDE16, SMEM, SHERMAN><<<grid, block, dyn, s->stream>>>(im->dev, im->mcc, cp, evq, s->d_ctl + 6);
    CU_OK(cudaGetLastError());
    s->stats.kernel_launches++;
    if (!report) return HS_SUCCESS;
    k_limex_report<<<(u32)s->sm_count, LX_THREADS, 0, s->stream>>>(im->dev, cp, ar, evq);
    CU_OK(cudaGetLastError());
    s->stats.kernel_launches++;
    return HS_SUCCESS;
}

hs_error_t launch_mcclellan(hs_gpu_scratch *s, const DbImage *im, const DevCorpus &cp, const DevArena &ar, bool report = true) {
    const DevMcc &m = im->mcc;
    if (!m.wide16) return m.in_smem ? launch_mcclellan_t<false, true, false>(s, im, cp, ar, report) : launch_mcclellan_t<false, false, false>(s, im, cp, ar, report);
    if (m.has_sherman) return m.in_smem ? launch_mcclellan_t<true, true, true>(s, im, cp, ar, report) : launch_mcclellan_t<true, false, true>(s, im, cp, ar, report);
    return m.in_smem ? launch_mcclellan_t<true, true, false>(s, im, cp, ar, report) : launch_mcclellan_t<true, false, false>(s, im, cp, ar, report);
}

// Full-Rose database of outfix engines (roseBlockExec with outfixes only, src/rose/block.c:345-422, catchup.c:612-672):
// every engine steps the whole batch (its own kernel launch; a block shorter than the engine's NFA.minWidth cannot
// match it), all of them append to ONE event queue, and one report pass runs the Rose report programs; the finalize
// sorts by (buffer, offset, id) and applies the dedupe keys, which is what the catch-up queue's offset order is for.
template <bool WIDE16>
hs_error_t launch_mcclellan_pack(hs_gpu_scratch *s, const DbImage *im, const DevMccPack &pk, size_t dyn, const DevCorpus &cp) {
    CU_OK(ensure_smem_optin(s, (const void *)k_mcclellan_pack<WIDE16>, dyn));
    u32 grid = s->opt_grid > 0 ? (u32)s->opt_grid : (u32)s->sm_count;
    const unsigned long long items = (unsigned long long)cp.count * pk.n;
    const unsigned long long need = (items + 31) / 32;
    if (grid > need) grid = need ? (u32)need : 1;
    CU_OK(s->recs2.reserve(s->lx_event_hint));
    const DevLxEvents evq{reinterpret_cast<LxEvent *>(s->recs2.p), s->d_ctl + 7, s->recs2.cap};
    k_mcclellan_pack<WIDE16><<<grid, MCC_THREADS, dyn, s->stream>>>(im->dev, pk, cp, evq, s->d_ctl + 6);
    CU_OK(cudaGetLastError());
    s->stats.kernel_launches++;
    return HS_SUCCESS;
}

// Full-Rose database of outfix engines (roseBlockExec with outfixes only, src/rose/block.c:345-422, catchup.c:612-672):
// every engine steps the whole batch (a block shorter than the engine's NFA.minWidth cannot match it), all of them append
// to ONE event queue, and one report pass runs the Rose report programs; the finalize sorts by (buffer, offset, id) and
// applies the dedupe keys, which is what the catch-up queue's offset order is for.  McClellan engines whose tables fit
// shared memory together share a launch (k_mcclellan_pack: work item = (buffer, engine)); option "pack" = 0 turns that off.
template <int D>
hs_error_t launch_dfa_pack(hs_gpu_scratch *s, const DbImage *im, const DevDfaPack &pk, const DevCorpus &cp) {
    const size_t dyn = dfapack_smem_bytes(pk);
    CU_OK(ensure_smem_optin(s, (const void *)k_dfa_pack<D>, dyn));
    u32 grid = s->opt_grid > 0 ? (u32)s->opt_grid : (u32)s->sm_count;
    const u32 need = (cp.count + 31) / 32;
    if (grid > need) grid = need ? need : 1;
    CU_OK(s->recs2.reserve(s->lx_event_hint));
    const DevLxEvents evq{reinterpret_cast<LxEvent *>(s->recs2.p), s->d_ctl + 7, s->recs2.cap};
    k_dfa_pack<D><<<grid, DP_THREADS, dyn, s->stream>>>(im->dev, pk, cp, evq, s->d_ctl + 6);
    CU_OK(cudaGetLastError());
    s->stats.kernel_launches++;
    return HS_SUCCESS;
}

hs_error_t launch_outfixes(hs_gpu_scratch *s, const DbImage *im, const DevCorpus &cp, const DevArena &ar) {
    bool first = true;
    const bool fused = s->opt_pack >= 2;
    if (fused)
        for (const DbImage::DfaPackImage &pk : im->packs) {
            if (!first) CU_OK(cudaMemsetAsync(s->d_ctl + 6, 0, sizeof(unsigned long long), s->stream));  // the work counter
            first = false;
            const hs_error_t rv = pk.D == 2 ? launch_dfa_pack<2>(s, im, pk.dev, cp) : pk.D == 4 ? launch_dfa_pack<4>(s, im, pk.dev, cp)
                                                                                                : launch_dfa_pack<8>(s, im, pk.dev, cp);
            if (rv != HS_SUCCESS) return rv;
        }
    for (size_t qi = 0; qi < im->subs.size();) {
        const DbImage *sub = im->subs[qi];
        if (fused && qi < im->packed.size() && im->packed[qi]) { qi++; continue; }
        if (!first) CU_OK(cudaMemsetAsync(s->d_ctl + 6, 0, sizeof(unsigned long long), s->stream));  // the work counter
        first = false;
        auto packable = [](const DbImage *e) {
            return e->view.kind == ENG_MCCLELLAN && e->mcc.in_smem && !e->mcc.has_sherman;
        };
        if (s->opt_pack && packable(sub)) {
            DevMccPack pk{};
            size_t tables = 0;
            size_t k = qi;
            while (k < im->subs.size() && pk.n < (u32)MCC_PACK_MAX && packable(im->subs[k]) &&
                   im->subs[k]->mcc.wide16 == sub->mcc.wide16 && !(fused && im->packed[k])) {
                const size_t tb = round_up((size_t)im->subs[k]->mcc.tab_bytes + 8, 16);
                if ((size_t)(pk.n + 1) * MCC_RM_BYTES + tables + tb > MCC_SMEM_MAX) break;
                pk.m[pk.n] = im->subs[k]->mcc;
                pk.tab_off[pk.n] = (u32)tables;  // relative to the end of the symbol tables; fixed up below
                tables += tb;
                pk.n++;
                k++;
            }
            if (pk.n >= 2) {
                for (u32 e = 0; e < pk.n; e++) pk.tab_off[e] += pk.n * (u32)MCC_RM_BYTES;
                const size_t dyn = (size_t)pk.n * MCC_RM_BYTES + tables;
                const hs_error_t rv = sub->mcc.wide16 ? launch_mcclellan_pack<true>(s, im, pk, dyn, cp)
                                                       : launch_mcclellan_pack<false>(s, im, pk, dyn, cp);
                if (rv != HS_SUCCESS) return rv;
                qi = k;
                continue;
            }
        }
        const hs_error_t rv = sub->view.kind == ENG_LIMEX ? launch_limex(s, sub, cp, ar, false) : launch_mcclellan(s, sub, cp, ar, false);
        if (rv != HS_SUCCESS) return rv;
        qi++;
    }
    const DevLxEvents evq{reinterpret_cast<LxEvent *>(s->recs2.p), s->d_ctl + 7, s->recs2.cap};
    k_limex_report<<<(u32)s->sm_count, LX_THREADS, 0, s->stream>>>(im->dev, cp, ar, evq);
    CU_OK(cudaGetLastError());
    s->stats.kernel_launches++;
    return HS_SUCCESS;
}

// Ordered mode: k_turbo / k_turbo2 append position-ordered records to per-warp segments, k_compact gathers them; no
// sort.  A TRAIN of n members (resident corpora, or the chunks of one host batch) is queued back to back on the
// scratch's stream -- member i may wait for ready[i] (its H2D copy on the copy stream) -- and ends in ONE host round
// trip; the members' records land in one array, buffer indices continuing from member to member (buf_base).
// Returns HS_INSUFFICIENT_SPACE when the general path must be used instead.
struct TrainMember {
    DevCorpus cp;
    u32 buf_base;
    cudaEvent_t ready;  // nullptr: no wait
};

hs_error_t scan_train_ordered(hs_gpu_scratch *s, DbImage *im, const TrainMember *mem, u32 n) {
    hs_gpu_stats_t &st = s->stats;
    const bool v2 = use_turbo2(s, im->dev);
    const u32 cta_warps = v2 ? turbo2_warps<true>(s, im->dev) : (u32)turbo_warps<true>();
    if (!cta_warps || !n) return HS_INSUFFICIENT_SPACE;
    u32 maxW = 0;
    for (u32 i = 0; i < n; i++) maxW = std::max(maxW, turbo_grid(s, mem[i].cp, cta_warps) * cta_warps);
    CU_OK(s->seg_count.reserve(maxW));
    // control blocks: two halves of train_cap blocks; a train uses one half, its k_compacts clear the other
    if (n > s->train_cap) {
        const u32 cap = n + n / 2 + 2;
        CU_OK(s->train_ctl.reserve((size_t)2 * cap * 8));
        if (!s->h_train_ctl.resize((size_t)cap * 8)) return HS_NOMEM;
        while (s->tev.size() < 3 * (size_t)cap) {
            cudaEvent_t e = nullptr;
            CU_OK(cudaEventCreate(&e));
            s->tev.push_back(e);
        }
        s->train_cap = cap;
        s->ctl_dirty = true;
    }
    for (int attempt = 0;; attempt++) {
        const u64 seg_cap = s->seg_cap_hint, out_cap = s->out_hint;
        if (seg_cap * maxW > (1ull << 28)) return HS_INSUFFICIENT_SPACE;  // > 4 GiB of segments: not this mode
        CU_OK(s->recs.reserve(seg_cap * maxW));
        CU_OK(s->recs2.reserve(out_cap));
        CU_OK(s->d_matches.reserve(out_cap));
        if (s->ctl_dirty) {
            CU_OK(cudaMemsetAsync(s->train_ctl.p, 0, (size_t)2 * s->train_cap * 64, s->stream));
            s->ctl_dirty = false;
        }
        unsigned long long *half = s->train_ctl.p + (size_t)s->ctl_flip * s->train_cap * 8;
        unsigned long long *other = s->train_ctl.p + (size_t)(s->ctl_flip ^ 1) * s->train_cap * 8;
        s->ctl_flip ^= 1;
        s->ctl_dirty = true;  // until this train's k_compacts are known to have run
        for (u32 i = 0; i < n; i++) {
            const DevCorpus &cp = mem[i].cp;
            unsigned long long *ctl = half + (size_t)i * 8;
            if (mem[i].ready) CU_OK(cudaStreamWaitEvent(s->stream, mem[i].ready, 0));
            CU_OK(cudaEventRecord(s->tev[3 * i], s->stream));  // the kernel's interval starts after the wait
            const u32 grid = turbo_grid(s, cp, cta_warps), W = grid * cta_warps;
            DevArena ar{s->recs.p, ctl + 0, ctl + 1, reinterpret_cast<u32 *>(ctl + 3), 0, s->seg_count.p, (u32)seg_cap, s->op_counts.p};
            hs_error_t rv;
            if (v2) rv = launch_turbo2_any<true>(s, im->dev, cp, ar, grid, cta_warps);
            else
            switch (im->dev.tK) {
            case 1: rv = launch_turbo<1, true>(s, im->dev, cp, ar); break;
            case 2: rv = launch_turbo<2, true>(s, im->dev, cp, ar); break;
            case 3: rv = launch_turbo<3, true>(s, im->dev, cp, ar); break;
            default: rv = launch_turbo<4, true>(s, im->dev, cp, ar); break;
            }
            if (rv != HS_SUCCESS) return rv;
            CU_OK(cudaEventRecord(s->tev[3 * i + 1], s->stream));
            k_compact<<<grid, 256, 0, s->stream>>>(s->recs.p, s->seg_count.p, (u32)seg_cap, W, s->recs2.p, s->d_matches.p, ctl,
                                                   other + (size_t)i * 8, i ? ctl - 8 : nullptr, mem[i].buf_base, out_cap);
            CU_OK(cudaGetLastError());
            CU_OK(cudaEventRecord(s->tev[3 * i + 2], s->stream));
            st.scan_launches++;
            st.kernel_launches += 2;
        }
        // device work ends here; the control-block read-back that follows is host synchronisation
        CU_OK(cudaMemcpyAsync(s->h_train_ctl.data(), half, (size_t)n * 64, cudaMemcpyDeviceToHost, s->stream));
        CU_OK(cudaStreamSynchronize(s->stream));
        s->ctl_dirty = false;
        u64 worst_seg = 0;
        u32 err = 0;
        for (u32 i = 0; i < n; i++) {
            err |= (u32)s->h_train_ctl[(size_t)i * 8 + 3];
            worst_seg = std::max<u64>(worst_seg, s->h_train_ctl[(size_t)i * 8 + 5]);
        }
        if (err & ERR_COLLECTOR) { im->ordered_ok = false; return HS_INSUFFICIENT_SPACE; }
        if (err) return HS_UNKNOWN_ERROR;
        const u64 written = s->h_train_ctl[(size_t)(n - 1) * 8 + 6];
        if (worst_seg <= seg_cap && written <= out_cap) break;
        if (attempt >= 3) return HS_INSUFFICIENT_SPACE;
        st.arena_retries++;
        if (worst_seg > seg_cap) s->seg_cap_hint = worst_seg + worst_seg / 2 + 16;
        if (written > out_cap) s->out_hint = written + written / 4 + 1024;
    }
    st.scan_ms = st.sort_ms = 0;
    st.candidates = st.confirms = 0;
    for (u32 i = 0; i < n; i++) {
        float a = 0, b = 0;
        cudaEventElapsedTime(&a, s->tev[3 * i], s->tev[3 * i + 1]);
        cudaEventElapsedTime(&b, s->tev[3 * i + 1], s->tev[3 * i + 2]);
        st.scan_ms += a;
        st.sort_ms += b;
        st.candidates += s->h_train_ctl[(size_t)i * 8 + 1];
        st.confirms += s->h_train_ctl[(size_t)i * 8 + 2];
    }
    cudaEventElapsedTime(&st.train_ms, s->tev[0], s->tev[3 * (n - 1) + 2]);
    st.kernel_ms = st.scan_ms;  // the k_turbo launches of the last (successful) attempt
    st.front_end = v2 ? 3u : 2u;
    st.raw_matches = st.matches = s->n_matches = s->h_train_ctl[(size_t)(n - 1) * 8 + 6];
    return HS_SUCCESS;
}

hs_error_t scan_device_ordered(hs_gpu_scratch *s, DbImage *im, const DevCorpus &cp) {
    const TrainMember m{cp, 0, nullptr};
    return scan_train_ordered(s, im, &m, 1);
}

inline bool ordered_eligible(const hs_gpu_scratch *s, const DbImage *im) {
    return im->dev.tK && s->opt_frontend != 1 && im->ordered_ok && im->turbo_tabs.all_simple && !im->view.dkey_count &&
           !im->view.ekey_count && s->opt_ordered;
}

// Device pass over a resident corpus: scan (+ retry on arena growth), sort, dedupe, convert.
hs_error_t scan_device(hs_gpu_scratch *s, DbImage *im, const hs_gpu_corpus &corpus) {
    hs_gpu_stats_t &st = s->stats;
    st.bytes = corpus.total;
    st.buffers = corpus.count;
    st.scan_launches = 0;
    st.kernel_launches = 0;
    st.kernel_ms = 0;
    st.arena_retries = 0;
    s->n_matches = 0;
    s->fetched = false;
    s->last_image = im;
    if (!corpus.total || !corpus.count) {
        st.matches = st.raw_matches = st.candidates = st.confirms = 0;
        st.scan_ms = st.sort_ms = st.kernel_ms = 0;
        return HS_SUCCESS;
    }
    DevCorpus cp = corpus.dev();
    if (s->stream_state_count) {
        // stream mode of a DFA engine: entry states up, exit states back behind the scan (fetch_matches waits for the stream)
        if (s->stream_state_count != corpus.count || (im->view.kind != ENG_MCCLELLAN && im->view.kind != ENG_SHENG)) return HS_INVALID;
        CU_OK(s->d_stream_state.reserve(2 * (size_t)corpus.count));
        CU_OK(cudaMemcpyAsync(s->d_stream_state.p, s->h_stream_state.data(), 2 * (size_t)corpus.count * sizeof(u32), cudaMemcpyHostToDevice, s->stream));
        cp.init_state = s->d_stream_state.p;
        cp.final_state = s->d_stream_state.p + corpus.count;
    }
    CU_OK(cudaEventRecord(s->ev[1], s->stream));  // what was queued before this point is the staging (h2d_ms = ev0 -> ev1)
    if (im->view.vectored) {
        // hs_scan_vector on an HS_MODE_VECTORED database (src/runtime.c:1106-1174): block i continues block i - 1 with
        // history carried over, so what is scanned is the concatenation -- which is how the corpus already lies in HBM.
        // Offsets and buffer lengths are 32-bit on the device: a stream of 4 GiB or more is refused.
        if (corpus.total >= 0xffffffffull - BACK_PAD) return HS_INVALID;
        if (!corpus.stream_ready) {
            const size_t ncells = (size_t)((corpus.total + CELL_BYTES - 1) / CELL_BYTES) + 1;
            CU_OK(corpus.stream_off.reserve(2));
            CU_OK(corpus.stream_cell.reserve(ncells));
            CU_OK(cudaMemsetAsync(corpus.stream_cell.p, 0, ncells * sizeof(u32), s->stream));
            k_stream_offsets<<<1, 1, 0, s->stream>>>(corpus.stream_off.p, corpus.total);
            CU_OK(cudaGetLastError());
            corpus.stream_ready = true;
        }
        cp = corpus.dev_stream();
        st.buffers = 1;
    }
    const bool has_engines = im->view.kind == ENG_LIMEX || im->view.kind == ENG_MCCLELLAN || im->view.kind == ENG_SHENG || !im->subs.empty();
    if (has_engines && s->opt_order && cp.count > 32) {
        if (!corpus.order_ready) {
            CU_OK(corpus.order.reserve(cp.count));
            CU_OK(s->order_hist.reserve(ORDER_BUCKETS));
            CU_OK(cudaMemsetAsync(s->order_hist.p, 0, ORDER_BUCKETS * sizeof(u32), s->stream));
            const unsigned g = (cp.count + 255) / 256;
            k_order_hist<<<g, 256, 0, s->stream>>>(cp.off, cp.count, s->order_hist.p);
            k_order_scan<<<1, 1024, 0, s->stream>>>(s->order_hist.p);
            k_order_scatter<<<g, 256, 0, s->stream>>>(cp.off, cp.count, s->order_hist.p, corpus.order.p);
            CU_OK(cudaGetLastError());
            corpus.order_ready = true;
        }
        cp.order = corpus.order.p;
    }
    // ---- ordered fast path: turbo front end, every program pre-decoded, no dedupe / exhaustion keys ----
    if (ordered_eligible(s, im)) {
        hs_error_t orv = scan_device_ordered(s, im, cp);
        if (orv != HS_INSUFFICIENT_SPACE) return orv;
        // fell out of the ordered mode's envelope (collector or segment bounds): use the general path below
    }
    CU_OK(s->recs.reserve(s->arena_hint));
    CU_OK(s->recs2.reserve(SMALL_FIN_MAX));     // before the launch: the event queue of the NFA/DFA kernels lives here too
    CU_OK(s->d_matches.reserve(SMALL_FIN_MAX));
    u64 raw = 0;
    bool small_done = false;
    for (int attempt = 0;; attempt++) {
        CU_OK(cudaMemsetAsync(s->d_ctl, 0, 8 * sizeof(unsigned long long), s->stream));
        DevArena ar{s->recs.p, s->d_ctl + 0, s->d_ctl + 1, reinterpret_cast<u32 *>(s->d_ctl + 3), s->recs.cap, nullptr, 0, s->op_counts.p};
        hs_error_t rv;
        CU_OK(cudaEventRecord(s->ev[6], s->stream));
        const bool use_turbo = im->dev.tK && s->opt_frontend != 1;
        const u32 t2w = use_turbo && use_turbo2(s, im->dev) ? turbo2_warps<false>(s, im->dev) : 0;
        if (t2w) {
            rv = launch_turbo2_any<false>(s, im->dev, cp, ar, turbo_grid(s, cp, t2w), t2w);
        } else if (use_turbo) {
            switch (im->dev.tK) {
            case 1: rv = launch_turbo<1, false>(s, im->dev, cp, ar); break;
            case 2: rv = launch_turbo<2, false>(s, im->dev, cp, ar); break;
            case 3: rv = launch_turbo<3, false>(s, im->dev, cp, ar); break;
            default: rv = launch_turbo<4, false>(s, im->dev, cp, ar); break;
            }
        } else
        switch (im->view.kind) {
        case ENG_FDR: rv = launch_scan<FdrNative, false>(s, im->dev, cp, ar); break;
        case ENG_TEDDY: rv = launch_scan<TeddyNative, false>(s, im->dev, cp, ar); break;
        case ENG_LIMEX: rv = launch_limex(s, im, cp, ar); break;
        case ENG_MCCLELLAN: case ENG_SHENG: rv = launch_mcclellan(s, im, cp, ar); break;
        case ENG_MULTI: rv = launch_outfixes(s, im, cp, ar); break;
        default: rv = launch_scan<Noodle, true>(s, im->dev, cp, ar); break;
        }
        if (rv != HS_SUCCESS) return rv;
        if (im->view.kind != ENG_MULTI && !im->subs.empty()) {  // mixed database: the engines add their matches to the same arena
            rv = launch_outfixes(s, im, cp, ar);
            if (rv != HS_SUCCESS) return rv;
        }
        st.front_end = !im->subs.empty() ? 6u : t2w ? 3u : use_turbo ? 2u : im->view.kind == ENG_LIMEX ? 4u
                     : (im->view.kind == ENG_MCCLELLAN || im->view.kind == ENG_SHENG) ? 5u : im->view.kind == ENG_MULTI ? 6u : 1u;
        CU_OK(cudaEventRecord(s->ev[7], s->stream));
        st.scan_launches++;
        st.kernel_launches++;
        // Few records are the rule for regex databases: the single-CTA finalize goes in right behind the scan and
        // decides on the device whether it applies, so the whole scan costs ONE host round trip.
        CU_OK(cudaEventRecord(s->ev[2], s->stream));
        // sized by the previous scan's count (fewer threads = fewer merge rounds and cheaper barriers)
        if (s->last_raw <= 256 * SMALL_FIN_ITEMS)
            CU_OK(launch_small_finalize<256>(s, s->recs.p, s->recs2.p, s->d_matches.p, s->d_ctl + 4, s->d_ctl + 0, s->recs.cap));
        else if (s->last_raw <= 512 * SMALL_FIN_ITEMS)
            CU_OK(launch_small_finalize<512>(s, s->recs.p, s->recs2.p, s->d_matches.p, s->d_ctl + 4, s->d_ctl + 0, s->recs.cap));
        else
            CU_OK(launch_small_finalize<1024>(s, s->recs.p, s->recs2.p, s->d_matches.p, s->d_ctl + 4, s->d_ctl + 0, s->recs.cap));
        CU_OK(cudaEventRecord(s->ev[3], s->stream));
        CU_OK(cudaMemcpyAsync(s->h_ctl, s->d_ctl, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s->stream));
        CU_OK(cudaStreamSynchronize(s->stream));
        if ((u32)s->h_ctl[3]) {
            if (getenv("HS_GPU_DEBUG")) fprintf(stderr, "hs_gpu: device error flags 0x%x\n", (u32)s->h_ctl[3]);
            return HS_UNKNOWN_ERROR;
        }
        raw = s->h_ctl[0];
        if ((im->view.kind == ENG_LIMEX || im->view.kind == ENG_MCCLELLAN || im->view.kind == ENG_SHENG || !im->subs.empty()) && s->h_ctl[7] > s->recs2.cap) {  // event queue overflow: grow, scan again
            if (attempt >= 3) return HS_NOMEM;
            st.arena_retries++;
            s->lx_event_hint = s->h_ctl[7] + s->h_ctl[7] / 8 + 1024;
            continue;
        }
        s->last_raw = raw;
        if (raw <= s->recs.cap) { small_done = s->h_ctl[4] != ~0ull; break; }
        if (attempt >= 3) return HS_NOMEM;
        st.arena_retries++;
        s->arena_hint = raw + raw / 8 + 1024;
        CU_OK(s->recs.reserve(s->arena_hint));
    }
    if (!small_done) CU_OK(cudaEventRecord(s->ev[2], s->stream));
    st.raw_matches = raw;
    st.candidates = s->h_ctl[1];
    st.confirms = s->h_ctl[2];
    u64 n_out = 0;
    if (small_done) {
        n_out = s->h_ctl[4];
        st.kernel_launches += 1;
    } else if (raw) {
        size_t tmp1 = 0, tmp2 = 0;
        CU_OK(s->recs2.reserve(raw));
        CU_OK(s->d_matches.reserve(raw));
        cub::DeviceMergeSort::SortKeys((void *)nullptr, tmp1, s->recs.p, (long long)raw, RecLess(), s->stream);
        CU_OK(s->keep.reserve(raw));
        cub::DeviceSelect::Flagged((void *)nullptr, tmp2, s->recs.p, s->keep.p, s->recs2.p,
                                   (unsigned long long *)nullptr, (long long)raw, s->stream);
        CU_OK(s->cub_tmp.reserve(std::max(tmp1, tmp2)));
        size_t tb = s->cub_tmp.bytes();
        CU_OK(cub::DeviceMergeSort::SortKeys(s->cub_tmp.p, tb, s->recs.p, (long long)raw, RecLess(), s->stream));
        tb = s->cub_tmp.bytes();
        k_flag_dups<<<(unsigned)((raw + 255) / 256), 256, 0, s->stream>>>(s->recs.p, s->keep.p, raw);
        CU_OK(cudaGetLastError());
        CU_OK(cub::DeviceSelect::Flagged(s->cub_tmp.p, tb, s->recs.p, s->keep.p, s->recs2.p, s->d_ctl + 4,
                                         (long long)raw, s->stream));
        CU_OK(cudaMemcpyAsync(s->h_ctl + 4, s->d_ctl + 4, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s->stream));
        CU_OK(cudaStreamSynchronize(s->stream));
        n_out = s->h_ctl[4];
        k_convert<<<(unsigned)((n_out + 255) / 256), 256, 0, s->stream>>>(s->recs2.p, s->d_matches.p, n_out);
        CU_OK(cudaGetLastError());
        st.kernel_launches += 2;  // k_flag_dups + k_convert (CUB's sort/select kernels are not counted)
    }
    if (im->view.ekey_count && n_out) {
        // records (sorted, deduped) are in recs2: resolve the exhaustion keys there, compact into recs, convert
        if (n_out >= 0xffffffffull) return HS_NOMEM;
        u32 H = 1024;
        while ((u64)H < 2 * n_out) H <<= 1;
        CU_OK(s->ekey_keys.reserve(H));
        CU_OK(s->ekey_first.reserve(H));
        CU_OK(s->keep.reserve(n_out));
        CU_OK(s->recs.reserve(n_out));
        CU_OK(s->d_matches.reserve(n_out));
        CU_OK(cudaMemsetAsync(s->ekey_keys.p, 0xff, (size_t)H * 8, s->stream));
        CU_OK(cudaMemsetAsync(s->ekey_first.p, 0xff, (size_t)H * 4, s->stream));
        const unsigned g = (unsigned)((n_out + 255) / 256);
        k_ekey_insert<<<g, 256, 0, s->stream>>>(s->recs2.p, n_out, s->ekey_keys.p, s->ekey_first.p, H - 1);
        k_ekey_filter<<<g, 256, 0, s->stream>>>(s->recs2.p, n_out, s->ekey_keys.p, s->ekey_first.p, H - 1, s->keep.p);
        CU_OK(cudaGetLastError());
        size_t tmp = 0;
        cub::DeviceSelect::Flagged((void *)nullptr, tmp, s->recs2.p, s->keep.p, s->recs.p, (unsigned long long *)nullptr,
                                   (long long)n_out, s->stream);
        CU_OK(s->cub_tmp.reserve(tmp));
        size_t tb = s->cub_tmp.bytes();
        CU_OK(cub::DeviceSelect::Flagged(s->cub_tmp.p, tb, s->recs2.p, s->keep.p, s->recs.p, s->d_ctl + 4, (long long)n_out, s->stream));
        CU_OK(cudaMemcpyAsync(s->h_ctl + 4, s->d_ctl + 4, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s->stream));
        CU_OK(cudaStreamSynchronize(s->stream));
        n_out = s->h_ctl[4];
        if (n_out) {
            k_convert<<<(unsigned)((n_out + 255) / 256), 256, 0, s->stream>>>(s->recs.p, s->d_matches.p, n_out);
            CU_OK(cudaGetLastError());
        }
        st.kernel_launches += 3;
        CU_OK(cudaEventRecord(s->ev[3], s->stream));
        CU_OK(cudaEventSynchronize(s->ev[3]));
    } else if (!small_done) {
        CU_OK(cudaEventRecord(s->ev[3], s->stream));
        CU_OK(cudaEventSynchronize(s->ev[3]));
    }
    cudaEventElapsedTime(&st.scan_ms, s->ev[1], s->ev[2]);
    cudaEventElapsedTime(&st.kernel_ms, s->ev[6], s->ev[7]);  // the last (successful) k_scan launch alone
    cudaEventElapsedTime(&st.sort_ms, s->ev[2], s->ev[3]);
    s->n_matches = n_out;
    st.matches = n_out;
    return HS_SUCCESS;
}

hs_error_t fetch_matches(hs_gpu_scratch *s, DbImage *im) {
    hs_gpu_stats_t &st = s->stats;
    st.d2h_ms = 0;
    if (s->fetched) return HS_SUCCESS;
    const u64 n = s->n_matches;
    if (n > s->h_matches_cap) {
        if (s->h_matches) cudaFreeHost(s->h_matches);
        s->h_matches = nullptr;
        s->h_matches_cap = 0;
        const size_t want = n + n / 4 + 1024;
        // room for either representation (both are 16 bytes)
        CU_OK(cudaHostAlloc((void **)&s->h_matches, want * sizeof(hs_gpu_match_t), cudaHostAllocDefault));
        s->h_matches_cap = want;
    }
    if (n) {
        CU_OK(cudaEventRecord(s->ev[4], s->stream));
        (void)im;
        CU_OK(cudaMemcpyAsync(s->h_matches, s->d_matches.p, n * sizeof(hs_gpu_match_t), cudaMemcpyDeviceToHost, s->stream));
        CU_OK(cudaEventRecord(s->ev[5], s->stream));
        CU_OK(cudaEventSynchronize(s->ev[5]));
        cudaEventElapsedTime(&st.d2h_ms, s->ev[4], s->ev[5]);
    }
    s->fetched = true;
    return HS_SUCCESS;
}

hs_error_t ensure_device(hs_gpu_scratch *s) {
    int cur = -1;
    if (cudaGetDevice(&cur) != cudaSuccess) return HS_ARCH_ERROR;
    if (cur != s->device && cudaSetDevice(s->device) != cudaSuccess) return HS_ARCH_ERROR;
    return HS_SUCCESS;
}

struct InUseGuard {
    hs_gpu_scratch *s;
    bool ok;
    explicit InUseGuard(hs_gpu_scratch *s_) : s(s_) { int z = 0; ok = s->in_use.compare_exchange_strong(z, 1); }
    ~InUseGuard() { if (ok) s->in_use.store(0); }
};

// Scattered buffers [first, last) -> one contiguous run at dst (buffer q lands at dst + off[q - base] - block_at).  A block
// of several MiB made of packet-sized pieces is copied by a few threads: one core's memcpy of 1.5 KB pieces (~7 GB/s)
// is far below what the PCIe copy behind it takes.
void gather_block(u8 *dst, const char *const *data, const unsigned int *length, const u64 *off, unsigned base, unsigned first,
                  unsigned last, u64 block_at) {
    auto part = [&](unsigned a, unsigned b) {
        for (unsigned q = a; q < b; q++)
            if (length[q]) std::memcpy(dst + (off[q - base] - block_at), data[q], length[q]);
    };
    const unsigned n = last - first;
    unsigned threads = std::min<unsigned>(8, std::max<unsigned>(1, std::thread::hardware_concurrency() / 2));
    if (n < 512 || threads < 2) { part(first, last); return; }
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < threads; t++) pool.emplace_back(part, first + (u64)n * t / threads, first + (u64)n * (t + 1) / threads);
    part(first, first + n / threads);
    for (std::thread &t : pool) t.join();
}

// Host buffers -> the staging corpus in HBM.  The buffer list is walked in blocks (<= 4096 buffers or about
// 8 MiB): a block whose buffers sit back to back in page-locked memory (hs_gpu_host_alloc) is copied straight
// from the caller's memory, any other block is gathered into the scratch's pinned staging area first; either
// way its cudaMemcpyAsync is queued as soon as the block has been walked, so the PCIe copy starts after a few
// microseconds of host work and the rest of the walk (and any gathering) runs underneath it.
hs_error_t stage_host_buffers(hs_gpu_scratch *s, const char *const *data, const unsigned int *length, unsigned count) {
    hs_gpu_corpus &cp = s->staging;
    if (!s->h_off.resize((size_t)count + 1)) return HS_NOMEM;
    u64 total = 0;
    for (unsigned i = 0; i < count; i++) total += length[i];
    cp.total = total;
    cp.count = count;
    cp.order_ready = false;
    cp.stream_ready = false;
    CU_OK(cp.text.reserve(FRONT_PAD + total + BACK_PAD));
    CU_OK(cp.off.reserve((size_t)count + 1));
    CU_OK(cudaEventRecord(s->ev[0], s->stream));
    // zero the pads (the front pad is read by look-behind, the back pad by the last rows)
    CU_OK(cudaMemsetAsync(cp.text.p, 0, FRONT_PAD, s->stream));
    CU_OK(cudaMemsetAsync(cp.text.p + FRONT_PAD + total, 0, BACK_PAD, s->stream));
    constexpr unsigned BLOCK_BUFS = 4096;
    constexpr u64 BLOCK_BYTES = 8u << 20;
    u64 pos = 0;
    for (unsigned first = 0; first < count;) {
        const u64 block_at = pos;
        const char *run = nullptr;  // address the next non-empty buffer must have for the block to be one run
        const char *run_begin = nullptr;
        bool contiguous = true;
        unsigned i = first;
        for (; i < count && i - first < BLOCK_BUFS && pos - block_at < BLOCK_BYTES; i++) {
            const unsigned len = length[i];
            s->h_off[i] = pos;
            if (len) {
                const char *p = data[i];
                if (!p) return HS_INVALID;
                if (!run_begin) run_begin = p; else if (p != run) contiguous = false;
                run = p + len;
                pos += len;
            }
        }
        const u64 bytes = pos - block_at;
        if (bytes) {
            const u8 *src = nullptr;
            if (contiguous) {
                cudaPointerAttributes at{};
                if (cudaPointerGetAttributes(&at, run_begin) == cudaSuccess && at.type == cudaMemoryTypeHost)
                    src = reinterpret_cast<const u8 *>(run_begin);
                else
                    cudaGetLastError();
            }
            if (!src) {
                if (total > s->h_stage_cap) {
                    // earlier blocks of this call never used the staging area (it would have been sized then)
                    CU_OK(cudaStreamSynchronize(s->stream));
                    if (s->h_stage) cudaFreeHost(s->h_stage);
                    s->h_stage = nullptr;
                    s->h_stage_cap = 0;
                    const size_t want = total + total / 4 + 4096;
                    CU_OK(cudaHostAlloc((void **)&s->h_stage, want, cudaHostAllocDefault));
                    s->h_stage_cap = want;
                }
                s->h_off[i] = pos;
                gather_block(s->h_stage + block_at, data, length, s->h_off.data(), 0, first, i, block_at);
                src = s->h_stage + block_at;
            }
            CU_OK(cudaMemcpyAsync(cp.text.p + FRONT_PAD + block_at, src, bytes, cudaMemcpyHostToDevice, s->stream));
        }
        first = i;
    }
    s->h_off[count] = total;
    CU_OK(cudaMemcpyAsync(cp.off.p, s->h_off.data(), ((size_t)count + 1) * 8, cudaMemcpyHostToDevice, s->stream));
    // the position -> buffer index is built while the text is on its way
    if (!build_cell_first(s->h_off.data(), count, total, s->h_cell_first, cp.nstrips)) return HS_NOMEM;
    CU_OK(cp.cell_first.reserve(s->h_cell_first.size()));
    CU_OK(cudaMemcpyAsync(cp.cell_first.p, s->h_cell_first.data(), s->h_cell_first.size() * 4,
                          cudaMemcpyHostToDevice, s->stream));
    return HS_SUCCESS;
}

// Pipelined form of the staging for databases the sort-free path serves: the batch is cut at buffer boundaries into a
// few chunks, each laid out as a corpus of its own (1 KiB-aligned, zero pads either side, offsets rebased) and copied
// on the scratch's COPY stream; `ready` events let the scan stream start chunk k while chunk k + 1 is still on the
// bus (the filters are position-local and buffers independent, so chunks need no halo).
hs_error_t stage_host_chunks(hs_gpu_scratch *s, const char *const *data, const unsigned int *length, unsigned count,
                             u64 total, std::vector<TrainMember> &mem) {
    hs_gpu_corpus &cp = s->staging;
    // Chunk sizes halve (1/2, 1/4, 1/8, 1/8 of the batch, none under 4 MiB): every launch has a fixed cost (148 CTAs
    // filling their tables), so few chunks; and the LAST one is small, because its scan is the only one nothing hides.
    constexpr u64 CHUNK_MIN = 4u << 20;
    constexpr unsigned CHUNKS_MAX = 4;
    struct Chunk { unsigned first, last; u64 bytes, text_at; size_t off_at, cell_at; u32 nstrips; };
    std::vector<Chunk> ch;
    {
        unsigned first = 0;
        u64 acc = 0, left = total;
        u64 want = std::max<u64>(CHUNK_MIN, total / 2);
        for (unsigned i = 0; i < count; i++) {
            acc += length[i];
            if (acc >= want || i + 1 == count) {
                ch.push_back(Chunk{first, i + 1, acc, 0, 0, 0, 0});
                first = i + 1;
                left -= acc;
                acc = 0;
                want = ch.size() + 1 >= CHUNKS_MAX ? left : std::max<u64>(CHUNK_MIN, left / 2);
                if (left < 2 * CHUNK_MIN) want = left;
                if (!left) want = ~0ull;  // only empty buffers remain: they ride in one last (empty) chunk
            }
        }
    }
    const unsigned nch = (unsigned)ch.size();
    u64 at = 0;
    size_t off_n = 0, cell_n = 0;
    for (Chunk &c : ch) {
        c.text_at = at + FRONT_PAD;   // `at` is 1 KiB-aligned and so is FRONT_PAD
        at = round_up(c.text_at + c.bytes + BACK_PAD, 1024);
        c.off_at = off_n;
        off_n += (size_t)(c.last - c.first) + 1;
        c.cell_at = cell_n;
        cell_n += (size_t)((c.bytes + CELL_BYTES - 1) / CELL_BYTES) + 1;
    }
    CU_OK(cp.text.reserve(at + FRONT_PAD));
    CU_OK(cp.off.reserve(off_n));
    CU_OK(cp.cell_first.reserve(cell_n));
    if (!s->h_off.resize(off_n) || !s->h_cell_first.resize(cell_n)) return HS_NOMEM;
    while (s->chunk_ready.size() < nch) {
        cudaEvent_t e = nullptr;
        CU_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        s->chunk_ready.push_back(e);
    }
    cp.total = total;
    cp.count = count;
    cp.order_ready = false;
    cp.stream_ready = false;
    // the copy stream must not overtake the previous call's scan of this staging area
    CU_OK(cudaEventRecord(s->ev[1], s->stream));
    CU_OK(cudaStreamWaitEvent(s->copy_stream, s->ev[1], 0));
    CU_OK(cudaEventRecord(s->ev[0], s->copy_stream));
    constexpr unsigned BLOCK_BUFS = 4096;
    constexpr u64 BLOCK_BYTES = 8u << 20;
    u64 host_pos = 0;  // running position in the pinned staging area (gathered blocks only)
    mem.clear();
    // The zero gaps first (front pad, between chunks, back pad): they depend on the chunk bounds alone.  Then the text,
    // block by block as the host walks the buffers, so the bus is busy from the first few microseconds on; a chunk's
    // offsets and position -> buffer table follow its text, and the chunk is ready when they have landed.
    for (unsigned k = 0; k <= nch; k++) {
        const u64 gap_from = k ? ch[k - 1].text_at + ch[k - 1].bytes : 0;
        const u64 gap_to = k < nch ? ch[k].text_at : gap_from + BACK_PAD;
        CU_OK(cudaMemsetAsync(cp.text.p + gap_from, 0, gap_to - gap_from, s->copy_stream));
    }
    for (unsigned k = 0; k < nch; k++) {
        Chunk &c = ch[k];
        u64 *off = s->h_off.data() + c.off_at;
        u64 pos = 0;
        for (unsigned first = c.first; first < c.last;) {
            const u64 block_at = pos;
            const char *run = nullptr, *run_begin = nullptr;
            bool contiguous = true;
            unsigned i = first;
            for (; i < c.last && i - first < BLOCK_BUFS && pos - block_at < BLOCK_BYTES; i++) {
                const unsigned len = length[i];
                off[i - c.first] = pos;
                if (len) {
                    const char *p = data[i];
                    if (!p) return HS_INVALID;
                    if (!run_begin) run_begin = p; else if (p != run) contiguous = false;
                    run = p + len;
                    pos += len;
                }
            }
            off[i - c.first] = pos;  // gather_block reads one offset past its last buffer's
            const u64 bytes = pos - block_at;
            if (bytes) {
                const u8 *src = nullptr;
                if (contiguous) {
                    cudaPointerAttributes pa{};
                    if (cudaPointerGetAttributes(&pa, run_begin) == cudaSuccess && pa.type == cudaMemoryTypeHost)
                        src = reinterpret_cast<const u8 *>(run_begin);
                    else
                        cudaGetLastError();
                }
                if (!src) {
                    if (total > s->h_stage_cap) {
                        CU_OK(cudaStreamSynchronize(s->copy_stream));
                        if (s->h_stage) cudaFreeHost(s->h_stage);
                        s->h_stage = nullptr;
                        s->h_stage_cap = 0;
                        const size_t want = total + total / 4 + 4096;
                        CU_OK(cudaHostAlloc((void **)&s->h_stage, want, cudaHostAllocDefault));
                        s->h_stage_cap = want;
                    }
                    gather_block(s->h_stage + host_pos, data, length, off, c.first, first, i, block_at);
                    src = s->h_stage + host_pos;
                    host_pos += bytes;
                }
                CU_OK(cudaMemcpyAsync(cp.text.p + c.text_at + block_at, src, bytes, cudaMemcpyHostToDevice, s->copy_stream));
            }
            first = i;
        }
        off[c.last - c.first] = pos;
        const u32 ncells = (u32)((c.bytes + CELL_BYTES - 1) / CELL_BYTES);
        u32 *cf = s->h_cell_first.data() + c.cell_at;
        const u32 cnt = c.last - c.first;
        u32 b = 0;
        for (u32 q = 0; q <= ncells; q++) {
            const u64 p = std::min<u64>((u64)q * CELL_BYTES, c.bytes ? c.bytes - 1 : 0);
            while (b + 1 < cnt && off[b + 1] <= p) b++;
            cf[q] = b;
        }
        c.nstrips = (u32)((c.bytes + STRIP_BYTES - 1) / STRIP_BYTES);
        CU_OK(cudaMemcpyAsync(cp.off.p + c.off_at, off, ((size_t)cnt + 1) * 8, cudaMemcpyHostToDevice, s->copy_stream));
        CU_OK(cudaMemcpyAsync(cp.cell_first.p + c.cell_at, cf, ((size_t)ncells + 1) * 4, cudaMemcpyHostToDevice, s->copy_stream));
        CU_OK(cudaEventRecord(s->chunk_ready[k], s->copy_stream));
        if (c.bytes)
            mem.push_back(TrainMember{DevCorpus{cp.text.p + c.text_at, cp.off.p + c.off_at, cp.cell_first.p + c.cell_at, c.bytes,
                                                cnt, c.nstrips, nullptr, nullptr, nullptr},
                                      c.first, s->chunk_ready[k]});
    }
    CU_OK(cudaEventRecord(s->ev[6], s->copy_stream));  // end of the H2D work (h2d_ms = ev0 -> ev6)
    return HS_SUCCESS;
}

hs_error_t scan_host_batch(const hs_database_t *db, const char *const *data, const unsigned int *length,
                           unsigned count, hs_gpu_scratch *s, DbImage **im_out) {
    hs_error_t rv = ensure_device(s);
    if (rv != HS_SUCCESS) return rv;
    cudaGetLastError();  // a non-sticky error left behind by an earlier call must not fail this one's launch checks
    DbImage *im = nullptr;
    rv = upload_database(s, db, &im);
    if (rv != HS_SUCCESS) return rv;
    *im_out = im;
    static const bool trace = getenv("HS_GPU_TRACE") != nullptr;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto us = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double, std::micro>(b - a).count();
    };
    const auto t0 = now();
    if (ordered_eligible(s, im) && s->opt_pipeline && count >= 2 && !im->view.vectored) {  // chunks have no halo: not for one stream
        u64 total = 0;
        bool one_run = true;  // every buffer begins where the one before it ended
        for (unsigned i = 0; i < count; i++) {
            if (i && length[i] && data[i] != data[i - 1] + length[i - 1]) one_run = false;
            total += length[i];
        }
        // Off by default.  Measured on C2 (64 MiB of packets, B200, PCIe 5): plain = one copy then one scan, wall 1.49 ms
        // (h2d 1.29); chunked (1/2, 1/4, 1/8, 1/8) wall 1.60 ms (h2d 1.36).  The scan + compact being hidden is 0.08 ms
        // of a 1.29 ms copy, kernels running beside the DMA take 2-3x longer and slow the DMA by 5 %, and the tail
        // (last chunk + control read-back + match copy) stays ~0.1 ms either way.  Gathered (pageable, scattered)
        // input: 14.8 GB/s plain, 13.4 GB/s chunked -- the host's gather is the bound.  "pipeline" = 1 chunks gathered
        // input only, 2 always.
        bool chunked = total >= (8u << 20);
        if (chunked && s->opt_pipeline == 1 && one_run) {
            cudaPointerAttributes pa{};
            if (cudaPointerGetAttributes(&pa, data[0]) == cudaSuccess && pa.type == cudaMemoryTypeHost) chunked = false;
            else cudaGetLastError();
        }
        if (chunked) {
            hs_gpu_stats_t &st = s->stats;
            st = hs_gpu_stats_t{};
            st.bytes = total;
            st.buffers = count;
            s->n_matches = 0;
            s->fetched = false;
            s->last_image = im;
            std::vector<TrainMember> mem;
            rv = stage_host_chunks(s, data, length, count, total, mem);
            if (rv != HS_SUCCESS) return rv;
            const auto t1 = now();
            rv = scan_train_ordered(s, im, mem.data(), (u32)mem.size());
            if (rv == HS_SUCCESS) {
                if (cudaEventElapsedTime(&st.h2d_ms, s->ev[0], s->ev[6]) != cudaSuccess) { st.h2d_ms = 0; cudaGetLastError(); }
                const auto t2 = now();
                rv = fetch_matches(s, im);
                if (trace)
                    for (size_t k = 0; k < mem.size(); k++) {
                        float a = 0, b = 0, c = 0;
                        cudaEventElapsedTime(&a, s->ev[0], s->tev[3 * k]);
                        cudaEventElapsedTime(&b, s->tev[3 * k], s->tev[3 * k + 1]);
                        cudaEventElapsedTime(&c, s->tev[3 * k + 1], s->tev[3 * k + 2]);
                        fprintf(stderr, "hs_gpu trace: chunk %zu: %llu bytes, scan starts at %.3f ms, scan %.3f ms, compact %.3f ms\n", k,
                                (unsigned long long)mem[k].cp.total, a, b, c);
                    }
                if (trace)
                    fprintf(stderr, "hs_gpu trace: stage (%zu chunks) %.1f us, scans (incl. wait for H2D) %.1f us, fetch %.1f us\n",
                            mem.size(), us(t0, t1), us(t1, t2), us(t2, now()));
                return rv;
            }
            if (rv != HS_INSUFFICIENT_SPACE) return rv;
            CU_OK(cudaStreamSynchronize(s->copy_stream));  // outside the sort-free path's envelope: the plain path below
        }
    }
    rv = stage_host_buffers(s, data, length, count);
    if (rv != HS_SUCCESS) return rv;
    const auto t1 = now();
    cudaEvent_t h2d_begin = s->ev[0];
    rv = scan_device(s, im, s->staging);
    if (rv != HS_SUCCESS) return rv;
    const auto t2 = now();
    // ev[1] was re-recorded by scan_device right after the copies were queued: h2d = ev0 -> ev1.  An empty batch
    // takes scan_device's early-out, which records nothing: no interval to read then.
    s->stats.h2d_ms = 0;
    if (s->staging.total && s->staging.count && cudaEventElapsedTime(&s->stats.h2d_ms, h2d_begin, s->ev[1]) != cudaSuccess) {
        s->stats.h2d_ms = 0;
        cudaGetLastError();
    }
    rv = fetch_matches(s, im);
    if (trace)
        fprintf(stderr, "hs_gpu trace: stage %.1f us, scan_device (incl. wait for H2D) %.1f us, fetch %.1f us\n",
                us(t0, t1), us(t1, t2), us(t2, now()));
    return rv;
}

hs_error_t validate_scan_args(const hs_database_t *db, hs_gpu_scratch *s) {
    if (!s || s->magic != 0x47505553) return HS_INVALID;
    if (!db || rd32(db, db::magic) != db::MAGIC) return HS_INVALID;
    if (rd32(db, db::version) != db::VERSION_5_4_11) return HS_DB_VERSION_ERROR;
    return HS_SUCCESS;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------
extern "C" {

const char *hs_gpu_version(void) { return HSG_VERSION; }

int hs_gpu_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

hs_error_t hs_gpu_set_device(int device) { return cudaSetDevice(device) == cudaSuccess ? HS_SUCCESS : HS_ARCH_ERROR; }

void *hs_gpu_host_alloc(size_t bytes) {
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void hs_gpu_host_free(void *p) { if (p) cudaFreeHost(p); }

hs_error_t hs_gpu_alloc_scratch(const hs_database_t *db, hs_gpu_scratch_t **scratch) {
    if (!db || !scratch) return HS_INVALID;
    // database sanity first, as hs_alloc_scratch does (dbIsValid)
    DbView probe;
    std::string why;
    int prv = parse_database(db, probe, why);
    if (prv != E_SUCCESS) {
        if (getenv("HS_GPU_DEBUG")) fprintf(stderr, "hs_gpu: database refused: %s\n", why.c_str());
        return prv;
    }
    if (hs_gpu_device_count() <= 0) return HS_ARCH_ERROR;  // no CPU fallback
    hs_gpu_scratch *s = *scratch;
    bool fresh = false;
    if (s) {
        if (s->magic != 0x47505553) return HS_INVALID;
        int z = 0;
        if (!s->in_use.compare_exchange_strong(z, 1)) return HS_SCRATCH_IN_USE;
    } else {
        s = new (std::nothrow) hs_gpu_scratch();
        if (!s) return HS_NOMEM;
        fresh = true;
        s->in_use.store(1);
        if (cudaGetDevice(&s->device) != cudaSuccess) { delete s; return HS_ARCH_ERROR; }
        cudaDeviceProp prop{};
        if (cudaGetDeviceProperties(&prop, s->device) == cudaSuccess) {
            s->sm_count = prop.multiProcessorCount;
            s->smem_optin_max = prop.sharedMemPerBlockOptin;
        }
        if (cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) != cudaSuccess ||
            cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking) != cudaSuccess) { delete s; return HS_UNKNOWN_ERROR; }
        for (auto &e : s->ev) cudaEventCreate(&e);
        if (cudaMalloc((void **)&s->d_ctl, 24 * sizeof(unsigned long long)) != cudaSuccess ||
            cudaHostAlloc((void **)&s->h_ctl, 8 * sizeof(unsigned long long), cudaHostAllocDefault) != cudaSuccess) {
            hs_gpu_free_scratch(s);
            return HS_NOMEM;
        }
        if (const char *g = getenv("HS_GPU_GRID")) s->opt_grid = atol(g);
    }
    DbImage *im = nullptr;
    hs_error_t rv = upload_database(s, db, &im);
    s->in_use.store(0);
    if (rv != HS_SUCCESS) {
        if (fresh) { hs_gpu_free_scratch(s); }
        return rv;
    }
    *scratch = s;
    return HS_SUCCESS;
}

hs_error_t hs_gpu_clone_scratch(const hs_gpu_scratch_t *src, hs_gpu_scratch_t **dest) {
    if (!src || !dest || src->magic != 0x47505553) return HS_INVALID;
    *dest = nullptr;
    hs_gpu_scratch *d = nullptr;
    for (DbImage *im : src->dbs) {
        hs_error_t rv = hs_gpu_alloc_scratch(reinterpret_cast<const hs_database_t *>(im->host_db), &d);
        if (rv != HS_SUCCESS) { if (d) hs_gpu_free_scratch(d); return rv; }
    }
    if (!d) return HS_INVALID;
    d->opt_frontend = src->opt_frontend;
    d->opt_pack = src->opt_pack;
    d->opt_grid = src->opt_grid;
    d->opt_warps = src->opt_warps;
    *dest = d;
    return HS_SUCCESS;
}

hs_error_t hs_gpu_scratch_size(const hs_gpu_scratch_t *s, size_t *host_bytes, size_t *device_bytes) {
    if (!s || s->magic != 0x47505553) return HS_INVALID;
    size_t dev = 8 * sizeof(unsigned long long);
    for (DbImage *im : s->dbs) dev += im->device_bytes;
    dev += s->staging.text.bytes() + s->staging.off.bytes() + s->staging.cell_first.bytes();
    dev += s->recs.bytes() + s->recs2.bytes() + s->d_matches.bytes() + s->cub_tmp.bytes();
    if (host_bytes) *host_bytes = sizeof(*s) + s->h_stage_cap + s->h_matches_cap * sizeof(hs_gpu_match_t);
    if (device_bytes) *device_bytes = dev;
    return HS_SUCCESS;
}

hs_error_t hs_gpu_free_scratch(hs_gpu_scratch_t *s) {
    if (!s) return HS_SUCCESS;
    if (s->magic != 0x47505553) return HS_INVALID;
    if (s->stream) cudaStreamSynchronize(s->stream);
    for (DbImage *im : s->dbs) { im->release(); delete im; }
    s->dbs.clear();
    s->staging.text.release();
    s->staging.off.release();
    s->staging.cell_first.release();
    s->recs.release();
    s->recs2.release();
    s->d_matches.release();
    s->cub_tmp.release();
    s->keep.release();
    s->seg_count.release();
    s->train_ctl.release();
    s->h_train_ctl.release();
    for (cudaEvent_t e : s->tev) cudaEventDestroy(e);
    s->tev.clear();
    for (cudaEvent_t e : s->chunk_ready) cudaEventDestroy(e);
    s->chunk_ready.clear();
    if (s->copy_stream) { cudaStreamSynchronize(s->copy_stream); cudaStreamDestroy(s->copy_stream); }
    s->h_off.release();
    s->h_cell_first.release();
    s->h_streams.release();
    s->h_stream_state.release();
    s->d_stream_state.release();
    if (s->h_stage) cudaFreeHost(s->h_stage);
    if (s->h_matches) cudaFreeHost(s->h_matches);
    if (s->h_ctl) cudaFreeHost(s->h_ctl);
    if (s->d_ctl) cudaFree(s->d_ctl);
    for (auto &e : s->ev) if (e) cudaEventDestroy(e);
    if (s->stream) cudaStreamDestroy(s->stream);
    s->magic = 0;
    delete s;
    return HS_SUCCESS;
}

// RoseEngine.mode of a database that validate_scan_args has accepted
static u32 database_mode(const hs_database_t *db) {
    const u8 *p = reinterpret_cast<const u8 *>(db);
    return rd32(p + rd32(p, db::bytecode), rose::mode);  // offsets as parse_database reads them (hsb_dbview.h)
}

// `per_buffer` (batches of independent block scans): a non-zero return of the callback ends the deliveries of THAT buffer,
// as it ends that packet's hs_scan in the loop this call replaces (src/runtime.c:459-461); the other buffers are still
// delivered and the call returns HS_SCAN_TERMINATED at the end.  Otherwise (one stream) it ends the call.
static hs_error_t deliver(hs_gpu_scratch *s, match_event_handler onEvent, void *context, void *const *contexts, bool per_buffer) {
    if (!onEvent) return HS_SUCCESS;  // matches suppressed, as with a NULL callback in the reference
    const hs_gpu_match_t *m = s->h_matches;
    bool stopped = false;
    for (u64 i = 0; i < s->n_matches; i++) {
        void *ctx = contexts ? contexts[m[i].buffer] : context;
        if (onEvent(m[i].id, 0, m[i].to, 0, ctx)) {
            if (!per_buffer) return HS_SCAN_TERMINATED;
            stopped = true;
            const u32 b = m[i].buffer;
            while (i + 1 < s->n_matches && m[i + 1].buffer == b) i++;   // matches are sorted by buffer
        }
    }
    return stopped ? HS_SCAN_TERMINATED : HS_SUCCESS;
}

hs_error_t hs_gpu_scan_batch(const hs_database_t *db, const char *const *data, const unsigned int *length,
                             unsigned int count, unsigned int flags, hs_gpu_scratch_t *scratch,
                             match_event_handler onEvent, void *const *contexts) {
    (void)flags;
    if (!scratch || !data || !length) return HS_INVALID;
    hs_error_t rv = validate_scan_args(db, scratch);
    if (rv != HS_SUCCESS) return rv;
    if (database_mode(db) != MODE_BLOCK) return HS_DB_MODE_ERROR;  // independent scans: block-mode databases only
    InUseGuard g(scratch);
    if (!g.ok) return HS_SCRATCH_IN_USE;
    DbImage *im = nullptr;
    rv = scan_host_batch(db, data, length, count, scratch, &im);
    if (rv != HS_SUCCESS) return rv;
    return deliver(scratch, onEvent, nullptr, contexts, true);
}

hs_error_t hs_gpu_scan_vector(const hs_database_t *db, const char *const *data, const unsigned int *length,
                              unsigned int count, unsigned int flags, hs_gpu_scratch_t *scratch,
                              match_event_handler onEvent, void *context) {
    (void)flags;
    if (!scratch || !data || !length) return HS_INVALID;
    hs_error_t rv = validate_scan_args(db, scratch);
    if (rv != HS_SUCCESS) return rv;
    if (database_mode(db) == MODE_STREAM) return HS_DB_MODE_ERROR;  // streams: hs_gpu_scan_stream(s)
    InUseGuard g(scratch);
    if (!g.ok) return HS_SCRATCH_IN_USE;
    DbImage *im = nullptr;
    rv = scan_host_batch(db, data, length, count, scratch, &im);
    if (rv != HS_SUCCESS) return rv;
    return deliver(scratch, onEvent, context, nullptr, database_mode(db) == MODE_BLOCK);
}

hs_error_t hs_gpu_scan_vector_ex(const hs_database_t *db, const char *const *data, const unsigned int *length,
                                 unsigned int count, unsigned int flags, hs_gpu_scratch_t *scratch,
                                 const hs_gpu_match_t **matches, uint64_t *n_matches) {
    (void)flags;
    if (!scratch || !data || !length || !matches || !n_matches) return HS_INVALID;
    hs_error_t rv = validate_scan_args(db, scratch);
    if (rv != HS_SUCCESS) return rv;
    if (database_mode(db) == MODE_STREAM) return HS_DB_MODE_ERROR;
    InUseGuard g(scratch);
    if (!g.ok) return HS_SCRATCH_IN_USE;
    DbImage *im = nullptr;
    rv = scan_host_batch(db, data, length, count, scratch, &im);
    if (rv != HS_SUCCESS) return rv;
    *matches = scratch->h_matches;
    *n_matches = scratch->n_matches;
    return HS_SUCCESS;
}

hs_error_t hs_gpu_scan(const hs_database_t *db, const char *data, unsigned int length, unsigned int flags,
                       hs_gpu_scratch_t *scratch, match_event_handler onEvent, void *context) {
    if (!scratch || !data) return HS_INVALID;
    hs_error_t rv = validate_scan_args(db, scratch);
    if (rv != HS_SUCCESS) return rv;
    if (database_mode(db) != MODE_BLOCK) return HS_DB_MODE_ERROR;  // src/runtime.c:335-337
    const char *one[1] = {data};
    unsigned int len[1] = {length};
    return hs_gpu_scan_vector(db, one, len, 1, flags, scratch, onEvent, context);
}

// ---- stream mode (pure-literal HS_MODE_STREAM databases) ----------------------------------------------------------
// hs_open_stream (src/runtime.c:543-580)
hs_error_t hs_gpu_open_stream(const hs_database_t *db, unsigned int flags, hs_gpu_stream_t **stream) {
    (void)flags;
    if (!stream) return HS_INVALID;
    *stream = nullptr;
    if (!db || rd32(db, db::magic) != db::MAGIC) return HS_INVALID;
    if (rd32(db, db::version) != db::VERSION_5_4_11) return HS_DB_VERSION_ERROR;
    if (database_mode(db) != MODE_STREAM) return HS_DB_MODE_ERROR;
    const u8 *p = reinterpret_cast<const u8 *>(db);
    const u8 *r = p + rd32(p, db::bytecode);
    if (rd8(r, rose::runtimeImpl) == rose::RUNTIME_FULL_ROSE) return HS_ARCH_ERROR;   // pure literal or one DFA engine
    hs_gpu_stream *st = new (std::nothrow) hs_gpu_stream();
    if (!st) return HS_NOMEM;
    st->db = db;
    st->crc = rd32(db, db::crc32);
    st->hreq = rd32(r, rose::historyRequired);
    *stream = st;
    return HS_SUCCESS;
}

// One block for each of `count` DIFFERENT streams of one database, in one device pass: block i is scanned behind the last
// historyRequired bytes of stream ids[i] (hs_scan_stream_internal, src/runtime.c:880-975: getHistory / populateCoreInfo
// put the history in front of the block, pureLiteralStreamExec runs the literal matcher over the block with it); the
// device sees `count` independent buffers history + block, the host drops what ends inside a history (reported by an
// earlier call), turns offsets into stream offsets, applies the stream's exhaustion keys and moves the stream on
// (maintainHistoryBuffer :483-508).  A callback that returns non-zero terminates ITS stream only.
// `closing` (hs_gpu_close_stream of a single-DFA stream whose engine accepts at end of data): no data is scanned -- the
// block is one dummy byte the kernel never looks at -- the entry state carries the close marker and the device raises the
// end-of-data accepts of that state (soleOutfixEodExec, src/runtime.c:872-900 -> nfaCheckFinalState); the stream does not move.
static hs_error_t scan_streams_impl(hs_gpu_stream_t *const *ids, const char *const *data, const unsigned int *length,
                                    unsigned int count, hs_gpu_scratch_t *scratch, match_event_handler onEvent,
                                    void *const *contexts, bool closing) {
    if (!ids || !data || !length || !scratch) return HS_INVALID;
    if (!count) return HS_SUCCESS;
    for (unsigned i = 0; i < count; i++)
        if (!ids[i] || ids[i]->magic != 0x5354524d || ids[i]->db != ids[0]->db || (length[i] && !data[i])) return HS_INVALID;
    const hs_database_t *db = ids[0]->db;
    hs_error_t rv = validate_scan_args(db, scratch);
    if (rv != HS_SUCCESS) return rv;
    if (rd32(db, db::crc32) != ids[0]->crc || database_mode(db) != MODE_STREAM) return HS_INVALID;
    InUseGuard g(scratch);
    if (!g.ok) return HS_SCRATCH_IN_USE;
    struct Release { hs_gpu_stream_t *const *ids; unsigned n; ~Release() { for (unsigned i = 0; i < n; i++) ids[i]->in_call = false; } } rel{ids, 0};
    for (unsigned i = 0; i < count; i++) {
        if (ids[i]->in_call) return HS_INVALID;   // the same stream twice in one call: its blocks would not be consecutive
        ids[i]->in_call = true;
        rel.n = i + 1;
    }
    if ((rv = ensure_device(scratch)) != HS_SUCCESS) return rv;
    DbImage *im = nullptr;
    if ((rv = upload_database(scratch, db, &im)) != HS_SUCCESS) return rv;
    // a single-DFA database carries the engine's state per stream instead of history (its historyRequired is 0)
    const bool engine = im->view.kind == ENG_MCCLELLAN || im->view.kind == ENG_SHENG;
    if (engine)
        for (unsigned i = 0; i < count; i++)
            if (!ids[i]->started) { ids[i]->started = true; ids[i]->engine_state = im->mcc.start; }
    // history + block, back to back (terminated streams and empty blocks contribute nothing: runtime.c:883-900), written
    // straight into page-locked memory of the scratch: the staging then sees one contiguous pinned run and copies it to
    // the device as it lies (stage_host_buffers), so the bytes are moved once on the host, by a few threads
    std::vector<u64> at((size_t)count + 1, 0);
    for (unsigned i = 0; i < count; i++) {
        const bool live = !ids[i]->terminated && length[i] && (!engine || ids[i]->engine_state != 0);
        at[i + 1] = at[i] + (live ? ids[i]->hist.size() + length[i] : 0);
        if (at[i + 1] - at[i] > 0xffffffffull) return HS_INVALID;
    }
    if (!scratch->h_streams.resize(at[count] + 64)) return HS_NOMEM;
    char *arena = reinterpret_cast<char *>(scratch->h_streams.data());
    std::vector<const char *> ptr(count);
    std::vector<unsigned int> len(count);
    {
        auto part = [&](unsigned a, unsigned b) {
            for (unsigned i = a; i < b; i++) {
                ptr[i] = arena + at[i];
                len[i] = (unsigned int)(at[i + 1] - at[i]);
                if (!len[i]) continue;
                std::memcpy(arena + at[i], ids[i]->hist.data(), ids[i]->hist.size());
                std::memcpy(arena + at[i] + ids[i]->hist.size(), data[i], length[i]);
            }
        };
        const unsigned threads = std::min<unsigned>(8, std::max<unsigned>(1, std::thread::hardware_concurrency() / 2));
        if (count < 512 || threads < 2) part(0, count);
        else {
            std::vector<std::thread> pool;
            for (unsigned t = 1; t < threads; t++) pool.emplace_back(part, (unsigned)((u64)count * t / threads), (unsigned)((u64)count * (t + 1) / threads));
            part(0, count / threads);
            for (std::thread &t : pool) t.join();
        }
    }
    struct StateGuard { hs_gpu_scratch *s; ~StateGuard() { s->stream_state_count = 0; } } state_guard{scratch};
    if (engine) {
        if (!scratch->h_stream_state.resize(2 * (size_t)count)) return HS_NOMEM;
        for (unsigned i = 0; i < count; i++) {
            scratch->h_stream_state[i] = len[i] ? (ids[i]->engine_state | (closing ? MCC_STREAM_CLOSE : 0u)) : 0u;
            scratch->h_stream_state[count + i] = 0;
        }
        scratch->stream_state_count = count;
    }
    rv = scan_host_batch(db, ptr.data(), len.data(), count, scratch, &im);
    if (rv != HS_SUCCESS) return rv;
    if (engine && at[count]) {   // (an all-empty call never reached the device)
        if (cudaMemcpy(scratch->h_stream_state.data() + count, scratch->d_stream_state.p + count, (size_t)count * sizeof(u32),
                       cudaMemcpyDeviceToHost) != cudaSuccess) { cudaGetLastError(); return HS_UNKNOWN_ERROR; }
        for (unsigned i = 0; i < count && !closing; i++)
            if (len[i]) ids[i]->engine_state = scratch->h_stream_state[count + i];
    }
    const std::vector<std::pair<u32, u32>> &ekey_of = im->view.stream_ekey_of;
    const bool any_ekey = im->view.stream_any_ekey;
    const hs_gpu_match_t *m = scratch->h_matches;
    bool any_stop = false;
    u64 delivered = 0;
    for (u64 k = 0; k < scratch->n_matches; k++) {
        hs_gpu_stream *st = ids[m[k].buffer];
        const u64 hlen = st->hist.size();
        if (st->terminated || (!closing && m[k].to <= hlen)) continue;   // (a closing pass reports at offset 0 of its data-less block)
        u32 ekey = ~0u;
        if (any_ekey) {
            auto it = std::lower_bound(ekey_of.begin(), ekey_of.end(), std::make_pair(m[k].id, 0u));
            if (it != ekey_of.end() && it->first == m[k].id) ekey = it->second;
        }
        if (ekey != ~0u) {
            auto pos = std::lower_bound(st->exhausted.begin(), st->exhausted.end(), ekey);
            if (pos != st->exhausted.end() && *pos == ekey) continue;   // exhausted by an earlier block
            st->exhausted.insert(pos, ekey);
        }
        delivered++;
        if (onEvent && onEvent(m[k].id, 0, st->offset - hlen + m[k].to, 0, contexts ? contexts[m[k].buffer] : nullptr)) {
            st->terminated = true;
            any_stop = true;
        }
    }
    scratch->stats.matches = delivered;
    for (unsigned i = 0; i < count && !closing; i++) {
        hs_gpu_stream *st = ids[i];
        if (st->terminated || !length[i]) continue;
        const u64 total = st->hist.size() + length[i];
        const u64 keep = std::min<u64>(st->hreq, total);
        const u8 *tail = reinterpret_cast<const u8 *>(arena) + at[i] + (total - keep);
        st->hist.assign(tail, tail + keep);   // (the arena is not the vector's own storage: no aliasing)
        st->offset += length[i];
    }
    return any_stop ? HS_SCAN_TERMINATED : HS_SUCCESS;
}

hs_error_t hs_gpu_scan_streams(hs_gpu_stream_t *const *ids, const char *const *data, const unsigned int *length,
                               unsigned int count, unsigned int flags, hs_gpu_scratch_t *scratch,
                               match_event_handler onEvent, void *const *contexts) {
    (void)flags;
    return scan_streams_impl(ids, data, length, count, scratch, onEvent, contexts, false);
}

// hs_scan_stream (src/runtime.c:980-996)
hs_error_t hs_gpu_scan_stream(hs_gpu_stream_t *id, const char *data, unsigned int length, unsigned int flags,
                              hs_gpu_scratch_t *scratch, match_event_handler onEvent, void *context) {
    if (!id || id->magic != 0x5354524d || !scratch || !data) return HS_INVALID;
    if (id->terminated) return HS_SCAN_TERMINATED;   // runtime.c:883-893
    hs_gpu_stream_t *ids[1] = {id};
    const char *d[1] = {data};
    unsigned int l[1] = {length};
    void *c[1] = {context};
    return hs_gpu_scan_streams(ids, d, l, 1, flags, scratch, onEvent, c);
}

// hs_close_stream (src/runtime.c:998-1033, report_eod_matches :636-700).  A pure-literal database has no end-of-data work
// (requiresEodCheck is 0).  A single-DFA database whose header asks for the check gets it when a handler is given
// (as in the reference, scratch and onEvent may be NULL: nothing is reported then): one device pass that raises the
// end-of-data accepts of the state the stream stands in, at the stream's length.
hs_error_t hs_gpu_close_stream(hs_gpu_stream_t *id, hs_gpu_scratch_t *scratch, match_event_handler onEvent, void *context) {
    if (!id || id->magic != 0x5354524d) return HS_INVALID;
    hs_error_t rv = HS_SUCCESS;
    const u8 *p = reinterpret_cast<const u8 *>(id->db);
    const bool wants_eod = rd8(p + rd32(p, db::bytecode), rose::requiresEodCheck) != 0;
    if (wants_eod && onEvent && scratch && id->started && id->engine_state && !id->terminated) {
        hs_gpu_stream_t *ids[1] = {id};
        const char dummy = 0, *d[1] = {&dummy};
        unsigned int l[1] = {1};
        void *c[1] = {context};
        rv = scan_streams_impl(ids, d, l, 1, scratch, onEvent, c, true);
        if (rv == HS_SCRATCH_IN_USE || rv == HS_INVALID) return rv;   // the stream stays open
    }
    id->magic = 0;
    delete id;
    return rv;
}

// hs_reset_stream (src/runtime.c:1035-1060)
hs_error_t hs_gpu_reset_stream(hs_gpu_stream_t *id, unsigned int flags, hs_gpu_scratch_t *scratch, match_event_handler onEvent,
                               void *context) {
    (void)flags; (void)scratch; (void)onEvent; (void)context;
    if (!id || id->magic != 0x5354524d) return HS_INVALID;
    id->offset = 0;
    id->hist.clear();
    id->exhausted.clear();
    id->terminated = false;
    id->started = false;
    id->engine_state = 0;
    return HS_SUCCESS;
}

// hs_copy_stream (src/runtime.c:711-737)
hs_error_t hs_gpu_copy_stream(hs_gpu_stream_t **to_id, const hs_gpu_stream_t *from_id) {
    if (!to_id) return HS_INVALID;
    *to_id = nullptr;
    if (!from_id || from_id->magic != 0x5354524d) return HS_INVALID;
    hs_gpu_stream *st = new (std::nothrow) hs_gpu_stream(*from_id);
    if (!st) return HS_NOMEM;
    st->in_call = false;
    *to_id = st;
    return HS_SUCCESS;
}

hs_error_t hs_gpu_corpus_create(const char *flat, const uint64_t *offsets, unsigned int count, hs_gpu_corpus_t **out) {
    if (!offsets || !out || !count || offsets[0] != 0) return HS_INVALID;
    for (unsigned i = 0; i < count; i++) {
        if (offsets[i + 1] < offsets[i] || offsets[i + 1] - offsets[i] > 0xffffffffull) return HS_INVALID;
    }
    const u64 total = offsets[count];
    if (total && !flat) return HS_INVALID;
    if (hs_gpu_device_count() <= 0) return HS_ARCH_ERROR;
    hs_gpu_corpus *c = new (std::nothrow) hs_gpu_corpus();
    if (!c) return HS_NOMEM;
    cudaGetDevice(&c->device);
    c->total = total;
    c->count = count;
    std::vector<u32> sf;
    build_cell_first(offsets, count, total, sf, c->nstrips);
    if (c->text.reserve(FRONT_PAD + total + BACK_PAD) != cudaSuccess || c->off.reserve((size_t)count + 1) != cudaSuccess ||
        c->cell_first.reserve(sf.size()) != cudaSuccess) {
        hs_gpu_corpus_free(c);
        return HS_NOMEM;
    }
    cudaMemset(c->text.p, 0, FRONT_PAD);
    cudaMemset(c->text.p + FRONT_PAD + total, 0, BACK_PAD);
    if (total) cudaMemcpy(c->text.p + FRONT_PAD, flat, total, cudaMemcpyHostToDevice);
    cudaMemcpy(c->off.p, offsets, ((size_t)count + 1) * 8, cudaMemcpyHostToDevice);
    cudaMemcpy(c->cell_first.p, sf.data(), sf.size() * 4, cudaMemcpyHostToDevice);
    if (cudaDeviceSynchronize() != cudaSuccess) { hs_gpu_corpus_free(c); return HS_UNKNOWN_ERROR; }
    *out = c;
    return HS_SUCCESS;
}

hs_error_t hs_gpu_corpus_free(hs_gpu_corpus_t *c) {
    if (!c) return HS_SUCCESS;
    if (c->magic != 0x43525053) return HS_INVALID;
    c->text.release();
    c->off.release();
    c->cell_first.release();
    c->order.release();
    c->magic = 0;
    delete c;
    return HS_SUCCESS;
}

hs_error_t hs_gpu_scan_corpus(const hs_database_t *db, const hs_gpu_corpus_t *corpus, hs_gpu_scratch_t *scratch,
                              uint64_t *n_matches) {
    if (!corpus || corpus->magic != 0x43525053) return HS_INVALID;
    hs_error_t rv = validate_scan_args(db, scratch);
    if (rv != HS_SUCCESS) return rv;
    if (database_mode(db) == MODE_STREAM) return HS_DB_MODE_ERROR;
    InUseGuard g(scratch);
    if (!g.ok) return HS_SCRATCH_IN_USE;
    if ((rv = ensure_device(scratch)) != HS_SUCCESS) return rv;
    DbImage *im = nullptr;
    if ((rv = upload_database(scratch, db, &im)) != HS_SUCCESS) return rv;
    scratch->stats.h2d_ms = 0;
    scratch->stats.d2h_ms = 0;
    rv = scan_device(scratch, im, *corpus);
    if (rv != HS_SUCCESS) return rv;
    if (n_matches) *n_matches = scratch->n_matches;
    return HS_SUCCESS;
}

hs_error_t hs_gpu_scan_corpora(const hs_database_t *db, const hs_gpu_corpus_t *const *corpora, unsigned int n,
                               hs_gpu_scratch_t *scratch, uint64_t *n_matches) {
    if (!corpora || !n) return HS_INVALID;
    for (unsigned i = 0; i < n; i++)
        if (!corpora[i] || corpora[i]->magic != 0x43525053) return HS_INVALID;
    if (n == 1) return hs_gpu_scan_corpus(db, corpora[0], scratch, n_matches);
    hs_error_t rv = validate_scan_args(db, scratch);
    if (rv != HS_SUCCESS) return rv;
    InUseGuard g(scratch);
    if (!g.ok) return HS_SCRATCH_IN_USE;
    if ((rv = ensure_device(scratch)) != HS_SUCCESS) return rv;
    DbImage *im = nullptr;
    if ((rv = upload_database(scratch, db, &im)) != HS_SUCCESS) return rv;
    hs_gpu_scratch *s = scratch;
    hs_gpu_stats_t &st = s->stats;
    if (!ordered_eligible(s, im) || im->view.vectored || im->view.streaming) return HS_ARCH_ERROR;  // a train needs the sort-free path (plain-report literal sets) and independent buffers
    st = hs_gpu_stats_t{};
    s->n_matches = 0;
    s->fetched = false;
    s->last_image = im;
    std::vector<TrainMember> mem;
    u64 bufs = 0;
    for (unsigned i = 0; i < n; i++) {
        st.bytes += corpora[i]->total;
        if (corpora[i]->total && corpora[i]->count) mem.push_back(TrainMember{corpora[i]->dev(), (u32)bufs, nullptr});
        bufs += corpora[i]->count;
        if (bufs > 0xffffffffull) return HS_INVALID;
    }
    st.buffers = bufs;
    if (!mem.empty()) {
        rv = scan_train_ordered(s, im, mem.data(), (u32)mem.size());
        if (rv == HS_INSUFFICIENT_SPACE) return HS_NOMEM;
        if (rv != HS_SUCCESS) return rv;
    }
    if (n_matches) *n_matches = s->n_matches;
    return HS_SUCCESS;
}

hs_error_t hs_gpu_fetch_matches(hs_gpu_scratch_t *scratch, const hs_gpu_match_t **matches, uint64_t *n_matches) {
    if (!scratch || scratch->magic != 0x47505553 || !matches || !n_matches) return HS_INVALID;
    InUseGuard g(scratch);
    if (!g.ok) return HS_SCRATCH_IN_USE;
    hs_error_t rv = fetch_matches(scratch, static_cast<DbImage *>(scratch->last_image));
    if (rv != HS_SUCCESS) return rv;
    *matches = scratch->h_matches;
    *n_matches = scratch->n_matches;
    return HS_SUCCESS;
}

hs_error_t hs_gpu_histogram(hs_gpu_scratch_t *scratch, void *device_hist_u64, uint32_t n_ids) {
    if (!scratch || scratch->magic != 0x47505553 || !device_hist_u64) return HS_INVALID;
    InUseGuard g(scratch);
    if (!g.ok) return HS_SCRATCH_IN_USE;
    const u64 n = scratch->n_matches;
    if (n) {
        k_histogram<<<(unsigned)((n + 255) / 256), 256, 0, scratch->stream>>>(
            scratch->d_matches.p, n, static_cast<unsigned long long *>(device_hist_u64), n_ids);
        CU_OK(cudaGetLastError());
        CU_OK(cudaStreamSynchronize(scratch->stream));
    }
    return HS_SUCCESS;
}

hs_error_t hs_gpu_database_mode(const hs_database_t *db, unsigned int *mode) {
    if (!db || !mode || rd32(db, db::magic) != db::MAGIC) return HS_INVALID;
    if (rd32(db, db::version) != db::VERSION_5_4_11) return HS_DB_VERSION_ERROR;
    *mode = database_mode(db);
    return HS_SUCCESS;
}

hs_error_t hs_gpu_last_stats(const hs_gpu_scratch_t *scratch, hs_gpu_stats_t *stats) {
    if (!scratch || scratch->magic != 0x47505553 || !stats) return HS_INVALID;
    *stats = scratch->stats;
    return HS_SUCCESS;
}

hs_error_t hs_gpu_database_opcodes(const hs_database_t *db, unsigned long long *hist, unsigned int n) {
    if (!db || !hist) return HS_INVALID;
    DbView v;
    std::string why;
    const int rv = parse_database(db, v, why);
    if (rv != E_SUCCESS) return (hs_error_t)rv;
    u64 h[128] = {0};
    u32 dummy = 0;
    for (u32 p : v.programs)
        if (census_program(v, p, why, dummy, h) != E_SUCCESS) return HS_INVALID;
    for (unsigned int i = 0; i < n && i < 128; i++) hist[i] = h[i];
    return HS_SUCCESS;
}

hs_error_t hs_gpu_opcode_counts(hs_gpu_scratch_t *scratch, unsigned long long *counts, unsigned int n) {
    if (!scratch || scratch->magic != 0x47505553 || !counts) return HS_INVALID;
    if (!scratch->op_counts.p) return HS_INVALID;  // option "count_ops" not set
    if (ensure_device(scratch) != HS_SUCCESS) return HS_UNKNOWN_ERROR;
    if (cudaStreamSynchronize(scratch->stream) != cudaSuccess) return HS_UNKNOWN_ERROR;
    if (n > 128) n = 128;
    return cudaMemcpy(counts, scratch->op_counts.p, n * sizeof(unsigned long long), cudaMemcpyDeviceToHost) == cudaSuccess
               ? HS_SUCCESS : HS_UNKNOWN_ERROR;
}

hs_error_t hs_gpu_set_option(hs_gpu_scratch_t *scratch, const char *name, long value) {
    if (!scratch || scratch->magic != 0x47505553 || !name) return HS_INVALID;
    if (!strcmp(name, "frontend")) { scratch->opt_frontend = value; return HS_SUCCESS; }
    if (!strcmp(name, "grid")) { scratch->opt_grid = value; return HS_SUCCESS; }
    if (!strcmp(name, "warps")) { scratch->opt_warps = value; return HS_SUCCESS; }
    if (!strcmp(name, "pipeline")) { scratch->opt_pipeline = value; return HS_SUCCESS; }
    if (!strcmp(name, "ordered")) { scratch->opt_ordered = value; return HS_SUCCESS; }
    if (!strcmp(name, "pack")) { scratch->opt_pack = value; return HS_SUCCESS; }
    if (!strcmp(name, "order")) { scratch->opt_order = value; return HS_SUCCESS; }
    if (!strcmp(name, "count_ops")) {
        if (!value) { scratch->op_counts.release(); return HS_SUCCESS; }
        if (ensure_device(scratch) != HS_SUCCESS || scratch->op_counts.reserve(128) != cudaSuccess) return HS_NOMEM;
        return cudaMemset(scratch->op_counts.p, 0, 128 * sizeof(unsigned long long)) == cudaSuccess ? HS_SUCCESS : HS_UNKNOWN_ERROR;
    }
    if (!strcmp(name, "segment")) { scratch->seg_cap_hint = value > 0 ? (u64)value : 1; return HS_SUCCESS; }
    if (!strcmp(name, "arena")) { scratch->arena_hint = value > 0 ? (u64)value : 1; return HS_SUCCESS; }
    return HS_INVALID;
}

}  // extern "C"
