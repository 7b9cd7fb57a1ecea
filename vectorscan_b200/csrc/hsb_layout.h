// hsb_layout.h -- byte layout of the Vectorscan 5.4.11 block-mode bytecode, as read (GPU runtime)
// and written (literal-set database writer) by this repository.
//
// This is an interop/wire-format description, not code taken from the reference: every constant is
// the byte offset or size the reference's x86-64 ABI gives the named field, and
// tests/test_layout.py pins each one against sizeof/offsetof printed by oracle/_ref/layout_probe
// (built from the reference's own headers).  All values are little-endian.
//
//   hs_database header        /root/reference/src/database.h:103-114
//   RoseEngine                /root/reference/src/rose/rose_internal.h:330-483
//   HWLM                      /root/reference/src/hwlm/hwlm_internal.h:48-57
//   FDR / Teddy headers       /root/reference/src/fdr/fdr_internal.h:69-86, teddy_internal.h:57-64
//   FDRConfirm / LitInfo      /root/reference/src/fdr/fdr_confirm.h:57-92
//   FDRFlood                  /root/reference/src/fdr/fdr_internal.h:50-62
//   noodTable                 /root/reference/src/hwlm/noodle_internal.h:38-49
//   Rose program opcodes      /root/reference/src/rose/rose_program.h:44-724
//   NFA header / NfaInfo      /root/reference/src/nfa/nfa_internal.h:89-128, src/rose/rose_internal.h:153-164
//   LimExNFA{32..512}         /root/reference/src/nfa/limex_internal.h:114-201
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define HSB_HD __host__ __device__ __forceinline__
#else
#define HSB_HD inline
#endif

namespace hsb {

using u8 = uint8_t;
using u16 = uint16_t;
using u32 = uint32_t;
using u64 = uint64_t;
using s32 = int32_t;

// ---- unaligned little-endian field access (host and device) --------------------------------
HSB_HD u8 rd8(const void *p, size_t off) { return reinterpret_cast<const u8 *>(p)[off]; }
HSB_HD u16 rd16(const void *p, size_t off) {
    const u8 *b = reinterpret_cast<const u8 *>(p) + off;
    return (u16)(b[0] | (b[1] << 8));
}
HSB_HD u32 rd32(const void *p, size_t off) {
    const u8 *b = reinterpret_cast<const u8 *>(p) + off;
    return (u32)b[0] | ((u32)b[1] << 8) | ((u32)b[2] << 16) | ((u32)b[3] << 24);
}
HSB_HD u64 rd64(const void *p, size_t off) {
    return (u64)rd32(p, off) | ((u64)rd32(p, off + 4) << 32);
}
inline void wr8(void *p, size_t off, u8 v) { reinterpret_cast<u8 *>(p)[off] = v; }
inline void wr16(void *p, size_t off, u16 v) { std::memcpy(reinterpret_cast<u8 *>(p) + off, &v, 2); }
inline void wr32(void *p, size_t off, u32 v) { std::memcpy(reinterpret_cast<u8 *>(p) + off, &v, 4); }
inline void wr64(void *p, size_t off, u64 v) { std::memcpy(reinterpret_cast<u8 *>(p) + off, &v, 8); }

constexpr size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
constexpr size_t CACHELINE = 64;

// ---- hs_database_t header ---------------------------------------------------------------------
namespace db {
constexpr u32 MAGIC = 0xdbdbdbdbu;
constexpr u32 VERSION_5_4_11 = (5u << 24) | (4u << 16) | (11u << 8);
constexpr size_t magic = 0, version = 4, length = 8, platform = 16, crc32 = 24, reserved0 = 28,
                 reserved1 = 32, bytecode = 36, padding = 40, bytes = 104, SIZE = 104;
// platform bits: a database stamped with all three is accepted by every x86 build of the reference
// runtime (src/database.c:115-124) and rules out the AVX2-only fat Teddy engines.
constexpr u64 NOAVX2 = 4u << 13, NOAVX512 = 8u << 13, NOAVX512VBMI = 0x10u << 13;
constexpr u64 PLATFORM_GENERIC = NOAVX2 | NOAVX512 | NOAVX512VBMI;
// serialized form (src/database.c:62-110,130-176): SIZE + length bytes; packed header
// {magic, version, length, platform(8), crc32, reserved0, reserved1} = 32 bytes, then the bytecode
constexpr size_t SER_HEADER = 32;
// in-memory images written by this repo: 64-byte aligned base, bytecode at +64 (inside `padding`)
constexpr size_t IMAGE_BYTECODE = 64;
}  // namespace db

constexpr u32 MODE_BLOCK = 1, MODE_STREAM = 2, MODE_VECTORED = 4;

// ---- RoseEngine -------------------------------------------------------------------------------
namespace rose {
constexpr size_t SIZE = 472;
constexpr size_t pureLiteral = 0, noFloatingRoots = 1, requiresEodCheck = 2, hasOutfixesInSmallBlock = 3,
                 runtimeImpl = 4, mpvTriggeredByLeaf = 5, canExhaust = 6, hasSom = 7, somHorizon = 8,
                 mode = 12, historyRequired = 16, ekeyCount = 20, lkeyCount = 24, lopCount = 28,
                 ckeyCount = 32, logicalTreeOffset = 36, combInfoMapOffset = 40, dkeyCount = 44,
                 dkeyLogSize = 48, invDkeyOffset = 52, somLocationCount = 56,
                 somLocationFatbitSize = 60, rolesWithStateCount = 64, stateSize = 68,
                 anchorStateSize = 72, tStateSize = 76, scratchStateSize = 80, smallWriteOffset = 84,
                 amatcherOffset = 88, ematcherOffset = 92, fmatcherOffset = 96, drmatcherOffset = 100,
                 sbmatcherOffset = 104, longLitTableOffset = 108, amatcherMinWidth = 112,
                 fmatcherMinWidth = 116, eodmatcherMinWidth = 120, amatcherMaxBiAnchoredWidth = 124,
                 fmatcherMaxBiAnchoredWidth = 128, reportProgramOffset = 132, reportProgramCount = 136,
                 delayProgramOffset = 140, anchoredProgramOffset = 144, activeArrayCount = 148,
                 activeLeftCount = 152, queueCount = 156, activeQueueArraySize = 160,
                 eagerIterOffset = 164, handledKeyCount = 168, handledKeyFatbitSize = 172,
                 leftOffset = 176, roseCount = 180, eodProgramOffset = 184, flushCombProgramOffset = 188,
                 lastFlushCombProgramOffset = 192, lastByteHistoryIterOffset = 196, minWidth = 200,
                 minWidthExcludingBoundaries = 204, maxBiAnchoredWidth = 208, anchoredDistance = 212,
                 anchoredMinDistance = 216, floatingDistance = 220, floatingMinDistance = 224,
                 smallBlockDistance = 228, floatingMinLiteralMatchOffset = 232, nfaInfoOffset = 236,
                 initialGroups = 240, floating_group_mask = 248, size = 256, delay_count = 260,
                 delay_fatbit_size = 264, anchored_count = 268, anchored_fatbit_size = 272,
                 maxFloatingDelayedMatch = 276, delayRebuildLength = 280, stateOffsets = 284,
                 boundary = 376, totalNumLiterals = 388, asize = 392, outfixBeginQueue = 396,
                 outfixEndQueue = 400, leftfixBeginQueue = 404, initMpvNfa = 408, rosePrefixCount = 412,
                 activeLeftIterOffset = 416, ematcherRegionSize = 420, somRevCount = 424,
                 somRevOffsetOffset = 428, longLitStreamState = 432, state_init = 436;
// RoseStateOffsets, relative to rose::stateOffsets
namespace so {
constexpr size_t history = 0, exhausted = 4, exhausted_size = 8, logicalVec = 12, logicalVec_size = 16,
                 combVec = 20, combVec_size = 24, activeLeafArray = 28, activeLeafArray_size = 32,
                 activeLeftArray = 36, activeLeftArray_size = 40, leftfixLagTable = 44, anchorState = 48,
                 groups = 52, groups_size = 56, longLitState = 60, longLitState_size = 64,
                 somLocation = 68, somValid = 72, somWritable = 76, somMultibit_size = 80,
                 nfaStateBegin = 84, end = 88, SIZE = 92;
}
// RoseBoundaryReports, relative to rose::boundary
namespace bd {
constexpr size_t reportEodOffset = 0, reportZeroOffset = 4, reportZeroEodOffset = 8;
}
constexpr u8 RUNTIME_FULL_ROSE = 0, RUNTIME_PURE_LITERAL = 1, RUNTIME_SINGLE_OUTFIX = 2;
constexpr u32 BOUND_INF = 0xffffffffu;
}  // namespace rose

// ---- HWLM -------------------------------------------------------------------------------------
namespace hwlm {
constexpr size_t type = 0, accel1_groups = 8, accel1 = 16, accel0 = 96, SIZE = 176;
constexpr size_t ACCEL_AUX_SIZE = 80;
constexpr size_t BODY = 192;  // ROUNDUP_CL(sizeof(HWLM)): engine body follows here
constexpr u8 ENGINE_NOOD = 16, ENGINE_FDR = 12;
}  // namespace hwlm

// ---- AccelAux (src/nfa/accel.h:72-142) -----------------------------------------------------------
namespace accel {
constexpr u8 NONE = 0, VERM = 1, VERM_NOCASE = 2, DVERM = 3, DVERM_NOCASE = 4, SHUFTI = 13,
             DSHUFTI = 14, TRUFFLE = 15, RED_TAPE = 16, DVERM_MASKED = 17;
// every variant starts {u8 accel_type; u8 offset;}
constexpr size_t type = 0, offset = 1;
constexpr size_t verm_c = 2;               // .verm {type, offset, c}
constexpr size_t dverm_c1 = 2, dverm_c2 = 3, dverm_m1 = 4, dverm_m2 = 5;
constexpr size_t shufti_lo = 16, shufti_hi = 32;       // m128 each (16-byte aligned)
constexpr size_t truffle_mask1 = 16, truffle_mask2 = 32;
}  // namespace accel

// ---- FDR / Teddy header -----------------------------------------------------------------------
namespace fdr {
constexpr size_t engineID = 0, size = 4, maxStringLen = 8, numStrings = 12, confOffset = 16,
                 floodOffset = 20, stride = 24, domain = 25, domainMask = 26, tabSize = 28, start = 32,
                 SIZE = 48;
constexpr size_t TABLE = 64;  // ROUNDUP_CL(sizeof(FDR)): u64 ft[1 << domain] follows here
constexpr u32 NUM_BUCKETS = 8;
}  // namespace fdr
namespace teddy {
constexpr size_t SIZE = 24;   // engineID..floodOffset shared with fdr::
constexpr size_t MASKS = 64;  // ROUNDUP_CL(sizeof(Teddy)): nibble masks follow here
constexpr u32 ENGINE_FIRST = 11, ENGINE_LAST = 18;  // 8-bucket Teddy, 1..4 masks, plain/packed
constexpr u32 FAT_FIRST = 3, FAT_LAST = 10;         // 16-bucket (AVX2-only) Teddy
constexpr size_t REINFORCE_ROWS = 257;              // (N_CHARS + 1) u64 per mask
}  // namespace teddy

namespace conf {  // struct FDRConfirm, then u32 litIndex[1 << nBits], then LitInfo[]
constexpr size_t andmsk = 0, mult = 8, nBits = 16, groups = 24, SIZE = 32;
constexpr u64 MULT = 0x0b4e0ef37bc32127ull;
}
namespace lit {  // struct LitInfo
constexpr size_t v = 0, msk = 8, groups = 16, id = 24, size = 28, flags = 29, next = 30, SIZE = 32;
constexpr u8 FLAG_NOREPEAT = 1;
}
namespace flood {  // u32 index[256], then FDRFlood[]
constexpr size_t allGroups = 0, suffix = 8, idCount = 12, ids = 16, groups = 80, SIZE = 208;
constexpr u32 MAX_IDS = 16;
}
namespace nood {  // struct noodTable
constexpr size_t id = 0, msk = 8, cmp = 16, msk_len = 24, key_offset = 25, nocase = 26, single = 27,
                 key0 = 28, key1 = 29, SIZE = 32;
}

// ---- Rose program instructions ------------------------------------------------------------------
// Every instruction is a C struct starting with `u8 code`, padded to ROSE_INSTR_MIN_ALIGN (8).
namespace op {
enum : u8 {
    END = 0, ANCHORED_DELAY = 1, CHECK_LIT_EARLY = 2, CHECK_GROUPS = 3, CHECK_ONLY_EOD = 4,
    CHECK_BOUNDS = 5, CHECK_NOT_HANDLED = 6, CHECK_SINGLE_LOOKAROUND = 7, CHECK_LOOKAROUND = 8,
    CHECK_MASK = 9, CHECK_MASK_32 = 10, CHECK_BYTE = 11, CHECK_SHUFTI_16x8 = 12,
    CHECK_SHUFTI_32x8 = 13, CHECK_SHUFTI_16x16 = 14, CHECK_SHUFTI_32x16 = 15, CHECK_INFIX = 16,
    CHECK_PREFIX = 17, PUSH_DELAYED = 18, DUMMY_NOP = 19, CATCH_UP = 20, CATCH_UP_MPV = 21,
    SOM_ADJUST = 22, SOM_LEFTFIX = 23, SOM_FROM_REPORT = 24, SOM_ZERO = 25, TRIGGER_INFIX = 26,
    TRIGGER_SUFFIX = 27, DEDUPE = 28, DEDUPE_SOM = 29, REPORT_CHAIN = 30, REPORT_SOM_INT = 31,
    REPORT_SOM_AWARE = 32, REPORT = 33, REPORT_EXHAUST = 34, REPORT_SOM = 35, REPORT_SOM_EXHAUST = 36,
    DEDUPE_AND_REPORT = 37, FINAL_REPORT = 38, CHECK_EXHAUSTED = 39, CHECK_MIN_LENGTH = 40,
    SET_STATE = 41, SET_GROUPS = 42, SQUASH_GROUPS = 43, CHECK_STATE = 44, SPARSE_ITER_BEGIN = 45,
    SPARSE_ITER_NEXT = 46, SPARSE_ITER_ANY = 47, ENGINES_EOD = 48, SUFFIXES_EOD = 49,
    MATCHER_EOD = 50, CHECK_LONG_LIT = 51, CHECK_LONG_LIT_NOCASE = 52, CHECK_MED_LIT = 53,
    CHECK_MED_LIT_NOCASE = 54, CLEAR_WORK_DONE = 55, MULTIPATH_LOOKAROUND = 56,
    CHECK_MULTIPATH_SHUFTI_16x8 = 57, CHECK_MULTIPATH_SHUFTI_32x8 = 58,
    CHECK_MULTIPATH_SHUFTI_32x16 = 59, CHECK_MULTIPATH_SHUFTI_64 = 60, INCLUDED_JUMP = 61,
    SET_LOGICAL = 62, SET_COMBINATION = 63, FLUSH_COMBINATION = 64, SET_EXHAUST = 65,
    LAST_FLUSH_COMBINATION = 66, CHECK_SHUFTI_64x8 = 67, CHECK_SHUFTI_64x16 = 68, CHECK_MASK_64 = 69,
    LAST = 69
};
constexpr size_t ALIGN = 8;
// operand offsets + padded sizes for the opcodes this runtime executes
namespace check_groups { constexpr size_t groups = 8, SIZE = 16; }
namespace check_mask { constexpr size_t and_mask = 8, cmp_mask = 16, neg_mask = 24, offset = 32, fail_jump = 36, SIZE = 40; }
namespace check_mask32 { constexpr size_t and_mask = 1, cmp_mask = 33, neg_mask = 68, offset = 72, fail_jump = 76, SIZE = 80; }
namespace check_byte { constexpr size_t and_mask = 1, cmp_mask = 2, negation = 3, offset = 4, fail_jump = 8, SIZE = 16; }
namespace check_shufti_16x8 { constexpr size_t nib_mask = 1, bucket_select_mask = 33, neg_mask = 52, offset = 56, fail_jump = 60, SIZE = 64; }
namespace check_shufti_32x8 { constexpr size_t hi_mask = 1, lo_mask = 17, bucket_select_mask = 33, neg_mask = 68, offset = 72, fail_jump = 76, SIZE = 80; }
namespace check_shufti_16x16 { constexpr size_t hi_mask = 1, lo_mask = 33, bucket_select_mask = 65, neg_mask = 100, offset = 104, fail_jump = 108, SIZE = 112; }
namespace check_shufti_32x16 { constexpr size_t hi_mask = 1, lo_mask = 33, bucket_select_mask_hi = 65, bucket_select_mask_lo = 97, neg_mask = 132, offset = 136, fail_jump = 140, SIZE = 144; }
namespace check_mask64 { constexpr size_t and_mask = 1, cmp_mask = 65, neg_mask = 136, offset = 144, fail_jump = 148, SIZE = 152; }
namespace check_shufti_64x8 { constexpr size_t hi_mask = 1, lo_mask = 65, bucket_select_mask = 129, neg_mask = 200, offset = 208, fail_jump = 212, SIZE = 216; }
namespace check_shufti_64x16 { constexpr size_t hi_mask_1 = 1, hi_mask_2 = 65, lo_mask_1 = 129, lo_mask_2 = 193, bucket_select_mask_hi = 257, bucket_select_mask_lo = 321, neg_mask = 392, offset = 400, fail_jump = 404, SIZE = 408; }
namespace dedupe { constexpr size_t quash_som = 1, dkey = 4, offset_adjust = 8, fail_jump = 12, SIZE = 16; }
namespace report { constexpr size_t onmatch = 4, offset_adjust = 8, SIZE = 16; }           // REPORT, FINAL_REPORT
namespace report_exhaust { constexpr size_t onmatch = 4, offset_adjust = 8, ekey = 12, SIZE = 16; }
namespace dedupe_and_report { constexpr size_t quash_som = 1, dkey = 4, onmatch = 8, offset_adjust = 12, fail_jump = 16, SIZE = 24; }
namespace check_exhausted { constexpr size_t ekey = 4, fail_jump = 8, SIZE = 16; }
namespace squash_groups { constexpr size_t groups = 8, SIZE = 16; }
namespace check_lit { constexpr size_t lit_offset = 4, lit_length = 8, fail_jump = 12, SIZE = 16; }  // CHECK_{MED,LONG}_LIT[_NOCASE]
namespace included_jump { constexpr size_t squash = 1, child_offset = 4, SIZE = 8; }
namespace set_exhaust { constexpr size_t ekey = 4, SIZE = 8; }
constexpr size_t SIZE_1 = 8;  // END, CATCH_UP, CLEAR_WORK_DONE, ... (code byte only)
}  // namespace op

// ---- NFA engines --------------------------------------------------------------------------------
namespace nfa {  // struct NFA (64 bytes, cache-line aligned); the implementation struct follows it
constexpr size_t flags = 0, length = 4, type = 8, rAccelType = 9, rAccelOffset = 10, maxBiAnchoredWidth = 11,
                 rAccelData = 12, queueIndex = 16, nPositions = 20, scratchStateSize = 24, streamStateSize = 28,
                 maxWidth = 32, minWidth = 36, maxOffset = 40, SIZE = 64;
constexpr u8 LIMEX_32 = 0, LIMEX_64 = 1, LIMEX_128 = 2, LIMEX_256 = 3, LIMEX_384 = 4, LIMEX_512 = 5;
constexpr u8 MCCLELLAN_8 = 6, MCCLELLAN_16 = 7, SHENG = 17, MCSHENG_8 = 19, MCSHENG_16 = 20, SHENG_32 = 21, SHENG_64 = 22, MCSHENG64_8 = 23, MCSHENG64_16 = 24;
constexpr u32 ACCEPTS_EOD = 1, ZOMBIE = 2;
}
namespace nfainfo {  // struct NfaInfo
constexpr size_t nfaOffset = 0, stateOffset = 4, fullStateOffset = 8, ekeyListOffset = 12, no_retrigger = 16,
                 in_sbmatcher = 17, eod = 18, SIZE = 20;
}
// struct mcclellan (src/nfa/mcclellan_internal.h:77-101) follows struct NFA; the successor table follows it
// (u8 or u16 per entry, row = state << alphaShift, column = remap[byte]); then mstate_aux[state], report lists
// {u32 count, ReportID[]}, accel structures and (16-bit only) 32-byte Sherman states.
namespace mcclellan {
constexpr size_t state_count = 0, length = 4, start_anchored = 8, start_floating = 10, aux_offset = 12,
                 sherman_offset = 16, sherman_end = 20, accel_limit_8 = 24, accept_limit_8 = 26, sherman_limit = 28,
                 wide_limit = 30, alphaShift = 32, flags = 33, has_accel = 34, has_wide = 35, remap = 36,
                 arb_report = 292, accel_offset = 296, haig_offset = 300, wide_offset = 304, SIZE = 308;
constexpr u32 FLAG_SINGLE = 1;
constexpr u32 ACCEPT_FLAG = 0x8000, ACCEL_FLAG = 0x4000, STATE_MASK = 0x3fff;
constexpr u32 SHERMAN_STATE = 1, SHERMAN_FIXED_SIZE = 32, SHERMAN_LEN_OFFSET = 1, SHERMAN_DADDY_OFFSET = 2,
              SHERMAN_CHARS_OFFSET = 4, MAX_SHERMAN_LIST_LEN = 9;
namespace aux { constexpr size_t accept = 0, accept_eod = 4, top = 8, accel_offset = 12, SIZE = 16; }
}
// struct mcsheng (src/nfa/mcsheng_internal.h:63-88): a McClellan DFA whose states 1 .. sheng_end - 1 (at most 16, none of
// them accepting) step through 256 byte-shuffle masks instead of the successor table: sheng_masks[c][s - 1] is the
// successor of state s on byte c, written as id - 1 for a state of the region, sheng_end - 1 for the dead state, and the
// state's own id (which must fit a byte) for a state outside the region.  The successor table that follows covers the
// states >= sheng_end only, indexed by id - sheng_end (mcsheng.c:380-388); mstate_aux[], report lists and Sherman states
// are McClellan's (namespace mcclellan).
namespace mcsheng {
constexpr size_t state_count = 0, length = 4, start_anchored = 8, start_floating = 10, aux_offset = 12,
                 sherman_offset = 16, sherman_end = 20, sheng_end = 24, sheng_accel_limit = 26, accel_limit_8 = 28,
                 accept_limit_8 = 30, sherman_limit = 32, alphaShift = 34, flags = 35, has_accel = 36, remap = 37,
                 arb_report = 296, accel_offset = 300, sheng_masks = 304, SIZE = 4400;
constexpr u32 MAX_SHENG_STATES = 16;
// struct mcsheng64 (mcsheng_internal.h:94-119, AVX512VBMI targets; NFA types 23 / 24): the same header, a region of up to
// 64 states, 64-byte masks (vpermb) at the next 64-byte boundary
constexpr size_t sheng_succ_masks64 = 320, SIZE64 = 320 + 256 * 64;
constexpr u32 MAX_SHENG64_STATES = 64;
}
// struct sheng (src/nfa/sheng_internal.h:68-79): a DFA of at most 16 states as 256 byte-shuffle masks; mask[c][s] is
// the successor of state s on byte c with flag bits (accept / dead / accel) beside the 4-bit state id.  sstate_aux[]
// and the report lists {u32 count, ReportID[]} follow; their offsets are relative to struct NFA.
namespace sheng {
constexpr size_t shuffle_masks = 0, length = 4096, aux_offset = 4100, report_offset = 4104, accel_offset = 4108,
                 n_states = 4112, anchored = 4113, floating = 4114, flags = 4115, report = 4116, SIZE = 4128;
constexpr u32 STATE_ACCEPT = 0x10, STATE_DEAD = 0x20, STATE_ACCEL = 0x40, STATE_MASK = 0xf;
constexpr u32 FLAG_SINGLE_REPORT = 1, FLAG_CAN_DIE = 2, FLAG_HAS_ACCEL = 4;
namespace aux { constexpr size_t accept = 0, accept_eod = 4, accel = 8, top = 12, SIZE = 16; }
// struct sheng32 / sheng64 (src/nfa/sheng_internal.h:81-107, AVX512VBMI targets): 256 succ_masks of 64 bytes, the same
// trailer, other flag bits.  One descriptor for the three shapes:
struct Shape {
    u32 max_states, stride, accept, dead, mask;  // mask entry = state id | flags
    size_t length, aux_offset, report_offset, accel_offset, n_states, anchored, floating, flags, report, SIZE;
    u8 nfa_type;
};
HSB_HD Shape shape(u32 nfa_type) {
    if (nfa_type == 21) return Shape{32, 64, 0x20, 0x40, 0x1f, 16384, 16388, 16392, 16396, 16400, 16401, 16402, 16403, 16404, 16448, 21};
    if (nfa_type == 22) return Shape{64, 64, 0x40, 0x80, 0x3f, 16384, 16388, 16392, 16396, 16400, 16401, 16402, 16403, 16404, 16448, 22};
    return Shape{16, 16, 0x10, 0x20, 0x0f, 4096, 4100, 4104, 4108, 4112, 4113, 4114, 4115, 4116, 4128, 17};
}
}
namespace nfaaccept {  // struct NFAAccept
constexpr size_t single_report = 0, reports = 4, squash = 8, SIZE = 12;
}
// LimExNFA<bits>: a fixed head, then twelve state-sized masks, shift[8], the shift bookkeeping and three m512
// exception-extraction masks.  State masks are stored as `bytes`-byte little-endian bit vectors aligned to
// `align` (u32, u64, m128, m256, m384 = 3 x m128, m512).
namespace limex {
constexpr size_t reachMap = 0, reachSize = 256, accelCount = 260, accelTableOffset = 264, accelAuxCount = 268,
                 accelAuxOffset = 272, acceptCount = 276, acceptOffset = 280, acceptEodCount = 284,
                 acceptEodOffset = 288, exceptionCount = 292, exceptionOffset = 296, repeatCount = 300,
                 repeatOffset = 304, squashOffset = 308, squashCount = 312, topCount = 316, topOffset = 320,
                 stateSize = 324, flags = 328, HEAD = 332;
constexpr u32 FLAG_COMPRESS_STATE = 1, FLAG_COMPRESS_MASKED = 2, FLAG_CANNOT_DIE = 4, FLAG_EXTRACT_EXP = 8;
constexpr u32 MAX_SHIFT_COUNT = 8, MAX_SHIFT_AMOUNT = 16;
constexpr u8 SQUASH_NONE = 0, SQUASH_CYCLIC = 1, SQUASH_TUG = 2, SQUASH_REPORT = 3;
struct Layout {
    u32 bits, bytes, align;
    size_t init, initDS, accept, acceptAtEOD, accel, accelPermute, accelCompare, accel_and_friends, compressMask,
        exceptionMask, repeatCyclicMask, zombieMask, shift, shiftCount, shiftAmount, exceptionShufMask,
        exceptionBitMask, exceptionAndMask, SIZE;
    // struct NFAException<bits>
    size_t e_squash, e_successors, e_reports, e_repeatOffset, e_hasSquash, e_trigger, E_SIZE;
};
HSB_HD Layout layout(u32 type) {
    const u32 bits_of[6] = {32, 64, 128, 256, 384, 512}, align_of[6] = {4, 8, 16, 32, 16, 64};
    Layout l{};
    l.bits = bits_of[type];
    l.bytes = l.bits / 8;
    l.align = align_of[type];
    const size_t S = l.bytes;
    l.init = (HEAD + l.align - 1) / l.align * l.align;
    l.initDS = l.init + S; l.accept = l.init + 2 * S; l.acceptAtEOD = l.init + 3 * S; l.accel = l.init + 4 * S;
    l.accelPermute = l.init + 5 * S; l.accelCompare = l.init + 6 * S; l.accel_and_friends = l.init + 7 * S;
    l.compressMask = l.init + 8 * S; l.exceptionMask = l.init + 9 * S; l.repeatCyclicMask = l.init + 10 * S;
    l.zombieMask = l.init + 11 * S; l.shift = l.init + 12 * S;
    l.shiftCount = l.shift + MAX_SHIFT_COUNT * S;
    l.shiftAmount = l.shiftCount + 4;
    l.exceptionShufMask = (l.shiftAmount + MAX_SHIFT_COUNT + 63) / 64 * 64;
    l.exceptionBitMask = l.exceptionShufMask + 64;
    l.exceptionAndMask = l.exceptionShufMask + 128;
    l.SIZE = l.exceptionShufMask + 192;
    l.e_squash = 0; l.e_successors = S; l.e_reports = 2 * S; l.e_repeatOffset = 2 * S + 4; l.e_hasSquash = 2 * S + 8;
    l.e_trigger = 2 * S + 9;
    l.E_SIZE = (2 * S + 10 + l.align - 1) / l.align * l.align;
    return l;
}
}  // namespace limex

constexpr u32 INVALID_EKEY = 0xffffffffu;
constexpr u32 INVALID_DKEY = 0xffffffffu;  // MO_INVALID_IDX

// src/util/multibit_build.cpp:49-73 -- bytes of a multibit holding total_bits keys
inline u32 mmbit_size(u32 total_bits) {
    if (total_bits <= 256) return (u32)(round_up(total_bits, 8) / 8);
    u64 level = 1, total = 0;
    while (level * 64 < total_bits) { total += level; level <<= 6; }
    total += ((u64)total_bits + 63) / 64;
    return (u32)(total * 8);
}
// src/util/fatbit_build.cpp:40-42 -- struct fatbit is at least 32 bytes
inline u32 fatbit_size(u32 total_bits) { u32 m = mmbit_size(total_bits); return m < 32 ? 32 : m; }

}  // namespace hsb
