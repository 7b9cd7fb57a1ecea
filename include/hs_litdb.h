/* hs_litdb.h -- C ABI of libhs_litdb.so: a host-side writer of pure-literal, block-mode databases
 * in the Vectorscan 5.4.11 bytecode format.
 *
 * It is the stand-in, for literal sets only, for the reference's host compiler entry point
 *     hs_compile_lit_multi()      /root/reference/src/hs_compile.h:640-648, src/hs.cpp:437-445
 * which cannot be built in this image (needs Boost + Ragel).  Where the real compiler is available
 * it is used unchanged and this library is not needed: the GPU runtime consumes any hs_database_t.
 * The emitted blob is accepted and scanned by the unmodified reference runtime (hs_scan).
 */
#ifndef HS_LITDB_H
#define HS_LITDB_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

struct hs_database;
#if !defined(HS_COMMON_H_) && !defined(HSB_HS_DATABASE_T) /* the reference's hs_common.h, or our other header, may have declared it */
#define HSB_HS_DATABASE_T
typedef struct hs_database hs_database_t; /* same opaque type as src/hs_common.h:56 */
#endif

#define HSB_LIT_SUCCESS 0
#define HSB_LIT_COMPILER_ERROR (-4) /* HS_COMPILER_ERROR, src/hs_common.h:503 */

/* per-literal flags: same values as HS_FLAG_CASELESS / HS_FLAG_SINGLEMATCH (src/hs_compile.h:869,911) */
#define HSB_FLAG_CASELESS 1u
#define HSB_FLAG_DOTALL 2u /* HS_FLAG_DOTALL, src/hs_compile.h:877 (regex writer only) */
#define HSB_FLAG_SINGLEMATCH 8u

#define HSB_MODE_BLOCK 1u /* HS_MODE_BLOCK, src/hs_compile.h:1156 */
#define HSB_MODE_STREAM 2u /* HS_MODE_STREAM, src/hs_compile.h:1166 */
#define HSB_MODE_VECTORED 4u /* HS_MODE_VECTORED, src/hs_compile.h:1171 */

/* literal-matcher choice; AUTO follows the reference's rules: 1 fragment -> noodle, <= 48 fragments
 * that pack into 8 buckets -> Teddy, else FDR (src/hwlm/hwlm_build.cpp:201-256). */
#define HSB_LIT_ENGINE_AUTO 0
#define HSB_LIT_ENGINE_FDR 1
#define HSB_LIT_ENGINE_TEDDY 2
#define HSB_LIT_ENGINE_NOODLE 3

/* The knobs the reference exposes through its Grey box / test hints (src/grey.h, fdr_compile.cpp:861-865);
 * zero-initialise for reference-default behaviour. */
typedef struct hsb_lit_options {
    int engine;      /* HSB_LIT_ENGINE_* */
    int fdr_domain;  /* 9..15, 0 = chooseEngine's pick */
    int fdr_stride;  /* 1, 2, 4, 0 = chooseEngine's pick */
    int teddy_masks; /* 1..4, 0 = chooseTeddyEngine's pick */
    int no_flood;    /* 1 = Grey::fdrAllowFlood off */
    int teddy_fat;   /* with engine = TEDDY: 1 = the 16-bucket "fat" Teddy an AVX2 target gets (engines 3-10,
                        src/fdr/teddy_engine_description.cpp:52-62; up to 96 fragments); needs an AVX2 runtime build */
} hsb_lit_options;

/* Same parameter meaning as hs_compile_lit_multi; `platform` is replaced by `options` (the blob is
 * always stamped for the generic x86 platform so every ISA build of the reference runtime accepts it),
 * `error` receives a malloc'ed message (hsb_free_error).  mode: HSB_MODE_BLOCK, or HSB_MODE_VECTORED / HSB_MODE_STREAM for
 * pure-literal sets -- the database the reference's hs_scan_vector (src/runtime.c:1129-1131) / hs_open_stream (:556-558)
 * wants: RoseEngine.mode = HS_MODE_VECTORED / HS_MODE_STREAM,
 * historyRequired = longest literal - 1 (or the deepest lookaround), stream-state layout with a history area
 * (rose_build_bytecode.cpp:381-440). */
int hsb_compile_lit_multi(const char *const *expressions, const unsigned *flags, const unsigned *ids,
                          const size_t *lens, unsigned elements, unsigned mode,
                          const hsb_lit_options *options, hs_database_t **db, char **error);
/* ---- extended literal form: what the reference compiler attaches to a literal beyond its text -------------------
 * Per-literal extras (all optional; a zeroed record = a plain literal):
 *  - `look`: byte constraints at negative offsets from the literal's END (offset -1 = its last byte), the
 *    "lookaround" the reference derives for mixed-case / masked literals (makeCheckLitMaskInstruction,
 *    src/rose/rose_build_program.cpp:1526-1558).  An entry is either and/cmp form -- ((byte & and_mask) == cmp_mask)
 *    != negate -- or a 256-bit class (cls != NULL: byte is in the class).  The writer picks the instruction the way
 *    makeLookaroundInstruction does (:1343-1523): CHECK_BYTE (one and/cmp entry), CHECK_MASK (span <= 8),
 *    CHECK_MASK_32 (<= 32), CHECK_MASK_64 (<= 64, AVX512 targets), and for classes CHECK_SHUFTI_16x8 / 32x8 / 16x16 /
 *    32x16 / 64x8 / 64x16.  An entry that lies before the start of the buffer fails the literal (program_runtime.c:604,651).
 *  - `group`: the literal's Rose group (0..62).  HSB_EXT_CHECK_GROUPS puts a CHECK_GROUPS at the head of its program,
 *    HSB_EXT_SQUASH_GROUP a SQUASH_GROUPS behind its report (only honoured where it is result-neutral: a SINGLEMATCH
 *    literal alone in its group, rose_build_groups.cpp isGroupSquasher), HSB_EXT_QUIET makes the report quiet
 *    (HS_FLAG_QUIET: no callback; with SINGLEMATCH a SET_EXHAUST, makeReport rose_build_program.cpp:533-640). */
typedef struct hsb_lit_look {
    int offset;                   /* -64 .. -1 */
    unsigned char and_mask, cmp_mask, negate, reserved;
    const unsigned char *cls;     /* 32-byte bitmap (bit c of byte c/8), or NULL for the and/cmp form */
} hsb_lit_look;
#define HSB_EXT_CHECK_GROUPS 1u
#define HSB_EXT_SQUASH_GROUP 2u
#define HSB_EXT_QUIET 4u
typedef struct hsb_lit_ext {
    const hsb_lit_look *look;
    unsigned look_count;
    unsigned group;
    unsigned flags; /* HSB_EXT_* */
} hsb_lit_ext;
/* database-level extras */
typedef struct hsb_lit_build_ext {
    int included_jump;                 /* FDR: INCLUDED_JUMP + bucket squash for suffix-included fragments
                                          (src/fdr/fdr_compile.cpp:654-812, rose_build_bytecode.cpp:2877-2950) */
    int target_avx512;                 /* allow the 64-byte instructions (only an AVX512 runtime build executes them) */
    unsigned long long groups_off;     /* group bits left out of RoseEngine.initialGroups */
    unsigned max_bi_anchored_width;    /* RoseEngine.maxBiAnchoredWidth; 0 = ROSE_BOUND_INF */
    unsigned reserved_pure_literal;    /* set by hsb_compile_lit_multi_ext; leave 0 */
} hsb_lit_build_ext;
int hsb_compile_lit_multi_ext(const char *const *expressions, const unsigned *flags, const unsigned *ids,
                              const size_t *lens, const hsb_lit_ext *ext, unsigned elements, unsigned mode,
                              const hsb_lit_options *options, const hsb_lit_build_ext *build_ext,
                              hs_database_t **db, char **error);
/* Mixed pattern sets -- literals AND regular expressions in ONE full-Rose database (RoseEngine.runtimeImpl =
 * ROSE_RUNTIME_FULL_ROSE): the literals sit in the floating literal matcher (FDR / Teddy / noodle) and their programs run
 * through the full interpreter roseRunProgram (src/rose/match.c:479-523, so class lookarounds -> CHECK_SHUFTI_* are
 * available); the regular expressions become outfix engines on their own queues exactly as hsb_compile_regex_rose lays
 * them out, and every literal program starts with CATCH_UP so that the engines' earlier matches are delivered first
 * (src/rose/block.c:345-422, catchup.c).  n_regexes = 0 gives a literal-only full-Rose database.  Report ids of
 * literals and of regexes must be disjoint. */
int hsb_compile_mixed(const char *const *literals, const unsigned *flags, const unsigned *ids,
                      const size_t *lens, const hsb_lit_ext *ext, unsigned n_literals,
                      const char *const *regexes, const unsigned *regex_flags, const unsigned *regex_ids, unsigned n_regexes,
                      unsigned mode, const hsb_lit_options *options, const hsb_lit_build_ext *build_ext,
                      unsigned regex_engine, unsigned dfa_states, hs_database_t **db, char **error);
/* Stand-in for hs_compile_multi() (/root/reference/src/hs_compile.h:450-456) for SMALL pattern sets in a
 * restricted dialect (vectorscan_b200/csrc/nfadb_writer.cpp): the whole set becomes ONE LimEx NFA inside a
 * single-outfix block-mode database -- the shape the reference compiler gives a pattern set that has no
 * useful literals -- which the unmodified reference runtime scans with hs_scan.  Flags: HSB_FLAG_CASELESS,
 * HSB_FLAG_DOTALL.  Fails (HSB_LIT_COMPILER_ERROR + message) on unsupported syntax, on patterns that match
 * the empty string and when the set needs more than 512 NFA states. */
int hsb_compile_regex_multi(const char *const *expressions, const unsigned *flags, const unsigned *ids,
                            unsigned elements, unsigned mode, hs_database_t **db, char **error);
/* Same front end, engine chosen by the caller.  HSB_REGEX_ENGINE_MCCLELLAN determinises the automaton (subset
 * construction + minimisation) and lays it out as the reference's McClellan DFA (src/nfa/mcclellan_internal.h;
 * 8-bit table for <= 256 states, 16-bit otherwise, at most 16 383 states) -- the engine the reference compiler
 * prefers for outfixes (src/grey.cpp:119).  ..._SHERMAN additionally stores states close to the start state or
 * to their parent as 32-byte Sherman states (mcclellancompile.cpp:1020-1200), 16-bit engines only. */
#define HSB_REGEX_ENGINE_LIMEX 0u
#define HSB_REGEX_ENGINE_MCCLELLAN 1u
#define HSB_REGEX_ENGINE_MCCLELLAN_SHERMAN 2u
#define HSB_REGEX_ENGINE_SHENG 3u /* DFA of <= 16 states as byte-shuffle masks (src/nfa/sheng_internal.h) */
#define HSB_REGEX_ENGINE_SHENG_WIDE 4u /* Sheng32 (<= 32 states) / Sheng64 (<= 64): the AVX512VBMI forms, 64-byte masks;
                                          only an AVX512VBMI build of the reference runtime executes them */
#define HSB_REGEX_ENGINE_MCSHENG 5u /* McClellan DFA whose hot core (<= 16 non-accepting states around the start state)
                                       steps through byte-shuffle masks: struct mcsheng, MCSHENG_NFA_8 / _16
                                       (src/nfa/mcsheng_internal.h:63-88) -- what the reference compiler builds for a DFA
                                       too large for Sheng (getDfa, src/rose/rose_build_bytecode.cpp) */
#define HSB_REGEX_ENGINE_MCCLELLAN_WIDE 6u /* McClellan with Sherman states AND wide states: runs of >= 8 single-entry states
                                              that leave the run the same way are stored as one "wide" state (16-bit
                                              engines; mcclellan_internal.h:53-61, mcclellancompile.cpp:773-838) */
#define HSB_REGEX_ENGINE_MCSHENG64 7u /* McSheng with a region of up to 64 states and 64-byte masks: struct mcsheng64,
                                         MCSHENG_64_NFA_8 / _16 (mcsheng_internal.h:94-119); only an AVX512VBMI build of the
                                         reference runtime executes it */
/* mode: HSB_MODE_BLOCK, or HSB_MODE_STREAM with a DFA engine (the engine's stream state is its state id; header
 * requiresEodCheck set when a pattern ends in \z, as hs_close_stream wants it: src/runtime.c:676-692). */
int hsb_compile_regex_multi_ex(const char *const *expressions, const unsigned *flags, const unsigned *ids,
                               unsigned elements, unsigned mode, unsigned engine, hs_database_t **db,
                               char **error);
/* The same front end for LARGE regex sets: a full-Rose database (RoseEngine.runtimeImpl = ROSE_RUNTIME_FULL_ROSE) whose
 * patterns are spread over many OUTFIX engines -- the shape hs_compile_multi gives patterns without usable literals
 * (src/rose/rose_build_bytecode.cpp:3608-3890, outfix queues :2078-2116).  `engine` = HSB_REGEX_ENGINE_MCCLELLAN[_SHERMAN]:
 * patterns join a group while its DFA stays within `dfa_states` states (0 = 1024); patterns whose own DFA is larger
 * become LimEx engines, as the reference falls back when DFA construction hits its limit.  HSB_REGEX_ENGINE_LIMEX: every
 * group is a LimEx NFA of <= 512 states.  DFAs of <= 16 states are laid out as Sheng engines.  At most 4096 engines.
 * Patterns ending in \z (end-of-data accepts) give the database an EOD program (ENGINES_EOD; at most 256 engines then).
 * Scanned by the unmodified reference runtime through roseBlockExec -> blockInitSufPQ -> roseCatchUpTo ->
 * roseBlockEodExec (src/rose/block.c:163-190,345-422, src/rose/catchup.c). */
int hsb_compile_regex_rose(const char *const *expressions, const unsigned *flags, const unsigned *ids,
                           unsigned elements, unsigned mode, unsigned engine, unsigned dfa_states,
                           hs_database_t **db, char **error);
void hsb_free_database(hs_database_t *db);
void hsb_free_error(char *error);

/* hs_database_size / hs_serialize_database / hs_deserialize_database (src/database.c:62-110,243-300):
 * byte-compatible with the reference's serialized form. */
int hsb_database_size(const hs_database_t *db, size_t *size);
int hsb_serialize_database(const hs_database_t *db, char **bytes, size_t *length);
int hsb_deserialize_database(const char *bytes, size_t length, hs_database_t **db);

unsigned hsb_crc32c(const void *p, size_t n);

#ifdef __cplusplus
}
#endif
#endif
