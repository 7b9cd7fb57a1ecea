/* hs_gpu.h -- C ABI of libhs_gpu.so: the B200 (sm_100a) block-mode scanning runtime for Vectorscan
 * databases.  Plain pointers and sizes only; no torch / CUDA types in any signature.
 *
 * Each entry point replaces, for block-mode databases, the reference runtime interface cited beside
 * it (paths under /root/reference).  Databases are the reference's own hs_database_t bytecode blobs
 * (from the unchanged host compiler, from hs_deserialize_database, or from hs_litdb.h).  Error codes are
 * the reference's (src/hs_common.h:474-584).  There is no CPU fallback: without a CUDA device every
 * scanning call returns HS_ARCH_ERROR.
 */
#ifndef HS_GPU_H
#define HS_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

struct hs_database;
#if !defined(HS_COMMON_H_) && !defined(HSB_HS_DATABASE_T) /* the reference's hs_common.h, or our other header, may have declared it */
#define HSB_HS_DATABASE_T
typedef struct hs_database hs_database_t; /* src/hs_common.h:56 */
#endif
struct hs_gpu_scratch;
typedef struct hs_gpu_scratch hs_gpu_scratch_t; /* stands where hs_scratch_t stands, src/hs_common.h:466 */
struct hs_gpu_corpus;
typedef struct hs_gpu_corpus hs_gpu_corpus_t;
/* stands where hs_stream_t stands, src/hs_common.h:48 */
typedef struct hs_gpu_stream hs_gpu_stream_t;

typedef int hs_error_t;
#define HS_SUCCESS 0
#define HS_INVALID (-1)
#define HS_NOMEM (-2)
#define HS_SCAN_TERMINATED (-3)
#define HS_COMPILER_ERROR (-4)
#define HS_DB_VERSION_ERROR (-5)
#define HS_DB_PLATFORM_ERROR (-6)
#define HS_DB_MODE_ERROR (-7)
#define HS_BAD_ALIGN (-8)
#define HS_BAD_ALLOC (-9)
#define HS_SCRATCH_IN_USE (-10)
#define HS_ARCH_ERROR (-11)
#define HS_INSUFFICIENT_SPACE (-12)
#define HS_UNKNOWN_ERROR (-13)

/* src/hs_runtime.h:125-129.  from is always 0 (no SOM), flags always 0; non-zero return stops delivery
 * and makes the scan call return HS_SCAN_TERMINATED. */
typedef int (*match_event_handler)(unsigned int id, unsigned long long from, unsigned long long to,
                                   unsigned int flags, void *context);

/* One reported match of a batch scan: sorted by (buffer, to, id). */
typedef struct hs_gpu_match {
    uint32_t buffer; /* index into the data[] array of the call */
    uint32_t id;     /* pattern id */
    uint64_t to;     /* offset one past the last byte of the match, relative to its buffer */
} hs_gpu_match_t;

/* Per-call counters and device timings (CUDA events on the scratch's stream). */
typedef struct hs_gpu_stats {
    uint64_t bytes;        /* sum of buffer lengths scanned */
    uint64_t buffers;
    uint64_t matches;      /* delivered (after dedupe / single-match resolution) */
    uint64_t raw_matches;  /* written to the device match arena */
    uint64_t candidates;   /* positions that survived the front-end filter */
    uint64_t confirms;     /* literal-confirm table probes */
    float h2d_ms;          /* host->device copy of the inputs (0 for a resident corpus) */
    float scan_ms;         /* front end + confirm + program + arena write */
    float sort_ms;         /* device sort / compaction of the arena */
    float d2h_ms;          /* device->host copy of the matches */
    uint32_t scan_launches; /* k_scan launches (1 + arena_retries) */
    uint32_t arena_retries; /* re-runs forced by match arena growth */
    float kernel_ms;        /* the (last) k_scan launch alone, CUDA events tight around it */
    uint32_t kernel_launches; /* kernels of this library launched by the call (CUB's are not counted) */
    float train_ms;           /* ordered path: first kernel's start to last kernel's end, gaps between members included */
    uint32_t front_end;       /* dominant kernel of the call: 1 k_scan, 2 k_turbo, 3 k_turbo2, 4 k_limex, 5 k_mcclellan, 6 several engines */
} hs_gpu_stats_t;

/* hs_alloc_scratch(db, &scratch)                       src/hs_runtime.h:555-556, src/scratch.c:244-389
 * *scratch == NULL allocates; otherwise the scratch is grown to also serve this database.  The database
 * is parsed, checked (block mode, supported engines and Rose opcodes only -> else HS_ARCH_ERROR /
 * HS_DB_MODE_ERROR) and uploaded to the device once. */
hs_error_t hs_gpu_alloc_scratch(const hs_database_t *db, hs_gpu_scratch_t **scratch);
/* hs_clone_scratch / hs_scratch_size / hs_free_scratch  src/hs_runtime.h:576,593,609 */
hs_error_t hs_gpu_clone_scratch(const hs_gpu_scratch_t *src, hs_gpu_scratch_t **dest);
hs_error_t hs_gpu_scratch_size(const hs_gpu_scratch_t *scratch, size_t *host_bytes, size_t *device_bytes);
hs_error_t hs_gpu_free_scratch(hs_gpu_scratch_t *scratch);

/* hs_scan(db, data, length, flags, scratch, onEvent, context)      src/hs_runtime.h:479-482, src/runtime.c:316-475
 * One buffer.  Matches are delivered in (to, id) order after the device pass completes. */
hs_error_t hs_gpu_scan(const hs_database_t *db, const char *data, unsigned int length, unsigned int flags,
                       hs_gpu_scratch_t *scratch, match_event_handler onEvent, void *context);

/* On a BLOCK-mode database: a batch of `count` INDEPENDENT block-mode scans -- the loop of hs_scan calls that
 * examples/pcapscan.cc:305-316 and tools/hsbench/main.cpp:486-512 run -- with the parameter list of hs_scan_vector
 * (src/hs_runtime.h:522-527).  Offsets are per buffer.  Matches are delivered buffer by buffer in index order.  A
 * non-zero return of the callback ends the deliveries of THAT buffer only -- as it would end that packet's hs_scan in the
 * loop -- and the call returns HS_SCAN_TERMINATED after the remaining buffers have been delivered.
 * On a VECTORED-mode database (HS_MODE_VECTORED, pure-literal databases): the reference's hs_scan_vector itself
 * (src/runtime.c:1106-1174) -- the buffers are consecutive blocks of ONE stream: a match may straddle blocks, offsets
 * run on across them, single-match ids stay exhausted to the end of the call; matches are delivered in (to, id) order
 * and hs_gpu_match_t.buffer is 0.  The stream must be shorter than 4 GiB (else HS_INVALID). */
hs_error_t hs_gpu_scan_vector(const hs_database_t *db, const char *const *data, const unsigned int *length,
                              unsigned int count, unsigned int flags, hs_gpu_scratch_t *scratch,
                              match_event_handler onEvent, void *context);
/* Same, with one callback context per buffer (contexts[i] may differ; NULL = no context). */
hs_error_t hs_gpu_scan_batch(const hs_database_t *db, const char *const *data, const unsigned int *length,
                             unsigned int count, unsigned int flags, hs_gpu_scratch_t *scratch,
                             match_event_handler onEvent, void *const *contexts);
/* Same, without callbacks: *matches points at the scratch-owned, sorted match array (valid until the next
 * call on this scratch), *n_matches is its length. */
hs_error_t hs_gpu_scan_vector_ex(const hs_database_t *db, const char *const *data, const unsigned int *length,
                                 unsigned int count, unsigned int flags, hs_gpu_scratch_t *scratch,
                                 const hs_gpu_match_t **matches, uint64_t *n_matches);

/* ---- stream mode: HS_MODE_STREAM pure-literal databases, and single-outfix databases holding ONE DFA engine ----------
 * (McClellan, McSheng, Sheng; not LimEx: HS_ARCH_ERROR at hs_gpu_alloc_scratch; their end-of-data matches are reported by
 * hs_gpu_close_stream when it is given a scratch and a handler, as by hs_close_stream)
 * hs_open_stream / hs_scan_stream / hs_close_stream / hs_reset_stream / hs_copy_stream  (src/hs_runtime.h:148-293,
 * src/runtime.c:543-580, 980-996, 998-1033, 1035-1060, 711-737): same parameter lists and error codes.  The stream object
 * lives on the host: its offset, its last historyRequired bytes, the exhaustion keys of the single-match reports it has
 * delivered.  Offsets in callbacks are stream offsets; a literal may straddle any number of blocks.  A non-zero callback
 * return terminates the stream: this and every later call on it return HS_SCAN_TERMINATED. */
hs_error_t hs_gpu_open_stream(const hs_database_t *db, unsigned int flags, hs_gpu_stream_t **stream);
hs_error_t hs_gpu_scan_stream(hs_gpu_stream_t *id, const char *data, unsigned int length, unsigned int flags,
                              hs_gpu_scratch_t *scratch, match_event_handler onEvent, void *context);
hs_error_t hs_gpu_close_stream(hs_gpu_stream_t *id, hs_gpu_scratch_t *scratch, match_event_handler onEvent, void *context);
hs_error_t hs_gpu_reset_stream(hs_gpu_stream_t *id, unsigned int flags, hs_gpu_scratch_t *scratch,
                               match_event_handler onEvent, void *context);
hs_error_t hs_gpu_copy_stream(hs_gpu_stream_t **to_id, const hs_gpu_stream_t *from_id);
/* The batch form, which is what fills a GPU: ONE block for each of `count` DIFFERENT streams of one database (the next
 * packet of many flows, examples/pcapscan.cc:290 scanStreams) in one device pass.  contexts[i] goes to the callbacks
 * of stream ids[i] (contexts may be NULL).  A stream may appear once per call (HS_INVALID otherwise).  A non-zero
 * callback return terminates ITS stream only; the call then returns HS_SCAN_TERMINATED after serving the others. */
hs_error_t hs_gpu_scan_streams(hs_gpu_stream_t *const *ids, const char *const *data, const unsigned int *length,
                               unsigned int count, unsigned int flags, hs_gpu_scratch_t *scratch,
                               match_event_handler onEvent, void *const *contexts);

/* Device-resident corpora: upload once, scan many times (what hsbench does with its in-memory corpus,
 * tools/hsbench/main.cpp:486-512).  The corpus is `count` buffers; `flat` holds them back to back and
 * offsets[i]..offsets[i+1] delimits buffer i (offsets has count+1 entries, offsets[0] == 0). */
hs_error_t hs_gpu_corpus_create(const char *flat, const uint64_t *offsets, unsigned int count,
                                hs_gpu_corpus_t **corpus);
hs_error_t hs_gpu_corpus_free(hs_gpu_corpus_t *corpus);
/* Scan a resident corpus; matches stay on the device until fetched.  Only device work is done.  With a vectored-mode
 * database the corpus' buffers are the consecutive blocks of one stream (see hs_gpu_scan_vector). */
hs_error_t hs_gpu_scan_corpus(const hs_database_t *db, const hs_gpu_corpus_t *corpus, hs_gpu_scratch_t *scratch,
                              uint64_t *n_matches);
/* Scan n resident corpora back to back as ONE logical batch (the hsbench loop over several in-memory blocks): the
 * kernels of all members are queued on the scratch's stream before the host waits once; buffer indices of the matches
 * continue from corpus to corpus.  Needs a database the sort-free path serves (plain-report literal sets; otherwise
 * HS_ARCH_ERROR, scan the corpora one by one). */
hs_error_t hs_gpu_scan_corpora(const hs_database_t *db, const hs_gpu_corpus_t *const *corpora, unsigned int n,
                               hs_gpu_scratch_t *scratch, uint64_t *n_matches);
hs_error_t hs_gpu_fetch_matches(hs_gpu_scratch_t *scratch, const hs_gpu_match_t **matches, uint64_t *n_matches);

/* RoseEngine.mode of a database (src/rose/rose_internal.h:  HS_MODE_BLOCK 1, HS_MODE_STREAM 2, HS_MODE_VECTORED 4) */
#define HS_GPU_MODE_BLOCK 1u
#define HS_GPU_MODE_STREAM 2u
#define HS_GPU_MODE_VECTORED 4u
hs_error_t hs_gpu_database_mode(const hs_database_t *db, unsigned int *mode);

hs_error_t hs_gpu_last_stats(const hs_gpu_scratch_t *scratch, hs_gpu_stats_t *stats);
/* Adds, for every match of the last device scan, 1 to device_hist_u64[id] (ids >= n_ids are skipped).
 * device_hist_u64 is a DEVICE pointer to n_ids uint64 counters owned by the caller -- the vector the
 * multi-GPU driver all-reduces over NCCL. */
hs_error_t hs_gpu_histogram(hs_gpu_scratch_t *scratch, void *device_hist_u64, uint32_t n_ids);

/* Host memory the library can DMA from without a staging copy (cudaHostAlloc). */
void *hs_gpu_host_alloc(size_t bytes);
void hs_gpu_host_free(void *p);

/* Runtime knobs.  front end: 0 = auto, 1 = "native" (the database's own FDR/Teddy tables and FDRConfirm
 * structures), 2 = "turbo" (tables re-encoded at upload for shared memory; identical results). */
hs_error_t hs_gpu_set_option(hs_gpu_scratch_t *scratch, const char *name, long value);
/* Test instrumentation: after hs_gpu_set_option(scratch, "count_ops", 1) the device interpreter of Rose programs
 * (roseRunProgram_l, src/rose/program_runtime.c:3083-3504) counts every instruction it executes by opcode
 * (enum RoseInstructionCode, src/rose/rose_program.h:48-212); this copies the first n (<= 128) counters out. */
/* Host-only (no device needed): how often each Rose opcode occurs in the programs the database's literal matcher /
 * engine can start -- the census hs_gpu_alloc_scratch runs, as a histogram. */
hs_error_t hs_gpu_database_opcodes(const hs_database_t *db, unsigned long long *hist, unsigned int n);
hs_error_t hs_gpu_opcode_counts(hs_gpu_scratch_t *scratch, unsigned long long *counts, unsigned int n);
int hs_gpu_device_count(void);
hs_error_t hs_gpu_set_device(int device);
const char *hs_gpu_version(void);

/* Byte-class primitives over a device-staged buffer, with the reference's return convention
 * (offset of the first byte in class, `len` if none; reverse variants: offset of the last, -1 if none):
 * shuftiExec/rshuftiExec  src/nfa/shufti.h ; truffleExec/rtruffleExec  src/nfa/truffle.h ;
 * vermicelliExec/rvermicelliExec/vermicelliDoubleExec  src/nfa/vermicelli.hpp */
long long hs_gpu_shufti(const uint8_t lo[16], const uint8_t hi[16], const uint8_t *buf, size_t len, int reverse);
long long hs_gpu_truffle(const uint8_t mask1[16], const uint8_t mask2[16], const uint8_t *buf, size_t len, int reverse);
long long hs_gpu_vermicelli(uint8_t c, int nocase, const uint8_t *buf, size_t len, int reverse, int negate);
long long hs_gpu_vermicelli_double(uint8_t c1, uint8_t c2, int nocase, const uint8_t *buf, size_t len);
/* shuftiDoubleExec (src/nfa/shufti.h, shufti_simd.hpp:196-285): first i with a pair bucket admitting byte i under
 * (lo1, hi1) and byte i + 1 under (lo2, hi2) -- masks with CLEAR bits for members; a trailing first byte counts. */
long long hs_gpu_shufti_double(const uint8_t lo1[16], const uint8_t hi1[16], const uint8_t lo2[16], const uint8_t hi2[16],
                               const uint8_t *buf, size_t len);
/* Measurement: stage the data on the device once, then search it `repeats` times (kind 0 shufti, 1 truffle,
 * 2 vermicelli, 3 negated vermicelli, 4 double vermicelli, 5 double shufti; masks = up to four 16-byte masks);
 * *kernel_ms = mean kernel time (CUDA events).  The shape of /root/reference/benchmarks/benchmarks.cpp:112-250. */
int hs_gpu_accel_stage(const uint8_t *buf, size_t len);
long long hs_gpu_accel_search_staged(int kind, const uint8_t *masks, uint8_t c1, uint8_t c2, int nocase, int reverse,
                                     int repeats, float *kernel_ms);

/* hs_gpu_shufti without any device allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same search, same masks, same return convention as hs_gpu_shufti (*result = offset of the first byte in
 * the class, `len` if none; reverse: offset of the last, -1 if none); `buf` is host memory.  The scratch the search needs
 * (result cell, class table, staged copy of the data) is a fixed file-scope device array of HS_GPU_NOALLOC_MAX_BYTES +
 * 1328 bytes that every GPU loading the library holds until the process exits; the call makes no cudaMalloc / cudaFree /
 * allocating library call.  len > HS_GPU_NOALLOC_MAX_BYTES -> HS_INSUFFICIENT_SPACE (and a line on stderr); no usable
 * device or cudaGetSymbolAddress failure -> HS_ARCH_ERROR.  Every call shares the one array (it keeps the previous call's
 * contents; each call rewrites what it reads): calls are serialised inside the process and finish before returning, and
 * anything that made two of them overlap on the GPU would race on it. */
#define HS_GPU_NOALLOC_MAX_BYTES (1u << 20)
hs_error_t hs_gpu_shufti_noalloc(const uint8_t lo[16], const uint8_t hi[16], const uint8_t *buf, size_t len, int reverse,
                                 long long *result);

#ifdef __cplusplus
}
#endif
#endif
