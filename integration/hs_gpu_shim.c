/* hs_gpu_shim.c -- link-time drop-in: the reference's block-mode runtime entry points, by their own names and
 * prototypes (/root/reference/src/hs_runtime.h:479-482 hs_scan, :555-556 hs_alloc_scratch, :576 hs_clone_scratch,
 * :593 hs_scratch_size, :609 hs_free_scratch, :522-527 hs_scan_vector; /root/reference/src/hs_runtime.h:125-129
 * match_event_handler), forwarding
 * to libhs_gpu.so.  An application that calls hs_alloc_scratch / hs_scan keeps its source unchanged and links this
 * library ahead of libhs for these symbols; hs_compile*, hs_(de)serialize_database, hs_free_database stay the
 * reference's own.  What src/dispatcher.c:50-87 does for ISA variants, done for one more backend.
 *
 * Build:   cc -shared -fPIC integration/hs_gpu_shim.c -Iinclude [-I<reference>/src -DHS_SHIM_WITH_REFERENCE_HEADER]
 *             -Lvectorscan_b200 -lhs_gpu -o libhs_gpushim.so
 * With -DHS_SHIM_WITH_REFERENCE_HEADER the reference's own hs_runtime.h is included, so a prototype that did not match
 * the reference's declaration would not compile (tests/test_shim.py does exactly that). */
#ifdef HS_SHIM_WITH_REFERENCE_HEADER
#include "hs_runtime.h" /* the reference's public header: declares every function defined below */
#endif
#include "hs_gpu.h"

#ifndef HS_CDECL
#define HS_CDECL
#endif
#ifndef HS_SHIM_WITH_REFERENCE_HEADER
typedef struct hs_scratch hs_scratch_t; /* opaque to the application, as in src/hs_runtime.h:48 */
typedef struct hs_stream hs_stream_t;   /* src/hs_common.h:48 */
#endif

hs_error_t HS_CDECL hs_alloc_scratch(const hs_database_t *db, hs_scratch_t **scratch) {
    /* grows an existing scratch for one more database, as hs_alloc_scratch does (src/scratch.c:244-389) */
    return hs_gpu_alloc_scratch(db, (hs_gpu_scratch_t **)scratch);
}

hs_error_t HS_CDECL hs_clone_scratch(const hs_scratch_t *src, hs_scratch_t **dest) {
    return hs_gpu_clone_scratch((const hs_gpu_scratch_t *)src, (hs_gpu_scratch_t **)dest);
}

hs_error_t HS_CDECL hs_scratch_size(const hs_scratch_t *scratch, size_t *scratch_size) {
    size_t host_bytes = 0, device_bytes = 0;
    const hs_error_t rv = hs_gpu_scratch_size((const hs_gpu_scratch_t *)scratch, &host_bytes, &device_bytes);
    if (rv == HS_SUCCESS && scratch_size) *scratch_size = host_bytes + device_bytes;
    return rv;
}

hs_error_t HS_CDECL hs_free_scratch(hs_scratch_t *scratch) {
    return hs_gpu_free_scratch((hs_gpu_scratch_t *)scratch);
}

hs_error_t HS_CDECL hs_scan(const hs_database_t *db, const char *data, unsigned int length, unsigned int flags,
                            hs_scratch_t *scratch, match_event_handler onEvent, void *context) {
    return hs_gpu_scan(db, data, length, flags, (hs_gpu_scratch_t *)scratch, onEvent, context);
}

/* src/runtime.c:1106-1174: vectored-mode databases only (HS_DB_MODE_ERROR otherwise, :1129-1131); the buffers are
 * consecutive blocks of one stream.  (hs_gpu_scan_vector also serves block-mode databases, as a batch of independent
 * scans -- an extension this entry point does not expose, so that it behaves exactly as the reference's.) */
hs_error_t HS_CDECL hs_scan_vector(const hs_database_t *db, const char *const *data, const unsigned int *length,
                                   unsigned int count, unsigned int flags, hs_scratch_t *scratch,
                                   match_event_handler onEvent, void *context) {
    unsigned int mode = 0;
    const hs_error_t rv = hs_gpu_database_mode(db, &mode);
    if (rv != HS_SUCCESS) return rv;
    if (mode != HS_GPU_MODE_VECTORED) return HS_DB_MODE_ERROR;
    return hs_gpu_scan_vector(db, data, length, count, flags, (hs_gpu_scratch_t *)scratch, onEvent, context);
}

/* Stream mode (pure-literal HS_MODE_STREAM databases): src/hs_runtime.h:148-293. */
hs_error_t HS_CDECL hs_open_stream(const hs_database_t *db, unsigned int flags, hs_stream_t **stream) {
    return hs_gpu_open_stream(db, flags, (hs_gpu_stream_t **)stream);
}

hs_error_t HS_CDECL hs_scan_stream(hs_stream_t *id, const char *data, unsigned int length, unsigned int flags,
                                   hs_scratch_t *scratch, match_event_handler onEvent, void *ctxt) {
    return hs_gpu_scan_stream((hs_gpu_stream_t *)id, data, length, flags, (hs_gpu_scratch_t *)scratch, onEvent, ctxt);
}

hs_error_t HS_CDECL hs_close_stream(hs_stream_t *id, hs_scratch_t *scratch, match_event_handler onEvent, void *ctxt) {
    return hs_gpu_close_stream((hs_gpu_stream_t *)id, (hs_gpu_scratch_t *)scratch, onEvent, ctxt);
}

hs_error_t HS_CDECL hs_reset_stream(hs_stream_t *id, unsigned int flags, hs_scratch_t *scratch, match_event_handler onEvent,
                                    void *context) {
    return hs_gpu_reset_stream((hs_gpu_stream_t *)id, flags, (hs_gpu_scratch_t *)scratch, onEvent, context);
}

hs_error_t HS_CDECL hs_copy_stream(hs_stream_t **to_id, const hs_stream_t *from_id) {
    return hs_gpu_copy_stream((hs_gpu_stream_t **)to_id, (const hs_gpu_stream_t *)from_id);
}
