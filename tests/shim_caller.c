/* shim_caller.c -- TEST INFRASTRUCTURE.  A caller written against the REFERENCE's runtime API only, shaped like the scan
 * half of /root/reference/examples/simplegrep.c:150-219 (allocate scratch, hs_scan with an event handler that prints
 * each match, free scratch).  The very same source is linked once against the unmodified reference runtime
 * (oracle/_ref) and once against integration/hs_gpu_shim.c + libhs_gpu.so; tests/test_shim.py compares the outputs.
 * The database arrives serialized (argv[1]); deserialization is the reference's hs_deserialize_database in the first
 * build and the byte-compatible hsb_deserialize_database (include/hs_litdb.h) in the second. */
#include <stdio.h>
#include <stdlib.h>

typedef struct hs_database hs_database_t;
typedef struct hs_scratch hs_scratch_t;
typedef int hs_error_t;
typedef int (*match_event_handler)(unsigned int id, unsigned long long from, unsigned long long to, unsigned int flags, void *ctx);
hs_error_t hs_alloc_scratch(const hs_database_t *db, hs_scratch_t **scratch);
hs_error_t hs_free_scratch(hs_scratch_t *scratch);
hs_error_t hs_scan(const hs_database_t *db, const char *data, unsigned int length, unsigned int flags,
                   hs_scratch_t *scratch, match_event_handler onEvent, void *context);
#ifdef SHIM_BUILD
int hsb_deserialize_database(const char *bytes, size_t length, hs_database_t **db);
#define DESERIALIZE hsb_deserialize_database
#else
hs_error_t hs_deserialize_database(const char *bytes, const size_t length, hs_database_t **db);
#define DESERIALIZE hs_deserialize_database
#endif

static char *slurp(const char *path, size_t *n) {
    FILE *f = fopen(path, "rb");
    if (!f) { perror(path); exit(2); }
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    char *p = malloc((size_t)sz + 1);
    if (fread(p, 1, (size_t)sz, f) != (size_t)sz) { perror("read"); exit(2); }
    fclose(f);
    *n = (size_t)sz;
    return p;
}

static int eventHandler(unsigned int id, unsigned long long from, unsigned long long to, unsigned int flags, void *ctx) {
    (void)from; (void)flags;
    printf("Match for pattern %u at offset %llu\n", id, to);
    return ctx && --*(int *)ctx == 0;   /* optional match budget: stop the scan when it is used up */
}

int main(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s <serialized database> <input file> [max matches]\n", argv[0]); return 2; }
    size_t dn, tn;
    char *dbytes = slurp(argv[1], &dn), *text = slurp(argv[2], &tn);
    hs_database_t *db = NULL;
    if (DESERIALIZE(dbytes, dn, &db) != 0) { fprintf(stderr, "ERROR: unable to load the database\n"); return 1; }
    hs_scratch_t *scratch = NULL;
    hs_error_t rv = hs_alloc_scratch(db, &scratch);
    if (rv != 0) { fprintf(stderr, "ERROR: Unable to allocate scratch space (%d). Exiting.\n", rv); return 1; }
    int budget = argc > 3 ? atoi(argv[3]) : 0;
    rv = hs_scan(db, text, (unsigned int)tn, 0, scratch, eventHandler, budget ? &budget : NULL);
    printf("hs_scan returned %d\n", rv);
    hs_free_scratch(scratch);
    return 0;
}
