/* oracle/cpu_bench.c -- TEST / BASELINE INFRASTRUCTURE ONLY (never on the product path).
 *
 * hsbench-style block-mode timing loop over the UNMODIFIED reference runtime (oracle/_ref/libhs_ref_*.so,
 * loaded with dlopen): the loop tools/hsbench/main.cpp:486-512 runs -- every thread owns a cloned scratch
 * and calls hs_scan once per buffer with a counting callback (engine_hyperscan.cpp:89-96) -- timed with
 * CLOCK_MONOTONIC around the whole pass, threads pinned one per core (main.cpp:154).
 * No matching logic lives here. */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <pthread.h>
#include <sched.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

typedef int (*match_cb)(unsigned, unsigned long long, unsigned long long, unsigned, void *);
typedef int (*fn_deser)(const char *, size_t, void **);
typedef int (*fn_alloc)(const void *, void **);
typedef int (*fn_clone)(const void *, void **);
typedef int (*fn_free)(void *);
typedef int (*fn_scan)(const void *, const char *, unsigned, unsigned, void *, match_cb, void *);

struct api { void *h; fn_deser deser; fn_alloc alloc; fn_clone clone; fn_free free_scratch, free_db; fn_scan scan; };

static int load_api(const char *lib, struct api *a) {
    a->h = dlopen(lib, RTLD_NOW | RTLD_LOCAL);
    if (!a->h) { fprintf(stderr, "cpu_bench: %s\n", dlerror()); return -1; }
    a->deser = (fn_deser)dlsym(a->h, "hs_deserialize_database");
    a->alloc = (fn_alloc)dlsym(a->h, "hs_alloc_scratch");
    a->clone = (fn_clone)dlsym(a->h, "hs_clone_scratch");
    a->free_scratch = (fn_free)dlsym(a->h, "hs_free_scratch");
    a->free_db = (fn_free)dlsym(a->h, "hs_free_database");
    a->scan = (fn_scan)dlsym(a->h, "hs_scan");
    return (a->deser && a->alloc && a->clone && a->free_scratch && a->free_db && a->scan) ? 0 : -1;
}

static int count_cb(unsigned id, unsigned long long from, unsigned long long to, unsigned flags, void *ctx) {
    (void)id; (void)from; (void)to; (void)flags;
    ++*(uint64_t *)ctx;
    return 0;
}

struct job {
    const struct api *a; const void *db; void *scratch;
    const char *flat; const uint64_t *off; uint32_t first, last; int repeats; int cpu;
    uint64_t matches; int err;
};

static void *worker(void *p) {
    struct job *j = (struct job *)p;
    if (j->cpu >= 0) {
        cpu_set_t set; CPU_ZERO(&set); CPU_SET(j->cpu, &set);
        pthread_setaffinity_np(pthread_self(), sizeof(set), &set);
    }
    for (int r = 0; r < j->repeats; r++)
        for (uint32_t b = j->first; b < j->last; b++) {
            int rv = j->a->scan(j->db, j->flat + j->off[b], (unsigned)(j->off[b + 1] - j->off[b]), 0, j->scratch,
                                count_cb, &j->matches);
            if (rv) j->err = rv;
        }
    return NULL;
}

/* Scans buffers [0,count) `repeats` times with `threads` threads (buffers split in contiguous, byte-balanced
 * ranges).  Returns seconds for the whole pass (<0 on error); *matches = matches counted over all repeats. */
double cpu_bench_run(const char *ref_lib, const char *serialized, size_t ser_len, const char *flat,
                     const uint64_t *off, uint32_t count, int threads, int repeats, uint64_t *matches) {
    struct api a;
    if (load_api(ref_lib, &a)) return -1.0;
    void *db = NULL, *proto = NULL;
    if (a.deser(serialized, ser_len, &db) || a.alloc(db, &proto)) return -2.0;
    if (threads < 1) threads = 1;
    struct job *jobs = calloc((size_t)threads, sizeof(*jobs));
    pthread_t *tids = calloc((size_t)threads, sizeof(*tids));
    uint64_t total = off[count], per = total / (uint64_t)threads + 1;
    uint32_t b = 0;
    for (int t = 0; t < threads; t++) {
        jobs[t].a = &a; jobs[t].db = db; jobs[t].flat = flat; jobs[t].off = off; jobs[t].repeats = repeats;
        jobs[t].cpu = t; jobs[t].first = b;
        uint64_t limit = (uint64_t)(t + 1) * per;
        while (b < count && off[b + 1] <= limit) b++;
        if (t == threads - 1) b = count;
        jobs[t].last = b;
        if (a.clone(proto, &jobs[t].scratch)) return -3.0;
    }
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int t = 0; t < threads; t++) pthread_create(&tids[t], NULL, worker, &jobs[t]);
    for (int t = 0; t < threads; t++) pthread_join(tids[t], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    uint64_t m = 0; int err = 0;
    for (int t = 0; t < threads; t++) { m += jobs[t].matches; err |= jobs[t].err; a.free_scratch(jobs[t].scratch); }
    a.free_scratch(proto); a.free_db(db);
    free(jobs); free(tids);
    if (matches) *matches = m;
    if (err) return -4.0;
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
