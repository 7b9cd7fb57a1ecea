/* oracle/cpu_bench.c -- TEST / BASELINE INFRASTRUCTURE ONLY (never on the product path).
 *
 * hsbench-shaped block-mode timing loop over the UNMODIFIED reference runtime (oracle/_ref/libhs_ref_*.so,
 * loaded with dlopen).  Shape of tools/hsbench/main.cpp:
 *   - one persistent worker per thread, pinned to one CPU of the process's affinity mask (main.cpp:154), each
 *     with its own cloned scratch (engine_hyperscan.cpp:131-163) and a counting callback (:89-96);
 *   - every worker scans the WHOLE corpus, `repeats` times, one hs_scan per buffer (benchBlock, main.cpp:486-512);
 *   - workers are created, pinned and parked on a barrier BEFORE the clock starts; the clock (CLOCK_MONOTONIC)
 *     brackets the passes only: release barrier -> passes -> completion barrier (main.cpp:560-590);
 *   - throughput = threads x repeats x corpus bytes / seconds (main.cpp:704-708).
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
    const char *flat; const uint64_t *off; uint32_t count; int repeats; int cpu;
    pthread_barrier_t *start, *done;
    uint64_t matches; int err;
    char pad[64];
};

static void *worker(void *p) {
    struct job *j = (struct job *)p;
    if (j->cpu >= 0) {
        cpu_set_t set; CPU_ZERO(&set); CPU_SET(j->cpu, &set);
        pthread_setaffinity_np(pthread_self(), sizeof(set), &set);
    }
    /* one untimed pass over a little of the corpus: page in the runtime's code and this worker's scratch */
    uint64_t warm = 0;
    for (uint32_t b = 0; b < j->count && b < 64; b++)
        j->a->scan(j->db, j->flat + j->off[b], (unsigned)(j->off[b + 1] - j->off[b]), 0, j->scratch, count_cb, &warm);
    pthread_barrier_wait(j->start);
    for (int r = 0; r < j->repeats; r++)
        for (uint32_t b = 0; b < j->count; b++) {
            int rv = j->a->scan(j->db, j->flat + j->off[b], (unsigned)(j->off[b + 1] - j->off[b]), 0, j->scratch,
                                count_cb, &j->matches);
            if (rv) j->err = rv;
        }
    pthread_barrier_wait(j->done);
    return NULL;
}

/* Every one of `threads` pinned workers scans buffers [0,count) `repeats` times.  Returns the seconds between
 * the release of the start barrier and the completion barrier (<0 on error); *matches = matches counted by all
 * workers over all repeats.  Bytes scanned = threads x repeats x off[count]. */
double cpu_bench_run(const char *ref_lib, const char *serialized, size_t ser_len, const char *flat,
                     const uint64_t *off, uint32_t count, int threads, int repeats, uint64_t *matches) {
    struct api a;
    if (load_api(ref_lib, &a)) return -1.0;
    void *db = NULL, *proto = NULL;
    if (a.deser(serialized, ser_len, &db) || a.alloc(db, &proto)) return -2.0;
    if (threads < 1) threads = 1;
    if (repeats < 1) repeats = 1;
    /* CPUs this process may run on (a container's mask need not start at 0) */
    cpu_set_t avail; CPU_ZERO(&avail);
    int ncpu = 0, cpus[1024];
    if (sched_getaffinity(0, sizeof(avail), &avail) == 0)
        for (int c = 0; c < CPU_SETSIZE && ncpu < 1024; c++) if (CPU_ISSET(c, &avail)) cpus[ncpu++] = c;
    struct job *jobs = calloc((size_t)threads, sizeof(*jobs));
    pthread_t *tids = calloc((size_t)threads, sizeof(*tids));
    pthread_barrier_t start, done;
    pthread_barrier_init(&start, NULL, (unsigned)threads + 1);
    pthread_barrier_init(&done, NULL, (unsigned)threads + 1);
    for (int t = 0; t < threads; t++) {
        jobs[t].a = &a; jobs[t].db = db; jobs[t].flat = flat; jobs[t].off = off; jobs[t].count = count;
        jobs[t].repeats = repeats; jobs[t].cpu = ncpu ? cpus[t % ncpu] : -1;
        jobs[t].start = &start; jobs[t].done = &done;
        if (a.clone(proto, &jobs[t].scratch)) return -3.0;
    }
    for (int t = 0; t < threads; t++) pthread_create(&tids[t], NULL, worker, &jobs[t]);
    struct timespec t0, t1;
    pthread_barrier_wait(&start);
    clock_gettime(CLOCK_MONOTONIC, &t0);
    pthread_barrier_wait(&done);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    for (int t = 0; t < threads; t++) pthread_join(tids[t], NULL);
    uint64_t m = 0; int err = 0;
    for (int t = 0; t < threads; t++) { m += jobs[t].matches; err |= jobs[t].err; a.free_scratch(jobs[t].scratch); }
    a.free_scratch(proto); a.free_db(db);
    pthread_barrier_destroy(&start); pthread_barrier_destroy(&done);
    free(jobs); free(tids);
    if (matches) *matches = m;
    if (err) return -4.0;
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
