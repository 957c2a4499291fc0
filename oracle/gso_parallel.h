/*
 * gso_parallel.h — minimal pthread parallel-for for the CPU oracle (this image has no
 * libgomp).  TEST INFRASTRUCTURE ONLY.
 */
#ifndef GSO_PARALLEL_H
#define GSO_PARALLEL_H
#include <pthread.h>
#include <stdatomic.h>
#include <stdlib.h>

typedef void (*gso_body_fn)(int lo, int hi, int tid, void* ctx);

extern int gso_num_threads; /* defined in gs_oracle.c; set with gso_set_threads() */

typedef struct {
    gso_body_fn fn;
    void* ctx;
    atomic_int* next;
    int hi, chunk, tid;
} gso_worker_arg;

static void* gso_worker(void* p) {
    gso_worker_arg* a = (gso_worker_arg*)p;
    for (;;) {
        int lo = atomic_fetch_add(a->next, a->chunk);
        if (lo >= a->hi) break;
        int hi = lo + a->chunk < a->hi ? lo + a->chunk : a->hi;
        a->fn(lo, hi, a->tid, a->ctx);
    }
    return NULL;
}

/* Runs fn over [lo, hi) in dynamically scheduled chunks on gso_num_threads threads. */
static void gso_parallel_for(int lo, int hi, int chunk, gso_body_fn fn, void* ctx) {
    int nt = gso_num_threads > 0 ? gso_num_threads : 1;
    if (chunk < 1) chunk = 1;
    if (nt == 1 || hi - lo <= chunk) {
        if (hi > lo) fn(lo, hi, 0, ctx);
        return;
    }
    atomic_int next;
    atomic_init(&next, lo);
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * nt);
    gso_worker_arg* args = (gso_worker_arg*)malloc(sizeof(gso_worker_arg) * nt);
    for (int t = 0; t < nt; t++) {
        args[t].fn = fn; args[t].ctx = ctx; args[t].next = &next;
        args[t].hi = hi; args[t].chunk = chunk; args[t].tid = t;
        if (t > 0) pthread_create(&th[t], NULL, gso_worker, &args[t]);
    }
    gso_worker(&args[0]);
    for (int t = 1; t < nt; t++) pthread_join(th[t], NULL);
    free(th);
    free(args);
}
#endif
