/*
 * rdst_pool.h — a small fork-join pool standing in for rayon's global pool.
 * TEST INFRASTRUCTURE ONLY (see rdst_oracle.h).
 *
 * rayon (1.12.0, Cargo.lock:299-311, not vendored) only schedules; no sort arithmetic
 * lives in it.  What the reference uses from it is: par_chunks / par_iter / par_bridge
 * .for_each (fork-join over independent items; call sites listed in SURVEY.md §2) and
 * current_num_threads().  pool_for(n, fn, ctx) runs fn(ctx, i) for i in 0..n on the
 * pool; the caller takes part in its own job; calls nest (a task may call pool_for),
 * idle workers pick items from any active job, like rayon's work stealing.
 */
#ifndef RDST_POOL_H
#define RDST_POOL_H
#include <pthread.h>
#include <sched.h>
#include <stdatomic.h>
#include <stdlib.h>
#include <unistd.h>

typedef void (*pool_fn)(void* ctx, size_t i);

typedef struct pool_job {
    pool_fn fn;
    void* ctx;
    size_t n;
    atomic_size_t next;
    atomic_size_t done;
    atomic_int refs; /* helpers currently holding a pointer to this job (taken under g_pool_mu) */
    struct pool_job* link;
} pool_job;

#define POOL_MAX_THREADS 512
static pthread_t g_pool_threads[POOL_MAX_THREADS];
static int g_pool_size = 0; /* workers besides the caller */
static int g_pool_threads_logical = 1;
static pthread_mutex_t g_pool_mu = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t g_pool_cv = PTHREAD_COND_INITIALIZER;
static pool_job* g_pool_jobs = NULL; /* active jobs with items left */
static int g_pool_stop = 0;

static int pool_run_one(pool_job* j) {
    size_t i = atomic_fetch_add(&j->next, 1);
    if (i >= j->n) return 0;
    j->fn(j->ctx, i);
    atomic_fetch_add(&j->done, 1);
    return 1;
}

static void pool_unlink_locked(pool_job* j) {
    pool_job** pp = &g_pool_jobs;
    while (*pp) {
        if (*pp == j) {
            *pp = j->link;
            return;
        }
        pp = &(*pp)->link;
    }
}

static void* pool_worker(void* arg) {
    (void)arg;
    pthread_mutex_lock(&g_pool_mu);
    for (;;) {
        pool_job* j = g_pool_jobs;
        while (j && atomic_load(&j->next) >= j->n) j = j->link;
        if (!j) {
            if (g_pool_stop) break;
            pthread_cond_wait(&g_pool_cv, &g_pool_mu);
            continue;
        }
        atomic_fetch_add(&j->refs, 1);
        pthread_mutex_unlock(&g_pool_mu);
        while (pool_run_one(j)) {
        }
        atomic_fetch_sub(&j->refs, 1); /* last touch of j */
        pthread_mutex_lock(&g_pool_mu);
    }
    pthread_mutex_unlock(&g_pool_mu);
    return NULL;
}

static void pool_shutdown(void) {
    pthread_mutex_lock(&g_pool_mu);
    g_pool_stop = 1;
    pthread_cond_broadcast(&g_pool_cv);
    pthread_mutex_unlock(&g_pool_mu);
    for (int i = 0; i < g_pool_size; ++i) pthread_join(g_pool_threads[i], NULL);
    g_pool_size = 0;
    g_pool_stop = 0;
}

/* (re)size the pool: `threads` logical threads = caller + threads-1 workers */
static void pool_set_threads(int threads) {
    if (threads < 1) threads = 1;
    if (threads > POOL_MAX_THREADS) threads = POOL_MAX_THREADS;
    if (threads == g_pool_threads_logical && (threads == 1 || g_pool_size == threads - 1)) return;
    if (g_pool_size) pool_shutdown();
    g_pool_threads_logical = threads;
    for (int i = 0; i < threads - 1; ++i) {
        if (pthread_create(&g_pool_threads[g_pool_size], NULL, pool_worker, NULL) == 0) ++g_pool_size;
    }
}

static int pool_threads(void) { return g_pool_threads_logical; }

static void pool_for(size_t n, pool_fn fn, void* ctx) {
    if (n == 0) return;
    if (n == 1 || g_pool_size == 0) {
        for (size_t i = 0; i < n; ++i) fn(ctx, i);
        return;
    }
    pool_job j;
    j.fn = fn;
    j.ctx = ctx;
    j.n = n;
    atomic_init(&j.next, 0);
    atomic_init(&j.done, 0);
    atomic_init(&j.refs, 0);
    pthread_mutex_lock(&g_pool_mu);
    j.link = g_pool_jobs;
    g_pool_jobs = &j;
    pthread_cond_broadcast(&g_pool_cv);
    pthread_mutex_unlock(&g_pool_mu);
    while (pool_run_one(&j)) {
    }
    pthread_mutex_lock(&g_pool_mu);
    pool_unlink_locked(&j);
    pthread_mutex_unlock(&g_pool_mu);
    /* items claimed by other workers may still be running; help other jobs meanwhile */
    while (atomic_load(&j.done) < n || atomic_load(&j.refs) != 0) {
        pool_job* o = NULL;
        pthread_mutex_lock(&g_pool_mu);
        o = g_pool_jobs;
        while (o && atomic_load(&o->next) >= o->n) o = o->link;
        if (o) atomic_fetch_add(&o->refs, 1);
        pthread_mutex_unlock(&g_pool_mu);
        if (o) {
            pool_run_one(o); /* help with one item of some other job, then re-check ours */
            atomic_fetch_sub(&o->refs, 1);
        } else {
            sched_yield();
        }
    }
}
#endif
