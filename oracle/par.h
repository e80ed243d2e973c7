/* par.h — tiny pthread parallel-for used by the oracle batch runners (contiguous slice per thread).
 * TEST INFRASTRUCTURE ONLY (see oracle.h). */
#ifndef POPORON_ORACLE_PAR_H
#define POPORON_ORACLE_PAR_H

#include <pthread.h>
#include <stddef.h>
#include <stdlib.h>

typedef void (*par_body_t)(void *arg, size_t lo, size_t hi);

typedef struct {
    par_body_t body;
    void *arg;
    size_t lo, hi;
} par_slice_t;

static void *par_tramp(void *v)
{
    par_slice_t *s = (par_slice_t *)v;
    s->body(s->arg, s->lo, s->hi);
    return NULL;
}

static void par_for(size_t n, int threads, par_body_t body, void *arg)
{
    if (threads <= 1 || n < 2) {
        body(arg, 0, n);
        return;
    }
    if ((size_t)threads > n)
        threads = (int)n;
    pthread_t *tid = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)threads);
    par_slice_t *sl = (par_slice_t *)malloc(sizeof(par_slice_t) * (size_t)threads);
    size_t chunk = (n + (size_t)threads - 1) / (size_t)threads;
    int started = 0;
    for (int t = 0; t < threads; t++) {
        size_t lo = (size_t)t * chunk, hi = lo + chunk;
        if (lo >= n)
            break;
        if (hi > n)
            hi = n;
        sl[t].body = body;
        sl[t].arg = arg;
        sl[t].lo = lo;
        sl[t].hi = hi;
        if (pthread_create(&tid[t], NULL, par_tramp, &sl[t]) != 0) {
            body(arg, lo, hi); /* fall back to running the slice inline */
            tid[t] = 0;
            sl[t].body = NULL;
        }
        started = t + 1;
    }
    for (int t = 0; t < started; t++)
        if (sl[t].body)
            pthread_join(tid[t], NULL);
    free(tid);
    free(sl);
}

#endif
