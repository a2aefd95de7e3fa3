/*
 * oracle/cpu_bench.c -- TEST/BENCH INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Times the reference's CPU path for the seal/open hot loop on host cores:
 * what @hyperswarm/secret-stream's `_write` does per frame (allocate len+20,
 * copy the chunk to offset 4, write the uint24-LE prefix, secretstream push in
 * place) and what `_incoming` does (pull in place) -- reached from
 * /root/reference/src/provider.ts:248 and :174.  The push/pull functions are
 * passed in as pointers so the same loop can drive
 *   - real libsodium (PyNaCl's _sodium.abi3.so, resolved by ctypes in the
 *     Python caller; the same C library sodium-native 4.1.1 wraps), or
 *   - the restatement in secretstream_ref.c (through the adapters below).
 * Streams are statically partitioned over `threads` pthreads; each thread
 * runs its streams' whole frame chain with no cross-thread synchronisation,
 * which is an upper bound on what the single-threaded Node reference can do.
 */
#define _GNU_SOURCE
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

typedef int (*push_fn)(void *state, unsigned char *c, unsigned long long *clen, const unsigned char *m,
                       unsigned long long mlen, const unsigned char *ad, unsigned long long adlen, unsigned char tag);
typedef int (*pull_fn)(void *state, unsigned char *m, unsigned long long *mlen, unsigned char *tag,
                       const unsigned char *c, unsigned long long clen, const unsigned char *ad,
                       unsigned long long adlen);

int ssref_push(uint8_t *, uint8_t *, const uint8_t *, uint64_t, const uint8_t *, uint64_t, uint8_t);
int ssref_pull(uint8_t *, uint8_t *, uint8_t *, const uint8_t *, uint64_t, const uint8_t *, uint64_t);

/* adapters giving the restatement libsodium's call shape */
int cpub_port_push(void *state, unsigned char *c, unsigned long long *clen, const unsigned char *m,
                   unsigned long long mlen, const unsigned char *ad, unsigned long long adlen, unsigned char tag) {
    if (clen) *clen = mlen + 17;
    return ssref_push((uint8_t *)state, c, m, mlen, ad, adlen, tag);
}
int cpub_port_pull(void *state, unsigned char *m, unsigned long long *mlen, unsigned char *tag, const unsigned char *c,
                   unsigned long long clen, const unsigned char *ad, unsigned long long adlen) {
    if (mlen && clen >= 17) *mlen = clen - 17;
    return ssref_pull((uint8_t *)state, m, tag, c, clen, ad, adlen);
}

typedef struct {
    push_fn push;
    pull_fn pull;
    uint8_t *tx_states, *rx_states; /* 52 B per stream */
    uint32_t n_streams;
    const uint8_t *plain;   /* step s, stream i at plain + (s % ring) * plain_step + plain_off[i] */
    const uint64_t *plain_off;
    const uint32_t *lens;   /* per stream, fixed across steps */
    uint64_t plain_step;
    uint8_t *wire;          /* step s, stream i at wire + (s % ring) * wire_step + wire_off[i]; len+20 bytes */
    const uint64_t *wire_off;
    uint64_t wire_step;
    uint32_t ring, steps, warmup;
    int do_seal, do_open;
    int threads;
    volatile int failures;
    int arrived;             /* workers that finished their warm-up */
    double *t_start, *t_end; /* per worker */
} cpub_job;

typedef struct { cpub_job *job; int tid; } cpub_arg;

static int run_steps(cpub_job *j, uint32_t lo, uint32_t hi, uint32_t first, uint32_t count) {
    int bad = 0;
    for (uint32_t s = first; s < first + count; s++) {
        uint64_t po = (uint64_t)(s % j->ring) * j->plain_step, wo = (uint64_t)(s % j->ring) * j->wire_step;
        for (uint32_t i = lo; i < hi; i++) {
            uint32_t len = j->lens[i];
            uint8_t *w = j->wire + wo + j->wire_off[i];
            if (j->do_seal) {
                uint32_t n = len + 17;
                memcpy(w + 4, j->plain + po + j->plain_off[i], len); /* T1: wrapped.set(data, 4) */
                w[0] = (uint8_t)n; w[1] = (uint8_t)(n >> 8); w[2] = (uint8_t)(n >> 16);
                bad |= j->push(j->tx_states + (size_t)i * 52, w + 3, NULL, w + 4, len, NULL, 0, 0);
            }
            if (j->do_open) {
                unsigned char tag;
                /* T1 `_incoming`: decrypt in place into message[1:len-16] */
                bad |= j->pull(j->rx_states + (size_t)i * 52, w + 4, NULL, &tag, w + 3, (unsigned long long)len + 17,
                               NULL, 0);
            }
        }
    }
    return bad;
}

/* The pool is up, has run the warm-up steps on its own streams (caches, branch predictors and the
 * chains are warm) and lines up at a spinning barrier before the timed steps: thread start-up and
 * futex wake-ups are set-up, not relay work, and must not be inside a region that may only last a
 * millisecond.  Every worker stamps its own start and end; the job's time is last end - first start. */
static double mono_now(void) {
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return (double)t.tv_sec + 1e-9 * (double)t.tv_nsec;
}

static void *worker(void *p) {
    cpub_arg *a = (cpub_arg *)p;
    cpub_job *j = a->job;
    uint32_t lo = (uint32_t)((uint64_t)j->n_streams * a->tid / j->threads);
    uint32_t hi = (uint32_t)((uint64_t)j->n_streams * (a->tid + 1) / j->threads);
    int bad = run_steps(j, lo, hi, 0, j->warmup);
    __atomic_add_fetch(&j->arrived, 1, __ATOMIC_SEQ_CST);
    while (__atomic_load_n(&j->arrived, __ATOMIC_ACQUIRE) < j->threads) __builtin_ia32_pause();
    j->t_start[a->tid] = mono_now();
    bad |= run_steps(j, lo, hi, j->warmup, j->steps);
    j->t_end[a->tid] = mono_now();
    if (bad) j->failures = 1;
    return NULL;
}

/* Returns elapsed seconds (wall, CLOCK_MONOTONIC) for `steps` passes after `warmup` untimed passes on
 * the same threads; <0 on failure. */
double cpub_run(void *push, void *pull, uint8_t *tx_states, uint8_t *rx_states, uint32_t n_streams,
                const uint8_t *plain, const uint64_t *plain_off, const uint32_t *lens, uint64_t plain_step,
                uint8_t *wire, const uint64_t *wire_off, uint64_t wire_step, uint32_t ring, uint32_t steps,
                uint32_t warmup, int do_seal, int do_open, int threads) {
    cpub_job job;
    memset(&job, 0, sizeof job);
    job.push = (push_fn)push; job.pull = (pull_fn)pull; job.tx_states = tx_states; job.rx_states = rx_states;
    job.n_streams = n_streams; job.plain = plain; job.plain_off = plain_off; job.lens = lens;
    job.plain_step = plain_step; job.wire = wire; job.wire_off = wire_off; job.wire_step = wire_step;
    job.ring = ring ? ring : 1; job.steps = steps; job.warmup = warmup; job.do_seal = do_seal; job.do_open = do_open;
    job.threads = threads < 1 ? 1 : threads;
    if ((uint32_t)job.threads > n_streams) job.threads = (int)n_streams;
    if (job.threads < 1) job.threads = 1;
    pthread_t *th = (pthread_t *)calloc((size_t)job.threads, sizeof *th);
    cpub_arg *args = (cpub_arg *)calloc((size_t)job.threads, sizeof *args);
    job.t_start = (double *)calloc((size_t)job.threads, sizeof(double));
    job.t_end = (double *)calloc((size_t)job.threads, sizeof(double));
    for (int t = 0; t < job.threads; t++) {
        args[t].job = &job; args[t].tid = t;
        pthread_create(&th[t], NULL, worker, &args[t]);
    }
    for (int t = 0; t < job.threads; t++) pthread_join(th[t], NULL);   /* the caller sleeps: every core belongs to a worker */
    double first = job.t_start[0], last = job.t_end[0];
    for (int t = 1; t < job.threads; t++) {
        if (job.t_start[t] < first) first = job.t_start[t];
        if (job.t_end[t] > last) last = job.t_end[t];
    }
    free(th); free(args); free(job.t_start); free(job.t_end);
    if (job.failures) return -1.0;
    return last - first;
}

/* BASELINE config 1 on one core: ONE connection, its frames pushed (secret-stream `_write`) and pulled
 * (`_incoming`) strictly in sequence, `reps` passes over the n chunks in `blob` (offsets off[0..n]).
 * scratch must hold the longest chunk + 20 bytes.  Returns seconds, <0 on a failed pull. */
double cpub_chain_run(void *push_p, void *pull_p, uint8_t *tx_state, uint8_t *rx_state, const uint8_t *blob,
                      const uint32_t *off, uint32_t n, uint32_t reps, uint8_t *scratch) {
    push_fn push = (push_fn)push_p;
    pull_fn pull = (pull_fn)pull_p;
    int bad = 0;
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (uint32_t r = 0; r < reps; r++)
        for (uint32_t i = 0; i < n; i++) {
            const uint32_t len = off[i + 1] - off[i], m = len + 17;
            unsigned char tag;
            memcpy(scratch + 4, blob + off[i], len);
            scratch[0] = (uint8_t)m; scratch[1] = (uint8_t)(m >> 8); scratch[2] = (uint8_t)(m >> 16);
            bad |= push(tx_state, scratch + 3, NULL, scratch + 4, len, NULL, 0, 0);
            bad |= pull(rx_state, scratch + 4, NULL, &tag, scratch + 3, (unsigned long long)len + 17, NULL, 0);
        }
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if (bad) return -1.0;
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
