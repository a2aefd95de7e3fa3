/*
 * tools/e2e_driver.c -- a C caller of the PUBLIC C ABI (include/symrelay.h), used by bench.py's
 * end-to-end leg and tools/e2e_probe.py so that the timed loop is the library + driver, not the
 * Python interpreter.  It makes exactly the calls an N-API addon would make per batch:
 * symrelay_batch_wait(ticket[i-depth]); ticket[i] = symrelay_batch_submit(...).  Step i reads
 * ring slot i % ring at in_base and writes to ring slot (i + depth) % ring at out_base (bench.py's
 * [plain | wire | recovered] slot layout).  Returns elapsed seconds for `steps` batches with the
 * pipeline drained on both sides, or a negative SYMRELAY_E* code.
 */
#define _GNU_SOURCE
#include <stdint.h>
#include <stdlib.h>
#include <time.h>

#include "../include/symrelay.h"

static double now(void) {
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return (double)t.tv_sec + 1e-9 * (double)t.tv_nsec;
}

/* desc_alt (may be NULL): a second descriptor array for the same frames (e.g. in another order), used on odd steps, so
 * that no two consecutive batches are byte-identical: the library's general path (scan, plan, per-stream stamps) every
 * time instead of its repeating-batch fast path. */
double e2e_ring_run2(symrelay_ctx *ctx, const symrelay_desc *desc, const symrelay_desc *desc_alt, uint32_t n, uint8_t *ring_base,
                     uint64_t slot_bytes, uint64_t in_bytes, uint64_t out_off, uint64_t out_bytes, uint32_t ring, uint32_t depth,
                     uint64_t first_step, uint32_t steps, uint32_t *status /* ring x n */, double *submit_seconds);

double e2e_ring_run(symrelay_ctx *ctx, const symrelay_desc *desc, uint32_t n, uint8_t *ring_base, uint64_t slot_bytes,
                    uint64_t in_bytes, uint64_t out_off, uint64_t out_bytes, uint32_t ring, uint32_t depth,
                    uint64_t first_step, uint32_t steps, uint32_t *status /* ring x n */, double *submit_seconds) {
    return e2e_ring_run2(ctx, desc, NULL, n, ring_base, slot_bytes, in_bytes, out_off, out_bytes, ring, depth, first_step, steps,
                         status, submit_seconds);
}

double e2e_ring_run2(symrelay_ctx *ctx, const symrelay_desc *desc, const symrelay_desc *desc_alt, uint32_t n, uint8_t *ring_base,
                     uint64_t slot_bytes, uint64_t in_bytes, uint64_t out_off, uint64_t out_bytes, uint32_t ring, uint32_t depth,
                     uint64_t first_step, uint32_t steps, uint32_t *status /* ring x n */, double *submit_seconds) {
    int64_t *tickets = (int64_t *)malloc(sizeof(int64_t) * depth);
    for (uint32_t d = 0; d < depth; d++) tickets[d] = -1;
    double sub = 0.0;
    const double t0 = now();
    for (uint64_t i = first_step; i < first_step + steps; i++) {
        int64_t *slot = &tickets[i % depth];
        if (*slot >= 0) {
            int rc = symrelay_batch_wait(ctx, *slot);
            if (rc) { free(tickets); return (double)rc; }
        }
        const double a = now();
        *slot = symrelay_batch_submit(ctx, (desc_alt && (i & 1)) ? desc_alt : desc, n, ring_base + (i % ring) * slot_bytes, in_bytes,
                                      ring_base + ((i + depth) % ring) * slot_bytes + out_off, out_bytes,
                                      status + (size_t)(i % ring) * n, NULL, NULL, 0);
        sub += now() - a;
        if (*slot < 0) { double e = (double)*slot; free(tickets); return e; }
    }
    for (uint32_t d = 0; d < depth; d++) {
        const uint64_t i = first_step + steps + d;   /* drain in submission order */
        int64_t t = tickets[i % depth];
        if (t >= 0) {
            int rc = symrelay_batch_wait(ctx, t);
            if (rc) { free(tickets); return (double)rc; }
        }
    }
    const double t1 = now();
    if (submit_seconds) *submit_seconds = sub;
    free(tickets);
    return t1 - t0;
}

/* Per-batch latency of the synchronous call: `steps` consecutive symrelay_batch_host calls on the same ring and
 * the same chains, one batch in flight; lat_ns[i] = wall time of call i (host arenas in -> results visible to the
 * host).  SURVEY.md 8(d), C2: "batch latency (p50/p99)".  Same argument meaning as e2e_ring_run. */
static uint64_t now_ns(void) {
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return (uint64_t)t.tv_sec * 1000000000ull + (uint64_t)t.tv_nsec;
}

int e2e_sync_latency(symrelay_ctx *ctx, const symrelay_desc *desc, uint32_t n, uint8_t *ring_base, uint64_t slot_bytes,
                     uint64_t in_bytes, uint64_t out_off, uint64_t out_bytes, uint32_t ring, uint32_t depth,
                     uint64_t first_step, uint32_t steps, uint32_t *status /* ring x n */, uint64_t *lat_ns) {
    for (uint64_t i = first_step; i < first_step + steps; i++) {
        const uint64_t a = now_ns();
        const int rc = symrelay_batch_host(ctx, desc, n, ring_base + (i % ring) * slot_bytes, in_bytes,
                                           ring_base + ((i + depth) % ring) * slot_bytes + out_off, out_bytes,
                                           status + (size_t)(i % ring) * n, NULL, NULL, 0);
        if (rc) return rc;
        lat_ns[i - first_step] = now_ns() - a;
    }
    return 0;
}

/* BASELINE config 1 restated for the library: ONE connection, its frames strictly in
 * sequence through the per-frame ABI exactly as the relay loop issues them (src/provider.ts:234-262):
 * symrelay_submit_seal(chunk[, F_EXTRACT]) -> flush -> wait for its completion -> release, then the next chunk.
 * chunks are concatenated in `blob` with offsets off[0..n]; extract[i] != 0 marks the body chunks.
 * Returns seconds for `reps` passes over the n chunks, or a negative SYMRELAY_E* code; *n_deltas counts delta bytes. */
double c1_run(symrelay_ctx *ctx, uint32_t slot, const uint8_t *blob, const uint32_t *off, const uint8_t *extract,
              uint32_t n, uint32_t reps, uint64_t *delta_bytes) {
    symrelay_completion c[4];
    uint64_t db = 0;
    const double t0 = now();
    for (uint32_t r = 0; r < reps; r++)
        for (uint32_t i = 0; i < n; i++) {
            const int64_t t = symrelay_submit_seal(ctx, slot, blob + off[i], off[i + 1] - off[i],
                                                   extract[i] ? SYMRELAY_F_EXTRACT : 0u);
            if (t < 0) return (double)t;
            int rc = symrelay_flush(ctx);
            if (rc) return (double)rc;
            int k = 0;
            while (k == 0) {
                k = symrelay_poll_wait(ctx, c, 4);
                if (k < 0) return (double)k;
            }
            if (k != 1 || c[0].ticket != t) return (double)SYMRELAY_EINVAL;
            const uint32_t code = SYMRELAY_STATUS_CODE(c[0].status);
            if (code != SYMRELAY_ST_OK) return -(double)(2000 + code);
            db += c[0].delta_len;
            symrelay_release(ctx, t);
        }
    const double t1 = now();
    if (delta_bytes) *delta_bytes = db;
    return t1 - t0;
}

/* The reference is ONE Node process (src/provider.ts:37-58: one SymmetryProvider, one event loop): this is
 * e2e_ring_run for a context that spans several GPUs, driven by one thread.  Per step it visits every GPU once:
 * wait for that GPU's batch of `depth` steps ago, submit its next one.  Arrays are indexed by GPU; every GPU has its
 * own descriptor array (its streams), ring and status block.  Returns seconds, or a negative SYMRELAY_E* code. */
double e2e_ring_run_multi(symrelay_ctx *ctx, int n_gpus, const symrelay_desc **desc, const uint32_t *n, uint8_t **ring_base,
                          const uint64_t *slot_bytes, const uint64_t *in_bytes, const uint64_t *out_off, const uint64_t *out_bytes,
                          uint32_t ring, uint32_t depth, uint64_t first_step, uint32_t steps, uint32_t **status) {
    int64_t *tickets = (int64_t *)malloc(sizeof(int64_t) * depth * (size_t)n_gpus);
    for (size_t d = 0; d < (size_t)depth * (size_t)n_gpus; d++) tickets[d] = -1;
    const double t0 = now();
    for (uint64_t i = first_step; i < first_step + steps; i++)
        for (int g = 0; g < n_gpus; g++) {
            int64_t *slot = &tickets[(size_t)g * depth + i % depth];
            if (*slot >= 0) {
                int rc = symrelay_batch_wait(ctx, *slot);
                if (rc) { free(tickets); return (double)rc; }
            }
            *slot = symrelay_batch_submit(ctx, desc[g], n[g], ring_base[g] + (i % ring) * slot_bytes[g], in_bytes[g],
                                          ring_base[g] + ((i + depth) % ring) * slot_bytes[g] + out_off[g], out_bytes[g],
                                          status[g] + (size_t)(i % ring) * n[g], NULL, NULL, 0);
            if (*slot < 0) { double e = (double)*slot; free(tickets); return e; }
        }
    for (uint32_t d = 0; d < depth; d++)
        for (int g = 0; g < n_gpus; g++) {
            const uint64_t i = first_step + steps + d;   /* drain in submission order */
            int64_t t = tickets[(size_t)g * depth + i % depth];
            if (t >= 0) {
                int rc = symrelay_batch_wait(ctx, t);
                if (rc) { free(tickets); return (double)rc; }
            }
        }
    const double t1 = now();
    free(tickets);
    return t1 - t0;
}
