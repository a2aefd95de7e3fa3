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

double e2e_ring_run(symrelay_ctx *ctx, const symrelay_desc *desc, uint32_t n, uint8_t *ring_base, uint64_t slot_bytes,
                    uint64_t in_bytes, uint64_t out_off, uint64_t out_bytes, uint32_t ring, uint32_t depth,
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
        *slot = symrelay_batch_submit(ctx, desc, n, ring_base + (i % ring) * slot_bytes, in_bytes,
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
