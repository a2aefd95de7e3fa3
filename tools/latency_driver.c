/*
 * tools/latency_driver.c -- added relay latency under Poisson load, driven from C through the PUBLIC
 * per-frame ABI (symrelay_submit_seal / symrelay_poll / symrelay_release), single producer thread as
 * in Node.  Frames of `frame_bytes` arrive as a Poisson process of aggregate rate `rate` frames/s on
 * streams picked uniformly from `slots`; latency = delivery time - SCHEDULED arrival time (so a
 * driver that falls behind is charged for it).  SURVEY.md section 8(d) "Latency".
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../include/symrelay.h"

static uint64_t now_ns(void) {
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return (uint64_t)t.tv_sec * 1000000000ull + (uint64_t)t.tv_nsec;
}
static uint64_t rng_state = 0x9e3779b97f4a7c15ull;
static uint64_t rng_next(void) {
    uint64_t x = rng_state;
    x ^= x << 13; x ^= x >> 7; x ^= x << 17;
    return rng_state = x;
}
static int cmp_u32(const void *a, const void *b) {
    const uint32_t x = *(const uint32_t *)a, y = *(const uint32_t *)b;
    return x < y ? -1 : x > y;
}

typedef struct {
    double p50_us, p99_us, p999_us, max_us, achieved_rate, eagain_per_frame;
    uint64_t frames;
} latency_result;

int latency_run(symrelay_ctx *ctx, const uint32_t *slots, uint32_t n_streams, uint32_t frame_bytes, double rate,
                double seconds, latency_result *res) {
    const uint64_t total = (uint64_t)(rate * seconds);
    uint32_t *lat = (uint32_t *)malloc(sizeof(uint32_t) * total);          /* latency in ns, saturating */
    uint64_t *arrival = (uint64_t *)malloc(sizeof(uint64_t) * total);      /* scheduled arrival, by frame id */
    uint8_t *payload = (uint8_t *)malloc(frame_bytes + 16);
    enum { MAXC = 4096 };
    symrelay_completion *comp = (symrelay_completion *)malloc(sizeof(symrelay_completion) * MAXC);
    /* ticket -> frame id: tickets are (batch seq << 20 | index); remember ids in a small open-addressing map */
    const uint64_t map_sz = 1ull << 22;
    int64_t *map_key = (int64_t *)malloc(sizeof(int64_t) * map_sz);
    uint32_t *map_val = (uint32_t *)malloc(sizeof(uint32_t) * map_sz);
    if (!lat || !arrival || !payload || !comp || !map_key || !map_val) return SYMRELAY_ENOMEM;
    for (uint64_t i = 0; i < map_sz; i++) map_key[i] = -1;
    memset(payload, 0x5a, frame_bytes + 16);
    const double mean_gap_ns = 1e9 / rate;
    uint64_t submitted = 0, done = 0, eagain = 0;
    const uint64_t t_start = now_ns();
    double next_arrival = (double)t_start;
    int rc = 0;
    while (done < total) {
        uint64_t t = now_ns();
        while (submitted < total && next_arrival <= (double)t) {
            const uint32_t slot = slots[rng_next() % n_streams];
            const int64_t ticket = symrelay_submit_seal(ctx, slot, payload, frame_bytes, 0);
            if (ticket == SYMRELAY_EAGAIN) { eagain++; break; }      /* `write()` false: poll, then retry */
            if (ticket < 0) { rc = (int)ticket; goto out; }
            uint64_t h = ((uint64_t)ticket * 0x9e3779b97f4a7c15ull) >> 42;
            while (map_key[h] != -1) h = (h + 1) & (map_sz - 1);
            map_key[h] = ticket; map_val[h] = (uint32_t)submitted;
            arrival[submitted++] = (uint64_t)next_arrival;
            const double u = ((double)(rng_next() >> 11) + 1.0) / 9007199254740993.0;
            next_arrival += -log(u) * mean_gap_ns;
            if ((submitted & 63) == 0) t = now_ns();
        }
        const int k = symrelay_poll(ctx, comp, MAXC);
        if (k < 0) { rc = k; goto out; }
        if (k) {
            const uint64_t td = now_ns();
            for (int i = 0; i < k; i++) {
                uint64_t h = ((uint64_t)comp[i].ticket * 0x9e3779b97f4a7c15ull) >> 42;
                while (map_key[h] != comp[i].ticket) h = (h + 1) & (map_sz - 1);
                const uint32_t id = map_val[h];
                map_key[h] = -2;   /* tombstone */
                const uint64_t d = td - arrival[id];
                lat[done++] = d > 0xffffffffull ? 0xffffffffu : (uint32_t)d;
                if (SYMRELAY_STATUS_CODE(comp[i].status) != SYMRELAY_ST_OK) { rc = SYMRELAY_EINVAL; goto out; }
                symrelay_release(ctx, comp[i].ticket);
            }
        }
        if (submitted == total) symrelay_flush(ctx);
    }
    {
        const uint64_t t_end = now_ns();
        qsort(lat, total, sizeof(uint32_t), cmp_u32);
        res->frames = total;
        res->p50_us = lat[total / 2] / 1e3;
        res->p99_us = lat[(uint64_t)(total * 0.99)] / 1e3;
        res->p999_us = lat[(uint64_t)(total * 0.999)] / 1e3;
        res->max_us = lat[total - 1] / 1e3;
        res->achieved_rate = (double)total / ((double)(t_end - t_start) * 1e-9);
        res->eagain_per_frame = (double)eagain / (double)total;
    }
out:
    free(lat); free(arrival); free(payload); free(comp); free(map_key); free(map_val);
    return rc;
}
