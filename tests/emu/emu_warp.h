// emu_warp.h -- TEST INFRASTRUCTURE ONLY.
// A 32-thread, barrier-synchronised stand-in for a CUDA warp so that
// symmetry_b200/csrc/ss_core.cuh (the product's device code) can be compiled with
// g++ and its lane logic unit-tested on a machine without a GPU.  Slow by design;
// never linked into, loaded by, or used as a fallback for the product library.
#pragma once
#include <pthread.h>
#include <stdint.h>

#define SS_DEV static inline

struct alignas(16) uint4 {
    uint32_t x, y, z, w;
};
struct alignas(8) uint2 {
    uint32_t x, y;
};
static inline uint2 make_uint2(uint32_t x, uint32_t y) { uint2 v; v.x = x; v.y = y; return v; }

struct EmuWarp {
    pthread_barrier_t bar;
    volatile uint32_t slot[32];
};
extern thread_local uint32_t emu_lane_id;
extern thread_local EmuWarp *emu_warp;

static inline uint32_t ss_lane() { return emu_lane_id; }
static inline void emu_sync() { pthread_barrier_wait(&emu_warp->bar); }
static inline uint32_t ss_shfl(uint32_t v, uint32_t src) {
    emu_warp->slot[emu_lane_id] = v;
    emu_sync();
    const uint32_t r = emu_warp->slot[src & 31u];
    emu_sync();
    return r;
}
static inline uint32_t ss_shfl_up(uint32_t v, uint32_t d) {
    emu_warp->slot[emu_lane_id] = v;
    emu_sync();
    const uint32_t r = emu_lane_id >= d ? emu_warp->slot[emu_lane_id - d] : v;
    emu_sync();
    return r;
}
static inline uint32_t ss_shfl_down(uint32_t v, uint32_t d) {
    emu_warp->slot[emu_lane_id] = v;
    emu_sync();
    const uint32_t r = emu_lane_id + d < 32u ? emu_warp->slot[emu_lane_id + d] : v;
    emu_sync();
    return r;
}
static inline uint32_t ss_reduce_or(uint32_t v) {
    emu_warp->slot[emu_lane_id] = v;
    emu_sync();
    uint32_t r = 0;
    for (int i = 0; i < 32; i++) r |= emu_warp->slot[i];
    emu_sync();
    return r;
}
static inline uint32_t ss_reduce_max(uint32_t v) {
    emu_warp->slot[emu_lane_id] = v;
    emu_sync();
    uint32_t r = 0;
    for (int i = 0; i < 32; i++) r = emu_warp->slot[i] > r ? emu_warp->slot[i] : r;
    emu_sync();
    return r;
}
static inline uint32_t ss_reduce_min(uint32_t v) {
    emu_warp->slot[emu_lane_id] = v;
    emu_sync();
    uint32_t r = 0xffffffffu;
    for (int i = 0; i < 32; i++) r = emu_warp->slot[i] < r ? emu_warp->slot[i] : r;
    emu_sync();
    return r;
}
static inline bool ss_any(bool p) { return ss_reduce_or(p ? 1u : 0u) != 0; }
static inline void ss_syncwarp() { emu_sync(); }
#define SS_STAGE_BYTES 2048u
static uint4 emu_stage[SS_STAGE_BYTES / 16u];   // one warp runs at a time
static inline uint4 *ss_stage_buffer() { return emu_stage; }
static uint32_t emu_powers[16 * 48];   // SS_MAX_SEG_FRAMES frames per warp
static inline uint32_t *ss_powers_buffer() { return emu_powers; }
// wide tiles (one frame per four-warp block) are never instantiated under the emulator: stubs so that the template parses
static inline uint32_t ss_wide_tid() { return emu_lane_id; }
static inline void ss_wide_sync() { emu_sync(); }
static uint32_t emu_wide[32];
static inline uint32_t *ss_wide_buffer() { return emu_wide; }
static inline uint32_t ss_popc(uint32_t v) { return (uint32_t)__builtin_popcount(v); }
static inline uint32_t ss_clz(uint32_t v) { return v ? (uint32_t)__builtin_clz(v) : 32u; }
static inline uint32_t emu_rotl(uint32_t x, int n) { return (x << n) | (x >> (32 - n)); }
static inline uint32_t ss_rotl16(uint32_t x) { return emu_rotl(x, 16); }
static inline uint32_t ss_rotl8(uint32_t x) { return emu_rotl(x, 8); }
static inline uint32_t ss_rotl12(uint32_t x) { return emu_rotl(x, 12); }
static inline uint32_t ss_rotl7(uint32_t x) { return emu_rotl(x, 7); }
static inline uint32_t ss_fsr(uint32_t lo, uint32_t hi, uint32_t s) {
    return (uint32_t)((((uint64_t)hi << 32) | lo) >> (s & 31u));
}
