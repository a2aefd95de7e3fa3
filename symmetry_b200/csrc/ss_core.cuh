// ss_core.cuh -- warp-level secretstream (XChaCha20-Poly1305) seal/open for sm_100a.
//
// What it computes: libsodium's crypto_secretstream_xchacha20poly1305_push/_pull
// plus @hyperswarm/secret-stream's uint24-LE framing -- the arithmetic behind
// `peer.write(chunk)` (/root/reference/src/provider.ts:248) and `peer.on('data')`
// (:174); spec in SURVEY.md section 8(a) rows A9/A10/A12.
//
// Mapping ("warp-per-stream, thread-per-block keystream"):
//   * one lane = one 64-byte ChaCha20 block: lane b==0 makes the Poly1305 key
//     (ic=0), b==1 the tag block (ic=1), b>=2 the payload keystream (ic=b);
//   * a PACKED tile puts several small frames side by side in one warp
//     (sum of block counts <= 32); a BULK tile is one frame whose lanes loop
//     with stride 32 over its blocks;
//   * Poly1305 runs lane-parallel: every lane Horner-reduces its own 64-byte
//     window of the MAC input with r, the lane partials are aligned with
//     r^E (per-lane square-and-multiply; bulk lanes first fold their strided
//     windows with r^128) and summed with a segmented warp-shuffle tree;
//   * the frame's leader lane finalises the tag, writes/compares it and
//     advances the per-stream nonce chain in the HBM state table.
//
// The same source is compiled (a) by nvcc for sm_100a -- the product -- and
// (b) by g++ under tests/emu/ with a 32-thread warp emulator, which exists only
// so the lane logic can be unit-tested on a machine without a GPU.  (b) is test
// infrastructure; nothing in the product links or loads it.
#pragma once
#include <stdint.h>

#ifdef SS_HOST_EMU
#include "emu_warp.h"  // SS_DEV, uint4, ss_lane(), collectives and bit intrinsics for the emulator
#else
#define SS_DEV __device__ __forceinline__
#define SS_NOINLINE __device__ __noinline__
SS_DEV uint32_t ss_lane() { return threadIdx.x & 31u; }
SS_DEV uint32_t ss_shfl(uint32_t v, uint32_t src) { return __shfl_sync(0xffffffffu, v, (int)src); }
SS_DEV uint32_t ss_shfl_up(uint32_t v, uint32_t d) { return __shfl_up_sync(0xffffffffu, v, d); }
SS_DEV uint32_t ss_shfl_down(uint32_t v, uint32_t d) { return __shfl_down_sync(0xffffffffu, v, d); }
SS_DEV uint32_t ss_reduce_or(uint32_t v) { return __reduce_or_sync(0xffffffffu, v); }
SS_DEV uint32_t ss_reduce_max(uint32_t v) { return __reduce_max_sync(0xffffffffu, v); }
SS_DEV uint32_t ss_reduce_min(uint32_t v) { return __reduce_min_sync(0xffffffffu, v); }
SS_DEV bool ss_any(bool p) { return __any_sync(0xffffffffu, p) != 0; }
SS_DEV void ss_syncwarp() { __syncwarp(); }
// strided tiles: r, r^2, r^3, r^4, r^(4*segw) of each of the warp's frames (up to 16 with 2 lanes per frame), each
// as 5 limbs followed by 4 limbs-times-5 (45 of 48 words)
#define SS_MAX_SEG_FRAMES 16u
SS_DEV uint32_t *ss_powers_buffer() {
    __shared__ uint32_t pw[4][SS_MAX_SEG_FRAMES * 48u];
    return pw[(threadIdx.x >> 5) & 3u];
}
#define SS_STAGE_BYTES 2048u
// per-warp staging window in shared memory (packed tiles, see ss_process_tile)
SS_DEV uint4 *ss_stage_buffer() {
    __shared__ __align__(16) uint4 stage[4][SS_STAGE_BYTES / 16u];
    return stage[(threadIdx.x >> 5) & 3u];
}
// wide tiles: one frame per 128-thread block (four warps walk its blocks with stride 128)
SS_DEV uint32_t ss_wide_tid() { return threadIdx.x & 127u; }
SS_DEV void ss_wide_sync() { __syncthreads(); }
SS_DEV uint32_t *ss_wide_buffer() {
    __shared__ uint32_t wb[32];   // [0..3] Poly1305 r words, [4] verdict, [8..27] the four warps' partial sums
    return wb;
}
SS_DEV uint32_t ss_popc(uint32_t v) { return (uint32_t)__popc(v); }
SS_DEV uint32_t ss_clz(uint32_t v) { return (uint32_t)__clz((int)v); }
SS_DEV uint32_t ss_rotl16(uint32_t x) { return __byte_perm(x, 0, 0x1032); }
SS_DEV uint32_t ss_rotl8(uint32_t x) { return __byte_perm(x, 0, 0x2103); }
SS_DEV uint32_t ss_rotl12(uint32_t x) { return __funnelshift_l(x, x, 12); }
SS_DEV uint32_t ss_rotl7(uint32_t x) { return __funnelshift_l(x, x, 7); }
// (lo, hi) >> s, low 32 bits, 0 <= s < 32
SS_DEV uint32_t ss_fsr(uint32_t lo, uint32_t hi, uint32_t s) { return __funnelshift_r(lo, hi, s); }
#endif

// Programmatic dependent launch (sm_90+): a kernel launched with the PDL attribute may start while
// its predecessor in the stream still runs; griddepcontrol.wait blocks until that predecessor has
// completed and its memory is visible.  Both are no-ops for ordinary launches.
#ifdef SS_HOST_EMU
static inline void ss_pdl_launch_dependents() {}
static inline void ss_pdl_wait() {}
#else
SS_DEV void ss_pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
SS_DEV void ss_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#endif

// Loads for the persistent kernel (VOL = true).  A resident grid re-reads the same host addresses batch after
// batch (the caller's arena ring, the control blocks) and the same stream states that other SMs have advanced in
// between; plain loads may be served from a stale L1/L2 line, which never happens to a kernel that is launched
// per batch.  ld.volatile fetches from the point of coherence every time.  Generic addressing: the payload
// pointer may also be the staged copy in shared memory.
#ifdef SS_HOST_EMU
template <bool VOL> static inline uint4 ss_ld16(const void *p) { return *reinterpret_cast<const uint4 *>(p); }
template <bool VOL> static inline uint32_t ss_ld4(const void *p) { return *reinterpret_cast<const uint32_t *>(p); }
template <bool VOL> static inline uint32_t ss_ld1(const void *p) { return *reinterpret_cast<const uint8_t *>(p); }
static inline void ss_dep_wait(const uint32_t *, uint32_t) {}
static inline uint32_t ss_ld_acquire_u32(const uint32_t *p) { return *p; }
static inline void ss_st_release_u32(uint32_t *p, uint32_t v) { *p = v; }
static inline void ss_nanosleep(uint32_t) {}
#else
SS_DEV uint32_t ss_ld_acquire_u32(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
SS_DEV void ss_st_release_u32(uint32_t *p, uint32_t v) { asm volatile("st.release.gpu.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
SS_DEV void ss_nanosleep(uint32_t ns) { __nanosleep(ns); }
template <bool VOL> SS_DEV uint4 ss_ld16(const void *p) {
    if (!VOL) return *reinterpret_cast<const uint4 *>(p);
    uint4 v;
    asm volatile("ld.volatile.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
template <bool VOL> SS_DEV uint32_t ss_ld4(const void *p) {
    if (!VOL) return *reinterpret_cast<const uint32_t *>(p);
    uint32_t v;
    asm volatile("ld.volatile.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
template <bool VOL> SS_DEV uint32_t ss_ld1(const void *p) {
    if (!VOL) return *reinterpret_cast<const uint8_t *>(p);
    uint32_t v;
    asm volatile("ld.volatile.u8 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// Persistent kernel: batch b may touch the stream states only after batch b-1 has left them (the MAC chain).
// *seq counts completed batches (device memory, released by the last warp of a batch).
SS_DEV void ss_dep_wait(const uint32_t *seq, uint32_t need) {
    if (!seq) return;
    uint32_t v;
    for (;;) {
        asm volatile("ld.acquire.gpu.u32 %0, [%1];" : "=r"(v) : "l"(seq) : "memory");
        if ((int32_t)(v - need) >= 0) break;
        __nanosleep(40);
    }
}
#endif

#define SS_ARG_INPUT_INDEPENDENT 1u  // SsArgs.flags: desc/tiles/in were not written by the preceding kernel
#define SS_ARG_STAGE_INPUT 2u        // SsArgs.flags: pull each packed tile's input through shared memory
#define SS_ARG_STAGE_OUTPUT 4u       // SsArgs.flags: assemble each packed tile's output in shared memory, store whole lines
#define SS_ARG_CHAIN 8u              // SsArgs.flags: resident kernel -- hand the stream states over through their stamps
#define SS_ARG_CHAIN_UNIFORM 16u     // ... and every frame expects the stamp ver_all (no per-descriptor array)
#define SS_ARG_VERDICT_LATE 32u      // SsArgs.flags: the kernel was launched as a programmatic dependent of the batch's extractor:
                                     // thread tiles seal SS_F_EXTRACT frames while it still runs and ask for its verdict last

#define SS_F_OPEN 0x04000000u
#define SS_F_EXTRACT 0x08000000u
#define SS_F_RUN 0x10000000u   // open frame that is part of a run of queued frames of one stream (see ss_runs.cuh)
#define SS_ST_OK 0u
#define SS_ST_BAD_MAC 1u
#define SS_ST_RETRY 5u
#define SS_RUN_CAP 64u   // frames of one stream per batch that can be opened as a run (== SS_RUN_MAX in ss_runs.cuh)

struct SsDesc {
    uint32_t slot, in_off, out_off, len_flags;
};
struct SsTile {
    uint32_t first, n_kind;
};
struct SsArgs {
    const SsDesc *desc;
    const SsTile *tiles;
    uint32_t n_tiles;
    const uint8_t *in;
    uint8_t *out;
    uint4 *states;  // [slot][dir]: 3 x uint4 = k[0..3] | k[4..7] | {frame counter, inonce0, inonce1, meta}
    uint32_t *status;
    uint32_t flags;
    const uint4 *run_nonce;   // scratch: per-descriptor nonce for SS_F_RUN frames (written by ss_run_nonce_kernel)
    // Write-back of the PREVIOUS batch (host pipeline only, see submit_locked): that batch left its wire
    // bytes and statuses in device staging; this kernel posts them to the caller's host memory once the
    // previous kernel is complete, and the stores land while this batch computes.
    const uint4 *wb_src = nullptr;
    uint4 *wb_dst = nullptr;
    uint32_t wb_n16 = 0;
    const uint32_t *wb_status_src = nullptr;
    uint32_t *wb_status_dst = nullptr;
    uint32_t wb_n_status = 0;
    unsigned long long *dbg_max = nullptr;  // SYMRELAY_TS=1, resident kernel: per batch, max over tiles of the four stamps
    unsigned long long *dbg_ts = nullptr;   // SYMRELAY_TS=1 (diagnostics): {kernel start, past the dependency wait, end} of warp 0
    // Persistent kernel only: wait until *dep_seq >= dep_need before touching the stream states.
    const uint32_t *dep_seq = nullptr;
    uint32_t dep_need = 0;
    // Resident kernel only: per-stream hand-over of the MAC chain between batches.  The upper 24 bits of a state's
    // meta word carry the stamp of the last resident-kernel batch that went through that (slot, direction);
    // ver[i] is the stamp descriptor i has to find there before it may read the state (0: no predecessor in
    // flight), new_ver the stamp this batch leaves.  Release/acquire on that one word orders the nonce and key
    // words written before it -- a frame waits for exactly its predecessor, not for a whole batch.
    const uint32_t *ver = nullptr;
    uint32_t new_ver = 0;
    SsDesc *desc_fill = nullptr;   // resident kernel: also leave a copy of the tile's descriptors here (device memory), for the
                                   // next batch that repeats them -- it then reads them from HBM instead of over PCIe
    uint32_t ver_all = 0;     // SS_ARG_CHAIN_UNIFORM: every descriptor expects this stamp (the batch repeats its predecessor's streams)
    const volatile uint32_t *verdicts = nullptr;   // fused extract+seal kernel: the 32 verdict words (shared memory) of this tile's chunks
};

SS_DEV void ss_dbg_stamp(const SsArgs &a, uint32_t tile_idx, uint32_t lane, uint32_t which) {
#ifndef SS_HOST_EMU
    if (a.dbg_ts && tile_idx == 0 && lane == 0 && which < 4u) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        a.dbg_ts[which] = t;
    }
    if (a.dbg_max && lane == 0) {   // resident kernel: the latest tile of the batch to reach each point
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        atomicMax(a.dbg_max + which, t);
    }
#endif
}

// Every warp of the grid takes a stripe.  Call after ss_pdl_wait().
SS_DEV void ss_write_back(const SsArgs &a, uint32_t warp, uint32_t n_warps, uint32_t lane) {
    if (!a.wb_n16) return;
    for (uint32_t i = warp * 32u + lane; i < a.wb_n16; i += n_warps * 32u) a.wb_dst[i] = a.wb_src[i];
    for (uint32_t i = warp * 32u + lane; i < a.wb_n_status; i += n_warps * 32u) a.wb_status_dst[i] = a.wb_status_src[i];
}

// ---------------------------------------------------------------- ChaCha20 --
// Double rounds unrolled per loop trip.  1 keeps the kernels' code inside the
// instruction caches (the fully unrolled kernels were 64 KB and stalled on
// instruction fetch with one warp per SM sub-partition); see DESIGN.md.
#ifndef SS_CHACHA_UNROLL
#define SS_CHACHA_UNROLL 1
#endif
#define SS_PRAGMA_(x) _Pragma(#x)
#define SS_UNROLL(n) SS_PRAGMA_(unroll n)
#define SS_QR(a, b, c, d)   \
    a += b; d ^= a; d = ss_rotl16(d); \
    c += d; b ^= c; b = ss_rotl12(b); \
    a += b; d ^= a; d = ss_rotl8(d);  \
    c += d; b ^= c; b = ss_rotl7(b);

// RFC 8439 block function; x <- keystream words for block counter `ic`.
SS_DEV void ss_chacha_block(const uint32_t k[8], uint32_t ic, uint32_t n0, uint32_t n1, uint32_t n2,
                            uint32_t x[16]) {
    x[0] = 0x61707865u; x[1] = 0x3320646eu; x[2] = 0x79622d32u; x[3] = 0x6b206574u;
#pragma unroll
    for (int i = 0; i < 8; i++) x[4 + i] = k[i];
    x[12] = ic; x[13] = n0; x[14] = n1; x[15] = n2;
    SS_UNROLL(SS_CHACHA_UNROLL)
    for (int r = 0; r < 10; r++) {
        SS_QR(x[0], x[4], x[8], x[12])
        SS_QR(x[1], x[5], x[9], x[13])
        SS_QR(x[2], x[6], x[10], x[14])
        SS_QR(x[3], x[7], x[11], x[15])
        SS_QR(x[0], x[5], x[10], x[15])
        SS_QR(x[1], x[6], x[11], x[12])
        SS_QR(x[2], x[7], x[8], x[13])
        SS_QR(x[3], x[4], x[9], x[14])
    }
    x[0] += 0x61707865u; x[1] += 0x3320646eu; x[2] += 0x79622d32u; x[3] += 0x6b206574u;
#pragma unroll
    for (int i = 0; i < 8; i++) x[4 + i] += k[i];
    x[12] += ic; x[13] += n0; x[14] += n1; x[15] += n2;
}

// ---------------------------------------------------------------- Poly1305 --
// 130-bit values in five 26-bit limbs; products with 32x32->64 multiply-adds.
SS_DEV void ss_pmul(uint32_t h[5], const uint32_t a[5], const uint32_t b[5]) {
    const uint32_t s1 = b[1] * 5u, s2 = b[2] * 5u, s3 = b[3] * 5u, s4 = b[4] * 5u;
    uint64_t d0 = (uint64_t)a[0] * b[0] + (uint64_t)a[1] * s4 + (uint64_t)a[2] * s3 + (uint64_t)a[3] * s2 + (uint64_t)a[4] * s1;
    uint64_t d1 = (uint64_t)a[0] * b[1] + (uint64_t)a[1] * b[0] + (uint64_t)a[2] * s4 + (uint64_t)a[3] * s3 + (uint64_t)a[4] * s2;
    uint64_t d2 = (uint64_t)a[0] * b[2] + (uint64_t)a[1] * b[1] + (uint64_t)a[2] * b[0] + (uint64_t)a[3] * s4 + (uint64_t)a[4] * s3;
    uint64_t d3 = (uint64_t)a[0] * b[3] + (uint64_t)a[1] * b[2] + (uint64_t)a[2] * b[1] + (uint64_t)a[3] * b[0] + (uint64_t)a[4] * s4;
    uint64_t d4 = (uint64_t)a[0] * b[4] + (uint64_t)a[1] * b[3] + (uint64_t)a[2] * b[2] + (uint64_t)a[3] * b[1] + (uint64_t)a[4] * b[0];
    uint32_t c;
    c = (uint32_t)(d0 >> 26); h[0] = (uint32_t)d0 & 0x3ffffffu; d1 += c;
    c = (uint32_t)(d1 >> 26); h[1] = (uint32_t)d1 & 0x3ffffffu; d2 += c;
    c = (uint32_t)(d2 >> 26); h[2] = (uint32_t)d2 & 0x3ffffffu; d3 += c;
    c = (uint32_t)(d3 >> 26); h[3] = (uint32_t)d3 & 0x3ffffffu; d4 += c;
    c = (uint32_t)(d4 >> 26); h[4] = (uint32_t)d4 & 0x3ffffffu;
    h[0] += c * 5u;
    c = h[0] >> 26; h[0] &= 0x3ffffffu; h[1] += c;
}

// d += a * P where P is stored as {p0..p4, 5*p1..5*p4} (9 words, shared memory in the bulk kernel)
SS_DEV void ss_pmac(uint64_t d[5], const uint32_t a[5], const uint32_t *P) {
    const uint32_t b0 = P[0], b1 = P[1], b2 = P[2], b3 = P[3], b4 = P[4], s1 = P[5], s2 = P[6], s3 = P[7], s4 = P[8];
    d[0] += (uint64_t)a[0] * b0 + (uint64_t)a[1] * s4 + (uint64_t)a[2] * s3 + (uint64_t)a[3] * s2 + (uint64_t)a[4] * s1;
    d[1] += (uint64_t)a[0] * b1 + (uint64_t)a[1] * b0 + (uint64_t)a[2] * s4 + (uint64_t)a[3] * s3 + (uint64_t)a[4] * s2;
    d[2] += (uint64_t)a[0] * b2 + (uint64_t)a[1] * b1 + (uint64_t)a[2] * b0 + (uint64_t)a[3] * s4 + (uint64_t)a[4] * s3;
    d[3] += (uint64_t)a[0] * b3 + (uint64_t)a[1] * b2 + (uint64_t)a[2] * b1 + (uint64_t)a[3] * b0 + (uint64_t)a[4] * s4;
    d[4] += (uint64_t)a[0] * b4 + (uint64_t)a[1] * b3 + (uint64_t)a[2] * b2 + (uint64_t)a[3] * b1 + (uint64_t)a[4] * b0;
}
SS_DEV void ss_store_power(uint32_t *P, const uint32_t v[5]) {
    P[0] = v[0]; P[1] = v[1]; P[2] = v[2]; P[3] = v[3]; P[4] = v[4];
    P[5] = v[1] * 5u; P[6] = v[2] * 5u; P[7] = v[3] * 5u; P[8] = v[4] * 5u;
}
// limbs of a full 16-byte block (2^128 bit set)
SS_DEV void ss_block_limbs(uint32_t t0, uint32_t t1, uint32_t t2, uint32_t t3, uint32_t l[5]) {
    l[0] = t0 & 0x3ffffffu;
    l[1] = ss_fsr(t0, t1, 26) & 0x3ffffffu;
    l[2] = ss_fsr(t1, t2, 20) & 0x3ffffffu;
    l[3] = ss_fsr(t2, t3, 14) & 0x3ffffffu;
    l[4] = (t3 >> 8) | (1u << 24);
}

// acc <- (acc + block) * r for the 16-byte block (t0..t3) with the given 2^128 bit.
SS_DEV void ss_pblock(uint32_t acc[5], uint32_t t0, uint32_t t1, uint32_t t2, uint32_t t3, uint32_t hibit,
                      const uint32_t r[5]) {
    uint32_t a[5];
    a[0] = acc[0] + (t0 & 0x3ffffffu);
    a[1] = acc[1] + (ss_fsr(t0, t1, 26) & 0x3ffffffu);
    a[2] = acc[2] + (ss_fsr(t1, t2, 20) & 0x3ffffffu);
    a[3] = acc[3] + (ss_fsr(t2, t3, 14) & 0x3ffffffu);
    a[4] = acc[4] + ((t3 >> 8) | hibit);
    ss_pmul(acc, a, r);
}

// Horner over the first `nbytes` (0..64) bytes of a 64-byte window.  Bytes past
// nbytes must already be zero except for the 0x01 terminator of a partial block.
// FULL_FAST: take a test-free path for full windows.  Only for bulk tiles, where all lanes of the warp
// have full windows together; in packed tiles the lanes differ and a second path would just be executed
// in addition to the first (measured: C2 5.1 -> 5.8 us per step).
template <bool FULL_FAST = false>
SS_DEV void ss_phorner(uint32_t acc[5], const uint32_t w[16], uint32_t nbytes, const uint32_t r[5]) {
    if (FULL_FAST && nbytes == 64u) {
#pragma unroll
        for (int i = 0; i < 4; i++) ss_pblock(acc, w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3], 1u << 24, r);
        return;
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if ((uint32_t)(16 * i) < nbytes) {
            const uint32_t hibit = ((uint32_t)(16 * i + 16) <= nbytes) ? (1u << 24) : 0u;
            ss_pblock(acc, w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3], hibit, r);
        }
    }
}

// OR the 4-byte little-endian value `val` into the window at byte offset `o`
// (any alignment, may hang over either edge).
SS_DEV void ss_insert(uint32_t w[16], int o, uint32_t val) {
    // branch-free: the value lands in word o>>2 (low part) and the next one (spill-over); selects, no
    // divergent branches -- this routine was the top stall site of the packed kernel in ncu's source view
    const int wi = o >> 2;
    const uint32_t sh = (uint32_t)(o & 3) * 8u;
    const uint32_t lo = val << sh;
    const uint32_t hi = sh ? (val >> (32u - sh)) : 0u;
#pragma unroll
    for (int i = 0; i < 16; i++) w[i] |= (i == wi ? lo : 0u) | (i == wi + 1 ? hi : 0u);
}

// Zero the bytes of the window at offsets >= n (n may be <= 0 or >= 64).
SS_DEV void ss_keep_bytes(uint32_t w[16], int n) {
    // branch-free: per word a mask of 0, a partial byte mask, or all ones (selects, no divergent branches)
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const int nb = n - 4 * i;
        const uint32_t part = (1u << (8 * (nb & 3))) - 1u;
        w[i] &= nb <= 0 ? 0u : (nb >= 4 ? 0xffffffffu : part);
    }
}

// Full reduction mod 2^130-5, + s, -> 16-byte tag as four words.
SS_DEV void ss_pfinal(const uint32_t hin[5], const uint32_t s[4], uint32_t mac[4]) {
    uint32_t h0 = hin[0], h1 = hin[1], h2 = hin[2], h3 = hin[3], h4 = hin[4], c;
    c = h0 >> 26; h0 &= 0x3ffffffu; h1 += c;
    c = h1 >> 26; h1 &= 0x3ffffffu; h2 += c;
    c = h2 >> 26; h2 &= 0x3ffffffu; h3 += c;
    c = h3 >> 26; h3 &= 0x3ffffffu; h4 += c;
    c = h4 >> 26; h4 &= 0x3ffffffu; h0 += c * 5u;
    c = h0 >> 26; h0 &= 0x3ffffffu; h1 += c;
    c = h1 >> 26; h1 &= 0x3ffffffu; h2 += c;
    c = h2 >> 26; h2 &= 0x3ffffffu; h3 += c;
    c = h3 >> 26; h3 &= 0x3ffffffu; h4 += c;
    c = h4 >> 26; h4 &= 0x3ffffffu; h0 += c * 5u;
    c = h0 >> 26; h0 &= 0x3ffffffu; h1 += c;
    uint32_t g0 = h0 + 5u; c = g0 >> 26; g0 &= 0x3ffffffu;
    uint32_t g1 = h1 + c; c = g1 >> 26; g1 &= 0x3ffffffu;
    uint32_t g2 = h2 + c; c = g2 >> 26; g2 &= 0x3ffffffu;
    uint32_t g3 = h3 + c; c = g3 >> 26; g3 &= 0x3ffffffu;
    uint32_t g4 = h4 + c - (1u << 26);
    const uint32_t m = (g4 >> 31) - 1u;  // all ones when h >= p
    h0 = (h0 & ~m) | (g0 & m); h1 = (h1 & ~m) | (g1 & m); h2 = (h2 & ~m) | (g2 & m);
    h3 = (h3 & ~m) | (g3 & m); h4 = (h4 & ~m) | (g4 & m);
    const uint32_t w0 = h0 | (h1 << 26), w1 = (h1 >> 6) | (h2 << 20), w2 = (h2 >> 12) | (h3 << 14),
                   w3 = (h3 >> 18) | (h4 << 8);
    uint64_t f;
    f = (uint64_t)w0 + s[0]; mac[0] = (uint32_t)f;
    f = (uint64_t)w1 + s[1] + (f >> 32); mac[1] = (uint32_t)f;
    f = (uint64_t)w2 + s[2] + (f >> 32); mac[2] = (uint32_t)f;
    f = (uint64_t)w3 + s[3] + (f >> 32); mac[3] = (uint32_t)f;
}

// Clamp r (RFC 8439 2.5) and split into limbs.
SS_DEV void ss_r_limbs(uint32_t t0, uint32_t t1, uint32_t t2, uint32_t t3, uint32_t r[5]) {
    t0 &= 0x0fffffffu; t1 &= 0x0ffffffcu; t2 &= 0x0ffffffcu; t3 &= 0x0ffffffcu;
    r[0] = t0 & 0x3ffffffu;
    r[1] = ss_fsr(t0, t1, 26) & 0x3ffffffu;
    r[2] = ss_fsr(t1, t2, 20) & 0x3ffffffu;
    r[3] = ss_fsr(t2, t3, 14) & 0x3ffffffu;
    r[4] = t3 >> 8;
}

// libsodium rekey (SURVEY 8(a) spec box): k, inonce <- (k ‖ inonce) ^ ChaCha20(k, n, ic=0)[0:40]; counter <- 1.
#ifndef SS_HOST_EMU
SS_NOINLINE
#else
static
#endif
void ss_rekey(uint32_t k[8], uint32_t nn[3]) {
    uint32_t x[16];
    ss_chacha_block(k, 0, nn[0], nn[1], nn[2], x);
#pragma unroll
    for (int i = 0; i < 8; i++) k[i] ^= x[i];
    nn[1] ^= x[8];
    nn[2] ^= x[9];
    nn[0] = 1;
}

// ------------------------------------------------------- payload load/store --
// Loads up to 64 bytes (16-byte pieces that start below nbytes) from a 16-aligned address.
template <bool FULL_FAST = false, bool VOL = false>
SS_DEV void ss_load64(const uint8_t *p, int nbytes, uint32_t m[16]) {
    const uint4 *q = reinterpret_cast<const uint4 *>(p);
    if (FULL_FAST && nbytes == 64) {    // full block: four unconditional vector loads
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint4 v = ss_ld16<VOL>(q + i);
            m[4 * i] = v.x; m[4 * i + 1] = v.y; m[4 * i + 2] = v.z; m[4 * i + 3] = v.w;
        }
        return;
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        uint4 v;
        v.x = v.y = v.z = v.w = 0;
        if (16 * i < nbytes) v = ss_ld16<VOL>(q + i);
        m[4 * i] = v.x; m[4 * i + 1] = v.y; m[4 * i + 2] = v.z; m[4 * i + 3] = v.w;
    }
}

// Stores exactly nbytes (1..64) bytes to a 16-aligned address: whole 16-byte
// pieces as one vector store, the ragged tail byte by byte (the MAC or the next
// frame's bytes may follow immediately).
template <bool FULL_FAST = false>
SS_DEV void ss_store64(uint8_t *p, int nbytes, const uint32_t m[16]) {
    uint4 *q = reinterpret_cast<uint4 *>(p);
    if (FULL_FAST && nbytes == 64) {    // full block: four vector stores, none of the ragged-tail tests
#pragma unroll
        for (int i = 0; i < 4; i++) {
            uint4 v;
            v.x = m[4 * i]; v.y = m[4 * i + 1]; v.z = m[4 * i + 2]; v.w = m[4 * i + 3];
            q[i] = v;
        }
        return;
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if (16 * i + 16 <= nbytes) {
            uint4 v;
            v.x = m[4 * i]; v.y = m[4 * i + 1]; v.z = m[4 * i + 2]; v.w = m[4 * i + 3];
            q[i] = v;
        } else if (16 * i < nbytes) {
            // ragged tail of 1..15 bytes as at most one 8-, 4-, 2- and 1-byte store (naturally aligned because
            // the piece is), instead of up to fifteen byte stores
            const uint32_t rem = (uint32_t)(nbytes - 16 * i);
            uint8_t *t = p + 16 * i;
            const uint32_t p0 = m[4 * i], p1 = m[4 * i + 1], p2 = m[4 * i + 2], p3 = m[4 * i + 3];
            uint32_t off = 0;
            if (rem & 8u) { *reinterpret_cast<uint2 *>(t) = make_uint2(p0, p1); off = 8; }
            if (rem & 4u) { *reinterpret_cast<uint32_t *>(t + off) = off ? p2 : p0; off += 4; }
            const uint32_t wsel = off >= 12u ? p3 : (off >= 8u ? p2 : (off >= 4u ? p1 : p0));   // word that holds byte `off`
            if (rem & 2u) { *reinterpret_cast<uint16_t *>(t + off) = (uint16_t)wsel; off += 2; }
            if (rem & 1u) t[off] = (uint8_t)(wsel >> (8u * (off & 3u)));
        }
    }
}

SS_DEV void ss_store_mac(uint8_t *p, const uint32_t mac[4]) {
    if ((reinterpret_cast<uintptr_t>(p) & 3u) == 0) {
        uint32_t *q = reinterpret_cast<uint32_t *>(p);
        q[0] = mac[0]; q[1] = mac[1]; q[2] = mac[2]; q[3] = mac[3];
    } else {
#pragma unroll
        for (int j = 0; j < 16; j++) p[j] = (uint8_t)(mac[j >> 2] >> (8 * (j & 3)));
    }
}

// The 16 tag bytes of a wire frame (any alignment) as four little-endian words.
template <bool VOL = false>
SS_DEV void ss_load_mac(const uint8_t *p, uint32_t w[4]) {
    if ((reinterpret_cast<uintptr_t>(p) & 3u) == 0) {
        const uint32_t *q = reinterpret_cast<const uint32_t *>(p);
        w[0] = ss_ld4<VOL>(q); w[1] = ss_ld4<VOL>(q + 1); w[2] = ss_ld4<VOL>(q + 2); w[3] = ss_ld4<VOL>(q + 3);
    } else {
        w[0] = w[1] = w[2] = w[3] = 0;
#pragma unroll
        for (int j = 0; j < 16; j++) w[j >> 2] |= ss_ld1<VOL>(p + j) << (8 * (j & 3));
    }
}

// --------------------------------------------------------------- the tile ---
// One warp processes one tile.  BULK == false: up to 16 small frames packed in the warp, one lane per 64-byte
// block, one pass.  BULK == true ("strided"): every frame of the tile gets segw lanes (2, 4, 8, 16 or all 32: bits
// 24..29 of n_kind, 0 = 32) which walk its blocks with stride segw -- one big frame per warp, or 32/segw mid-size
// frames of one block count side by side.
// `given`: the tile itself when the launch carried the tile list in its parameters (small host batches:
// one dependent read over PCIe fewer), else it is read from a.tiles.
// SEG (with BULK): the tile may hold several frames with fewer than 32 lanes each; SEG == false keeps the one-frame-per-
// warp case free of the bookkeeping (its register budget is tight: 72, seven blocks of 128 threads per SM).
// WIDE (with BULK, not SEG): one frame per 128-thread BLOCK -- the four warps walk the frame's blocks with stride 128
// and combine their partial MACs through shared memory.  For the few frames of 8 KiB and more in a ragged batch:
// one warp needs nine rounds for 16 KiB, which is what such a batch then waits for (BASELINE config 3); four warps
// need three.  A batch of many big frames (config 4) is throughput-bound and keeps one warp per frame.
template <bool BULK, bool VOL = false, bool SEG = false, bool WIDE = false>
SS_DEV void ss_process_tile(const SsArgs &a, uint32_t tile_idx, const SsTile *given = nullptr) {
    const uint32_t lane = ss_lane();
    const uint32_t fl = WIDE ? ss_wide_tid() : 0u;   // wide tiles: this thread's lane number among the frame's 128
    ss_dbg_stamp(a, tile_idx, lane, 0);
    ss_pdl_launch_dependents();   // the next batch's kernel may start its own prologue now
    if (!(a.flags & SS_ARG_INPUT_INDEPENDENT)) ss_pdl_wait();   // descriptors may come from the previous kernel
    const SsTile tile = given ? *given : a.tiles[tile_idx];
    const uint32_t nfr = (BULK && !SEG) ? 1u : (tile.n_kind & 0xffffu);
    const uint32_t segw = (BULK && SEG) ? (((tile.n_kind >> 24) & 63u) ? ((tile.n_kind >> 24) & 63u) : 32u) : 32u;   // lanes per frame (strided tiles)

    // ---- lane -> (frame f, first lane st of that frame) ---------------------
    uint32_t d_slot = 0, d_in = 0, d_out = 0, d_lf = 0, my_nc = 0, d_skip = 0, d_ver = 0;
    if (lane < nfr) {
        const uint4 d = ss_ld16<VOL>(a.desc + tile.first + lane);
        d_slot = d.x; d_in = d.y; d_out = d.z; d_lf = d.w;
        my_nc = 2u + (((d_lf & 0xffffffu) + 63u) >> 6);
        if (!BULK && a.desc_fill) *reinterpret_cast<uint4 *>(a.desc_fill + tile.first + lane) = d;
        if (!BULK && (a.flags & SS_ARG_CHAIN)) d_ver = (a.flags & SS_ARG_CHAIN_UNIFORM) ? a.ver_all : a.ver[tile.first + lane];
    }
    uint32_t f = 0, st = 0, my_start = 0;
    bool valid = true;
    if (BULK && SEG) {
        f = lane / segw;           // segw is a power of two
        st = f * segw;
        valid = f < nfr;
        if (!valid) f = 0;
    }
    if (!BULK) {
        uint32_t incl = my_nc;
#pragma unroll
        for (uint32_t o = 1; o < 32; o <<= 1) {
            const uint32_t t = ss_shfl_up(incl, o);
            if (lane >= o) incl += t;
        }
        my_start = incl - my_nc;   // first lane of frame #lane (meaningful for lane < nfr)
        const uint32_t startmask = ss_reduce_or(lane < nfr ? (1u << (incl - my_nc)) : 0u);
        const uint32_t total = ss_shfl(incl, 31);
        const uint32_t below = startmask & (0xffffffffu >> (31u - lane));
        f = ss_popc(below) - 1u;
        st = 31u - ss_clz(below);
        valid = lane < total;
    }
    const uint32_t slot = ss_shfl(d_slot, f);
    const uint32_t in_off = ss_shfl(d_in, f);
    const uint32_t out_off = ss_shfl(d_out, f);
    const uint32_t lf = ss_shfl(d_lf, f);
    const uint32_t exp_ver = (!BULK && (a.flags & SS_ARG_CHAIN)) ? ss_shfl(d_ver, f) : 0u;   // (the resident kernel runs packed and thread tiles only)

    const uint32_t mlen = lf & 0xffffffu;
    const bool is_open = (lf & SS_F_OPEN) != 0;
    const uint32_t nc = 2u + ((mlen + 63u) >> 6);          // ChaCha20 blocks of this frame
    const uint32_t T = 80u + mlen + (mlen & 15u);          // MAC input bytes (libsodium pad quirk)
    const uint32_t NB = (T + 15u) >> 4;                    // Poly1305 blocks
    const uint32_t V = 64u + mlen;                         // LE64(64+mlen) sits at stream offset T-8
    const uint8_t *src = a.in + in_off + (is_open ? 4u : 0u);   // payload in  (16-aligned)  [may be re-pointed at the staged copy]
    uint8_t *dst = a.out + out_off + (is_open ? 0u : 4u);       // payload out (16-aligned)
    uint4 *sp = a.states + ((size_t)slot * 2u + (is_open ? 1u : 0u)) * 3u;

    // Packed tiles fetch everything they read from the arenas up front so the latency (HBM, or
    // PCIe when the arenas are pinned host memory) hides behind the ~1,000 ALU instructions of
    // ChaCha20.  When the caller vouches that the inputs were not produced by the preceding
    // kernel, that fetch even happens before the programmatic-dependency wait, i.e. while the
    // previous batch's kernel is still running.
    uint32_t m[16], hw = 0, wmac[4] = {0, 0, 0, 0};
    const uint32_t b_first = lane - st;
    const bool pre_payload = !BULK && valid && b_first >= 2u && b_first < nc;
    const int pre_cbytes = pre_payload ? ((int)(mlen - 64u * (b_first - 2u)) < 64 ? (int)(mlen - 64u * (b_first - 2u)) : 64) : 0;
    const bool early = !BULK && (a.flags & SS_ARG_INPUT_INDEPENDENT);
    // Zero-copy batches read the arena over PCIe, where the 16-byte-per-lane pattern below turns into
    // 32-byte requests and the link, not the SMs, sets the pace.  When all frames of the tile sit in one
    // window of <= 2 KiB (they do when the scheduler wrote them back to back) the warp fetches that window
    // with fully coalesced 512-byte loads into shared memory and every later read comes from there.
    const uint8_t *in_base = a.in;
    if (!BULK && (a.flags & SS_ARG_STAGE_INPUT)) {
        const uint32_t f_lo = (is_open ? in_off - 12u : in_off) & ~15u;
        const uint32_t f_hi = in_off + mlen + (is_open ? 20u : 0u);
        const uint32_t w_lo = ss_reduce_min(valid ? f_lo : 0xffffffffu);
        const uint32_t w_hi = ss_reduce_max(valid ? f_hi : 0u);
        if (w_hi > w_lo && w_hi - w_lo <= SS_STAGE_BYTES) {   // warp-uniform
            uint4 *stage = ss_stage_buffer();
            const uint4 *g = reinterpret_cast<const uint4 *>(a.in + w_lo);
            const uint32_t pieces = (w_hi - w_lo + 15u) >> 4;
#pragma unroll
            for (uint32_t i = 0; i < SS_STAGE_BYTES / 512u; i++) {
                const uint32_t p = lane + 32u * i;
                if (p < pieces) stage[p] = ss_ld16<VOL>(g + p);
            }
            ss_syncwarp();
            in_base = reinterpret_cast<const uint8_t *>(stage) - w_lo;
            src = in_base + in_off + (is_open ? 4u : 0u);
        }
    }
    if (early) {
        ss_load64<false, VOL>(src + 64u * (size_t)(pre_payload ? b_first - 2u : 0u), pre_cbytes, m);
        if (valid && is_open && b_first == 1u) hw = ss_ld4<VOL>(in_base + in_off);
        if (valid && is_open && b_first == 0u) ss_load_mac<VOL>(src + mlen, wmac);
    }
    ss_dbg_stamp(a, tile_idx, lane, 1);
    ss_pdl_wait();   // from here on we touch what the previous kernel may have written
    if (!BULK) ss_dep_wait(a.dep_seq, a.dep_need);   // (batch-level dependency; unused when a.ver is set)
    if (!BULK && (a.flags & SS_ARG_CHAIN)) {
        // resident kernel: every lane waits until its frame's predecessor has left the state (warp-uniform loop)
        const uint32_t *vp = reinterpret_cast<const uint32_t *>(sp + 2) + 3;
        bool ready = !valid || exp_ver == 0u;
        for (;;) {
            if (!ready) ready = (ss_ld_acquire_u32(vp) >> 8) == exp_ver;
            if (!ss_any(!ready)) break;
            ss_nanosleep(40);
        }
    }
    ss_dbg_stamp(a, tile_idx, lane, 2);
    if (!BULK) ss_write_back(a, tile_idx, a.n_tiles, lane);

    // A frame the extractor refused (TypeError in the reference, or a shape it leaves to the host)
    // is not sealed: provider.ts:243-248 throws before peer.write(chunk).
    if (lane < nfr && (d_lf & SS_F_EXTRACT) && !(d_lf & SS_F_OPEN)) d_skip = a.status[tile.first + lane] != SS_ST_OK;
    // A run frame without a speculative nonce -- the entry point got no scratch, or the frame sits beyond
    // SS_RUN_MAX in its run (the pre-pass flags those) -- is not opened: RETRY, state untouched.
    if (lane < nfr && (d_lf & (SS_F_RUN | SS_F_OPEN)) == (SS_F_RUN | SS_F_OPEN)) {
        if (!a.run_nonce || a.run_nonce[tile.first + lane].w >= SS_RUN_CAP) {
            d_skip = 1;
            a.status[tile.first + lane] = SS_ST_RETRY;
        }
    }
    if (ss_shfl(d_skip, f)) valid = false;

    uint32_t k[8], nn[3], meta = 0;
    {
        uint4 s0, s1, s2;
        s0.x = s0.y = s0.z = s0.w = 0; s1 = s0; s2 = s0;
        if (valid) { s0 = ss_ld16<VOL>(sp); s1 = ss_ld16<VOL>(sp + 1); s2 = ss_ld16<VOL>(sp + 2); }
        k[0] = s0.x; k[1] = s0.y; k[2] = s0.z; k[3] = s0.w;
        k[4] = s1.x; k[5] = s1.y; k[6] = s1.z; k[7] = s1.w;
        nn[0] = s2.x; nn[1] = s2.y; nn[2] = s2.z; meta = s2.w;
    }
    // A frame of a run uses the nonce the pre-pass derived from the MACs of the frames queued before
    // it (SURVEY 8(a) "parallel open"); the stream state itself is advanced later by the commit pass.
    const bool in_run = (lf & SS_F_RUN) != 0 && is_open;
    if (valid && in_run) {
        const uint4 o = a.run_nonce[tile.first + f];
        nn[0] = o.x; nn[1] = o.y; nn[2] = o.z;
    }

    if (!early) {
        if (!BULK) {
            ss_load64<false, VOL>(src + 64u * (size_t)(pre_payload ? b_first - 2u : 0u), pre_cbytes, m);
            if (valid && is_open && b_first == 1u) hw = ss_ld4<VOL>(in_base + in_off);
        }
        if (!BULK && valid && is_open && lane == st) ss_load_mac<VOL>(src + mlen, wmac);   // the wire tag, long before it is compared
    }

    // Output side of the same idea: over PCIe, 16-byte stores at a 64-byte stride (plus 4-byte headers,
    // tags and status words) leave as tiny write transactions.  When the tile's outputs share one
    // <= 2 KiB window the lanes assemble them in shared memory (the input copy is no longer needed: every
    // lane holds its input in registers by now) and the warp stores the window in whole 512-byte rows.
    // Bytes between the frames of that window are written as zeros, which the host-arena entry points
    // allow (they never promised to preserve gaps).
    uint8_t *out_base = a.out;
    bool out_staged = false;
    uint32_t ow_lo = 0, ow_pieces = 0;
    if (!BULK && (a.flags & SS_ARG_STAGE_OUTPUT)) {
        const uint32_t o_lo = (is_open ? out_off : out_off - 12u) & ~15u;
        const uint32_t o_hi = out_off + mlen + (is_open ? 0u : 20u);
        const uint32_t w_lo = ss_reduce_min(valid ? o_lo : 0xffffffffu);
        const uint32_t w_hi = ss_reduce_max(valid ? o_hi : 0u);
        if (w_hi > w_lo && w_hi - w_lo <= SS_STAGE_BYTES) {   // warp-uniform
            uint4 *stage = ss_stage_buffer();
            ow_lo = w_lo;
            ow_pieces = (w_hi - w_lo + 15u) >> 4;
            ss_syncwarp();   // all reads of the staged input are done
            uint4 zero;
            zero.x = zero.y = zero.z = zero.w = 0;
#pragma unroll
            for (uint32_t i = 0; i < SS_STAGE_BYTES / 512u; i++)
                if (lane + 32u * i < ow_pieces) stage[lane + 32u * i] = zero;
            ss_syncwarp();
            out_base = reinterpret_cast<uint8_t *>(stage) - w_lo;
            dst = out_base + out_off + (is_open ? 0u : 4u);
            out_staged = true;
        }
    }

    uint32_t A[5] = {0, 0, 0, 0, 0};   // lane partial, to be scaled by r^E
    uint32_t HX[5] = {0, 0, 0, 0, 0};  // bulk: partial of the final window (exponent 0)
    uint32_t E = 0;
    uint32_t r[5] = {0, 0, 0, 0, 0}, r128[5] = {0, 0, 0, 0, 0}, sw[4] = {0, 0, 0, 0};
    uint32_t tagval = 0;               // lane b==1: seal -> wire tag byte, open -> decrypted tag
    uint32_t keep[16];                 // packed open: plaintext held back until the MAC verdict
    bool have_A = false;
    // strided tiles walk virtual blocks 0..nc of every frame; the loop is warp-uniform (it holds shuffles), so it runs
    // for the longest frame of the tile (the planner groups frames of one block count)
    const uint32_t niter = !BULK ? 1u : (SEG ? ss_reduce_max(valid ? (nc + segw) / segw : 0u)
                                             : (WIDE ? ((nc + 1u + 127u) >> 7) : ((nc + 1u + 31u) >> 5)));
    const uint32_t stride = WIDE ? 128u : (BULK ? segw : 32u);

#pragma unroll 1
    for (uint32_t it = 0; it < niter; ++it) {
        const uint32_t b = (WIDE ? fl : lane - st) + stride * it;
        const bool has_block = valid && b < nc;
        const bool need_ks = has_block;
        // (Bulk tiles load after the rounds: they have enough warps in flight, and holding the
        // payload registers across ChaCha20 cost them occupancy -- 65.8 % -> 62.9 % of the ALU peak.)
        const bool has_payload = has_block && b >= 2u;
        const int cbytes = has_payload ? ((int)(mlen - 64u * (b - 2u)) < 64 ? (int)(mlen - 64u * (b - 2u)) : 64) : 0;
        uint32_t x[16];
#pragma unroll
        for (int i = 0; i < 16; i++) x[i] = 0;
        if (need_ks) ss_chacha_block(k, b, nn[0], nn[1], nn[2], x);
        if (BULK) {
            ss_load64<true>(src + 64u * (size_t)(has_payload ? b - 2u : 0u), cbytes, m);
            if (valid && b == 1u && is_open) hw = *reinterpret_cast<const uint32_t *>(a.in + in_off);
        }

        if (it == 0) {
            // Poly1305 key: r to every lane of the frame, s stays on the leader.
            uint32_t t0, t1, t2, t3;
            if (WIDE) {   // block 0 was computed by thread 0 of the block: hand r to the other warps through shared memory
                uint32_t *wb = ss_wide_buffer();
                if (fl == 0u) { wb[0] = x[0]; wb[1] = x[1]; wb[2] = x[2]; wb[3] = x[3]; }
                ss_wide_sync();
                t0 = wb[0]; t1 = wb[1]; t2 = wb[2]; t3 = wb[3];
            } else {
                t0 = ss_shfl(x[0], st); t1 = ss_shfl(x[1], st); t2 = ss_shfl(x[2], st); t3 = ss_shfl(x[3], st);
            }
            ss_r_limbs(t0, t1, t2, t3, r);
            sw[0] = x[4]; sw[1] = x[5]; sw[2] = x[6]; sw[3] = x[7];
            if (BULK) {
                uint32_t t[5], t3[5];
                uint32_t *pw = ss_powers_buffer() + 48u * f;
                ss_pmul(t, r, r);        // r^2
                ss_pmul(t3, t, r);       // r^3
                ss_pmul(r128, t, t);     // r^4
                ss_syncwarp();           // (the previous tile of this warp is done with the buffer)
                if (valid && lane == st) {   // every lane of a frame holds the same values: one writer is enough
                    ss_store_power(pw, r); ss_store_power(pw + 9, t); ss_store_power(pw + 18, t3); ss_store_power(pw + 27, r128);
                }
#pragma unroll 1
                for (uint32_t w = 1; w < (WIDE ? 128u : segw); w <<= 1) {  // r^8 .. r^(4*stride): the step between a lane's consecutive windows
                    ss_pmul(t, r128, r128);
#pragma unroll
                    for (int j = 0; j < 5; j++) r128[j] = t[j];
                }
                if (valid && lane == st) ss_store_power(pw + 36, r128);
                ss_syncwarp();
            }
        }

        // ---- this lane's 64-byte window of the MAC input --------------------
        // stream = B(64) ‖ c(mlen) ‖ 0^(mlen%16) ‖ LE64(0) ‖ LE64(64+mlen); window q covers [64q, 64q+64)
        bool has_win = false;
        uint32_t q = 0;
        if (valid) {
            if (b == 0) {
                if (!BULK) { has_win = true; q = nc - 1u; }       // packed: the key lane takes the overflow window
#pragma unroll
                for (int i = 0; i < 16; i++) x[i] = 0;
            } else if (b == 1) {
                has_win = true; q = 0;
                if (!is_open) {
                    x[0] ^= (lf >> 24) & 3u;                        // (tag ‖ 0^63) ^ keystream
                    tagval = x[0] & 0xffu;
                    *reinterpret_cast<uint32_t *>(out_base + out_off) = (mlen + 17u) | (tagval << 24);
                } else {
                    const uint32_t c0 = hw >> 24;
                    tagval = (x[0] ^ c0) & 0xffu;                  // decrypted tag
                    x[0] = (x[0] & 0xffffff00u) | c0;              // MAC covers the wire byte
                }
            } else if (b < nc) {
                has_win = true; q = b - 1u;
                const uint32_t j = b - 2u;
                if (!is_open) {
#pragma unroll
                    for (int i = 0; i < 16; i++) x[i] ^= m[i];     // ciphertext
                    ss_store64<BULK>(dst + 64u * (size_t)j, cbytes, x);
                } else {
                    if (!BULK) {
#pragma unroll
                        for (int i = 0; i < 16; i++) keep[i] = x[i] ^ m[i];
                    } else {
                        // Bulk frames cannot hold their plaintext in registers until the verdict, and a second
                        // pass re-read the whole ciphertext from HBM (ncu: 3.28 GB read for 2.15 GB of input).
                        // So decrypt now, in the same pass as the MAC, and wipe the output if the tag turns out
                        // wrong: no unauthenticated byte is ever visible to the caller.
#pragma unroll
                        for (int i = 0; i < 16; i++) x[i] ^= m[i];
                        ss_store64<true>(dst + 64u * (size_t)j, cbytes, x);
                    }
#pragma unroll
                    for (int i = 0; i < 16; i++) x[i] = m[i];      // MAC is over the ciphertext
                }
                if (cbytes < 64) ss_keep_bytes(x, cbytes);
            } else if (BULK && b == nc) {
                has_win = true; q = nc - 1u;                        // bulk: virtual block = overflow window
            }
        }
        const int so = (int)(64u * q);
        int nbytes = has_win ? (int)T - so : 0;
        nbytes = nbytes < 0 ? 0 : (nbytes > 64 ? 64 : nbytes);
        if (nbytes > 0) {
            // Only the windows at the very end of the MAC input carry the length word / the terminator;
            // the full payload windows of a bulk frame skip ~250 instructions of fix-up here.
            const int vo = (int)T - 8 - so;
            if (vo < 64 && vo > -4) ss_insert(x, vo, V);
            if ((T & 15u) && (int)T - so < 64) ss_insert(x, (int)T - so, 1u);   // Poly1305 terminator of the partial last block
            const bool final_win = (uint32_t)so + 64u >= T;
            if (BULK && nbytes == 64 && !final_win) {
                // Full window of a bulk frame: A <- A*r^128 + b0*r^4 + b1*r^3 + b2*r^2 + b3*r as FIVE products
                // accumulated in 64-bit and ONE carry pass (instead of five multiply-reduce rounds); the
                // powers come from shared memory so they cost no registers.
                const uint32_t *pw = ss_powers_buffer() + 48u * f;
                uint64_t d[5] = {0, 0, 0, 0, 0};
                uint32_t l[5];
                if (have_A) ss_pmac(d, A, pw + 36);
                ss_block_limbs(x[0], x[1], x[2], x[3], l);     ss_pmac(d, l, pw + 27);
                ss_block_limbs(x[4], x[5], x[6], x[7], l);     ss_pmac(d, l, pw + 18);
                ss_block_limbs(x[8], x[9], x[10], x[11], l);   ss_pmac(d, l, pw + 9);
                ss_block_limbs(x[12], x[13], x[14], x[15], l); ss_pmac(d, l, pw);
                uint32_t c;
                c = (uint32_t)(d[0] >> 26); A[0] = (uint32_t)d[0] & 0x3ffffffu; d[1] += c;
                c = (uint32_t)(d[1] >> 26); A[1] = (uint32_t)d[1] & 0x3ffffffu; d[2] += c;
                c = (uint32_t)(d[2] >> 26); A[2] = (uint32_t)d[2] & 0x3ffffffu; d[3] += c;
                c = (uint32_t)(d[3] >> 26); A[3] = (uint32_t)d[3] & 0x3ffffffu; d[4] += c;
                const uint64_t c4 = d[4] >> 26; A[4] = (uint32_t)d[4] & 0x3ffffffu;
                const uint64_t h0 = (uint64_t)A[0] + c4 * 5u;
                A[0] = (uint32_t)h0 & 0x3ffffffu; A[1] += (uint32_t)(h0 >> 26);
                have_A = true;
                E = NB - 4u * q - 4u;
                continue;
            }
            uint32_t C[5] = {0, 0, 0, 0, 0};
            ss_phorner<BULK>(C, x, (uint32_t)nbytes, r);
            if (BULK && final_win) {
#pragma unroll
                for (int i = 0; i < 5; i++) HX[i] = C[i];
            } else {
                if (BULK && have_A) {
                    uint32_t t[5];
                    ss_pmul(t, A, r128);
#pragma unroll
                    for (int i = 0; i < 5; i++) A[i] = t[i] + C[i];
                } else {
#pragma unroll
                    for (int i = 0; i < 5; i++) A[i] = C[i];
                }
                have_A = true;
                E = final_win ? 0u : NB - 4u * q - 4u;             // Poly1305 blocks after this window
            }
        }
    }

    // ---- align lane partials: A <- A * r^E ----------------------------------
    {
        const uint32_t maxe = ss_reduce_max(E);
        const uint32_t bits = 32u - ss_clz(maxe);  // clz(0) == 32
        uint32_t base[5];
#pragma unroll
        for (int i = 0; i < 5; i++) base[i] = r[i];
#pragma unroll 1
        for (uint32_t bit = 0; bit < bits; ++bit) {
            const bool mine = (E >> bit) & 1u;
            if (ss_any(mine)) {   // warp-uniform branch, lanes select: no divergent region to reconverge from
                uint32_t t[5];
                ss_pmul(t, A, base);
#pragma unroll
                for (int i = 0; i < 5; i++) A[i] = mine ? t[i] : A[i];
            }
            if (bit + 1u < bits) {
                uint32_t t[5];
                ss_pmul(t, base, base);
#pragma unroll
                for (int i = 0; i < 5; i++) base[i] = t[i];
            }
        }
    }
    if (BULK) {
#pragma unroll
        for (int i = 0; i < 5; i++) A[i] += HX[i];
    }

    // ---- segmented sum of the limb vectors onto the leader lane -------------
    const uint32_t endl = BULK ? st + segw - 1u : st + nc - 1u;
#pragma unroll
    for (uint32_t o = 1; o < 32; o <<= 1) {
#pragma unroll
        for (int i = 0; i < 5; i++) {
            const uint32_t t = ss_shfl_down(A[i], o);
            if (lane + o <= endl) A[i] += t;
        }
    }
    const uint32_t tag_b1 = ss_shfl(tagval, st + 1u);  // st+1 <= 31 always (every frame has >= 2 lanes)
    if (WIDE) {   // the four warps' sums meet in shared memory; thread 0 of the block is the frame's leader
        uint32_t *wb = ss_wide_buffer();
        if (lane == 0u) {
            // a warp's sum of 32 partials uses up to 31 bits per limb: carry it back under 2^26 before four of them meet
            uint32_t c;
            c = A[0] >> 26; A[0] &= 0x3ffffffu; A[1] += c;
            c = A[1] >> 26; A[1] &= 0x3ffffffu; A[2] += c;
            c = A[2] >> 26; A[2] &= 0x3ffffffu; A[3] += c;
            c = A[3] >> 26; A[3] &= 0x3ffffffu; A[4] += c;
            c = A[4] >> 26; A[4] &= 0x3ffffffu; A[0] += c * 5u;
#pragma unroll
            for (int i = 0; i < 5; i++) wb[8u + 5u * (fl >> 5) + (uint32_t)i] = A[i];
        }
        ss_wide_sync();
        if (fl == 0u) {
#pragma unroll
            for (int i = 0; i < 5; i++) A[i] = wb[8 + i] + wb[13 + i] + wb[18 + i] + wb[23 + i];
        }
    }

    // ---- leader: tag, verdict, nonce chain -----------------------------------
    const bool leader = valid && (WIDE ? fl == 0u : lane == st);
    uint32_t ok = 1;
    uint32_t mac[4] = {0, 0, 0, 0};
    if (leader) {
        ss_pfinal(A, sw, mac);
        if (!is_open) ss_store_mac(dst + mlen, mac);
        else {
            if (BULK) ss_load_mac(src + mlen, wmac);   // bulk tiles keep it out of the loop's live registers
            ok = ((wmac[0] ^ mac[0]) | (wmac[1] ^ mac[1]) | (wmac[2] ^ mac[2]) | (wmac[3] ^ mac[3])) == 0 ? 1u : 0u;
        }
    }
    if (WIDE) {
        uint32_t *wb = ss_wide_buffer();
        if (fl == 0u) wb[4] = ok;
        ss_wide_sync();
        ok = wb[4];
    } else
    ok = ss_shfl(ok, st);

    if (is_open && ok) {
        if (!BULK) {
            const uint32_t b = lane - st;
            if (valid && b >= 2u && b < nc) {
                const uint32_t j = b - 2u;
                const int cbytes = (int)(mlen - 64u * j) < 64 ? (int)(mlen - 64u * j) : 64;
                ss_store64(dst + 64u * (size_t)j, cbytes, keep);
            }
        }
    }
    if (BULK) {
        if (valid && is_open && !ok) {   // forged frame: wipe what was decrypted speculatively
            uint32_t z[16];
#pragma unroll
            for (int i = 0; i < 16; i++) z[i] = 0;
#pragma unroll 1
            for (uint32_t it = 0; it < niter; ++it) {
                const uint32_t b = (WIDE ? fl : lane - st) + stride * it;
                if (b >= 2u && b < nc) {
                    const uint32_t j = b - 2u;
                    const int cbytes = (int)(mlen - 64u * j) < 64 ? (int)(mlen - 64u * j) : 64;
                    ss_store64(dst + 64u * (size_t)j, cbytes, z);
                }
            }
        }
    }

    uint32_t stword = SS_ST_BAD_MAC;
    if (leader) {
        if (ok && in_run) {
            stword = SS_ST_OK | (tag_b1 << 8);   // state untouched: ss_run_commit_kernel owns it
        } else if (ok) {
            const uint32_t tag = is_open ? tag_b1 : ((lf >> 24) & 3u);
            nn[1] ^= mac[0];
            nn[2] ^= mac[1];
            nn[0] += 1u;
            if ((tag & 2u) || nn[0] == 0u) {
                ss_rekey(k, nn);
                uint4 s0, s1;
                s0.x = k[0]; s0.y = k[1]; s0.z = k[2]; s0.w = k[3];
                s1.x = k[4]; s1.y = k[5]; s1.z = k[6]; s1.w = k[7];
                sp[0] = s0; sp[1] = s1;
            }
            uint4 s2;
            s2.x = nn[0]; s2.y = nn[1]; s2.z = nn[2]; s2.w = meta;
            sp[2] = s2;
            stword = SS_ST_OK | (tag << 8);
        }
        // resident kernel: hand the state on -- also when the frame failed, its successor is waiting for the stamp
        if (!BULK && (a.flags & SS_ARG_CHAIN)) ss_st_release_u32(reinterpret_cast<uint32_t *>(sp + 2) + 3, (meta & 0xffu) | (a.new_ver << 8));
        if (BULK) a.status[tile.first + f] = stword;
    }
    ss_dbg_stamp(a, tile_idx, lane, 5);
    if (!BULK) {
        // lane i collects the status of frame i from that frame's leader: one coalesced store per tile
        const uint32_t v = ss_shfl(stword, lane < nfr ? my_start : 0u);
        if (lane < nfr && !d_skip) a.status[tile.first + lane] = v;
        if (out_staged) {
            ss_syncwarp();   // every lane's pieces are in shared memory
            const uint4 *stage = ss_stage_buffer();
            uint4 *g = reinterpret_cast<uint4 *>(a.out + ow_lo);
#pragma unroll
            for (uint32_t i = 0; i < SS_STAGE_BYTES / 512u; i++)
                if (lane + 32u * i < ow_pieces) g[lane + 32u * i] = stage[lane + 32u * i];
        }
    }
    ss_dbg_stamp(a, tile_idx, lane, 3);
}

// ------------------------------------------------------ thread-per-frame tile --
// For large batches of SMALL frames (<= 4 payload blocks) the packed mapping leaves throughput on
// the table: its lane-parallel Poly1305 (window fix-ups, r^E alignment, shuffle tree) and per-warp
// bookkeeping cost about as many instructions as ChaCha20 itself, and the block counts rarely fill
// all 32 lanes.  Here one LANE owns one whole frame: its ChaCha20 blocks run back to back and
// Poly1305 is plain Horner with r -- no collectives at all, every lane busy when the frames of a
// warp have the same block count (the planner only forms these tiles from such frames).  Latency per
// frame is higher (blocks are serial), so the planner uses it only when a batch is big enough to be
// throughput-bound; small batches (BASELINE config 2) stay on packed tiles.
SS_DEV void ss_thread_window(uint32_t acc[5], uint32_t w[16], uint32_t q, uint32_t T, uint32_t V, const uint32_t r[5]) {
    const int so = (int)(64u * q);
    int nbytes = (int)T - so;
    nbytes = nbytes < 0 ? 0 : (nbytes > 64 ? 64 : nbytes);
    if (nbytes <= 0) return;
    const int vo = (int)T - 8 - so;
    if (vo < 64 && vo > -4) ss_insert(w, vo, V);
    if ((T & 15u) && (int)T - so < 64) ss_insert(w, (int)T - so, 1u);
    ss_phorner(acc, w, (uint32_t)nbytes, r);
}

// LATE = 1: the instantiation that understands SS_ARG_VERDICT_LATE (kept out of the kernels that never see it: the extra
// live values cost ss_thread_kernel a spill and BASELINE config 3 two microseconds).  LATE = 2: the tile runs inside
// the fused extract+seal kernel: same order of work, but the verdict of frame `lane` arrives in a.verdicts[lane]
// (shared memory, SS_VERDICT_EMPTY until the chunk's bookkeeping thread in the same block has written it).
#define SS_VERDICT_EMPTY 0xffffffffu
template <bool VOL = false, int LATE = 0>
SS_DEV void ss_process_thread_tile(const SsArgs &a, uint32_t tile_idx, const SsTile *given = nullptr) {
    const uint32_t lane = ss_lane();
    ss_pdl_launch_dependents();
    // SS_ARG_VERDICT_LATE: the kernel this one depends on is the batch's extractor, which writes status words and deltas
    // only.  Both kernels are one serial chain per frame with a few warps per SM, so they run side by side: this one seals
    // while the verdicts are being worked out and waits for them at the very end (a refused chunk is then taken back).
    const bool late_ok = LATE == 2 || (LATE == 1 && (a.flags & SS_ARG_VERDICT_LATE) != 0);
#ifdef SS_PROBE_TIMELINE
    if (LATE && tile_idx == 0 && lane == 0) PROBE_SET(2);
    if (LATE && lane == 0) PROBE_MAX(8);
#endif
    if (!(a.flags & SS_ARG_INPUT_INDEPENDENT) && !late_ok) ss_pdl_wait();
    const SsTile tile = given ? *given : a.tiles[tile_idx];
    const uint32_t nfr = tile.n_kind & 0xffffu;
    bool valid = lane < nfr;
    SsDesc d;
    d.slot = d.in_off = d.out_off = d.len_flags = 0;
    if (valid) {
        const uint4 dv = ss_ld16<VOL>(a.desc + tile.first + lane);
        d.slot = dv.x; d.in_off = dv.y; d.out_off = dv.z; d.len_flags = dv.w;
        if (a.desc_fill) *reinterpret_cast<uint4 *>(a.desc_fill + tile.first + lane) = dv;
    }
    if (!late_ok) ss_pdl_wait();
    ss_dep_wait(a.dep_seq, a.dep_need);
    ss_write_back(a, tile_idx, a.n_tiles, lane);
    const bool chain = (a.flags & SS_ARG_CHAIN) != 0;
    if (chain) {   // resident kernel: wait for this frame's predecessor (see ss_process_tile)
        const uint32_t exp_ver = !valid ? 0u : ((a.flags & SS_ARG_CHAIN_UNIFORM) ? a.ver_all : a.ver[tile.first + lane]);
        const uint32_t *vp = reinterpret_cast<const uint32_t *>(a.states + ((size_t)d.slot * 2u + ((d.len_flags & SS_F_OPEN) ? 1u : 0u)) * 3u + 2u) + 3;
        bool ready = !valid || exp_ver == 0u;
        for (;;) {
            if (!ready) ready = (ss_ld_acquire_u32(vp) >> 8) == exp_ver;
            if (!ss_any(!ready)) break;
            ss_nanosleep(40);
        }
    }
    const uint32_t lf = d.len_flags;
    const bool wants_verdict = (lf & SS_F_EXTRACT) && !(lf & SS_F_OPEN);
    if (valid && wants_verdict && !late_ok && a.status[tile.first + lane] != SS_ST_OK) valid = false;
    if (valid && (lf & (SS_F_RUN | SS_F_OPEN)) == (SS_F_RUN | SS_F_OPEN) &&
        (!a.run_nonce || a.run_nonce[tile.first + lane].w >= SS_RUN_CAP)) {   // see ss_process_tile
        a.status[tile.first + lane] = SS_ST_RETRY;
        valid = false;
    }
    if (!valid) return;
    const uint32_t mlen = lf & 0xffffffu;
    const bool is_open = (lf & SS_F_OPEN) != 0;
    const uint32_t npay = (mlen + 63u) >> 6;
    const uint32_t nc = 2u + npay;
    const uint32_t T = 80u + mlen + (mlen & 15u);
    const uint32_t V = 64u + mlen;
    const uint8_t *src = a.in + d.in_off + (is_open ? 4u : 0u);
    uint8_t *dst = a.out + d.out_off + (is_open ? 0u : 4u);
    uint4 *sp = a.states + ((size_t)d.slot * 2u + (is_open ? 1u : 0u)) * 3u;
    const uint4 s0 = ss_ld16<VOL>(sp), s1 = ss_ld16<VOL>(sp + 1), s2 = ss_ld16<VOL>(sp + 2);
    uint32_t k[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
    uint32_t nn[3] = {s2.x, s2.y, s2.z};
    const bool in_run = (lf & SS_F_RUN) != 0 && is_open;
    if (in_run) {
        const uint4 o = a.run_nonce[tile.first + lane];
        nn[0] = o.x; nn[1] = o.y; nn[2] = o.z;
    }
    uint32_t x[16], r[5], sw[4], acc[5] = {0, 0, 0, 0, 0};
    ss_chacha_block(k, 0, nn[0], nn[1], nn[2], x);            // Poly1305 key block
    ss_r_limbs(x[0], x[1], x[2], x[3], r);
    sw[0] = x[4]; sw[1] = x[5]; sw[2] = x[6]; sw[3] = x[7];
    ss_chacha_block(k, 1, nn[0], nn[1], nn[2], x);            // tag block
    uint32_t tag;
    if (!is_open) {
        tag = (lf >> 24) & 3u;
        x[0] ^= tag;
        *reinterpret_cast<uint32_t *>(a.out + d.out_off) = (mlen + 17u) | ((x[0] & 0xffu) << 24);
    } else {
        const uint32_t c0 = ss_ld4<VOL>(a.in + d.in_off) >> 24;
        tag = (x[0] ^ c0) & 0xffu;
        x[0] = (x[0] & 0xffffff00u) | c0;
    }
    ss_thread_window(acc, x, 0, T, V, r);
    // payload windows: seal encrypts and MACs in one pass; open MACs the ciphertext first
#pragma unroll 1
    for (uint32_t j = 0; j < npay; j++) {
        const int cbytes = (int)(mlen - 64u * j) < 64 ? (int)(mlen - 64u * j) : 64;
        uint32_t m[16];
        ss_load64<false, VOL>(src + 64u * (size_t)j, cbytes, m);
        if (!is_open) {
            ss_chacha_block(k, 2u + j, nn[0], nn[1], nn[2], x);
#pragma unroll
            for (int i = 0; i < 16; i++) m[i] ^= x[i];
            ss_store64(dst + 64u * (size_t)j, cbytes, m);
        }
        if (cbytes < 64) ss_keep_bytes(m, cbytes);
        ss_thread_window(acc, m, 1u + j, T, V, r);
    }
    if (64u * (nc - 1u) < T) {                                 // overflow window: the tail of the length block
#pragma unroll
        for (int i = 0; i < 16; i++) x[i] = 0;
        ss_thread_window(acc, x, nc - 1u, T, V, r);
    }
    uint32_t mac[4];
    ss_pfinal(acc, sw, mac);
    bool ok = true;
    if (!is_open) ss_store_mac(dst + mlen, mac);
    else {
        uint32_t wmac[4];
        ss_load_mac<VOL>(src + mlen, wmac);
        ok = ((wmac[0] ^ mac[0]) | (wmac[1] ^ mac[1]) | (wmac[2] ^ mac[2]) | (wmac[3] ^ mac[3])) == 0;
        if (ok) {
#pragma unroll 1
            for (uint32_t j = 0; j < npay; j++) {              // verified: decrypt (ciphertext is still in L1/L2)
                const int cbytes = (int)(mlen - 64u * j) < 64 ? (int)(mlen - 64u * j) : 64;
                uint32_t m[16];
                ss_load64<false, VOL>(src + 64u * (size_t)j, cbytes, m);
                ss_chacha_block(k, 2u + j, nn[0], nn[1], nn[2], x);
#pragma unroll
                for (int i = 0; i < 16; i++) m[i] ^= x[i];
                ss_store64(dst + 64u * (size_t)j, cbytes, m);
            }
        }
    }
    const uint32_t frame_idx = tile.first + lane;
#ifdef SS_PROBE_TIMELINE
    if (LATE && lane == 0) PROBE_MAX(5);
#endif
    if (late_ok) {
        uint32_t verdict = SS_ST_OK;
        if (LATE == 2) {
            if (wants_verdict) {
                verdict = a.verdicts[lane];
                for (uint32_t spins = 0; verdict == SS_VERDICT_EMPTY && spins < (1u << 24); spins++) {
                    ss_nanosleep(200);
                    verdict = a.verdicts[lane];
                }
            }
        } else {
            ss_pdl_wait();   // the extractor has finished: its status words are visible
            if (wants_verdict) verdict = *reinterpret_cast<const volatile uint32_t *>(a.status + frame_idx);
        }
        if (LATE == 2 && verdict == SS_VERDICT_EMPTY) a.status[frame_idx] = SS_ST_RETRY;   // (never seen: the bookkeeping thread always delivers)
        if (verdict != SS_ST_OK) {
            // refused (the reference throws before peer.write, or the host must extract): the stream state stays as it
            // was, the status word is the extractor's, and the wire bytes sealed ahead of the verdict are wiped
            uint32_t z[16];
#pragma unroll
            for (int i = 0; i < 16; i++) z[i] = 0;
            *reinterpret_cast<uint32_t *>(a.out + d.out_off) = 0u;
#pragma unroll 1
            for (uint32_t j = 0; j < npay; j++) {
                const int cbytes = (int)(mlen - 64u * j) < 64 ? (int)(mlen - 64u * j) : 64;
                ss_store64(dst + 64u * (size_t)j, cbytes, z);
            }
            ss_store_mac(dst + mlen, z);
            return;
        }
    }
    uint32_t *verp = reinterpret_cast<uint32_t *>(sp + 2) + 3;
    const uint32_t ver_word = (s2.w & 0xffu) | (a.new_ver << 8);
    if (!ok) { a.status[frame_idx] = SS_ST_BAD_MAC; if (chain) ss_st_release_u32(verp, ver_word); return; }
    if (in_run) { a.status[frame_idx] = SS_ST_OK | (tag << 8); return; }
    nn[1] ^= mac[0];
    nn[2] ^= mac[1];
    nn[0] += 1u;
    if ((tag & 2u) || nn[0] == 0u) {
        ss_rekey(k, nn);
        uint4 t0, t1;
        t0.x = k[0]; t0.y = k[1]; t0.z = k[2]; t0.w = k[3];
        t1.x = k[4]; t1.y = k[5]; t1.z = k[6]; t1.w = k[7];
        sp[0] = t0; sp[1] = t1;
    }
    uint4 t2;
    t2.x = nn[0]; t2.y = nn[1]; t2.z = nn[2]; t2.w = s2.w;
    sp[2] = t2;
    if (chain) ss_st_release_u32(verp, ver_word);
    a.status[frame_idx] = SS_ST_OK | (tag << 8);
#ifdef SS_PROBE_TIMELINE
    if (LATE && lane == 0) PROBE_MAX(3);
#endif
}
