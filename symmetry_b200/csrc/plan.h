// plan.h -- host-side packing of frame descriptors into warp tiles.
//
// The reference handles one frame per `peer.write` call
// (/root/reference/src/provider.ts:248); here a batch of frames from many
// connections is cut into warp-sized tiles: consecutive small frames whose
// ChaCha20 block counts (2 + ceil(mlen/64) each) sum to <= 32 share a warp,
// a frame with more than 30 payload blocks gets a bulk tile of its own.
#pragma once
#include <stdint.h>

#include "../../include/symrelay.h"

static inline uint32_t symrelay_nblocks(uint32_t mlen) { return 2u + ((mlen + 63u) >> 6); }

static inline int symrelay_desc_check_inline(const symrelay_desc *d) {
    const uint32_t mlen = d->len_flags & SYMRELAY_LEN_MASK;
    if (mlen > SYMRELAY_MAX_FRAME) return SYMRELAY_EINVAL;
    if (d->len_flags & SYMRELAY_F_OPEN) {
        if ((d->in_off & 15u) != 12u || (d->out_off & 15u) != 0u) return SYMRELAY_EINVAL;
    } else {
        if ((d->in_off & 15u) != 0u || (d->out_off & 15u) != 12u) return SYMRELAY_EINVAL;
    }
    return 0;
}

#ifndef SYMRELAY_THREAD_TILE_MIN_FRAMES
#define SYMRELAY_THREAD_TILE_MIN_FRAMES 8192u /* batches smaller than this are latency-bound: packed tiles only */
#endif
#define SYMRELAY_THREAD_TILE_MAX_BLOCKS 6u   /* frames of up to 4 payload blocks (256 B) */

// Small tiles first (packed, or thread-per-frame tiles: bit 17 of n_kind, 32 consecutive frames with
// the same block count, one per lane), bulk tiles after them.  Runs on the host's per-batch critical
// path, so pass 1 (validation + block counts) is a branch-free loop the compiler vectorises and the
// tile cutting only walks the block-count array.
#include <vector>

static inline int64_t symrelay_plan_inline(const symrelay_desc *desc, uint32_t n, symrelay_tile *tiles,
                                           uint32_t max_tiles, uint32_t *n_bulk_out) {
    static thread_local std::vector<uint32_t> ncv;
    if (ncv.size() < n) ncv.resize(n);
    uint32_t *nc = ncv.data();
    uint32_t bad = 0;
    for (uint32_t i = 0; i < n; i++) {
        const uint32_t lf = desc[i].len_flags;
        const uint32_t mlen = lf & SYMRELAY_LEN_MASK;
        const uint32_t op = (lf >> 26) & 1u;   // SYMRELAY_F_OPEN
        bad |= (uint32_t)(mlen > SYMRELAY_MAX_FRAME) | (uint32_t)((desc[i].in_off & 15u) != 12u * op) |
               (uint32_t)((desc[i].out_off & 15u) != 12u * (1u - op));
        nc[i] = 2u + ((mlen + 63u) >> 6);
    }
    if (bad) return SYMRELAY_EINVAL;
    uint64_t n_tiles = 0;
    uint32_t n_bulk = 0;
    uint32_t run_first = 0, run_n = 0, run_blocks = 0;
    const bool thread_tiles = n >= SYMRELAY_THREAD_TILE_MIN_FRAMES;
    for (uint32_t i = 0; i < n; i++) {
        const uint32_t c = nc[i];
        if (thread_tiles && c <= SYMRELAY_THREAD_TILE_MAX_BLOCKS && i + 32u <= n) {
            uint32_t j = 1;   // 32 consecutive frames with one block count -> one lane each, no lane idles
            while (j < 32u && nc[i + j] == c) j++;
            if (j == 32u) {
                if (run_n) {   // close the packed tile in progress first
                    if (tiles && n_tiles < max_tiles) { tiles[n_tiles].first = run_first; tiles[n_tiles].n_kind = run_n; }
                    n_tiles++;
                    run_n = 0; run_blocks = 0;
                }
                if (tiles && n_tiles < max_tiles) { tiles[n_tiles].first = i; tiles[n_tiles].n_kind = 32u | (1u << 17); }
                n_tiles++;
                i += 31u;
                continue;
            }
        }
        const bool bulk = c > 32u;
        if (bulk || run_blocks + c > 32u || run_n == 16u) {
            if (run_n) {
                if (tiles && n_tiles < max_tiles) { tiles[n_tiles].first = run_first; tiles[n_tiles].n_kind = run_n; }
                n_tiles++;
            }
            run_n = 0; run_blocks = 0;
        }
        if (bulk) { n_bulk++; continue; }
        if (run_n == 0) run_first = i;
        run_n++; run_blocks += c;
    }
    if (run_n) {
        if (tiles && n_tiles < max_tiles) { tiles[n_tiles].first = run_first; tiles[n_tiles].n_kind = run_n; }
        n_tiles++;
    }
    if (n_bulk) {
        for (uint32_t i = 0; i < n; i++) {
            if (nc[i] > 32u) {
                if (tiles && n_tiles < max_tiles) { tiles[n_tiles].first = i; tiles[n_tiles].n_kind = 1u | (1u << 16); }
                n_tiles++;
            }
        }
    }
    if (n_bulk_out) *n_bulk_out = n_bulk;
    return (int64_t)n_tiles;
}
