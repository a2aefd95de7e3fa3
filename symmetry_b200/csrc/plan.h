// plan.h -- host-side packing of frame descriptors into warp tiles.
//
// The reference handles one frame per `peer.write` call
// (/root/reference/src/provider.ts:248); here a batch of frames from many
// connections is cut into warp-sized tiles: consecutive small frames whose
// ChaCha20 block counts (2 + ceil(mlen/64) each) sum to <= 32 share a warp,
// a frame with more than 30 payload blocks gets a bulk tile of its own.
#pragma once
#include <stdint.h>
#include <stdlib.h>

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
#ifndef SYMRELAY_THREAD_TILE_MAX_BLOCKS
#ifndef SYMRELAY_THREAD_TILE_MAX_BLOCKS
#define SYMRELAY_THREAD_TILE_MAX_BLOCKS 6u   /* frames of up to 4 payload blocks (256 B) */
#endif
#endif

// Strided tiles (bit 16 of n_kind, bits 24..29 = lanes per frame): frames of 7..32 blocks (385 B .. 1,920 B) whose bucket
// is too small for thread tiles to be throughput-bound.  k lanes walk a frame's blocks with stride k, 32/k frames of
// one block count side by side; k is the smallest power of two that leaves a lane at most five blocks, so a tile
// takes about as long as five blocks whatever the frame size.  They run in the bulk kernel with the big frames
// (one frame per warp is the k = 32 case).
static inline uint32_t symrelay_seg_lanes(uint32_t nc) {
    uint32_t k = 2u;
    while ((nc + k) / k > 5u) k <<= 1;   // ceil((nc + 1) / k) virtual blocks per lane
    return k > 32u ? 32u : k;
}
#define SYMRELAY_THREAD_TILE_BIG_BUCKET 8192u   /* a size class (<= 32 blocks) with this many frames fills the GPU with thread tiles */
/* Wide tiles: a frame of 128 blocks and more (8 KiB payload; size class 7 and up) gets a whole 128-thread block -- when the
 * batch holds only a few such frames, so that it would wait for single warps walking them (BASELINE config 3).  Many
 * big frames (config 4) are throughput-bound: one warp each is the efficient shape, the wide form adds block barriers. */
#define SYMRELAY_WIDE_MIN_BLOCKS 127u
#define SYMRELAY_WIDE_MAX_FRAMES 2048u
/* ... and only while the whole batch is small enough to be waiting for its longest frame: with a block or more of work for
 * every thread the GPU can hold (148 SMs x 2,048) the batch is bound by total work, and the wide form's barriers and extra
 * launch cost more than the shorter tail saves (BASELINE config 3 on one GPU: 71 -> 74 us with it, measured). */
#define SYMRELAY_WIDE_MAX_BATCH_BLOCKS 262144u
/* ... and at most this many descriptors: the device-resident entry points get their tile list in device memory and cannot
 * look for wide tiles; with this cap they know from the descriptor count alone when the wide kernel has nothing to do. */
#define SYMRELAY_WIDE_MAX_BATCH_FRAMES 32768u
static inline bool symrelay_wide_ok(uint32_t n_big, uint64_t batch_blocks, uint32_t n_frames) {
    static const int mode = getenv("SYMRELAY_WIDE") ? atoi(getenv("SYMRELAY_WIDE")) : 1;   // 0: never (kill switch); 2: whatever the batch size (measurements)
    if (mode == 2) return n_big != 0u && n_big <= SYMRELAY_WIDE_MAX_FRAMES;
    return mode != 0 && n_big != 0u && n_big <= SYMRELAY_WIDE_MAX_FRAMES && batch_blocks < SYMRELAY_WIDE_MAX_BATCH_BLOCKS &&
           n_frames <= SYMRELAY_WIDE_MAX_BATCH_FRAMES;
}
#define SYMRELAY_THREAD_TILE_HUGE_BATCH 98304u  /* a uniform batch of BIG frames with this many of them (3,072 warps: ~20 per SM) does too */

// Small tiles first (packed, or thread-per-frame tiles: bit 17 of n_kind, 32 consecutive frames with
// the same block count, one per lane), bulk tiles after them.  Runs on the host's per-batch critical
// path, so pass 1 (validation + block counts) is a branch-free loop the compiler vectorises and the
// tile cutting only walks the block-count array.
#include <vector>

// Optional bounds checks folded into the planner's first pass (the host batch entry points use them so
// that the descriptors are walked once, not three times).
struct symrelay_plan_check {
    uint64_t in_bytes, out_bytes;
    uint32_t max_streams;
    // results
    uint32_t out_of_bounds;   // some frame leaves its arena or names a slot >= max_streams
    uint32_t any_flags;       // OR of all len_flags
    uint32_t out_overlap;     // output rows (16 B) not strictly increasing in descriptor order
    // set by symrelay_plan_inline when every frame has one block count and the tiles follow in closed form:
    // `uniform_thread` thread-per-frame tiles of 32 frames, then packed tiles of `uniform_per` frames (0: not uniform)
    uint32_t uniform_per, uniform_thread;
};

// Pass 1 of the planner.  No loop-carried dependency and all 32-bit (offsets are uint32 and arenas are
// capped just below 4 GiB, so `off > size - need` cannot wrap).  out: bad, oob, any_flags, nc_min, nc_max,
// overlap (output rows of 16 bytes not strictly increasing in descriptor order).
static inline void symrelay_plan_scan_scalar(const symrelay_desc *desc, uint32_t n, uint32_t *nc, uint32_t *row_lo,
                                             uint32_t *row_hi, uint32_t in_bytes, uint32_t out_bytes, uint32_t max_streams,
                                             uint32_t out[6]) {
    uint32_t bad = 0, oob = 0, any_flags = 0, nc_min = 0xffffffffu, nc_max = 0;
    for (uint32_t i = 0; i < n; i++) {
        const uint32_t lf = desc[i].len_flags, in_off = desc[i].in_off, out_off = desc[i].out_off;
        const uint32_t mlen = lf & SYMRELAY_LEN_MASK;
        const uint32_t op = (lf >> 26) & 1u;   // SYMRELAY_F_OPEN
        const uint32_t c = 2u + ((mlen + 63u) >> 6);
        const uint32_t need_in = mlen + 20u * op, need_out = mlen + 20u - 20u * op;
        bad |= (uint32_t)(mlen > SYMRELAY_MAX_FRAME) | (uint32_t)((in_off & 15u) != 12u * op) |
               (uint32_t)((out_off & 15u) != 12u - 12u * op);
        oob |= (uint32_t)(need_in > in_bytes) | (uint32_t)(in_off > in_bytes - need_in) |
               (uint32_t)(need_out > out_bytes) | (uint32_t)(out_off > out_bytes - need_out) |
               (uint32_t)(desc[i].slot >= max_streams);
        any_flags |= lf;
        nc[i] = c;
        nc_min = c < nc_min ? c : nc_min;
        nc_max = c > nc_max ? c : nc_max;
        row_lo[i] = out_off >> 4;
        row_hi[i] = (out_off >> 4) + (((out_off & 15u) + need_out + 15u) >> 4);
    }
    uint32_t overlap = 0;   // plain arrays: vectorises without gathers
    for (uint32_t i = 1; i < n; i++) overlap |= (uint32_t)(row_lo[i] < row_hi[i - 1]);
    out[0] = bad; out[1] = oob; out[2] = any_flags; out[3] = nc_min; out[4] = nc_max; out[5] = overlap;
}

// The library links plan_scan_avx2.cpp (SYMRELAY_HAVE_AVX2_SCAN) and uses it where the CPU has AVX2.
#ifdef SYMRELAY_HAVE_AVX2_SCAN
extern "C" void symrelay_plan_scan_avx2(const symrelay_desc *desc, uint32_t n, uint32_t *nc, uint32_t *row_lo, uint32_t *row_hi,
                                        uint32_t in_bytes, uint32_t out_bytes, uint32_t max_streams, uint32_t out[6]);
extern "C" int symrelay_plan_scan_has_avx2(void);
#endif
static inline void symrelay_plan_scan(const symrelay_desc *desc, uint32_t n, uint32_t *nc, uint32_t *row_lo, uint32_t *row_hi,
                                      uint32_t in_bytes, uint32_t out_bytes, uint32_t max_streams, uint32_t out[6]) {
#ifdef SYMRELAY_HAVE_AVX2_SCAN
    static const int have = symrelay_plan_scan_has_avx2() && !getenv("SYMRELAY_NO_AVX2");   // the variable is a kill switch
    if (have) { symrelay_plan_scan_avx2(desc, n, nc, row_lo, row_hi, in_bytes, out_bytes, max_streams, out); return; }
#endif
    symrelay_plan_scan_scalar(desc, n, nc, row_lo, row_hi, in_bytes, out_bytes, max_streams, out);
}

static inline int64_t symrelay_plan_inline(const symrelay_desc *desc, uint32_t n, symrelay_tile *tiles,
                                           uint32_t max_tiles, uint32_t *n_bulk_out,
                                           symrelay_plan_check *chk = nullptr) {
    static thread_local std::vector<uint32_t> ncv;
    if (ncv.size() < 3u * (size_t)n) ncv.resize(3u * (size_t)n);
    uint32_t *nc = ncv.data();
    uint32_t *row_lo = nc + n, *row_hi = nc + 2u * (size_t)n;   // output rows in 16-byte units
    // pass 1: validation, block counts, output rows
    const uint64_t cap = 0xffffffffull;
    uint32_t scan[6];
    symrelay_plan_scan(desc, n, nc, row_lo, row_hi, (uint32_t)(chk && chk->in_bytes < cap ? chk->in_bytes : cap),
                       (uint32_t)(chk && chk->out_bytes < cap ? chk->out_bytes : cap), chk ? chk->max_streams : 0xffffffffu, scan);
    const uint32_t bad = scan[0], oob = scan[1], any_flags = scan[2], nc_min = scan[3], nc_max = scan[4];
    if (chk) { chk->out_of_bounds = oob; chk->any_flags = any_flags; chk->out_overlap = scan[5]; chk->uniform_per = chk->uniform_thread = 0; }
    if (bad) return SYMRELAY_EINVAL;
    uint64_t n_tiles = 0;
    uint32_t n_bulk = 0;
    if (n && nc_min == nc_max && (nc_max <= 32u || n >= SYMRELAY_THREAD_TILE_HUGE_BATCH)) {
        // every frame has the same block count (token streaming, BASELINE config 2): the tiles follow in
        // closed form -- no per-descriptor walk
        // (frames of more than 32 blocks get here only in a huge uniform batch: one lane per frame, throughput-bound)
        const bool thread_ok = (n >= SYMRELAY_THREAD_TILE_MIN_FRAMES && nc_max <= SYMRELAY_THREAD_TILE_MAX_BLOCKS) ||
                               n >= SYMRELAY_THREAD_TILE_BIG_BUCKET;
        if (!thread_ok && nc_max > SYMRELAY_THREAD_TILE_MAX_BLOCKS) {
            // mid-size frames, not enough of them for thread tiles: strided tiles, all in the bulk class
            const uint32_t k = symrelay_seg_lanes(nc_max), per_s = 32u / k;
            for (uint32_t i = 0; i < n; i += per_s) {
                const uint32_t cnt = n - i < per_s ? n - i : per_s;
                if (tiles && n_tiles < max_tiles) { tiles[n_tiles].first = i; tiles[n_tiles].n_kind = cnt | (1u << 16) | (k << 24); }
                n_tiles++;
            }
            if (n_bulk_out) *n_bulk_out = (uint32_t)n_tiles;
            return (int64_t)n_tiles;
        }
        uint32_t per = 32u / nc_max;
        if (per > 16u) per = 16u;
        uint32_t i = 0;
        if (thread_ok)
            for (; i + 32u <= n; i += 32u) {
                if (tiles && n_tiles < max_tiles) { tiles[n_tiles].first = i; tiles[n_tiles].n_kind = 32u | (1u << 17); }
                n_tiles++;
            }
        if (nc_max > 32u) {   // huge batch of big frames: the last < 32 of them as ordinary bulk tiles
            const uint32_t rest = n - i;
            for (; i < n; i++) {
                if (tiles && n_tiles < max_tiles) { tiles[n_tiles].first = i; tiles[n_tiles].n_kind = 1u | (1u << 16); }
                n_tiles++;
            }
            if (n_bulk_out) *n_bulk_out = rest;
            return (int64_t)n_tiles;
        }
        for (; i < n; i += per) {
            const uint32_t k = n - i < per ? n - i : per;
            if (tiles && n_tiles < max_tiles) { tiles[n_tiles].first = i; tiles[n_tiles].n_kind = k; }
            n_tiles++;
        }
        if (n_bulk_out) *n_bulk_out = 0;
        if (chk) { chk->uniform_per = per; chk->uniform_thread = thread_ok ? n / 32u : 0u; }
        return (int64_t)n_tiles;
    }
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
        uint32_t n_big = 0;
        uint64_t blocks = 0;
        for (uint32_t i = 0; i < n; i++) { n_big += nc[i] > SYMRELAY_WIDE_MIN_BLOCKS ? 1u : 0u; blocks += nc[i]; }
        const bool wide = symrelay_wide_ok(n_big, blocks, n);
        for (uint32_t i = 0; i < n; i++) {
            if (nc[i] > 32u) {
                if (tiles && n_tiles < max_tiles) {
                    tiles[n_tiles].first = i;
                    tiles[n_tiles].n_kind = 1u | (1u << 16) | ((wide && nc[i] > SYMRELAY_WIDE_MIN_BLOCKS) ? (1u << 18) : 0u);
                }
                n_tiles++;
            }
        }
    }
    if (n_bulk_out) *n_bulk_out = n_bulk;
    return (int64_t)n_tiles;
}

// A four-way fork-join for the planner of very large batches (BASELINE config 3 is 131,072 descriptors per batch and its
// host path was bound by ~1 ms of single-threaded planning per batch against ~60 us of kernels).  Three workers are
// started with the first large batch and sleep on a condition variable between batches; the caller runs share 0 itself.
// If the pool is busy (another GPU's batch is being planned), cannot start, or the process has forked since, the caller
// simply runs all four shares one after the other: the result is the same either way.
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <string.h>
#include <unistd.h>
#ifndef SYMRELAY_PLAN_PAR_MIN
#define SYMRELAY_PLAN_PAR_MIN 32768u   /* descriptors; below this the wake-ups cost more than they save */
#endif
class SymrelayPlanPool {
public:
    // Never destroyed: the workers sleep until the process ends (a destructor would have to join them from whatever
    // thread runs exit handlers -- and in a forked child, where they do not exist, there is nothing to join).
    static SymrelayPlanPool &get() { static SymrelayPlanPool *p = new SymrelayPlanPool; return *p; }
    template <class F> void run(F &&f) {
        std::unique_lock<std::mutex> busy(busy_, std::defer_lock);
        if (!usable() || !busy.try_lock()) { for (int k = 0; k < 4; k++) f(k); return; }
        job_ = [&f](int k) { f(k); };
        { std::lock_guard<std::mutex> g(m_); pending_ = 3; gen_++; }
        cv_go_.notify_all();
        f(0);
        std::unique_lock<std::mutex> g(m_);
        cv_done_.wait(g, [this] { return pending_ == 0; });
    }
private:
    SymrelayPlanPool() {
        if (getenv("SYMRELAY_PLAN_THREADS") && atoi(getenv("SYMRELAY_PLAN_THREADS")) == 0) return;   // kill switch
        pid_ = getpid();
        try {
            for (int k = 0; k < 3; k++) th_[k] = std::thread([this, k] { worker(k + 1); });
            ok_ = true;
        } catch (...) {   // no threads to be had: whatever did start is told to leave, and the caller plans alone
            { std::lock_guard<std::mutex> g(m_); quit_ = true; gen_++; }
            cv_go_.notify_all();
            for (auto &t : th_) if (t.joinable()) t.join();
        }
    }
    SymrelayPlanPool(const SymrelayPlanPool &) = delete;
    bool usable() const { return ok_ && pid_ == getpid(); }
    void worker(int k) {
        unsigned long long seen = 0;
        for (;;) {
            std::unique_lock<std::mutex> g(m_);
            cv_go_.wait(g, [&] { return gen_ != seen; });
            seen = gen_;
            if (quit_) return;
            g.unlock();
            job_(k);
            g.lock();
            if (--pending_ == 0) cv_done_.notify_one();
        }
    }
    std::thread th_[3];
    std::mutex m_, busy_;
    std::condition_variable cv_go_, cv_done_;
    std::function<void(int)> job_;
    unsigned long long gen_ = 0;
    int pending_ = 0;
    bool quit_ = false, ok_ = false;
    pid_t pid_ = 0;
};

// Bucketed planning for large host batches.  Frames that arrive in any order pack badly (BASELINE config 3:
// 19,441 tiles in arrival order against 9,068 when grouped by size, no thread tiles) and the greedy cut above
// mispredicts on every tile boundary (1.5 ms for 131,072 descriptors).  A batch holds at most one frame per
// (slot, direction), so its descriptors may be processed in any order: this form sorts them by block count
// while copying them to `sorted` (a counting sort; the host path copies them into its control block anyway),
// tiles every bucket in closed form and returns the permutation so that the caller can hand statuses back in
// the caller's order (status_of_caller[perm[k]] = status_of_kernel[k]).  Batches with SYMRELAY_F_RUN frames
// (whose order is the chain) are refused with -2: plan them with symrelay_plan_inline.
// tiles may be NULL to size; sorted/perm are written only when tiles is non-NULL.
static inline int64_t symrelay_plan_bucketed(const symrelay_desc *desc, uint32_t n, symrelay_tile *tiles, uint32_t max_tiles,
                                             uint32_t *n_bulk_out, symrelay_plan_check *chk, symrelay_desc *sorted,
                                             uint32_t *perm) {
    static thread_local std::vector<uint32_t> ncv;
    if (ncv.size() < 3u * (size_t)n) ncv.resize(3u * (size_t)n);
    uint32_t *nc = ncv.data(), *row_lo = nc + n, *row_hi = nc + 2u * (size_t)n;
    const uint64_t cap = 0xffffffffull;
    const uint32_t in_cap = (uint32_t)(chk && chk->in_bytes < cap ? chk->in_bytes : cap),
                   out_cap = (uint32_t)(chk && chk->out_bytes < cap ? chk->out_bytes : cap),
                   max_streams = chk ? chk->max_streams : 0xffffffffu;
    // The batch is cut into four consecutive quarters; every pass below works quarter by quarter -- on four threads when
    // the batch is large (SymrelayPlanPool), one after the other otherwise -- and a quarter's frames of a bucket are
    // placed behind the earlier quarters' frames of that bucket, so the sorted order is the one a single walk gives.
    constexpr uint32_t PQ = 4;
    const uint32_t quarter = (((n + PQ - 1u) / PQ) + 7u) & ~7u;
    const bool parallel = n >= SYMRELAY_PLAN_PAR_MIN;
    auto run4 = [parallel](auto &&f) {   // (the pool and its threads come into being with the first large batch)
        if (parallel) SymrelayPlanPool::get().run(f);
        else for (int k = 0; k < 4; k++) f(k);
    };
    // bucket 0..30: 2..32 blocks; bucket 31: bulk.  Inside the bulk bucket the frames are ordered by size class
    // (power-of-two ranges of the block count), LARGEST FIRST: a 16 KiB frame keeps its warp nine times as long as a
    // 2 KiB one, so the long ones must start first or they finish last.
    uint32_t qscan[PQ][6], qcount[PQ][32], qclass[PQ][20];
    uint64_t qblocks[PQ];
    run4([&](int k) {
        const uint32_t lo = (uint32_t)k * quarter < n ? (uint32_t)k * quarter : n, hi = lo + quarter < n ? lo + quarter : n;
        memset(qcount[k], 0, sizeof(qcount[k]));
        memset(qclass[k], 0, sizeof(qclass[k]));
        symrelay_plan_scan(desc + lo, hi - lo, nc + lo, row_lo + lo, row_hi + lo, in_cap, out_cap, max_streams, qscan[k]);
        uint64_t blocks = 0;
        for (uint32_t i = lo; i < hi; i++) {
            const uint32_t c = nc[i];
            blocks += c;
            qcount[k][(c > 32u ? 33u : c) - 2u]++;
            if (c > 32u) qclass[k][31 - __builtin_clz(c)]++;   // c in [33, 2^18 + 2]: classes 5..18
        }
        qblocks[k] = blocks;
    });
    uint32_t scan[6] = {0, 0, 0, 0xffffffffu, 0, 0};
    for (uint32_t k = 0; k < PQ; k++) {
        const uint32_t lo = k * quarter;
        if (lo >= n) break;
        scan[0] |= qscan[k][0]; scan[1] |= qscan[k][1]; scan[2] |= qscan[k][2]; scan[5] |= qscan[k][5];
        scan[3] = qscan[k][3] < scan[3] ? qscan[k][3] : scan[3];
        scan[4] = qscan[k][4] > scan[4] ? qscan[k][4] : scan[4];
        if (k) scan[5] |= (uint32_t)(row_lo[lo] < row_hi[lo - 1u]);   // the output rows across the cut
    }
    if (chk) { chk->out_of_bounds = scan[1]; chk->any_flags = scan[2]; chk->out_overlap = scan[5]; }
    if (scan[0]) return SYMRELAY_EINVAL;
    if (scan[2] & SYMRELAY_F_RUN) return -2;
    uint32_t count[32], bulk_class[20];
    uint64_t batch_blocks = 0;
    for (uint32_t b = 0; b < 32u; b++) count[b] = qcount[0][b] + qcount[1][b] + qcount[2][b] + qcount[3][b];
    for (uint32_t b = 0; b < 20u; b++) bulk_class[b] = qclass[0][b] + qclass[1][b] + qclass[2][b] + qclass[3][b];
    for (uint32_t k = 0; k < PQ; k++) batch_blocks += qblocks[k];
    const bool thread_ok = n >= SYMRELAY_THREAD_TILE_MIN_FRAMES;
    uint32_t base[32], at = 0;
    uint64_t n_tiles = 0;
    uint32_t n_strided = 0;
    uint32_t sbase[32], scount[32];   // per bucket: the part that becomes strided tiles (second pass, bulk class)
    for (uint32_t b = 0; b < 32u; b++) { base[b] = at; at += count[b]; sbase[b] = 0; scount[b] = 0; }
    // Tiles are emitted with the LONGEST first (six-block frames before two-block ones): the tile list's order is the
    // order in which warps are started, a batch of this size needs about two rounds of resident warps, and whatever
    // starts in the second round decides when the step ends -- with the short tiles there it ends sooner (BASELINE
    // config 3, measured with the timeline probe: the six-block thread tiles started at 30 us and ended at 61).
    for (uint32_t b = 31u; b-- > 0u;) {   // (bucket 31 = bulk tiles: appended after every small tile, below)
        uint32_t m = count[b], s = base[b];
        const uint32_t c = b + 2u;
        const bool threads = thread_ok && (c <= SYMRELAY_THREAD_TILE_MAX_BLOCKS || m >= SYMRELAY_THREAD_TILE_BIG_BUCKET);
        if (!threads && c > SYMRELAY_THREAD_TILE_MAX_BLOCKS) { sbase[b] = s; scount[b] = m; continue; }
        if (threads)
            for (; m >= 32u; m -= 32u, s += 32u) {
                if (tiles && n_tiles < max_tiles) { tiles[n_tiles].first = s; tiles[n_tiles].n_kind = 32u | (1u << 17); }
                n_tiles++;
            }
        if (c > SYMRELAY_THREAD_TILE_MAX_BLOCKS) { sbase[b] = s; scount[b] = m; continue; }   // the tail of a big bucket (< 32 frames)
        uint32_t per = 32u / c;
        if (per > 16u) per = 16u;
        for (; m; ) {
            const uint32_t k = m < per ? m : per;
            if (tiles && n_tiles < max_tiles) { tiles[n_tiles].first = s; tiles[n_tiles].n_kind = k; }
            n_tiles++;
            m -= k; s += k;
        }
    }
    for (uint32_t b = 31u; b-- > 0u;) {
        const uint32_t c = b + 2u, k = symrelay_seg_lanes(c), per_s = 32u / k;
        for (uint32_t m = scount[b], s = sbase[b]; m; ) {
            const uint32_t cnt = m < per_s ? m : per_s;
            if (tiles && n_tiles < max_tiles) { tiles[n_tiles].first = s; tiles[n_tiles].n_kind = cnt | (1u << 16) | (k << 24); }
            n_tiles++; n_strided++;
            m -= cnt; s += cnt;
        }
    }
    {   // big frames: largest size class first (see the sort below), so the wide ones are the first n_big of them
        uint32_t n_big = 0;
        for (uint32_t k = 7u; k < 20u; k++) n_big += bulk_class[k];   // classes 7..: more than SYMRELAY_WIDE_MIN_BLOCKS blocks
        const uint32_t n_wide = symrelay_wide_ok(n_big, batch_blocks, n) ? n_big : 0u;
        for (uint32_t k = 0; k < count[31]; k++) {
            if (tiles && n_tiles < max_tiles) {
                tiles[n_tiles].first = base[31] + k;
                tiles[n_tiles].n_kind = 1u | (1u << 16) | (k < n_wide ? (1u << 18) : 0u);
            }
            n_tiles++;
        }
    }
    if (n_bulk_out) *n_bulk_out = count[31] + n_strided;
    if (tiles && sorted && perm) {
        uint32_t next[PQ][32], bnext[PQ][20];
        for (uint32_t b = 0; b < 32u; b++)
            for (uint32_t k = 0, at2 = base[b]; k < PQ; k++) { next[k][b] = at2; at2 += qcount[k][b]; }
        for (uint32_t c = 20u, at2 = base[31]; c-- > 0u;)   // largest class first
            for (uint32_t k = 0; k < PQ; k++) { bnext[k][c] = at2; at2 += qclass[k][c]; }
        run4([&](int k) {
            const uint32_t lo = (uint32_t)k * quarter < n ? (uint32_t)k * quarter : n, hi = lo + quarter < n ? lo + quarter : n;
            for (uint32_t i = lo; i < hi; i++) {
                const uint32_t c = nc[i];
                const uint32_t pos = c > 32u ? bnext[k][31 - __builtin_clz(c)]++ : next[k][c - 2u]++;
                sorted[pos] = desc[i];
                perm[pos] = i;
            }
        });
    }
    return (int64_t)n_tiles;
}
