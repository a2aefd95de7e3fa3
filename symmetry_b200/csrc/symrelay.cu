// symrelay.cu -- kernels + C ABI of the batched secretstream relay (sm_100a).
//
// Replaces, for all connections at once, the per-frame work that the reference
// does on Node's main thread for every `peer.write(chunk)`
// (/root/reference/src/provider.ts:248) and every inbound frame (:174):
// secret-stream framing + libsodium secretstream push/pull.  See
// include/symrelay.h for the boundary and DESIGN.md for layouts and rooflines.
// There is deliberately no CPU path in this file: every entry point that
// computes needs a CUDA device.
#include <cuda_runtime.h>
#include <emmintrin.h>

#include <atomic>
#include <condition_variable>
#include <thread>
#include <sys/eventfd.h>
#include <unistd.h>
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <deque>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/symrelay.h"
#include "plan.h"
#ifdef SS_PROBE_TIMELINE   // measurement builds only (tools/probes/c5_timeline.py): when do the two kernels of an extract batch run?
__device__ unsigned long long g_probe[32];
__device__ __forceinline__ unsigned long long probe_now() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define PROBE_SET(i) (g_probe[i] = probe_now())
#define PROBE_MAX(i) atomicMax(&g_probe[i], probe_now())
extern "C" __attribute__((visibility("default"))) int symrelay_probe_timeline(unsigned long long *out, int reset) {
    if (out && cudaMemcpyFromSymbol(out, g_probe, sizeof(g_probe)) != cudaSuccess) return -1;
    if (reset) { unsigned long long z[32] = {0}; if (cudaMemcpyToSymbol(g_probe, z, sizeof(z)) != cudaSuccess) return -1; }
    return 0;
}
#else
#define PROBE_SET(i) ((void)0)
#define PROBE_MAX(i) ((void)0)
#endif

#include "ss_core.cuh"
#include "ss_extract.cuh"
#include "ss_runs.cuh"
#include "ss_blake2b.cuh"

static_assert(sizeof(symrelay_desc) == sizeof(SsDesc) && sizeof(symrelay_tile) == sizeof(SsTile), "ABI layout");

// ------------------------------------------------------------------ kernels --
// One warp per tile.  Packed tiles: several small frames per warp.
#ifndef SS_PACKED_MIN_BLOCKS
#define SS_PACKED_MIN_BLOCKS 1
#endif
__global__ void __launch_bounds__(128, SS_PACKED_MIN_BLOCKS) ss_packed_kernel(SsArgs a) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (warp >= a.n_tiles) return;
    if (a.tiles[warp].n_kind & (1u << 17)) ss_process_thread_tile<false, 1>(a, warp);   // warp-uniform: whole tile is one kind
    else ss_process_tile<false>(a, warp);
}

// Large batches split the two small-tile kinds over two kernels (each launched over the whole small-tile range and taking
// only its own kind): the thread-per-frame path needs far fewer registers than the packed one, and BASELINE config 3 is
// 3,750 thread tiles that want to be resident all at once.
#ifndef SS_THREAD_MIN_BLOCKS
#define SS_THREAD_MIN_BLOCKS 5   // 90 registers, no spills, 20 warps per SM.  Measured on config 3 / the 512 B and 4 KiB points of the size
                                 // sweep: 4 -> 59.3 us; 5 -> 57.2 us, 88.9 %, 87.4 %; 6 (80 registers) -> 57.9 us, 87.6 %, 85.7 %; 7 (72, spills) -> 60.9 us
#endif
__global__ void __launch_bounds__(128, SS_THREAD_MIN_BLOCKS) ss_thread_kernel(SsArgs a) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (warp >= a.n_tiles) return;
    const SsTile tile = a.tiles[warp];
    if (!(tile.n_kind & (1u << 17))) return;
    if ((threadIdx.x & 31u) == 0u) { if (warp == 0) PROBE_SET(22); PROBE_MAX(28); }   // (tools/probes/c3_timeline.py)
    ss_process_thread_tile(a, warp, &tile);
    if ((threadIdx.x & 31u) == 0u) PROBE_MAX(23);
}
// ... and the form for batches with an extractor pass (SS_ARG_VERDICT_LATE, ss_core.cuh)
__global__ void __launch_bounds__(128, SS_THREAD_MIN_BLOCKS) ss_thread_late_kernel(SsArgs a) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (warp >= a.n_tiles) return;
    const SsTile tile = a.tiles[warp];
    if (!(tile.n_kind & (1u << 17))) return;
    ss_process_thread_tile<false, 1>(a, warp, &tile);
}
__global__ void __launch_bounds__(128, SS_PACKED_MIN_BLOCKS) ss_packed_only_kernel(SsArgs a) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (warp >= a.n_tiles) return;
    const SsTile tile = a.tiles[warp];
    if (tile.n_kind & (1u << 17)) return;
    if ((threadIdx.x & 31u) == 0u) PROBE_MAX(20);
    ss_process_tile<false>(a, warp, &tile);
    if ((threadIdx.x & 31u) == 0u) PROBE_MAX(21);
}

// The same kernel for small host batches: the tile list rides in the launch parameters (constant bank), so a
// warp does not start with a read of host memory that only tells it where its descriptors are.  Zero-copy
// batches begin with three dependent reads over PCIe (tile -> descriptors -> payload); this removes the first.
#define SS_INLINE_TILES 448u   // 3.5 KB: with SsArgs under the 4 KB parameter limit
struct SsTileBlock {
    SsTile t[SS_INLINE_TILES];
};
__global__ void __launch_bounds__(128, SS_PACKED_MIN_BLOCKS) ss_packed_inline_kernel(SsArgs a, SsTileBlock tb) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (warp >= a.n_tiles) return;
    const SsTile tile = tb.t[warp];
    if (tile.n_kind & (1u << 17)) ss_process_thread_tile<false, 1>(a, warp, &tile);
    else ss_process_tile<false>(a, warp, &tile);
}

// Bulk tiles: one frame per warp, lanes stride over its 64-byte blocks.
// a.tiles already points at the first bulk tile.
#ifndef SS_BULK_MIN_BLOCKS
#define SS_BULK_MIN_BLOCKS 7   // <= 72 registers, 28 warps per SM: measured best on 64 KiB frames (5: 60.7 %, 6: 64.4 %, 7: 65.4 %, 8: 65.4 % of the ALU peak)
#endif
__global__ void __launch_bounds__(128, SS_BULK_MIN_BLOCKS) ss_bulk_kernel(SsArgs a) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (warp >= a.n_tiles) return;
    const SsTile tile = a.tiles[warp];
    if (((tile.n_kind >> 24) & 63u) || (tile.n_kind & (1u << 18))) return;   // a strided or a wide tile: another kernel's
    if ((threadIdx.x & 31u) == 0u) PROBE_MAX(29);
    ss_process_tile<true>(a, warp, &tile);
    if ((threadIdx.x & 31u) == 0u) PROBE_MAX(27);
}

// Wide tiles (plan.h): one frame of more than 8 KiB per 128-thread BLOCK, for batches that hold only a few of them and
// would otherwise wait for single warps walking those frames.  a.tiles points at the first bulk-class tile of the range
// the caller picked; blocks that find another kind's tile leave.
#ifndef SS_WIDE_MIN_BLOCKS
#define SS_WIDE_MIN_BLOCKS 6
#endif
__global__ void __launch_bounds__(128, SS_WIDE_MIN_BLOCKS) ss_wide_kernel(SsArgs a) {
    const uint32_t t = blockIdx.x;   // (block-uniform: every thread of a block takes the same branch)
    if (t >= a.n_tiles) return;
    const SsTile tile = a.tiles[t];
    if (!(tile.n_kind & (1u << 18))) return;
    ss_process_tile<true, false, false, true>(a, t, &tile);
}

// Strided tiles (plan.h): 32/k mid-size frames of one block count per warp, k lanes each.  Launched over the same
// bulk-class tile range as ss_bulk_kernel; each of the two kernels takes its own kind and leaves the other's alone
// (the entry points get one count for the class, and the tile list may live in device memory).
__global__ void __launch_bounds__(128, 4) ss_strided_kernel(SsArgs a) {
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (warp >= a.n_tiles) return;
    const SsTile tile = a.tiles[warp];
    if (!((tile.n_kind >> 24) & 63u)) return;
    if ((threadIdx.x & 31u) == 0u) PROBE_MAX(30);
    ss_process_tile<true, false, true>(a, warp, &tile);
    if ((threadIdx.x & 31u) == 0u) PROBE_MAX(25);
}

// Posts a deferred batch's results to the caller's host memory when no later batch carries them.
__global__ void __launch_bounds__(128) ss_write_back_kernel(SsArgs a) {
    ss_write_back(a, (blockIdx.x * blockDim.x + threadIdx.x) >> 5, (gridDim.x * blockDim.x) >> 5, threadIdx.x & 31u);
}

// Runs of queued frames of one stream (see ss_runs.cuh): nonce pre-pass and state commit.
// run_nonce is SCRATCH memory: 16 bytes per descriptor of the batch, valid only between these passes.
__global__ void __launch_bounds__(128) ss_run_nonce_kernel(const SsDesc *desc, uint32_t n, const uint8_t *in,
                                                           const uint4 *states, uint4 *run_nonce) {
    ss_run_nonce(desc, n, in, states, run_nonce, blockIdx.x * blockDim.x + threadIdx.x);
}
__global__ void __launch_bounds__(128) ss_run_commit_kernel(const SsDesc *desc, uint32_t n, const uint8_t *in,
                                                            uint4 *states, uint32_t *status) {
    ss_run_commit(desc, n, in, states, status, blockIdx.x * blockDim.x + threadIdx.x);
}

// Delta extraction for the frames that carry SYMRELAY_F_EXTRACT.  Runs before (and, see
// SS_ARG_VERDICT_LATE, beside) the seal kernel of the same batch: it leaves its verdict in status[i] (frames whose
// verdict is not OK are not sealed) and the delta bytes at delta_arena + in_off (the plaintext layout, so no
// allocation or scratch is needed: every frame slot holds >= max(16, mlen) bytes).
// A block takes 64 chunks and gives each chunk to THREE threads in three different warps -- grammar, event recording,
// bookkeeping (ss_extract.cuh: "the two passes as three roles") -- that hand their results on through small queues in
// shared memory, so the stages of one chunk run side by side instead of one after the other.  Warps 2 and 3 run the
// grammar (the long chain): with one block per SM each of them has a scheduler to itself; the lighter roles share
// the other two schedulers.
#define SS_EXTRACT_CHUNKS 64u    // chunks per block
#define SS_EXTRACT_THREADS 192u  // two warps per role
#define SS_EXTRACT_STAGE 384u    // chunks up to this size are staged in shared memory and take the three-role extractor
#define SS_EXTRACT_EVENTS 95u    // events per chunk; a chunk with more falls back to the single pass
// per chunk, in words: the staged chunk (+4 bytes the prefix search may read) | container stack | trailer | flag queue |
// event queue.  307 words: the lanes of a warp hit distinct banks
#define SS_EXTRACT_TEXT_W ((SS_EXTRACT_STAGE + 4u) / 4u)
#define SS_EXTRACT_STRIDE_W (SS_EXTRACT_TEXT_W + SX_FAST_STACK_WORDS + 1u + SX_FAST_FLAG_WORDS(SS_EXTRACT_STAGE) + SS_EXTRACT_EVENTS + 2u)
#define SS_EXTRACT_SMEM (SX_FAST_ENTRIES * 8u + 256u + SS_EXTRACT_CHUNKS * SS_EXTRACT_STRIDE_W * 4u)
static_assert(SS_EXTRACT_STRIDE_W % 2u == 1u, "an odd stride keeps the lanes on distinct banks");

// The two-pass tables (grammar by byte class, then the 256 class codes) live in a cudaMalloc'd block per
// GPU, not in file-scope __device__ arrays: 15 KB more of module-global data moved later allocations and
// cost the C2 step 8 % (4.92 -> 5.31 us, same kernel code, measured with and without the arrays).
#define SX_FAST_BLOCK_BYTES (SX_FAST_ENTRIES * 8u + 256u)                 /* what the kernel stages in shared memory */
#define SX_TABLES_BYTES (SX_FAST_BLOCK_BYTES + SX_TABLE_ENTRIES * 2u)   /* + the single-pass table (fallback, long chunks) */
// role: 0 grammar, 1 recording, 2 bookkeeping, 3 none (a seal warp of the fused kernel: it only helps to set up).
// verdicts: fused kernel only -- the block's 64 verdict words in shared memory (written for every chunk; the status word
// of an accepted chunk is then left to the seal thread, which writes the tag into it).
__device__ __forceinline__ void ss_extract_block(const SsDesc *desc, uint32_t n, const uint8_t *in, const uint4 *states,
                                                 uint32_t *status, symrelay_delta *delta, uint8_t *delta_arena,
                                                 const uint4 *fast_block, uint32_t role, uint32_t group,
                                                 volatile uint32_t *verdicts) {
    // The grammar consumes its chunk one byte at a time, each step depending on the last: read from global memory
    // that is ~600 cycles per byte.  So the chunk is first pulled into shared memory with 16-byte loads that are all
    // in flight together, and the byte loop then runs against shared memory (chunk, class table and grammar table).
    extern __shared__ __align__(16) uint8_t sx_smem[];
    uint64_t *fast = reinterpret_cast<uint64_t *>(sx_smem);
    uint8_t *cls = sx_smem + SX_FAST_ENTRIES * 8u;
    uint32_t *sx_stage = reinterpret_cast<uint32_t *>(cls + 256u);
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t c = group * 32u + lane;                              // this thread's chunk within the block
    const uint32_t i = blockIdx.x * SS_EXTRACT_CHUNKS + c;
    // Three things come from global memory and two of them are a dependent chain (descriptor -> chunk and
    // provider kind): start the chain first, copy the tables while it is in flight.
    SsDesc d;
    d.slot = 0; d.in_off = 0; d.out_off = 0; d.len_flags = 0;
    if (i < n) d = desc[i];
    const bool mine_to_do = (d.len_flags & SS_F_EXTRACT) && !(d.len_flags & SS_F_OPEN);
    const uint32_t mlen = d.len_flags & 0xffffffu;
    const uint8_t *text = in + d.in_off;
    uint32_t *base = sx_stage + (size_t)c * SS_EXTRACT_STRIDE_W;
    uint8_t *mine = reinterpret_cast<uint8_t *>(base);
    uint32_t *stack = base + SS_EXTRACT_TEXT_W, *trailer = stack + SX_FAST_STACK_WORDS, *fl = trailer + 1u,
             *ev = fl + SX_FAST_FLAG_WORDS(SS_EXTRACT_STAGE);
    const bool staged = mine_to_do && mlen <= SS_EXTRACT_STAGE;
    uint32_t kind = 0;
    if (staged && role == 0u) {
        const uint4 *src = reinterpret_cast<const uint4 *>(text);
        const uint32_t n16 = (mlen + 15u) >> 4;
#pragma unroll 4
        for (uint32_t b = 0; b < n16; b++) {
            const uint4 v = src[b];
            base[4 * b] = v.x; base[4 * b + 1] = v.y; base[4 * b + 2] = v.z; base[4 * b + 3] = v.w;
        }
    }
    if (staged && role == 1u) {   // the queues start empty
        const uint32_t nw = SX_FAST_FLAG_WORDS(mlen);
        *trailer = SX_PIPE_EMPTY;
        for (uint32_t w = 0; w < nw; w++) fl[w] = SX_PIPE_EMPTY;
    }
    if (mine_to_do && role == 2u) {
        kind = states[((size_t)d.slot * 2u) * 3u + 2u].w & 0xffu;   // provider kind: low byte of the tx state's meta word
        if (staged)
            for (uint32_t w = 0; w <= SS_EXTRACT_EVENTS; w++) ev[w] = SX_PIPE_EMPTY;
    }
    {
        uint4 *dst = reinterpret_cast<uint4 *>(fast);   // the class codes follow the grammar table in both places
        for (uint32_t j = threadIdx.x; j < SX_FAST_BLOCK_BYTES / 16u; j += blockDim.x) dst[j] = fast_block[j];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) PROBE_SET(0);
    __syncthreads();
    if (lane == 0) PROBE_MAX(6);
    if (role == 3u) return;
    if (!mine_to_do) {
        if (verdicts && role == 2u) verdicts[c] = SS_ST_OK;   // nothing to extract: nothing to hold the seal back for
        return;
    }
    if (role == 0u) {
        if (staged) sx::fast_grammar(mine, mlen, fast, cls, stack, fl, trailer);
        if (lane == 0) PROBE_MAX(4);
        return;
    }
    if (role == 1u) {
        if (staged) sx::fast_record(mlen, sx::fast_prefixed(mine, mlen), fl, trailer, ev, SS_EXTRACT_EVENTS);
        if (lane == 0) PROBE_MAX(7);
        return;
    }
    const uint32_t cap = mlen < 16u ? 16u : ((mlen + 15u) & ~15u);
    sx::Result r;
    bool done = false;
    if (staged) {
        text = mine;
        done = sx::fast_bookkeep(mine, mlen, sx::fast_prefixed(mine, mlen), kind, delta_arena + d.in_off, cap, ev, trailer,
                                 SS_EXTRACT_EVENTS, &r);
    }
    if (!done)
        r = sx::extract_chunk(text, mlen, kind, delta_arena + d.in_off, cap,
                              reinterpret_cast<const uint16_t *>(reinterpret_cast<const uint8_t *>(fast_block) + SX_FAST_BLOCK_BYTES));
    if (!verdicts || r.status != SS_ST_OK) status[i] = r.status;
    delta[i].off = d.in_off;
    delta[i].len = r.len;
    if (verdicts) verdicts[c] = r.status;
    if (lane == 0) PROBE_MAX(1);
}

__global__ void __launch_bounds__(SS_EXTRACT_THREADS) ss_extract_kernel(const SsDesc *desc, uint32_t n, const uint8_t *in,
                                                                       const uint4 *states, uint32_t *status,
                                                                       symrelay_delta *delta, uint8_t *delta_arena,
                                                                       const uint4 *fast_block) {
    ss_pdl_launch_dependents();   // the batch's seal kernel may start beside this one (SS_ARG_VERDICT_LATE, ss_core.cuh)
    const uint32_t warp = threadIdx.x >> 5;
    // warps 2,3: grammar (the long chain; with one block per SM each has a scheduler to itself); 0,1: recording; 4,5: bookkeeping
    ss_extract_block(desc, n, in, states, status, delta, delta_arena, fast_block, (warp & 2u) ? 0u : (warp < 2u ? 1u : 2u),
                     warp & 1u, nullptr);
}

// Extractor and seal of one batch in ONE kernel, for batches whose frames all sit in thread tiles of 32 consecutive
// descriptors (many chunks of one size class: BASELINE config 5).  A block takes 64 frames: six warps extract (three
// roles x two groups of 32 chunks) and two warps seal the same 64 frames, one lane per frame, while the verdicts are
// being worked out; a seal lane then polls its chunk's verdict word in shared memory and commits or takes the frame
// back.  Two kernels side by side (SS_ARG_VERDICT_LATE) do the same in principle, but the block scheduler packs the
// seal kernel's blocks onto the SMs the extractor left emptiest -- measured: the seal took 26 us instead of 13 -- here
// every SM gets the same mix by construction.
#define SS_FUSED_THREADS 256u
#define SS_FUSED_SMEM (SS_EXTRACT_SMEM + SS_EXTRACT_CHUNKS * 4u)
__global__ void __launch_bounds__(SS_FUSED_THREADS, 2) ss_extract_seal_kernel(SsArgs a, uint32_t n, symrelay_delta *delta,
                                                                             uint8_t *delta_arena, const uint4 *fast_block) {
    extern __shared__ __align__(16) uint8_t sx_smem[];
    volatile uint32_t *verdicts = reinterpret_cast<volatile uint32_t *>(sx_smem + SS_EXTRACT_SMEM);
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u, group = warp & 1u;
    // warps 0,1: grammar; 2,3: seal; 4,5: recording; 6,7: bookkeeping -- per scheduler (warp % 4): a grammar warp with a
    // recording warp (which mostly sleeps), a seal warp with a bookkeeping warp
    const uint32_t role = warp < 2u ? 0u : (warp < 4u ? 3u : (warp < 6u ? 1u : 2u));
    if (role == 3u) verdicts[group * 32u + lane] = SS_VERDICT_EMPTY;
    ss_extract_block(a.desc, n, a.in, a.states, a.status, delta, delta_arena, fast_block, role, group, verdicts);
    if (role != 3u) return;
    const uint32_t first = blockIdx.x * SS_EXTRACT_CHUNKS + group * 32u;
    if (first >= n) return;
    SsTile tile;
    tile.first = first;
    tile.n_kind = (n - first < 32u ? n - first : 32u) | (1u << 17);
    a.verdicts = verdicts + group * 32u;
    ss_process_thread_tile<false, 2>(a, blockIdx.x * 2u + group, &tile);
}


// ------------------------------------------------------ persistent relay kernel --
// Small host batches (the token-streaming shape, BASELINE config 2) used to cost one kernel launch each:
// 2.5 us of driver call, and 5.7 us between the end of kernel j and kernel j+1 being allowed past its
// dependency wait, because a grid only counts as complete once its host-visible stores have been flushed
// over PCIe.  This kernel stays resident instead.  The host writes a batch header into a
// ring in mapped pinned memory and bumps a doorbell; one poller warp copies new headers into device memory.
// The SMs are split into two roles:
//   * COMPUTE warps draw tile tickets from a device counter, fetch their tile's descriptors and payload from
//     the host arenas (prefetching batches ahead), wait until each frame's PREDECESSOR on its stream has left
//     the stream state (the MAC chain, handed over per stream through a stamp in the state itself -- neither a
//     grid boundary nor a barrier between batches), seal/open, and leave wire bytes and status words in
//     DEVICE staging;
//   * DRAIN warps, on other SMs, copy a finished batch's staging to the caller's host memory in 512-byte rows,
//     fence to system scope and raise that batch's completion flag in host memory.
// The split is what keeps the chain short: an SM that streams stores over PCIe acknowledges its other stores
// late (measured: the device-scope fence of a computing warp took up to 7.7 us while results were draining from
// the same SMs, 0.4 us otherwise), and that fence sits between every two batches of a stream.
// The kernel exits when it has been idle for `idle_ns` (or the host asks), so cudaDeviceSynchronize()/cudaFree()
// by other code in the process still return; the host relaunches it with the next batch.  Replaces the same
// per-frame libsodium calls as every other batch path (/root/reference/src/provider.ts:248, :174).
#define PK_RING 32u        // batches posted and not yet reaped, at most (>= the host pipeline depth)
#define PK_RUNNING 1u
#define PK_EXITING 2u
#define PK_EXITED 3u
#define PK_CHUNK 4096u     // bytes of results one drain ticket moves to the host

struct PkSlot {            // 128 bytes; written by the host, read by the poller warp
    const SsDesc *desc;
    const SsTile *tiles;
    const uint8_t *in;         // caller's arena (pinned host memory)
    uint8_t *out;              // device staging for the results
    uint32_t *status;          // device staging for the status words
    uint32_t n_desc, n_tiles;
    uint32_t flags, uni_per, uni_thread, out_n16;   // out_n16: 16-byte units of the out arena
    uint8_t *host_out;         // where the drain warps put them
    uint32_t *host_status;
    const uint32_t *ver;       // per descriptor: the stamp its stream state must carry before the frame may use it
    uint32_t new_ver, ver_all; // the stamp this batch leaves in every state it goes through; SS_ARG_CHAIN_UNIFORM: the one every frame expects
    SsDesc *desc_fill;         // non-NULL: tiles also copy their descriptors here (device memory) for the next batch that repeats them
    uint64_t pad[3];
};
static_assert(sizeof(PkSlot) == 128, "slot layout");

struct PkHost {            // mapped pinned host memory
    volatile unsigned long long doorbell;   // batches posted so far (host writes)
    volatile uint32_t quit;                 // host: exit as soon as everything posted is done
    uint32_t pad0[13];
    volatile uint32_t state;                // PK_* (the GPU writes; the host writes PK_RUNNING before a launch)
    uint32_t pad1[15];
    volatile unsigned long long exit_installed;   // batches the exiting kernel had taken (= all completed)
    uint64_t pad2[7];
    struct { volatile unsigned long long v; uint64_t pad[7]; } done[PK_RING];   // [seq % PK_RING] = seq + 1: results of batch seq are in host memory
    PkSlot slot[PK_RING];
};

struct PkEntry {           // device memory, one per ring slot
    unsigned long long range;    // first tile ticket << 24 | tiles, published last (one 8-byte store)
    unsigned long long drange;   // first drain ticket << 24 | chunks
    unsigned long long seq;
    uint32_t done, sysdone;
    PkSlot s;
    uint64_t pad[12];
};
static_assert(sizeof(PkEntry) == 256, "entry layout");
#define PK_MAX_CTAS 1024u
struct PkDev {
    unsigned long long ticket;        uint64_t pad0[15];
    unsigned long long drain_ticket;  uint64_t pad1[15];
    unsigned long long sys_completed; uint64_t pad2[15];
    // One copy per CTA of the words every idle or dependent warp polls, 128 bytes apart: with a single copy
    // hundreds of polling warps hammer one L2 slice and the store they are waiting for queues behind their loads.
    struct { unsigned long long tiles, chunks; uint64_t pad[14]; } posted[PK_MAX_CTAS];   // posted so far; bit 63 of tiles: the grid exits
    PkEntry e[PK_RING];
};
#define PK_TS_WORDS 16u     // SYMRELAY_TS=1, per batch: [0] doorbell seen, [1..4] tile 0 {start, loads issued, dependency met, done}, [5] tile 0
                            // signalled, [7] batch in host memory, [8..13] max over the batch's tiles of stamps 0..5

__device__ __forceinline__ unsigned long long pk_ld_u64(const volatile unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.volatile.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long pk_ld_acquire_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long pk_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// release/acquire fences (not the sequentially consistent ones __threadfence*() compile to)
__device__ __forceinline__ void pk_fence_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
__device__ __forceinline__ void pk_fence_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }

// Warp 0 of block 0: doorbell -> device-side batch table; decides when the grid exits.  Warp-uniform: every
// lane issues the same loads (one transaction, broadcast), lane 0 writes the entry, all lanes write the copies.
__device__ void pk_poller(PkHost *host, PkDev *dev, unsigned long long first_seq, unsigned long long idle_ns,
                          unsigned long long *ts) {
    const uint32_t lane = threadIdx.x & 31u;
    unsigned long long installed = first_seq, cum = 0, dcum = 0;
    unsigned long long t_last = pk_now();
    for (;;) {
        const uint4 db4 = ss_ld16<true>(const_cast<const unsigned long long *>(&host->doorbell));   // doorbell ‖ quit in one read
        const unsigned long long db = (unsigned long long)db4.x | ((unsigned long long)db4.y << 32);
        const uint32_t quit = db4.z;
        if (db > installed) {
            pk_fence_sys();   // acquire: what the host wrote before the doorbell (headers, control blocks, payload)
            const unsigned long long t_seen = pk_now();
            while (installed < db) {
                // Lane L installs batch installed + L: the headers of everything new are read at once (one PCIe round
                // trip, not one per batch -- under load a round trip is 5-10 us and eight batches queued up cost 37 us).
                const unsigned long long m = db - installed < 32ull ? db - installed : 32ull;
                const bool mine = lane < m;
                const unsigned long long my_seq = installed + lane;
                uint4 q[7];
#pragma unroll
                for (int i = 0; i < 7; i++) q[i] = make_uint4(0, 0, 0, 0);
                if (mine) {
                    const uint4 *hq = reinterpret_cast<const uint4 *>(&host->slot[my_seq % PK_RING]);   // plain loads, ordered by the acquire fence above
#pragma unroll
                    for (int i = 0; i < 7; i++) q[i] = hq[i];
                }
                const unsigned long long n_tiles = mine ? q[2].w : 0u;   // PkSlot::n_tiles
                const unsigned long long n_chunks = mine ? ((unsigned long long)q[3].w * 16ull + PK_CHUNK - 1u) / PK_CHUNK +
                                                               ((unsigned long long)q[2].z * 4ull + PK_CHUNK - 1u) / PK_CHUNK : 0u;   // out arena + status words
                unsigned long long it = n_tiles, ic = n_chunks;   // inclusive prefix sums over the lanes
#pragma unroll
                for (uint32_t o = 1; o < 32u; o <<= 1) {
                    const unsigned long long a = __shfl_up_sync(0xffffffffu, it, o), c = __shfl_up_sync(0xffffffffu, ic, o);
                    if (lane >= o) { it += a; ic += c; }
                }
                PkEntry *e = &dev->e[my_seq % PK_RING];
                if (mine) {
                    uint4 *es = reinterpret_cast<uint4 *>(&e->s);
#pragma unroll
                    for (int i = 0; i < 7; i++) es[i] = q[i];
                    e->seq = my_seq;
                    e->done = 0; e->sysdone = 0;
                    if (ts) {
                        unsigned long long *w = ts + (my_seq % 4096ull) * PK_TS_WORDS;
                        w[0] = t_seen; w[8] = w[9] = w[10] = w[11] = w[12] = w[13] = 0;
                    }
                }
                pk_fence_gpu();   // fields before the range words
                if (mine) {
                    *reinterpret_cast<volatile unsigned long long *>(&e->drange) = ((dcum + ic - n_chunks) << 24) | n_chunks;
                    *reinterpret_cast<volatile unsigned long long *>(&e->range) = ((cum + it - n_tiles) << 24) | n_tiles;
                }
                cum += __shfl_sync(0xffffffffu, it, 31);
                dcum += __shfl_sync(0xffffffffu, ic, 31);
                installed += m;
            }
            __syncwarp();
            pk_fence_gpu();   // release: one fence, then relaxed stores of the copies (a st.release each would fence each time)
            for (uint32_t c = lane; c < gridDim.x; c += 32u) {
                asm volatile("st.relaxed.gpu.u64 [%0], %1;" ::"l"(&dev->posted[c].chunks), "l"(dcum) : "memory");
                asm volatile("st.relaxed.gpu.u64 [%0], %1;" ::"l"(&dev->posted[c].tiles), "l"(cum) : "memory");
            }
            t_last = pk_now();
            continue;
        }
        const unsigned long long sc = pk_ld_u64(&dev->sys_completed);
        const unsigned long long now = pk_now();
        if (sc != installed - first_seq) { t_last = now; continue; }   // batches in flight: not idle
        if (quit || now - t_last > idle_ns) {
            // Exit hand-shake (the host does: doorbell store, full fence, state load).  The state store is a posted
            // PCIe write and the doorbell re-read a PCIe read, which cannot pass it: either the host sees EXITING
            // and waits for the verdict, or this re-read sees its new batch.
            if (lane == 0) host->state = PK_EXITING;
            __threadfence_system();
            const unsigned long long db2 = pk_ld_u64(&host->doorbell);
            if (db2 != installed) {
                if (lane == 0) host->state = PK_RUNNING;
                __threadfence_system();
                t_last = pk_now();
                continue;
            }
            if (lane == 0) host->exit_installed = installed;
            __threadfence_system();
            if (lane == 0) host->state = PK_EXITED;
            __threadfence_system();
            for (uint32_t c = lane; c < gridDim.x; c += 32u)
                asm volatile("st.release.gpu.u64 [%0], %1;" ::"l"(&dev->posted[c].tiles), "l"(cum | (1ull << 63)) : "memory");
            return;
        }
    }
}

// ticket -> ring entry: lane j looks at entry j; `which`: 0 tile range, 1 drain range.  Returns the entry index and
// the ticket's offset inside it.
__device__ __forceinline__ uint32_t pk_find(PkDev *dev, unsigned long long t, uint32_t lane, int which, uint32_t *k) {
    const unsigned long long r = pk_ld_u64(which ? &dev->e[lane].drange : &dev->e[lane].range);
    const unsigned long long start = r >> 24, cnt = r & 0xffffffull;
    const uint32_t hits = __ballot_sync(0xffffffffu, t >= start && t < start + cnt);
    const uint32_t j = (uint32_t)__ffs((int)hits) - 1u;   // exactly one entry holds a posted ticket
    *k = (uint32_t)(t - __shfl_sync(0xffffffffu, start, (int)j));
    return j;
}

// DRAIN role: staging -> the caller's host memory, PK_CHUNK bytes per ticket.
__device__ void pk_drain(PkHost *host, PkDev *dev, unsigned long long first_seq, unsigned long long *ts) {
    const uint32_t lane = threadIdx.x & 31u;
    for (;;) {
        unsigned long long t = 0;
        if (lane == 0) t = atomicAdd(&dev->drain_ticket, 1ull);
        t = __shfl_sync(0xffffffffu, t, 0);
        for (;;) {
            if (pk_ld_acquire_u64(&dev->posted[blockIdx.x].chunks) > t) break;
            if (pk_ld_acquire_u64(&dev->posted[blockIdx.x].tiles) >> 63) return;
            __nanosleep(200);
        }
        uint32_t k;
        PkEntry *e = &dev->e[pk_find(dev, t, lane, 1, &k)];
        pk_fence_gpu();
        const PkSlot s = e->s;
        const unsigned long long seq = e->seq;
        while (ss_ld_acquire_u32(&e->done) != s.n_tiles) __nanosleep(100);   // every tile of the batch has stored its results
        const uint32_t out_chunks = (uint32_t)(((unsigned long long)s.out_n16 * 16ull + PK_CHUNK - 1u) / PK_CHUNK);
        const uint4 *src;
        uint4 *dst;
        uint32_t n16;
        if (k < out_chunks) {
            src = reinterpret_cast<const uint4 *>(s.out) + (size_t)k * (PK_CHUNK / 16u);
            dst = reinterpret_cast<uint4 *>(s.host_out) + (size_t)k * (PK_CHUNK / 16u);
            n16 = s.out_n16 - k * (PK_CHUNK / 16u);
        } else {
            const uint32_t kk = k - out_chunks, n16s = (s.n_desc + 3u) >> 2;   // status words, 16 bytes at a time (both buffers are padded)
            src = reinterpret_cast<const uint4 *>(s.status) + (size_t)kk * (PK_CHUNK / 16u);
            dst = reinterpret_cast<uint4 *>(s.host_status) + (size_t)kk * (PK_CHUNK / 16u);
            n16 = n16s - kk * (PK_CHUNK / 16u);
        }
        if (n16 > PK_CHUNK / 16u) n16 = PK_CHUNK / 16u;
        uint4 v[PK_CHUNK / 512u];
#pragma unroll
        for (uint32_t i = 0; i < PK_CHUNK / 512u; i++)
            if (lane + 32u * i < n16) v[i] = src[lane + 32u * i];
#pragma unroll
        for (uint32_t i = 0; i < PK_CHUNK / 512u; i++)
            if (lane + 32u * i < n16) dst[lane + 32u * i] = v[i];
        pk_fence_sys();
        __syncwarp();
        if (lane == 0 && atomicAdd(&e->sysdone, 1u) + 1u == (uint32_t)(pk_ld_u64(&e->drange) & 0xffffffull)) {
            // Every drain warp fenced its rows to system scope BEFORE it counted itself (release); this is the
            // acquire side of that device-scope counter, so a device-scope fence completes the chain: the flag below
            // is ordered after all of the batch's rows for any observer, the host included.  (A second system-scope
            // fence here cost 3-5 us of every batch's latency.)
            pk_fence_gpu();
            host->done[seq % PK_RING].v = seq + 1ull;
            if (ts) ts[(seq % 4096ull) * PK_TS_WORDS + 7] = pk_now();
            atomicAdd(&dev->sys_completed, 1ull);
        }
    }
}

__global__ void __launch_bounds__(128, 3) ss_persistent_kernel(PkHost *host, PkDev *dev, uint4 *states,
                                                                 unsigned long long first_seq, unsigned long long idle_ns,
                                                                 unsigned long long *ts, uint32_t drain_every) {
    const uint32_t lane = threadIdx.x & 31u;
    if (blockIdx.x == 0 && (threadIdx.x >> 5) == 0) { pk_poller(host, dev, first_seq, idle_ns, ts); return; }
    uint32_t smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    if (smid % drain_every == drain_every - 1u) { pk_drain(host, dev, first_seq, ts); return; }   // role by SM, not by block
    const unsigned long long *my_posted = &dev->posted[blockIdx.x].tiles;
    for (;;) {
        unsigned long long t = 0;
        if (lane == 0) t = atomicAdd(&dev->ticket, 1ull);
        t = __shfl_sync(0xffffffffu, t, 0);
        for (;;) {   // wait until the ticket names a posted tile
            const unsigned long long p = pk_ld_acquire_u64(my_posted);
            if ((p & ~(1ull << 63)) > t) break;
            if (p >> 63) return;
            __nanosleep(200);
        }
        uint32_t k;
        PkEntry *e = &dev->e[pk_find(dev, t, lane, 0, &k)];
        pk_fence_gpu();   // the entry's fields were written before its range word
        const PkSlot s = e->s;
        const unsigned long long seq = e->seq;
        SsArgs a;
        a.desc = s.desc;
        a.tiles = s.tiles;
        a.in = s.in;
        a.out = s.out;
        a.status = s.status;
        a.n_tiles = s.n_tiles;
        a.states = states;
        a.flags = s.flags | SS_ARG_INPUT_INDEPENDENT | SS_ARG_CHAIN;
        a.run_nonce = nullptr;
        a.ver = s.ver;           // the MAC chain is handed over per stream (ss_core.cuh), not per batch
        a.new_ver = s.new_ver;
        a.ver_all = s.ver_all;
        a.desc_fill = s.desc_fill;
        unsigned long long *my_ts = (ts && k == 0) ? ts + (seq % 4096ull) * PK_TS_WORDS : nullptr;
        a.dbg_ts = my_ts ? my_ts + 1 : nullptr;     // ss_process_tile stamps [1..4] for tile index 0
        a.dbg_max = ts ? ts + (seq % 4096ull) * PK_TS_WORDS + 8 : nullptr;   // [8..]: the batch's LAST tile to get there
        SsTile tile;
        if (s.uni_per) {   // every frame has one block count: the tile follows from its index
            if (k < s.uni_thread) { tile.first = 32u * k; tile.n_kind = 32u | (1u << 17); }
            else {
                tile.first = 32u * s.uni_thread + (k - s.uni_thread) * s.uni_per;
                tile.n_kind = s.n_desc - tile.first < s.uni_per ? s.n_desc - tile.first : s.uni_per;
            }
        } else {
            tile = s.tiles[k];
        }
        // Plain loads throughout: the acquire loads above (and in the dependency wait) order them after the host's
        // and the previous batch's stores, and pinned host memory is not kept in the GPU's caches (tools/probes).
        if (tile.n_kind & (1u << 17)) ss_process_thread_tile<false>(a, k, &tile);
        else ss_process_tile<false, false>(a, k, &tile);
        // this tile's results are in device staging: count it; the batch's last tile releases the drain
        pk_fence_gpu();
        __syncwarp();
        if (lane == 0) atomicAdd(&e->done, 1u);
        ss_dbg_stamp(a, k, lane, 4);
        if (my_ts && lane == 0) my_ts[5] = pk_now();
    }
}

// Connection set-up in batch, one thread per new connection: the two stream ids of
// secret-stream's first message and a fresh secretstream header for the outbound direction.
//   id_local  = BLAKE2b-256(key = handshake hash, msg = namespace of OUR role)     (crypto_generichash)
//   id_remote = ... of the peer's role: what the first inbound message must carry
//   header    = 24 bytes of ChaCha20(seed, nonce = connection index) -- the seed is 32 fresh random bytes per call
// rec: hash[64] ‖ initiator flag (u32) ‖ pad; out: id_local[32] ‖ id_remote[32] ‖ header[24] ‖ pad[8]
struct SsConnRec { uint8_t hash[64]; uint32_t initiator; uint32_t pad[3]; };
struct SsConnOut { uint8_t id_local[32], id_remote[32], header[24], pad[8]; };
static_assert(sizeof(SsConnRec) == 80 && sizeof(SsConnOut) == 96, "connect record layout");
__global__ void ss_connect_kernel(const SsConnRec *recs, uint32_t n, const uint4 *ns_pair /* initiator[32] ‖ responder[32] */,
                                  const uint4 *seed /* 32 bytes */, uint32_t call_no, SsConnOut *out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t *kw = reinterpret_cast<const uint64_t *>(recs[i].hash);
    uint64_t key[8], ns_i[4], ns_r[4], idl[4], idr[4];
#pragma unroll
    for (int j = 0; j < 8; j++) key[j] = kw[j];
    const uint64_t *nw = reinterpret_cast<const uint64_t *>(ns_pair);
#pragma unroll
    for (int j = 0; j < 4; j++) { ns_i[j] = nw[j]; ns_r[j] = nw[4 + j]; }
    const bool init = recs[i].initiator != 0;
    sb_blake2b256_keyed(key, init ? ns_i : ns_r, idl);
    sb_blake2b256_keyed(key, init ? ns_r : ns_i, idr);
    uint64_t *o = reinterpret_cast<uint64_t *>(&out[i]);
#pragma unroll
    for (int j = 0; j < 4; j++) { o[j] = idl[j]; o[4 + j] = idr[j]; }
    uint32_t k[8], x[16];
    const uint32_t *sw = reinterpret_cast<const uint32_t *>(seed);
#pragma unroll
    for (int j = 0; j < 8; j++) k[j] = sw[j];
    ss_chacha_block(k, 0, i, call_no, 0x73796d72u, x);
    uint32_t *ho = reinterpret_cast<uint32_t *>(out[i].header);
#pragma unroll
    for (int j = 0; j < 6; j++) ho[j] = x[j];
}

// Stream set-up (libsodium init_push/init_pull): one thread per
// (slot, direction).  rec = key[32] ‖ init[init_len]; init_len 24 derives the
// sub-key with HChaCha20, 52 adopts an existing libsodium state, 0 clears.
struct SsInitRec {
    uint32_t slot_dir;  // slot * 2 + dir
    uint32_t init_len;
    uint32_t meta;
    uint32_t pad;
    uint8_t key[32];
    uint8_t init[64];
};
static_assert(sizeof(SsInitRec) == 112, "init record layout");

__global__ void ss_init_kernel(const SsInitRec *recs, uint32_t n, uint4 *states) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const SsInitRec *r = recs + i;
    const uint32_t *kw = reinterpret_cast<const uint32_t *>(r->key);
    const uint32_t *iw = reinterpret_cast<const uint32_t *>(r->init);
    uint32_t k[8], nn[3];
    if (r->init_len == 24u) {
        // HChaCha20(key, header[0:16]): ChaCha20 rounds without the feed-forward, words 0..3 and 12..15
        uint32_t x[16];
        x[0] = 0x61707865u; x[1] = 0x3320646eu; x[2] = 0x79622d32u; x[3] = 0x6b206574u;
#pragma unroll
        for (int j = 0; j < 8; j++) x[4 + j] = kw[j];
#pragma unroll
        for (int j = 0; j < 4; j++) x[12 + j] = iw[j];
#pragma unroll 1
        for (int rr = 0; rr < 10; rr++) {
            SS_QR(x[0], x[4], x[8], x[12])
            SS_QR(x[1], x[5], x[9], x[13])
            SS_QR(x[2], x[6], x[10], x[14])
            SS_QR(x[3], x[7], x[11], x[15])
            SS_QR(x[0], x[5], x[10], x[15])
            SS_QR(x[1], x[6], x[11], x[12])
            SS_QR(x[2], x[7], x[8], x[13])
            SS_QR(x[3], x[4], x[9], x[14])
        }
        k[0] = x[0]; k[1] = x[1]; k[2] = x[2]; k[3] = x[3];
        k[4] = x[12]; k[5] = x[13]; k[6] = x[14]; k[7] = x[15];
        nn[0] = 1u; nn[1] = iw[4]; nn[2] = iw[5];
    } else if (r->init_len == 52u) {
#pragma unroll
        for (int j = 0; j < 8; j++) k[j] = iw[j];
        nn[0] = iw[8]; nn[1] = iw[9]; nn[2] = iw[10];
    } else {
#pragma unroll
        for (int j = 0; j < 8; j++) k[j] = 0;
        nn[0] = nn[1] = nn[2] = 0;
    }
    uint4 *sp = states + (size_t)r->slot_dir * 3u;
    sp[0] = make_uint4(k[0], k[1], k[2], k[3]);
    sp[1] = make_uint4(k[4], k[5], k[6], k[7]);
    sp[2] = make_uint4(nn[0], nn[1], nn[2], r->meta);
}

// Integer-throughput probes: the roofline denominator MEASURED_PEAKS.json lacks
// (SURVEY.md section 8(d)).  ARX: ChaCha20's own ADD : XOR : ROTATE = 1:1:1 mix
// on 16 register words with no memory traffic (4 independent quarter-rounds
// in flight per thread); IMAD: independent 32x32->64 multiply-add chains as in
// the Poly1305 limb products.
__global__ void __launch_bounds__(256) ss_probe_arx_kernel(uint32_t *out, int double_rounds) {
    uint32_t x[16];
#pragma unroll
    for (int i = 0; i < 16; i++) x[i] = threadIdx.x * 0x9e3779b9u + blockIdx.x + i;
#pragma unroll 1
    for (int r = 0; r < double_rounds; r++) {
        SS_QR(x[0], x[4], x[8], x[12])
        SS_QR(x[1], x[5], x[9], x[13])
        SS_QR(x[2], x[6], x[10], x[14])
        SS_QR(x[3], x[7], x[11], x[15])
        SS_QR(x[0], x[5], x[10], x[15])
        SS_QR(x[1], x[6], x[11], x[12])
        SS_QR(x[2], x[7], x[8], x[13])
        SS_QR(x[3], x[4], x[9], x[14])
    }
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) acc ^= x[i];
    if (acc == 0x12345u) out[0] = acc;  // keep the work alive
}

__global__ void __launch_bounds__(256) ss_probe_imad_kernel(uint32_t *out, int iters) {
    uint64_t d[8];
    uint32_t a = threadIdx.x * 2654435761u + 1u, b = blockIdx.x * 40503u + 7u;
#pragma unroll
    for (int i = 0; i < 8; i++) d[i] = i;
#pragma unroll 1
    for (int r = 0; r < iters; r++) {
#pragma unroll
        for (int i = 0; i < 8; i++) d[i] += (uint64_t)(a + i) * (uint32_t)(b ^ (uint32_t)d[(i + 1) & 7]);
    }
    uint64_t acc = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) acc ^= d[i];
    if (acc == 0x12345ull) out[0] = (uint32_t)acc;
}

// ---- run batches without any allocation -------------------------------------------------------
// Who needs this: a provider usually shares its GPU with the model server it relays for (Ollama,
// llama.cpp ...), which has claimed nearly all of HBM before the relay receives its first bulk
// segment; an allocation for run scratch made at that moment can simply fail.  And the caller is
// Node's event-loop thread: cudaMalloc/cudaFree synchronise the whole device, which would stall
// every connection's batch in flight.  For those callers the nonce scratch of a run batch lives in
// this fixed array instead of caller- or library-allocated memory.  Cost: 1 MiB of HBM on every
// GPU that loads the library, held until the process exits, and a hard cap on the batch size
// (SYMRELAY_STATIC_RUN_MAX descriptors, the one definition in include/symrelay.h).
// In host code `g_run_nonce_static` is only a handle for the runtime, not an address a kernel can
// use: shard_init() looks the device address up once per GPU, right after making that GPU current,
// and keeps it in Shard::d_run_static.
static __device__ __align__(256) uint4 g_run_nonce_static[SYMRELAY_STATIC_RUN_MAX];
static_assert(sizeof(uint4) * SYMRELAY_STATIC_RUN_MAX == (1u << 20), "the header promises 1 MiB per GPU");

// --------------------------------------------------------------- host side --
#define SLOT_CLOSED 0xffffffffu   /* slot_stamp entry of a slot that is not open */
namespace {

struct Stage {
    uint8_t *d_ctrl = nullptr;   // tiles ‖ descs
    uint8_t *h_ctrl = nullptr;   // pinned mirror
    uint8_t *d_in = nullptr, *d_out = nullptr;
    uint32_t *d_status = nullptr, *h_status = nullptr;
    symrelay_delta *d_delta = nullptr, *h_delta = nullptr;
    uint8_t *d_delta_arena = nullptr;   // allocated on first use (extraction batches only)
    uint4 *d_run_nonce = nullptr;       // scratch for runs, allocated on first use
    symrelay_delta *user_delta = nullptr;
    cudaEvent_t ev_in = nullptr, ev_k = nullptr, ev_out = nullptr;
    // in-flight batch bookkeeping
    int64_t ticket = -1;
    uint32_t n = 0;
    uint32_t *user_status = nullptr;
    bool busy = false;
    // Deferred write-back (symrelay_batch_submit pipeline, see submit_locked): the kernel left this batch's
    // results in d_out / d_status; the NEXT kernel on the stream (stage `wb_by`) posts them to the caller's
    // host memory, or a small flush kernel does if nothing follows.
    bool deferred = false;
    int wb_by = -1;               // stage whose ev_out covers the write-back; -1: not issued yet
    uint8_t *user_out = nullptr;
    uint64_t out_bytes = 0;
    // Bucketed planning (large staged batches): the kernels saw the descriptors sorted by block count;
    // perm[k] = the caller's index of sorted descriptor k, applied to statuses and deltas on completion.
    std::vector<uint32_t> perm;
    bool permuted = false;
    int64_t pk_seq = -1;          // >= 0: this batch went to the persistent kernel as its batch number pk_seq
    uint64_t tmpl_gen = 0;        // h_ctrl holds tiles ‖ descriptors of this template generation (0: something else)
    uint64_t dev_tmpl_gen = 0;    // d_ctrl holds the descriptors of this template generation (left there by the tiles of an earlier batch)
};

struct Shard {
    std::mutex mu;                         // everything below and this GPU's Sched: one lock per GPU, none shared between GPUs
    int dev = -1;
    uint4 *d_states = nullptr;
    uint4 *d_sx_fast = nullptr;            // extractor tables (SX_TABLES_BYTES), allocated last in shard_init
    uint4 *d_run_static = nullptr;         // device address of g_run_nonce_static on this GPU (symrelay_batch_device_runs_static)
    // persistent kernel (small zero-copy host batches)
    PkHost *pk_host = nullptr;             // mapped pinned: doorbell, state, completion flags, batch headers
    PkDev *pk_dev = nullptr;
    uint64_t pk_posted = 0;                // batches posted so far
    int pk_grid = 0;                       // CTAs of the resident grid (all co-resident)
    unsigned long long *d_pk_ts = nullptr; // SYMRELAY_TS=1: PK_TS_WORDS stamps per batch, 4096 batches round-robin
    bool pk_quit_req = false;              // the kernel was asked to exit because ordinary launches follow on s_k
    uint32_t pk_floor = 0;                 // batch stamps up to here are settled (an ordinary launch on s_k followed them)
    bool stamp_dirty = false;              // the last batch stamped the slot table but never ran (refused / failed)
    // Template: the descriptors of the last batch that took the general path to the resident kernel, with what
    // planning and checking them produced.  A batch whose descriptors (and arena sizes) are byte-identical, and
    // which directly follows a batch of that template, needs none of that again: same tiles, same streams, every
    // frame's predecessor is the previous batch.  Token streaming produces exactly this (BASELINE config 2).
    std::vector<symrelay_desc> tmpl_desc;
    uint64_t tmpl_in_bytes = 0, tmpl_out_bytes = 0, tmpl_sum = 0, last_sum = 0;
    uint32_t tmpl_n = 0, tmpl_n_tiles = 0, tmpl_uni_per = 0, tmpl_uni_thread = 0, tmpl_flags = 0;
    size_t tmpl_tiles_bytes = 0;
    uint64_t tmpl_gen = 0;                 // identity of the template (0: none); Stage::tmpl_gen says whose control block a stage holds
    uint64_t slot_gen = 0, tmpl_slot_gen = 0;   // bumped by every stream open/close on this GPU
    bool tmpl_streak = false;              // the previous batch on this GPU was this template, through the resident kernel
    bool tmpl_lazy = false;                // slot_stamp entries of the template's streams are behind by (stamp - tmpl_stamped) batches
    std::vector<uint8_t> tmpl_tiles;       // the template's tile list (non-uniform batches)
    cudaStream_t s_init = nullptr;         // stream set-up / export: must not queue behind the resident kernel on s_k
    std::vector<uint32_t> slot_stamp;      // [slot*2+dir] -> stamp of the last host batch that named it (duplicate check)
    uint32_t stamp = 0;
    std::vector<uint8_t> h_sx_fast;
    cudaStream_t s_in = nullptr, s_k = nullptr, s_out = nullptr;
    int pending_wb = -1;                           // stage whose results still sit in device staging
    unsigned long long *d_ts = nullptr;            // SYMRELAY_TS=1: 4 time stamps per batch, 4096 batches round-robin
    uint64_t ts_n = 0;
    cudaStream_t s_aux = nullptr, s_aux2 = nullptr, s_aux3 = nullptr, s_aux4 = nullptr;   // side streams: the tile kinds of one batch run next to each other
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_join2 = nullptr, ev_join3 = nullptr, ev_join4 = nullptr;
    std::vector<Stage> stages;
    SsInitRec *h_init = nullptr, *d_init = nullptr;  // staging for stream_open
    uint32_t init_cap = 0;
    int64_t next_ticket_seq = 0;
};

}  // namespace

#include "scheduler.inc"

struct symrelay_ctx {
    int n_gpus = 0;
    Shard shard[8];
    Sched sched[8];
    symrelay_opts opts{};
    std::vector<uint8_t> slot_state;      // 0 free, 1 open
    std::vector<uint8_t> slot_kind;       // provider kind given at open (extractor statistics)
    std::vector<uint32_t> free_slots[8];  // per GPU (slot % n_gpus == gpu)
    std::mutex mu;        // the slot table (slot_state, free_slots) only; taken before a shard's lock, never after
    std::mutex err_mu;    // err
    std::mutex pin_mu;    // writers of pins[] (leaf lock)
    // Completion wake-up (symrelay_event_fd): one watcher thread per context, started on first use
    int evfd = -1;
    std::thread watcher;
    std::atomic<bool> watcher_stop{false};
    std::atomic<uint32_t> watcher_work{0};   // bumped by the front end whenever a frame is accepted or a batch goes out
    std::mutex watch_mu;
    std::condition_variable watch_cv;
    std::string err;
    std::atomic<uint64_t> launches{0};
    int64_t ticket_counter = 0;
    // arenas handed out by symrelay_alloc_pinned (and the front end's own): read without a lock on every batch
    struct Pin { std::atomic<uintptr_t> base{0}; std::atomic<uint64_t> bytes{0}; };
    Pin pins[2048];
    std::atomic<uint32_t> n_pins{0};
    uint64_t zero_copy_max = 1ull << 20;                   // in+out bytes up to which a batch runs zero-copy
    uint64_t zero_copy_batches = 0, staged_batches = 0;
    bool pdl = true;   // SYMRELAY_PDL=0 turns programmatic dependent launch off
    bool inline_tiles = true;   // SYMRELAY_INLINE_TILES=0: small host batches read their tile list from memory like all others
    bool bucket = true;     // SYMRELAY_BUCKET=0: large staged batches are planned in the caller's descriptor order
    bool persist = true;    // SYMRELAY_PERSIST=0: small host batches are launched one kernel per batch (the round-1 path)
    uint64_t pk_idle_ns = 200000;   // SYMRELAY_PERSIST_IDLE_US: the resident kernel exits after this long without work
    uint64_t pk_batches = 0, pk_launches = 0, tmpl_hits = 0;
    uint32_t pk_ctas_per_sm = 1;    // SYMRELAY_PERSIST_CTAS_PER_SM
    uint32_t pk_drain_every = 4;    // SYMRELAY_PERSIST_DRAIN_EVERY: every n-th SM copies results to the host, the others compute
    bool defer_wb = true;   // SYMRELAY_DEFER=0: pipelined zero-copy batches write their results to the host themselves
    bool stage_input = true;       // SYMRELAY_STAGE=0: zero-copy batches read the arena lane by lane
    uint64_t inject_refusal = 0;   // SYMRELAY_INJECT_REFUSAL=k: the front end's k-th batch is refused (tests of the error path)
    bool extract_fused = true;     // SYMRELAY_EXTRACT_FUSED=0: never the fused extract+seal kernel (two kernels, see extract_overlap)
    bool extract_overlap = true;   // SYMRELAY_EXTRACT_OVERLAP=0: the seal kernel of an extract batch starts only when the extractor is done
    uint32_t device_arg_flags = 0; // SYMRELAY_STAGE_DEV=1: also stage on the device-resident entry points (experiment)
    // SYMRELAY_PROF=1: where a host batch spends its time (printed by symrelay_destroy)
    bool prof = false;
    double prof_t[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint64_t prof_n = 0;
};

static inline double prof_now() {
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return (double)t.tv_sec * 1e6 + 1e-3 * (double)t.tv_nsec;
}
#define PROF_MARK(ctx, k)                                        \
    do {                                                         \
        if ((ctx)->prof) {                                       \
            const double prof_t1 = prof_now();                   \
            (ctx)->prof_t[k] += prof_t1 - prof_t0;               \
            prof_t0 = prof_t1;                                   \
        }                                                        \
    } while (0)

static bool is_library_pinned(const symrelay_ctx *ctx, const void *p, uint64_t bytes) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const uint32_t n = ctx->n_pins.load(std::memory_order_acquire);
    for (uint32_t i = 0; i < n; i++) {
        const uintptr_t base = ctx->pins[i].base.load(std::memory_order_relaxed);
        if (base && a >= base && a + bytes <= base + ctx->pins[i].bytes.load(std::memory_order_relaxed)) return true;
    }
    return false;
}
static bool pin_add(symrelay_ctx *ctx, void *p, uint64_t bytes) {
    std::lock_guard<std::mutex> lk(ctx->pin_mu);
    const uint32_t n = ctx->n_pins.load(std::memory_order_relaxed);
    uint32_t at = n;
    for (uint32_t i = 0; i < n; i++)
        if (!ctx->pins[i].base.load(std::memory_order_relaxed)) { at = i; break; }
    if (at >= 2048u) return false;
    ctx->pins[at].bytes.store(bytes, std::memory_order_relaxed);
    ctx->pins[at].base.store(reinterpret_cast<uintptr_t>(p), std::memory_order_release);
    if (at == n) ctx->n_pins.store(n + 1u, std::memory_order_release);
    return true;
}
static void pin_remove(symrelay_ctx *ctx, void *p) {
    std::lock_guard<std::mutex> lk(ctx->pin_mu);
    const uint32_t n = ctx->n_pins.load(std::memory_order_relaxed);
    for (uint32_t i = 0; i < n; i++)
        if (ctx->pins[i].base.load(std::memory_order_relaxed) == reinterpret_cast<uintptr_t>(p)) {
            ctx->pins[i].base.store(0, std::memory_order_release);
            break;
        }
}

static thread_local char g_err[256];

static int fail(symrelay_ctx *ctx, int code, const char *fmt, ...) {
    char buf[256];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) { std::lock_guard<std::mutex> lk(ctx->err_mu); ctx->err = buf; }
    snprintf(g_err, sizeof g_err, "%s", buf);
    return code;
}

#define CU(ctx, call)                                                                                      \
    do {                                                                                                   \
        cudaError_t e__ = (call);                                                                          \
        if (e__ != cudaSuccess)                                                                            \
            return fail(ctx, SYMRELAY_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, \
                        __LINE__);                                                                         \
    } while (0)

static inline uint64_t align_up(uint64_t v, uint64_t a) { return (v + a - 1) / a * a; }

// cudaSetDevice on every call showed up in the per-batch host cost, so it is skipped when the runtime
// says the device is current already.  The question goes to the runtime every time (cudaGetDevice is a
// thread-local read): other CUDA users on the thread -- torch, another addon -- switch devices behind us.
static inline cudaError_t set_device(int dev) {
    int cur = -1;
    if (cudaGetDevice(&cur) == cudaSuccess && cur == dev) return cudaSuccess;
    return cudaSetDevice(dev);
}

// Launch with the programmatic-dependent-launch attribute so that consecutive batches overlap:
// the next kernel's launch latency and its descriptor / payload fetches run under the previous
// kernel; it blocks at griddepcontrol.wait before touching the stream states.
template <typename K>
static cudaError_t launch_pdl(K kernel, uint32_t grid, uint32_t threads, cudaStream_t stream, bool pdl, SsArgs a) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(threads);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, a);
}

static int launch_tiles(symrelay_ctx *ctx, const symrelay_desc *d_desc, uint32_t n, const symrelay_tile *d_tiles,
                        uint32_t n_tiles, uint32_t n_bulk, const uint8_t *d_in, uint8_t *d_out, uint4 *d_states,
                        uint32_t *d_status, symrelay_delta *d_delta, uint8_t *d_delta_arena, cudaStream_t stream,
                        uint32_t arg_flags = 0, uint4 *d_run_nonce = nullptr, Shard *fork = nullptr,
                        const Stage *write_back = nullptr, const symrelay_tile *host_tiles = nullptr) {
    const uint32_t n_packed = n_tiles - n_bulk;
    if (d_run_nonce) {
        ss_run_nonce_kernel<<<(n + 127) / 128, 128, 0, stream>>>(reinterpret_cast<const SsDesc *>(d_desc), n, d_in,
                                                                 d_states, d_run_nonce);
        ctx->launches++;
        arg_flags &= ~SS_ARG_INPUT_INDEPENDENT;   // the tile kernel now depends on the pre-pass
    }
    // Every frame in a thread tile of 32 (the planner's choice for many small frames of one size class), nothing else in
    // the batch: extractor and seal go into one kernel (ss_extract_seal_kernel), which cuts the descriptors into groups of
    // 32 itself -- any frame can be a thread tile's lane, so the tile list is only the planner's opinion that this is the
    // throughput shape.
    const bool fused = d_delta && d_delta_arena && ctx->extract_fused && !d_run_nonce && !write_back && n_bulk == 0 &&
                       n_tiles != 0 && (uint64_t)n_tiles * 32u == n;
    const uint4 *fast_block = nullptr;   // the extractor tables of the GPU that owns this state table
    if (d_delta && d_delta_arena) {
        for (int g = 0; g < ctx->n_gpus; g++)
            if (ctx->shard[g].d_states == d_states) fast_block = ctx->shard[g].d_sx_fast;
        if (!fast_block) return fail(ctx, SYMRELAY_EINVAL, "extract: no extractor tables for this state table");
    }
    if (d_delta && d_delta_arena && !fused) {
        ss_extract_kernel<<<(n + SS_EXTRACT_CHUNKS - 1) / SS_EXTRACT_CHUNKS, SS_EXTRACT_THREADS,
                            SS_EXTRACT_SMEM, stream>>>(
            reinterpret_cast<const SsDesc *>(d_desc), n, d_in, d_states, d_status, d_delta, d_delta_arena, fast_block);
        ctx->launches++;
    }
    SsArgs a;
    a.desc = reinterpret_cast<const SsDesc *>(d_desc);
    a.in = d_in;
    a.out = d_out;
    a.states = d_states;
    a.status = d_status;
    a.flags = arg_flags;
    a.run_nonce = d_run_nonce;
    if (fused) {
        a.tiles = nullptr;
        a.n_tiles = n_tiles;
        a.flags &= ~(SS_ARG_STAGE_INPUT | SS_ARG_STAGE_OUTPUT);
        ss_extract_seal_kernel<<<(n + SS_EXTRACT_CHUNKS - 1) / SS_EXTRACT_CHUNKS, SS_FUSED_THREADS, SS_FUSED_SMEM, stream>>>(
            a, n, d_delta, d_delta_arena, fast_block);
        ctx->launches++;
        CU(ctx, cudaGetLastError());
        return 0;
    }
    // PDL only where it was measured to help: the host zero-copy path (16.2 -> 10.9 us per C2 batch).
    // Device-resident launches replayed from a CUDA graph got slower with it (6.3 -> 10.3 us).
    const bool pdl = ctx->pdl && (arg_flags & SS_ARG_INPUT_INDEPENDENT);
    // The tile kinds of one batch touch disjoint frames, so each kind gets its own kernel and, when several are
    // present, its own side stream next to the caller's (fork/join with events; BASELINE config 3: thread tiles
    // 24 us, packed 16, strided 15, big frames 36 when run one after the other):
    //   thread tiles (and the combined small-tile kernel of small batches)  -> the caller's stream
    //   big frames, one per warp                                           -> side stream 1
    //   strided tiles (mid-size frames)                                    -> side stream 2
    //   packed tiles of a large batch                                      -> side stream 3
    const bool big = n_packed > 148u * 8u && !write_back;   // large batch: split the small tiles by kind (see ss_thread_kernel)
    bool has_thread = n_packed != 0, has_packed = big, has_strided = n_bulk != 0, has_bulk = n_bulk != 0,
         has_wide = n_bulk != 0 && (n <= SYMRELAY_WIDE_MAX_BATCH_FRAMES ||   // (the planner forms no wide tiles in larger batches ...
                                    (getenv("SYMRELAY_WIDE") && atoi(getenv("SYMRELAY_WIDE")) == 2));   // ... unless told to, for measurements)
    uint32_t wide_lo = 0, wide_hi = n_bulk;   // the range of bulk-class tiles the wide kernel is launched over
    if (host_tiles) {   // the tile list is in host memory: launch only what is there
        has_strided = has_bulk = has_wide = false;
        wide_lo = n_bulk; wide_hi = 0;
        for (uint32_t i = n_packed; i < n_tiles; i++) {
            if ((host_tiles[i].n_kind >> 24) & 63u) has_strided = true;
            else if (host_tiles[i].n_kind & (1u << 18)) {
                has_wide = true;
                if (i - n_packed < wide_lo) wide_lo = i - n_packed;
                wide_hi = i - n_packed + 1u;
            } else has_bulk = true;
        }
        if (big) {
            has_thread = has_packed = false;
            for (uint32_t i = 0; i < n_packed && !(has_thread && has_packed); i++) {
                if (host_tiles[i].n_kind & (1u << 17)) has_thread = true; else has_packed = true;
            }
        }
    }
    const int kinds = (n_packed ? (big ? (int)has_thread + (int)has_packed : 1) : 0) + (int)has_strided + (int)has_bulk + (int)has_wide;
    const bool side = fork && kinds > 1;
    // A batch with an extractor pass whose small tiles all go to one kernel right behind it on this stream: that kernel is
    // launched as the extractor's programmatic dependent and its thread tiles ask for the verdicts last (ss_core.cuh).
    const bool late = d_delta && d_delta_arena && ctx->extract_overlap && !d_run_nonce && !write_back && kinds == 1 && n_packed;
    if (late) a.flags |= SS_ARG_VERDICT_LATE;
    if (side) CU(ctx, cudaEventRecord(fork->ev_fork, stream));
    cudaEvent_t joins[4];
    int n_joins = 0;
    // the stream for a kind that has a side stream: that stream (made to wait on the fork) when the batch forks at all
    auto side_stream = [&](cudaStream_t s_side, cudaEvent_t ev) -> cudaStream_t {
        if (!side) return stream;
        cudaStreamWaitEvent(s_side, fork->ev_fork, 0);
        joins[n_joins++] = ev;
        return s_side;
    };
    // Launch order = the order in which the block scheduler hands out SM resources: the kinds with the longest warps
    // first (a 16 KiB frame keeps its warp for nine rounds), the throughput kernels fill what is left.
    if (n_bulk) {
        SsArgs b = a;
        b.tiles = reinterpret_cast<const SsTile *>(d_tiles) + n_packed;
        b.n_tiles = n_bulk;
        // (One warp per block, so that a warp that finds the other kind's tile gives its registers back at once, was
        // measured and is worse: 71 -> 80 us on BASELINE config 3.)
        const uint32_t threads = n_bulk <= 148u * 8u ? 32u : 128u;
        const uint32_t wpb = threads / 32u;
        if (has_wide) {   // the longest frames of the batch: first
            SsArgs w = b;
            w.tiles = b.tiles + wide_lo;
            w.n_tiles = wide_hi - wide_lo;
            cudaStream_t s = side_stream(fork ? fork->s_aux4 : nullptr, fork ? fork->ev_join4 : nullptr);
            CU(ctx, launch_pdl(ss_wide_kernel, w.n_tiles, 128, s, false, w));
            ctx->launches++;
            if (s != stream) CU(ctx, cudaEventRecord(fork->ev_join4, s));
        }
        if (has_bulk) {
            cudaStream_t s = side_stream(fork ? fork->s_aux : nullptr, fork ? fork->ev_join : nullptr);
            CU(ctx, launch_pdl(ss_bulk_kernel, (n_bulk + wpb - 1) / wpb, threads, s, pdl && !side, b));
            ctx->launches++;
            if (s != stream) CU(ctx, cudaEventRecord(fork->ev_join, s));
        }
        if (has_strided) {
            cudaStream_t s = side_stream(fork ? fork->s_aux2 : nullptr, fork ? fork->ev_join2 : nullptr);
            CU(ctx, launch_pdl(ss_strided_kernel, (n_bulk + wpb - 1) / wpb, threads, s, false, b));
            ctx->launches++;
            if (s != stream) CU(ctx, cudaEventRecord(fork->ev_join2, s));
        }
    }
    if (n_packed) {
        a.tiles = reinterpret_cast<const SsTile *>(d_tiles);
        a.n_tiles = n_packed;
        if (fork && fork->d_ts) a.dbg_ts = fork->d_ts + 4u * (size_t)(fork->ts_n++ % 4096u);
        if (write_back) {   // this kernel also posts the previous batch's results to the host
            a.wb_src = reinterpret_cast<const uint4 *>(write_back->d_out);
            a.wb_dst = reinterpret_cast<uint4 *>(write_back->user_out);
            a.wb_n16 = (uint32_t)(write_back->out_bytes >> 4);
            a.wb_status_src = write_back->d_status;
            a.wb_status_dst = write_back->h_status;
            a.wb_n_status = write_back->n;
        }
        const uint32_t threads = n_packed <= 148u * 8u ? 32u : 128u;  // few tiles: one warp per SM sub-partition
        const uint32_t wpb = threads / 32u;
        const uint32_t grid = (n_packed + wpb - 1) / wpb;
        if (big) {
            if (has_packed) {
                cudaStream_t s = has_thread ? side_stream(fork ? fork->s_aux3 : nullptr, fork ? fork->ev_join3 : nullptr) : stream;
                CU(ctx, launch_pdl(ss_packed_only_kernel, grid, threads, s, false, a));
                ctx->launches++;
                if (s != stream) CU(ctx, cudaEventRecord(fork->ev_join3, s));
            }
            if (has_thread) {   // the caller's stream
                if (late) CU(ctx, launch_pdl(ss_thread_late_kernel, grid, threads, stream, true, a));
                else CU(ctx, launch_pdl(ss_thread_kernel, grid, threads, stream, false, a));
                ctx->launches++;
            }
        } else if (host_tiles && n_packed <= SS_INLINE_TILES && ctx->inline_tiles) {
            static thread_local SsTileBlock tb;
            memcpy(tb.t, host_tiles, (size_t)n_packed * sizeof(SsTile));
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(grid);
            cfg.blockDim = dim3(threads);
            cfg.stream = stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            cfg.numAttrs = ((pdl && !side) || late) ? 1 : 0;
            CU(ctx, cudaLaunchKernelEx(&cfg, ss_packed_inline_kernel, a, tb));
            ctx->launches++;
        } else {
            CU(ctx, launch_pdl(ss_packed_kernel, grid, threads, stream, (pdl && !side) || late, a));
            ctx->launches++;
        }
    }
    for (int j = 0; j < n_joins; j++) CU(ctx, cudaStreamWaitEvent(stream, joins[j], 0));
    if (d_run_nonce) {
        ss_run_commit_kernel<<<(n + 127) / 128, 128, 0, stream>>>(reinterpret_cast<const SsDesc *>(d_desc), n, d_in,
                                                                  d_states, d_status);
        ctx->launches++;
    }
    CU(ctx, cudaGetLastError());
    return 0;
}

extern "C" {

int symrelay_abi_version(void) { return SYMRELAY_ABI_VERSION; }

const char *symrelay_strerror(int code) {
    switch (code) {
        case SYMRELAY_OK: return "ok";
        case SYMRELAY_EINVAL: return "invalid argument";
        case SYMRELAY_EAGAIN: return "ring full, try again";
        case SYMRELAY_ENOMEM: return "out of memory / slots";
        case SYMRELAY_ENOENT: return "no such slot or ticket";
        case SYMRELAY_EPROTO: return "protocol violation by the peer";
        case SYMRELAY_ECUDA: return "CUDA failure";
        default: return "unknown error";
    }
}

const char *symrelay_last_error(symrelay_ctx *ctx) {
    if (ctx) {   // a copy for this thread: another thread's failure may replace the context's string at any time
        std::lock_guard<std::mutex> lk(ctx->err_mu);
        snprintf(g_err, sizeof g_err, "%s", ctx->err.c_str());
    }
    return g_err;
}

uint64_t symrelay_launch_count(symrelay_ctx *ctx) { return ctx ? ctx->launches.load() : 0; }

static int shard_init(symrelay_ctx *ctx, Shard &sh, int dev) {
    sh.dev = dev;
    CU(ctx, cudaSetDevice(dev));
    {   // with `dev` current: the address of the fixed run scratch on this GPU (no allocation involved)
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_run_nonce_static);
        if (e != cudaSuccess || !p)
            return fail(ctx, SYMRELAY_ECUDA, "cudaGetSymbolAddress(run scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
        sh.d_run_static = reinterpret_cast<uint4 *>(p);
    }
    sh.slot_stamp.assign((size_t)ctx->opts.max_streams * 2u, SLOT_CLOSED);
    const size_t table_bytes = (size_t)ctx->opts.max_streams * 2u * 48u;
    CU(ctx, cudaMalloc(&sh.d_states, table_bytes));
    CU(ctx, cudaMemset(sh.d_states, 0, table_bytes));
    CU(ctx, cudaStreamCreateWithFlags(&sh.s_in, cudaStreamNonBlocking));
    CU(ctx, cudaStreamCreateWithFlags(&sh.s_k, cudaStreamNonBlocking));
    CU(ctx, cudaStreamCreateWithFlags(&sh.s_out, cudaStreamNonBlocking));
    {   // the side streams of the long-running tile kinds get the highest priority: their blocks are placed first
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);
        CU(ctx, cudaStreamCreateWithPriority(&sh.s_aux, cudaStreamNonBlocking, hi));
        CU(ctx, cudaStreamCreateWithPriority(&sh.s_aux2, cudaStreamNonBlocking, hi));
    }
    CU(ctx, cudaStreamCreateWithFlags(&sh.s_aux3, cudaStreamNonBlocking));
    CU(ctx, cudaStreamCreateWithFlags(&sh.s_aux4, cudaStreamNonBlocking));
    CU(ctx, cudaStreamCreateWithFlags(&sh.s_init, cudaStreamNonBlocking));
    CU(ctx, cudaHostAlloc(&sh.pk_host, sizeof(PkHost), cudaHostAllocMapped | cudaHostAllocPortable));
    memset(sh.pk_host, 0, sizeof(PkHost));
    sh.pk_host->state = PK_EXITED;
    CU(ctx, cudaMalloc(&sh.pk_dev, sizeof(PkDev)));
    CU(ctx, cudaEventCreateWithFlags(&sh.ev_fork, cudaEventDisableTiming));
    CU(ctx, cudaEventCreateWithFlags(&sh.ev_join, cudaEventDisableTiming));
    CU(ctx, cudaEventCreateWithFlags(&sh.ev_join2, cudaEventDisableTiming));
    CU(ctx, cudaEventCreateWithFlags(&sh.ev_join3, cudaEventDisableTiming));
    CU(ctx, cudaEventCreateWithFlags(&sh.ev_join4, cudaEventDisableTiming));
    {
        std::vector<uint16_t> table(SX_TABLE_ENTRIES);
        sx::build_table(table.data());
        std::vector<uint64_t> fast(SX_FAST_ENTRIES);
        uint8_t cls[256];
        if (!sx::build_fast_tables(fast.data(), cls)) return fail(ctx, SYMRELAY_EINVAL, "extractor grammar has more than 32 byte classes");
        sh.h_sx_fast.assign(SX_TABLES_BYTES, 0);
        memcpy(sh.h_sx_fast.data() + SX_FAST_BLOCK_BYTES, table.data(), SX_TABLE_ENTRIES * sizeof(uint16_t));
        memcpy(sh.h_sx_fast.data(), fast.data(), SX_FAST_ENTRIES * sizeof(uint64_t));
        memcpy(sh.h_sx_fast.data() + SX_FAST_ENTRIES * sizeof(uint64_t), cls, 256);
        CU(ctx, cudaFuncSetAttribute(ss_extract_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SS_EXTRACT_SMEM));
        // Two of its blocks fill 190 KB of an SM; left to itself the runtime then configures 196 KB of shared memory and
        // the seal kernel that is meant to run beside it (SS_ARG_VERDICT_LATE) finds no room for its 8 KB blocks.
        CU(ctx, cudaFuncSetAttribute(ss_extract_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
        CU(ctx, cudaFuncSetAttribute(ss_extract_seal_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SS_FUSED_SMEM));
        CU(ctx, cudaFuncSetAttribute(ss_extract_seal_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
        // The tile kernels of one batch run side by side on the same SMs, and an SM keeps one shared-memory configuration
        // for everything resident on it.  Left to the runtime each kernel asks for its own (0 KB for thread tiles, ~100 KB
        // for seven blocks of bulk tiles ...) and which SMs can take which kernel then depends on what ran before:
        // BASELINE config 3 took 58 us on a fresh context and 63 us right behind config 4.  One common configuration that
        // is large enough for all of them makes every SM eligible for every kind.
        if (!getenv("SYMRELAY_CARVEOUT") || atoi(getenv("SYMRELAY_CARVEOUT")) != 0) {
            const int pct = getenv("SYMRELAY_CARVEOUT") ? atoi(getenv("SYMRELAY_CARVEOUT")) : 40;   // 100 KB: seven blocks of bulk tiles; measured 40/50/75/100 %: 57.8/57.9/58.1/58.9 us
            // (bit per kernel, measurements.  Not the wide kernel: its few blocks are the batch's longest warps, and with a
            // configuration of their own they land on SMs they share with little else -- one eighth of config 3: 25.4 us
            // against 27.2 with it included.  Leaving out any OTHER kernel is what must not happen: 39.6-41.6 us.)
            const int which = getenv("SYMRELAY_CARVEOUT_SET") ? atoi(getenv("SYMRELAY_CARVEOUT_SET")) : 31;
            if (which & 1) CU(ctx, cudaFuncSetAttribute(ss_packed_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
            if (which & 2) CU(ctx, cudaFuncSetAttribute(ss_packed_only_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
            if (which & 4) CU(ctx, cudaFuncSetAttribute(ss_thread_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
            if (which & 8) CU(ctx, cudaFuncSetAttribute(ss_strided_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
            if (which & 16) CU(ctx, cudaFuncSetAttribute(ss_bulk_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
            if (which & 32) CU(ctx, cudaFuncSetAttribute(ss_wide_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
        }
    }
    sh.init_cap = 4096;
    CU(ctx, cudaHostAlloc(&sh.h_init, sizeof(SsInitRec) * sh.init_cap, cudaHostAllocDefault));
    CU(ctx, cudaMalloc(&sh.d_init, sizeof(SsInitRec) * sh.init_cap));
    sh.stages.resize(ctx->opts.pipeline_depth);
    const size_t ctrl_bytes = (size_t)ctx->opts.max_batch_frames * (sizeof(symrelay_desc) + sizeof(symrelay_tile) + 4u) + 64u;
    for (Stage &st : sh.stages) {
        CU(ctx, cudaMalloc(&st.d_ctrl, ctrl_bytes));
        CU(ctx, cudaHostAlloc(&st.h_ctrl, ctrl_bytes, cudaHostAllocMapped));
        CU(ctx, cudaMalloc(&st.d_in, ctx->opts.stage_bytes));
        CU(ctx, cudaMalloc(&st.d_out, ctx->opts.stage_bytes));
        CU(ctx, cudaMalloc(&st.d_status, (size_t)ctx->opts.max_batch_frames * 4u + 16u));   // +16: the drain moves status words 16 bytes at a time
        CU(ctx, cudaHostAlloc(&st.h_status, (size_t)ctx->opts.max_batch_frames * 4u + 16u, cudaHostAllocMapped));
        CU(ctx, cudaMalloc(&st.d_delta, (size_t)ctx->opts.max_batch_frames * sizeof(symrelay_delta)));
        CU(ctx, cudaHostAlloc(&st.h_delta, (size_t)ctx->opts.max_batch_frames * sizeof(symrelay_delta), cudaHostAllocMapped));
        CU(ctx, cudaEventCreateWithFlags(&st.ev_in, cudaEventDisableTiming));
        CU(ctx, cudaEventCreateWithFlags(&st.ev_k, cudaEventDisableTiming));
        CU(ctx, cudaEventCreateWithFlags(&st.ev_out, cudaEventDisableTiming));
    }
    if (getenv("SYMRELAY_TS")) {
        CU(ctx, cudaMalloc(&sh.d_ts, 4096u * 4u * 8u)); CU(ctx, cudaMemset(sh.d_ts, 0, 4096u * 4u * 8u));
        CU(ctx, cudaMalloc(&sh.d_pk_ts, 4096u * PK_TS_WORDS * 8u)); CU(ctx, cudaMemset(sh.d_pk_ts, 0, 4096u * PK_TS_WORDS * 8u));
    }
    CU(ctx, cudaMalloc(&sh.d_sx_fast, SX_TABLES_BYTES));
    CU(ctx, cudaMemcpy(sh.d_sx_fast, sh.h_sx_fast.data(), SX_TABLES_BYTES, cudaMemcpyHostToDevice));
    return 0;
}

static void shard_free(Shard &sh) {
    if (sh.dev < 0) return;
    set_device(sh.dev);
    if (sh.pk_host) sh.pk_host->quit = 1;   // the resident kernel leaves as soon as it is drained
    if (sh.s_k) cudaStreamSynchronize(sh.s_k);
    if (sh.s_init) cudaStreamSynchronize(sh.s_init);
    if (sh.d_ts && sh.ts_n > 64) {   // diagnostics: how consecutive host batches overlap on the GPU
        std::vector<unsigned long long> ts(4096u * 4u);
        cudaMemcpy(ts.data(), sh.d_ts, ts.size() * 8u, cudaMemcpyDeviceToHost);
        const uint64_t last = sh.ts_n - 1, first = last >= 2048 ? last - 2047 : 1;
        double pre = 0, wait = 0, post = 0, period = 0, lead = 0;
        uint64_t cnt = 0;
        for (uint64_t k = first + 1; k <= last; k++) {
            const unsigned long long *c = &ts[4u * (k % 4096u)], *p = &ts[4u * ((k - 1) % 4096u)];
            if (!c[0] || !p[0] || c[3] < c[0]) continue;
            pre += (double)(c[1] - c[0]); wait += (double)(c[2] - c[1]); post += (double)(c[3] - c[2]);
            period += (double)(c[3] - p[3]); lead += (double)((long long)(p[3] - c[0]));
            cnt++;
        }
        if (cnt)
            fprintf(stderr, "symrelay kernel timeline over %llu batches (us, warp 0): reads before the wait %.2f, waiting %.2f, after the wait %.2f; "
                    "end to end %.2f per batch; a kernel starts %.2f before its predecessor ends\n",
                    (unsigned long long)cnt, pre / cnt * 1e-3, wait / cnt * 1e-3, post / cnt * 1e-3, period / cnt * 1e-3, lead / cnt * 1e-3);
    }
    cudaFree(sh.d_ts);
    if (sh.d_pk_ts && sh.pk_posted > 8) {   // diagnostics: where a batch spends its time inside the resident kernel
        std::vector<unsigned long long> ts(4096u * PK_TS_WORDS);
        cudaMemcpy(ts.data(), sh.d_pk_ts, ts.size() * 8u, cudaMemcpyDeviceToHost);
        if (const char *path = getenv("SYMRELAY_TS_DUMP")) {   // raw stamps of the last 4096 batches, for offline analysis
            if (FILE *f = fopen(path, "wb")) {
                const unsigned long long hdr[2] = {sh.pk_posted, PK_TS_WORDS};
                fwrite(hdr, 8, 2, f);
                fwrite(ts.data(), 8, ts.size(), f);
                fclose(f);
            }
        }
        const uint64_t last = sh.pk_posted - 1, first = last >= 2048 ? last - 2047 : 1;
        double d[8] = {0, 0, 0, 0, 0, 0, 0, 0}, m[6] = {0, 0, 0, 0, 0, 0}, mm[4] = {0, 0, 0, 0};
        uint64_t cnt = 0;
        for (uint64_t k = first + 1; k <= last; k++) {
            const unsigned long long *c = &ts[PK_TS_WORDS * (k % 4096u)], *p = &ts[PK_TS_WORDS * ((k - 1) % 4096u)];
            if (!c[0] || !c[1] || !c[7] || !p[7] || c[7] < c[0]) continue;
            d[0] += (double)(c[1] - c[0]);   // doorbell seen -> tile 0 starts
            d[1] += (double)(c[2] - c[1]);   // descriptors and loads issued
            d[2] += (double)(c[3] - c[2]);   // waiting for the previous batch
            d[3] += (double)(c[4] - c[3]);   // ChaCha20/Poly1305 + stores
            d[4] += (double)(c[5] - c[4]);   // device-scope fence + count
            d[5] += (double)(c[7] - c[5]);   // tile 0 signalled -> the whole batch drained to the host
            d[6] += (double)(c[7] - c[0]);   // doorbell seen -> batch's results in host memory
            d[7] += (double)(c[7] - p[7]);   // period
            m[0] += (double)((long long)(c[8] - c[0]));     // doorbell seen -> the batch's last tile starts
            m[1] += (double)((long long)(c[9] - c[8]));     // ... has its loads issued
            m[2] += (double)((long long)(c[10] - p[11]));   // previous batch's last tile done -> this batch's last tile past the dependency
            m[3] += (double)((long long)(c[11] - c[10]));   // ... -> last tile done
            m[4] += (double)((long long)(c[12] - p[12]));   // chain link: last signal to last signal
            m[3] = m[3];
            mm[0] += (double)((long long)(c[10] - p[12]));  // previous batch's last signal -> this batch's last tile past the dependency
            mm[1] += (double)((long long)(c[13] - c[10]));  // -> last tile computed (states written)
            mm[2] += (double)((long long)(c[12] - c[13]));  // -> last tile signalled
            mm[3] += (double)((long long)(c[7] - c[12]));   // -> the batch's results are in host memory (drain warps)
            m[5] += (double)((long long)(c[3] - p[11]));    // previous batch's last tile done -> tile 0 past the dependency
            cnt++;
        }
        if (cnt)
            fprintf(stderr, "symrelay resident-kernel timeline over %llu batches (us, tile 0): doorbell->start %.2f, loads issued %.2f, "
                    "waiting for the previous batch %.2f, compute+stores %.2f, signal %.2f, then until the batch is drained %.2f; doorbell->results "
                    "in host memory %.2f; period %.2f\n", (unsigned long long)cnt, d[0] / cnt * 1e-3, d[1] / cnt * 1e-3, d[2] / cnt * 1e-3,
                    d[3] / cnt * 1e-3, d[4] / cnt * 1e-3, d[5] / cnt * 1e-3, d[6] / cnt * 1e-3, d[7] / cnt * 1e-3);
        if (cnt)
            fprintf(stderr, "  latest tile of a batch: starts %.2f after the doorbell was seen, loads issued %.2f later; previous batch's last "
                    "tile done -> tile 0 past the dependency %.2f, -> last tile past it %.2f, -> last tile done +%.2f; chain link %.2f\n",
                    m[0] / cnt * 1e-3, m[1] / cnt * 1e-3, m[5] / cnt * 1e-3, m[2] / cnt * 1e-3, m[3] / cnt * 1e-3, m[4] / cnt * 1e-3);
        if (cnt)
            fprintf(stderr, "  chain (max over the tiles of a batch): previous batch's last signal -> last tile past the dependency %.2f -> states "
                    "written +%.2f -> signalled +%.2f -> drained to host memory +%.2f\n", mm[0] / cnt * 1e-3, mm[1] / cnt * 1e-3,
                    mm[2] / cnt * 1e-3, mm[3] / cnt * 1e-3);
    }
    cudaFree(sh.d_pk_ts);
    if (sh.s_out) cudaStreamSynchronize(sh.s_out);
    for (Stage &st : sh.stages) {
        cudaFree(st.d_ctrl); cudaFreeHost(st.h_ctrl); cudaFree(st.d_in); cudaFree(st.d_out);
        cudaFree(st.d_status); cudaFreeHost(st.h_status);
        cudaFree(st.d_delta); cudaFreeHost(st.h_delta); cudaFree(st.d_delta_arena); cudaFree(st.d_run_nonce);
        if (st.ev_in) cudaEventDestroy(st.ev_in);
        if (st.ev_k) cudaEventDestroy(st.ev_k);
        if (st.ev_out) cudaEventDestroy(st.ev_out);
    }
    cudaFree(sh.d_states); cudaFreeHost(sh.h_init); cudaFree(sh.d_init); cudaFree(sh.d_sx_fast);
    cudaFree(sh.pk_dev); cudaFreeHost(sh.pk_host);
    if (sh.s_init) cudaStreamDestroy(sh.s_init);
    if (sh.s_in) cudaStreamDestroy(sh.s_in);
    if (sh.s_k) cudaStreamDestroy(sh.s_k);
    if (sh.s_out) cudaStreamDestroy(sh.s_out);
    if (sh.s_aux) cudaStreamDestroy(sh.s_aux);
    if (sh.s_aux2) cudaStreamDestroy(sh.s_aux2);
    if (sh.s_aux3) cudaStreamDestroy(sh.s_aux3);
    if (sh.s_aux4) cudaStreamDestroy(sh.s_aux4);
    if (sh.ev_join2) cudaEventDestroy(sh.ev_join2);
    if (sh.ev_join3) cudaEventDestroy(sh.ev_join3);
    if (sh.ev_join4) cudaEventDestroy(sh.ev_join4);
    if (sh.ev_fork) cudaEventDestroy(sh.ev_fork);
    if (sh.ev_join) cudaEventDestroy(sh.ev_join);
}

symrelay_ctx *symrelay_create(const int *gpu_ids, int n_gpus, const symrelay_opts *opts) {
    if (n_gpus < 1 || n_gpus > 8 || !gpu_ids) {
        fail(nullptr, SYMRELAY_EINVAL, "symrelay_create: n_gpus must be 1..8");
        return nullptr;
    }
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count < 1) {
        fail(nullptr, SYMRELAY_ECUDA, "symrelay_create: no CUDA device (%s); this library has no CPU path",
             e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
        return nullptr;
    }
    symrelay_ctx *ctx = new (std::nothrow) symrelay_ctx();
    if (!ctx) return nullptr;
    if (opts) ctx->opts = *opts;
    if (!ctx->opts.max_streams) ctx->opts.max_streams = 65536;
    if (!ctx->opts.pipeline_depth) ctx->opts.pipeline_depth = 4;
    if (!ctx->opts.stage_bytes) ctx->opts.stage_bytes = 64ull << 20;
    if (!ctx->opts.max_batch_frames) ctx->opts.max_batch_frames = 131072;
    ctx->opts.stage_bytes = align_up(ctx->opts.stage_bytes, 256);
    if (ctx->opts.flags & SYMRELAY_OPT_NO_ZERO_COPY) ctx->zero_copy_max = 0;
    { const char *e = getenv("SYMRELAY_PROF"); ctx->prof = e && e[0] == '1'; }
    { const char *e = getenv("SYMRELAY_DEFER"); if (e && e[0] == '0') ctx->defer_wb = false; }
    { const char *e = getenv("SYMRELAY_PERSIST"); if (e && e[0] == '0') ctx->persist = false; }
    if (const char *m = getenv("SYMRELAY_PERSIST_CTAS_PER_SM")) ctx->pk_ctas_per_sm = (uint32_t)atoi(m);
    if (const char *m = getenv("SYMRELAY_PERSIST_DRAIN_EVERY")) { const int v = atoi(m); if (v >= 2) ctx->pk_drain_every = (uint32_t)v; }
    if (const char *m = getenv("SYMRELAY_PERSIST_IDLE_US")) ctx->pk_idle_ns = strtoull(m, nullptr, 10) * 1000ull;
    { const char *e = getenv("SYMRELAY_BUCKET"); if (e && e[0] == '0') ctx->bucket = false; }
    { const char *e = getenv("SYMRELAY_INLINE_TILES"); if (e && e[0] == '0') ctx->inline_tiles = false; }
    if (const char *m = getenv("SYMRELAY_INJECT_REFUSAL")) ctx->inject_refusal = strtoull(m, nullptr, 10);
    if (const char *m = getenv("SYMRELAY_PDL")) ctx->pdl = atoi(m) != 0;
    if (const char *m = getenv("SYMRELAY_STAGE")) ctx->stage_input = atoi(m) != 0;
    if (const char *m = getenv("SYMRELAY_EXTRACT_OVERLAP")) ctx->extract_overlap = atoi(m) != 0;
    if (const char *m = getenv("SYMRELAY_EXTRACT_FUSED")) ctx->extract_fused = atoi(m) != 0;
    if (const char *m = getenv("SYMRELAY_STAGE_DEV")) ctx->device_arg_flags = atoi(m) ? (SS_ARG_STAGE_INPUT | SS_ARG_STAGE_OUTPUT) : 0u;
    ctx->n_gpus = n_gpus;
    for (int g = 0; g < n_gpus; g++) {
        if (gpu_ids[g] < 0 || gpu_ids[g] >= count) {
            fail(nullptr, SYMRELAY_EINVAL, "symrelay_create: gpu id %d out of range (have %d)", gpu_ids[g], count);
            symrelay_destroy(ctx);
            return nullptr;
        }
        if (shard_init(ctx, ctx->shard[g], gpu_ids[g]) != 0) {
            symrelay_destroy(ctx);
            return nullptr;
        }
    }
    ctx->slot_state.assign(ctx->opts.max_streams, 0);
    ctx->slot_kind.assign(ctx->opts.max_streams, 0);
    for (uint32_t s = ctx->opts.max_streams; s-- > 0;) ctx->free_slots[s % (uint32_t)n_gpus].push_back(s);
    return ctx;
}

static void sched_free(symrelay_ctx *ctx, int g);

void symrelay_destroy(symrelay_ctx *ctx) {
    if (!ctx) return;
    if (ctx->watcher.joinable()) {
        ctx->watcher_stop = true;
        { std::lock_guard<std::mutex> lk(ctx->watch_mu); ctx->watch_cv.notify_all(); }
        ctx->watcher.join();
    }
    if (ctx->evfd >= 0) close(ctx->evfd);
    if (ctx->prof && ctx->prof_n) {
        const char *names[8] = {"recycle", "plan", "checks", "ctrl copy", "launch", "event record", "wait", ""};
        fprintf(stderr, "symrelay host batch profile over %llu batches (us per batch):", (unsigned long long)ctx->prof_n);
        for (int k = 0; k < 7; k++) fprintf(stderr, " %s %.2f", names[k], ctx->prof_t[k] / (double)ctx->prof_n);
        fprintf(stderr, "\n");
    }
    for (int g = 0; g < ctx->n_gpus; g++) { shard_free(ctx->shard[g]); sched_free(ctx, g); }
    delete ctx;
}

// ---- streams ------------------------------------------------------------------
static int finish_stage(symrelay_ctx *ctx, Shard &sh, Stage &st);
static int sched_slot_busy(symrelay_ctx *ctx, uint32_t slot);     // scheduler_impl.inc
static void sched_slot_reset(symrelay_ctx *ctx, uint32_t slot);
// A slot changes hands: its states are rewritten by the init kernel (stamp 0), so nobody may wait for an old stamp.
static void tmpl_flush_stamps(Shard &sh);
static void stamp_reset(symrelay_ctx *ctx, uint32_t slot, bool open) {
    Shard &sh = ctx->shard[slot % (uint32_t)ctx->n_gpus];
    tmpl_flush_stamps(sh);   // first: a late flush must not write over what is set here
    sh.tmpl_streak = false;
    sh.slot_gen++;           // a cached template's "all slots open" verdict no longer holds
    sh.slot_stamp[(size_t)slot * 2u] = sh.slot_stamp[(size_t)slot * 2u + 1u] = open ? 0u : SLOT_CLOSED;
}
static int push_init_recs(symrelay_ctx *ctx, Shard &sh, uint32_t n) {
    CU(ctx, set_device(sh.dev));
    // On its own stream: s_k may be occupied by the resident kernel for as long as batches keep coming, and a new
    // or closing connection touches no slot that a batch in flight names.
    CU(ctx, cudaMemcpyAsync(sh.d_init, sh.h_init, sizeof(SsInitRec) * n, cudaMemcpyHostToDevice, sh.s_init));
    ss_init_kernel<<<(n + 127) / 128, 128, 0, sh.s_init>>>(sh.d_init, n, sh.d_states);
    ctx->launches++;
    CU(ctx, cudaGetLastError());
    CU(ctx, cudaStreamSynchronize(sh.s_init));  // h_init is reused by the next call; the states are visible to later batches
    return 0;
}

static void fill_rec(SsInitRec *r, uint32_t slot, uint32_t dir, const uint8_t *key, const uint8_t *init, int init_len,
                     uint32_t meta) {
    memset(r, 0, sizeof *r);
    r->slot_dir = slot * 2u + dir;
    r->init_len = key ? (uint32_t)init_len : 0u;
    r->meta = meta;
    if (key) {
        memcpy(r->key, key, 32);
        memcpy(r->init, init, (size_t)init_len);
    }
}

int symrelay_stream_open_batch(symrelay_ctx *ctx, uint32_t n, const uint8_t *tx_keys, const uint8_t *tx_inits,
                               const uint8_t *rx_keys, const uint8_t *rx_inits, int init_len, int provider_kind,
                               uint32_t *slots_out) {
    if (!ctx || !slots_out) return SYMRELAY_EINVAL;
    if (init_len != 24 && init_len != 52) return fail(ctx, SYMRELAY_EINVAL, "init_len must be 24 or 52");
    if ((tx_keys && !tx_inits) || (rx_keys && !rx_inits)) return fail(ctx, SYMRELAY_EINVAL, "key without init");
    std::lock_guard<std::mutex> lk(ctx->mu);
    // allocate slots round-robin over the GPUs
    uint32_t avail = 0;
    for (int g = 0; g < ctx->n_gpus; g++) avail += (uint32_t)ctx->free_slots[g].size();
    if (avail < n) return fail(ctx, SYMRELAY_ENOMEM, "slot table full (%u free, %u wanted)", avail, n);
    int g = 0;
    for (uint32_t i = 0; i < n; i++) {
        while (ctx->free_slots[g].empty()) g = (g + 1) % ctx->n_gpus;
        slots_out[i] = ctx->free_slots[g].back();
        ctx->free_slots[g].pop_back();
        ctx->slot_state[slots_out[i]] = 1;
        ctx->slot_kind[slots_out[i]] = (uint8_t)(provider_kind < 0 || provider_kind > 2 ? 0 : provider_kind);
        g = (g + 1) % ctx->n_gpus;
    }
    for (int gi = 0; gi < ctx->n_gpus; gi++) {
        Shard &sh = ctx->shard[gi];
        std::lock_guard<std::mutex> slk(sh.mu);
        uint32_t fill = 0;
        for (uint32_t i = 0; i < n; i++) {
            if ((int)(slots_out[i] % (uint32_t)ctx->n_gpus) != gi) continue;
            sched_slot_reset(ctx, slots_out[i]);
            stamp_reset(ctx, slots_out[i], true);
            fill_rec(&sh.h_init[fill++], slots_out[i], 0, tx_keys ? tx_keys + 32ull * i : nullptr,
                     tx_inits ? tx_inits + (size_t)init_len * i : nullptr, init_len, (uint32_t)provider_kind);
            fill_rec(&sh.h_init[fill++], slots_out[i], 1, rx_keys ? rx_keys + 32ull * i : nullptr,
                     rx_inits ? rx_inits + (size_t)init_len * i : nullptr, init_len, (uint32_t)provider_kind);
            if (fill + 2 > sh.init_cap) {
                int rc = push_init_recs(ctx, sh, fill);
                if (rc) return rc;
                fill = 0;
            }
        }
        if (fill) {
            int rc = push_init_recs(ctx, sh, fill);
            if (rc) return rc;
        }
    }
    return 0;
}

int symrelay_stream_open(symrelay_ctx *ctx, const uint8_t *tx_key, const uint8_t *tx_init, int tx_init_len,
                         const uint8_t *rx_key, const uint8_t *rx_init, int rx_init_len, int provider_kind,
                         uint32_t *slot_out) {
    if (!ctx || !slot_out) return SYMRELAY_EINVAL;
    if (tx_key && tx_init_len != 24 && tx_init_len != 52) return fail(ctx, SYMRELAY_EINVAL, "tx_init_len must be 24 or 52");
    if (rx_key && rx_init_len != 24 && rx_init_len != 52) return fail(ctx, SYMRELAY_EINVAL, "rx_init_len must be 24 or 52");
    if ((tx_key && !tx_init) || (rx_key && !rx_init)) return fail(ctx, SYMRELAY_EINVAL, "key without init");
    std::lock_guard<std::mutex> lk(ctx->mu);
    int g = -1;
    size_t best = 0;
    for (int i = 0; i < ctx->n_gpus; i++)
        if (ctx->free_slots[i].size() > best) { best = ctx->free_slots[i].size(); g = i; }
    if (g < 0) return fail(ctx, SYMRELAY_ENOMEM, "slot table full");
    const uint32_t slot = ctx->free_slots[g].back();
    ctx->free_slots[g].pop_back();
    ctx->slot_state[slot] = 1;
    ctx->slot_kind[slot] = (uint8_t)(provider_kind < 0 || provider_kind > 2 ? 0 : provider_kind);
    Shard &sh = ctx->shard[g];
    std::lock_guard<std::mutex> slk(sh.mu);
    sched_slot_reset(ctx, slot);
    stamp_reset(ctx, slot, true);
    fill_rec(&sh.h_init[0], slot, 0, tx_key, tx_init, tx_init_len, (uint32_t)provider_kind);
    fill_rec(&sh.h_init[1], slot, 1, rx_key, rx_init, rx_init_len, (uint32_t)provider_kind);
    int rc = push_init_recs(ctx, sh, 2);
    if (rc) return rc;
    *slot_out = slot;
    return 0;
}

int symrelay_stream_close(symrelay_ctx *ctx, uint32_t slot) {
    if (!ctx) return SYMRELAY_EINVAL;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (slot >= ctx->opts.max_streams || !ctx->slot_state[slot]) return fail(ctx, SYMRELAY_ENOENT, "slot %u not open", slot);
    Shard &sh = ctx->shard[slot % (uint32_t)ctx->n_gpus];
    std::lock_guard<std::mutex> slk(sh.mu);
    if (sched_slot_busy(ctx, slot))
        return fail(ctx, SYMRELAY_EAGAIN, "slot %u still has frames in a batch that is filling or in flight: flush, poll, then close", slot);
    fill_rec(&sh.h_init[0], slot, 0, nullptr, nullptr, 0, 0);
    fill_rec(&sh.h_init[1], slot, 1, nullptr, nullptr, 0, 0);
    int rc = push_init_recs(ctx, sh, 2);  // wipes the key material on the device
    if (rc) return rc;
    sched_slot_reset(ctx, slot);   // a half-received frame of the old connection must not eat the next one's bytes
    stamp_reset(ctx, slot, false);
    ctx->slot_state[slot] = 0;
    ctx->free_slots[slot % (uint32_t)ctx->n_gpus].push_back(slot);
    return 0;
}

int symrelay_stream_export(symrelay_ctx *ctx, uint32_t slot, uint8_t tx_state[SYMRELAY_STATEBYTES],
                           uint8_t rx_state[SYMRELAY_STATEBYTES]) {
    if (!ctx) return SYMRELAY_EINVAL;
    if (slot >= ctx->opts.max_streams || !ctx->slot_state[slot]) return fail(ctx, SYMRELAY_ENOENT, "slot %u not open", slot);
    Shard &sh = ctx->shard[slot % (uint32_t)ctx->n_gpus];
    std::lock_guard<std::mutex> lk(sh.mu);
    CU(ctx, set_device(sh.dev));
    for (Stage &st : sh.stages) {   // "live" means after every host batch submitted so far
        const int rc = finish_stage(ctx, sh, st);
        if (rc) return rc;
    }
    uint8_t buf[96];
    CU(ctx, cudaMemcpyAsync(buf, sh.d_states + (size_t)slot * 6u, 96, cudaMemcpyDeviceToHost, sh.s_init));
    CU(ctx, cudaStreamSynchronize(sh.s_init));
    if (tx_state) { memcpy(tx_state, buf, 44); memset(tx_state + 44, 0, 8); }
    if (rx_state) { memcpy(rx_state, buf + 48, 44); memset(rx_state + 44, 0, 8); }
    return 0;
}

// ---- connection set-up in batch -------------------------------------------------------------------
int symrelay_connect_batch(symrelay_ctx *ctx, int gpu_index, uint32_t n, const uint8_t *handshake_hashes, const uint8_t *initiator,
                           const uint8_t ns_initiator[32], const uint8_t ns_responder[32], const uint8_t *seed32,
                           uint8_t *local_ids, uint8_t *remote_ids, uint8_t *headers, uint8_t *first_messages) {
    if (!ctx || gpu_index < 0 || gpu_index >= ctx->n_gpus) return SYMRELAY_EINVAL;
    if (n == 0) return 0;
    if (!handshake_hashes || !initiator || !ns_initiator || !ns_responder)
        return fail(ctx, SYMRELAY_EINVAL, "symrelay_connect_batch: null argument");
    Shard &sh = ctx->shard[gpu_index];
    std::lock_guard<std::mutex> lk(sh.mu);
    CU(ctx, set_device(sh.dev));
    uint8_t seed[32];
    if (seed32) memcpy(seed, seed32, 32);
    else {   // 32 fresh random bytes from the kernel's generator
        size_t got = 0;
        if (FILE *f = fopen("/dev/urandom", "rb")) { got = fread(seed, 1, 32, f); fclose(f); }
        if (got != 32) return fail(ctx, SYMRELAY_EINVAL, "symrelay_connect_batch: no entropy source for the headers");
    }
    const uint32_t chunk = 16384;
    uint8_t *d_buf = nullptr;   // recs ‖ outs ‖ ns pair ‖ seed (set-up path: allocation here is fine)
    const size_t recs_b = (size_t)chunk * sizeof(SsConnRec), outs_b = (size_t)chunk * sizeof(SsConnOut);
    CU(ctx, cudaMalloc(&d_buf, recs_b + outs_b + 96));
    std::vector<SsConnRec> recs(chunk);
    std::vector<SsConnOut> outs(chunk);
    uint8_t nsseed[96];
    memcpy(nsseed, ns_initiator, 32); memcpy(nsseed + 32, ns_responder, 32); memcpy(nsseed + 64, seed, 32);
    cudaError_t e = cudaMemcpyAsync(d_buf + recs_b + outs_b, nsseed, 96, cudaMemcpyHostToDevice, sh.s_init);
    for (uint32_t at = 0; at < n && e == cudaSuccess; at += chunk) {
        const uint32_t m = n - at < chunk ? n - at : chunk;
        for (uint32_t i = 0; i < m; i++) {
            memcpy(recs[i].hash, handshake_hashes + 64ull * (at + i), 64);
            recs[i].initiator = initiator[at + i] ? 1u : 0u;
        }
        e = cudaMemcpyAsync(d_buf, recs.data(), (size_t)m * sizeof(SsConnRec), cudaMemcpyHostToDevice, sh.s_init);
        if (e != cudaSuccess) break;
        ss_connect_kernel<<<(m + 127) / 128, 128, 0, sh.s_init>>>(reinterpret_cast<const SsConnRec *>(d_buf), m,
                                                                  reinterpret_cast<const uint4 *>(d_buf + recs_b + outs_b),
                                                                  reinterpret_cast<const uint4 *>(d_buf + recs_b + outs_b + 64),
                                                                  (uint32_t)(at / chunk), reinterpret_cast<SsConnOut *>(d_buf + recs_b));
        ctx->launches++;
        e = cudaMemcpyAsync(outs.data(), d_buf + recs_b, (size_t)m * sizeof(SsConnOut), cudaMemcpyDeviceToHost, sh.s_init);
        if (e == cudaSuccess) e = cudaStreamSynchronize(sh.s_init);
        if (e != cudaSuccess) break;
        for (uint32_t i = 0; i < m; i++) {
            const size_t k = at + i;
            if (local_ids) memcpy(local_ids + 32 * k, outs[i].id_local, 32);
            if (remote_ids) memcpy(remote_ids + 32 * k, outs[i].id_remote, 32);
            if (headers) memcpy(headers + 24 * k, outs[i].header, 24);
            if (first_messages) {   // uint24-LE(56) ‖ streamId ‖ header: secret-stream's first frame, sent in the clear
                uint8_t *fm = first_messages + 59 * k;
                fm[0] = 56; fm[1] = 0; fm[2] = 0;
                memcpy(fm + 3, outs[i].id_local, 32);
                memcpy(fm + 35, outs[i].header, 24);
            }
        }
    }
    cudaFree(d_buf);
    if (e != cudaSuccess) return fail(ctx, SYMRELAY_ECUDA, "symrelay_connect_batch: %s", cudaGetErrorString(e));
    return 0;
}

// ---- planning -------------------------------------------------------------------
int64_t symrelay_plan(const symrelay_desc *desc, uint32_t n, symrelay_tile *tiles, uint32_t max_tiles,
                      uint32_t *n_bulk_out) {
    if (!desc && n) return SYMRELAY_EINVAL;
    return symrelay_plan_inline(desc, n, tiles, max_tiles, n_bulk_out);
}

int64_t symrelay_plan_sorted(const symrelay_desc *desc, uint32_t n, symrelay_tile *tiles, uint32_t max_tiles,
                             uint32_t *n_bulk_out, symrelay_desc *sorted_out, uint32_t *perm_out) {
    if (!desc && n) return SYMRELAY_EINVAL;
    if (tiles && (!sorted_out || !perm_out)) return SYMRELAY_EINVAL;
    const int64_t r = symrelay_plan_bucketed(desc, n, tiles, max_tiles, n_bulk_out, nullptr, sorted_out, perm_out);
    return r == -2 ? (int64_t)SYMRELAY_EINVAL : r;
}

int symrelay_desc_check(const symrelay_desc *d) { return d ? symrelay_desc_check_inline(d) : SYMRELAY_EINVAL; }

// ---- batches --------------------------------------------------------------------
int symrelay_batch_device(symrelay_ctx *ctx, int gpu_index, const symrelay_desc *d_desc, uint32_t n,
                          const symrelay_tile *d_tiles, uint32_t n_tiles, uint32_t n_bulk_tiles,
                          const uint8_t *d_in, uint8_t *d_out, uint32_t *d_status, symrelay_delta *d_delta,
                          uint8_t *d_delta_arena, void *cuda_stream) {
    if (!ctx || gpu_index < 0 || gpu_index >= ctx->n_gpus) return SYMRELAY_EINVAL;
    if (n == 0 || n_tiles == 0) return 0;
    if (!d_desc || !d_tiles || !d_in || !d_out || !d_status || n_bulk_tiles > n_tiles)
        return fail(ctx, SYMRELAY_EINVAL, "symrelay_batch_device: null pointer or bad tile counts");
    Shard &sh = ctx->shard[gpu_index];
    CU(ctx, set_device(sh.dev));
    return launch_tiles(ctx, d_desc, n, d_tiles, n_tiles, n_bulk_tiles, d_in, d_out, sh.d_states, d_status, d_delta,
                        d_delta_arena, (cudaStream_t)cuda_stream, ctx->device_arg_flags, nullptr, &sh);
}

int symrelay_batch_device_runs(symrelay_ctx *ctx, int gpu_index, const symrelay_desc *d_desc, uint32_t n,
                               const symrelay_tile *d_tiles, uint32_t n_tiles, uint32_t n_bulk_tiles,
                               const uint8_t *d_in, uint8_t *d_out, uint32_t *d_status, void *d_scratch,
                               uint64_t scratch_bytes, void *cuda_stream) {
    if (!ctx || gpu_index < 0 || gpu_index >= ctx->n_gpus) return SYMRELAY_EINVAL;
    if (n == 0 || n_tiles == 0) return 0;
    if (!d_desc || !d_tiles || !d_in || !d_out || !d_status || !d_scratch || n_bulk_tiles > n_tiles)
        return fail(ctx, SYMRELAY_EINVAL, "symrelay_batch_device_runs: null pointer or bad tile counts");
    if (scratch_bytes < (uint64_t)n * 16u || (reinterpret_cast<uintptr_t>(d_scratch) & 15u))
        return fail(ctx, SYMRELAY_EINVAL, "symrelay_batch_device_runs: scratch must be 16-byte aligned and hold 16 bytes per descriptor");
    Shard &sh = ctx->shard[gpu_index];
    CU(ctx, set_device(sh.dev));
    return launch_tiles(ctx, d_desc, n, d_tiles, n_tiles, n_bulk_tiles, d_in, d_out, sh.d_states, d_status, nullptr,
                        nullptr, (cudaStream_t)cuda_stream, 0u, reinterpret_cast<uint4 *>(d_scratch));
}

int symrelay_batch_device_runs_static(symrelay_ctx *ctx, int gpu_index, const symrelay_desc *d_desc, uint32_t n,
                                      const symrelay_tile *d_tiles, uint32_t n_tiles, uint32_t n_bulk_tiles,
                                      const uint8_t *d_in, uint8_t *d_out, uint32_t *d_status, void *cuda_stream) {
    if (!ctx || gpu_index < 0 || gpu_index >= ctx->n_gpus) return SYMRELAY_EINVAL;
    if (n == 0 || n_tiles == 0) return 0;
    if (n > SYMRELAY_STATIC_RUN_MAX)
        return fail(ctx, SYMRELAY_EINVAL,
                    "symrelay_batch_device_runs_static: %u descriptors exceed the built-in scratch of %u; split the "
                    "batch or use symrelay_batch_device_runs with your own scratch", n, SYMRELAY_STATIC_RUN_MAX);
    if (!d_desc || !d_tiles || !d_in || !d_out || !d_status || n_bulk_tiles > n_tiles)
        return fail(ctx, SYMRELAY_EINVAL, "symrelay_batch_device_runs_static: null pointer or bad tile counts");
    Shard &sh = ctx->shard[gpu_index];
    CU(ctx, set_device(sh.dev));
    uint4 *scratch = sh.d_run_static;   // resolved by symrelay_create with this GPU current
    // Nothing in the array has to start at zero: the nonce pre-pass of this call writes every entry that
    // this call's tile kernels read (entries of non-run descriptors keep stale values and are never read).
    return launch_tiles(ctx, d_desc, n, d_tiles, n_tiles, n_bulk_tiles, d_in, d_out, sh.d_states, d_status, nullptr,
                        nullptr, (cudaStream_t)cuda_stream, 0u, scratch);
}

// ---- persistent kernel: host side ------------------------------------------------------------
// Ordinary launches are about to go to s_k: the resident kernel must leave first (stream order then keeps every
// batch in submission order).  No waiting here -- the launches simply queue behind it.
static inline uint32_t pk_version(uint32_t stamp) { return (stamp & 0x7fffffu) | 0x800000u; }   // 24 bits, never 0, repeats every 8 M batches

// Waits until every batch handed to the resident kernel has finished (rare paths only).
static int pk_drain_host(symrelay_ctx *ctx, Shard &sh) {
    for (Stage &st : sh.stages)
        if (st.busy && st.pk_seq >= 0) {
            const int rc = finish_stage(ctx, sh, st);
            if (rc) return rc;
        }
    return 0;
}

// The fast path does not touch the stamp table; before anyone reads it again, give the template's streams the
// stamp of the last batch that went through them.
static void tmpl_flush_stamps(Shard &sh) {
    if (!sh.tmpl_lazy) return;
    uint32_t *stamp = sh.slot_stamp.data();
    const uint32_t now = sh.stamp;
    for (const symrelay_desc &d : sh.tmpl_desc) {
        uint32_t &e = stamp[(size_t)d.slot * 2u + ((d.len_flags >> 26) & 1u)];
        if (e != SLOT_CLOSED) e = now;
    }
    sh.tmpl_lazy = false;
}

static void pk_quiesce(Shard &sh) {
    if (!sh.pk_host || sh.pk_quit_req) return;
    if (sh.pk_host->state == PK_EXITED && sh.pk_posted == sh.pk_host->exit_installed) return;   // not running, nothing pending
    sh.pk_host->quit = 1;
    sh.pk_quit_req = true;
}

static int pk_launch(symrelay_ctx *ctx, Shard &sh, uint64_t first_seq) {
    sh.pk_host->state = PK_RUNNING;
    std::atomic_thread_fence(std::memory_order_seq_cst);
    CU(ctx, cudaMemsetAsync(sh.pk_dev, 0, sizeof(PkDev), sh.s_k));
    // Resident CTAs per SM: a warp is held from its tile's first PCIe read until its results have been fenced to the
    // host (~15 us), of which ~4 us compute; with one CTA per SM the batches behind the one that computes cannot all
    // prefetch.  Never more CTAs than fit at once: every warp of the grid must be able to run (they wait on each other).
    if (!sh.pk_grid) {
        int sms = 148, per_sm = 1;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, sh.dev);
        CU(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ss_persistent_kernel, 128, 0));
        if (per_sm > (int)ctx->pk_ctas_per_sm) per_sm = (int)ctx->pk_ctas_per_sm;
        if (per_sm < 1) per_sm = 1;
        sms *= per_sm;
        if (sms > (int)PK_MAX_CTAS) sms = (int)PK_MAX_CTAS;
        sh.pk_grid = sms;
    }
    const int sms = sh.pk_grid;
    ss_persistent_kernel<<<(unsigned)sms, 128, 0, sh.s_k>>>(sh.pk_host, sh.pk_dev, sh.d_states, (unsigned long long)first_seq,
                                                            (unsigned long long)ctx->pk_idle_ns, sh.d_pk_ts, ctx->pk_drain_every);
    ctx->launches++;
    ctx->pk_launches++;
    CU(ctx, cudaGetLastError());
    return 0;
}

// Spin until pred() holds; every few microseconds make sure the kernel has not died under us.
extern "C++" {
template <typename P>
static int pk_spin(symrelay_ctx *ctx, Shard &sh, P pred, const char *what) {
    for (uint32_t spins = 1;; spins++) {
        if (pred()) return 0;
        __builtin_ia32_pause();
        if ((spins & 0x3fffu) == 0) {
            const cudaError_t e = cudaStreamQuery(sh.s_k);
            if (e != cudaSuccess && e != cudaErrorNotReady)
                return fail(ctx, SYMRELAY_ECUDA, "resident relay kernel failed while %s: %s", what, cudaGetErrorString(e));
            if (e == cudaSuccess && !pred()) {
                // nothing is running on s_k although work is pending: the kernel left (idle exit raced this batch
                // in a way the hand-shake did not cover, or it was never started) -- start it on what is pending
                std::atomic_thread_fence(std::memory_order_seq_cst);
                if (pred()) return 0;
                if (sh.pk_host->state == PK_EXITED && sh.pk_host->exit_installed < sh.pk_posted) {
                    const int rc = pk_launch(ctx, sh, sh.pk_host->exit_installed);
                    if (rc) return rc;
                }
            }
        }
    }
}
}  // extern "C++"

// Hands one planned batch to the resident kernel.  st.h_ctrl holds tiles ‖ descs already.
static int pk_post(symrelay_ctx *ctx, Shard &sh, Stage &st, const uint8_t *ctrl, size_t tiles_bytes, uint32_t n, uint32_t n_tiles,
                   const uint8_t *in, uint8_t *out, uint64_t out_bytes, uint32_t flags, uint32_t uni_per, uint32_t uni_thread,
                   const uint32_t *ver, uint32_t new_ver, uint32_t ver_all = 0, const SsDesc *desc_dev = nullptr, SsDesc *desc_fill = nullptr) {
    PkHost *H = sh.pk_host;
    if (sh.pk_quit_req) {
        // Launches were queued behind the kernel since; a batch posted now would overtake them.  Wait for the kernel
        // to be gone (it leaves as soon as it is drained), then the relaunch below queues behind those launches.
        const int rc = pk_spin(ctx, sh, [&] { return H->state == PK_EXITED; }, "waiting for it to exit");
        if (rc) return rc;
        sh.pk_quit_req = false;
        H->quit = 0;
    }
    const uint64_t seq = sh.pk_posted;
    PkSlot &ps = H->slot[seq % PK_RING];
    ps.desc = desc_dev ? desc_dev : reinterpret_cast<const SsDesc *>(ctrl + tiles_bytes);
    ps.desc_fill = desc_fill;
    ps.tiles = reinterpret_cast<const SsTile *>(ctrl);
    ps.in = in; ps.out = st.d_out; ps.status = st.d_status;   // results: device staging first, the drain warps move them
    ps.n_desc = n; ps.n_tiles = n_tiles;
    ps.flags = flags; ps.uni_per = uni_per; ps.uni_thread = uni_thread; ps.out_n16 = (uint32_t)(out_bytes >> 4);
    ps.host_out = out; ps.host_status = st.h_status;
    ps.ver = ver; ps.new_ver = new_ver; ps.ver_all = ver_all;
    _mm_sfence();   // header, control block (streaming stores) and the caller's payload before the doorbell
    std::atomic_thread_fence(std::memory_order_release);
    H->doorbell = seq + 1;
    sh.pk_posted = seq + 1;
    st.pk_seq = (int64_t)seq;
    ctx->pk_batches++;
    std::atomic_thread_fence(std::memory_order_seq_cst);   // doorbell store before the state load (see pk_poller)
    uint32_t state = H->state;
    if (state == PK_EXITING) {
        const int rc = pk_spin(ctx, sh, [&] { return H->state != PK_EXITING; }, "settling its exit");
        if (rc) return rc;
        state = H->state;
    }
    if (state == PK_EXITED) return pk_launch(ctx, sh, H->exit_installed);
    return 0;
}

static inline bool pk_batch_done(const Shard &sh, int64_t seq) {
    return sh.pk_host->done[(uint64_t)seq % PK_RING].v >= (unsigned long long)seq + 1ull;
}

// Has the batch in this stage finished (results in host memory)?  Never blocks.
static bool stage_ready(const Shard &sh, const Stage &st) {
    if (!st.busy) return true;
    if (st.pk_seq >= 0) return pk_batch_done(sh, st.pk_seq);
    if (st.deferred && st.wb_by < 0) return false;   // its results still sit in device staging
    return cudaEventQuery(st.deferred ? sh.stages[(size_t)st.wb_by].ev_out : st.ev_out) == cudaSuccess;
}

// The results of stage `idx` still sit in device staging and nothing that follows will carry them.
static int flush_write_back(symrelay_ctx *ctx, Shard &sh, int idx) {
    Stage &p = sh.stages[(size_t)idx];
    pk_quiesce(sh);
    SsArgs a;
    memset(&a, 0, sizeof a);
    a.wb_src = reinterpret_cast<const uint4 *>(p.d_out);
    a.wb_dst = reinterpret_cast<uint4 *>(p.user_out);
    a.wb_n16 = (uint32_t)(p.out_bytes >> 4);
    a.wb_status_src = p.d_status;
    a.wb_status_dst = p.h_status;
    a.wb_n_status = p.n;
    const uint32_t blocks = (a.wb_n16 + 127u) / 128u;
    ss_write_back_kernel<<<blocks < 1u ? 1u : (blocks > 592u ? 592u : blocks), 128, 0, sh.s_k>>>(a);
    ctx->launches++;
    CU(ctx, cudaGetLastError());
    CU(ctx, cudaEventRecord(p.ev_out, sh.s_k));
    p.wb_by = idx;
    sh.pending_wb = -1;
    return 0;
}

static int finish_stage(symrelay_ctx *ctx, Shard &sh, Stage &st) {
    if (!st.busy) return 0;
    CU(ctx, set_device(sh.dev));
    if (st.pk_seq >= 0) {
        const int64_t seq = st.pk_seq;
        const int rc = pk_spin(ctx, sh, [&] { return pk_batch_done(sh, seq); }, "waiting for a batch");
        if (rc) return rc;
        std::atomic_thread_fence(std::memory_order_acquire);
        st.pk_seq = -1;
    } else
    if (st.deferred) {
        const int idx = (int)(&st - sh.stages.data());
        if (sh.pending_wb == idx) {
            const int rc = flush_write_back(ctx, sh, idx);
            if (rc) return rc;
        }
        // the carrier's event; if that stage has been reused since, its event is later in the stream, which is still enough
        CU(ctx, cudaEventSynchronize(sh.stages[(size_t)st.wb_by].ev_out));
        st.deferred = false;
    } else
    CU(ctx, cudaEventSynchronize(st.ev_out));
    if (st.permuted) {
        const uint32_t *perm = st.perm.data();
        if (st.user_status) for (uint32_t k = 0; k < st.n; k++) st.user_status[perm[k]] = st.h_status[k];
        if (st.user_delta) for (uint32_t k = 0; k < st.n; k++) st.user_delta[perm[k]] = st.h_delta[k];
        st.permuted = false;
    } else {
        if (st.user_status) memcpy(st.user_status, st.h_status, (size_t)st.n * 4u);
        if (st.user_delta) memcpy(st.user_delta, st.h_delta, (size_t)st.n * sizeof(symrelay_delta));
    }
    st.busy = false;
    return 0;
}

static int64_t submit_locked(symrelay_ctx *ctx, int gpu_index, const symrelay_desc *desc, uint32_t n,
                             const uint8_t *in, uint64_t in_bytes, uint8_t *out, uint64_t out_bytes,
                             uint32_t *status, symrelay_delta *delta, uint8_t *delta_arena,
                             uint64_t delta_arena_bytes, bool may_defer = false) {
    Shard &sh = ctx->shard[gpu_index];
    if (n > ctx->opts.max_batch_frames) return fail(ctx, SYMRELAY_EINVAL, "batch of %u frames exceeds max_batch_frames %u", n, ctx->opts.max_batch_frames);
    if (in_bytes > ctx->opts.stage_bytes || out_bytes > ctx->opts.stage_bytes)
        return fail(ctx, SYMRELAY_EINVAL, "arena (%llu in / %llu out bytes) exceeds stage_bytes %llu",
                    (unsigned long long)in_bytes, (unsigned long long)out_bytes, (unsigned long long)ctx->opts.stage_bytes);
    if ((in_bytes & 15u) || (out_bytes & 15u)) return fail(ctx, SYMRELAY_EINVAL, "arena sizes must be multiples of 16");
    double prof_t0 = ctx->prof ? prof_now() : 0.0;
    CU(ctx, set_device(sh.dev));
    const int64_t seq = sh.next_ticket_seq++;
    Stage &st = sh.stages[(size_t)(seq % (int64_t)sh.stages.size())];
    int rc = finish_stage(ctx, sh, st);  // recycle: the batch that used this stage P submissions ago
    if (rc) return rc;
    PROF_MARK(ctx, 0);

    // Fast path: the batch repeats the template (see Shard) and directly follows a batch of it.
    if (sh.tmpl_streak && n == sh.tmpl_n && in_bytes == sh.tmpl_in_bytes && out_bytes == sh.tmpl_out_bytes &&
        sh.slot_gen == sh.tmpl_slot_gen && !sh.stamp_dirty && sh.stamp < 0xfffffd00u &&
        memcmp(desc, sh.tmpl_desc.data(), (size_t)n * sizeof(symrelay_desc)) == 0 &&
        is_library_pinned(ctx, in, in_bytes) && is_library_pinned(ctx, out, out_bytes)) {
        PROF_MARK(ctx, 1);
        if (st.tmpl_gen != sh.tmpl_gen) {   // this stage's control block holds something else: (re)write tiles ‖ descriptors once
            if (!sh.tmpl_uni_per) memcpy(st.h_ctrl, sh.tmpl_tiles.data(), (size_t)sh.tmpl_n_tiles * sizeof(symrelay_tile));
            memcpy(st.h_ctrl + sh.tmpl_tiles_bytes, desc, (size_t)n * sizeof(symrelay_desc));
            st.tmpl_gen = sh.tmpl_gen;
        }
        PROF_MARK(ctx, 3);
        const uint32_t prev = sh.stamp++;   // the stamp table itself is brought up to date when the streak ends (tmpl_flush_stamps)
        sh.tmpl_lazy = true;
        // Descriptors: the first batch of a template through this stage leaves a copy in the stage's device control
        // block (every tile stores the 160 bytes it has just read); later ones read that copy instead of the host's.
        SsDesc *dev_descs = reinterpret_cast<SsDesc *>(st.d_ctrl);
        const bool cached = st.dev_tmpl_gen == sh.tmpl_gen;
        st.dev_tmpl_gen = sh.tmpl_gen;
        rc = pk_post(ctx, sh, st, st.h_ctrl, sh.tmpl_tiles_bytes, n, sh.tmpl_n_tiles, in, out, out_bytes,
                     sh.tmpl_flags | SS_ARG_CHAIN_UNIFORM, sh.tmpl_uni_per, sh.tmpl_uni_thread, nullptr, pk_version(sh.stamp),
                     pk_version(prev), cached ? dev_descs : nullptr, cached ? nullptr : dev_descs);
        if (rc) { sh.stamp_dirty = true; sh.tmpl_streak = false; return rc; }
        PROF_MARK(ctx, 5);
        st.deferred = false; st.wb_by = -1; st.permuted = false;
        ctx->prof_n++;
        ctx->zero_copy_batches++;
        ctx->tmpl_hits++;
        st.busy = true;
        st.n = n;
        st.user_status = status;
        st.user_delta = nullptr;
        st.ticket = (seq << 3) | gpu_index;
        return st.ticket;
    }
    tmpl_flush_stamps(sh);     // the general path reads and writes the stamp table: bring it up to date first
    sh.tmpl_streak = false;
    st.tmpl_gen = 0;

    // control block: tiles ‖ descs, planned straight into pinned memory
    symrelay_tile *h_tiles = reinterpret_cast<symrelay_tile *>(st.h_ctrl);
    uint32_t n_bulk = 0;
    symrelay_plan_check chk;
    chk.in_bytes = in_bytes; chk.out_bytes = out_bytes; chk.max_streams = ctx->opts.max_streams;
    chk.uniform_per = chk.uniform_thread = 0;
    // Large batches bound for the DMA pipeline are planned bucketed (plan.h): descriptors sorted by block count
    // on their way into the control block.  The tile area is then sized for the worst case (one tile per frame).
    int64_t n_tiles = -2;
    bool bucketed = false;
    const size_t bucket_tiles_bytes = align_up((size_t)n * sizeof(symrelay_tile), 16);
    if (ctx->bucket && n >= 4096u && in_bytes + out_bytes > ctx->zero_copy_max) {
        if (st.perm.size() < n) st.perm.resize(n);
        n_tiles = symrelay_plan_bucketed(desc, n, h_tiles, ctx->opts.max_batch_frames, &n_bulk, &chk,
                                         reinterpret_cast<symrelay_desc *>(st.h_ctrl + bucket_tiles_bytes), st.perm.data());
        bucketed = n_tiles >= 0;
    }
    if (n_tiles == -2) n_tiles = symrelay_plan_inline(desc, n, h_tiles, ctx->opts.max_batch_frames, &n_bulk, &chk);
    if (n_tiles < 0) return fail(ctx, SYMRELAY_EINVAL, "bad descriptor (alignment or length) in batch");
    PROF_MARK(ctx, 1);
    const size_t tiles_bytes = bucketed ? bucket_tiles_bytes : align_up((size_t)n_tiles * sizeof(symrelay_tile), 16);
    const size_t ver_off = tiles_bytes + (size_t)n * sizeof(symrelay_desc);   // control block: tiles ‖ descriptors ‖ expected stamps
    // bounds / ownership checks (computed by the planner's first pass over the descriptors)
    uint32_t bad = chk.out_of_bounds;
    const uint32_t any_flags = chk.any_flags, out_overlap = chk.out_overlap;
    const uint32_t ng = (uint32_t)ctx->n_gpus;
    if (ng > 1)
        for (uint32_t i = 0; i < n; i++) bad |= (uint32_t)((int)(desc[i].slot % ng) != gpu_index);
    if (bad) return fail(ctx, SYMRELAY_EINVAL, "a descriptor points outside its arena, names a slot beyond max_streams, or a slot that is not on gpu %d", gpu_index);
    {   // a closed slot holds an all-zero key: sealing with it must be impossible, not merely wrong.  The same
        // walk enforces the chain rule of the header: at most one frame per (slot, direction) in a batch --
        // a second one would race the first on the nonce chain in HBM and both would "succeed" -- except
        // inbound runs (consecutive F_OPEN|F_RUN descriptors of one slot), and those at most 64 long.
        uint32_t *stamp = sh.slot_stamp.data();   // [slot*2+dir]: stamp of the last batch that named it; SLOT_CLOSED for a slot that is not open
        if (sh.stamp_dirty) {
            // the previous batch left its stamp in the table but never ran: let everything in flight finish and
            // declare all stamps so far settled, so that nobody waits for a stamp that is never written
            rc = pk_drain_host(ctx, sh);
            if (rc) return rc;
            sh.pk_floor = sh.stamp;
        }
        sh.stamp_dirty = true;   // until this batch has been posted or launched
        if (++sh.stamp >= 0xfffffe00u) {   // (once in ~4 billion batches) start the stamps over, with nothing in flight
            rc = pk_drain_host(ctx, sh);
            if (rc) return rc;
            for (uint32_t &v : sh.slot_stamp) v = v == SLOT_CLOSED ? SLOT_CLOSED : 0u;
            sh.stamp = 1u; sh.pk_floor = 0u;
        }
        const uint32_t now = sh.stamp, floor = sh.pk_floor, runf = SYMRELAY_F_OPEN | SYMRELAY_F_RUN;
        // ver[i]: the stamp frame i's stream state has to carry before the resident kernel may use it = the last
        // resident-kernel batch that went through that (slot, direction), 0 when that one is long done (stamps up
        // to pk_floor were followed by an ordinary launch on the same stream).  Written behind the descriptors in the
        // control block.
        uint32_t *ver = reinterpret_cast<uint32_t *>(st.h_ctrl + ver_off);
        uint32_t closed = 0, dup = 0, too_long = 0;
        if (!(any_flags & SYMRELAY_F_RUN)) {
            // no runs in the batch (every outbound batch, most inbound ones): the short form of the walk below
            for (uint32_t i = 0; i < n; i++) {
                const uint32_t idx = desc[i].slot * 2u + ((desc[i].len_flags >> 26) & 1u);
                const uint32_t before = stamp[idx];
                closed |= (uint32_t)(before == SLOT_CLOSED);
                dup |= (uint32_t)(before == now);
                stamp[idx] = before == SLOT_CLOSED ? SLOT_CLOSED : now;
                ver[i] = before > floor ? pk_version(before) : 0u;
            }
        } else {
            uint32_t run_len = 1, prev_slot = 0xffffffffu, prev_lf = 0;
            for (uint32_t i = 0; i < n; i++) {
                const uint32_t slot = desc[i].slot, lf = desc[i].len_flags;
                const uint32_t idx = slot * 2u + ((lf >> 26) & 1u);
                const bool cont = slot == prev_slot && (lf & runf) == runf && (prev_lf & runf) == runf;
                const uint32_t before = stamp[idx];
                closed |= (uint32_t)(before == SLOT_CLOSED);
                dup |= (uint32_t)(before == now && !cont);
                stamp[idx] = before == SLOT_CLOSED ? SLOT_CLOSED : now;
                ver[i] = before > floor ? pk_version(before) : 0u;
                run_len = cont ? run_len + 1u : 1u;
                too_long |= (uint32_t)(run_len > SS_RUN_MAX);
                prev_slot = slot; prev_lf = lf;
            }
        }
        if (closed) return fail(ctx, SYMRELAY_ENOENT, "a descriptor names a slot that is not open");
        if (dup) return fail(ctx, SYMRELAY_EINVAL, "two frames of one (slot, direction) in a batch: a stream's frames are MAC-chained, put them in consecutive batches (or mark inbound ones SYMRELAY_F_RUN)");
        if (too_long) return fail(ctx, SYMRELAY_EINVAL, "a run of more than %u SYMRELAY_F_RUN frames of one slot", SS_RUN_MAX);
    }
    bool extract = false;
    const bool runs = (any_flags & SYMRELAY_F_RUN) != 0;
    if (any_flags & SYMRELAY_F_EXTRACT) {
        for (uint32_t i = 0; i < n; i++) {
            const uint32_t mlen = desc[i].len_flags & SYMRELAY_LEN_MASK;
            if ((desc[i].len_flags & SYMRELAY_F_EXTRACT) && !(desc[i].len_flags & SYMRELAY_F_OPEN)) {
                extract = true;
                const uint64_t cap = mlen < 16u ? 16u : ((mlen + 15u) & ~15u);
                if ((uint64_t)desc[i].in_off + cap > delta_arena_bytes || (uint64_t)desc[i].in_off + cap > in_bytes)
                    return fail(ctx, SYMRELAY_EINVAL, "descriptor %u: delta slot (in_off + max(16, align16(len))) outside the arenas", i);
            }
        }
    }
    if (extract && (!delta || !delta_arena))
        return fail(ctx, SYMRELAY_EINVAL, "SYMRELAY_F_EXTRACT needs delta and delta_arena");
    if (extract && !st.d_delta_arena) CU(ctx, cudaMalloc(&st.d_delta_arena, ctx->opts.stage_bytes));
    if (runs && !st.d_run_nonce) CU(ctx, cudaMalloc(&st.d_run_nonce, (size_t)ctx->opts.max_batch_frames * sizeof(uint4)));
    PROF_MARK(ctx, 2);
    if (!bucketed) memcpy(st.h_ctrl + tiles_bytes, desc, (size_t)n * sizeof(symrelay_desc));
    st.permuted = bucketed;
    const size_t ctrl_bytes = tiles_bytes + (size_t)n * sizeof(symrelay_desc);
    PROF_MARK(ctx, 3);

    // Small batches whose arenas are library-pinned run ZERO-COPY: one launch reads descriptors and
    // payload straight from pinned host memory over PCIe and writes wire bytes and status straight back
    // -- no staging copies, a third of the driver calls, and a much shorter path for latency.  Larger
    // batches keep the DMA pipeline below (copy engines overlap the kernels of neighbouring batches).
    // (Not for inbound frames of more than 384 bytes: strided, bulk and wide tiles decrypt in the same pass as the MAC
    // and wipe a forged frame afterwards -- in device memory nobody sees the interval, in the caller's host memory the
    // application could.  Such batches are staged: their results are copied out after the kernel has finished.)
    const bool spec_open = n_bulk != 0 && (any_flags & SYMRELAY_F_OPEN) != 0;
    const bool zero_copy = !spec_open && in_bytes + out_bytes <= ctx->zero_copy_max && is_library_pinned(ctx, in, in_bytes) &&
                           is_library_pinned(ctx, out, out_bytes) &&
                           (!extract || is_library_pinned(ctx, delta_arena, in_bytes < delta_arena_bytes ? in_bytes : delta_arena_bytes));
    if (zero_copy && ctx->persist && n_bulk == 0 && !extract && !runs && n_tiles > 0 && sh.stages.size() <= PK_RING) {
        // Packed-tile batches go to the resident kernel: no launch, no grid boundary between consecutive batches.
        if (sh.pending_wb >= 0) { rc = flush_write_back(ctx, sh, sh.pending_wb); if (rc) return rc; }
        const uint32_t flags = ctx->stage_input ? SS_ARG_STAGE_INPUT : 0u;   // outputs go to device staging: no PCIe-shaped stores needed
        PROF_MARK(ctx, 4);
        rc = pk_post(ctx, sh, st, st.h_ctrl, tiles_bytes, n, (uint32_t)n_tiles, in, out, out_bytes, flags,
                     chk.uniform_per, chk.uniform_thread, reinterpret_cast<const uint32_t *>(st.h_ctrl + ver_off), pk_version(sh.stamp));
        if (rc) return rc;
        sh.stamp_dirty = false;
        {   // Template bookkeeping.  A cheap checksum decides whether this batch looks like the previous one; only
            // then are its descriptors kept (a copy per batch would tax traffic that never repeats).
            const uint64_t *w = reinterpret_cast<const uint64_t *>(desc);
            // four interleaved Fletcher-style sums (a += word; b += a): depends on the ORDER of the descriptors too -- the
            // same frames in another order are another template -- and costs ~0.3 us; memcmp has the last word anyway
            uint64_t a0 = n, a1 = 0, a2 = 0, a3 = 0, b0 = 0, b1 = 0, b2 = 0, b3 = 0;
            for (uint32_t i = 0; i + 4u <= 2u * n; i += 4u) {
                a0 += w[i]; b0 += a0; a1 += w[i + 1]; b1 += a1; a2 += w[i + 2]; b2 += a2; a3 += w[i + 3]; b3 += a3;
            }
            if (n & 1u) { a0 += w[2u * n - 2u]; b0 += a0; a1 += w[2u * n - 1u]; b1 += a1; }
            const uint64_t sum = ((b0 ^ (b1 << 7) ^ (b2 << 13) ^ (b3 << 29)) * 0x9e3779b97f4a7c15ull) ^ (a0 + a1 + a2 + a3) ^ 1ull;
            if (sh.tmpl_gen && sum == sh.tmpl_sum && n == sh.tmpl_n && in_bytes == sh.tmpl_in_bytes && out_bytes == sh.tmpl_out_bytes &&
                memcmp(desc, sh.tmpl_desc.data(), (size_t)n * sizeof(symrelay_desc)) == 0) {
                sh.tmpl_streak = true;    // same as the installed template: the next identical batch takes the fast path
                sh.tmpl_slot_gen = sh.slot_gen;
                st.tmpl_gen = (tiles_bytes == sh.tmpl_tiles_bytes) ? sh.tmpl_gen : 0;
            } else if (sum == sh.last_sum) {
                sh.tmpl_desc.assign(desc, desc + n);
                sh.tmpl_n = n; sh.tmpl_in_bytes = in_bytes; sh.tmpl_out_bytes = out_bytes; sh.tmpl_sum = sum;
                sh.tmpl_n_tiles = (uint32_t)n_tiles; sh.tmpl_uni_per = chk.uniform_per; sh.tmpl_uni_thread = chk.uniform_thread;
                sh.tmpl_flags = flags; sh.tmpl_tiles_bytes = tiles_bytes;
                sh.tmpl_tiles.assign(st.h_ctrl, st.h_ctrl + (chk.uniform_per ? 0 : (size_t)n_tiles * sizeof(symrelay_tile)));
                sh.tmpl_gen++;
                sh.tmpl_slot_gen = sh.slot_gen;
                sh.tmpl_streak = true;
                st.tmpl_gen = sh.tmpl_gen;
            }
            sh.last_sum = sum;
        }
        PROF_MARK(ctx, 5);
        st.deferred = false;
        st.wb_by = -1;
        ctx->prof_n++;
        ctx->zero_copy_batches++;
        st.busy = true;
        st.n = n;
        st.user_status = status;
        st.user_delta = nullptr;
        st.ticket = (seq << 3) | gpu_index;
        return st.ticket;
    }
    pk_quiesce(sh);   // ordinary launches on s_k from here on: whatever the resident kernel was given is settled by stream order
    sh.pk_floor = sh.stamp;
    sh.stamp_dirty = false;   // (a failure below leaves stamps behind, but at or below the floor)
    if (zero_copy) {
        // (Copying the control block or the input with cudaMemcpyAsync on the kernel's stream instead was
        // measured and is much worse: 26-36 us per batch against 16 us, small DMA copies have a high fixed cost.)
        const uint8_t *k_ctrl = st.h_ctrl;
        const uint8_t *k_in = in;
        const int cur = (int)(&st - sh.stages.data());
        // Deferred write-back (pipelined callers only): the next batch of a stream cannot start before this
        // kernel is complete, and complete includes its results having crossed PCIe.  The state chain only
        // needs the nonces, so the kernel leaves the results in device staging and the NEXT kernel on the
        // stream posts them to the host as its first act; they land while that batch computes (measured:
        // 12.7 -> 10.4 us per 2,048-frame batch).  Only for batches that are one packed-tile kernel.
        const bool defer = may_defer && ctx->defer_wb && n_bulk == 0 && !extract && !runs && (uint32_t)n_tiles > 0;
        // ... and this batch carries the previous one's results if it launches a packed kernel at all
        const Stage *carry = nullptr;
        if (sh.pending_wb >= 0) {
            if ((uint32_t)n_tiles > n_bulk) carry = &sh.stages[(size_t)sh.pending_wb];
            else { rc = flush_write_back(ctx, sh, sh.pending_wb); if (rc) return rc; }
        }
        uint32_t flags = (!extract ? SS_ARG_INPUT_INDEPENDENT : 0u) | (ctx->stage_input ? SS_ARG_STAGE_INPUT : 0u);
        if (ctx->stage_input && !out_overlap && !defer) flags |= SS_ARG_STAGE_OUTPUT;   // device staging needs no PCIe-shaped stores
        rc = launch_tiles(ctx, reinterpret_cast<const symrelay_desc *>(k_ctrl + tiles_bytes), n,
                          reinterpret_cast<const symrelay_tile *>(k_ctrl), (uint32_t)n_tiles, n_bulk, k_in,
                          defer ? st.d_out : out, sh.d_states, defer ? st.d_status : st.h_status,
                          extract ? st.h_delta : nullptr, extract ? delta_arena : nullptr, sh.s_k, flags,
                          runs ? st.d_run_nonce : nullptr, &sh, carry, reinterpret_cast<const symrelay_tile *>(k_ctrl));
        if (rc) return rc;
        PROF_MARK(ctx, 4);
        CU(ctx, cudaEventRecord(st.ev_out, sh.s_k));
        PROF_MARK(ctx, 5);
        if (carry) { sh.stages[(size_t)sh.pending_wb].wb_by = cur; sh.pending_wb = -1; }
        st.deferred = defer;
        st.wb_by = -1;
        st.user_out = out;
        st.out_bytes = out_bytes;
        if (defer) sh.pending_wb = cur;
        ctx->prof_n++;
        ctx->zero_copy_batches++;
        st.busy = true;
        st.n = n;
        st.user_status = status;
        st.user_delta = extract ? delta : nullptr;
        st.ticket = (seq << 3) | gpu_index;
        return st.ticket;
    }
    if (sh.pending_wb >= 0) { rc = flush_write_back(ctx, sh, sh.pending_wb); if (rc) return rc; }
    ctx->staged_batches++;
    st.dev_tmpl_gen = 0;   // d_ctrl is overwritten
    CU(ctx, cudaMemcpyAsync(st.d_ctrl, st.h_ctrl, ctrl_bytes, cudaMemcpyHostToDevice, sh.s_in));
    if (in_bytes) CU(ctx, cudaMemcpyAsync(st.d_in, in, in_bytes, cudaMemcpyHostToDevice, sh.s_in));
    CU(ctx, cudaEventRecord(st.ev_in, sh.s_in));
    CU(ctx, cudaStreamWaitEvent(sh.s_k, st.ev_in, 0));
    rc = launch_tiles(ctx, reinterpret_cast<const symrelay_desc *>(st.d_ctrl + tiles_bytes), n,
                      reinterpret_cast<const symrelay_tile *>(st.d_ctrl), (uint32_t)n_tiles, n_bulk, st.d_in, st.d_out,
                      sh.d_states, st.d_status, extract ? st.d_delta : nullptr, extract ? st.d_delta_arena : nullptr,
                      sh.s_k, 0u, runs ? st.d_run_nonce : nullptr, &sh);
    if (rc) return rc;
    CU(ctx, cudaEventRecord(st.ev_k, sh.s_k));
    CU(ctx, cudaStreamWaitEvent(sh.s_out, st.ev_k, 0));
    if (out_bytes) CU(ctx, cudaMemcpyAsync(out, st.d_out, out_bytes, cudaMemcpyDeviceToHost, sh.s_out));
    CU(ctx, cudaMemcpyAsync(st.h_status, st.d_status, (size_t)n * 4u, cudaMemcpyDeviceToHost, sh.s_out));
    if (extract) {
        CU(ctx, cudaMemcpyAsync(st.h_delta, st.d_delta, (size_t)n * sizeof(symrelay_delta), cudaMemcpyDeviceToHost, sh.s_out));
        CU(ctx, cudaMemcpyAsync(delta_arena, st.d_delta_arena, in_bytes < delta_arena_bytes ? in_bytes : delta_arena_bytes,
                                cudaMemcpyDeviceToHost, sh.s_out));
    }
    CU(ctx, cudaEventRecord(st.ev_out, sh.s_out));
    st.deferred = false;
    st.busy = true;
    st.n = n;
    st.user_status = status;
    st.user_delta = extract ? delta : nullptr;
    st.ticket = (seq << 3) | gpu_index;
    return st.ticket;
}

int64_t symrelay_batch_submit(symrelay_ctx *ctx, const symrelay_desc *desc, uint32_t n, const uint8_t *in,
                              uint64_t in_bytes, uint8_t *out, uint64_t out_bytes, uint32_t *status,
                              symrelay_delta *delta, uint8_t *delta_arena, uint64_t delta_arena_bytes) {
    if (!ctx || !desc || n == 0 || !in || !out) return fail(ctx, SYMRELAY_EINVAL, "symrelay_batch_submit: null/empty argument");
    const int gpu_index = (int)(desc[0].slot % (uint32_t)ctx->n_gpus);
    std::lock_guard<std::mutex> lk(ctx->shard[gpu_index].mu);
    return submit_locked(ctx, gpu_index, desc, n, in, in_bytes, out, out_bytes, status, delta, delta_arena,
                         delta_arena_bytes, /*may_defer=*/true);
}

int symrelay_batch_wait(symrelay_ctx *ctx, int64_t ticket) {
    if (!ctx || ticket < 0) return SYMRELAY_EINVAL;
    const int gpu_index = (int)(ticket & 7);
    if (gpu_index >= ctx->n_gpus) return fail(ctx, SYMRELAY_ENOENT, "bad ticket");
    Shard &sh = ctx->shard[gpu_index];
    std::lock_guard<std::mutex> lk(sh.mu);
    const int64_t seq = ticket >> 3;
    if (seq >= sh.next_ticket_seq) return fail(ctx, SYMRELAY_ENOENT, "ticket not issued");
    Stage &st = sh.stages[(size_t)(seq % (int64_t)sh.stages.size())];
    if (st.busy && st.ticket == ticket) {
        double prof_t0 = ctx->prof ? prof_now() : 0.0;
        const int rc = finish_stage(ctx, sh, st);
        PROF_MARK(ctx, 6);
        return rc;
    }
    return 0;  // already recycled => already finished
}

int symrelay_batch_host(symrelay_ctx *ctx, const symrelay_desc *desc, uint32_t n, const uint8_t *in,
                        uint64_t in_bytes, uint8_t *out, uint64_t out_bytes, uint32_t *status,
                        symrelay_delta *delta, uint8_t *delta_arena, uint64_t delta_arena_bytes) {
    if (n == 0) return 0;
    if (!ctx || !desc || !in || !out) return fail(ctx, SYMRELAY_EINVAL, "symrelay_batch_host: null argument");
    int64_t t;
    {   // a synchronous caller waits for this very batch: let the kernel write to the host itself
        std::lock_guard<std::mutex> lk(ctx->shard[desc[0].slot % (uint32_t)ctx->n_gpus].mu);
        t = submit_locked(ctx, (int)(desc[0].slot % (uint32_t)ctx->n_gpus), desc, n, in, in_bytes, out, out_bytes, status,
                          delta, delta_arena, delta_arena_bytes, /*may_defer=*/false);
    }
    if (t < 0) return (int)t;
    return symrelay_batch_wait(ctx, t);
}

#include "scheduler_impl.inc"

int symrelay_probe_int_peak(symrelay_ctx *ctx, int gpu_index, double *arx_tops, double *imad_tops) {
    if (!ctx || gpu_index < 0 || gpu_index >= ctx->n_gpus) return SYMRELAY_EINVAL;
    Shard &sh = ctx->shard[gpu_index];
    std::lock_guard<std::mutex> lk(sh.mu);
    CU(ctx, set_device(sh.dev));
    pk_quiesce(sh);
    uint32_t *d_out = reinterpret_cast<uint32_t *>(sh.d_init);  // never written in practice
    cudaEvent_t e0, e1;
    CU(ctx, cudaEventCreate(&e0));
    CU(ctx, cudaEventCreate(&e1));
    const int blocks = 148 * 8, threads = 256;
    float best_arx = 1e30f, best_imad = 1e30f;
    const int dr = 4000, it = 8000;
    for (int rep = 0; rep < 4; rep++) {
        CU(ctx, cudaEventRecord(e0, sh.s_k));
        ss_probe_arx_kernel<<<blocks, threads, 0, sh.s_k>>>(d_out, dr);
        CU(ctx, cudaEventRecord(e1, sh.s_k));
        CU(ctx, cudaEventSynchronize(e1));
        float ms = 0;
        CU(ctx, cudaEventElapsedTime(&ms, e0, e1));
        if (rep && ms < best_arx) best_arx = ms;
        CU(ctx, cudaEventRecord(e0, sh.s_k));
        ss_probe_imad_kernel<<<blocks, threads, 0, sh.s_k>>>(d_out, it);
        CU(ctx, cudaEventRecord(e1, sh.s_k));
        CU(ctx, cudaEventSynchronize(e1));
        CU(ctx, cudaEventElapsedTime(&ms, e0, e1));
        if (rep && ms < best_imad) best_imad = ms;
        ctx->launches += 2;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    const double nthreads = (double)blocks * threads;
    if (arx_tops) *arx_tops = nthreads * dr * 96.0 / (best_arx * 1e-3) / 1e12;   // 8 quarter-rounds x 12 ops
    if (imad_tops) *imad_tops = nthreads * it * 8.0 / (best_imad * 1e-3) / 1e12;  // multiply-adds only
    return 0;
}

void *symrelay_alloc_pinned(symrelay_ctx *ctx, uint64_t bytes) {
    void *p = nullptr;
    if (!ctx || bytes == 0) return nullptr;
    set_device(ctx->shard[0].dev);
    if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable | cudaHostAllocMapped) != cudaSuccess) {
        fail(ctx, SYMRELAY_ENOMEM, "cudaHostAlloc(%llu) failed", (unsigned long long)bytes);
        return nullptr;
    }
    if (!pin_add(ctx, p, bytes)) {
        cudaFreeHost(p);
        fail(ctx, SYMRELAY_ENOMEM, "too many pinned arenas (2048)");
        return nullptr;
    }
    return p;
}

void symrelay_free_pinned(symrelay_ctx *ctx, void *p) {
    if (!p) return;
    if (ctx) pin_remove(ctx, p);
    cudaFreeHost(p);
}

}  // extern "C"
