// emu_main.cpp -- TEST INFRASTRUCTURE ONLY (see emu_warp.h).
// Runs the product's tile code (ss_core.cuh) and host planner (plan.h) on CPU
// threads so tests/test_emu_kernel_logic.py can compare them with the oracle
// without a GPU.
#define SS_HOST_EMU 1
#include "emu_warp.h"

#include "../../symmetry_b200/csrc/plan.h"
#include "../../symmetry_b200/csrc/ss_core.cuh"
#include "../../symmetry_b200/csrc/ss_extract.cuh"
#include "../../symmetry_b200/csrc/ss_runs.cuh"

#include <cstring>
#include <thread>
#include <vector>

static uint32_t emu_arg_flags = 0;   // SsArgs.flags for the next batches (tests flip SS_ARG_STAGE_INPUT)
extern "C" void emu_set_arg_flags(uint32_t f) { emu_arg_flags = f; }
thread_local uint32_t emu_lane_id;
thread_local EmuWarp *emu_warp;

struct Job {
    EmuWarp *warp;
    SsArgs args;
    uint32_t n_packed, n_bulk;
    uint32_t lane;
};

static void *lane_main(void *p) {
    Job *j = (Job *)p;
    emu_lane_id = j->lane;
    emu_warp = j->warp;
    for (uint32_t t = 0; t < j->n_packed; t++) {
        if (j->args.tiles[t].n_kind & (1u << 17)) ss_process_thread_tile(j->args, t);
        else ss_process_tile<false>(j->args, t);
    }
    for (uint32_t t = 0; t < j->n_bulk; t++) {
        if ((j->args.tiles[j->n_packed + t].n_kind >> 24) & 63u) ss_process_tile<true, false, true>(j->args, j->n_packed + t);   // strided
        else ss_process_tile<true>(j->args, j->n_packed + t);
    }
    return nullptr;
}

#include "../../symmetry_b200/csrc/ss_blake2b.cuh"
// the product's BLAKE2b (stream ids) for a CPU check against libsodium's crypto_generichash
extern "C" void emu_blake2b256_keyed(const uint8_t key[64], const uint8_t msg[32], uint8_t out[32]) {
    uint64_t k[8], m[4], o[4];
    memcpy(k, key, 64); memcpy(m, msg, 32);
    sb_blake2b256_keyed(k, m, o);
    memcpy(out, o, 32);
}

extern "C" int64_t emu_plan(const symrelay_desc *desc, uint32_t n, symrelay_tile *tiles, uint32_t max_tiles,
                            uint32_t *n_bulk) {
    return symrelay_plan_inline(desc, n, tiles, max_tiles, n_bulk);
}

extern "C" int emu_run_batch(const symrelay_desc *desc, const symrelay_tile *tiles, uint32_t n_tiles, uint32_t n_bulk,
                             const uint8_t *in, uint8_t *out, void *states, uint32_t *status) {
    // runs: nonce pre-pass (thread-local code, run serially here), tiles, then the commit pass
    uint32_t n_desc = 0;
    for (uint32_t t = 0; t < n_tiles; t++) {
        const uint32_t end = tiles[t].first + (tiles[t].n_kind & 0xffffu);
        if (end > n_desc) n_desc = end;
    }
    std::vector<uint4> run_nonce(n_desc ? n_desc : 1);
    for (uint32_t i = 0; i < n_desc; i++)
        ss_run_nonce((const SsDesc *)desc, n_desc, in, (const uint4 *)states, run_nonce.data(), i);
    EmuWarp warp;
    pthread_barrier_init(&warp.bar, nullptr, 32);
    Job jobs[32];
    pthread_t th[32];
    for (uint32_t l = 0; l < 32; l++) {
        jobs[l].warp = &warp;
        jobs[l].args = SsArgs{(const SsDesc *)desc, (const SsTile *)tiles, n_tiles, in, out, (uint4 *)states, status, emu_arg_flags, run_nonce.data()};
        jobs[l].n_packed = n_tiles - n_bulk;
        jobs[l].n_bulk = n_bulk;
        jobs[l].lane = l;
        pthread_create(&th[l], nullptr, lane_main, &jobs[l]);
    }
    for (uint32_t l = 0; l < 32; l++) pthread_join(th[l], nullptr);
    pthread_barrier_destroy(&warp.bar);
    for (uint32_t i = 0; i < n_desc; i++) ss_run_commit((const SsDesc *)desc, n_desc, in, (uint4 *)states, status, i);
    return 0;
}

static uint64_t emu_fast_calls = 0, emu_fast_taken = 0;
extern "C" void emu_extract_fast_counts(uint64_t *calls, uint64_t *taken) { *calls = emu_fast_calls; *taken = emu_fast_taken; }

// The extractor is thread-local code (no warp collectives): call it directly.
extern "C" uint32_t emu_extract(const uint8_t *text, uint32_t n, uint32_t kind, uint8_t *out, uint32_t cap,
                                uint32_t *len_out) {
    static uint16_t table[SX_TABLE_ENTRIES];
    static bool built = false;
    if (!built) { sx::build_table(table); built = true; }
    // run on a 16-byte aligned copy (as the kernel sees its arena) and once more at an odd address, so
    // both the word-skipping and the byte-only paths are compared with the oracle
    static std::vector<uint8_t> buf;
    buf.assign((size_t)n + 64, 0);
    uint8_t *al = buf.data() + ((16 - (reinterpret_cast<uintptr_t>(buf.data()) & 15)) & 15);
    memcpy(al, text, n);
    const sx::Result r = sx::extract_chunk(al, n, kind, out, cap, table);
    {
        static std::vector<uint8_t> buf2, out2;
        buf2.assign((size_t)n + 64, 0);
        out2.assign((size_t)cap + 16, 0);
        uint8_t *od = buf2.data() + ((16 - (reinterpret_cast<uintptr_t>(buf2.data()) & 15)) & 15) + 1;
        memcpy(od, text, n);
        const sx::Result r2 = sx::extract_chunk(od, n, kind, out2.data(), cap, table);
        if (r2.status != r.status || r2.len != r.len || (r.status == 0 && memcmp(out2.data(), out, r.len < cap ? r.len : cap)))
            return 0xdeadu;   // the two paths disagree: fail the test loudly
    }
    {
        // the two-pass form the kernel runs on staged chunks, with the kernel's event capacity: whenever
        // it accepts the chunk it has to agree with the single pass byte for byte
        static uint64_t fast[SX_FAST_ENTRIES];
        static uint8_t cls[256];
        static uint32_t n_cls = 0;
        if (!n_cls) { n_cls = sx::build_fast_tables(fast, cls); if (!n_cls) return 0xdeaeu; }
        if (n <= SX_FAST_MAX_LEN) {
            static std::vector<uint8_t> out3;
            static uint32_t ev[SX_FAST_SCRATCH_WORDS_N(95u, SX_FAST_MAX_LEN)];   // 95 events, as in the kernel
            out3.assign((size_t)cap + 16, 0);
            sx::Result r3;
            emu_fast_calls++;
            if (sx::extract_fast(al, n, kind, out3.data(), cap, fast, cls, ev, 95, &r3)) {
                emu_fast_taken++;
                if (r3.status != r.status || r3.len != r.len || (r.status == 0 && memcmp(out3.data(), out, r.len < cap ? r.len : cap)))
                    return 0xdeafu;
                for (uint32_t i = cap; i < cap + 16; i++) if (out3[i]) return 0xdeb0u;
            }
        }
    }
    {
        // ... and the same three roles as the GPU runs them: three threads at once, queues pre-filled with the "empty"
        // pattern, each consumer polling for its next word (real interleavings of the hand-over, on CPU threads)
        static uint64_t fast2[SX_FAST_ENTRIES];
        static uint8_t cls2[256];
        static bool built2 = false;
        if (!built2) { if (!sx::build_fast_tables(fast2, cls2)) return 0xdeb1u; built2 = true; }
        if (n <= 384u) {   // the kernel stages chunks up to this size
            static std::vector<uint8_t> out4;
            out4.assign((size_t)cap + 16, 0);
            static uint32_t scratch[SX_FAST_SCRATCH_WORDS_N(95u, 384u)];
            uint32_t *stack = scratch, *trailer = scratch + SX_FAST_STACK_WORDS, *fl = trailer + 1u,
                     *evq = fl + SX_FAST_FLAG_WORDS(n);
            *trailer = SX_PIPE_EMPTY;
            for (uint32_t w = 0; w < SX_FAST_FLAG_WORDS(n); w++) fl[w] = SX_PIPE_EMPTY;
            for (uint32_t w = 0; w <= 95u; w++) evq[w] = SX_PIPE_EMPTY;
            const bool prefixed = sx::fast_prefixed(al, n);
            sx::Result r4;
            bool done4 = false;
            std::thread ta([&] { sx::fast_grammar(al, n, fast2, cls2, stack, fl, trailer); });
            std::thread tb([&] { sx::fast_record(n, prefixed, fl, trailer, evq, 95u); });
            std::thread tc([&] { done4 = sx::fast_bookkeep(al, n, prefixed, kind, out4.data(), cap, evq, trailer, 95u, &r4); });
            ta.join(); tb.join(); tc.join();
            if (done4 && (r4.status != r.status || r4.len != r.len || (r.status == 0 && memcmp(out4.data(), out, r.len < cap ? r.len : cap))))
                return 0xdeb2u;
        }
    }
    *len_out = r.len;
    return r.status;
}
