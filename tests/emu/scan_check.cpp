// Test infrastructure: the planner's AVX2 descriptor scan (symmetry_b200/csrc/plan_scan_avx2.cpp) against the
// scalar loop of plan.h on random batches, valid and invalid.  Exit code 0 = identical everywhere, 77 = no AVX2.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>
#define SYMRELAY_HAVE_AVX2_SCAN
#include "../../symmetry_b200/csrc/plan.h"

int main() {
    if (!symrelay_plan_scan_has_avx2()) { printf("no avx2\n"); return 77; }
    std::mt19937 rng(11);
    for (int it = 0; it < 20000; it++) {
        const uint32_t n = it < 200 ? (uint32_t)it : 1u + rng() % 300u;
        std::vector<symrelay_desc> d(n);
        const int mode = (int)(rng() % 8u);
        uint32_t off = 0;
        for (uint32_t i = 0; i < n; i++) {
            uint32_t len = rng() % 5u == 0 ? rng() % 70000u : rng() % 300u;
            const bool op = rng() & 1u;
            d[i].slot = rng() % 5000u;
            d[i].len_flags = len | (op ? SYMRELAY_F_OPEN : 0u) | ((rng() & 3u) << 24) | (rng() % 7u == 0 ? SYMRELAY_F_EXTRACT : 0u);
            d[i].in_off = off + (op ? 12u : 0u);
            d[i].out_off = off + (op ? 0u : 12u);
            off += ((len + 32u + 15u) & ~15u) + (rng() % 3u) * 16u;
            if (mode == 1 && rng() % 40u == 0) d[i].in_off += 1u + rng() % 15u;          // misaligned
            if (mode == 2 && rng() % 40u == 0) d[i].out_off ^= 4u;
            if (mode == 3 && rng() % 40u == 0) d[i].len_flags |= 0x00ffffffu;             // too long
            if (mode == 4 && rng() % 40u == 0) d[i].in_off = 0xfffffff0u + (op ? 12u : 0u); // out of bounds
            if (mode == 5 && rng() % 40u == 0) d[i].out_off = (rng() % (off + 1u)) & ~15u | (op ? 0u : 12u);   // overlap
            if (mode == 6 && rng() % 40u == 0) d[i].slot = 0xffffffffu - rng() % 3u;
        }
        const uint32_t in_bytes = mode == 7 ? rng() % (off + 64u) : off + 64u, out_bytes = rng() % 9u == 0 ? rng() % (off + 64u) : off + 64u;
        const uint32_t streams[5] = {0u, 1u, 4096u, 5000u, 0xffffffffu};
        const uint32_t max_streams = streams[rng() % 5u];
        std::vector<uint32_t> a(3 * n + 8), b(3 * n + 8);
        uint32_t oa[6], ob[6];
        symrelay_plan_scan_scalar(d.data(), n, a.data(), a.data() + n, a.data() + 2 * n, in_bytes, out_bytes, max_streams, oa);
        symrelay_plan_scan_avx2(d.data(), n, b.data(), b.data() + n, b.data() + 2 * n, in_bytes, out_bytes, max_streams, ob);
        bool same = memcmp(a.data(), b.data(), 3 * (size_t)n * 4) == 0 && (oa[0] != 0) == (ob[0] != 0) && (oa[1] != 0) == (ob[1] != 0) &&
                    oa[2] == ob[2] && (n == 0 || (oa[3] == ob[3] && oa[4] == ob[4])) && (oa[5] != 0) == (ob[5] != 0);
        if (!same) {
            printf("mismatch it=%d n=%u mode=%d max_streams=%u: scalar %u %u %x %u %u %u, avx2 %u %u %x %u %u %u\n", it, n, mode, max_streams,
                   oa[0], oa[1], oa[2], oa[3], oa[4], oa[5], ob[0], ob[1], ob[2], ob[3], ob[4], ob[5]);
            return 1;
        }
    }
    printf("identical on 20000 batches\n");
    return 0;
}
