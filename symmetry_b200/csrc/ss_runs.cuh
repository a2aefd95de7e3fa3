// ss_runs.cuh -- opening several queued frames of ONE stream in the same batch.
//
// In the reference, secret-stream's `_incoming` hands frames to libsodium one at a time
// (/root/reference/src/provider.ts:174 sits behind it), because frame i+1's nonce is frame i's nonce
// XOR mac_i[0:8] with the counter bumped.  On the receive side the MACs are already on the wire, so
// with no REKEY in between frame p of a run can be opened independently
// ("parallel open"):
//        nonce_p = LE32(counter + p) ‖ (inonce ^ mac_0[0:8] ^ ... ^ mac_{p-1}[0:8])
// A run = consecutive descriptors of one slot with SYMRELAY_F_OPEN | SYMRELAY_F_RUN.
//   1. ss_run_nonce_kernel   one thread per run frame: walk back over the run, XOR the wire MACs,
//                            write the speculative nonce into the per-batch scratch array;
//   2. the ordinary tile kernels open every frame with that nonce and record OK|tag or BAD_MAC,
//      leaving the stream state alone;
//   3. ss_run_commit_kernel  one thread per run: accept the longest prefix of verified frames,
//                            stop after a failed frame or a REKEY/counter wrap (the speculation for
//                            the frames behind it was wrong: they get SS_ST_RETRY and are resubmitted),
//                            and advance the state exactly as that many sequential pulls would.
#pragma once
#include "ss_core.cuh"

#define SS_RUN_MAX SS_RUN_CAP   // frames of one stream per batch (64)

SS_DEV void ss_wire_mac8(const uint8_t *in, const SsDesc &d, uint32_t &m0, uint32_t &m1) {
    const uint8_t *p = in + d.in_off + 4u + (d.len_flags & 0xffffffu);
    m0 = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
    m1 = (uint32_t)p[4] | ((uint32_t)p[5] << 8) | ((uint32_t)p[6] << 16) | ((uint32_t)p[7] << 24);
}

SS_DEV bool ss_same_run(const SsDesc &a, const SsDesc &b) {
    const uint32_t f = SS_F_OPEN | SS_F_RUN;
    return (a.len_flags & f) == f && (b.len_flags & f) == f && a.slot == b.slot;
}

// thread i: nonce of descriptor i if it is a run frame
SS_DEV void ss_run_nonce(const SsDesc *desc, uint32_t n, const uint8_t *in, const uint4 *states, uint4 *run_nonce,
                         uint32_t i) {
    if (i >= n) return;
    const SsDesc d = desc[i];
    const uint32_t f = SS_F_OPEN | SS_F_RUN;
    if ((d.len_flags & f) != f) return;
    const uint4 s2 = states[((size_t)d.slot * 2u + 1u) * 3u + 2u];
    uint32_t n1 = s2.y, n2 = s2.z, p = 0;
    for (uint32_t j = i; j > 0 && p < SS_RUN_MAX; j--) {
        const SsDesc prev = desc[j - 1];
        if (!ss_same_run(prev, d)) break;
        uint32_t m0, m1;
        ss_wire_mac8(in, prev, m0, m1);
        n1 ^= m0; n2 ^= m1;
        p++;
    }
    // p == SS_RUN_MAX: at least that many frames precede this one, its nonce would be truncated.  The tile
    // kernels read .w, leave such a frame alone and report RETRY (resubmit it in the next batch).
    uint4 o;
    o.x = s2.x + p; o.y = n1; o.z = n2; o.w = p;
    run_nonce[i] = o;
}

// thread i: if descriptor i starts a run, commit its verified prefix
SS_DEV void ss_run_commit(const SsDesc *desc, uint32_t n, const uint8_t *in, uint4 *states, uint32_t *status,
                          uint32_t i) {
    if (i >= n) return;
    const SsDesc d = desc[i];
    const uint32_t f = SS_F_OPEN | SS_F_RUN;
    if ((d.len_flags & f) != f) return;
    if (i > 0 && ss_same_run(desc[i - 1], d)) return;   // not the first frame of its run
    uint4 *sp = states + ((size_t)d.slot * 2u + 1u) * 3u;
    uint4 s0 = sp[0], s1 = sp[1], s2 = sp[2];
    uint32_t k[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
    uint32_t nn[3] = {s2.x, s2.y, s2.z};
    bool live = true, rekeyed = false;
    for (uint32_t j = i, p = 0; j < n; j++, p++) {
        const SsDesc e = desc[j];
        if (j != i && !ss_same_run(desc[j - 1], e)) break;
        if (!live || p >= SS_RUN_MAX) { status[j] = SS_ST_RETRY; live = false; continue; }
        const uint32_t st = status[j];
        if ((st & 0xffu) != SS_ST_OK) { live = false; continue; }   // keeps its BAD_MAC; the stream is dead there
        uint32_t m0, m1;
        ss_wire_mac8(in, e, m0, m1);
        nn[1] ^= m0; nn[2] ^= m1; nn[0] += 1u;
        if (((st >> 8) & 2u) || nn[0] == 0u) {   // REKEY tag or counter wrap: later frames used the wrong key
            ss_rekey(k, nn);
            rekeyed = true;
            live = false;
        }
    }
    if (rekeyed) {
        s0.x = k[0]; s0.y = k[1]; s0.z = k[2]; s0.w = k[3];
        s1.x = k[4]; s1.y = k[5]; s1.z = k[6]; s1.w = k[7];
        sp[0] = s0; sp[1] = s1;
    }
    s2.x = nn[0]; s2.y = nn[1]; s2.z = nn[2];
    sp[2] = s2;
}
