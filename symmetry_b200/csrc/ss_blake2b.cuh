// ss_blake2b.cuh -- keyed BLAKE2b-256 of a short message, one thread per connection.
//
// What it replaces: @hyperswarm/secret-stream computes a 32-byte stream id per direction when a connection
// is set up -- libsodium crypto_generichash(out[32], namespace[32], key = Noise handshake hash[64]) -- and
// puts it, with the 24-byte secretstream header, into the first message of the connection
// (uint24(56) ‖ streamId ‖ header; the receive side is typed at /root/reference/src/types.ts:156-159, the
// hook is hyperswarm's 'connection' event, src/provider.ts:55-58).  SURVEY.md 8(f) N3: for connection churn
// that is done for many new connections at once, next to the HChaCha20 sub-key derivation (ss_init_kernel).
// RFC 7693, sequential mode, no salt/personalisation (libsodium's crypto_generichash defaults).
//
// Compiled by nvcc for the product and by g++ under tests/emu (SS_HOST_EMU) for a CPU check against libsodium.
#pragma once
#include <stdint.h>

#ifdef SS_HOST_EMU
#define SB_DEV static inline
#else
#define SB_DEV __device__ __forceinline__
#endif

SB_DEV uint64_t sb_rotr(uint64_t x, int n) { return (x >> n) | (x << (64 - n)); }

#define SB_G(a, b, c, d, x, y)                  \
    a = a + b + (x); d = sb_rotr(d ^ a, 32);    \
    c = c + d;       b = sb_rotr(b ^ c, 24);    \
    a = a + b + (y); d = sb_rotr(d ^ a, 16);    \
    c = c + d;       b = sb_rotr(b ^ c, 63);

// One compression: h <- F(h, m, t, last)
SB_DEV void sb_compress(uint64_t h[8], const uint64_t m[16], uint64_t t, bool last) {
    const uint64_t IV[8] = {0x6a09e667f3bcc908ull, 0xbb67ae8584caa73bull, 0x3c6ef372fe94f82bull, 0xa54ff53a5f1d36f1ull,
                            0x510e527fade682d1ull, 0x9b05688c2b3e6c1full, 0x1f83d9abfb41bd6bull, 0x5be0cd19137e2179ull};
    const uint8_t SIGMA[12][16] = {
        {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15}, {14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3},
        {11, 8, 12, 0, 5, 2, 15, 13, 10, 14, 3, 6, 7, 1, 9, 4}, {7, 9, 3, 1, 13, 12, 11, 14, 2, 6, 5, 10, 4, 0, 15, 8},
        {9, 0, 5, 7, 2, 4, 10, 15, 14, 1, 11, 12, 6, 8, 3, 13}, {2, 12, 6, 10, 0, 11, 8, 3, 4, 13, 7, 5, 15, 14, 1, 9},
        {12, 5, 1, 15, 14, 13, 4, 10, 0, 7, 6, 3, 9, 2, 8, 11}, {13, 11, 7, 14, 12, 1, 3, 9, 5, 0, 15, 4, 8, 6, 2, 10},
        {6, 15, 14, 9, 11, 3, 0, 8, 12, 2, 13, 7, 1, 4, 10, 5}, {10, 2, 8, 4, 7, 6, 1, 5, 15, 11, 9, 14, 3, 12, 13, 0},
        {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15}, {14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3}};
    uint64_t v[16];
    for (int i = 0; i < 8; i++) { v[i] = h[i]; v[8 + i] = IV[i]; }
    v[12] ^= t;                 // (messages here are far below 2^64 bytes: the high counter word stays 0)
    if (last) v[14] = ~v[14];
    for (int r = 0; r < 12; r++) {
        const uint8_t *s = SIGMA[r];
        SB_G(v[0], v[4], v[8], v[12], m[s[0]], m[s[1]])
        SB_G(v[1], v[5], v[9], v[13], m[s[2]], m[s[3]])
        SB_G(v[2], v[6], v[10], v[14], m[s[4]], m[s[5]])
        SB_G(v[3], v[7], v[11], v[15], m[s[6]], m[s[7]])
        SB_G(v[0], v[5], v[10], v[15], m[s[8]], m[s[9]])
        SB_G(v[1], v[6], v[11], v[12], m[s[10]], m[s[11]])
        SB_G(v[2], v[7], v[8], v[13], m[s[12]], m[s[13]])
        SB_G(v[3], v[4], v[9], v[14], m[s[14]], m[s[15]])
    }
    for (int i = 0; i < 8; i++) h[i] ^= v[i] ^ v[8 + i];
}

// out[4 words of 64 bits] = BLAKE2b-256(key = key64 (64 bytes), message = msg32 (32 bytes)); words little-endian.
SB_DEV void sb_blake2b256_keyed(const uint64_t key64[8], const uint64_t msg32[4], uint64_t out[4]) {
    uint64_t h[8] = {0x6a09e667f3bcc908ull, 0xbb67ae8584caa73bull, 0x3c6ef372fe94f82bull, 0xa54ff53a5f1d36f1ull,
                     0x510e527fade682d1ull, 0x9b05688c2b3e6c1full, 0x1f83d9abfb41bd6bull, 0x5be0cd19137e2179ull};
    h[0] ^= 0x01010000ull ^ (64ull << 8) ^ 32ull;   // depth 1, fan-out 1, key length 64, digest length 32
    uint64_t m[16];
    for (int i = 0; i < 8; i++) { m[i] = key64[i]; m[8 + i] = 0; }
    sb_compress(h, m, 128, false);                  // the key, padded to a block
    for (int i = 0; i < 4; i++) m[i] = msg32[i];
    for (int i = 4; i < 16; i++) m[i] = 0;
    sb_compress(h, m, 128 + 32, true);
    for (int i = 0; i < 4; i++) out[i] = h[i];
}
