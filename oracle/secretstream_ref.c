/*
 * oracle/secretstream_ref.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement of the arithmetic behind the reference's
 * `peer.write(chunk)` / `peer.on('data')` path:
 *     /root/reference/src/provider.ts:248 (seal), :174 (open),
 *     state shape /root/reference/src/types.ts:168-177.
 * The arithmetic itself is NOT in /root/reference: it lives in the un-vendored
 * npm dependency chain hyperswarm 4.8.0 -> @hyperswarm/secret-stream 6.6.1 ->
 * sodium-secretstream 1.1.0 -> sodium-native 4.1.1 -> libsodium
 * (/root/reference/package-lock.json:763-777, 5725-5758).  This file restates
 * libsodium's *published* construction (crypto_secretstream_xchacha20poly1305,
 * RFC 8439 ChaCha20/Poly1305, draft-irtf-cfrg-xchacha-03 HChaCha20) as written
 * out in SURVEY.md section 8(a) "Spec box for A9-A11".
 *
 * Pinning: tests/test_oracle_*.py check every function here against
 *   (1) RFC 8439 2.3.2 / 2.4.2 / 2.5.2 and XChaCha-draft 2.2.1 known answers,
 *   (2) the golden chain in SURVEY.md 8(c) and tests/golden/ (generated from the
 *       real libsodium 1.0.20 bundled with PyNaCl; script committed),
 *   (3) live differential runs against that same libsodium.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library.  The CUDA product never does.
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>

#define SS_STATEBYTES 52
#define SS_HEADERBYTES 24
#define SS_KEYBYTES 32
#define SS_ABYTES 17
#define SS_TAG_REKEY 0x02

static inline uint32_t ld32(const uint8_t *p) {
    return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}
static inline void st32(uint8_t *p, uint32_t v) {
    p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24);
}
static inline void st64(uint8_t *p, uint64_t v) {
    for (int i = 0; i < 8; i++) p[i] = (uint8_t)(v >> (8 * i));
}
static inline uint32_t rotl(uint32_t x, int n) { return (x << n) | (x >> (32 - n)); }

#define QR(a, b, c, d)                       \
    do {                                     \
        a += b; d ^= a; d = rotl(d, 16);     \
        c += d; b ^= c; b = rotl(b, 12);     \
        a += b; d ^= a; d = rotl(d, 8);      \
        c += d; b ^= c; b = rotl(b, 7);      \
    } while (0)

static void chacha_rounds(uint32_t x[16]) {
    for (int i = 0; i < 10; i++) {
        QR(x[0], x[4], x[8], x[12]);
        QR(x[1], x[5], x[9], x[13]);
        QR(x[2], x[6], x[10], x[14]);
        QR(x[3], x[7], x[11], x[15]);
        QR(x[0], x[5], x[10], x[15]);
        QR(x[1], x[6], x[11], x[12]);
        QR(x[2], x[7], x[8], x[13]);
        QR(x[3], x[4], x[9], x[14]);
    }
}

static const uint32_t SIGMA[4] = {0x61707865u, 0x3320646eu, 0x79622d32u, 0x6b206574u};

/* RFC 8439 2.3: one 64-byte ChaCha20 block, 32-bit counter, 96-bit nonce. */
void ssref_chacha20_block(const uint8_t key[32], uint32_t counter, const uint8_t nonce[12], uint8_t out[64]) {
    uint32_t in[16], x[16];
    for (int i = 0; i < 4; i++) in[i] = SIGMA[i];
    for (int i = 0; i < 8; i++) in[4 + i] = ld32(key + 4 * i);
    in[12] = counter;
    for (int i = 0; i < 3; i++) in[13 + i] = ld32(nonce + 4 * i);
    memcpy(x, in, sizeof x);
    chacha_rounds(x);
    for (int i = 0; i < 16; i++) st32(out + 4 * i, x[i] + in[i]);
}

/* draft-irtf-cfrg-xchacha-03 2.2: HChaCha20(key, 16-byte input) -> 32 bytes. */
void ssref_hchacha20(const uint8_t key[32], const uint8_t in16[16], uint8_t out[32]) {
    uint32_t x[16];
    for (int i = 0; i < 4; i++) x[i] = SIGMA[i];
    for (int i = 0; i < 8; i++) x[4 + i] = ld32(key + 4 * i);
    for (int i = 0; i < 4; i++) x[12 + i] = ld32(in16 + 4 * i);
    chacha_rounds(x);
    for (int i = 0; i < 4; i++) st32(out + 4 * i, x[i]);
    for (int i = 0; i < 4; i++) st32(out + 16 + 4 * i, x[12 + i]);
}

/* crypto_stream_chacha20_ietf_xor_ic: out = in ^ keystream starting at block `ic`.
 * `in` may be NULL (pure keystream). */
void ssref_chacha20_ietf_xor_ic(uint8_t *out, const uint8_t *in, uint64_t len, const uint8_t nonce[12],
                                uint32_t ic, const uint8_t key[32]) {
    uint8_t ks[64];
    uint64_t off = 0;
    while (off < len) {
        ssref_chacha20_block(key, ic++, nonce, ks);
        uint64_t n = len - off < 64 ? len - off : 64;
        for (uint64_t i = 0; i < n; i++) out[off + i] = (in ? in[off + i] : 0) ^ ks[i];
        off += n;
    }
}

/* ---- Poly1305 (RFC 8439 2.5), incremental, 26-bit limbs ------------------ */
typedef struct {
    uint32_t r[5], h[5], pad[4];
    uint8_t buf[16];
    size_t fill;
} ssref_poly;

static void poly_blocks(ssref_poly *st, const uint8_t *m, size_t n, uint32_t hibit) {
    const uint32_t r0 = st->r[0], r1 = st->r[1], r2 = st->r[2], r3 = st->r[3], r4 = st->r[4];
    const uint32_t s1 = r1 * 5, s2 = r2 * 5, s3 = r3 * 5, s4 = r4 * 5;
    uint32_t h0 = st->h[0], h1 = st->h[1], h2 = st->h[2], h3 = st->h[3], h4 = st->h[4];
    while (n >= 16) {
        h0 += ld32(m) & 0x3ffffff;
        h1 += (ld32(m + 3) >> 2) & 0x3ffffff;
        h2 += (ld32(m + 6) >> 4) & 0x3ffffff;
        h3 += (ld32(m + 9) >> 6) & 0x3ffffff;
        h4 += (ld32(m + 12) >> 8) | hibit;
        uint64_t d0 = (uint64_t)h0 * r0 + (uint64_t)h1 * s4 + (uint64_t)h2 * s3 + (uint64_t)h3 * s2 + (uint64_t)h4 * s1;
        uint64_t d1 = (uint64_t)h0 * r1 + (uint64_t)h1 * r0 + (uint64_t)h2 * s4 + (uint64_t)h3 * s3 + (uint64_t)h4 * s2;
        uint64_t d2 = (uint64_t)h0 * r2 + (uint64_t)h1 * r1 + (uint64_t)h2 * r0 + (uint64_t)h3 * s4 + (uint64_t)h4 * s3;
        uint64_t d3 = (uint64_t)h0 * r3 + (uint64_t)h1 * r2 + (uint64_t)h2 * r1 + (uint64_t)h3 * r0 + (uint64_t)h4 * s4;
        uint64_t d4 = (uint64_t)h0 * r4 + (uint64_t)h1 * r3 + (uint64_t)h2 * r2 + (uint64_t)h3 * r1 + (uint64_t)h4 * r0;
        uint32_t c;
        c = (uint32_t)(d0 >> 26); h0 = (uint32_t)d0 & 0x3ffffff; d1 += c;
        c = (uint32_t)(d1 >> 26); h1 = (uint32_t)d1 & 0x3ffffff; d2 += c;
        c = (uint32_t)(d2 >> 26); h2 = (uint32_t)d2 & 0x3ffffff; d3 += c;
        c = (uint32_t)(d3 >> 26); h3 = (uint32_t)d3 & 0x3ffffff; d4 += c;
        c = (uint32_t)(d4 >> 26); h4 = (uint32_t)d4 & 0x3ffffff; h0 += c * 5;
        c = h0 >> 26; h0 &= 0x3ffffff; h1 += c;
        m += 16; n -= 16;
    }
    st->h[0] = h0; st->h[1] = h1; st->h[2] = h2; st->h[3] = h3; st->h[4] = h4;
}

void ssref_poly_init(ssref_poly *st, const uint8_t key[32]) {
    st->r[0] = ld32(key) & 0x3ffffff;
    st->r[1] = (ld32(key + 3) >> 2) & 0x3ffff03;
    st->r[2] = (ld32(key + 6) >> 4) & 0x3ffc0ff;
    st->r[3] = (ld32(key + 9) >> 6) & 0x3f03fff;
    st->r[4] = (ld32(key + 12) >> 8) & 0x00fffff;
    for (int i = 0; i < 5; i++) st->h[i] = 0;
    for (int i = 0; i < 4; i++) st->pad[i] = ld32(key + 16 + 4 * i);
    st->fill = 0;
}

void ssref_poly_update(ssref_poly *st, const uint8_t *m, size_t n) {
    if (st->fill) {
        size_t want = 16 - st->fill;
        if (want > n) want = n;
        memcpy(st->buf + st->fill, m, want);
        st->fill += want; m += want; n -= want;
        if (st->fill < 16) return;
        poly_blocks(st, st->buf, 16, 1u << 24);
        st->fill = 0;
    }
    if (n >= 16) {
        size_t full = n & ~(size_t)15;
        poly_blocks(st, m, full, 1u << 24);
        m += full; n -= full;
    }
    if (n) { memcpy(st->buf, m, n); st->fill = n; }
}

void ssref_poly_final(ssref_poly *st, uint8_t mac[16]) {
    if (st->fill) {
        st->buf[st->fill] = 1;
        for (size_t i = st->fill + 1; i < 16; i++) st->buf[i] = 0;
        poly_blocks(st, st->buf, 16, 0);
    }
    uint32_t h0 = st->h[0], h1 = st->h[1], h2 = st->h[2], h3 = st->h[3], h4 = st->h[4], c;
    c = h1 >> 26; h1 &= 0x3ffffff; h2 += c;
    c = h2 >> 26; h2 &= 0x3ffffff; h3 += c;
    c = h3 >> 26; h3 &= 0x3ffffff; h4 += c;
    c = h4 >> 26; h4 &= 0x3ffffff; h0 += c * 5;
    c = h0 >> 26; h0 &= 0x3ffffff; h1 += c;
    /* g = h + 5 - 2^130 ; select g if h >= p */
    uint32_t g0 = h0 + 5; c = g0 >> 26; g0 &= 0x3ffffff;
    uint32_t g1 = h1 + c; c = g1 >> 26; g1 &= 0x3ffffff;
    uint32_t g2 = h2 + c; c = g2 >> 26; g2 &= 0x3ffffff;
    uint32_t g3 = h3 + c; c = g3 >> 26; g3 &= 0x3ffffff;
    uint32_t g4 = h4 + c - (1u << 26);
    uint32_t mask = (g4 >> 31) - 1; /* all-ones if g4 did not go negative */
    h0 = (h0 & ~mask) | (g0 & mask);
    h1 = (h1 & ~mask) | (g1 & mask);
    h2 = (h2 & ~mask) | (g2 & mask);
    h3 = (h3 & ~mask) | (g3 & mask);
    h4 = (h4 & ~mask) | (g4 & mask);
    uint32_t w0 = h0 | (h1 << 26), w1 = (h1 >> 6) | (h2 << 20), w2 = (h2 >> 12) | (h3 << 14), w3 = (h3 >> 18) | (h4 << 8);
    uint64_t f;
    f = (uint64_t)w0 + st->pad[0]; st32(mac, (uint32_t)f);
    f = (uint64_t)w1 + st->pad[1] + (f >> 32); st32(mac + 4, (uint32_t)f);
    f = (uint64_t)w2 + st->pad[2] + (f >> 32); st32(mac + 8, (uint32_t)f);
    f = (uint64_t)w3 + st->pad[3] + (f >> 32); st32(mac + 12, (uint32_t)f);
}

void ssref_poly1305(uint8_t mac[16], const uint8_t *m, size_t n, const uint8_t key[32]) {
    ssref_poly st;
    ssref_poly_init(&st, key);
    ssref_poly_update(&st, m, n);
    ssref_poly_final(&st, mac);
}

/* ---- secretstream state machine (SURVEY 8(a) spec box) -------------------
 * state layout (52 B): k[32] | nonce[12] = LE32 counter ‖ inonce[8] | pad[8]=0 */
void ssref_init(uint8_t state[SS_STATEBYTES], const uint8_t header[SS_HEADERBYTES], const uint8_t key[SS_KEYBYTES]) {
    ssref_hchacha20(key, header, state);
    st32(state + 32, 1);
    memcpy(state + 36, header + 16, 8);
    memset(state + 44, 0, 8);
}

void ssref_rekey(uint8_t state[SS_STATEBYTES]) {
    uint8_t t[40];
    memcpy(t, state, 32);
    memcpy(t + 32, state + 36, 8);
    ssref_chacha20_ietf_xor_ic(t, t, 40, state + 32, 0, state);
    memcpy(state, t, 32);
    memcpy(state + 36, t + 32, 8);
    st32(state + 32, 1);
}

static void ss_mac(uint8_t mac[16], const uint8_t polykey[32], const uint8_t *ad, uint64_t adlen,
                   const uint8_t block[64], const uint8_t *c, uint64_t mlen) {
    static const uint8_t zeros[16] = {0};
    uint8_t lens[16];
    ssref_poly st;
    ssref_poly_init(&st, polykey);
    ssref_poly_update(&st, ad, (size_t)adlen);
    ssref_poly_update(&st, zeros, (size_t)((0x10 - adlen) & 0xf));
    ssref_poly_update(&st, block, 64);
    ssref_poly_update(&st, c, (size_t)mlen);
    /* libsodium quirk: pad length is (0x10 - 64 + mlen) & 0xf == mlen mod 16 */
    ssref_poly_update(&st, zeros, (size_t)((0x10 - 64 + mlen) & 0xf));
    st64(lens, adlen);
    st64(lens + 8, 64 + mlen);
    ssref_poly_update(&st, lens, 16);
    ssref_poly_final(&st, mac);
}

static void ss_advance(uint8_t state[SS_STATEBYTES], const uint8_t mac[16], uint8_t tag) {
    for (int i = 0; i < 8; i++) state[36 + i] ^= mac[i];
    uint32_t ctr = ld32(state + 32) + 1;
    st32(state + 32, ctr);
    if ((tag & SS_TAG_REKEY) || ctr == 0) ssref_rekey(state);
}

/* out must hold mlen + 17 bytes.  Returns 0. */
int ssref_push(uint8_t state[SS_STATEBYTES], uint8_t *out, const uint8_t *m, uint64_t mlen, const uint8_t *ad,
               uint64_t adlen, uint8_t tag) {
    uint8_t polykey[64], block[64], mac[16];
    ssref_chacha20_ietf_xor_ic(polykey, NULL, 64, state + 32, 0, state);
    memset(block, 0, 64);
    block[0] = tag;
    ssref_chacha20_ietf_xor_ic(block, block, 64, state + 32, 1, state);
    out[0] = block[0];
    ssref_chacha20_ietf_xor_ic(out + 1, m, mlen, state + 32, 2, state);
    ss_mac(mac, polykey, ad, adlen, block, out + 1, mlen);
    memcpy(out + 1 + mlen, mac, 16);
    ss_advance(state, mac, tag);
    return 0;
}

/* m must hold inlen - 17 bytes.  Returns 0, or -1 (nothing modified) on a short
 * frame or MAC mismatch. */
int ssref_pull(uint8_t state[SS_STATEBYTES], uint8_t *m, uint8_t *tag_out, const uint8_t *in, uint64_t inlen,
               const uint8_t *ad, uint64_t adlen) {
    if (inlen < SS_ABYTES) return -1;
    uint64_t mlen = inlen - SS_ABYTES;
    uint8_t polykey[64], block[64], mac[16];
    ssref_chacha20_ietf_xor_ic(polykey, NULL, 64, state + 32, 0, state);
    memset(block, 0, 64);
    block[0] = in[0];
    ssref_chacha20_ietf_xor_ic(block, block, 64, state + 32, 1, state);
    uint8_t tag = block[0];
    block[0] = in[0];
    ss_mac(mac, polykey, ad, adlen, block, in + 1, mlen);
    uint8_t diff = 0;
    for (int i = 0; i < 16; i++) diff |= mac[i] ^ in[1 + mlen + i];
    if (diff) return -1;
    ssref_chacha20_ietf_xor_ic(m, in + 1, mlen, state + 32, 2, state);
    ss_advance(state, mac, tag);
    if (tag_out) *tag_out = tag;
    return 0;
}

/* ---- framing (@hyperswarm/secret-stream 6.6.1 `_write`) ---
 * wire = uint24-LE(mlen + 17) ‖ push(m).  wire must hold mlen + 20 bytes. */
int ssref_seal_frame(uint8_t state[SS_STATEBYTES], uint8_t *wire, const uint8_t *m, uint32_t mlen, uint8_t tag) {
    uint32_t n = mlen + SS_ABYTES;
    if (n > 0xffffffu) return -1;
    wire[0] = (uint8_t)n; wire[1] = (uint8_t)(n >> 8); wire[2] = (uint8_t)(n >> 16);
    return ssref_push(state, wire + 3, m, mlen, NULL, 0, tag);
}

int ssref_open_frame(uint8_t state[SS_STATEBYTES], uint8_t *m, uint8_t *tag_out, const uint8_t *wire, uint32_t wirelen) {
    if (wirelen < 3) return -1;
    uint32_t n = (uint32_t)wire[0] | ((uint32_t)wire[1] << 8) | ((uint32_t)wire[2] << 16);
    if (n + 3 != wirelen) return -1;
    return ssref_pull(state, m, tag_out, wire + 3, n, NULL, 0);
}

/* ---- batch drivers used by bench.py's cpu_baseline "port" leg ------------
 * One frame per stream, frames laid out with fixed strides. */
void ssref_seal_batch(uint8_t *states, uint32_t n, uint8_t *wire, uint32_t wire_stride, const uint8_t *plain,
                      uint32_t plain_stride, const uint32_t *lens) {
    for (uint32_t i = 0; i < n; i++)
        ssref_seal_frame(states + (size_t)i * SS_STATEBYTES, wire + (size_t)i * wire_stride,
                         plain + (size_t)i * plain_stride, lens[i], 0);
}

int ssref_open_batch(uint8_t *states, uint32_t n, uint8_t *plain, uint32_t plain_stride, const uint8_t *wire,
                     uint32_t wire_stride, const uint32_t *lens) {
    int bad = 0;
    for (uint32_t i = 0; i < n; i++) {
        uint8_t tag;
        if (ssref_open_frame(states + (size_t)i * SS_STATEBYTES, plain + (size_t)i * plain_stride, &tag,
                             wire + (size_t)i * wire_stride, lens[i] + 20))
            bad++;
    }
    return bad;
}
