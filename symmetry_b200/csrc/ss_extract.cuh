// ss_extract.cuh -- the reference's per-chunk delta extraction as one thread-local pass.
//
// Restates, with ECMAScript semantics, what the relay loop computes before
// every `peer.write(chunk)` (/root/reference/src/provider.ts:243-246):
//     completion += getChatDataFromProvider(provider, safeParseStreamResponse(chunk.toString()))
//   src/utils.ts:16-18  startsWith('data:')
//   src/utils.ts:20-31  JSON.parse(s.split('data:')[1]) | JSON.parse(s), any throw -> undefined
//   src/utils.ts:33-52  provider switch (src/constants.ts:22-29) with its optional chains
// One thread handles one chunk: a strict JSON pushdown automaton (ECMA-404
// grammar, whitespace SP/TAB/LF/CR only, no raw control characters in strings,
// duplicate keys -> last wins) that tracks just the values the three provider
// branches read -- root, root.choices, choices[0], .delta, .content and
// root.content -- then applies the optional-chain / truthiness / `+=` coercion
// rules.  Outcomes: OK + WTF-8 delta bytes, THROW (the reference raises a
// TypeError here, so the chunk must not be sealed), or UNSUPPORTED for value
// shapes whose ToString is not worth a GPU implementation (numbers with an
// exponent or > 15 digits, non-empty arrays, nesting deeper than 64): the host
// then runs the reference's own helper for that chunk.  Checked against
// oracle/extract.py on tests/golden/extract_cases.json.
#pragma once
#include <stdint.h>

#include <initializer_list>

#ifndef SS_DEV
#define SS_DEV static inline
#endif

#define SX_OK 0u
#define SX_THROW 3u
#define SX_UNSUPPORTED 4u

#define SX_KIND_DEFAULT 0u
#define SX_KIND_OLLAMA 1u
#define SX_KIND_LLAMACPP 2u

namespace sx {

enum : uint32_t { V_ABSENT = 0, V_UNDEF, V_NULL, V_FALSE, V_TRUE, V_NUMBER, V_STRING, V_OBJECT, V_ARRAY };
enum : uint32_t { NF_FRAC = 1, NF_EXP = 2, NF_NONZERO = 4, AF_EMPTY = 8 };
enum : uint32_t { KEY_NONE = 0, KEY_CHOICES, KEY_CONTENT, KEY_DELTA, KEY_ZERO };
enum : uint32_t { T_ROOT = 0, T_CHOICES, T_ELEM0, T_DELTA, T_CONTENT, T_RCONTENT, T_NONE };

struct Val {
    uint32_t kind, s, e, flags;
};

SS_DEV bool is_ws(uint32_t c) { return c == 0x20 || c == 0x09 || c == 0x0a || c == 0x0d; }
SS_DEV bool is_digit(uint32_t c) { return c - 0x30u < 10u; }
SS_DEV int hexval(uint32_t c) {
    if (c - 0x30u < 10u) return (int)(c - 0x30u);
    c |= 0x20u;
    if (c - 0x61u < 6u) return (int)(c - 0x61u + 10u);
    return -1;
}
SS_DEV uint32_t byte_at(const uint8_t *p, uint32_t i) { return p[i]; }

// names compared against decoded key characters
SS_DEV uint32_t name_char(uint32_t which, uint32_t pos) {
    // "choices", "content", "delta", "0" packed little-endian
    const uint64_t n = which == KEY_CHOICES ? 0x0073656369'6f6863ull
                     : which == KEY_CONTENT ? 0x00746e65'746e6f63ull
                     : which == KEY_DELTA   ? 0x000000'61746c6564ull
                                            : 0x30ull;
    return pos < 8 ? (uint32_t)(n >> (8 * pos)) & 0xffu : 0u;
}
SS_DEV uint32_t name_len(uint32_t which) {
    return which == KEY_CHOICES ? 7u : which == KEY_CONTENT ? 7u : which == KEY_DELTA ? 5u : 1u;
}

struct Out {
    uint8_t *p;
    uint32_t n, cap;
    bool overflow;
};
SS_DEV void put(Out &o, uint32_t b) {
    if (o.n < o.cap) o.p[o.n] = (uint8_t)b;
    else o.overflow = true;
    o.n++;
}
SS_DEV void put_cp(Out &o, uint32_t cp) {  // generalized UTF-8 (lone surrogates allowed: WTF-8)
    if (cp < 0x80) put(o, cp);
    else if (cp < 0x800) { put(o, 0xc0 | (cp >> 6)); put(o, 0x80 | (cp & 0x3f)); }
    else if (cp < 0x10000) { put(o, 0xe0 | (cp >> 12)); put(o, 0x80 | ((cp >> 6) & 0x3f)); put(o, 0x80 | (cp & 0x3f)); }
    else { put(o, 0xf0 | (cp >> 18)); put(o, 0x80 | ((cp >> 12) & 0x3f)); put(o, 0x80 | ((cp >> 6) & 0x3f)); put(o, 0x80 | (cp & 0x3f)); }
}
SS_DEV void put_lit(Out &o, const char *s) {
    for (int i = 0; s[i]; i++) put(o, (uint8_t)s[i]);
}

// Decodes a JSON string body text[s,e) (already validated by the scan) the way
// `chunk.toString()` + JSON.parse see it: UTF-8 with U+FFFD per maximal invalid
// subpart, escapes, \uD8xx\uDCxx pairs joined.  Emits WTF-8; returns whether the
// decoded string === 'undefined' (src/utils.ts:46).
SS_DEV bool decode_string(const uint8_t *t, uint32_t s, uint32_t e, Out &o) {
    const char *und = "undefined";
    uint32_t um = 0;       // matched prefix length of "undefined", 0xff = mismatch
    uint32_t i = s;
    while (i < e) {
        uint32_t c = t[i];
        uint32_t cp;
        if (c == 0x5c) {  // backslash
            const uint32_t x = t[i + 1];
            i += 2;
            if (x == 'u') {
                cp = ((uint32_t)hexval(t[i]) << 12) | ((uint32_t)hexval(t[i + 1]) << 8) |
                     ((uint32_t)hexval(t[i + 2]) << 4) | (uint32_t)hexval(t[i + 3]);
                i += 4;
                if (cp - 0xd800u < 0x400u && i + 6 <= e && t[i] == 0x5c && t[i + 1] == 'u') {
                    const uint32_t lo = ((uint32_t)hexval(t[i + 2]) << 12) | ((uint32_t)hexval(t[i + 3]) << 8) |
                                        ((uint32_t)hexval(t[i + 4]) << 4) | (uint32_t)hexval(t[i + 5]);
                    if (lo - 0xdc00u < 0x400u) {
                        cp = 0x10000u + ((cp - 0xd800u) << 10) + (lo - 0xdc00u);
                        i += 6;
                    }
                }
            } else {
                cp = x == 'b' ? 8u : x == 'f' ? 12u : x == 'n' ? 10u : x == 'r' ? 13u : x == 't' ? 9u : x;
            }
        } else if (c < 0x80) {
            cp = c;
            i++;
        } else {
            // WHATWG UTF-8 decoder: lower/upper bound of the next continuation byte, bytes still needed
            uint32_t need = 0, lo = 0x80, hi = 0xbf;
            cp = 0xfffd;
            if (c >= 0xc2 && c <= 0xdf) { need = 1; cp = c & 0x1f; }
            else if (c >= 0xe0 && c <= 0xef) { need = 2; cp = c & 0x0f; if (c == 0xe0) lo = 0xa0; if (c == 0xed) hi = 0x9f; }
            else if (c >= 0xf0 && c <= 0xf4) { need = 3; cp = c & 0x07; if (c == 0xf0) lo = 0x90; if (c == 0xf4) hi = 0x8f; }
            i++;
            bool bad = need == 0;
            while (need) {
                if (i >= e) { bad = true; break; }
                const uint32_t b = t[i];
                if (b < lo || b > hi) { bad = true; break; }  // offending byte is reprocessed as a new start
                lo = 0x80; hi = 0xbf;
                cp = (cp << 6) | (b & 0x3f);
                i++;
                need--;
            }
            if (bad) cp = 0xfffd;
        }
        put_cp(o, cp);
        if (um != 0xff) um = (um < 9 && cp == (uint32_t)(uint8_t)und[um]) ? um + 1 : 0xff;
    }
    return um == 9;
}

// Number::toString for the subset the GPU formats itself: -?int[.frac] with at
// most 15 significant digits and no exponent (such decimals round-trip exactly).
SS_DEV bool emit_number(const uint8_t *t, const Val &v, Out &o) {
    if (!(v.flags & NF_NONZERO)) { put(o, '0'); return true; }   // 0, -0, 0.0, 0e5 ... all print "0"
    if (v.flags & NF_EXP) return false;
    uint32_t i = v.s;
    const bool neg = t[i] == '-';
    if (neg) i++;
    uint32_t dot = v.e;
    for (uint32_t j = i; j < v.e; j++) if (t[j] == '.') dot = j;
    uint32_t fe = v.e;                       // strip trailing fraction zeros
    if (dot < v.e) { while (fe > dot + 1 && t[fe - 1] == '0') fe--; if (fe == dot + 1) fe = dot; }
    // significant digits
    uint32_t sig = 0;
    bool lead = true;
    uint32_t frac_lead_zeros = 0;
    for (uint32_t j = i; j < fe; j++) {
        if (t[j] == '.') continue;
        if (lead && t[j] == '0') { if (j > dot) frac_lead_zeros++; continue; }
        lead = false;
        sig++;
    }
    if (sig > 15) return false;
    const bool int_zero = (dot - i == 1) && t[i] == '0';
    if (int_zero && frac_lead_zeros > 5) return false;            // < 1e-6 prints in exponent form
    if (neg) put(o, '-');
    for (uint32_t j = i; j < fe; j++) put(o, t[j]);
    return true;
}

// ---- the automaton ----------------------------------------------------------------------------
// One table look-up per input byte: T[state][byte] -> next state (bits 0..5) | action (bits 6..10).
// Actions fire only at structural characters and at the first/last character of a value, so the
// inner loop is the same handful of instructions for every lane whatever its chunk contains (the
// first version was a switch over ~20 states: the lanes of a warp, each in its own JSON position,
// serialised the switch arms and the kernel took 72 us for 16,384 chunks).
enum : uint32_t {
    S_VALUE = 0, S_ARR_FIRST, S_AFTER, S_OBJ_FIRST, S_OBJ_KEY, S_COLON,
    S_STR, S_ESC, S_U1, S_U2, S_U3, S_U4,
    S_KSTR, S_KESC, S_KU1, S_KU2, S_KU3, S_KU4,
    S_NMINUS, S_NZERO, S_NINT, S_NDOT, S_NFRAC, S_NE, S_NESIGN, S_NEXP,
    S_T1, S_T2, S_T3, S_F1, S_F2, S_F3, S_F4, S_N1, S_N2, S_N3,
    S_ERR, S_COUNT
};
enum : uint32_t {
    A_NONE = 0, A_OPEN_OBJ, A_OPEN_ARR, A_STR_START, A_NUM_START, A_LIT_T, A_LIT_F, A_LIT_N,
    A_KEY_START, A_KEY_ESC, A_KEY_END, A_STR_END, A_CLOSE_OBJ, A_CLOSE_ARR, A_CLOSE_ARR_EMPTY, A_COMMA,
    A_NUM_END, A_NUM_END_COMMA, A_NUM_END_CLOSE_OBJ, A_NUM_END_CLOSE_ARR
};
#define SX_TABLE_ENTRIES (sx::S_COUNT * 256u)

// Fills T (SX_TABLE_ENTRIES uint16) with the ECMA-404 grammar.  Host code: the CUDA library builds it
// once per device and the kernel stages it in shared memory.
static inline void build_table(uint16_t *T) {
    auto set = [&](uint32_t st, uint32_t c, uint32_t next, uint32_t act) { T[st * 256u + c] = (uint16_t)(next | (act << 6)); };
    auto is_hex = [](uint32_t c) { return (c >= '0' && c <= '9') || ((c | 0x20u) >= 'a' && (c | 0x20u) <= 'f'); };
    for (uint32_t st = 0; st < S_COUNT; st++)
        for (uint32_t c = 0; c < 256; c++) set(st, c, S_ERR, A_NONE);
    const uint8_t ws[4] = {0x20, 0x09, 0x0a, 0x0d};
    for (uint32_t st : {(uint32_t)S_VALUE, (uint32_t)S_ARR_FIRST}) {
        for (uint8_t w : ws) set(st, w, st, A_NONE);
        set(st, '{', S_OBJ_FIRST, A_OPEN_OBJ);
        set(st, '[', S_ARR_FIRST, A_OPEN_ARR);
        set(st, '"', S_STR, A_STR_START);
        set(st, '-', S_NMINUS, A_NUM_START);
        set(st, '0', S_NZERO, A_NUM_START);
        for (uint32_t c = '1'; c <= '9'; c++) set(st, c, S_NINT, A_NUM_START);
        set(st, 't', S_T1, A_LIT_T);
        set(st, 'f', S_F1, A_LIT_F);
        set(st, 'n', S_N1, A_LIT_N);
    }
    set(S_ARR_FIRST, ']', S_AFTER, A_CLOSE_ARR_EMPTY);
    for (uint8_t w : ws) set(S_AFTER, w, S_AFTER, A_NONE);
    set(S_AFTER, ',', S_VALUE, A_COMMA);
    set(S_AFTER, '}', S_AFTER, A_CLOSE_OBJ);
    set(S_AFTER, ']', S_AFTER, A_CLOSE_ARR);
    for (uint32_t st : {(uint32_t)S_OBJ_FIRST, (uint32_t)S_OBJ_KEY}) {
        for (uint8_t w : ws) set(st, w, st, A_NONE);
        set(st, '"', S_KSTR, A_KEY_START);
    }
    set(S_OBJ_FIRST, '}', S_AFTER, A_CLOSE_OBJ);
    for (uint8_t w : ws) set(S_COLON, w, S_COLON, A_NONE);
    set(S_COLON, ':', S_VALUE, A_NONE);
    // strings (value strings and key strings have the same grammar; keys report their end differently)
    for (int k = 0; k < 2; k++) {
        const uint32_t str = k ? S_KSTR : S_STR, esc = k ? S_KESC : S_ESC, u1 = k ? S_KU1 : S_U1;
        for (uint32_t c = 0x20; c < 256; c++) set(str, c, str, A_NONE);
        set(str, '"', k ? S_COLON : S_AFTER, k ? A_KEY_END : A_STR_END);
        set(str, 0x5c, esc, k ? A_KEY_ESC : A_NONE);
        for (char c : {'"', '\\', '/', 'b', 'f', 'n', 'r', 't'}) set(esc, (uint8_t)c, str, A_NONE);
        set(esc, 'u', u1, A_NONE);
        for (uint32_t j = 0; j < 4; j++)
            for (uint32_t c = 0; c < 256; c++)
                if (is_hex(c)) set(u1 + j, c, j == 3 ? str : u1 + j + 1, A_NONE);
    }
    // numbers
    set(S_NMINUS, '0', S_NZERO, A_NONE);
    for (uint32_t c = '1'; c <= '9'; c++) set(S_NMINUS, c, S_NINT, A_NONE);
    for (uint32_t c = '0'; c <= '9'; c++) {
        set(S_NINT, c, S_NINT, A_NONE);
        set(S_NDOT, c, S_NFRAC, A_NONE);
        set(S_NFRAC, c, S_NFRAC, A_NONE);
        set(S_NE, c, S_NEXP, A_NONE);
        set(S_NESIGN, c, S_NEXP, A_NONE);
        set(S_NEXP, c, S_NEXP, A_NONE);
    }
    set(S_NZERO, '.', S_NDOT, A_NONE);
    set(S_NINT, '.', S_NDOT, A_NONE);
    for (uint32_t st : {(uint32_t)S_NZERO, (uint32_t)S_NINT, (uint32_t)S_NFRAC}) { set(st, 'e', S_NE, A_NONE); set(st, 'E', S_NE, A_NONE); }
    set(S_NE, '+', S_NESIGN, A_NONE);
    set(S_NE, '-', S_NESIGN, A_NONE);
    for (uint32_t st : {(uint32_t)S_NZERO, (uint32_t)S_NINT, (uint32_t)S_NFRAC, (uint32_t)S_NEXP}) {   // a number ends at a delimiter
        for (uint8_t w : ws) set(st, w, S_AFTER, A_NUM_END);
        set(st, ',', S_VALUE, A_NUM_END_COMMA);
        set(st, '}', S_AFTER, A_NUM_END_CLOSE_OBJ);
        set(st, ']', S_AFTER, A_NUM_END_CLOSE_ARR);
    }
    // literals
    set(S_T1, 'r', S_T2, A_NONE); set(S_T2, 'u', S_T3, A_NONE); set(S_T3, 'e', S_AFTER, A_NONE);
    set(S_F1, 'a', S_F2, A_NONE); set(S_F2, 'l', S_F3, A_NONE); set(S_F3, 's', S_F4, A_NONE); set(S_F4, 'e', S_AFTER, A_NONE);
    set(S_N1, 'u', S_N2, A_NONE); set(S_N2, 'l', S_N3, A_NONE); set(S_N3, 'l', S_AFTER, A_NONE);
}

// Which of the four names the path reads does the key text[s,e) spell?  Escapes are legal in keys
// ("choices"), so a key that contains a backslash is decoded character by character.
SS_DEV uint32_t match_key(const uint8_t *t, uint32_t s, uint32_t e, bool has_esc) {
    if (!has_esc) {
        const uint32_t len = e - s;
        if (len != 1u && len != 5u && len != 7u) return KEY_NONE;
        uint32_t cand = len == 7u ? ((1u << KEY_CHOICES) | (1u << KEY_CONTENT)) : len == 5u ? (1u << KEY_DELTA) : (1u << KEY_ZERO);
        for (uint32_t i = 0; i < len; i++) {
            const uint32_t c = t[s + i];
#pragma unroll
            for (uint32_t k = KEY_CHOICES; k <= KEY_ZERO; k++)
                if (name_char(k, i) != c) cand &= ~(1u << k);
        }
#pragma unroll
        for (uint32_t k = KEY_CHOICES; k <= KEY_ZERO; k++)
            if ((cand >> k) & 1u) return k;
        return KEY_NONE;
    }
    uint32_t cand = 0x1e, kpos = 0, i = s;
    while (i < e) {
        uint32_t c = t[i++];
        if (c == 0x5c) {
            const uint32_t x = t[i++];
            if (x == 'u') {
                c = ((uint32_t)hexval(t[i]) << 12) | ((uint32_t)hexval(t[i + 1]) << 8) | ((uint32_t)hexval(t[i + 2]) << 4) |
                    (uint32_t)hexval(t[i + 3]);
                i += 4;
                if (c == 0) c = 0xffffu;   // never equal to the zero padding of a name
            } else {
                c = x == 'b' ? 8u : x == 'f' ? 12u : x == 'n' ? 10u : x == 'r' ? 13u : x == 't' ? 9u : x;
            }
        }
#pragma unroll
        for (uint32_t k = KEY_CHOICES; k <= KEY_ZERO; k++)
            if (name_char(k, kpos) != c) cand &= ~(1u << k);
        kpos++;
    }
#pragma unroll
    for (uint32_t k = KEY_CHOICES; k <= KEY_ZERO; k++)
        if (((cand >> k) & 1u) && kpos == name_len(k)) return k;
    return KEY_NONE;
}

SS_DEV uint32_t number_flags(const uint8_t *t, uint32_t s, uint32_t e) {
    uint32_t f = 0;
    bool mant = true;
    for (uint32_t i = s; i < e; i++) {
        const uint32_t c = t[i];
        if (c == '.') f |= NF_FRAC;
        else if (c == 'e' || c == 'E') { f |= NF_EXP; mant = false; }
        else if (mant && c >= '1' && c <= '9') f |= NF_NONZERO;
    }
    return f;
}

struct Result {
    uint32_t status, len;
};

// text = chunk bytes, n = length, kind = SX_KIND_*, out = at least `cap` writable bytes, T = build_table()
SS_DEV Result extract_chunk(const uint8_t *text, uint32_t n, uint32_t kind, uint8_t *out, uint32_t cap, const uint16_t *T) {
    // ---- utils.ts:16-27: prefix test and split('data:')[1] ------------------
    uint32_t pos = 0, end = n;
    const bool prefixed = n >= 5 && text[0] == 'd' && text[1] == 'a' && text[2] == 't' && text[3] == 'a' && text[4] == ':';
    if (prefixed) pos = 5;

    Val tv[6];
#pragma unroll
    for (int i = 0; i < 6; i++) { tv[i].kind = V_ABSENT; tv[i].s = tv[i].e = tv[i].flags = 0; }
    uint32_t state = S_VALUE, depth = 0;
    uint64_t stack = 0;                  // bit d-1: container at depth d is an object
    uint32_t lastkey = KEY_NONE;
    bool act2 = false, act3 = false, act4 = false, first2 = false;
    bool key_esc = false;
    uint32_t cur = T_NONE;               // tracked slot of the string/number being scanned
    uint32_t last_started = T_NONE;      // tracked slot of the most recently started value (array emptiness)
    uint32_t vstart = 0;
    bool too_deep = false;

    for (; pos < end; pos++) {
        const uint32_t c = text[pos];
        if (c == 'd' && prefixed && pos + 5 <= n && text[pos + 1] == 'a' && text[pos + 2] == 't' && text[pos + 3] == 'a' &&
            text[pos + 4] == ':') {
            end = pos;                   // the second 'data:' ends split()[1], wherever it falls
            break;
        }
        const uint32_t ent = T[state * 256u + c];
        uint32_t next = ent & 63u;
        uint32_t act = ent >> 6;
        if (act != A_NONE) {
            // a number that ends here first closes itself, then the delimiter acts as it would after any value
            if (act >= A_NUM_END) {
                if (cur != T_NONE) { tv[cur].e = pos; tv[cur].flags = number_flags(text, tv[cur].s, pos); }
                act = act == A_NUM_END ? A_NONE : act == A_NUM_END_COMMA ? A_COMMA : act == A_NUM_END_CLOSE_OBJ ? A_CLOSE_OBJ : A_CLOSE_ARR;
            }
            if (act >= A_OPEN_OBJ && act <= A_LIT_N) {
                // ---- a value starts: is it one of the values the provider branches read? ----
                const uint32_t slot =
                    depth == 0 ? T_ROOT
                    : depth == 1 && ((stack >> 0) & 1) && lastkey == KEY_CHOICES ? T_CHOICES
                    : depth == 1 && ((stack >> 0) & 1) && lastkey == KEY_CONTENT ? T_RCONTENT
                    : depth == 2 && act2 && (((stack >> 1) & 1) ? lastkey == KEY_ZERO : first2) ? T_ELEM0
                    : depth == 3 && act3 && lastkey == KEY_DELTA ? T_DELTA
                    : depth == 4 && act4 && lastkey == KEY_CONTENT ? T_CONTENT
                    : T_NONE;
                const uint32_t vk = act == A_OPEN_OBJ ? V_OBJECT : act == A_OPEN_ARR ? V_ARRAY : act == A_STR_START ? V_STRING
                                  : act == A_NUM_START ? V_NUMBER : act == A_LIT_T ? V_TRUE : act == A_LIT_F ? V_FALSE : V_NULL;
                if (slot != T_NONE) {      // (re)assigned: it and everything below start over (duplicate keys: last wins)
                    tv[slot].kind = vk; tv[slot].s = pos; tv[slot].e = pos; tv[slot].flags = 0;
                    if (slot == T_CHOICES) { tv[T_ELEM0].kind = tv[T_DELTA].kind = tv[T_CONTENT].kind = V_ABSENT; act2 = act3 = act4 = false; }
                    if (slot == T_ELEM0) { tv[T_DELTA].kind = tv[T_CONTENT].kind = V_ABSENT; act3 = act4 = false; }
                    if (slot == T_DELTA) { tv[T_CONTENT].kind = V_ABSENT; act4 = false; }
                }
                if (depth == 2 && act2 && !((stack >> 1) & 1)) first2 = false;
                last_started = slot;
                cur = slot;
                if (act == A_OPEN_OBJ || act == A_OPEN_ARR) {
                    const bool obj = act == A_OPEN_OBJ;
                    if (depth >= 64) { too_deep = true; state = S_ERR; break; }
                    stack = (stack & ~(1ull << depth)) | ((uint64_t)(obj ? 1 : 0) << depth);
                    depth++;
                    if (slot == T_CHOICES) { act2 = true; first2 = !obj; }
                    if (slot == T_ELEM0) act3 = obj;
                    if (slot == T_DELTA) act4 = obj;
                } else if (act == A_STR_START) {
                    vstart = pos + 1;
                }
            } else if (act == A_KEY_START) {
                vstart = pos + 1; key_esc = false;
            } else if (act == A_KEY_ESC) {
                key_esc = true;
            } else if (act == A_KEY_END) {
                lastkey = match_key(text, vstart, pos, key_esc);
            } else if (act == A_STR_END) {
                if (cur != T_NONE) { tv[cur].s = vstart; tv[cur].e = pos; }
            } else if (act == A_COMMA) {
                if (depth == 0) next = S_ERR;
                else next = ((stack >> (depth - 1)) & 1) ? S_OBJ_KEY : S_VALUE;
            } else if (act == A_CLOSE_OBJ || act == A_CLOSE_ARR || act == A_CLOSE_ARR_EMPTY) {
                const bool want_obj = act == A_CLOSE_OBJ;
                if (depth == 0 || (((stack >> (depth - 1)) & 1) != 0) != want_obj) next = S_ERR;
                else {
                    if (act == A_CLOSE_ARR_EMPTY && last_started != T_NONE && tv[last_started].kind == V_ARRAY)
                        tv[last_started].flags |= AF_EMPTY;
                    depth--;
                    if (depth < 2) act2 = false;
                    if (depth < 3) act3 = false;
                    if (depth < 4) act4 = false;
                    next = S_AFTER;
                }
            }
        }
        state = next;
        if (state == S_ERR) break;
    }
    // a number may also end with the input
    if (state == S_NZERO || state == S_NINT || state == S_NFRAC || state == S_NEXP) {
        if (cur != T_NONE) { tv[cur].e = end; tv[cur].flags = number_flags(text, tv[cur].s, end); }
        state = S_AFTER;
    }

    Result res;
    res.status = SX_OK;
    res.len = 0;
    if (too_deep) { res.status = SX_UNSUPPORTED; return res; }
    const bool parsed = state == S_AFTER && depth == 0;   // else JSON.parse threw -> data === undefined

    Out o;
    o.p = out; o.n = 0; o.cap = cap; o.overflow = false;

    // ---- utils.ts:33-52 ------------------------------------------------------
    Val v;          // the value handed to `completion +=`
    v.kind = V_UNDEF; v.s = v.e = v.flags = 0;
    bool coerce_all = false;   // llamacpp appends whatever it got; the others append '' unless truthy
    const uint32_t root = parsed ? tv[T_ROOT].kind : V_UNDEF;
    if (kind == SX_KIND_LLAMACPP) {
        coerce_all = true;
        if (root == V_OBJECT && tv[T_RCONTENT].kind != V_ABSENT) v = tv[T_RCONTENT];   // data?.content
    } else {
        if (root == V_UNDEF || root == V_NULL) { res.len = 0; return res; }            // data?. short-circuits -> ''
        // data.choices
        uint32_t ck = root == V_OBJECT ? tv[T_CHOICES].kind : V_UNDEF;
        if (ck == V_ABSENT) ck = V_UNDEF;
        if (ck == V_UNDEF || ck == V_NULL) { res.status = SX_THROW; return res; }      // undefined[0]
        // choices[0]
        uint32_t ek;
        if (ck == V_ARRAY || ck == V_OBJECT) ek = tv[T_ELEM0].kind == V_ABSENT ? V_UNDEF : tv[T_ELEM0].kind;
        else if (ck == V_STRING) ek = tv[T_CHOICES].e > tv[T_CHOICES].s ? V_STRING : V_UNDEF;   // "abc"[0] === "a"
        else ek = V_UNDEF;                                                              // number / boolean
        if (ek == V_UNDEF || ek == V_NULL) { res.status = SX_THROW; return res; }      // undefined.delta
        // .delta
        uint32_t dk = (ek == V_OBJECT && ck != V_STRING) ? (tv[T_DELTA].kind == V_ABSENT ? V_UNDEF : tv[T_DELTA].kind) : V_UNDEF;
        if (dk == V_UNDEF || dk == V_NULL) {
            if (kind == SX_KIND_DEFAULT) { res.status = SX_THROW; return res; }        // delta.content (utils.ts:46)
            res.len = 0; return res;                                                    // delta?.content -> ''
        }
        if (dk == V_OBJECT && tv[T_CONTENT].kind != V_ABSENT) v = tv[T_CONTENT];
    }

    // ---- truthiness and ToString ------------------------------------------------
    bool truthy = false;
    switch (v.kind) {
        case V_TRUE: case V_OBJECT: case V_ARRAY: truthy = true; break;
        case V_NUMBER: truthy = (v.flags & NF_NONZERO) != 0; break;
        case V_STRING: truthy = v.e > v.s; break;
        default: truthy = false; break;
    }
    if (!coerce_all && !truthy) { res.len = 0; return res; }
    bool ok = true;
    switch (v.kind) {
        case V_UNDEF: case V_ABSENT: put_lit(o, "undefined"); break;
        case V_NULL: put_lit(o, "null"); break;
        case V_TRUE: put_lit(o, "true"); break;
        case V_FALSE: put_lit(o, "false"); break;
        case V_OBJECT: put_lit(o, "[object Object]"); break;
        case V_ARRAY: ok = (v.flags & AF_EMPTY) != 0; break;        // [] -> ''; others: Array#join on the host
        case V_NUMBER:
            if ((v.flags & NF_NONZERO) && (v.flags & NF_EXP)) ok = false;   // may under/overflow: host decides
            else ok = emit_number(text, v, o);
            break;
        case V_STRING: {
            const bool is_undefined_word = decode_string(text, v.s, v.e, o);
            if (kind == SX_KIND_DEFAULT && is_undefined_word) o.n = 0;      // utils.ts:46
            break;
        }
        default: break;
    }
    if (!ok || o.overflow) { res.status = SX_UNSUPPORTED; res.len = 0; return res; }
    res.len = o.n;
    return res;
}

}  // namespace sx
