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
    // (A separate copy loop for plain ASCII deltas was measured and dropped: among the 32 chunks of a warp one nearly
    // always holds an escape or a multi-byte character, so the warp ran both loops: config 5 24.8 -> 25.1 us.)
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

// What the scan leaves behind: the kinds of the five values on the path and, for the two values that can
// reach `completion +=`, where their text sits.
struct Tracked {
    uint32_t root, choices, elem0, delta;
    bool choices_nonempty;      // data.choices is a string with at least one character ("abc"[0])
    Val content, rcontent;      // choices[0].delta.content and root.content
};

// ---- utils.ts:33-52: the provider switch, truthiness and ToString -------------------------------
SS_DEV Result finish(const uint8_t *text, bool parsed, bool too_deep, const Tracked &tr, uint32_t kind, uint8_t *out,
                     uint32_t cap) {
    Result res;
    res.status = SX_OK;
    res.len = 0;
    if (too_deep) { res.status = SX_UNSUPPORTED; return res; }

    Out o;
    o.p = out; o.n = 0; o.cap = cap; o.overflow = false;

    Val v;          // the value handed to `completion +=`
    v.kind = V_UNDEF; v.s = v.e = v.flags = 0;
    bool coerce_all = false;   // llamacpp appends whatever it got; the others append '' unless truthy
    const uint32_t root = parsed ? tr.root : V_UNDEF;   // JSON.parse threw -> data === undefined
    if (kind == SX_KIND_LLAMACPP) {
        coerce_all = true;
        if (root == V_OBJECT && tr.rcontent.kind != V_ABSENT) v = tr.rcontent;          // data?.content
    } else {
        if (root == V_UNDEF || root == V_NULL) { res.len = 0; return res; }            // data?. short-circuits -> ''
        // data.choices
        uint32_t ck = root == V_OBJECT ? tr.choices : V_UNDEF;
        if (ck == V_ABSENT) ck = V_UNDEF;
        if (ck == V_UNDEF || ck == V_NULL) { res.status = SX_THROW; return res; }      // undefined[0]
        // choices[0]
        uint32_t ek;
        if (ck == V_ARRAY || ck == V_OBJECT) ek = tr.elem0 == V_ABSENT ? V_UNDEF : tr.elem0;
        else if (ck == V_STRING) ek = tr.choices_nonempty ? V_STRING : V_UNDEF;         // "abc"[0] === "a"
        else ek = V_UNDEF;                                                              // number / boolean
        if (ek == V_UNDEF || ek == V_NULL) { res.status = SX_THROW; return res; }      // undefined.delta
        // .delta
        uint32_t dk = (ek == V_OBJECT && ck != V_STRING) ? (tr.delta == V_ABSENT ? V_UNDEF : tr.delta) : V_UNDEF;
        if (dk == V_UNDEF || dk == V_NULL) {
            if (kind == SX_KIND_DEFAULT) { res.status = SX_THROW; return res; }        // delta.content (utils.ts:46)
            res.len = 0; return res;                                                    // delta?.content -> ''
        }
        if (dk == V_OBJECT && tr.content.kind != V_ABSENT) v = tr.content;
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

    const bool word_ok = (reinterpret_cast<uintptr_t>(text) & 3u) == 0;
    for (; pos < end; pos++) {
        // Inside a string nothing happens until a quote, a backslash or a control character: skip four
        // bytes at a time while a word holds none of them (and no 'd', which could start the second
        // "data:" that ends split()[1]).  Most bytes of a chat-completion chunk sit inside strings.
        if ((state == S_STR || state == S_KSTR) && word_ok) {
            while ((pos & 3u) == 0 && pos + 4u <= end) {
                const uint32_t w = *reinterpret_cast<const uint32_t *>(text + pos);
                const uint32_t q = w ^ 0x22222222u, b = w ^ 0x5c5c5c5cu, d = w ^ 0x64646464u, ctl = w & 0xe0e0e0e0u;
                const uint32_t hit = ((q - 0x01010101u) & ~q) | ((b - 0x01010101u) & ~b) | ((d - 0x01010101u) & ~d) |
                                     ((ctl - 0x01010101u) & ~ctl);
                if (hit & 0x80808080u) break;
                pos += 4u;
            }
            if (pos >= end) break;
        }
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

    Tracked tr;
    tr.root = tv[T_ROOT].kind; tr.choices = tv[T_CHOICES].kind; tr.elem0 = tv[T_ELEM0].kind; tr.delta = tv[T_DELTA].kind;
    tr.choices_nonempty = tv[T_CHOICES].e > tv[T_CHOICES].s;
    tr.content = tv[T_CONTENT]; tr.rcontent = tv[T_RCONTENT];
    return finish(text, state == S_AFTER && depth == 0, too_deep, tr, kind, out, cap);
}


// ---- the two-pass form the kernel runs on staged chunks ---------------------------------------------
// The single pass above puts the bookkeeping (which value is this? which key was that?) inside the byte
// loop.  One thread owns one chunk, so in every iteration some lane of the warp is at a structural
// character and the whole warp walks the bookkeeping code: measured 60 us for 16,384 chunks of ~216 B,
// about 550 cycles per byte.  Here the byte loop only runs the grammar: one class look-up and one table
// look-up per byte, the same few instructions for every lane, and what the grammar meets on the way (a
// key, a string, a number, a bracket) is appended to a per-thread event list.  The bookkeeping then runs
// once per event (~30 per chunk) with the lanes in step.
//
//  - The container kind is not on a stack the byte loop has to consult: the states after a value are
//    split by container (top level / object / array) and `ctx` holds, one per byte, the four states that
//    depend on the container: after a value, after a comma, after '}' and after ']'.  A table entry
//    either names its next state or picks one of those four, and a single byte-permute does the pick, so
//    the loop-carried chain is: table load -> permute -> address -> table load.
//  - Bytes are first mapped to classes (columns of the full table that are identical in every state);
//    the class table is derived from the grammar, not written by hand.
enum : uint32_t {
    Q_ERR = 0,                   // zero: a permute digit that replicates a clear sign bit yields this state
    Q_VALUE, Q_ARR_FIRST, Q_AFTER_T, Q_AFTER_A, Q_AFTER_O, Q_OBJ_FIRST, Q_OBJ_KEY, Q_COLON,
    Q_STR, Q_ESC, Q_U1, Q_U2, Q_U3, Q_U4,
    Q_KSTR, Q_KSTR2, Q_KESC, Q_KU1, Q_KU2, Q_KU3, Q_KU4,   // KSTR2: a key in which a backslash has been seen
    Q_NMINUS, Q_NZERO, Q_NINT, Q_NDOT, Q_NFRAC, Q_NE, Q_NESIGN, Q_NEXP,
    Q_T1, Q_T2, Q_T3, Q_F1, Q_F2, Q_F3, Q_F4, Q_N1, Q_N2, Q_N3,
    // the second "data:" (src/utils.ts:22 keeps what lies between the first and the second):
    Q_STR_D1, Q_STR_D2, Q_STR_D3, Q_STR_D4,         // a string body that ends in "d", "da", "dat", "data"
    Q_KSTR_D1, Q_KSTR_D2, Q_KSTR_D3, Q_KSTR_D4,
    Q_KSTR2_D1, Q_KSTR2_D2, Q_KSTR2_D3, Q_KSTR2_D4,
    Q_U4_D, Q_KU4_D,                                // the third hex digit of a \u escape was 'd'
    Q_X1, Q_X2, Q_X3, Q_X4, Q_DONE,                 // after a complete top-level value: "d", "da", "dat", "data", "data:"
    Q_CUT,                                          // "data:" inside a string: the piece before it ends mid-string
    Q_COUNT
};
enum : uint32_t {   // events, values first
    E_NONE = 0, E_OPEN_OBJ, E_OPEN_ARR, E_STR, E_NUM, E_NUM_CLOSE, E_LIT_T, E_LIT_F, E_LIT_N,
    E_KEY, E_KEY_ESC, E_CLOSE, E_CLOSE_EMPTY
};
// byte 3 of an entry's low word: event in bits 0..3, QF_START, QF_SPLIT (see build_fast_tables) and in the top two bits the change of
// nesting depth as a signed number (01 = a container opens, 11 = one closes).  QF_OBJ only tells
// fast_entry() which container opens and is not stored.
enum : uint32_t { QF_START = 0x10, QF_SPLIT = 0x20, QF_PUSH = 0x40, QF_POP = 0xc0, QF_OBJ = 0x100 };
enum : uint32_t { QSEL_AFTER = 0, QSEL_COMMA = 1, QSEL_CLOSE_OBJ = 2, QSEL_CLOSE_ARR = 3, QSEL_TABLE = 6 };
#define SX_FAST_CLASSES 32u
#define SX_FAST_ENTRIES (sx::Q_COUNT * SX_FAST_CLASSES)
#define SX_FAST_MAX_LEN 4095u       /* positions are stored in 12 bits */
#define SX_FAST_MAX_DEPTH 13u       /* deeper chunks are left to the single pass */
#define SX_FAST_STACK_WORDS (SX_FAST_MAX_DEPTH + 3u)
#define SX_CTX_TOP ((uint32_t)sx::Q_AFTER_T)                          /* comma, '}' and ']' are errors at top level */
#define SX_CTX_OBJ ((uint32_t)sx::Q_AFTER_O | (sx::Q_OBJ_KEY << 8))   /* | state after this object's '}' << 16 */
#define SX_CTX_ARR ((uint32_t)sx::Q_AFTER_A | (sx::Q_VALUE << 8))     /* | state after this array's ']' << 24 */

// entry, low word: bits 0..15 the permute selector, 16..23 the next state, 24..31 event | QF_*.  The
// selector's low digit picks a byte of `ctx` (0..3) or byte 2 of the entry itself (6); the other three
// digits pick the sign of that byte, i.e. zero.
// entry, high word: how this byte changes `ctx`, again as a permute of (ctx, word): the identity for
// most bytes; an opening bracket builds the new container's four states from bytes 2 and 3 of the word
// and moves the old "after a value" state into the slot of the matching closing bracket.
static inline uint64_t fast_entry(uint32_t next, uint32_t sel, uint32_t flags) {
    const uint32_t lo = (0xeee0u | sel) | (next << 16) | ((flags & 0xffu) << 24);
    uint32_t hi = 0x3210u;
    if ((flags & 0xc0u) == QF_PUSH) hi = (flags & QF_OBJ) ? (0xe076u | (SX_CTX_OBJ << 16)) : (0x0e76u | (SX_CTX_ARR << 16));
    return lo | ((uint64_t)hi << 32);
}

// Fills F[SX_FAST_ENTRIES] and cls8[256] (class * 8: the byte offset of the class inside a table row).
// Returns the number of byte classes (<= SX_FAST_CLASSES), or 0 if the grammar no longer fits.
static inline uint32_t build_fast_tables(uint64_t *F, uint8_t *cls8) {
    static uint64_t G[Q_COUNT][256];
    auto set = [&](uint32_t st, uint32_t c, uint32_t next, uint32_t flags) { G[st][c] = fast_entry(next, QSEL_TABLE, flags); };
    auto pick = [&](uint32_t st, uint32_t c, uint32_t sel, uint32_t flags) { G[st][c] = fast_entry(Q_ERR, sel, flags); };
    auto is_hex = [](uint32_t c) { return (c >= '0' && c <= '9') || ((c | 0x20u) >= 'a' && (c | 0x20u) <= 'f'); };
    for (uint32_t st = 0; st < Q_COUNT; st++)
        for (uint32_t c = 0; c < 256; c++) set(st, c, Q_ERR, 0);
    const uint8_t ws[4] = {0x20, 0x09, 0x0a, 0x0d};
    for (uint32_t st : {(uint32_t)Q_VALUE, (uint32_t)Q_ARR_FIRST}) {
        for (uint8_t w : ws) set(st, w, st, 0);
        set(st, '{', Q_OBJ_FIRST, E_OPEN_OBJ | QF_PUSH | QF_OBJ);
        set(st, '[', Q_ARR_FIRST, E_OPEN_ARR | QF_PUSH);
        set(st, '"', Q_STR, QF_START);
        set(st, '-', Q_NMINUS, QF_START);
        set(st, '0', Q_NZERO, QF_START);
        for (uint32_t c = '1'; c <= '9'; c++) set(st, c, Q_NINT, QF_START);
        set(st, 't', Q_T1, E_LIT_T);
        set(st, 'f', Q_F1, E_LIT_F);
        set(st, 'n', Q_N1, E_LIT_N);
    }
    pick(Q_ARR_FIRST, ']', QSEL_CLOSE_ARR, E_CLOSE_EMPTY | QF_POP);
    for (uint32_t st : {(uint32_t)Q_AFTER_T, (uint32_t)Q_AFTER_O, (uint32_t)Q_AFTER_A})
        for (uint8_t w : ws) set(st, w, st, 0);
    set(Q_AFTER_O, ',', Q_OBJ_KEY, 0);
    pick(Q_AFTER_O, '}', QSEL_CLOSE_OBJ, E_CLOSE | QF_POP);
    set(Q_AFTER_A, ',', Q_VALUE, 0);
    pick(Q_AFTER_A, ']', QSEL_CLOSE_ARR, E_CLOSE | QF_POP);
    for (uint32_t st : {(uint32_t)Q_OBJ_FIRST, (uint32_t)Q_OBJ_KEY}) {
        for (uint8_t w : ws) set(st, w, st, 0);
        set(st, '"', Q_KSTR, QF_START);
    }
    pick(Q_OBJ_FIRST, '}', QSEL_CLOSE_OBJ, E_CLOSE | QF_POP);
    for (uint8_t w : ws) set(Q_COLON, w, Q_COLON, 0);
    set(Q_COLON, ':', Q_VALUE, 0);
    // value strings
    for (uint32_t c = 0x20; c < 256; c++) set(Q_STR, c, Q_STR, 0);
    pick(Q_STR, '"', QSEL_AFTER, E_STR);
    set(Q_STR, 0x5c, Q_ESC, 0);
    for (char c : {'"', '\\', '/', 'b', 'f', 'n', 'r', 't'}) set(Q_ESC, (uint8_t)c, Q_STR, 0);
    set(Q_ESC, 'u', Q_U1, 0);
    for (uint32_t j = 0; j < 4; j++)
        for (uint32_t c = 0; c < 256; c++)
            if (is_hex(c)) set(Q_U1 + j, c, j == 3 ? (uint32_t)Q_STR : Q_U1 + j + 1, 0);
    // key strings: the same grammar; a key that held an escape ends in a different event
    for (uint32_t st : {(uint32_t)Q_KSTR, (uint32_t)Q_KSTR2}) {
        for (uint32_t c = 0x20; c < 256; c++) set(st, c, st, 0);
        set(st, '"', Q_COLON, st == Q_KSTR ? E_KEY : E_KEY_ESC);
        set(st, 0x5c, Q_KESC, 0);
    }
    for (char c : {'"', '\\', '/', 'b', 'f', 'n', 'r', 't'}) set(Q_KESC, (uint8_t)c, Q_KSTR2, 0);
    set(Q_KESC, 'u', Q_KU1, 0);
    for (uint32_t j = 0; j < 4; j++)
        for (uint32_t c = 0; c < 256; c++)
            if (is_hex(c)) set(Q_KU1 + j, c, j == 3 ? (uint32_t)Q_KSTR2 : Q_KU1 + j + 1, 0);
    // numbers
    set(Q_NMINUS, '0', Q_NZERO, 0);
    for (uint32_t c = '1'; c <= '9'; c++) set(Q_NMINUS, c, Q_NINT, 0);
    for (uint32_t c = '0'; c <= '9'; c++) {
        set(Q_NINT, c, Q_NINT, 0);
        set(Q_NDOT, c, Q_NFRAC, 0);
        set(Q_NFRAC, c, Q_NFRAC, 0);
        set(Q_NE, c, Q_NEXP, 0);
        set(Q_NESIGN, c, Q_NEXP, 0);
        set(Q_NEXP, c, Q_NEXP, 0);
    }
    set(Q_NZERO, '.', Q_NDOT, 0);
    set(Q_NINT, '.', Q_NDOT, 0);
    for (uint32_t st : {(uint32_t)Q_NZERO, (uint32_t)Q_NINT, (uint32_t)Q_NFRAC}) { set(st, 'e', Q_NE, 0); set(st, 'E', Q_NE, 0); }
    set(Q_NE, '+', Q_NESIGN, 0);
    set(Q_NE, '-', Q_NESIGN, 0);
    for (uint32_t st : {(uint32_t)Q_NZERO, (uint32_t)Q_NINT, (uint32_t)Q_NFRAC, (uint32_t)Q_NEXP}) {   // a number ends at a delimiter
        for (uint8_t w : ws) pick(st, w, QSEL_AFTER, E_NUM);
        pick(st, ',', QSEL_COMMA, E_NUM);
        pick(st, '}', QSEL_CLOSE_OBJ, E_NUM_CLOSE | QF_POP);
        pick(st, ']', QSEL_CLOSE_ARR, E_NUM_CLOSE | QF_POP);
    }
    // literals
    set(Q_T1, 'r', Q_T2, 0); set(Q_T2, 'u', Q_T3, 0); pick(Q_T3, 'e', QSEL_AFTER, 0);
    set(Q_F1, 'a', Q_F2, 0); set(Q_F2, 'l', Q_F3, 0); set(Q_F3, 's', Q_F4, 0); pick(Q_F4, 'e', QSEL_AFTER, 0);
    set(Q_N1, 'u', Q_N2, 0); set(Q_N2, 'l', Q_N3, 0); pick(Q_N3, 'l', QSEL_AFTER, 0);

    // ---- the second "data:" ----
    // For a chunk that starts with "data:" the reference parses only what lies before the next "data:".
    // Where that next one follows a complete top-level value (two events in one chunk) the grammar ends
    // there: Q_DONE swallows the rest.  Inside a string body the four states per string kind follow "d",
    // "da", "dat", "data" and the ':' ends in Q_CUT: the piece before the split stops inside a string, so
    // JSON.parse throws.  Anywhere else the table only notices that a split may be there (QF_SPLIT: the
    // chunk is then left to the single pass).  In strings (a \\u escape can supply the 'd' or the "da"
    // with its last hex digits; earlier in an escape, or right after the backslash, what follows "d" or
    // "da" is an error with or without the split).  Outside strings the pattern would have to start at a
    // 'd' met in a state that is not Q_ERR -- after Q_ERR the piece before the split holds the same error.
    {
        const uint32_t body[3] = {Q_STR, Q_KSTR, Q_KSTR2}, d1[3] = {Q_STR_D1, Q_KSTR_D1, Q_KSTR2_D1};
        for (uint32_t st = Q_VALUE; st <= Q_N3; st++) {
            const bool in_string = (st >= Q_STR && st <= Q_KU4);
            if (!in_string) G[st]['d'] |= (uint64_t)QF_SPLIT << 24;
        }
        const uint32_t u3[2] = {Q_U3, Q_KU3}, u4[2] = {Q_U4, Q_KU4}, u4d[2] = {Q_U4_D, Q_KU4_D}, ud1[2] = {Q_STR_D1, Q_KSTR2_D1};
        for (uint32_t k = 0; k < 2; k++) {
            for (uint32_t c = 0; c < 256; c++) G[u4d[k]][c] = G[u4[k]][c];
            set(u3[k], 'd', u4d[k], 0);
            set(u4[k], 'd', ud1[k], 0);
            set(u4d[k], 'd', ud1[k], 0);
            set(u4d[k], 'a', ud1[k] + 1, 0);
        }
        for (uint32_t k = 0; k < 3; k++) {
            for (uint32_t j = 0; j < 4; j++)
                for (uint32_t c = 0; c < 256; c++) G[d1[k] + j][c] = G[body[k]][c];
            for (uint32_t j = 0; j < 5; j++) set(j ? d1[k] + j - 1 : body[k], 'd', d1[k], 0);
            set(d1[k], 'a', d1[k] + 1, 0);
            set(d1[k] + 1, 't', d1[k] + 2, 0);
            set(d1[k] + 2, 'a', d1[k] + 3, 0);
            set(d1[k] + 3, ':', Q_CUT, 0);
        }
        set(Q_AFTER_T, 'd', Q_X1, 0);
        set(Q_X1, 'a', Q_X2, 0); set(Q_X2, 't', Q_X3, 0); set(Q_X3, 'a', Q_X4, 0); set(Q_X4, ':', Q_DONE, 0);
        for (uint32_t c = 0; c < 256; c++) { set(Q_DONE, c, Q_DONE, 0); set(Q_CUT, c, Q_CUT, 0); }
    }

    // byte classes: bytes whose columns agree in every state
    uint32_t n_cls = 0;
    uint8_t rep[256];
    for (uint32_t c = 0; c < 256; c++) {
        uint32_t k = 0;
        for (; k < n_cls; k++) {
            bool same = true;
            for (uint32_t st = 0; st < Q_COUNT && same; st++) same = G[st][c] == G[st][rep[k]];
            if (same) break;
        }
        if (k == n_cls) {
            if (n_cls == SX_FAST_CLASSES) return 0;
            rep[n_cls++] = (uint8_t)c;
        }
        cls8[c] = (uint8_t)(k * 8u);
    }
    for (uint32_t st = 0; st < Q_COUNT; st++)
        for (uint32_t k = 0; k < SX_FAST_CLASSES; k++) F[st * SX_FAST_CLASSES + k] = k < n_cls ? G[st][rep[k]] : fast_entry(Q_ERR, QSEL_TABLE, 0);
    return n_cls;
}

SS_DEV uint32_t permute(uint32_t a, uint32_t b, uint32_t sel) {
#ifdef __CUDA_ARCH__
    uint32_t r;   // PTX prmt: bit 3 of a selector digit replicates the sign of the chosen byte (__byte_perm does not promise that)
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
    return r;
#else
    const uint64_t v = ((uint64_t)b << 32) | a;
    uint32_t r = 0;
    for (uint32_t i = 0; i < 4; i++) {
        const uint32_t d = (sel >> (4 * i)) & 15u;
        uint32_t byte = (uint32_t)(v >> (8 * (d & 7u))) & 0xffu;
        if (d & 8u) byte = (byte & 0x80u) ? 0xffu : 0u;
        r |= byte << (8 * i);
    }
    return r;
#endif
}

// t must be readable up to t[s + 7] (every caller's text is readable 8 bytes past its end).  No loop: the lanes of the
// bookkeeping warp run this for every key of every chunk, and eight independent byte loads cost less than a loop
// that shifts a 64-bit word into place byte by byte.
SS_DEV uint32_t match_key_plain(const uint8_t *t, uint32_t s, uint32_t e) {
    const uint32_t len = e - s;
    if (len != 1u && len != 5u && len != 7u) return KEY_NONE;
    const uint8_t *q = t + s;
    const uint32_t lo = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
    const uint32_t hi = (uint32_t)q[4] | ((uint32_t)q[5] << 8) | ((uint32_t)q[6] << 16);
    if (len == 1u) return (lo & 0xffu) == 0x30u ? KEY_ZERO : KEY_NONE;
    if (len == 5u) return (lo == 0x746c6564u && (hi & 0xffu) == 0x61u) ? KEY_DELTA : KEY_NONE;            // "delt" "a"
    return (lo == 0x696f6863u && hi == 0x00736563u) ? KEY_CHOICES                                          // "choi" "ces"
         : (lo == 0x746e6f63u && hi == 0x00746e65u) ? KEY_CONTENT : KEY_NONE;                              // "cont" "ent"
}

// ---- the two passes as three roles ------------------------------------------------------------------
// One thread per chunk runs ~12,000 dependent instructions (25 us for 16,384 chunks of ~216 B with one warp per
// scheduler).  The work splits into three stages that only talk through small queues, so the kernel gives a chunk
// to three threads in three different warps of a block and the stages run side by side:
//   A  fast_grammar   the byte loop proper: class look-up, table entry, the two permutes, the container stack.
//                     Per four input bytes it publishes ONE word holding the four entries' flag bytes.
//   B  fast_record    turns flag bytes into the event list (value start positions, one word per event).
//   C  fast_bookkeep  the per-event bookkeeping (which value is this?), then the provider switch / ToString.
// Queues are arrays in shared memory pre-filled with SX_PIPE_EMPTY: a consumer polls the word it needs next, so a
// word is its own "ready" flag and nothing needs a fence (no real flag word or event word has all bits set).  A
// publishes a trailer word last (final state, "ended inside a number"); B ends the event list with an SX_E_END
// word that also carries the deepest nesting and the "second data: somewhere unusual" flag.  A never waits for anyone; B and C give up after SX_PIPE_SPIN_LIMIT empty polls (the chunk then takes the
// single pass), so no interleaving can hang.  Run one after the other by one thread (extract_fast below, the
// emulator's form) the same code needs no pre-fill: every poll finds its word.
#define SX_PIPE_EMPTY 0xffffffffu
#define SX_PIPE_SPIN_LIMIT (1u << 20)
#define SX_E_END 13u                /* event type of the terminator word: | deepest nesting << 8 | SX_END_* */
#define SX_TRAILER_VALID 0x80000000u
#define SX_END_OVERFLOW 1u             /* terminator word: the event list may have overflowed (or B gave up) */
#define SX_END_SPLIT 2u                /* terminator word: a second "data:" somewhere unusual (QF_SPLIT seen) */
#define SX_TRAILER_NUMEND 0x00020000u  /* the input ended inside a number: one more E_NUM event at `end` */
#define SX_FAST_FLAG_WORDS(n) (((n) + 3u) >> 2)
#define SX_FAST_SCRATCH_WORDS_N(ev_cap, n) (SX_FAST_STACK_WORDS + 1u + SX_FAST_FLAG_WORDS(n) + (ev_cap) + 1u)

SS_DEV bool fast_prefixed(const uint8_t *text, uint32_t n) {
    return n >= 5 && text[0] == 'd' && text[1] == 'a' && text[2] == 't' && text[3] == 'a' && text[4] == ':';
}
// Waiting: the first few empty polls follow each other directly (the producer is usually a fraction of a microsecond
// away), after that each is followed by a short sleep so that a consumer far ahead of its producer leaves the issue
// slots to the warps that have work.  (A sleep after EVERY empty poll made the consumers the bottleneck: the lanes of
// a warp move on together, so one late word puts 32 chunks to sleep; 0.8 us per poll turned ~20 rounds into 16 us.)
#ifdef __CUDA_ARCH__
SS_DEV void pipe_backoff(uint32_t spins) { if (spins >= 8u) __nanosleep(100); }
#else
SS_DEV void pipe_backoff(uint32_t) {}
#endif

// Role A.  text: 4-byte aligned, n <= SX_FAST_MAX_LEN, readable up to text + n + 8; stack: SX_FAST_STACK_WORDS words
// of scratch; fl: SX_FAST_FLAG_WORDS(n) words (the queue to B); trailer: one word.
SS_DEV void fast_grammar(const uint8_t *text, uint32_t n, const uint64_t *F, const uint8_t *cls8, uint32_t *stack,
                         volatile uint32_t *fl, volatile uint32_t *trailer) {
    uint32_t pos = fast_prefixed(text, n) ? 5u : 0u;
    const uint32_t end = n;
    // Written without branches: each lane of a warp is at a different place in its own chunk, so any branch here is
    // taken by some lane in nearly every iteration.  A warp of these runs alone on its scheduler and most instructions
    // wait for the one before: what counts is how few instructions a byte takes.
    //  - saved[d] holds the ctx of the container d levels up.  Every byte stores the current ctx at its own depth
    //    (harmless unless the byte opens a container, when that slot becomes live) and reloads the slot below, so that
    //    a closing bracket finds its parent's ctx in a register.
    uint32_t *saved = stack + 1;                       // saved[-1] is read at depth 0
    uint32_t state = Q_VALUE, ctx = SX_CTX_TOP, popctx = SX_CTX_TOP, depth = 0;
    const char *Fb = reinterpret_cast<const char *>(F);
    auto step = [&](const char *col) -> uint32_t {   // col: the byte's column of the table; returns the entry's flag byte
        const uint64_t ent = *reinterpret_cast<const uint64_t *>(col + state * (SX_FAST_CLASSES * 8u));
        const uint32_t lo = (uint32_t)ent, hi = (uint32_t)(ent >> 32);
        state = permute(ctx, lo, lo);              // loop-carried: table load -> permute -> address -> table load
        const uint32_t dclamp = depth < SX_FAST_MAX_DEPTH + 1u ? depth : SX_FAST_MAX_DEPTH + 1u;   // also catches a wrapped depth
        saved[dclamp] = ctx;
        ctx = (int32_t)lo < 0 ? popctx : permute(ctx, hi, hi);   // a bracket of the wrong kind lands in Q_ERR, where nothing matters any more
        depth += (uint32_t)((int32_t)lo >> 30);
        const uint32_t dnew = depth < SX_FAST_MAX_DEPTH + 1u ? depth : SX_FAST_MAX_DEPTH + 1u;
        popctx = saved[(int32_t)dnew - 1];
        return lo >> 24;      // (the deepest nesting and the QF_SPLIT bit are B's to collect: both are in the flag bytes)
    };
    // the class of a byte does not depend on the state: fetch four bytes ahead of the chain
    const char *k[4], *kn[4];
#pragma unroll
    for (uint32_t j = 0; j < 4; j++) k[j] = Fb + cls8[text[pos + j]];
    uint32_t g = 0;
    while (pos + 4u <= end) {
#pragma unroll
        for (uint32_t j = 0; j < 4; j++) kn[j] = Fb + cls8[text[pos + 4u + j]];
        uint32_t w = 0;
#pragma unroll
        for (uint32_t j = 0; j < 4; j++) w |= step(k[j]) << (8u * j);
        fl[g++] = w;
#pragma unroll
        for (uint32_t j = 0; j < 4; j++) k[j] = kn[j];
        pos += 4u;
    }
    if (pos < end) {
        uint32_t w = 0;
#pragma unroll
        for (uint32_t j = 0; j < 3; j++)
            if (pos + j < end) w |= step(k[j]) << (8u * j);
        fl[g++] = w;
    }
    uint32_t t = SX_TRAILER_VALID;
    if (state == Q_NZERO || state == Q_NINT || state == Q_NFRAC || state == Q_NEXP) {   // a number may end with the input
        t |= SX_TRAILER_NUMEND;
        state = ctx & 0xffu;
    }
    *trailer = t | state;
}

// Polls *q until it holds something; false after SX_PIPE_SPIN_LIMIT empty polls.
SS_DEV bool pipe_take(volatile const uint32_t *q, uint32_t *w) {
    uint32_t v = *q;
    for (uint32_t spins = 0; v == SX_PIPE_EMPTY; spins++) {
        if (spins >= SX_PIPE_SPIN_LIMIT) return false;
        pipe_backoff(spins);
        v = *q;
    }
    *w = v;
    return true;
}

// Role B.  ev: ev_cap + 1 words (the queue to C).  Every flag byte is handled by the same few instructions; an
// event is one predicated store.
SS_DEV void fast_record(uint32_t n, bool prefixed, volatile const uint32_t *fl, volatile const uint32_t *trailer,
                        volatile uint32_t *ev, uint32_t ev_cap) {
    const uint32_t pos0 = prefixed ? 5u : 0u;
    const uint32_t groups = n > pos0 ? (n - pos0 + 3u) >> 2 : 0u;
    uint32_t vs12 = 0, nev = 0, depth = 0, max_depth = 0, seen = 0;
    bool dead = false;
    for (uint32_t g = 0; g < groups && !dead; g++) {
        uint32_t w;
        if (!pipe_take(fl + g, &w)) { dead = true; break; }
        seen |= w;
#pragma unroll
        for (uint32_t j = 0; j < 4; j++) {
            const uint32_t f = (w >> (8u * j)) & 0xffu, p = pos0 + 4u * g + j;
            depth += (uint32_t)((int32_t)(f << 24) >> 30);   // the flag byte's top two bits: +1 a container opens, -1 one closes
            max_depth = depth > max_depth && (int32_t)depth > 0 ? depth : max_depth;
            vs12 = (f & QF_START) ? (p << 12) : vs12;
            if ((f & 15u) && nev < ev_cap) ev[nev] = ((f & 15u) << 24) | vs12 | p;
            nev += (f & 15u) != 0u;
        }
    }
    uint32_t t = 0;
    if (!dead && !pipe_take(trailer, &t)) dead = true;   // (the trailer slot starts as SX_PIPE_EMPTY too)
    if (!dead && (t & SX_TRAILER_NUMEND)) {
        if (nev < ev_cap) ev[nev] = (E_NUM << 24) | vs12 | n;
        nev++;
    }
    const bool overflow = dead || nev >= ev_cap;   // as in the one-thread form: a full list may have overflowed
    const bool split = (seen & (QF_SPLIT * 0x01010101u)) != 0u;
    ev[nev < ev_cap ? nev : ev_cap] = (SX_E_END << 24) | ((max_depth < 255u ? max_depth : 255u) << 8) |
                                     (overflow ? SX_END_OVERFLOW : 0u) | (split ? SX_END_SPLIT : 0u);
}

// The bookkeeping of one event (what the single pass does inside its byte loop).
struct Book {
    Tracked tr;
    uint32_t d2, lastkey, last_started;
    bool root_obj, d2_obj, act2, act3, act4, first2;
};
SS_DEV void book_init(Book &b) {
    b.tr.root = b.tr.choices = b.tr.elem0 = b.tr.delta = V_ABSENT;
    b.tr.choices_nonempty = false;
    b.tr.content.kind = b.tr.rcontent.kind = V_ABSENT;
    b.tr.content.s = b.tr.content.e = b.tr.content.flags = 0;
    b.tr.rcontent.s = b.tr.rcontent.e = b.tr.rcontent.flags = 0;
    b.d2 = 0; b.lastkey = KEY_NONE; b.last_started = T_NONE;
    b.root_obj = b.d2_obj = b.act2 = b.act3 = b.act4 = b.first2 = false;
}
SS_DEV void book_event(Book &b, const uint8_t *text, uint32_t e) {
    Tracked &tr = b.tr;
    const uint32_t type = (e >> 24) & 15u, vs = (e >> 12) & 0xfffu, p = e & 0xfffu;
    if (type == E_KEY || type == E_KEY_ESC) {
        b.lastkey = type == E_KEY ? match_key_plain(text, vs + 1u, p) : match_key(text, vs + 1u, p, true);
        return;
    }
    if (type <= E_LIT_N) {
        const uint32_t slot =
            b.d2 == 0 ? T_ROOT
            : b.d2 == 1 && b.root_obj && b.lastkey == KEY_CHOICES ? T_CHOICES
            : b.d2 == 1 && b.root_obj && b.lastkey == KEY_CONTENT ? T_RCONTENT
            : b.d2 == 2 && b.act2 && (b.d2_obj ? b.lastkey == KEY_ZERO : b.first2) ? T_ELEM0
            : b.d2 == 3 && b.act3 && b.lastkey == KEY_DELTA ? T_DELTA
            : b.d2 == 4 && b.act4 && b.lastkey == KEY_CONTENT ? T_CONTENT
            : T_NONE;
        const uint32_t vk = type == E_OPEN_OBJ ? V_OBJECT : type == E_OPEN_ARR ? V_ARRAY : type == E_STR ? V_STRING
                          : type <= E_NUM_CLOSE ? V_NUMBER : type == E_LIT_T ? V_TRUE : type == E_LIT_F ? V_FALSE : V_NULL;
        // (re)assigned: it and everything below start over (duplicate keys: last wins)
        if (slot == T_ROOT) tr.root = vk;
        if (slot == T_CHOICES) {
            tr.choices = vk; tr.choices_nonempty = type == E_STR && p > vs + 1u;
            tr.elem0 = tr.delta = tr.content.kind = V_ABSENT; b.act2 = b.act3 = b.act4 = false;
        }
        if (slot == T_ELEM0) { tr.elem0 = vk; tr.delta = tr.content.kind = V_ABSENT; b.act3 = b.act4 = false; }
        if (slot == T_DELTA) { tr.delta = vk; tr.content.kind = V_ABSENT; b.act4 = false; }
        if (slot == T_CONTENT || slot == T_RCONTENT) {
            Val v;
            v.kind = vk; v.s = v.e = v.flags = 0;
            if (type == E_STR) { v.s = vs + 1u; v.e = p; }
            if (vk == V_NUMBER) { v.s = vs; v.e = p; v.flags = number_flags(text, vs, p); }
            if (slot == T_CONTENT) tr.content = v; else tr.rcontent = v;
        }
        if (b.d2 == 2 && b.act2 && !b.d2_obj) b.first2 = false;
        b.last_started = slot;
        if (type == E_OPEN_OBJ || type == E_OPEN_ARR) {
            const bool obj = type == E_OPEN_OBJ;
            if (b.d2 == 0) b.root_obj = obj;
            if (b.d2 == 1) b.d2_obj = obj;
            b.d2++;
            if (slot == T_CHOICES) { b.act2 = true; b.first2 = !obj; }
            if (slot == T_ELEM0) b.act3 = obj;
            if (slot == T_DELTA) b.act4 = obj;
        }
    }
    if (type == E_NUM_CLOSE || type >= E_CLOSE) {
        if (type == E_CLOSE_EMPTY) {   // "[]": the value that opened it is the most recent one
            if (b.last_started == T_CONTENT && tr.content.kind == V_ARRAY) tr.content.flags |= AF_EMPTY;
            if (b.last_started == T_RCONTENT && tr.rcontent.kind == V_ARRAY) tr.rcontent.flags |= AF_EMPTY;
        }
        b.d2--;
        if (b.d2 < 2) b.act2 = false;
        if (b.d2 < 3) b.act3 = false;
        if (b.d2 < 4) b.act4 = false;
    }
}

// Role C.  Returns false if the chunk is left to extract_chunk: more events than the list holds, nesting deeper than
// SX_FAST_MAX_DEPTH, a second "data:" somewhere unusual, or a producer that never delivered.  The bookkeeping runs
// on every chunk as its events arrive -- whether the chunk parsed is known only at the end, and a chunk that did not
// parse discards what was tracked (the events of such a chunk may be unbalanced: nothing here minds).
SS_DEV bool fast_bookkeep(const uint8_t *text, uint32_t n, bool prefixed, uint32_t kind, uint8_t *out, uint32_t cap,
                          volatile const uint32_t *ev, volatile const uint32_t *trailer, uint32_t ev_cap, Result *res) {
    Book b;
    book_init(b);
    uint32_t e = 0;
    for (uint32_t k = 0;; k++) {
        if (k > ev_cap || !pipe_take(ev + k, &e)) return false;
#ifdef SS_PROBE_TIMELINE
        if ((threadIdx.x & 31u) == 0u && k == 0) PROBE_MAX(10);
        if ((threadIdx.x & 31u) == 0u && k == 10) PROBE_MAX(11);
        if ((threadIdx.x & 31u) == 0u && k == 20) PROBE_MAX(12);
#endif
        if (((e >> 24) & 15u) == SX_E_END) break;
        book_event(b, text, e);
    }
#ifdef SS_PROBE_TIMELINE
    if ((threadIdx.x & 31u) == 0u) PROBE_MAX(9);
#endif
    if (e & SX_END_OVERFLOW) return false;                               // the list may have overflowed (or B gave up)
    uint32_t t;
    if (!pipe_take(trailer, &t)) return false;                           // (B has seen it already)
    const uint32_t state = t & 0xffu, max_depth = (e >> 8) & 0xffu;
    if (max_depth > SX_FAST_MAX_DEPTH) return false;
    if (prefixed && (e & SX_END_SPLIT)) return false;                    // a second "data:" somewhere unusual
    if (!prefixed && state == Q_CUT) return false;                       // no split without the prefix: parse it all
    const bool parsed = state == Q_AFTER_T || (prefixed && state == Q_DONE);   // Q_AFTER_T is only reachable outside every container
    if (!parsed) book_init(b);
#ifdef SS_PROBE_TIMELINE
    if ((threadIdx.x & 31u) == 0u) PROBE_MAX(13);
#endif
    *res = finish(text, parsed, false, b.tr, kind, out, cap);
#ifdef SS_PROBE_TIMELINE
    if ((threadIdx.x & 31u) == 0u) PROBE_MAX(14);
#endif
    return true;
}

// The three roles one after the other in one thread (the emulator's form; the kernel runs them in three warps).
// Returns false if the chunk is left to extract_chunk.  text: 4-byte aligned, n <= SX_FAST_MAX_LEN, readable up to
// text + n + 8; F, cls8: build_fast_tables(); scratch: SX_FAST_SCRATCH_WORDS_N(ev_cap, n) words.
SS_DEV bool extract_fast(const uint8_t *text, uint32_t n, uint32_t kind, uint8_t *out, uint32_t cap, const uint64_t *F,
                         const uint8_t *cls8, uint32_t *scratch, uint32_t ev_cap, Result *res) {
    uint32_t *stack = scratch, *trailer = scratch + SX_FAST_STACK_WORDS, *fl = trailer + 1u,
             *ev = fl + SX_FAST_FLAG_WORDS(n);
    const bool prefixed = fast_prefixed(text, n);
    fast_grammar(text, n, F, cls8, stack, fl, trailer);
    fast_record(n, prefixed, fl, trailer, ev, ev_cap);
    return fast_bookkeep(text, n, prefixed, kind, out, cap, ev, trailer, ev_cap, res);
}

}  // namespace sx
