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
enum : uint32_t {
    S_VALUE = 0, S_AFTER, S_OBJ_FIRST, S_OBJ_KEY, S_COLON, S_ARR_FIRST,
    S_STR, S_ESC, S_U1, S_U2, S_U3, S_U4,
    S_NMINUS, S_NZERO, S_NINT, S_NDOT, S_NFRAC, S_NE, S_NESIGN, S_NEXP,
    S_LIT, S_ERR
};
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

struct Result {
    uint32_t status, len;
};

// text = chunk bytes, n = length, kind = SX_KIND_*, out = at least `cap` writable bytes
SS_DEV Result extract_chunk(const uint8_t *text, uint32_t n, uint32_t kind, uint8_t *out, uint32_t cap) {
    // ---- utils.ts:16-27: prefix test and split('data:')[1] ------------------
    uint32_t pos = 0, end = n;
    bool prefixed = n >= 5 && text[0] == 'd' && text[1] == 'a' && text[2] == 't' && text[3] == 'a' && text[4] == ':';
    if (prefixed) pos = 5;

    Val tv[6];
#pragma unroll
    for (int i = 0; i < 6; i++) { tv[i].kind = V_ABSENT; tv[i].s = tv[i].e = tv[i].flags = 0; }
    uint32_t state = S_VALUE, depth = 0;
    uint64_t stack = 0;                  // bit d-1: container at depth d is an object
    uint32_t lastkey = KEY_NONE;
    bool act2 = false, act3 = false, act4 = false, first2 = false;
    bool str_is_key = false;
    uint32_t cand = 0, kpos = 0, ucode = 0;   // key matching
    uint32_t cur = T_NONE;               // tracked slot of the value being scanned (string/number/literal)
    uint32_t last_started = T_NONE;      // tracked slot of the most recently started value (array emptiness)
    uint32_t lit_word = 0, lit_pos = 0;  // literal matching: 0 true, 1 false, 2 null
    uint32_t nflags = 0, vstart = 0;
    bool too_deep = false;

    // which tracked slot does a value that starts now belong to?
#define SX_SLOT_OF_NEW_VALUE()                                                                         \
    (depth == 0 ? T_ROOT                                                                               \
     : depth == 1 && ((stack >> 0) & 1) && lastkey == KEY_CHOICES ? T_CHOICES                          \
     : depth == 1 && ((stack >> 0) & 1) && lastkey == KEY_CONTENT ? T_RCONTENT                         \
     : depth == 2 && act2 && (((stack >> 1) & 1) ? lastkey == KEY_ZERO : first2) ? T_ELEM0             \
     : depth == 3 && act3 && lastkey == KEY_DELTA ? T_DELTA                                            \
     : depth == 4 && act4 && lastkey == KEY_CONTENT ? T_CONTENT                                        \
     : T_NONE)
    // a tracked value is (re)assigned: it and everything below it start over (duplicate keys: last wins)
#define SX_BEGIN_TRACKED(slot, k)                                                                       \
    do {                                                                                               \
        if ((slot) != T_NONE) {                                                                        \
            tv[slot].kind = (k); tv[slot].s = pos; tv[slot].e = pos; tv[slot].flags = 0;               \
            if ((slot) == T_CHOICES) { tv[T_ELEM0].kind = tv[T_DELTA].kind = tv[T_CONTENT].kind = V_ABSENT; act2 = act3 = act4 = false; } \
            if ((slot) == T_ELEM0) { tv[T_DELTA].kind = tv[T_CONTENT].kind = V_ABSENT; act3 = act4 = false; } \
            if ((slot) == T_DELTA) { tv[T_CONTENT].kind = V_ABSENT; act4 = false; }                    \
        }                                                                                              \
        if (depth == 2 && act2 && !((stack >> 1) & 1)) first2 = false;                                 \
        last_started = (slot);                                                                         \
    } while (0)

    for (;; pos++) {
        uint32_t c;
        bool at_end = pos >= end;
        if (!at_end && prefixed && pos + 5 <= n && text[pos] == 'd' && text[pos + 1] == 'a' && text[pos + 2] == 't' &&
            text[pos + 3] == 'a' && text[pos + 4] == ':') {
            at_end = true;               // the second 'data:' ends split()[1], wherever it falls
            end = pos;
        }
        c = at_end ? 0x100u : text[pos];

        // a number ends at the first character that cannot continue it
        if (state == S_NZERO || state == S_NINT || state == S_NFRAC || state == S_NEXP) {
            bool cont = false;
            if (!at_end) {
                if (state == S_NINT && is_digit(c)) cont = true;
                else if ((state == S_NZERO || state == S_NINT) && c == '.') cont = true;
                else if ((state == S_NZERO || state == S_NINT || state == S_NFRAC) && (c == 'e' || c == 'E')) cont = true;
                else if ((state == S_NFRAC || state == S_NEXP) && is_digit(c)) cont = true;
            }
            if (!cont) {
                if (cur != T_NONE) { tv[cur].e = pos; tv[cur].flags = nflags; }
                state = S_AFTER;
            }
        }
        if (at_end) break;

        switch (state) {
            case S_VALUE:
            case S_ARR_FIRST: {
                if (is_ws(c)) break;
                if (state == S_ARR_FIRST && c == ']') {
                    if (last_started != T_NONE && tv[last_started].kind == V_ARRAY) tv[last_started].flags |= AF_EMPTY;
                    depth--;
                    if (depth < 2) act2 = false;
                    if (depth < 3) act3 = false;
                    if (depth < 4) act4 = false;
                    state = S_AFTER;
                    break;
                }
                const uint32_t slot = SX_SLOT_OF_NEW_VALUE();
                if (c == '{' || c == '[') {
                    const bool obj = c == '{';
                    SX_BEGIN_TRACKED(slot, obj ? V_OBJECT : V_ARRAY);
                    if (depth >= 64) { too_deep = true; state = S_ERR; break; }
                    stack = (stack & ~(1ull << depth)) | ((uint64_t)(obj ? 1 : 0) << depth);
                    depth++;
                    if (slot == T_CHOICES) { act2 = true; first2 = !obj; }
                    if (slot == T_ELEM0) act3 = obj;
                    if (slot == T_DELTA) act4 = obj;
                    state = obj ? S_OBJ_FIRST : S_ARR_FIRST;
                } else if (c == '"') {
                    SX_BEGIN_TRACKED(slot, V_STRING);
                    cur = slot; str_is_key = false; vstart = pos + 1;
                    state = S_STR;
                } else if (c == '-' || is_digit(c)) {
                    SX_BEGIN_TRACKED(slot, V_NUMBER);
                    cur = slot; nflags = (c >= '1' && c <= '9') ? NF_NONZERO : 0;
                    state = c == '-' ? S_NMINUS : c == '0' ? S_NZERO : S_NINT;
                } else if (c == 't' || c == 'f' || c == 'n') {
                    lit_word = c == 't' ? 0 : c == 'f' ? 1 : 2;
                    SX_BEGIN_TRACKED(slot, lit_word == 0 ? V_TRUE : lit_word == 1 ? V_FALSE : V_NULL);
                    cur = slot; lit_pos = 1;
                    state = S_LIT;
                } else {
                    state = S_ERR;
                }
                break;
            }
            case S_LIT: {
                const char *w = lit_word == 0 ? "true" : lit_word == 1 ? "false" : "null";
                const uint32_t wl = lit_word == 1 ? 5u : 4u;
                if (c != (uint32_t)(uint8_t)w[lit_pos]) { state = S_ERR; break; }
                lit_pos++;
                if (lit_pos == wl) state = S_AFTER;
                break;
            }
            case S_AFTER: {
                if (is_ws(c)) break;
                if (depth == 0) { state = S_ERR; break; }
                const bool in_obj = (stack >> (depth - 1)) & 1;
                if (c == ',') state = in_obj ? S_OBJ_KEY : S_VALUE;
                else if ((c == '}' && in_obj) || (c == ']' && !in_obj)) {
                    depth--;
                    if (depth < 2) act2 = false;
                    if (depth < 3) act3 = false;
                    if (depth < 4) act4 = false;
                } else state = S_ERR;
                break;
            }
            case S_OBJ_FIRST:
            case S_OBJ_KEY: {
                if (is_ws(c)) break;
                if (c == '"') { str_is_key = true; cand = 0x1e; kpos = 0; state = S_STR; }
                else if (c == '}' && state == S_OBJ_FIRST) {
                    depth--;
                    if (depth < 2) act2 = false;
                    if (depth < 3) act3 = false;
                    if (depth < 4) act4 = false;
                    state = S_AFTER;
                } else state = S_ERR;
                break;
            }
            case S_COLON: {
                if (is_ws(c)) break;
                state = c == ':' ? S_VALUE : S_ERR;
                break;
            }
            case S_STR: {
                if (c == '"') {
                    if (str_is_key) {
                        lastkey = KEY_NONE;
#pragma unroll
                        for (uint32_t k = KEY_CHOICES; k <= KEY_ZERO; k++)
                            if (((cand >> k) & 1) && kpos == name_len(k)) lastkey = k;
                        state = S_COLON;
                    } else {
                        if (cur != T_NONE) { tv[cur].s = vstart; tv[cur].e = pos; }
                        state = S_AFTER;
                    }
                } else if (c == 0x5c) state = S_ESC;
                else if (c < 0x20) state = S_ERR;
                else if (str_is_key) {
#pragma unroll
                    for (uint32_t k = KEY_CHOICES; k <= KEY_ZERO; k++)
                        if (name_char(k, kpos) != c) cand &= ~(1u << k);
                    kpos++;
                }
                break;
            }
            case S_ESC: {
                uint32_t dc = 0xffffffffu;
                if (c == '"' || c == 0x5c || c == '/') dc = c;
                else if (c == 'b') dc = 8; else if (c == 'f') dc = 12; else if (c == 'n') dc = 10;
                else if (c == 'r') dc = 13; else if (c == 't') dc = 9;
                else if (c == 'u') { ucode = 0; state = S_U1; break; }
                if (dc == 0xffffffffu) { state = S_ERR; break; }
                if (str_is_key) {
#pragma unroll
                    for (uint32_t k = KEY_CHOICES; k <= KEY_ZERO; k++)
                        if (name_char(k, kpos) != dc) cand &= ~(1u << k);
                    kpos++;
                }
                state = S_STR;
                break;
            }
            case S_U1: case S_U2: case S_U3: case S_U4: {
                const int h = hexval(c);
                if (h < 0) { state = S_ERR; break; }
                ucode = (ucode << 4) | (uint32_t)h;
                if (state == S_U4) {
                    if (str_is_key) {
#pragma unroll
                        for (uint32_t k = KEY_CHOICES; k <= KEY_ZERO; k++)
                            if (name_char(k, kpos) != ucode || ucode == 0) cand &= ~(1u << k);
                        kpos++;
                    }
                    state = S_STR;
                } else state = state + 1;
                break;
            }
            case S_NMINUS:
                if (c == '0') state = S_NZERO;
                else if (c >= '1' && c <= '9') { nflags |= NF_NONZERO; state = S_NINT; }
                else state = S_ERR;
                break;
            case S_NZERO:   // continuation characters only ('.', e/E) reach here
            case S_NINT:
                if (is_digit(c)) { /* S_NINT digit */ }
                else if (c == '.') { nflags |= NF_FRAC; state = S_NDOT; }
                else { nflags |= NF_EXP; state = S_NE; }
                break;
            case S_NDOT:
                if (is_digit(c)) { if (c != '0') nflags |= NF_NONZERO; state = S_NFRAC; }
                else state = S_ERR;
                break;
            case S_NFRAC:
                if (is_digit(c)) { if (c != '0') nflags |= NF_NONZERO; }
                else { nflags |= NF_EXP; state = S_NE; }
                break;
            case S_NE:
                if (c == '+' || c == '-') state = S_NESIGN;
                else if (is_digit(c)) state = S_NEXP;
                else state = S_ERR;
                break;
            case S_NESIGN:
                state = is_digit(c) ? S_NEXP : S_ERR;
                break;
            case S_NEXP:
                break;  // digits only reach here
            default:
                state = S_ERR;
                break;
        }
        if (state == S_ERR) break;
    }
#undef SX_SLOT_OF_NEW_VALUE
#undef SX_BEGIN_TRACKED

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
