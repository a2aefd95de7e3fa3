"""Delta-extraction oracle -- TEST INFRASTRUCTURE ONLY ("parity unpinned").

CPU restatement, with ECMAScript semantics, of what the reference does to each
upstream chunk before ``peer.write(chunk)``:

    /root/reference/src/provider.ts:243-246   completion += getChatDataFromProvider(
                                                  provider, safeParseStreamResponse(chunk.toString()))
    /root/reference/src/utils.ts:16-18        isStreamWithDataPrefix
    /root/reference/src/utils.ts:20-31        safeParseStreamResponse
    /root/reference/src/utils.ts:33-52        getChatDataFromProvider
    /root/reference/src/constants.ts:22-29    apiProviders
    /root/reference/src/provider.ts:270-274   a thrown TypeError aborts that stream's relay

No JavaScript engine exists in this image, and the
reference's own tests pin nothing here (section 4), so this restatement IS the
oracle and its table-driven cases (tests/golden/extract_cases.json) are the
fixtures: parity for the extractor is "unpinned" in the task's sense.

Values: a parsed JSON value is modelled as dict / list / str / float / bool /
None, with ``UNDEFINED`` standing for JS ``undefined``.  Strings are Python
``str`` that may hold lone surrogates (JS strings are UTF-16); the delta is
returned as WTF-8 bytes (UTF-8, lone surrogates encoded as 3-byte sequences).
"""
from __future__ import annotations

import math

PROVIDERS = ("litellm", "llamacpp", "lmstudio", "ollama", "oobabooga", "openwebui")  # constants.ts:22-29

OK, THROW = 0, 1  # THROW = the TypeError case of SURVEY row A4 (stream aborted, chunk not written)


class _Undefined:
    def __repr__(self):
        return "undefined"


UNDEFINED = _Undefined()


class JsTypeError(Exception):
    """TypeError: Cannot read properties of undefined/null."""


class JsSyntaxError(Exception):
    pass


# ----------------------------------------------------------------------------
# chunk.toString()  (Buffer -> JS string, UTF-8 with U+FFFD per maximal subpart;
# the BOM is NOT stripped by Buffer#toString)
def buffer_to_string(chunk: bytes) -> str:
    return chunk.decode("utf-8", "replace")


# ----------------------------------------------------------------------------
# JSON.parse (ECMA-262 25.5.1 / ECMA-404): strict grammar, duplicate keys -> last wins
_WS = " \t\n\r"
_ESC = {'"': '"', "\\": "\\", "/": "/", "b": "\b", "f": "\f", "n": "\n", "r": "\r", "t": "\t"}
_HEX = "0123456789abcdefABCDEF"


def json_parse(text: str):
    pos = 0
    n = len(text)

    def skip_ws():
        nonlocal pos
        while pos < n and text[pos] in _WS:
            pos += 1

    def parse_string() -> str:
        nonlocal pos
        assert text[pos] == '"'
        pos += 1
        out = []
        while True:
            if pos >= n:
                raise JsSyntaxError("unterminated string")
            ch = text[pos]
            if ch == '"':
                pos += 1
                break
            if ord(ch) < 0x20:
                raise JsSyntaxError("control character in string")
            if ch == "\\":
                pos += 1
                if pos >= n:
                    raise JsSyntaxError("bad escape")
                e = text[pos]
                if e in _ESC:
                    out.append(_ESC[e])
                    pos += 1
                elif e == "u":
                    hx = text[pos + 1 : pos + 5]
                    if len(hx) != 4 or any(c not in _HEX for c in hx):
                        raise JsSyntaxError("bad \\u escape")
                    out.append(chr(int(hx, 16)))
                    pos += 5
                else:
                    raise JsSyntaxError("bad escape")
            else:
                out.append(ch)
                pos += 1
        return _join_utf16("".join(out))

    def parse_number() -> float:
        nonlocal pos
        start = pos
        if pos < n and text[pos] == "-":
            pos += 1
        if pos < n and text[pos] == "0":
            pos += 1
        elif pos < n and text[pos] in "123456789":
            while pos < n and text[pos] in "0123456789":
                pos += 1
        else:
            raise JsSyntaxError("bad number")
        if pos < n and text[pos] == ".":
            pos += 1
            if not (pos < n and text[pos] in "0123456789"):
                raise JsSyntaxError("bad fraction")
            while pos < n and text[pos] in "0123456789":
                pos += 1
        if pos < n and text[pos] in "eE":
            pos += 1
            if pos < n and text[pos] in "+-":
                pos += 1
            if not (pos < n and text[pos] in "0123456789"):
                raise JsSyntaxError("bad exponent")
            while pos < n and text[pos] in "0123456789":
                pos += 1
        return float(text[start:pos])

    def parse_value(depth: int):
        nonlocal pos
        skip_ws()
        if pos >= n:
            raise JsSyntaxError("unexpected end")
        ch = text[pos]
        if ch == "{":
            pos += 1
            obj = {}
            skip_ws()
            if pos < n and text[pos] == "}":
                pos += 1
                return obj
            while True:
                skip_ws()
                if not (pos < n and text[pos] == '"'):
                    raise JsSyntaxError("expected key")
                k = parse_string()
                skip_ws()
                if not (pos < n and text[pos] == ":"):
                    raise JsSyntaxError("expected colon")
                pos += 1
                v = parse_value(depth + 1)
                obj.pop(k, None)  # last duplicate wins (insertion order is irrelevant here)
                obj[k] = v
                skip_ws()
                if pos < n and text[pos] == ",":
                    pos += 1
                    continue
                if pos < n and text[pos] == "}":
                    pos += 1
                    return obj
                raise JsSyntaxError("expected , or }")
        if ch == "[":
            pos += 1
            arr = []
            skip_ws()
            if pos < n and text[pos] == "]":
                pos += 1
                return arr
            while True:
                arr.append(parse_value(depth + 1))
                skip_ws()
                if pos < n and text[pos] == ",":
                    pos += 1
                    continue
                if pos < n and text[pos] == "]":
                    pos += 1
                    return arr
                raise JsSyntaxError("expected , or ]")
        if ch == '"':
            return parse_string()
        if ch == "-" or ch in "0123456789":
            return parse_number()
        for lit, val in (("true", True), ("false", False), ("null", None)):
            if text.startswith(lit, pos):
                pos += len(lit)
                return val
        raise JsSyntaxError("unexpected token")

    v = parse_value(0)
    skip_ws()
    if pos != n:
        raise JsSyntaxError("trailing characters")
    return v


def _join_utf16(s: str) -> str:
    """Combine escaped surrogate PAIRS into one code point (JS strings are UTF-16,
    so "\\ud83d\\ude00" is one astral character); lone surrogates stay."""
    if not any(0xD800 <= ord(c) <= 0xDFFF for c in s):
        return s
    out = []
    i = 0
    while i < len(s):
        c = ord(s[i])
        if 0xD800 <= c <= 0xDBFF and i + 1 < len(s) and 0xDC00 <= ord(s[i + 1]) <= 0xDFFF:
            out.append(chr(0x10000 + ((c - 0xD800) << 10) + (ord(s[i + 1]) - 0xDC00)))
            i += 2
        else:
            out.append(s[i])
            i += 1
    return "".join(out)


# ----------------------------------------------------------------------------
# utils.ts
def is_stream_with_data_prefix(s: str) -> bool:  # utils.ts:16-18
    return s.startswith("data:")


def safe_parse_stream_response(s: str):  # utils.ts:20-31
    try:
        if is_stream_with_data_prefix(s):
            return json_parse(s.split("data:")[1])
        return json_parse(s)
    except (JsSyntaxError, RecursionError):
        return UNDEFINED


def _get(v, name: str):
    """JS property read `v.name` / `v[name]` for the names this path uses."""
    if v is UNDEFINED or v is None:
        raise JsTypeError(f"Cannot read properties of {v!r} (reading '{name}')")
    if isinstance(v, dict):
        return v.get(name, UNDEFINED)
    if name == "0":
        if isinstance(v, list):
            return v[0] if v else UNDEFINED
        if isinstance(v, str):
            return _first_utf16_unit(v) if v else UNDEFINED
    return UNDEFINED  # numbers, booleans, other names on arrays/strings


def _first_utf16_unit(s: str) -> str:
    c = ord(s[0])
    if c >= 0x10000:
        return chr(0xD800 + ((c - 0x10000) >> 10))
    return s[0]


def _opt(v, name: str):
    """`v?.name`"""
    if v is UNDEFINED or v is None:
        return UNDEFINED
    return _get(v, name)


def _truthy(v) -> bool:
    if v is UNDEFINED or v is None or v is False:
        return False
    if isinstance(v, bool):
        return v
    if isinstance(v, float):
        return not (v == 0.0 or math.isnan(v))
    if isinstance(v, str):
        return len(v) > 0
    return True


def get_chat_data_from_provider(provider: str, data):  # utils.ts:33-52
    if provider in ("ollama", "openwebui"):
        if data is UNDEFINED or data is None:
            return ""
        c = _opt(_get(_get(_get(data, "choices"), "0"), "delta"), "content")
        return c if _truthy(c) else ""
    if provider == "llamacpp":
        return _opt(data, "content")
    # litellm + default (lmstudio, oobabooga, anything else)
    if data is UNDEFINED or data is None:
        return ""  # both optional chains short-circuit; undefined !== 'undefined'
    delta = _get(_get(_get(data, "choices"), "0"), "delta")
    if _get(delta, "content") == "undefined" and isinstance(_get(delta, "content"), str):
        return ""
    c = _opt(delta, "content")
    return c if _truthy(c) else ""


def number_to_string(x: float) -> str:
    """Number::toString(x), radix 10 (ECMA-262 6.1.6.1.20)."""
    if math.isnan(x):
        return "NaN"
    if x == 0:
        return "0"
    if math.isinf(x):
        return "Infinity" if x > 0 else "-Infinity"
    if x < 0:
        return "-" + number_to_string(-x)
    r = repr(x)  # shortest round-trip digits, same digit string as the spec's
    if "e" in r or "E" in r:
        mant, exp = r.lower().split("e")
        e = int(exp)
    else:
        mant, e = r, 0
    if "." in mant:
        ip, fp = mant.split(".")
    else:
        ip, fp = mant, ""
    digits = (ip + fp).lstrip("0")
    point = len(ip) + e  # decimal point position relative to start of ip+fp
    lead = len(ip + fp) - len((ip + fp).lstrip("0"))
    point -= lead
    digits = digits.rstrip("0") or "0"
    k = len(digits)
    nn = point  # value = 0.digits * 10^nn
    if k <= nn <= 21:
        return digits + "0" * (nn - k)
    if 0 < nn <= 21:
        return digits[:nn] + "." + digits[nn:]
    if -6 < nn <= 0:
        return "0." + "0" * (-nn) + digits
    ex = nn - 1
    sign = "+" if ex >= 0 else "-"
    if k == 1:
        return digits + "e" + sign + str(abs(ex))
    return digits[0] + "." + digits[1:] + "e" + sign + str(abs(ex))


def to_js_string(v) -> str:
    """ToString as applied by `completion += v`."""
    if v is UNDEFINED:
        return "undefined"
    if v is None:
        return "null"
    if v is True:
        return "true"
    if v is False:
        return "false"
    if isinstance(v, str):
        return v
    if isinstance(v, float):
        return number_to_string(v)
    if isinstance(v, list):  # Array.prototype.toString -> join(',')
        return ",".join("" if (e is None or e is UNDEFINED) else to_js_string(e) for e in v)
    if isinstance(v, dict):
        return "[object Object]"
    raise TypeError(type(v))


def wtf8(s: str) -> bytes:
    return s.encode("utf-8", "surrogatepass")


def extract(provider: str, chunk: bytes):
    """One iteration of provider.ts:243-246.  Returns (status, delta_bytes).

    status THROW means the reference throws a TypeError here: the chunk is NOT
    written to the peer and the relay for this stream ends (provider.ts:270-274).
    """
    s = buffer_to_string(chunk)
    data = safe_parse_stream_response(s)
    try:
        v = get_chat_data_from_provider(provider, data)
    except JsTypeError:
        return THROW, b""
    return OK, wtf8(to_js_string(v))
