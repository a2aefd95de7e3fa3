"""ctypes front-end of oracle/secretstream_ref.c -- TEST INFRASTRUCTURE ONLY.

The C file restates libsodium's secretstream (the arithmetic behind
/root/reference/src/provider.ts:248 ``peer.write(chunk)``); see its header for
citations and how it is pinned.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libssref.so")

STATEBYTES, HEADERBYTES, KEYBYTES, ABYTES = 52, 24, 32, 17
TAG_MESSAGE, TAG_PUSH, TAG_REKEY, TAG_FINAL = 0, 1, 2, 3


def build(force: bool = False) -> str:
    """Compile the restatement with gcc (oracle/Makefile)."""
    srcs = [os.path.join(_HERE, f) for f in ("secretstream_ref.c", "cpu_bench.c", "Makefile")]
    stale = (not os.path.exists(_SO)) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "-s"] + (["-B"] if force else []), check=True)
    return _SO


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        u8p, u32, u64 = C.c_char_p, C.c_uint32, C.c_uint64
        L.ssref_chacha20_block.argtypes = [u8p, u32, u8p, u8p]
        L.ssref_hchacha20.argtypes = [u8p, u8p, u8p]
        L.ssref_chacha20_ietf_xor_ic.argtypes = [u8p, u8p, u64, u8p, u32, u8p]
        L.ssref_poly1305.argtypes = [u8p, u8p, C.c_size_t, u8p]
        L.ssref_init.argtypes = [u8p, u8p, u8p]
        L.ssref_rekey.argtypes = [u8p]
        L.ssref_push.argtypes = [u8p, u8p, u8p, u64, u8p, u64, C.c_uint8]
        L.ssref_push.restype = C.c_int
        L.ssref_pull.argtypes = [u8p, u8p, u8p, u8p, u64, u8p, u64]
        L.ssref_pull.restype = C.c_int
        L.ssref_seal_frame.argtypes = [u8p, u8p, u8p, u32, C.c_uint8]
        L.ssref_seal_frame.restype = C.c_int
        L.ssref_open_frame.argtypes = [u8p, u8p, u8p, u8p, u32]
        L.ssref_open_frame.restype = C.c_int
        L.cpub_run.restype = C.c_double
        L.cpub_run.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, u32, C.c_void_p, C.c_void_p,
                               C.c_void_p, u64, C.c_void_p, C.c_void_p, u64, u32, u32, u32, C.c_int, C.c_int, C.c_int]
        L.cpub_chain_run.restype = C.c_double
        L.cpub_chain_run.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, u32, u32,
                                     C.c_void_p]
        _lib = L
    return _lib


def chacha20_block(key: bytes, counter: int, nonce: bytes) -> bytes:
    out = C.create_string_buffer(64)
    lib().ssref_chacha20_block(key, counter, nonce, out)
    return out.raw


def hchacha20(key: bytes, inp16: bytes) -> bytes:
    out = C.create_string_buffer(32)
    lib().ssref_hchacha20(key, inp16, out)
    return out.raw


def chacha20_ietf_xor_ic(data: bytes, nonce: bytes, ic: int, key: bytes) -> bytes:
    out = C.create_string_buffer(max(1, len(data)))
    lib().ssref_chacha20_ietf_xor_ic(out, data, len(data), nonce, ic, key)
    return out.raw[: len(data)]


def poly1305(msg: bytes, key: bytes) -> bytes:
    out = C.create_string_buffer(16)
    lib().ssref_poly1305(out, msg, len(msg), key)
    return out.raw


class State:
    """52-byte secretstream state: k[32] | LE32 counter | inonce[8] | pad[8]."""

    def __init__(self, raw: bytes | None = None):
        self.buf = C.create_string_buffer(STATEBYTES)
        if raw is not None:
            assert len(raw) == STATEBYTES
            self.buf.raw = raw

    @classmethod
    def init(cls, key: bytes, header: bytes) -> "State":
        assert len(key) == KEYBYTES and len(header) == HEADERBYTES
        s = cls()
        lib().ssref_init(s.buf, header, key)
        return s

    @property
    def raw(self) -> bytes:
        return self.buf.raw

    def rekey(self) -> None:
        lib().ssref_rekey(self.buf)

    def push(self, m: bytes, tag: int = TAG_MESSAGE, ad: bytes = b"") -> bytes:
        out = C.create_string_buffer(len(m) + ABYTES)
        lib().ssref_push(self.buf, out, m, len(m), ad or None, len(ad), tag)
        return out.raw

    def pull(self, c: bytes, ad: bytes = b""):
        """Returns (plaintext, tag) or None on failure (state untouched)."""
        m = C.create_string_buffer(max(1, len(c)))
        tag = C.create_string_buffer(1)
        rc = lib().ssref_pull(self.buf, m, tag, c, len(c), ad or None, len(ad))
        if rc != 0:
            return None
        return m.raw[: len(c) - ABYTES], tag.raw[0]

    def seal_frame(self, m: bytes, tag: int = TAG_MESSAGE) -> bytes:
        """uint24-LE(len+17) ‖ push(m): what secret-stream `_write` puts on the wire."""
        out = C.create_string_buffer(len(m) + 20)
        rc = lib().ssref_seal_frame(self.buf, out, m, len(m), tag)
        assert rc == 0
        return out.raw

    def open_frame(self, wire: bytes):
        m = C.create_string_buffer(max(1, len(wire)))
        tag = C.create_string_buffer(1)
        rc = lib().ssref_open_frame(self.buf, m, tag, wire, len(wire))
        if rc != 0:
            return None
        return m.raw[: len(wire) - 20], tag.raw[0]
