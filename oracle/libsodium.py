"""Real libsodium 1.0.20 (PyNaCl's bundled build) -- TEST INFRASTRUCTURE ONLY.

This is the same C library the reference reaches through
sodium-universal 4.0.1 -> sodium-native 4.1.1
(/root/reference/package-lock.json:5725-5758) for every ``peer.write`` /
``peer.on('data')`` (/root/reference/src/provider.ts:248, :174).  It pins
oracle/secretstream_ref.c and generates tests/golden/.  PyNaCl is part of the
image, so this module also works on the GPU box.
"""
from __future__ import annotations

import ctypes as C
import os

import nacl
import nacl.bindings as nb

STATEBYTES = nb.crypto_secretstream_xchacha20poly1305_STATEBYTES
HEADERBYTES = nb.crypto_secretstream_xchacha20poly1305_HEADERBYTES
KEYBYTES = nb.crypto_secretstream_xchacha20poly1305_KEYBYTES
ABYTES = nb.crypto_secretstream_xchacha20poly1305_ABYTES

_SO = os.path.join(os.path.dirname(nacl.__file__), "_sodium.abi3.so")
_cdll = None


def cdll() -> C.CDLL:
    """The shared object itself, for function pointers (bench) and primitives."""
    global _cdll
    if _cdll is None:
        _cdll = C.CDLL(_SO)
    return _cdll


def fn_ptr(name: str) -> int:
    return C.cast(getattr(cdll(), name), C.c_void_p).value


class State:
    """libsodium secretstream state with a deterministic header.

    ``init_pull(state, header, key)`` yields exactly the state ``init_push``
    would for that header (SURVEY.md 7.2 [VERIFIED]), so seal states are
    reproducible without libsodium's RNG.
    """

    def __init__(self, key: bytes, header: bytes):
        self.st = nb.crypto_secretstream_xchacha20poly1305_state()
        nb.crypto_secretstream_xchacha20poly1305_init_pull(self.st, header, key)

    @property
    def raw(self) -> bytes:
        return bytes(self.st.statebuf[0:STATEBYTES])

    def push(self, m: bytes, tag: int = 0, ad: bytes | None = None) -> bytes:
        return nb.crypto_secretstream_xchacha20poly1305_push(self.st, m, ad, tag)

    def pull(self, c: bytes, ad: bytes | None = None):
        try:
            m, tag = nb.crypto_secretstream_xchacha20poly1305_pull(self.st, c, ad)
        except Exception:
            return None
        return m, tag

    def rekey(self) -> None:
        nb.crypto_secretstream_xchacha20poly1305_rekey(self.st)

    def seal_frame(self, m: bytes, tag: int = 0) -> bytes:
        n = len(m) + ABYTES
        return bytes((n & 0xFF, (n >> 8) & 0xFF, (n >> 16) & 0xFF)) + self.push(m, tag)

    def open_frame(self, wire: bytes):
        n = wire[0] | (wire[1] << 8) | (wire[2] << 16)
        if n + 3 != len(wire) or n < ABYTES:
            return None
        return self.pull(wire[3:])


def hchacha20(key: bytes, inp16: bytes) -> bytes:
    out = C.create_string_buffer(32)
    rc = cdll().crypto_core_hchacha20(out, inp16, key, None)
    assert rc == 0
    return out.raw


def chacha20_ietf_xor_ic(data: bytes, nonce: bytes, ic: int, key: bytes) -> bytes:
    out = C.create_string_buffer(max(1, len(data)))
    rc = cdll().crypto_stream_chacha20_ietf_xor_ic(out, data, C.c_ulonglong(len(data)), nonce, C.c_uint32(ic), key)
    assert rc == 0
    return out.raw[: len(data)]


def poly1305(msg: bytes, key: bytes) -> bytes:
    out = C.create_string_buffer(16)
    rc = cdll().crypto_onetimeauth_poly1305(out, msg, C.c_ulonglong(len(msg)), key)
    assert rc == 0
    return out.raw
