"""CPU check of the PRODUCT's extractor source (symmetry_b200/csrc/ss_extract.cuh, compiled with
g++ in tests/emu/) against oracle/extract.py: the frozen quirk matrix and a mutation fuzz.
UNSUPPORTED is an allowed answer only for the documented shapes (the host then runs the JS helper);
any OK/THROW answer must equal the oracle's."""
import ctypes as C
import os
import random
import subprocess

import pytest

from oracle import extract as ex

HERE = os.path.dirname(os.path.abspath(__file__))
KIND = {"ollama": 1, "openwebui": 1, "llamacpp": 2}


@pytest.fixture(scope="module")
def lib():
    subprocess.run(["make", "-C", os.path.join(HERE, "emu"), "-s"], check=True)
    L = C.CDLL(os.path.join(HERE, "emu", "_build", "libss_emu.so"))
    L.emu_extract.restype = C.c_uint32
    L.emu_extract.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.c_char_p, C.c_uint32, C.POINTER(C.c_uint32)]
    return L


def run(lib, provider, chunk):
    cap = max(16, (len(chunk) + 15) & ~15)
    out = C.create_string_buffer(cap + 8)
    ln = C.c_uint32(0)
    st = lib.emu_extract(chunk, len(chunk), KIND.get(provider, 0), out, cap, C.byref(ln))
    assert out.raw[cap:] == bytes(8), "wrote past its delta slot"
    return st, out.raw[: ln.value]


def check(lib, provider, chunk):
    st, got = run(lib, provider, chunk)
    want_st, want = ex.extract(provider, chunk)
    if st == 4:
        return "unsupported"
    assert st == (3 if want_st == ex.THROW else 0), (provider, chunk, st, want_st)
    if st == 0:
        assert got == want, (provider, chunk, got, want)
    return "ok"


def test_frozen_matrix(lib, extract_cases):
    n_uns = 0
    for c in extract_cases:
        chunk = bytes.fromhex(c["chunk"])
        st, got = run(lib, c["provider"], chunk)
        if st == 4:
            n_uns += 1
            # only numbers with an exponent and non-empty arrays are left to the host in this table
            assert b"1e21" in chunk or b"[1," in chunk, chunk
            continue
        assert st == (3 if c["status"] == 1 else 0), (c["provider"], chunk)
        if st == 0:
            assert got.hex() == c["delta"], (c["provider"], chunk)
    assert n_uns == 13


SEEDS = [
    b'data: {"id":"chatcmpl-1","object":"chat.completion.chunk","created":1730000000,"model":"llama3:8b",'
    b'"system_fingerprint":"fp_ollama","choices":[{"index":0,"delta":{"role":"assistant","content":" tok"},'
    b'"finish_reason":null}]}\n\n',
    b'data: {"choices":[{"index":0,"delta":{"content":"\\u00e9\\ud83d\\ude00 \\"q\\" \\\\ "}}]}\n\n',
    b'{"content":"llama.cpp tok","stop":false,"n":12,"x":[1,2,{"a":null}],"y":{"content":"inner"}}',
    b'data: {"choices":{"0":{"delta":{"content":true}}},"choices":[{"delta":{"content":-12.50}}]}',
    'data: {"choices":[{"delta":{"content":"é中😀"}}]}'.encode(),
]
ALPHABET = b'{}[]":,\\ \n\tdata:choicesdeltacontent0123456789.-+eEtruefalsenull\xff\xe4\xb8\xc3\xa9u'


@pytest.mark.parametrize("seed", range(6))
def test_mutation_fuzz(lib, seed):
    rng = random.Random(seed)
    counts = {"ok": 0, "unsupported": 0}
    for _ in range(2500):
        b = bytearray(rng.choice(SEEDS))
        for _m in range(rng.choice([0, 1, 1, 2, 3, 6])):
            op = rng.randrange(4)
            pos = rng.randrange(len(b) + 1)
            if op == 0 and b:
                del b[min(pos, len(b) - 1)]
            elif op == 1:
                b.insert(pos, rng.choice(ALPHABET))
            elif op == 2 and b:
                b[min(pos, len(b) - 1)] = rng.choice(ALPHABET)
            else:
                j = rng.randrange(len(b) + 1)
                lo, hi = min(pos, j), max(pos, j)
                b[lo:lo] = b[lo:hi][:40]  # duplicate a slice (duplicate keys, repeated events)
        if rng.random() < 0.1:
            b = b[: rng.randrange(len(b) + 1)]  # chunk cut mid-JSON
        for prov in ("ollama", "llamacpp", "litellm"):
            counts[check(lib, prov, bytes(b))] += 1
    assert counts["ok"] > 6000 and counts["unsupported"] < counts["ok"] // 10
