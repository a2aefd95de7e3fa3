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


def mutated_chunks(seed, count):
    """Seed chunks with a few random byte edits each (shared with the GPU fuzz in test_gpu_extract.py)."""
    rng = random.Random(seed)
    for _ in range(count):
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
        yield bytes(b)


SPLIT_PIECES = [b"data:", b"data: ", b"data", b"dat", b"da", b"d", b"ddata:", b"dadata:", b"datdata:", b"datadata:",
                b"\\u00da", b"\\u00data:", b"ta:", b"a:", b'"}\n\ndata: {"content":"z"}']


def split_chunks(seed, count):
    """Seed chunks with pieces of "data:" dropped anywhere."""
    rng = random.Random(1000 + seed)
    for _ in range(count):
        b = bytearray(rng.choice(SEEDS))
        for _m in range(rng.choice([1, 1, 2, 3])):
            pos = rng.randrange(len(b) + 1)
            b[pos:pos] = rng.choice(SPLIT_PIECES)
        if rng.random() < 0.15:
            b = b[: rng.randrange(len(b) + 1)]
        yield bytes(b)


@pytest.mark.parametrize("seed", range(6))
def test_mutation_fuzz(lib, seed):
    counts = {"ok": 0, "unsupported": 0}
    for b in mutated_chunks(seed, 2500):
        for prov in ("ollama", "llamacpp", "litellm"):
            counts[check(lib, prov, b)] += 1
    assert counts["ok"] > 6000 and counts["unsupported"] < counts["ok"] // 10


def test_two_pass_form_covers_the_matrix_and_falls_back(lib, extract_cases):
    """emu_extract runs the kernel's two-pass extractor next to the single pass on every call and fails
    with 0xdea* if they disagree; here: the matrix is handled by the two-pass form itself, and chunks
    with more events than a thread's list holds fall back without changing the answer."""
    calls, taken = C.c_uint64(), C.c_uint64()
    lib.emu_extract_fast_counts(C.byref(calls), C.byref(taken))
    c0, t0 = calls.value, taken.value
    for c in extract_cases:
        run(lib, c["provider"], bytes.fromhex(c["chunk"]))
    lib.emu_extract_fast_counts(C.byref(calls), C.byref(taken))
    assert calls.value - c0 == len(extract_cases)
    # deeply nested cases and a second "data:" in an unusual place go to the single pass
    assert taken.value - t0 >= len(extract_cases) - 50
    c0, t0 = calls.value, taken.value
    busy = [b"data: [" + b"[]," * 120 + b"[]]", b'{"content":[' + b"1," * 200 + b"1]}",
            b'data: {"choices":[{"delta":{"content":"ok"},"x":[' + b"0," * 150 + b"0]}]}"]
    for chunk in busy:
        for prov in ("openai", "ollama", "llamacpp"):
            assert check(lib, prov, chunk) in ("ok", "unsupported")
    lib.emu_extract_fast_counts(C.byref(calls), C.byref(taken))
    assert calls.value - c0 == 9 and taken.value == t0      # every one of them overflowed the event list
    st, got = run(lib, "ollama", busy[2])
    assert (st, got) == (0, b"ok")
    # two events in one chunk are split by the grammar itself; "data:" inside a string is not
    for chunk, fast in ((b'data: {"content":"a"}\n\ndata: {"content":"b"}\n\n', 1), (b'data: {"content":"a data: b"}', 1), (b'{"content":"a data: b"}', 0),
                        (b'data: 12data: 13', 0), (b'data: truedata:', 1), (b'data: {"content":"\\u00data:"}', 1)):
        lib.emu_extract_fast_counts(C.byref(calls), C.byref(taken))
        t0 = taken.value
        assert check(lib, "llamacpp", chunk) == "ok"
        lib.emu_extract_fast_counts(C.byref(calls), C.byref(taken))
        assert taken.value - t0 == fast, chunk
    # nesting: the two-pass form takes 13 levels, the single pass 64, 65 are left to the host
    for depth, want, fast in ((13, 0, 1), (14, 0, 0), (64, 0, 0), (65, 4, 0)):
        lib.emu_extract_fast_counts(C.byref(calls), C.byref(taken))
        t0 = taken.value
        st, got = run(lib, "llamacpp", b"data: " + b"[" * depth + b"]" * depth)
        lib.emu_extract_fast_counts(C.byref(calls), C.byref(taken))
        assert st == want and taken.value - t0 == fast


@pytest.mark.parametrize("seed", range(3))
def test_second_prefix_fuzz(lib, seed):
    """Pieces of "data:" dropped anywhere in a chunk: the split the reference does (src/utils.ts:22) is
    handled by the grammar table in the two-pass form and by a look-ahead in the single pass; both have
    to agree with the oracle wherever the pieces land (inside strings, escapes, numbers, between events)."""
    n = 0
    for b in split_chunks(seed, 2500):
        for prov in ("ollama", "llamacpp", "litellm"):
            n += check(lib, prov, b) == "ok"
    assert n > 7000
