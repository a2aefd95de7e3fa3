"""GPU parity for the fused extractor: SYMRELAY_F_EXTRACT frames through the
C ABI.  Deltas and THROW verdicts must equal oracle/extract.py; frames the reference would not have
written (TypeError) or that the GPU hands back (UNSUPPORTED) must leave the stream state untouched,
all others must be sealed bit-exactly (libsodium)."""
import numpy as np
import pytest

import batch_util as bu
from oracle import extract as ex
from oracle import libsodium as ls

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=["zero_copy", "staged"])
def relay(request):
    """Both host paths: small pinned batches run zero-copy (kernels read/write the pinned host
    arenas over PCIe); with zero_copy off the same batches go through the staged DMA pipeline."""
    import symmetry_b200 as sb

    r = sb.SymRelay(gpu_ids=(0,), max_streams=1 << 16, zero_copy=request.param == "zero_copy")
    r.test_mode = request.param
    yield r
    r.close()


def run_extract_batch(relay, slots, chunks):
    import symmetry_b200 as sb

    lens = [len(c) for c in chunks]
    p_offs, p_bytes = bu.plain_layout(lens)
    w_offs, w_bytes = bu.wire_layout(lens)
    plain = relay.alloc_pinned(p_bytes)
    plain[:] = 0
    for c, o in zip(chunks, p_offs):
        plain[o : o + len(c)] = np.frombuffer(c, dtype=np.uint8)
    wire = relay.alloc_pinned(w_bytes)
    wire[:] = 0
    darena = relay.alloc_pinned(p_bytes)
    delta = np.zeros(len(chunks), dtype=sb.DELTA_DTYPE)
    d = bu.make_descs(slots, lens, p_offs, w_offs, extract=True)
    st = relay.batch_host(d, plain, wire, delta=delta, delta_arena=darena)
    deltas = [bytes(darena[int(x["off"]) : int(x["off"]) + int(x["len"])]) for x in delta]
    return st, deltas, wire, w_offs


def test_quirk_matrix_through_the_abi(relay, extract_cases):
    import symmetry_b200 as sb

    rng = np.random.default_rng(5)
    by_provider = {}
    for c in extract_cases:
        by_provider.setdefault(c["provider"], []).append(c)
    for prov, cases in by_provider.items():
        n = len(cases)
        keys = rng.integers(0, 256, (n, 32), dtype=np.uint8)
        hdrs = rng.integers(0, 256, (n, 24), dtype=np.uint8)
        slots = relay.open_streams(keys, hdrs, None, None, provider=sb.provider_kind(prov))
        chunks = [bytes.fromhex(c["chunk"]) for c in cases]
        st, deltas, wire, w_offs = run_extract_batch(relay, slots, chunks)
        n_uns = 0
        for i, c in enumerate(cases):
            code = int(st[i]) & 0xFF
            ref = ls.State(bytes(keys[i]), bytes(hdrs[i]))
            if code == sb.relay.ST_EXTRACT_UNSUPPORTED:
                n_uns += 1
                assert relay.export(int(slots[i]))[0] == ref.raw, "unsupported frame must not be sealed"
                continue
            if c["status"] == ex.THROW:
                assert code == sb.relay.ST_EXTRACT_THROW, (prov, chunks[i])
                assert relay.export(int(slots[i]))[0] == ref.raw, "TypeError chunk must not be sealed"
                continue
            assert code == 0, (prov, chunks[i], code)
            assert deltas[i].hex() == c["delta"], (prov, chunks[i])
            want = ref.seal_frame(chunks[i])
            assert bytes(wire[w_offs[i] : w_offs[i] + len(want)]) == want
            assert relay.export(int(slots[i]))[0] == ref.raw
        assert n_uns <= 2, (prov, n_uns)
        for s in slots:
            relay.close_stream(int(s))


def test_config5_chat_completions_accumulate(relay):
    """BASELINE config 5 in miniature: many streams of synthetic llama3:8b completions, apiProvider
    ollama; per batch one chunk per stream; wire frames and per-stream `completion` strings."""
    import symmetry_b200 as sb
    from symmetry_b200 import synth

    n, n_chunks = 2048, 6
    keys, hdrs = synth.stream_keys(n)
    slots = relay.open_streams(keys, hdrs, None, None, provider=sb.PROVIDER_OLLAMA)
    streams = [synth.chat_chunks(n_chunks, seed=77, stream_id=i) for i in range(n)]
    completion = [b""] * n
    want_completion = [b""] * n
    refs = {i: ls.State(bytes(keys[i]), bytes(hdrs[i])) for i in range(0, n, 97)}
    for step in range(n_chunks + 2):
        chunks = [s[step] for s in streams]
        st, deltas, wire, w_offs = run_extract_batch(relay, slots, chunks)
        assert (st == 0).all()
        for i in range(n):
            completion[i] += deltas[i]
            code, dl = ex.extract("ollama", chunks[i])
            assert code == ex.OK
            want_completion[i] += dl
        for i, ref in refs.items():
            want = ref.seal_frame(chunks[i])
            assert bytes(wire[w_offs[i] : w_offs[i] + len(want)]) == want
    assert completion == want_completion
    assert all(len(c) > 0 for c in completion)


@pytest.mark.parametrize("provider", ["ollama", "llamacpp", "litellm"])
def test_fuzzed_chunks_match_the_oracle(relay, provider):
    """The emulator tests' mutation and second-"data:" fuzz, here compiled by nvcc and run on the GPU:
    6,000 chunks in one batch, one stream each, every verdict and delta against oracle/extract.py."""
    import symmetry_b200 as sb
    from test_emu_extract import mutated_chunks, split_chunks

    chunks = list(mutated_chunks(11, 4000)) + list(split_chunks(11, 2000))
    chunks = [c if c else b" " for c in chunks]
    n = len(chunks)
    rng = np.random.default_rng(23)
    keys = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    hdrs = rng.integers(0, 256, (n, 24), dtype=np.uint8)
    slots = relay.open_streams(keys, hdrs, None, None, provider=sb.provider_kind(provider))
    st, deltas, wire, w_offs = run_extract_batch(relay, slots, chunks)
    n_ok = n_throw = n_uns = 0
    for i, c in enumerate(chunks):
        code = int(st[i]) & 0xFF
        want_code, want = ex.extract(provider, c)
        if code == sb.relay.ST_EXTRACT_UNSUPPORTED:
            n_uns += 1
            continue
        if want_code == ex.THROW:
            assert code == sb.relay.ST_EXTRACT_THROW, (provider, c)
            n_throw += 1
            continue
        assert code == 0 and deltas[i] == want, (provider, c, code, deltas[i], want)
        n_ok += 1
    assert n_ok > 3000 and n_uns < n // 10, (n_ok, n_throw, n_uns)
    for i in range(0, n, 61):   # sealed frames of accepted chunks are libsodium's; rejected ones left the state alone
        ref = ls.State(bytes(keys[i]), bytes(hdrs[i]))
        if int(st[i]) & 0xFF == 0:
            want = ref.seal_frame(chunks[i])
            assert bytes(wire[w_offs[i] : w_offs[i] + len(want)]) == want
        assert relay.export(int(slots[i]))[0] == ref.raw
    # a refused chunk leaves no ciphertext behind (the seal kernel may run beside the extractor and takes back what it sealed
    # ahead of a bad verdict): its wire slot is all zeros, as the arena was.  (Staged batches copy the whole device arena
    # back, gaps and refused slots included, so there the slot holds whatever an earlier batch left in the staging.)
    for i, c in enumerate(chunks):
        if relay.test_mode == "zero_copy" and int(st[i]) & 0xFF != 0:
            assert not wire[w_offs[i] : w_offs[i] + len(c) + 20].any(), (i, c)
    for s_ in slots:
        relay.close_stream(int(s_))


def test_seal_beside_the_extractor_takes_back_refused_chunks(relay):
    """8,192 chunks of one size class (thread tiles, one kernel): the seal kernel is launched as the extractor's
    programmatic dependent and asks for the verdicts last.  Every fifth chunk is one the reference throws on or the GPU
    hands back: status from the extractor, stream state untouched, no ciphertext left in its wire slot; all others
    sealed bit-exactly with the right delta.  A second batch on the same streams shows the chains moved (or not)."""
    import json

    import symmetry_b200 as sb

    n = 8192
    rng = np.random.default_rng(99)
    keys = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    hdrs = rng.integers(0, 256, (n, 24), dtype=np.uint8)
    slots = relay.open_streams(keys, hdrs, None, None, provider=sb.provider_kind("litellm"))
    refs = [ls.State(bytes(keys[i]), bytes(hdrs[i])) for i in range(n)]
    bad = [b'data: {"choices":[]}', b'data: {"choices":[{"delta":{"content":1e300}}]}', b'data: {"id":1}',
           b'data: {"choices":[{"index":0}]}']
    for step in range(2):
        chunks = []
        for i in range(n):
            if (i + step) % 5 == 0:
                c = bad[(i // 5) % len(bad)]
            else:
                ev = {"id": f"chatcmpl-{i:06d}", "object": "chat.completion.chunk", "model": "llama3:8b",
                      "choices": [{"index": 0, "delta": {"role": "assistant", "content": f"tok{i}s{step}"}, "finish_reason": None}]}
                c = b"data: " + json.dumps(ev, separators=(",", ":")).encode()
            assert len(c) <= 250
            chunks.append(c + b" " * (200 + (i * 7) % 51 - len(c)) if len(c) < 200 + (i * 7) % 51 else c)
        assert all(193 <= len(c) <= 256 for c in chunks)   # four payload blocks each: one thread tile per 32 frames
        st, deltas, wire, w_offs = run_extract_batch(relay, slots, chunks)
        n_bad = 0
        for i, c in enumerate(chunks):
            code = int(st[i]) & 0xFF
            want_code, want = ex.extract("litellm", c)
            if (i + step) % 5 == 0:
                n_bad += 1
                assert code in (sb.relay.ST_EXTRACT_THROW, sb.relay.ST_EXTRACT_UNSUPPORTED), (i, c, code)
                assert code == sb.relay.ST_EXTRACT_UNSUPPORTED or want_code == ex.THROW
                if relay.test_mode == "zero_copy":
                    assert not wire[w_offs[i] : w_offs[i] + len(c) + 20].any(), (i, c)
            else:
                assert code == 0 and want_code == ex.OK and deltas[i] == want, (i, c, code)
                frame = refs[i].seal_frame(c)
                assert bytes(wire[w_offs[i] : w_offs[i] + len(frame)]) == frame, i
            if i % 13 == 0 or (i + step) % 5 == 0:
                assert relay.export(int(slots[i]))[0] == refs[i].raw, i
        assert n_bad > 1600
    for s_ in slots:
        relay.close_stream(int(s_))


def test_extract_flag_needs_delta_buffers(relay):
    import symmetry_b200 as sb

    d = bu.make_descs([0], [16], [0], [12], extract=True)
    with pytest.raises(sb.RelayError):
        relay.batch_host(d, bu.aligned_zeros(64), bu.aligned_zeros(64))
