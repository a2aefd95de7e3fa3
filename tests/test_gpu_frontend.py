"""GPU tests of the per-frame front end (symrelay_submit_seal/_open/flush/poll/release) and of the
host mirror of the reference's relay loop (provider.ts:195-275).  An "unmodified reference client"
is modelled by a plain libsodium init_pull + pull loop with uint24 de-framing."""
import json
import random

import numpy as np
import pytest

from oracle import extract as ex
from oracle import libsodium as ls

pytestmark = pytest.mark.gpu


@pytest.fixture()
def relay():
    import symmetry_b200 as sb

    r = sb.SymRelay(gpu_ids=(0,), max_streams=8192, flush_us=0)
    yield r
    r.close()


def test_per_stream_fifo_and_chain_with_interleaved_submits(relay):
    rng = random.Random(1)
    n = 200
    keys = [rng.randbytes(32) for _ in range(n)]
    hdrs = [rng.randbytes(24) for _ in range(n)]
    slots = [relay.open_stream(k, h, k, h) for k, h in zip(keys, hdrs)]
    refs = [ls.State(k, h) for k, h in zip(keys, hdrs)]
    import symmetry_b200 as sb

    want = {}
    got_order = {i: [] for i in range(n)}
    n_eagain = 0

    def collect(wait):
        for c in relay.poll(wait=wait):
            i, m = want.pop(c["ticket"])
            assert c["status"] == 0 and not c["open"]
            got_order[i].append((m, c["data"]))

    def write(i, m):
        nonlocal n_eagain
        while True:
            try:
                want[relay.submit_seal(slots[i], m)] = (i, m)
                return
            except sb.RelayError as e:      # `peer.write()` returned false: await 'drain' (provider.ts:250-252)
                assert e.code == sb.relay.EAGAIN
                n_eagain += 1
                collect(wait=True)

    # several frames per stream submitted back to back: the front end must split them over batches
    for rnd in range(5):
        for i in rng.sample(range(n), n):
            write(i, rng.randbytes(rng.choice([0, 1, 64, 200, 3000])))
            if rng.random() < 0.3:
                write(i, rng.randbytes(rng.choice([5, 64])))
    relay.flush()
    while want:
        collect(wait=True)
        relay.flush()
    assert n_eagain > 0   # this pattern closes a batch per repeated stream and must hit back-pressure
    for i in range(n):
        for m, wire in got_order[i]:          # completions of a stream come in submit order = chain order
            assert wire == refs[i].seal_frame(m)
    st = relay.sched_stats()
    assert st["frames"] >= 1000 and st["batches"] >= 2


def test_open_side_and_host_rejections(relay):
    rng = random.Random(2)
    k, h = rng.randbytes(32), rng.randbytes(24)
    slot = relay.open_stream(k, h, k, h)
    peer = ls.State(k, h)   # the remote end's push state (same key/header as our rx)
    frames = [peer.seal_frame(rng.randbytes(L)) for L in (0, 1, 63, 64, 500)]
    msgs = []
    mirror = ls.State(k, h)
    for f in frames:
        msgs.append(mirror.open_frame(f)[0])
    tickets = []
    import symmetry_b200 as sb

    res = {}
    for f in frames:
        tickets.append(relay.submit_open(slot, f))
        while True:
            try:
                relay.flush()
                break
            except sb.RelayError as e:      # every pipeline stage holds an unfinished batch: drain one and retry
                assert e.code == sb.relay.EAGAIN
                for c in relay.poll(wait=True):
                    res[c["ticket"]] = c
    tickets.append(relay.submit_open(slot, b"\x05\x00\x00abcde"))            # < 17 bytes after the prefix
    tickets.append(relay.submit_open(slot, b"\x30\x00\x00" + bytes(30)))      # prefix disagrees with the length
    forged = bytearray(peer.seal_frame(b"forged?"))
    forged[10] ^= 1
    tickets.append(relay.submit_open(slot, bytes(forged)))
    relay.flush()
    while len(res) < len(tickets):
        for c in relay.poll(wait=True):
            res[c["ticket"]] = c
    for t, m in zip(tickets[:5], msgs):
        assert res[t]["status"] == 0 and res[t]["open"] and res[t]["data"] == m
    assert res[tickets[5]]["status"] == 2 and res[tickets[6]]["status"] == 2
    assert res[tickets[7]]["status"] == 1 and res[tickets[7]]["data"] == b""
    assert relay.export(slot)[1] == mirror.raw   # rejected and forged frames left the rx state alone


def test_ring_full_returns_eagain_then_drains():
    import symmetry_b200 as sb

    with sb.SymRelay(gpu_ids=(0,), max_streams=64, sched_batches=2, sched_batch_frames=4) as r:
        k, h = bytes(range(32)), bytes(24)
        slots = [r.open_stream(k, h, None, None) for _ in range(16)]
        accepted, refused = [], 0
        for s in slots:
            try:
                accepted.append(r.submit_seal(s, b"x" * 40))
            except sb.RelayError as e:
                assert e.code == sb.relay.EAGAIN
                refused += 1
        assert refused > 0 and len(accepted) >= 4     # `write()` false: the loop must await 'drain'
        r.flush()
        done = 0
        while done < len(accepted):
            done += len(r.poll(wait=True))
            r.flush()
        t = r.submit_seal(slots[-1], b"after drain")   # space again
        r.flush()
        assert [c["status"] for c in r.poll(wait=True)] == [0] and t > 0


def test_relay_loop_mirror_against_reference_semantics(relay):
    """provider.ts:195-275 for many connections: preamble, chunks (+delta), trailer; a TypeError chunk
    ends that connection's relay without a trailer; an unmodified client decrypts what was sent."""
    import symmetry_b200 as sb
    from symmetry_b200 import provider, synth

    rng = random.Random(3)
    n = 300
    keys = [rng.randbytes(32) for _ in range(n)]
    hdrs = [rng.randbytes(24) for _ in range(n)]
    peers = []
    for i in range(n):
        slot = relay.open_stream(keys[i], hdrs[i], None, None, provider=sb.PROVIDER_OLLAMA)
        chunks = synth.chat_chunks(rng.randrange(1, 9), seed=9, stream_id=i)
        if i % 10 == 3:   # OpenAI usage chunk mid-stream: `"choices":[]` -> TypeError in the reference
            chunks.insert(rng.randrange(len(chunks)), b'data: {"choices":[],"usage":{"total_tokens":3}}\n\n')
        if i % 10 == 5:   # a value shape the GPU hands back to the host helper
            chunks.insert(1, b'data: {"choices":[{"delta":{"content":1e21}}]}\n\n')
        peers.append(provider.PeerRelayState(slot, f"key-{i}", chunks))

    def host_extract(p, chunk):   # stands in for the reference's own JS helper
        code, delta = ex.extract("ollama", chunk)
        return code == ex.OK, delta

    provider.handle_inference_requests(relay, peers, host_extract=host_extract)
    for i, p in enumerate(peers):
        client = ls.State(keys[i], hdrs[i])                  # unmodified client: init_pull + pull loop
        got = [client.open_frame(w) for w in p.wire]
        assert all(g is not None and g[1] == 0 for g in got)
        plain = [g[0] for g in got]
        # what the reference would have done, restated with the oracle
        want_frames = [json.dumps({"symmetryEmitterKey": f"key-{i}"}, separators=(",", ":")).encode()]
        want_completion, aborted = b"", False
        for ch in p.frames[1:-1]:
            code, delta = ex.extract("ollama", ch)
            if code == ex.THROW:
                aborted = True
                break
            want_completion += delta
            want_frames.append(ch)
        if not aborted:
            want_frames.append(provider.createMessage("inferenceEnded", f"key-{i}").encode())
        assert p.aborted == aborted
        assert plain == want_frames
        assert p.completion == want_completion
        if i % 10 == 5 and not aborted:
            assert p.unsupported == [2]
    assert any(p.aborted for p in peers) and any(not p.aborted for p in peers)
    # the library counted how the fused extractor fared, per provider kind: every connection with the 1e21 chunk that
    # got that far cost one UNSUPPORTED, every aborted one a THROW
    st = relay.extract_stats()[sb.PROVIDER_OLLAMA]
    assert st["unsupported"] == sum(1 for p in peers if p.unsupported) and st["throw"] == sum(1 for p in peers if p.aborted)
    assert st["ok"] > 0 and st["other"] == 0 and relay.extract_stats()[sb.PROVIDER_DEFAULT]["ok"] == 0


def test_deadline_flush_and_latency_stats():
    import time

    import symmetry_b200 as sb

    with sb.SymRelay(gpu_ids=(0,), max_streams=4096, flush_us=100) as r:
        k, h = bytes(range(32)), bytes(24)
        slots = [r.open_stream(k, h, None, None) for _ in range(1000)]
        got = 0
        for s in slots:
            while True:
                try:
                    r.submit_seal(s, b"tok " * 16)
                    break
                except sb.RelayError as e:
                    assert e.code == sb.relay.EAGAIN
                    got += len(r.poll())
        deadline = time.time() + 5
        while got < 1000 and time.time() < deadline:   # no explicit flush: the deadline closes batches
            got += len(r.poll())
        assert got == 1000
        st = r.sched_stats()
        assert st["frames"] == 1000 and 0 < st["p50_us"] < 50000 and st["p99_us"] >= st["p50_us"]


def test_frames_already_queued_behind_a_rekey_run_come_back_retry_not_forged(relay):
    """A run stops at a REKEY-tagged frame (the frames behind it used the wrong key: RETRY).  Frames of the same
    stream that were fed BEFORE the caller saw those completions were opened against the lagging state; they are
    not forgeries and must come back RETRY too, so that a caller who resubmits every RETRY in arrival order ends up
    with the whole stream -- the reference, one pull at a time, never sees any of this."""
    import symmetry_b200 as sb

    rng = random.Random(15)
    k, h = rng.randbytes(32), rng.randbytes(24)
    slot = relay.open_stream(None, None, k, h)
    sender = ls.State(k, h)
    msgs = [rng.randbytes(n) for n in (10, 70, 0, 300, 64, 5, 1000, 33)]
    tags = [0, 2, 0, 0, 0, 0, 0, 0]                       # frame 1 carries TAG_REKEY
    frames = [sender.seal_frame(m, t) for m, t in zip(msgs, tags)]
    tk1, used = relay.rx_feed(slot, b"".join(frames[:4]))           # one segment: a run of four
    assert used == sum(map(len, frames[:4])) and len(tk1) == 4
    relay.flush()
    tk2, used = relay.rx_feed(slot, b"".join(frames[4:6]))          # the next segment, before any completion was read
    relay.flush()
    tk3, used = relay.rx_feed(slot, b"".join(frames[6:]))
    relay.flush()
    tickets = tk1 + tk2 + tk3
    got = {}
    while len(got) < len(tickets):
        for c in relay.poll(wait=True):
            got[c["ticket"]] = c
    st = [got[t]["status"] for t in tickets]
    assert st[:2] == [0, 0] and got[tickets[0]]["data"] == msgs[0] and got[tickets[1]]["tag"] == 2
    assert st[2:] == [sb.relay.ST_RETRY] * 6, st            # none of them BAD_MAC
    # resubmit in arrival order (a few per batch): everything opens, the chain is libsodium's
    redo = []
    for i in range(2, 8):
        redo.append((relay.submit_open(slot, frames[i]), i))
        if i % 2:
            relay.flush()
    relay.flush()
    got = {}
    while len(got) < len(redo):
        for c in relay.poll(wait=True):
            got[c["ticket"]] = c
        relay.flush()
    for t, i in redo:
        assert got[t]["status"] == 0 and got[t]["data"] == msgs[i], i
    mirror = ls.State(k, h)
    for f in frames:
        mirror.open_frame(f)
    assert relay.export(slot)[1] == mirror.raw


def test_transcript_written_from_gpu_extractor_output(relay, tmp_path):
    """`saveCompletion` (provider.ts:264-269, 277-297) fed with the deltas the GPU extractor produced
    while sealing -- split surrogate pairs across chunks, a lone surrogate, escapes -- writes the file the reference
    would have written from its JS string (restated with the oracle's deltas)."""
    import symmetry_b200 as sb
    from symmetry_b200 import provider, synth

    rng = random.Random(14)
    peers, keys = [], []
    for i in range(40):
        k, h = rng.randbytes(32), rng.randbytes(24)
        keys.append(k)
        slot = relay.open_stream(k, h, None, None, provider=sb.PROVIDER_OLLAMA)
        chunks = synth.chat_chunks(12, seed=21, stream_id=i)
        chunks.insert(3, b'data: {"choices":[{"delta":{"content":"\\ud83d"}}]}\n\n')     # halves of one pair in two chunks
        chunks.insert(4, b'data: {"choices":[{"delta":{"content":"\\ude00 \\ud800! \\"q\\" \\\\ \\n"}}]}\n\n')
        peers.append(provider.PeerRelayState(slot, f"key-{i}", chunks))
    provider.handle_inference_requests(relay, peers)
    msgs = [{"role": "user", "content": "hi"}]
    for i, p in enumerate(peers):
        assert not p.aborted
        want = b"".join(ex.extract("ollama", c)[1] for c in p.frames[1:-1])
        assert p.completion == want, "GPU deltas differ from the oracle's"
        a = provider.saveCompletion(p.completion, keys[i], i, msgs, str(tmp_path))
        b = provider.saveCompletion(want, keys[i], i + 1000, msgs, str(tmp_path))
        ta, tb = open(a, encoding="utf-8").read(), open(b, encoding="utf-8").read()
        assert ta == tb and "\U0001F600" in ta and "\\ud800!" in ta
        assert json.loads(ta)[1]["role"] == "assistant"


def test_event_fd_wakes_the_loop_and_flushes_on_deadline():
    """the event loop must never block on the GPU nor poll for it.  The context's eventfd becomes readable
    when a batch has finished, and the watcher thread performs the deadline flush, so a loop that only select()s on
    the descriptor gets its completions: no symrelay_flush, no symrelay_poll before the wake-up."""
    import os
    import select

    import symmetry_b200 as sb

    rng = random.Random(8)
    with sb.SymRelay(gpu_ids=(0,), max_streams=256, flush_us=300) as r:
        fd = r.event_fd()
        assert fd >= 0 and r.event_fd() == fd
        keys = [rng.randbytes(32) for _ in range(20)]
        hdrs = [rng.randbytes(24) for _ in range(20)]
        slots = [r.open_stream(k, h, None, None) for k, h in zip(keys, hdrs)]
        refs = [ls.State(k, h) for k, h in zip(keys, hdrs)]
        for rnd in range(3):
            msgs = [rng.randbytes(rng.choice([0, 64, 300])) for _ in slots]
            want = {r.submit_seal(s, m): (i, m) for i, (s, m) in enumerate(zip(slots, msgs))}
            got = {}
            for _ in range(50):
                ready, _, _ = select.select([fd], [], [], 1.0)
                assert ready, "the event descriptor never became readable: no deadline flush or no wake-up"
                assert int.from_bytes(os.read(fd, 8), "little") >= 1
                for c in r.poll():
                    got[c["ticket"]] = c
                if len(got) == len(want):
                    break
            assert len(got) == len(want)
            for t, (i, m) in want.items():
                assert got[t]["status"] == 0 and got[t]["data"] == refs[i].seal_frame(m)
        assert not select.select([fd], [], [], 0.05)[0], "readable with nothing finished"


def test_receive_side_reassembly_and_runs(relay):
    """secret-stream `_onrawdata` semantics: frames arrive as a byte stream cut at arbitrary points
    (inside the uint24 length too); all frames completed by a segment are opened in one batch as a run."""
    rng = random.Random(4)
    n = 40
    keys = [rng.randbytes(32) for _ in range(n)]
    hdrs = [rng.randbytes(24) for _ in range(n)]
    slots = [relay.open_stream(None, None, k, h) for k, h in zip(keys, hdrs)]
    senders = [ls.State(k, h) for k, h in zip(keys, hdrs)]
    expect = {i: [] for i in range(n)}
    streams = []
    for i in range(n):
        blob = b""
        for _ in range(rng.randrange(1, 30)):
            m = rng.randbytes(rng.choice([0, 0, 1, 40, 64, 200, 200, 1500, 4000]))   # 0 = keep-alive frame
            expect[i].append(m)
            blob += senders[i].seal_frame(m)
        streams.append(blob)
    # cut every connection's byte stream into random segments and feed them round-robin
    cursors = [0] * n
    tickets = {i: [] for i in range(n)}
    got = {}
    while any(c < len(s) for c, s in zip(cursors, streams)):
        for i in rng.sample(range(n), n):
            if cursors[i] >= len(streams[i]):
                continue
            seg = streams[i][cursors[i] : cursors[i] + rng.choice([1, 2, 3, 5, 19, 100, 700, 5000])]
            tk, used = relay.rx_feed(slots[i], seg)
            tickets[i] += tk
            cursors[i] += used
            if used < len(seg):                      # ring full: drain, the rest is fed again next round
                for c in relay.poll(wait=True):
                    got[c["ticket"]] = c
        relay.flush()
        for c in relay.poll():
            got[c["ticket"]] = c
    relay.flush()
    want_total = sum(len(v) for v in expect.values())
    while len(got) < want_total:
        for c in relay.poll(wait=True):
            got[c["ticket"]] = c
        relay.flush()
    for i in range(n):
        assert len(tickets[i]) == len(expect[i])
        for t, m in zip(tickets[i], expect[i]):
            assert got[t]["status"] == 0 and got[t]["open"] and got[t]["data"] == m
        ref = ls.State(keys[i], hdrs[i])
        off = 0
        blob = streams[i]
        while off < len(blob):
            ln = blob[off] | (blob[off + 1] << 8) | (blob[off + 2] << 16)
            assert ref.open_frame(blob[off : off + 3 + ln]) is not None
            off += 3 + ln
        assert relay.export(slots[i])[1] == ref.raw
    assert relay.sched_stats()["frames"] >= want_total


def test_two_gpus_in_one_context_shard_by_slot():
    """A Node provider is ONE process: slots are sharded slot % n_gpus inside one context, each GPU
    keeps its own state table and pipeline, and no state ever crosses GPUs."""
    import torch

    import symmetry_b200 as sb

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    rng = random.Random(6)
    n = 400
    keys = [rng.randbytes(32) for _ in range(n)]
    hdrs = [rng.randbytes(24) for _ in range(n)]
    with sb.SymRelay(gpu_ids=(0, 1), max_streams=4096) as r:
        karr = np.frombuffer(b"".join(keys), dtype=np.uint8).reshape(n, 32)
        harr = np.frombuffer(b"".join(hdrs), dtype=np.uint8).reshape(n, 24)
        slots = [int(s) for s in r.open_streams(karr, harr, karr, harr)]
        assert {s % 2 for s in slots} == {0, 1}
        refs = [ls.State(k, h) for k, h in zip(keys, hdrs)]
        want = {}
        for rnd in range(3):
            for i in range(n):
                m = rng.randbytes(rng.choice([0, 64, 200, 3000]))
                want[r.submit_seal(slots[i], m)] = (i, m)
            r.flush()
            assert len(want) == n, "tickets must be unique across the context's GPUs"
            got = {}
            for _ in range(10000):
                for c in r.poll(wait=True):
                    got[c["ticket"]] = c
                if len(got) >= n:
                    break
            assert len(got) == n
            for t, (i, m) in want.items():
                assert got[t]["status"] == 0 and got[t]["data"] == refs[i].seal_frame(m)
                # and the peer's frames come back through the same slot's rx state on the same GPU
            want = {}
        for i in (0, 1, n - 2, n - 1):
            assert r.export(slots[i])[0] == refs[i].raw
        # the extractor's tables are per GPU: chunks of streams on either GPU give the oracle's deltas
        import batch_util as bu
        from oracle import extract as ex
        from symmetry_b200 import synth

        ekeys, ehdrs = synth.stream_keys(256, first=5000)
        eslots = [int(s) for s in r.open_streams(ekeys, ehdrs, None, None, provider=sb.PROVIDER_OLLAMA)]
        chunks = [synth.chat_chunks(3, seed=9, stream_id=i)[1] for i in range(256)]
        for g in (0, 1):
            idx = [i for i in range(256) if eslots[i] % 2 == g]
            assert idx
            lens = [len(chunks[i]) for i in idx]
            p_offs, p_bytes = bu.plain_layout(lens)
            w_offs, w_bytes = bu.wire_layout(lens)
            plain, wire, darena = r.alloc_pinned(p_bytes), r.alloc_pinned(w_bytes), r.alloc_pinned(p_bytes)
            plain[:] = 0
            for i, o in zip(idx, p_offs):
                plain[o : o + len(chunks[i])] = np.frombuffer(chunks[i], dtype=np.uint8)
            delta = np.zeros(len(idx), dtype=sb.DELTA_DTYPE)
            d = bu.make_descs([eslots[i] for i in idx], lens, p_offs, w_offs, extract=True)
            st = r.batch_host(d, plain, wire, delta=delta, delta_arena=darena)
            assert (st == 0).all()
            for k, i in enumerate(idx):
                off, ln = int(delta[k]["off"]), int(delta[k]["len"])
                assert bytes(darena[off : off + ln]) == ex.extract("ollama", chunks[i])[1]
