"""GPU tests of connection set-up in batch and of the receive side's first message and keep-alive rules
(SURVEY.md 8(f) N1 / N3).  In the reference all of this happens inside @hyperswarm/secret-stream when hyperswarm
raises 'connection' (/root/reference/src/provider.ts:55-58; state typed at src/types.ts:156-159, 178-179): each
side sends uint24(56) ‖ streamId ‖ header in the clear, checks the peer's id, derives the secretstream state from
the peer's header, and swallows empty keep-alive messages.  The oracle for the ids is libsodium's
crypto_generichash (PyNaCl), for the states libsodium's init_pull."""
import random

import nacl.bindings as nb
import numpy as np
import pytest

from oracle import libsodium as ls

pytestmark = pytest.mark.gpu

NS_I = bytes(range(32))            # stand-ins for hypercore-crypto's namespace('hyperswarm/secret-stream', 3)[0..1]
NS_R = bytes(range(100, 132))


@pytest.fixture()
def relay():
    import symmetry_b200 as sb

    r = sb.SymRelay(gpu_ids=(0,), max_streams=4096, flush_us=0)
    yield r
    r.close()


def test_connect_batch_ids_and_headers(relay):
    rng = random.Random(1)
    n = 700
    hh = np.frombuffer(rng.randbytes(64 * n), dtype=np.uint8).reshape(n, 64)
    ini = np.array([rng.random() < 0.5 for _ in range(n)], dtype=np.uint8)
    lid, rid, hdr, fm = relay.connect_batch(hh, ini, NS_I, NS_R, seed=bytes(32))
    for i in range(n):
        mine, theirs = (NS_I, NS_R) if ini[i] else (NS_R, NS_I)
        assert bytes(lid[i]) == nb.crypto_generichash_blake2b_salt_personal(mine, digest_size=32, key=bytes(hh[i]))
        assert bytes(rid[i]) == nb.crypto_generichash_blake2b_salt_personal(theirs, digest_size=32, key=bytes(hh[i]))
        assert bytes(fm[i]) == b"\x38\x00\x00" + bytes(lid[i]) + bytes(hdr[i])
    assert len({bytes(h) for h in hdr}) == n, "headers must all differ"
    lid2, _, hdr2, _ = relay.connect_batch(hh, ini, NS_I, NS_R, seed=bytes(32))
    assert (lid2 == lid).all() and (hdr2 == hdr).all(), "same seed, same headers (the seed is the only randomness)"
    _, _, hdr3, _ = relay.connect_batch(hh, ini, NS_I, NS_R)           # OS entropy
    assert not (hdr3 == hdr).all()
    # the headers initialise outbound streams: sealed frames open with libsodium's init_pull on the same header
    keys = np.frombuffer(rng.randbytes(32 * 16), dtype=np.uint8).reshape(16, 32)
    slots = relay.open_streams(keys, hdr3[:16], None, None)
    for i, s in enumerate(slots):
        t = relay.submit_seal(int(s), b"hello %d" % i)
        relay.flush()
        (c,) = relay.poll(wait=True)
        assert c["ticket"] == t and c["data"] == ls.State(bytes(keys[i]), bytes(hdr3[i])).seal_frame(b"hello %d" % i)


def test_first_message_then_frames_and_keepalive(relay):
    import symmetry_b200 as sb

    rng = random.Random(2)
    hh = np.frombuffer(rng.randbytes(64 * 3), dtype=np.uint8).reshape(3, 64)
    # our side is the responder of all three; the peers are initiators
    _, want_remote, _, _ = relay.connect_batch(hh, np.zeros(3, np.uint8), NS_I, NS_R)
    peer_ids = [nb.crypto_generichash_blake2b_salt_personal(NS_I, digest_size=32, key=bytes(hh[i])) for i in range(3)]
    assert [bytes(x) for x in want_remote] == peer_ids
    keys = [rng.randbytes(32) for _ in range(3)]
    hdrs = [rng.randbytes(24) for _ in range(3)]
    slots = [relay.open_stream(None, None, None, None) for _ in range(3)]
    for s, k, rid in zip(slots, keys, want_remote):
        relay.rx_expect_header(s, k, bytes(rid))
    relay.set_keepalive(slots[0], True)
    # connection 0: first message and frames in one segment, cut in odd places; keep-alive frames are swallowed
    sender = ls.State(keys[0], hdrs[0])
    msgs = [b"one", b"", b"two" * 100, b"", b"", b"three"]
    blob = b"\x38\x00\x00" + peer_ids[0] + hdrs[0] + b"".join(sender.seal_frame(m) for m in msgs)
    tickets, pos = [], 0
    for cut in (2, 40, 59, 61, 200, len(blob)):
        tk, used = relay.rx_feed(slots[0], blob[pos:cut])
        assert used == cut - pos
        tickets += tk
        pos = cut
    assert len(tickets) == len(msgs)
    relay.flush()
    got = {}
    for _ in range(100):
        for c in relay.poll(wait=True):
            got[c["ticket"]] = c
        relay.flush()
        if len(got) == 3:
            break
    delivered = [got[t]["data"] for t in tickets if t in got]
    assert delivered == [b"one", b"two" * 100, b"three"] and all(c["status"] == 0 for c in got.values())
    mirror = ls.State(keys[0], hdrs[0])
    for m in msgs:
        mirror.seal_frame(m)
    assert relay.export(slots[0])[1] == mirror.raw, "swallowed keep-alives still advance the chain"
    # connection 1: keep-alive off -> the empty frame is delivered like any other
    sender = ls.State(keys[1], hdrs[1])
    blob = b"\x38\x00\x00" + peer_ids[1] + hdrs[1] + sender.seal_frame(b"") + sender.seal_frame(b"x")
    tk, used = relay.rx_feed(slots[1], blob)
    assert used == len(blob) and len(tk) == 2
    relay.flush()
    got = {}
    while len(got) < 2:
        for c in relay.poll(wait=True):
            got[c["ticket"]] = c
    assert [got[t]["data"] for t in tk] == [b"", b"x"]
    # connection 2: wrong stream id -> the connection is refused, nothing is opened
    with pytest.raises(sb.RelayError) as e:
        relay.rx_feed(slots[2], b"\x38\x00\x00" + bytes(32) + hdrs[2])
    assert e.value.code == sb.relay.EPROTO
    with pytest.raises(sb.RelayError) as e:
        relay.rx_feed(slots[2], b"\x10\x00\x00" + bytes(16))      # a first message of the wrong size
    assert e.value.code == sb.relay.EPROTO
