"""Parity scenarios shared by the emulator test (CPU) and the GPU tests.

A backend hides where the stream states live and how a batch is run:
  open_streams(tx_states52, rx_states52) -> slots
  run(descs, arena_in, arena_out) -> status (uint32 per frame)
  export(slot) -> (tx52, rx52)
Every expectation comes from the real libsodium (oracle/libsodium.py)."""
import random

import numpy as np

import batch_util as bu
from oracle import libsodium as ls


def power_law_lens(rng: random.Random, n: int, alpha: float = 1.2, cap: int = 16384):
    """BASELINE config 3 sizes: L = min(cap, floor(32 * U^(-1/alpha))), U ~ (0,1]."""
    return [min(cap, int(32 * (1 - rng.random()) ** (-1 / alpha))) for _ in range(n)]


def seal_open_roundtrip(be, lens, tags, seed, chain=2, corrupt=()):
    rng = random.Random(seed)
    n = len(lens)
    keys = [rng.randbytes(32) for _ in range(n)]
    hdrs = [rng.randbytes(24) for _ in range(n)]
    ref_tx = [ls.State(k, h) for k, h in zip(keys, hdrs)]
    ref_rx = [ls.State(k, h) for k, h in zip(keys, hdrs)]
    slots = be.open_streams([s.raw for s in ref_tx], [s.raw for s in ref_rx])
    p_offs, p_bytes = bu.plain_layout(lens)
    w_offs, w_bytes = bu.wire_layout(lens)
    for step in range(chain):
        msgs = [rng.randbytes(L) for L in lens]
        plain = bu.aligned_zeros(p_bytes)
        for m, o in zip(msgs, p_offs):
            plain[o : o + len(m)] = np.frombuffer(m, dtype=np.uint8)
        wire = bu.aligned_zeros(w_bytes)
        wire[:] = 0xAA  # canary: nothing outside the frames may be touched
        d = bu.make_descs(slots, lens, p_offs, w_offs, tags=tags)
        status = be.run(d, plain, wire)
        expect = [s.seal_frame(m, t) for s, m, t in zip(ref_tx, msgs, tags)]
        canary = np.ones(w_bytes, dtype=bool)
        for i, (o, e) in enumerate(zip(w_offs, expect)):
            got = bytes(wire[o : o + len(e)])
            assert got == e, f"step {step} frame {i} len {lens[i]} tag {tags[i]}"
            canary[o : o + len(e)] = False
            assert be.export(slots[i])[0] == ref_tx[i].raw, f"tx state frame {i} len {lens[i]}"
            assert status[i] == (tags[i] << 8)
        if getattr(be, "preserves_gaps", True):  # host-arena batches overwrite the whole out arena
            assert (wire[canary] == 0xAA).all(), "seal wrote outside its frames"
        # open what was sealed (optionally corrupted)
        bad = set(corrupt) if step == 0 else set()
        for i in bad:
            pos = w_offs[i] + 3 + rng.randrange(lens[i] + 17)
            wire[pos] ^= 1 << rng.randrange(8)
        out = bu.aligned_zeros(p_bytes)
        out[:] = 0x55
        d2 = bu.make_descs(slots, lens, w_offs, p_offs, open_=True)
        before = [be.export(s)[1] for s in slots]
        status = be.run(d2, wire, out)
        for i, (m, o) in enumerate(zip(msgs, p_offs)):
            if i in bad:
                assert status[i] == 1, f"forged frame {i} accepted"
                assert be.export(slots[i])[1] == before[i]
                if getattr(be, "preserves_gaps", True):
                    assert (out[o : o + max(16, bu.align16(lens[i]))] == 0x55).all(), "plaintext released on bad MAC"
                elif lens[i] >= 8:
                    assert bytes(out[o : o + len(m)]) != m, "plaintext released on bad MAC"
                continue
            assert status[i] == (tags[i] << 8), f"open status frame {i} len {lens[i]}: {status[i]:#x}"
            assert bytes(out[o : o + len(m)]) == m
            assert ref_rx[i].open_frame(expect[i]) == (m, tags[i])
            assert be.export(slots[i])[1] == ref_rx[i].raw
        if bad:
            # pull left the state untouched, so the genuine frame must still open
            for o, e in zip(w_offs, expect):
                wire[o : o + len(e)] = np.frombuffer(e, dtype=np.uint8)
            idx = sorted(bad)
            d3 = bu.make_descs([slots[i] for i in idx], [lens[i] for i in idx], [w_offs[i] for i in idx],
                               [p_offs[i] for i in idx], open_=True)
            status = be.run(d3, wire, out)
            for j, i in enumerate(idx):
                assert status[j] == (tags[i] << 8)
                assert ref_rx[i].open_frame(expect[i]) == (msgs[i], tags[i])
                assert be.export(slots[i])[1] == ref_rx[i].raw


def golden_chains(be, ss_vectors, max_frames=10**9):
    """Replays tests/golden/secretstream_vectors.json (libsodium-generated) frame by frame."""
    import hashlib

    from golden_util import frame_message

    n = 0
    for ch in ss_vectors:
        key, hdr = bytes.fromhex(ch["key"]), bytes.fromhex(ch["header"])
        st0 = ls.State(key, hdr).raw
        assert st0.hex() == ch["state0"]
        (slot,) = be.open_streams([st0], [st0])
        for fr in ch["frames"][:max_frames]:
            m = frame_message(fr)
            plain = bu.aligned_zeros(max(16, bu.align16(len(m))))
            plain[: len(m)] = np.frombuffer(m, dtype=np.uint8)
            wire = bu.aligned_zeros(bu.align16(12 + len(m) + 20))
            d = bu.make_descs([slot], [len(m)], [0], [12], tags=[fr["tag"]])
            st = be.run(d, plain, wire)
            w = bytes(wire[12 : 12 + len(m) + 20])
            assert st[0] == fr["tag"] << 8
            assert hashlib.sha256(w).hexdigest() == fr["wire_sha256"], (fr["len"], fr["tag"])
            if "wire" in fr:
                assert w.hex() == fr["wire"]
            assert be.export(slot)[0].hex() == fr["state_after"]
            back = bu.aligned_zeros(max(16, bu.align16(len(m))))
            d2 = bu.make_descs([slot], [len(m)], [12], [0], open_=True)
            st = be.run(d2, wire, back)
            assert st[0] == fr["tag"] << 8 and bytes(back[: len(m)]) == m
            assert be.export(slot)[1].hex() == fr["state_after"]
            n += 1
    return n
