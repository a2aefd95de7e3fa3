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


def seal_open_roundtrip(be, lens, tags, seed, chain=2, corrupt=(), speculative_min_len=1921):
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
                    region = out[o : o + max(16, bu.align16(lens[i]))]
                    if lens[i] < speculative_min_len:   # packed tiles hold the plaintext back: the output is untouched
                        assert (region == 0x55).all(), "plaintext released on bad MAC"
                    else:                 # bulk and strided tiles decrypt while they MAC and wipe on a bad tag
                        assert (region[: lens[i]] == 0).all() and (region[lens[i] :] == 0x55).all()
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


def run_open_scenario(be, seed=0, n_streams=24):
    """Runs: several frames queued on one stream are opened in ONE batch (SYMRELAY_F_RUN).  Expected
    behaviour = that many sequential libsodium pulls: a forged frame kills the stream there (frames
    behind it: RETRY, state advanced only over the verified prefix), a REKEY tag or counter wrap is
    honoured (frames behind it: RETRY, and they open fine when resubmitted)."""
    rng = random.Random(seed)
    keys = [rng.randbytes(32) for _ in range(n_streams)]
    hdrs = [rng.randbytes(24) for _ in range(n_streams)]
    senders = [ls.State(k, h) for k, h in zip(keys, hdrs)]      # the remote peers
    mirrors = [ls.State(k, h) for k, h in zip(keys, hdrs)]      # what sequential pulls would do
    slots = be.open_streams([s.raw for s in senders], [s.raw for s in mirrors])
    plan = []   # (stream, [(msg, tag, forged)])
    for i in range(n_streams):
        k = rng.choice([1, 2, 3, 5, 9, 20])
        frames = []
        for j in range(k):
            L = rng.choice([0, 1, 17, 64, 200, 1000, 1920, 1921, 5000])
            frames.append([rng.randbytes(L), 0, False])
        if i % 6 == 1 and k > 2:
            frames[k // 2][2] = True            # forged in the middle
        if i % 6 == 2 and k > 2:
            frames[k // 2][1] = 2               # REKEY tag in the middle
        if i % 6 == 3:
            frames[0][1] = 3                    # FINAL (rekeys too: tag & 2)
        plan.append((i, frames))
    # wire bytes
    wires, lens, owner = [], [], []
    for i, frames in plan:
        for m, tag, forged in frames:
            w = bytearray(senders[i].seal_frame(m, tag))
            if forged:
                w[3 + rng.randrange(len(m) + 1)] ^= 0x40     # tag byte or ciphertext: the MAC bytes stay genuine
            wires.append(bytes(w)); lens.append(len(m)); owner.append(i)
    w_offs, w_bytes = bu.wire_layout(lens)
    p_offs, p_bytes = bu.plain_layout(lens)
    arena = bu.aligned_zeros(w_bytes)
    for w, o in zip(wires, w_offs):
        arena[o : o + len(w)] = np.frombuffer(w, dtype=np.uint8)
    out = bu.aligned_zeros(p_bytes)
    d = bu.make_descs([slots[i] for i in owner], lens, w_offs, p_offs, open_=True, run=True)
    status = be.run(d, arena, out)
    idx = 0
    retry = []
    for i, frames in plan:
        alive = True
        for m, tag, forged in frames:
            code = int(status[idx]) & 0xFF
            if not alive:
                assert code == 5, (i, idx, code)            # RETRY
                retry.append(idx)
            elif forged:
                assert code == 1, (i, idx, code)            # BAD_MAC, as pull would report
                alive = False
            else:
                assert int(status[idx]) == (tag << 8), (i, idx, status[idx])
                assert bytes(out[p_offs[idx] : p_offs[idx] + len(m)]) == m
                assert mirrors[i].open_frame(wires[idx]) == (m, tag)
                if tag & 2:
                    alive = False                           # speculation ends at a rekey
            idx += 1
        assert be.export(slots[i])[1] == mirrors[i].raw, f"rx state of stream {i}"
    # resubmit what was held back behind a REKEY: it must now open (streams with a forged frame stay dead)
    dead = {i for i, frames in plan if any(f[2] for f in frames)}
    again = [j for j in retry if owner[j] not in dead]
    if again:
        d2 = bu.make_descs([slots[owner[j]] for j in again], [lens[j] for j in again], [w_offs[j] for j in again],
                           [p_offs[j] for j in again], open_=True, run=True)
        st2 = be.run(d2, arena, out)
        for k, j in enumerate(again):
            i = owner[j]
            got = mirrors[i].open_frame(wires[j])
            assert got is not None and int(st2[k]) == (got[1] << 8)
            assert bytes(out[p_offs[j] : p_offs[j] + lens[j]]) == got[0]
        for i in {owner[j] for j in again}:
            assert be.export(slots[i])[1] == mirrors[i].raw
    return len(wires), len(retry)
