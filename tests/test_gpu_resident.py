"""GPU tests of the resident relay kernel (small pinned host batches: symrelay_batch_submit/_wait/_host).

The reference seals and opens one frame per libsodium call (/root/reference/src/provider.ts:248, :174), so
"frame i+1 of a stream uses the nonce frame i left" holds trivially there.  Here consecutive batches are in
flight together inside one resident grid and the chain is handed over per stream through a stamp in the state;
these tests drive every hand-over there is -- batch to batch, template fast path in and out, resident kernel to
launched kernels and back, idle exit and relaunch, slot reuse, forged frames in the middle of a pipeline -- and
compare every byte with libsodium.  SYMRELAY_PERSIST=0 runs the same scenarios through the launch-per-batch
path (the round-1 path, kept for A/B)."""
import random
import time

import numpy as np
import pytest

import batch_util as bu
from oracle import libsodium as ls

pytestmark = pytest.mark.gpu


@pytest.fixture(params=["resident", "launch_per_batch"])
def relay(request, monkeypatch):
    import symmetry_b200 as sb

    monkeypatch.setenv("SYMRELAY_PERSIST", "1" if request.param == "resident" else "0")
    monkeypatch.setenv("SYMRELAY_PERSIST_IDLE_US", "150")
    r = sb.SymRelay(gpu_ids=(0,), max_streams=8192, pipeline_depth=8)
    r.mode = request.param
    yield r
    r.close()


def _streams(relay, rng, n, both=True):
    keys = [rng.randbytes(32) for _ in range(n)]
    hdrs = [rng.randbytes(24) for _ in range(n)]
    karr = np.frombuffer(b"".join(keys), dtype=np.uint8).reshape(n, 32)
    harr = np.frombuffer(b"".join(hdrs), dtype=np.uint8).reshape(n, 24)
    slots = relay.open_streams(karr, harr, karr if both else None, harr if both else None)
    return keys, hdrs, [int(s) for s in slots]


class Pipe:
    """Pipelined seal batches over a ring of pinned arenas, checked against libsodium in submission order."""

    def __init__(self, relay, keys, hdrs, slots, ring=16):
        self.r, self.slots, self.ring = relay, slots, ring
        self.refs = [ls.State(k, h) for k, h in zip(keys, hdrs)]
        self.pending = []          # (ticket, idx, lens, plain copy, wire arena, status, w_offs)
        self.arenas = {}
        self.k = 0

    def _arena(self, nbytes, which):
        key = (self.k % self.ring, which)
        a = self.arenas.get(key)
        if a is None or a.nbytes < nbytes:
            a = self.r.alloc_pinned(max(nbytes, 1 << 16))
            self.arenas[key] = a
        return a

    def submit(self, idx, lens, rng, sync=False):
        p_offs, p_bytes = bu.plain_layout(lens)
        w_offs, w_bytes = bu.wire_layout(lens)
        if len(self.pending) >= self.ring - 1:
            self.reap(len(self.pending) - self.ring + 2)
        plain, wire = self._arena(p_bytes, "p"), self._arena(w_bytes, "w")
        plain[:p_bytes] = np.frombuffer(rng.randbytes(p_bytes), dtype=np.uint8)
        stat = np.full(len(idx), 0xFF, dtype=np.uint32)
        descs = bu.make_descs([self.slots[i] for i in idx], lens, p_offs, w_offs)
        keep = bytes(plain[:p_bytes])
        if sync:
            self.r.batch_host(descs, plain[:p_bytes], wire[:w_bytes], stat)
            t = None
        else:
            t = self.r.submit(descs, plain, p_bytes, wire, w_bytes, stat)
        self.pending.append((t, idx, lens, keep, wire, stat, w_offs, p_offs))
        self.k += 1

    def reap(self, count=None):
        count = len(self.pending) if count is None else count
        for _ in range(count):
            t, idx, lens, keep, wire, stat, w_offs, p_offs = self.pending.pop(0)
            if t is not None:
                self.r.wait(t)
            assert (stat == 0).all(), stat
            for j, i in enumerate(idx):
                want = self.refs[i].seal_frame(keep[p_offs[j] : p_offs[j] + lens[j]])
                assert bytes(wire[w_offs[j] : w_offs[j] + lens[j] + 20]) == want, (self.k, i)


def test_repeating_and_changing_batches_keep_every_chain(relay):
    """The template fast path (byte-identical descriptors batch after batch) entered, left and re-entered: the
    stamps the fast path skips must be right again the moment a different batch names the same streams."""
    rng = random.Random(100)
    n = 300
    keys, hdrs, slots = _streams(relay, rng, n, both=False)
    pipe = Pipe(relay, keys, hdrs, slots)
    all_idx = list(range(n))
    same = [64] * n
    for _ in range(9):                        # template installs, then the fast path
        pipe.submit(all_idx, same, rng)
    sub = rng.sample(all_idx, 120)            # a different batch over a subset: general path, waits per stream
    pipe.submit(sub, [rng.choice([0, 1, 64, 100, 700]) for _ in sub], rng)
    for _ in range(5):                        # back to the template
        pipe.submit(all_idx, same, rng)
    pipe.submit(list(reversed(all_idx)), same, rng)      # same streams, other order: another template
    for _ in range(4):
        pipe.submit(list(reversed(all_idx)), same, rng)
    pipe.reap()
    for i in (0, 7, n - 1):
        assert relay.export(slots[i])[0] == pipe.refs[i].raw


def test_resident_kernel_and_launched_kernels_alternate(relay):
    """Small packed batches (resident kernel) between batches that must be launched: bulk frames, pageable
    arenas (staged DMA path), synchronous calls.  Stream order on the library's stream is the only thing that
    orders them; every frame must still see its predecessor's nonce."""
    rng = random.Random(101)
    n = 64
    keys, hdrs, slots = _streams(relay, rng, n, both=False)
    pipe = Pipe(relay, keys, hdrs, slots)
    idx = list(range(n))
    small = lambda: [rng.choice([0, 5, 64, 64, 200]) for _ in idx]
    bulk = lambda: [rng.choice([64, 2500, 5000]) for _ in idx]
    for kind in ["s", "s", "b", "s", "s", "s", "b", "b", "s", "sync", "s", "s", "sync", "b", "s"]:
        if kind == "s":
            pipe.submit(idx, small(), rng)
        elif kind == "b":
            pipe.submit(idx, bulk(), rng)
        else:
            pipe.submit(idx, small(), rng, sync=True)
    pipe.reap()
    # pageable arenas take the staged path
    lens = small()
    p_offs, p_bytes = bu.plain_layout(lens)
    w_offs, w_bytes = bu.wire_layout(lens)
    plain, wire = bu.aligned_zeros(p_bytes), bu.aligned_zeros(w_bytes)
    plain[:] = np.frombuffer(rng.randbytes(p_bytes), dtype=np.uint8)
    st = relay.batch_host(bu.make_descs(slots, lens, p_offs, w_offs), plain, wire)
    assert (st == 0).all()
    for i in idx:
        assert bytes(wire[w_offs[i] : w_offs[i] + lens[i] + 20]) == pipe.refs[i].seal_frame(bytes(plain[p_offs[i] : p_offs[i] + lens[i]]))
    pipe.submit(idx, small(), rng)
    pipe.reap()


def test_idle_exit_and_relaunch(relay):
    """The resident kernel leaves after 150 us without work (so that cudaDeviceSynchronize and cudaFree elsewhere in
    the process return) and is started again by the next batch; nothing is lost in either hand-shake."""
    import torch

    rng = random.Random(102)
    n = 128
    keys, hdrs, slots = _streams(relay, rng, n, both=False)
    pipe = Pipe(relay, keys, hdrs, slots)
    idx = list(range(n))
    for burst in range(6):
        for _ in range(rng.choice([1, 2, 5])):
            pipe.submit(idx, [64] * n, rng)
        if burst % 2:
            pipe.reap()
            time.sleep(0.002)                 # idle: the kernel exits
        else:
            torch.cuda.synchronize()          # must return although batches were just posted
            pipe.reap()
    pipe.reap()
    launches_before = relay.launches
    pipe.submit(idx, [64] * n, rng)
    pipe.reap()
    if relay.mode == "resident":
        assert relay.launches > launches_before   # a relaunch (memset is not a kernel; the resident kernel is)


def test_more_batches_in_flight_than_the_header_ring(relay):
    """Host pipelines deeper than the stages recycle stages; tickets stay valid and in order."""
    rng = random.Random(103)
    n = 40
    keys, hdrs, slots = _streams(relay, rng, n, both=False)
    pipe = Pipe(relay, keys, hdrs, slots, ring=48)
    idx = list(range(n))
    for k in range(200):
        pipe.submit(idx, [rng.choice([64, 65]) if k % 7 == 0 else 64 for _ in idx], rng)
    pipe.reap()


def test_forged_frames_inside_a_pipeline(relay):
    """Open batches in flight together; some frames are forged.  A forged frame fails alone, leaves its stream's
    state untouched -- and must still hand the stream on, or its successor in the next batch would wait for ever;
    resubmitting the true frame later continues the chain."""
    rng = random.Random(104)
    n = 100
    keys, hdrs, slots = _streams(relay, rng, n, both=True)
    senders = [ls.State(k, h) for k, h in zip(keys, hdrs)]
    mirrors = [ls.State(k, h) for k, h in zip(keys, hdrs)]
    depth = 6
    lens = [64] * n
    w_offs, w_bytes = bu.wire_layout(lens)
    p_offs, p_bytes = bu.plain_layout(lens)
    descs = bu.make_descs(slots, lens, w_offs, p_offs, open_=True)
    true_frames = [[senders[i].seal_frame(rng.randbytes(64)) for i in range(n)] for _ in range(depth)]
    # stream i's frame of batch 2 is forged for i % 10 == 3: batches 3.. of those streams then fail too (their
    # nonce chain is behind), exactly as sequential pulls would
    arenas, outs, stats, tickets = [], [], [], []
    for k in range(depth):
        a, o = relay.alloc_pinned(w_bytes), relay.alloc_pinned(p_bytes)
        a[:] = 0
        o[:] = 0xCC
        for i in range(n):
            f = bytearray(true_frames[k][i])
            if k == 2 and i % 10 == 3:
                f[9] ^= 0x40
            a[w_offs[i] : w_offs[i] + len(f)] = np.frombuffer(bytes(f), dtype=np.uint8)
        st = np.full(n, 0xFF, dtype=np.uint32)
        arenas.append(a); outs.append(o); stats.append(st)
        tickets.append(relay.submit(descs, a, w_bytes, o, p_bytes, st))
    for t in tickets:
        relay.wait(t)
    for k in range(depth):
        for i in range(n):
            dead = i % 10 == 3 and k >= 2
            if dead:
                assert int(stats[k][i]) & 0xFF == 1, (k, i)
            else:
                m, tag = mirrors[i].open_frame(true_frames[k][i])
                assert int(stats[k][i]) == 0 and bytes(outs[k][p_offs[i] : p_offs[i] + 64]) == m, (k, i)
    # the stream is where the forged frame found it: the true frames, resubmitted in order, open
    idx = [i for i in range(n) if i % 10 == 3]
    for k in range(2, depth):
        sub_lens = [64] * len(idx)
        wo, wb = bu.wire_layout(sub_lens)
        po, pb = bu.plain_layout(sub_lens)
        a, o = relay.alloc_pinned(wb), relay.alloc_pinned(pb)
        for j, i in enumerate(idx):
            a[wo[j] : wo[j] + 84] = np.frombuffer(true_frames[k][i], dtype=np.uint8)
        st = relay.batch_host(bu.make_descs([slots[i] for i in idx], sub_lens, wo, po, open_=True), a, o)
        for j, i in enumerate(idx):
            m, tag = mirrors[i].open_frame(true_frames[k][i])
            assert int(st[j]) == 0 and bytes(o[po[j] : po[j] + 64]) == m
    for i in (3, 13, 4):
        assert relay.export(slots[i])[1] == mirrors[i].raw


def test_slot_reuse_between_batches(relay):
    """A connection closes, another takes its slot (LIFO), and the batch shape does not change: the new stream must
    not wait for the old one's stamp, and must not be sealed with its key."""
    rng = random.Random(105)
    n = 50
    keys, hdrs, slots = _streams(relay, rng, n, both=False)
    pipe = Pipe(relay, keys, hdrs, slots)
    idx = list(range(n))
    for _ in range(5):
        pipe.submit(idx, [64] * n, rng)
    pipe.reap()
    victim = 17
    relay.close_stream(slots[victim])
    k2, h2 = rng.randbytes(32), rng.randbytes(24)
    new_slot = relay.open_stream(k2, h2, None, None)
    assert new_slot == slots[victim]
    pipe.refs[victim] = ls.State(k2, h2)
    for _ in range(5):                         # byte-identical descriptors as before the close
        pipe.submit(idx, [64] * n, rng)
    pipe.reap()
    assert relay.export(new_slot)[0] == pipe.refs[victim].raw


def test_refused_batch_between_resident_batches(relay):
    """A batch that is refused after the slot table was stamped (two frames of one stream) must not leave a stamp
    behind that a later frame waits for."""
    import symmetry_b200 as sb

    rng = random.Random(106)
    n = 30
    keys, hdrs, slots = _streams(relay, rng, n, both=False)
    pipe = Pipe(relay, keys, hdrs, slots)
    idx = list(range(n))
    for _ in range(3):
        pipe.submit(idx, [64] * n, rng)
    lens = [64] * (n + 1)
    p_offs, p_bytes = bu.plain_layout(lens)
    w_offs, w_bytes = bu.wire_layout(lens)
    a, o = relay.alloc_pinned(p_bytes), relay.alloc_pinned(w_bytes)
    with pytest.raises(sb.RelayError, match="two frames of one"):
        relay.submit(bu.make_descs(slots + [slots[0]], lens, p_offs, w_offs), a, p_bytes, o, w_bytes,
                     np.zeros(n + 1, dtype=np.uint32))
    for _ in range(3):
        pipe.submit(idx, [rng.choice([64, 100]) for _ in idx], rng)
    pipe.reap()
