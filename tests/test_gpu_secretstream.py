"""GPU parity tests: the CUDA path, called through the C ABI (ctypes), against
the real libsodium on the same seeded inputs, the libsodium-generated golden
chains, and -- at BASELINE.json's full sizes -- size-independent properties
(seal -> open round trip, state chain equality on a sample, forged frames
rejected).  Bit-exact: bytes, tags, statuses and 52-byte states."""
import random

import numpy as np
import pytest

import batch_util as bu
import parity_cases as pc
from oracle import libsodium as ls

pytestmark = pytest.mark.gpu


class GpuBackend:
    """Host-arena batches through symrelay_batch_host.  The whole out arena is copied back from
    the device staging buffer, so bytes between frames are unspecified (preserves_gaps False);
    GpuDeviceBackend below checks the exact-footprint property on device buffers."""

    preserves_gaps = False

    def __init__(self, relay):
        self.relay = relay

    def open_streams(self, tx_states, rx_states):
        n = len(tx_states)
        keys = np.zeros((n, 32), dtype=np.uint8)
        tx = np.frombuffer(b"".join(tx_states), dtype=np.uint8).reshape(n, 52)
        rx = np.frombuffer(b"".join(rx_states), dtype=np.uint8).reshape(n, 52)
        return [int(s) for s in self.relay.open_streams(keys, tx, keys, rx)]

    def run(self, descs, arena_in, arena_out):
        return self.relay.batch_host(descs, arena_in, arena_out)

    def export(self, slot):
        return self.relay.export(slot)


class GpuDeviceBackend(GpuBackend):
    """Same scenarios through symrelay_batch_device on torch tensors: the kernels must touch
    nothing outside the frames they write."""

    preserves_gaps = True

    def run(self, descs, arena_in, arena_out):
        import torch

        dev = torch.device("cuda:0")
        tiles, n_bulk = self.relay.plan(descs)
        t_desc = torch.from_numpy(descs.view(np.uint8).copy()).to(dev)
        t_tiles = torch.from_numpy(tiles.view(np.uint8).copy()).to(dev)
        t_in = torch.from_numpy(np.ascontiguousarray(arena_in)).to(dev)
        t_out = torch.from_numpy(np.ascontiguousarray(arena_out)).to(dev)
        t_status = torch.full((len(descs),), 0xFF, dtype=torch.int32, device=dev)
        if (descs["len_flags"] & bu.F_RUN).any():
            scratch = torch.empty(16 * len(descs), dtype=torch.uint8, device=dev)   # caller-provided scratch
            self.relay.batch_device_runs(t_desc.data_ptr(), len(descs), t_tiles.data_ptr(), len(tiles), n_bulk,
                                         t_in.data_ptr(), t_out.data_ptr(), t_status.data_ptr(), scratch.data_ptr(),
                                         scratch.numel(), stream=torch.cuda.current_stream().cuda_stream)
        else:
            self.relay.batch_device(t_desc.data_ptr(), len(descs), t_tiles.data_ptr(), len(tiles), n_bulk,
                                    t_in.data_ptr(), t_out.data_ptr(), t_status.data_ptr(),
                                    stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        arena_out[:] = t_out.cpu().numpy()
        return t_status.cpu().numpy().astype(np.uint32)


@pytest.fixture(scope="module")
def relay():
    import symmetry_b200 as sb

    r = sb.SymRelay(gpu_ids=(0,), max_streams=1 << 17, stage_bytes=96 << 20)
    yield r
    r.close()


@pytest.fixture()
def be(relay):
    return GpuBackend(relay)


def test_header_init_matches_libsodium(relay):
    rng = random.Random(3)
    for _ in range(8):
        tk, th, rk, rh = rng.randbytes(32), rng.randbytes(24), rng.randbytes(32), rng.randbytes(24)
        slot = relay.open_stream(tk, th, rk, rh)
        tx, rx = relay.export(slot)
        assert tx == ls.State(tk, th).raw and rx == ls.State(rk, rh).raw
        relay.close_stream(slot)
    n = 300
    tks = np.frombuffer(rng.randbytes(32 * n), dtype=np.uint8).reshape(n, 32)
    ths = np.frombuffer(rng.randbytes(24 * n), dtype=np.uint8).reshape(n, 24)
    slots = relay.open_streams(tks, ths, tks, ths)
    assert len(set(int(s) for s in slots)) == n
    for i in (0, 1, 150, 299):
        want = ls.State(bytes(tks[i]), bytes(ths[i])).raw
        assert relay.export(int(slots[i])) == (want, want)


@pytest.fixture()
def dbe(relay):
    return GpuDeviceBackend(relay)


def test_device_path_touches_only_its_frames(dbe):
    lens = list(range(0, 200)) + [1919, 1920, 1921, 2048, 4097, 65537]
    pc.seal_open_roundtrip(dbe, lens, [i % 4 for i in range(len(lens))], seed=12, chain=2, corrupt=(3, 64, 200, 205))


def test_runs_of_queued_frames_host_path(be):
    for seed in (3, 4):
        n, n_retry = pc.run_open_scenario(be, seed=seed, n_streams=60)
        assert n > 100 and n_retry > 0


def test_runs_of_queued_frames_device_path(dbe):
    n, n_retry = pc.run_open_scenario(dbe, seed=5, n_streams=60)
    assert n > 100 and n_retry > 0


def test_golden_chains(be, ss_vectors):
    assert pc.golden_chains(be, ss_vectors) >= 200


def test_packed_every_small_length(be):
    lens = list(range(0, 260))
    pc.seal_open_roundtrip(be, lens, [0] * len(lens), seed=1, chain=3)


def test_packed_tags_and_rekey(be):
    rng = random.Random(5)
    lens = [rng.choice([0, 1, 15, 16, 17, 63, 64, 65, 100, 200, 500, 1000, 1919, 1920]) for _ in range(400)]
    pc.seal_open_roundtrip(be, lens, [i % 4 for i in range(400)], seed=2, chain=4)


def test_bulk_lengths(be):
    lens = [1921, 1984, 2048, 4095, 4096, 4097, 16383, 16384, 65535, 65536, 65537, 64 * 1022, 64 * 1023 + 1,
            1 << 20, (1 << 20) + 3]
    pc.seal_open_roundtrip(be, lens, [0, 1, 2, 3, 0, 0, 0, 0, 2, 0, 0, 1, 0, 0, 3], seed=3, chain=2)


def test_config3_ragged_power_law_with_forgeries(be):
    lens = pc.power_law_lens(random.Random(11), 3000) + [3000, 0, 0, 64]
    pc.seal_open_roundtrip(be, lens, [0] * len(lens), seed=4, chain=2, corrupt=(0, 5, 17, 1000, 2999, 3000, 3001, 3003))


def test_config2_1024_streams_x64B_chain(be):
    pc.seal_open_roundtrip(be, [64] * 1024, [0] * 1024, seed=6, chain=8)


def test_thread_per_frame_tiles_on_a_large_uniform_batch(be):
    """>= 8,192 frames with runs of equal block counts: the planner switches those to thread-per-frame
    tiles (one lane = one frame).  Same bit-exact bar, incl. tags, rekey and forged frames."""
    lens = [64] * 6000 + [200] * 5000 + [0] * 640 + [256] * 320 + [1500, 5000, 30]
    tiles, _ = be.relay.plan(bu.make_descs(np.arange(len(lens)), lens, bu.plain_layout(lens)[0], bu.wire_layout(lens)[0]))
    assert sum(1 for t in tiles if int(t["n_kind"]) >> 17) == 6000 // 32 + 5000 // 32 + 640 // 32 + 320 // 32
    tags = [(i % 4) if i % 11 == 0 else 0 for i in range(len(lens))]
    pc.seal_open_roundtrip(be, lens, tags, seed=31, chain=2, corrupt=(0, 31, 32, 6001, 11000, 11700, 11960))


@pytest.mark.parametrize("length", [385, 448, 449, 512, 640, 700, 1000, 1024, 1025, 1280, 1500, 1856, 1857, 1919, 1920])
def test_strided_tiles_mid_size_frames(dbe, length):
    """Frames of 7..32 ChaCha20 blocks in a batch that is all one size class: 32/k frames per warp, k lanes each walking
    the frame's blocks with stride k (2, 4 or 8 lanes here), partial MACs combined with r^(4k).  Every lane count, the
    tail tile with fewer frames than fit, tags incl. REKEY, forged frames (decrypted speculatively, wiped on a bad
    tag, state untouched, the true frame opens afterwards)."""
    n = 37
    tags = [0, 1, 2, 3] * 9 + [0]
    pc.seal_open_roundtrip(dbe, [length] * n, tags, seed=1000 + length, chain=2, corrupt=(0, 5, 36), speculative_min_len=385)


@pytest.mark.parametrize("lens", [
    [8000, 8001, 8064, 8065, 8192, 12345, 16192, 16193, 16256, 16257, 65536, 200000],   # 8,001 bytes = 128 blocks: the first wide count; 16,193: a third round
    [8192] * 5 + [100, 700, 3000, 5000] * 3 + [24500, 8127],               # mixed with packed, strided and one-warp tiles
])
def test_wide_tiles_few_big_frames(dbe, lens):
    """A batch with only a few frames of 8 KiB and more: each gets a 128-thread block (four warps, stride 128, partial
    MACs through shared memory, r^512 between a thread's windows).  Block counts around multiples of 128, tags incl.
    REKEY, forged frames wiped, two frames per stream chained over two batches."""
    tags = [(i * 7 + 1) & 3 for i in range(len(lens))]
    pc.seal_open_roundtrip(dbe, lens, tags, seed=4242 + len(lens), chain=2, corrupt=(0, 4, len(lens) - 1), speculative_min_len=385)


def test_counter_wrap_rekeys(be):
    key, hdr = bytes(range(32)), bytes(24)
    raw = bytearray(ls.State(key, hdr).raw)
    raw[32:36] = (0xFFFFFFFF).to_bytes(4, "little")
    ref = ls.State(key, hdr)
    ref.st.statebuf[0:52] = bytes(raw)
    (slot,) = be.open_streams([bytes(raw)], [bytes(raw)])
    plain = bu.aligned_zeros(16)
    plain[:4] = np.frombuffer(b"wrap", dtype=np.uint8)
    wire = bu.aligned_zeros(48)
    be.run(bu.make_descs([slot], [4], [0], [12]), plain, wire)
    assert bytes(wire[12:36]) == ref.seal_frame(b"wrap")
    assert be.export(slot)[0] == ref.raw and ref.raw[32:36] == (1).to_bytes(4, "little")


def _full_size_roundtrip(relay, lens, seed, sample):
    """Seal then open n frames through host arenas; check every plaintext comes back, and a
    sample of frames/states against libsodium."""
    rng = np.random.default_rng(seed)
    n = len(lens)
    keys = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    hdrs = rng.integers(0, 256, (n, 24), dtype=np.uint8)
    slots = relay.open_streams(keys, hdrs, keys, hdrs)
    p_offs, p_bytes = bu.plain_layout(lens)
    w_offs, w_bytes = bu.wire_layout(lens)
    plain = relay.alloc_pinned(p_bytes)
    plain[:] = rng.integers(0, 256, p_bytes, dtype=np.uint8)
    wire = relay.alloc_pinned(w_bytes)
    back = relay.alloc_pinned(p_bytes)
    st = relay.batch_host(bu.make_descs(slots, lens, p_offs, w_offs), plain, wire)
    assert (st == 0).all()
    for i in sample:
        ref = ls.State(bytes(keys[i]), bytes(hdrs[i]))
        want = ref.seal_frame(bytes(plain[p_offs[i] : p_offs[i] + lens[i]]))
        assert bytes(wire[w_offs[i] : w_offs[i] + lens[i] + 20]) == want, (i, lens[i])
        assert relay.export(int(slots[i]))[0] == ref.raw
    st = relay.batch_host(bu.make_descs(slots, lens, w_offs, p_offs, open_=True), wire, back)
    assert (st == 0).all()
    ok = True
    for o, L in zip(p_offs, lens):
        ok &= bool((back[o : o + L] == plain[o : o + L]).all())
    assert ok
    for i in sample:
        assert relay.export(int(slots[i]))[0] == relay.export(int(slots[i]))[1]
    return slots


def test_config3_full_size_65536_streams(relay):
    lens = pc.power_law_lens(random.Random(0x5EED0001), 65536)
    _full_size_roundtrip(relay, lens, seed=7, sample=[0, 1, 2, 100, 4097, 30000, 65535] +
                         [int(i) for i in np.argsort(lens)[-4:]])


def test_config4_bulk_64KiB_frames(relay):
    _full_size_roundtrip(relay, [65536] * 512, seed=8, sample=[0, 255, 511])


def test_device_resident_batch_on_torch_stream(relay):
    """symrelay_batch_device: buffers in HBM (torch tensors), launch on torch's current stream."""
    import torch

    rng = random.Random(9)
    lens = [64] * 1024 + pc.power_law_lens(rng, 500) + [65536, 5000]
    n = len(lens)
    keys = [rng.randbytes(32) for _ in range(n)]
    hdrs = [rng.randbytes(24) for _ in range(n)]
    karr = np.frombuffer(b"".join(keys), dtype=np.uint8).reshape(n, 32)
    harr = np.frombuffer(b"".join(hdrs), dtype=np.uint8).reshape(n, 24)
    slots = relay.open_streams(karr, harr, karr, harr)
    p_offs, p_bytes = bu.plain_layout(lens)
    w_offs, w_bytes = bu.wire_layout(lens)
    plain = np.frombuffer(rng.randbytes(p_bytes), dtype=np.uint8).copy()
    descs = bu.make_descs(slots, lens, p_offs, w_offs)
    tiles, n_bulk = relay.plan(descs)
    dev = torch.device("cuda:0")
    t_desc = torch.from_numpy(descs.view(np.uint8).copy()).to(dev)
    t_tiles = torch.from_numpy(tiles.view(np.uint8).copy()).to(dev)
    t_in = torch.from_numpy(plain).to(dev)
    t_out = torch.zeros(w_bytes, dtype=torch.uint8, device=dev)
    t_status = torch.full((n,), 0xFF, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    relay.batch_device(t_desc.data_ptr(), n, t_tiles.data_ptr(), len(tiles), n_bulk, t_in.data_ptr(),
                       t_out.data_ptr(), t_status.data_ptr(), stream=stream)
    torch.cuda.synchronize()
    wire = t_out.cpu().numpy()
    assert (t_status.cpu().numpy() == 0).all()
    for i in list(range(0, n, 97)) + [n - 2, n - 1]:
        want = ls.State(keys[i], hdrs[i]).seal_frame(bytes(plain[p_offs[i] : p_offs[i] + lens[i]]))
        assert bytes(wire[w_offs[i] : w_offs[i] + lens[i] + 20]) == want
    # open in place on the device: wire arena -> plaintext written into a second arena
    d2 = bu.make_descs(slots, lens, w_offs, p_offs, open_=True)
    tiles2, n_bulk2 = relay.plan(d2)
    t_desc2 = torch.from_numpy(d2.view(np.uint8).copy()).to(dev)
    t_tiles2 = torch.from_numpy(tiles2.view(np.uint8).copy()).to(dev)
    t_back = torch.zeros(p_bytes, dtype=torch.uint8, device=dev)
    relay.batch_device(t_desc2.data_ptr(), n, t_tiles2.data_ptr(), len(tiles2), n_bulk2, t_out.data_ptr(),
                       t_back.data_ptr(), t_status.data_ptr(), stream=stream)
    torch.cuda.synchronize()
    back = t_back.cpu().numpy()
    assert (t_status.cpu().numpy() == 0).all()
    for o, L in zip(p_offs, lens):
        assert (back[o : o + L] == plain[o : o + L]).all()


def test_sorting_planner_on_the_device_path(relay):
    """symrelay_plan_sorted: descriptors grouped by size class for the device entry points; wire bytes are
    libsodium's and the status of sorted descriptor k belongs to the caller's frame perm[k] (forged frames)."""
    import torch

    rng = random.Random(19)
    lens = pc.power_law_lens(rng, 9000) + [64] * 300 + [40000, 70000]
    rng.shuffle(lens)
    n = len(lens)
    karr = np.frombuffer(rng.randbytes(32 * n), dtype=np.uint8).reshape(n, 32)
    harr = np.frombuffer(rng.randbytes(24 * n), dtype=np.uint8).reshape(n, 24)
    slots = relay.open_streams(karr, harr, karr, harr)
    p_offs, p_bytes = bu.plain_layout(lens)
    w_offs, w_bytes = bu.wire_layout(lens)
    plain = np.frombuffer(rng.randbytes(p_bytes), dtype=np.uint8).copy()
    dev = torch.device("cuda:0")
    stream = torch.cuda.current_stream().cuda_stream

    def run(descs, t_in, out_bytes):
        tiles, n_bulk, sorted_descs, perm = relay.plan_sorted(descs)
        assert sorted(perm.tolist()) == list(range(n)) and (sorted_descs == descs[perm]).all()
        assert len(tiles) < len(relay.plan(descs)[0])            # it packs better than the caller's order
        t_desc = torch.from_numpy(sorted_descs.view(np.uint8).copy()).to(dev)
        t_tiles = torch.from_numpy(tiles.view(np.uint8).copy()).to(dev)
        t_out = torch.zeros(out_bytes, dtype=torch.uint8, device=dev)
        t_status = torch.full((n,), 0xFF, dtype=torch.int32, device=dev)
        relay.batch_device(t_desc.data_ptr(), n, t_tiles.data_ptr(), len(tiles), n_bulk, t_in.data_ptr(), t_out.data_ptr(),
                           t_status.data_ptr(), stream=stream)
        torch.cuda.synchronize()
        status = np.empty(n, dtype=np.int64)
        status[perm] = t_status.cpu().numpy() & 0xFF          # back to the caller's order
        return t_out, status

    t_wire, st = run(bu.make_descs(slots, lens, p_offs, w_offs), torch.from_numpy(plain).to(dev), w_bytes)
    assert (st == 0).all()
    wire = t_wire.cpu().numpy()
    for i in list(range(0, n, 131)) + [lens.index(70000), lens.index(40000)]:
        want = ls.State(bytes(karr[i]), bytes(harr[i])).seal_frame(bytes(plain[p_offs[i] : p_offs[i] + lens[i]]))
        assert bytes(wire[w_offs[i] : w_offs[i] + lens[i] + 20]) == want, i
    forged = sorted(rng.sample(range(n), 25))
    for i in forged:
        wire[w_offs[i] + 3] ^= 1
    t_back, st = run(bu.make_descs(slots, lens, w_offs, p_offs, open_=True), torch.from_numpy(wire).to(dev), p_bytes)
    assert [i for i in range(n) if st[i] != 0] == forged
    back = t_back.cpu().numpy()
    for i in range(0, n, 53):
        if i not in forged:
            assert (back[p_offs[i] : p_offs[i] + lens[i]] == plain[p_offs[i] : p_offs[i] + lens[i]]).all()
    for s_ in slots:
        relay.close_stream(int(s_))


def test_pipelined_submit_keeps_chain_order(relay):
    """Consecutive batches carry consecutive frames of the same streams (MAC chain)."""
    rng = random.Random(10)
    n, depth = 256, 12
    keys = [rng.randbytes(32) for _ in range(n)]
    hdrs = [rng.randbytes(24) for _ in range(n)]
    karr = np.frombuffer(b"".join(keys), dtype=np.uint8).reshape(n, 32)
    harr = np.frombuffer(b"".join(hdrs), dtype=np.uint8).reshape(n, 24)
    slots = relay.open_streams(karr, harr, None, None)
    lens = [rng.choice([0, 17, 64, 200, 3000]) for _ in range(n)]
    p_offs, p_bytes = bu.plain_layout(lens)
    w_offs, w_bytes = bu.wire_layout(lens)
    descs = bu.make_descs(slots, lens, p_offs, w_offs)
    plains = [relay.alloc_pinned(p_bytes) for _ in range(depth)]
    wires = [relay.alloc_pinned(w_bytes) for _ in range(depth)]
    stats = [np.full(n, 0xFF, dtype=np.uint32) for _ in range(depth)]
    tickets = []
    for k in range(depth):
        plains[k][:] = np.frombuffer(rng.randbytes(p_bytes), dtype=np.uint8)
        tickets.append(relay.submit(descs, plains[k], p_bytes, wires[k], w_bytes, stats[k]))
    for t in tickets:
        relay.wait(t)
    refs = [ls.State(k, h) for k, h in zip(keys, hdrs)]
    for k in range(depth):
        assert (stats[k] == 0).all()
        for i in range(0, n, 5):
            want = refs[i].seal_frame(bytes(plains[k][p_offs[i] : p_offs[i] + lens[i]]))
            assert bytes(wires[k][w_offs[i] : w_offs[i] + lens[i] + 20]) == want, (k, i)


def test_deferred_write_back_in_mixed_pipelines(relay):
    """A pipelined small batch leaves its results on the device until the next kernel on the stream (or the
    wait) posts them to the host.  Here every hand-over is exercised: small after small (carried), small
    before a batch with bulk tiles (carried by its packed kernel), before a bulk-only batch (flushed), before
    a batch on pageable memory (flushed, staged DMA path), waits in and out of submission order, and a wait
    with nothing behind it.  Every wire byte must be libsodium's, in chain order."""
    rng = random.Random(44)
    n = 96
    keys = [rng.randbytes(32) for _ in range(n)]
    hdrs = [rng.randbytes(24) for _ in range(n)]
    karr = np.frombuffer(b"".join(keys), dtype=np.uint8).reshape(n, 32)
    harr = np.frombuffer(b"".join(hdrs), dtype=np.uint8).reshape(n, 24)
    slots = relay.open_streams(karr, harr, None, None)
    refs = [ls.State(k, h) for k, h in zip(keys, hdrs)]
    kinds = {
        "small": lambda: [rng.choice([0, 1, 63, 64, 65, 200, 700]) for _ in range(n)],
        "mixed": lambda: [rng.choice([64, 200, 2500, 5000]) for _ in range(n)],
        "bulk": lambda: [rng.choice([2000, 3000, 4096]) for _ in range(n)],
    }
    plan = [("small", True, "later"), ("small", True, "later"), ("mixed", True, "later"), ("small", True, "now"),
            ("small", True, "later"), ("bulk", True, "later"), ("small", True, "later"), ("small", False, "later"),
            ("small", True, "later"), ("small", True, "later"), ("small", True, "reverse"), ("small", True, "now")]
    inflight = []   # (ticket, lens, plain, wire, stat, w_offs, p_offs)

    def check(item):
        _, lens, plain, wire, stat, w_offs, p_offs = item
        assert (stat == 0).all()
        for i in range(n):
            want = refs[i].seal_frame(bytes(plain[p_offs[i] : p_offs[i] + lens[i]]))
            assert bytes(wire[w_offs[i] : w_offs[i] + lens[i] + 20]) == want

    done = []
    for kind, pinned, when in plan:
        lens = kinds[kind]()
        p_offs, p_bytes = bu.plain_layout(lens)
        w_offs, w_bytes = bu.wire_layout(lens)
        if pinned:
            plain, wire = relay.alloc_pinned(p_bytes), relay.alloc_pinned(w_bytes)
        else:
            plain, wire = bu.aligned_zeros(p_bytes), bu.aligned_zeros(w_bytes)
        plain[:] = np.frombuffer(rng.randbytes(p_bytes), dtype=np.uint8)
        wire[:] = 0xEE
        stat = np.full(n, 0xFF, dtype=np.uint32)
        t = relay.submit(bu.make_descs(slots, lens, p_offs, w_offs), plain, p_bytes, wire, w_bytes, stat)
        inflight.append((t, lens, plain, wire, stat, w_offs, p_offs))
        if when == "now":          # everything submitted so far, oldest first
            for item in inflight:
                relay.wait(item[0])
            done += inflight
            inflight = []
        elif when == "reverse":    # newest first: the older tickets are then already complete
            for item in reversed(inflight):
                relay.wait(item[0])
            done += inflight
            inflight = []
    assert not inflight
    for item in done:              # reference chains advance in submission order
        check(item)


def test_bucketed_host_batches_hand_statuses_back_in_caller_order(relay):
    """Large staged batches are sorted by size inside the library; statuses (and the bytes) must still belong
    to the caller's descriptor i: frames in arrival order, forged frames at known indices on the open side."""
    rng = random.Random(77)
    n = 12000
    keys = np.frombuffer(rng.randbytes(32 * n), dtype=np.uint8).reshape(n, 32)
    hdrs = np.frombuffer(rng.randbytes(24 * n), dtype=np.uint8).reshape(n, 24)
    slots = relay.open_streams(keys, hdrs, keys, hdrs)
    lens = [rng.choice([0, 5, 64, 64, 64, 130, 200, 700, 1900, 2500, 9000]) for _ in range(n)]
    p_offs, p_bytes = bu.plain_layout(lens)
    w_offs, w_bytes = bu.wire_layout(lens)
    assert p_bytes + w_bytes > (1 << 20)
    plain, wire, back = relay.alloc_pinned(p_bytes), relay.alloc_pinned(w_bytes), relay.alloc_pinned(p_bytes)
    plain[:] = np.frombuffer(rng.randbytes(p_bytes), dtype=np.uint8)
    st = relay.batch_host(bu.make_descs(slots, lens, p_offs, w_offs), plain, wire)
    assert (st == 0).all()
    for i in range(0, n, 211):
        ref = ls.State(bytes(keys[i]), bytes(hdrs[i]))
        want = ref.seal_frame(bytes(plain[p_offs[i] : p_offs[i] + lens[i]]))
        assert bytes(wire[w_offs[i] : w_offs[i] + lens[i] + 20]) == want, i
    forged = sorted(rng.sample(range(n), 40))
    for i in forged:
        wire[w_offs[i] + 3 + (lens[i] + 17) // 2] ^= 0x20     # somewhere in tag | ciphertext | mac
    st = relay.batch_host(bu.make_descs(slots, lens, w_offs, p_offs, open_=True), wire, back)
    bad = [i for i in range(n) if (int(st[i]) & 0xFF) != 0]
    assert bad == forged
    for i in range(0, n, 97):
        if i not in forged:
            assert (back[p_offs[i] : p_offs[i] + lens[i]] == plain[p_offs[i] : p_offs[i] + lens[i]]).all(), i
    for s_ in slots:
        relay.close_stream(int(s_))


def test_zero_copy_and_staged_paths_agree():
    """The same pinned-arena batches through both host paths give identical wire bytes and states."""
    import symmetry_b200 as sb

    rng = random.Random(21)
    lens = [64] * 300 + [rng.choice([0, 1, 17, 200, 1000, 1920, 1921, 5000]) for _ in range(200)]
    n = len(lens)
    keys = np.frombuffer(rng.randbytes(32 * n), dtype=np.uint8).reshape(n, 32)
    hdrs = np.frombuffer(rng.randbytes(24 * n), dtype=np.uint8).reshape(n, 24)
    p_offs, p_bytes = bu.plain_layout(lens)
    w_offs, w_bytes = bu.wire_layout(lens)
    assert p_bytes + w_bytes < (1 << 20)
    payload = [np.frombuffer(rng.randbytes(p_bytes), dtype=np.uint8) for _ in range(3)]
    results = []
    for zc in (True, False):
        with sb.SymRelay(gpu_ids=(0,), max_streams=4096, zero_copy=zc) as r:
            slots = r.open_streams(keys, hdrs, keys, hdrs)
            plain, wire, back = r.alloc_pinned(p_bytes), r.alloc_pinned(w_bytes), r.alloc_pinned(p_bytes)
            wires = []
            for k in range(3):
                plain[:] = payload[k]
                st = r.batch_host(bu.make_descs(slots, lens, p_offs, w_offs), plain, wire)
                assert (st == 0).all()
                wires.append(bytes(b"".join(bytes(wire[o : o + L + 20]) for o, L in zip(w_offs, lens))))
                st = r.batch_host(bu.make_descs(slots, lens, w_offs, p_offs, open_=True), wire, back)
                assert (st == 0).all()
                for o, L in zip(p_offs, lens):
                    assert (back[o : o + L] == plain[o : o + L]).all()
            results.append((wires, [r.export(int(s)) for s in slots[::37]]))
    assert results[0] == results[1]
    ref = ls.State(bytes(keys[0]), bytes(hdrs[0]))
    for k in range(3):
        w = ref.seal_frame(bytes(payload[k][p_offs[0] : p_offs[0] + lens[0]]))
        assert results[0][0][k][: len(w)] == w


def test_bad_arguments_are_rejected(relay):
    import symmetry_b200 as sb

    d = bu.make_descs([0], [64], [0], [12])
    small = bu.aligned_zeros(64)
    with pytest.raises(sb.RelayError):
        relay.batch_host(d, small, small)  # wire frame does not fit the out arena
    d = bu.make_descs([0], [64], [4], [12])
    with pytest.raises(sb.RelayError):
        relay.batch_host(d, bu.aligned_zeros(128), bu.aligned_zeros(128))  # misaligned plaintext
    with pytest.raises(sb.RelayError):
        relay.export((1 << 17) - 1)  # not open (the last slot of this relay; the tests above never get that far)
    k, h = bytes(32), bytes(24)
    slot = relay.open_stream(k, h, k, h)
    relay.close_stream(slot)
    d = bu.make_descs([slot], [64], [0], [12])
    with pytest.raises(sb.RelayError, match="not open"):
        relay.batch_host(d, bu.aligned_zeros(128), bu.aligned_zeros(128))  # a closed slot has no key
