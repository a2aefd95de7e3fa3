"""GPU tests for caller mistakes and connection churn: things the reference cannot get wrong because it
handles one frame of one connection per call (/root/reference/src/provider.ts:248, :174), and which a
batched replacement must refuse or survive instead of silently corrupting a stream's MAC chain."""
import random

import numpy as np
import pytest

import batch_util as bu
from oracle import libsodium as ls

pytestmark = pytest.mark.gpu


@pytest.fixture()
def relay():
    import symmetry_b200 as sb

    r = sb.SymRelay(gpu_ids=(0,), max_streams=2048, flush_us=0)
    yield r
    r.close()


def _open_arena(wires, lens):
    w_offs, w_bytes = bu.wire_layout(lens)
    p_offs, p_bytes = bu.plain_layout(lens)
    arena = bu.aligned_zeros(w_bytes)
    for w, o in zip(wires, w_offs):
        arena[o : o + len(w)] = np.frombuffer(w, dtype=np.uint8)
    return arena, w_offs, p_offs, p_bytes


def test_two_frames_of_one_stream_in_a_host_batch_are_refused(relay):
    import symmetry_b200 as sb

    k, h = bytes(range(32)), bytes(range(24))
    a = relay.open_stream(k, h, k, h)
    b = relay.open_stream(k, h, k, h)
    lens = [64, 64, 64]
    p_offs, p_bytes = bu.plain_layout(lens)
    w_offs, w_bytes = bu.wire_layout(lens)
    plain, wire = bu.aligned_zeros(p_bytes), bu.aligned_zeros(w_bytes)
    before = relay.export(a)[0]
    with pytest.raises(sb.RelayError, match="two frames of one"):
        relay.batch_host(bu.make_descs([a, b, a], lens, p_offs, w_offs), plain, wire)
    assert relay.export(a)[0] == before, "a refused batch must not advance any chain"
    # one seal and one open of the same slot are different directions: allowed
    ref = ls.State(k, h)
    w = ref.seal_frame(b"x" * 64)
    arena, wo, po, pb = _open_arena([w], [64])
    st = relay.batch_host(bu.make_descs([b], [64], wo, po, open_=True), arena, bu.aligned_zeros(pb))
    assert (st == 0).all()


def test_runs_longer_than_64_are_refused_on_host_batches_and_retried_on_device_batches(relay):
    import torch

    import symmetry_b200 as sb

    k, h = random.Random(3).randbytes(32), random.Random(4).randbytes(24)
    sender = ls.State(k, h)
    n = 70
    msgs = [bytes([i]) * (i % 5 * 16) for i in range(n)]
    wires = [sender.seal_frame(m) for m in msgs]
    lens = [len(m) for m in msgs]
    arena, w_offs, p_offs, p_bytes = _open_arena(wires, lens)
    slot = relay.open_stream(None, None, k, h)
    descs = bu.make_descs([slot] * n, lens, w_offs, p_offs, open_=True, run=True)
    with pytest.raises(sb.RelayError, match="run of more than 64"):
        relay.batch_host(descs, arena, bu.aligned_zeros(p_bytes))
    # device entry point: descriptors live in HBM, the library cannot read them -- frames 64.. come back RETRY,
    # the first 64 are opened and committed, and the resubmitted tail continues the chain
    dev = torch.device("cuda:0")
    tiles, nb = relay.plan(descs)
    t_desc = torch.from_numpy(descs.view(np.uint8).copy()).to(dev)
    t_tiles = torch.from_numpy(tiles.view(np.uint8).copy()).to(dev)
    t_in = torch.from_numpy(np.ascontiguousarray(arena)).to(dev)
    t_out = torch.zeros(p_bytes, dtype=torch.uint8, device=dev)
    t_st = torch.full((n,), 0xFF, dtype=torch.int32, device=dev)
    scratch = torch.empty(16 * n, dtype=torch.uint8, device=dev)
    relay.batch_device_runs(t_desc.data_ptr(), n, t_tiles.data_ptr(), len(tiles), nb, t_in.data_ptr(), t_out.data_ptr(),
                            t_st.data_ptr(), scratch.data_ptr(), scratch.numel())
    torch.cuda.synchronize()
    st = t_st.cpu().numpy().astype(np.uint32) & 0xFF
    assert (st[:64] == 0).all() and (st[64:] == sb.relay.ST_RETRY).all(), st
    out = t_out.cpu().numpy()
    for j in range(64):
        assert bytes(out[p_offs[j] : p_offs[j] + lens[j]]) == msgs[j]
    mirror = ls.State(k, h)
    for w in wires[:64]:
        mirror.open_frame(w)
    assert relay.export(slot)[1] == mirror.raw
    tail = list(range(64, n))
    arena2, wo2, po2, pb2 = _open_arena([wires[j] for j in tail], [lens[j] for j in tail])
    out2 = bu.aligned_zeros(pb2)
    st2 = relay.batch_host(bu.make_descs([slot] * len(tail), [lens[j] for j in tail], wo2, po2, open_=True, run=True),
                           arena2, out2)
    assert (st2 == 0).all()
    for k2, j in enumerate(tail):
        assert bytes(out2[po2[k2] : po2[k2] + lens[j]]) == msgs[j]


def test_run_frames_without_scratch_come_back_retry(relay):
    """symrelay_batch_device has no scratch for speculative nonces: F_RUN frames are left alone, not dereferenced."""
    import torch

    import symmetry_b200 as sb

    k, h = bytes(32), bytes(24)
    sender = ls.State(k, h)
    wires = [sender.seal_frame(b"abc"), sender.seal_frame(b"defg" * 30)]
    lens = [3, 120]
    arena, w_offs, p_offs, p_bytes = _open_arena(wires, lens)
    slot = relay.open_stream(None, None, k, h)
    before = relay.export(slot)[1]
    descs = bu.make_descs([slot, slot], lens, w_offs, p_offs, open_=True, run=True)
    tiles, nb = relay.plan(descs)
    dev = torch.device("cuda:0")
    t = [torch.from_numpy(np.ascontiguousarray(x).view(np.uint8).copy()).to(dev) for x in (descs, tiles, arena)]
    t_out = torch.zeros(p_bytes, dtype=torch.uint8, device=dev)
    t_st = torch.full((2,), 0xFF, dtype=torch.int32, device=dev)
    relay.batch_device(t[0].data_ptr(), 2, t[1].data_ptr(), len(tiles), nb, t[2].data_ptr(), t_out.data_ptr(),
                       t_st.data_ptr())
    torch.cuda.synchronize()
    assert (t_st.cpu().numpy() == sb.relay.ST_RETRY).all()
    assert relay.export(slot)[1] == before and not t_out.cpu().numpy().any()


def test_close_forgets_a_half_received_frame(relay):
    """A peer sends a 3-byte prefix that promises a huge frame, a few bytes, and disconnects.  Slots are
    reused LIFO, so the next connection gets that slot: its first bytes must not be eaten as the old
    frame's remainder (secret-stream's `_onrawdata` state dies with its stream, src/types.ts:156-159)."""
    rng = random.Random(9)
    k1, h1, k2, h2 = rng.randbytes(32), rng.randbytes(24), rng.randbytes(32), rng.randbytes(24)
    slot = relay.open_stream(None, None, k1, h1)
    tk, used = relay.rx_feed(slot, bytes([0xFF, 0xFF, 0x7F]) + b"junk" * 10)
    assert tk == [] and used == 43
    relay.close_stream(slot)
    slot2 = relay.open_stream(None, None, k2, h2)
    assert slot2 == slot, "LIFO reuse is what makes the stale state dangerous"
    sender = ls.State(k2, h2)
    msgs = [rng.randbytes(n) for n in (5, 300, 0, 64)]
    blob = b"".join(sender.seal_frame(m) for m in msgs)
    tk, used = relay.rx_feed(slot2, blob[:7])
    tk2, used2 = relay.rx_feed(slot2, blob[7:])
    assert used == 7 and used2 == len(blob) - 7
    relay.flush()
    got = {}
    while len(got) < len(msgs):
        for c in relay.poll(wait=True):
            got[c["ticket"]] = c
    for t, m in zip(tk + tk2, msgs):
        assert got[t]["status"] == 0 and got[t]["data"] == m


def test_close_waits_for_queued_frames(relay):
    import symmetry_b200 as sb

    rng = random.Random(11)
    keys = [rng.randbytes(32) for _ in range(3)]
    hdrs = [rng.randbytes(24) for _ in range(3)]
    slots = [relay.open_stream(k, h, k, h) for k, h in zip(keys, hdrs)]
    t0 = relay.submit_seal(slots[0], b"first")
    with pytest.raises(sb.RelayError) as e:
        relay.close_stream(slots[0])          # its frame sits in the filling batch
    assert e.value.code == sb.relay.EAGAIN
    relay.flush()
    got = {c["ticket"]: c for c in relay.poll(wait=True)}
    assert got[t0]["status"] == 0
    relay.close_stream(slots[0])              # drained: now it closes


def test_a_refused_batch_is_retired_and_the_ring_keeps_moving(monkeypatch):
    """One connection's mistake (or a CUDA failure) must not wedge every other stream of the GPU: the batch
    that was refused completes with ST_FAILED per frame and the next batch goes through."""
    import symmetry_b200 as sb

    monkeypatch.setenv("SYMRELAY_INJECT_REFUSAL", "2")      # the front end's second batch is refused
    rng = random.Random(13)
    keys = [rng.randbytes(32) for _ in range(4)]
    hdrs = [rng.randbytes(24) for _ in range(4)]
    with sb.SymRelay(gpu_ids=(0,), max_streams=64) as r:
        slots = [r.open_stream(k, h, k, h) for k, h in zip(keys, hdrs)]
        refs = [ls.State(k, h) for k, h in zip(keys, hdrs)]
        for rnd, want_status in enumerate([sb.relay.ST_OK, sb.relay.ST_FAILED, sb.relay.ST_OK, sb.relay.ST_OK]):
            msgs = [rng.randbytes(50 + i) for i in range(4)]
            tk = [r.submit_seal(s, m) for s, m in zip(slots, msgs)]
            if want_status == sb.relay.ST_FAILED:
                with pytest.raises(sb.RelayError, match="INJECT"):
                    r.flush()
            else:
                r.flush()
            got = {}
            for _ in range(1000):
                for c in r.poll(wait=True):
                    got[c["ticket"]] = c
                if len(got) == 4:
                    break
            for i, t in enumerate(tk):
                assert got[t]["status"] == want_status, (rnd, got[t])
                if want_status == sb.relay.ST_OK:
                    assert got[t]["data"] == refs[i].seal_frame(msgs[i])   # the refused frames never touched the chain
                else:
                    assert got[t]["data"] == b""


def test_foreign_cuda_users_on_the_thread_do_not_misdirect_the_library():
    """torch (or another addon) may switch the current device between two calls on the same thread."""
    import torch

    import symmetry_b200 as sb

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    rng = random.Random(12)
    k, h = rng.randbytes(32), rng.randbytes(24)
    with sb.SymRelay(gpu_ids=(1,), max_streams=256) as r:
        torch.cuda.set_device(0)
        torch.zeros(1, device="cuda:0")
        slot = r.open_stream(k, h, k, h)
        ref = ls.State(k, h)
        for i in range(4):
            torch.cuda.set_device(i % 2)
            m = rng.randbytes(200)
            plain, wire = r.alloc_pinned(208), r.alloc_pinned(256)
            plain[:200] = np.frombuffer(m, dtype=np.uint8)
            st = r.batch_host(bu.make_descs([slot], [200], [0], [12]), plain, wire)
            assert (st == 0).all() and bytes(wire[12:232]) == ref.seal_frame(m)
        torch.cuda.set_device(0)
        assert r.export(slot)[0] == ref.raw
        # the allocation-free run entry point resolves its scratch on the context's GPU, not the current one
        sender = ls.State(k, h)
        wires = [sender.seal_frame(rng.randbytes(n)) for n in (10, 70)]
        arena, w_offs, p_offs, p_bytes = _open_arena(wires, [10, 70])
        descs = bu.make_descs([slot, slot], [10, 70], w_offs, p_offs, open_=True, run=True)
        tiles, nb = r.plan(descs)
        d1 = torch.device("cuda:1")
        t = [torch.from_numpy(np.ascontiguousarray(x).view(np.uint8).copy()).to(d1) for x in (descs, tiles, arena)]
        t_out = torch.zeros(p_bytes, dtype=torch.uint8, device=d1)
        t_st = torch.full((2,), 0xFF, dtype=torch.int32, device=d1)
        torch.cuda.set_device(0)
        r.batch_device_runs_static(t[0].data_ptr(), 2, t[1].data_ptr(), len(tiles), nb, t[2].data_ptr(),
                                   t_out.data_ptr(), t_st.data_ptr(), stream=torch.cuda.current_stream(d1).cuda_stream)
        torch.cuda.synchronize(d1)
        assert (t_st.cpu().numpy() == 0).all()
