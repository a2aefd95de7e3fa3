"""The step after the path: sealed frames leave as datagrams, many per system call.

The reference's transport is udx-native (/root/reference/package-lock.json:6241-6253, UDXSocket / UDXStream typed at
src/types.ts:91-122), one packet per system call.  symrelay_udp_send_batch is the send primitive of a batched
transport; here a loopback receiver puts every connection's byte stream back together from the datagrams and gets
byte-identical frames -- on the CPU with frames sealed by libsodium, on the GPU through the whole relay:
symrelay_submit_seal -> poll -> sendmmsg -> recv -> symrelay_rx_feed -> plaintext."""
import ctypes as C
import random
import socket
import struct

import numpy as np
import pytest

from oracle import libsodium as ls

MTU = 1200


def _sockets():
    rx = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
    rx.setsockopt(socket.SOL_SOCKET, socket.SO_RCVBUF, 8 << 20)
    rx.bind(("127.0.0.1", 0))
    rx.settimeout(2.0)
    tx = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
    tx.connect(rx.getsockname())          # connected socket: dst == NULL
    return tx, rx


def _collect(rx, n_datagrams):
    """conn -> byte stream, slices ordered by their offset (loopback keeps order, a network would not)."""
    parts = {}
    for _ in range(n_datagrams):
        d = rx.recv(65536)
        magic, conn, off = struct.unpack_from("<IIQ", d)
        assert magic == 0x554D5953 and len(d) <= MTU
        parts.setdefault(conn, []).append((off, d[16:]))
    out = {}
    for conn, sl in parts.items():
        sl.sort()
        pos, blob = sl[0][0], b""
        for off, b in sl:
            assert off == pos, "a slice is missing or overlaps"
            blob += b
            pos += len(b)
        out[conn] = (sl[0][0], blob)
    return out


def _send(lib, _lib, tx, comps, conn_ids, offs, flags=0):
    n = len(comps)
    arr = (_lib.SymrelayCompletion * n)(*comps)
    cid = (C.c_uint32 * n)(*conn_ids)
    so = (C.c_uint64 * n)(*offs)
    st = _lib.SymrelayUdpStats()
    rc = lib.symrelay_udp_send_batch(tx.fileno(), arr, n, None, None, cid, so, MTU, flags, C.byref(st))
    assert rc >= 0, rc
    return rc, list(so), st


def test_frames_survive_packetisation_and_sendmmsg_batches_the_calls():
    from symmetry_b200 import _lib

    lib = _lib.load()
    rng = random.Random(1)
    tx, rx = _sockets()
    n = 40
    senders = [ls.State(rng.randbytes(32), rng.randbytes(24)) for _ in range(n)]
    offs = [rng.randrange(0, 1 << 40) for _ in range(n)]
    start = list(offs)
    streams = {i: b"" for i in range(n)}
    for rnd in range(3):
        frames, bufs, comps = [], [], []
        for i in range(n):
            w = senders[i].seal_frame(rng.randbytes(rng.choice([0, 64, 200, 1163, 1164, 1165, 5000])))
            buf = np.frombuffer(w, dtype=np.uint8).copy()
            bufs.append(buf)
            c = _lib.SymrelayCompletion()
            c.ticket, c.slot, c.status, c.data, c.len, c.flags = i, i, 0, buf.ctypes.data, len(w), 0
            comps.append(c)
            streams[i] += w
        bad = _lib.SymrelayCompletion()          # a failed frame and an opened one never travel
        bad.status, bad.data, bad.len = 1, bufs[0].ctypes.data, 10
        opened = _lib.SymrelayCompletion()
        opened.status, opened.data, opened.len, opened.flags = 0, bufs[0].ctypes.data, 10, 0x04000000
        sent, offs2, st = _send(lib, _lib, tx, comps + [bad, opened], list(range(n)) + [999, 998], offs + [0, 0],
                                flags=1 if rnd == 2 else 0)
        offs = offs2[:n]
        want = sum((len(b) + MTU - 17) // (MTU - 16) for b in bufs)
        assert sent == want == st.datagrams and st.frames == n
        assert st.syscalls == (want if rnd == 2 else 1), "one sendmmsg for the whole batch (one per datagram with the flag)"
        got = _collect(rx, sent)
        for i in range(n):
            first, blob = got[i]
            assert first + len(blob) == offs[i]
            assert streams[i].endswith(blob)
    assert [o - s for o, s in zip(offs, start)] == [len(streams[i]) for i in range(n)]


@pytest.mark.gpu
def test_seal_send_receive_open_through_the_relay():
    import symmetry_b200 as sb
    from symmetry_b200 import _lib

    rng = random.Random(2)
    tx, rx = _sockets()
    n = 64
    keys = [rng.randbytes(32) for _ in range(n)]
    hdrs = [rng.randbytes(24) for _ in range(n)]
    with sb.SymRelay(gpu_ids=(0,), max_streams=1024) as r:
        out_slots = [r.open_stream(k, h, None, None) for k, h in zip(keys, hdrs)]       # the provider's side
        in_slots = [r.open_stream(None, None, k, h) for k, h in zip(keys, hdrs)]        # the client's side
        offs = {s: 0 for s in out_slots}
        for rnd in range(3):
            msgs = {s: rng.randbytes(rng.choice([0, 64, 300, 1400, 4000])) for s in out_slots}
            for s, m in msgs.items():
                r.submit_seal(s, m)
            r.flush()
            n_c = 0
            cbuf = (_lib.SymrelayCompletion * 256)()
            while n_c < n:                                      # completions stay in the pinned ring until released
                k = r.lib.symrelay_poll_wait(r.ctx, C.cast(C.byref(cbuf, n_c * C.sizeof(_lib.SymrelayCompletion)),
                                                           C.POINTER(_lib.SymrelayCompletion)), 256 - n_c)
                assert k >= 0
                n_c += k
            comps = [cbuf[i] for i in range(n_c)]
            conn = [c.slot for c in comps]
            so = [offs[c.slot] for c in comps]
            sent, so2, st = _send(r.lib, _lib, tx, comps, conn, so)
            for c, o in zip(comps, so2):
                offs[c.slot] = o
                r.release(c.ticket)
            got = _collect(rx, sent)                            # ... the network ...
            tickets = {}
            for s_out, s_in in zip(out_slots, in_slots):
                first, blob = got[s_out]
                tk, used = r.rx_feed(s_in, blob)                # secret-stream `_onrawdata` on the other side
                assert used == len(blob) and len(tk) == 1
                tickets[tk[0]] = msgs[s_out]
            r.flush()
            done = {}
            while len(done) < n:
                for c in r.poll(wait=True):
                    done[c["ticket"]] = c
            for t, m in tickets.items():
                assert done[t]["status"] == 0 and done[t]["open"] and done[t]["data"] == m
