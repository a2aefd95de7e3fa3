"""Send-path cost after the relay: wire frames of one C2 batch (2,048 x 84 B) to a loopback UDP socket,
one system call per datagram (what the reference's transport does per frame) against one sendmmsg per batch.

    python tools/udp_bench.py [--frames 2048] [--bytes 84] [--batches 200]
"""
import argparse
import ctypes as C
import json
import os
import socket
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from symmetry_b200 import _lib  # noqa: E402


def run(frames=2048, nbytes=84, batches=200, mtu=1200):
    lib = _lib.load()
    rx = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
    rx.bind(("127.0.0.1", 0))                      # nobody reads: loopback drops what does not fit, the sender never blocks
    tx = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
    tx.connect(rx.getsockname())
    tx.setblocking(False)
    buf = np.random.default_rng(0).integers(0, 256, frames * 96, dtype=np.uint8)
    comps = (_lib.SymrelayCompletion * frames)()
    for i in range(frames):
        comps[i].slot, comps[i].status, comps[i].data, comps[i].len = i, 0, buf.ctypes.data + 96 * i, nbytes
    cid = (C.c_uint32 * frames)(*range(frames))
    out = {}
    for name, flags in (("sendmsg_per_datagram", 1), ("sendmmsg_per_batch", 0)):
        so = (C.c_uint64 * frames)()
        st = _lib.SymrelayUdpStats()
        for _ in range(5):
            lib.symrelay_udp_send_batch(tx.fileno(), comps, frames, None, None, cid, so, mtu, flags, C.byref(st))
        st = _lib.SymrelayUdpStats()
        t0 = time.perf_counter()
        done = 0
        for _ in range(batches):
            rc = lib.symrelay_udp_send_batch(tx.fileno(), comps, frames, None, None, cid, so, mtu, flags, C.byref(st))
            done += max(rc, 0)
        dt = time.perf_counter() - t0
        out[name] = {"frames_per_s": st.frames / dt, "us_per_frame": dt / max(st.frames, 1) * 1e6, "datagrams": st.datagrams,
                     "syscalls": st.syscalls}
    out["speedup"] = out["sendmmsg_per_batch"]["frames_per_s"] / out["sendmsg_per_datagram"]["frames_per_s"]
    out["what"] = (f"{frames} wire frames of {nbytes} B per batch to a loopback UDP socket, one host thread; the send primitive only "
                   "(16-byte header + gather from the ring), no reliability layer")
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=2048)
    ap.add_argument("--bytes", type=int, default=84)
    ap.add_argument("--batches", type=int, default=200)
    a = ap.parse_args()
    print(json.dumps(run(a.frames, a.bytes, a.batches)))
