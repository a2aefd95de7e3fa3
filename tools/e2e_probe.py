"""Host-path probe: per-call cost and round-trip latency of symrelay_batch_submit/_wait (C2 shape)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import batch_util as bu  # noqa: E402
import symmetry_b200 as sb  # noqa: E402
from symmetry_b200 import synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
mlen = int(sys.argv[2]) if len(sys.argv) > 2 else 64
for zc in (True, False):
    DEPTH = int(os.environ.get("PROBE_DEPTH", "4"))
    with sb.SymRelay(gpu_ids=(0,), max_streams=4 * n, zero_copy=zc, pipeline_depth=max(DEPTH, 2)) as r:
        keys, hdrs = synth.stream_keys(n)
        slots = r.open_streams(keys, hdrs, keys, hdrs)
        lens = [mlen] * n
        p_offs, P = bu.plain_layout(lens)
        w_offs, Wb = bu.wire_layout(lens)
        d = bu.make_descs(slots, lens, p_offs, w_offs)
        R = 8
        plain = [r.alloc_pinned(P) for _ in range(R)]
        wire = [r.alloc_pinned(Wb) for _ in range(R)]
        stat = [np.zeros(n, dtype=np.uint32) for _ in range(R)]
        for p in plain:
            p[:] = 7
        # round-trip latency, one batch in flight
        lat = []
        for i in range(300):
            t0 = time.perf_counter()
            t = r.submit(d, plain[i % R], P, wire[i % R], Wb, stat[i % R])
            t1 = time.perf_counter()
            r.wait(t)
            t2 = time.perf_counter()
            lat.append((t1 - t0, t2 - t0))
        lat = np.array(lat[50:]) * 1e6
        # pipelined throughput, depth 4
        tick = {}
        t0 = time.perf_counter()
        K = 5000
        sub = 0.0
        for i in range(K):
            if i - 4 in tick:
                r.wait(tick.pop(i - 4))
            a = time.perf_counter()
            tick[i] = r.submit(d, plain[i % R], P, wire[i % R], Wb, stat[i % R])
            sub += time.perf_counter() - a
        for k in sorted(tick):
            r.wait(tick[k])
        dt = time.perf_counter() - t0
        # the same loop from C (tools/e2e_driver.c): no interpreter in the way
        import ctypes as C
        from symmetry_b200 import build
        drv = C.CDLL(build.build_driver())
        drv.e2e_ring_run.restype = C.c_double
        drv.e2e_ring_run.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64,
                                     C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32, C.c_void_p,
                                     C.POINTER(C.c_double)]
        slot_bytes = P + Wb
        ring = r.alloc_pinned(16 * slot_bytes)
        ring[:] = 9
        dd = bu.make_descs(slots, lens, p_offs, w_offs)
        st2 = np.zeros(16 * n, dtype=np.uint32)
        subs = C.c_double(0)
        secs = drv.e2e_ring_run(r.ctx, dd.ctypes.data, n, ring.ctypes.data, slot_bytes, P, P, Wb, 16, DEPTH, 0, K,
                                st2.ctypes.data, C.byref(subs))
        assert secs > 0, secs
        print(f"   C driver (depth {DEPTH}): pipelined {secs/K*1e6:.1f} us/batch (submit calls {subs.value/K*1e6:.1f} us/batch)")
        print(f"zero_copy={zc} n={n} mlen={mlen}: submit call p50 {np.median(lat[:,0]):.1f} us, round trip p50 "
              f"{np.median(lat[:,1]):.1f} us p99 {np.percentile(lat[:,1],99):.1f} us; pipelined {dt/K*1e6:.1f} us/batch "
              f"(submit calls {sub/K*1e6:.1f} us/batch)")
