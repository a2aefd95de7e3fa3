"""p50/p99 added relay latency at N concurrent streams under Poisson load (SURVEY 8(d) "Latency").

    python tools/latency_bench.py [--streams 65536] [--rate 2e6] [--bytes 200] [--seconds 2] [--flush-us 100]
"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import symmetry_b200 as sb  # noqa: E402
from symmetry_b200 import build, synth  # noqa: E402


class Result(C.Structure):
    _fields_ = [("p50_us", C.c_double), ("p99_us", C.c_double), ("p999_us", C.c_double), ("max_us", C.c_double),
                ("achieved_rate", C.c_double), ("eagain_per_frame", C.c_double), ("frames", C.c_uint64)]


def run(streams, rate, nbytes, seconds, flush_us, batches=16):
    drv = C.CDLL(build.build_driver())
    drv.latency_run.restype = C.c_int
    drv.latency_run.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_double, C.c_double,
                                C.POINTER(Result)]
    with sb.SymRelay(gpu_ids=(0,), max_streams=max(4096, 2 * streams), flush_us=flush_us, sched_batches=batches) as r:
        keys, hdrs = synth.stream_keys(streams)
        slots = r.open_streams(keys, hdrs, None, None)
        res = Result()
        # warm-up (first launches, pinned arena first touch)
        assert drv.latency_run(r.ctx, slots.ctypes.data, streams, nbytes, rate, 0.2, C.byref(res)) == 0
        rc = drv.latency_run(r.ctx, slots.ctypes.data, streams, nbytes, rate, seconds, C.byref(res))
        assert rc == 0, (rc, r.lib.symrelay_last_error(r.ctx))
        st = r.sched_stats()
    return {"streams": streams, "offered_frames_per_s": rate, "frame_bytes": nbytes, "flush_us": flush_us,
            "frames": res.frames, "p50_added_latency_us": res.p50_us, "p99_us": res.p99_us, "p999_us": res.p999_us,
            "max_us": res.max_us, "achieved_frames_per_s": res.achieved_rate, "eagain_per_frame": res.eagain_per_frame,
            "mean_batch_fill": st["mean_batch_fill"], "zero_copy_batches": st["zero_copy_batches"],
            "staged_batches": st["staged_batches"]}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--streams", type=int, default=65536)
    ap.add_argument("--rate", type=float, default=2e6)
    ap.add_argument("--bytes", type=int, default=200)
    ap.add_argument("--seconds", type=float, default=2.0)
    ap.add_argument("--flush-us", type=int, default=100)
    a = ap.parse_args()
    print(json.dumps(run(a.streams, a.rate, a.bytes, a.seconds, a.flush_us)))
