"""When do the kernels of a config-3 step run next to each other?  Needs a library built with -DSS_PROBE_TIMELINE
(see c5_timeline.py).  Prints, relative to the thread-tile kernel's first warp: when the last tile of each kind started
and when the last one finished."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "tests"))
import bench  # noqa: E402
from symmetry_b200 import _lib  # noqa: E402

job = bench.Job()
L = _lib.load()
L.symrelay_probe_timeline.argtypes = [C.c_void_p, C.c_int]
for rep in range(3):
    L.symrelay_probe_timeline(None, 1)
    bench.measure_device(job, "c3", 2, 3, True)
    out = (C.c_ulonglong * 32)()
    L.symrelay_probe_timeline(out, 1)
    t0 = out[22]
    rows = [("thread tiles: last tile starts", 28), ("thread tiles: last tile done", 23),
            ("packed tiles: last tile starts", 20), ("packed tiles: last tile done", 21),
            ("strided tiles: last tile starts", 30), ("strided tiles: last tile done", 25),
            ("big frames: last tile starts", 29), ("big frames: last tile done", 27)]
    print(f"run {rep}:")
    for name, k in rows:
        print(f"   {name:34s} {(out[k] - t0) / 1e3 if out[k] else float('nan'):8.2f} us")
