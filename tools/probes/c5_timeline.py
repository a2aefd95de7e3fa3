"""When do the two kernels of an extract batch run?  Needs a library built with -DSS_PROBE_TIMELINE:
    python -c "from symmetry_b200 import build; build.build(force=True, extra=['-DSS_PROBE_TIMELINE'], out='/root/repo/symmetry_b200/libsymrelay_probe.so')"
    SYMRELAY_SO=symmetry_b200/libsymrelay_probe.so python tools/probes/c5_timeline.py
Prints, relative to the extractor's first block: the times at which its roles finished and the seal kernel's thread
tiles started, finished computing (before the late wait) and finished."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import batch_util as bu  # noqa: E402
import symmetry_b200 as sb  # noqa: E402
from symmetry_b200 import _lib, synth  # noqa: E402

n = 16384
relay = sb.SymRelay(gpu_ids=(0,), max_streams=1 << 16)
L = _lib.load()
keys, hdrs = synth.stream_keys(n)
slots = relay.open_streams(keys, hdrs, None, None, provider=sb.PROVIDER_OLLAMA)
streams = [synth.chat_chunks(8, seed=5, stream_id=i) for i in range(n)]
for step in range(6):
    chunks = [s[step] for s in streams]
    lens = [len(c) for c in chunks]
    p_offs, p_bytes = bu.plain_layout(lens)
    w_offs, w_bytes = bu.wire_layout(lens)
    plain = relay.alloc_pinned(p_bytes)
    for c, o in zip(chunks, p_offs):
        plain[o:o + len(c)] = np.frombuffer(c, dtype=np.uint8)
    wire = relay.alloc_pinned(w_bytes)
    darena = relay.alloc_pinned(p_bytes)
    delta = np.zeros(n, dtype=sb.DELTA_DTYPE)
    d = bu.make_descs(slots, lens, p_offs, w_offs, extract=True)
    L.symrelay_probe_timeline.argtypes = [C.c_void_p, C.c_int]
    L.symrelay_probe_timeline(None, 1)
    st = relay.batch_host(d, plain, wire, delta=delta, delta_arena=darena)
    assert (st == 0).all()
    out = (C.c_ulonglong * 16)()
    L.symrelay_probe_timeline(out, 1)
    t0 = out[0]
    names = {6: "extract: last block past its set-up", 4: "extract: last grammar warp done", 7: "extract: last recording warp done",
             10: "extract: bookkeeping has event 0", 11: "extract: bookkeeping has event 10", 12: "extract: bookkeeping has event 20",
             9: "extract: bookkeeping saw the end of the list", 13: "extract: before finish()", 14: "extract: after finish()", 1: "extract: last bookkeeping warp done", 2: "seal: tile 0 starts", 8: "seal: last tile starts",
             5: "seal: last tile done computing", 3: "seal: last tile done"}
    print(f"step {step} (mean chunk {np.mean(lens):.0f} B):")
    for k, v in names.items():
        print(f"   {v:42s} {(out[k] - t0) / 1e3 if out[k] else float('nan'):8.2f} us")
