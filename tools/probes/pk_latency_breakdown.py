"""Where a synchronous batch (one in flight) spends its time inside the resident kernel.
    SYMRELAY_TS=1 SYMRELAY_TS_DUMP=/tmp/pk.bin python bench.py --steps 20 --warmup 5 --sub none --no-latency --no-cpu
    python tools/probes/pk_latency_breakdown.py /tmp/pk.bin 1000
The last N batches of such a run are bench.py's batch-latency leg."""
import sys

import numpy as np

path, last_n = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 1000
raw = np.fromfile(path, dtype=np.uint64)
posted, words = int(raw[0]), int(raw[1])
ts = raw[2:].reshape(4096, words).astype(np.int64)
rows = []
for k in range(posted - last_n, posted):
    c = ts[k % 4096]
    if c[0] and c[1] and c[7] and c[7] >= c[0]:
        rows.append(c)
c = np.array(rows)
def us(x):
    return f"{np.median(x) / 1e3:6.2f} (p90 {np.percentile(x, 90) / 1e3:6.2f})"
print(f"{len(rows)} batches")
print("doorbell seen -> tile 0 starts          ", us(c[:, 1] - c[:, 0]))
print("tile 0: loads issued                    ", us(c[:, 2] - c[:, 1]))
print("tile 0: dependency met                  ", us(c[:, 3] - c[:, 2]))
print("tile 0: compute + stores                ", us(c[:, 4] - c[:, 3]))
print("tile 0: fence + count                   ", us(c[:, 5] - c[:, 4]))
print("tile 0 signalled -> batch drained       ", us(c[:, 7] - c[:, 5]))
print("doorbell seen -> last tile starts       ", us(c[:, 8] - c[:, 0]))
print("last tile: start -> loads issued        ", us(c[:, 9] - c[:, 8]))
print("last tile: loads issued -> dependency   ", us(c[:, 10] - c[:, 9]))
print("last tile: dependency -> computed       ", us(c[:, 13] - c[:, 10]))
print("last tile: computed -> signalled        ", us(c[:, 12] - c[:, 13]))
print("last signal -> results in host memory   ", us(c[:, 7] - c[:, 12]))
print("doorbell seen -> results in host memory ", us(c[:, 7] - c[:, 0]))
print("previous batch drained -> doorbell seen ", us(c[1:, 0] - c[:-1, 7]))
