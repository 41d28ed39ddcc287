"""Reference-order Graph.Add (b200hnsw_add_batch) on a graph of `rows` rows: nodes per second for each window of the
speculative schedule (1 = the one-node-at-a-time kernel).  The base graph is wave-built (fast); every leg adds `batch` new
rows to it in the reference's order, so the row count grows by `batch` per leg.
usage: k2s_probe.py rows dim batch [windows, comma separated; 0 = automatic]"""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import hnsw_b200
from benchlib.data import draw_levels, gen

rows, dim, batch = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
windows = [int(w) for w in (sys.argv[4] if len(sys.argv) > 4 else "1,4,8,16,0").split(",")]
M, EF = 16, 64
total = rows + batch * len(windows)
X = gen("uniform", total, dim, 42)
levels = draw_levels(total, 0.25, 44)
g = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=EF, Distance=hnsw_b200.EuclideanDistance, capacity=total)
t0 = time.time()
g.add_arrays(np.arange(rows, dtype=np.int64), X[:rows], levels[:rows], wave_max=16384)
print(f"base graph {rows} x {dim} (waves): {time.time() - t0:.1f} s", flush=True)
at = rows
for w in windows:
    g.set_add_window(w)
    a0, r0, w0 = g.add_window_stats()
    t0 = time.time()
    g.add_arrays(np.arange(at, at + batch, dtype=np.int64), X[at:at + batch], levels[at:at + batch], wave_max=0)
    wall = time.time() - t0
    a1, r1, w1 = g.add_window_stats()
    st = g.last_stats
    print(f"rows {at:8d} window {w:3d}: {batch / wall:9.0f} vec/s wall, kernel {st['kernel_ms']:8.1f} ms "
          f"({batch / st['kernel_ms'] * 1e3:9.0f} vec/s), accepted {a1 - a0} run again {r1 - r0} repaired {w1 - w0}, launches {st['launches']}",
          flush=True)
    at += batch
