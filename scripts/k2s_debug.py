"""Windowed Add from an empty graph, with the trace on: what the kernel reports per launch."""
import os, sys
os.environ["B200HNSW_TRACE_ADD"] = "1"
sys.path.insert(0, ".")
import numpy as np
import hnsw_b200
from benchlib.data import draw_levels
n, d, M, ef = 600, 128, 16, 20
rng = np.random.default_rng(7)
X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
lv = draw_levels(n, 0.25, 44)
for w in (1, 2, 7):
    g = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=hnsw_b200.EuclideanDistance)
    g.set_add_window(w)
    g.add_arrays(np.arange(n, dtype=np.int64), X, lv, wave_max=0)
    print("window", w, "topography", g.Topography(), "entries", list(g.entries()), "stats", g.add_window_stats(), g.last_stats, flush=True)
    g.close()
