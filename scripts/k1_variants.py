"""Exploration: K1 variants (B200HNSW_K1_VARIANT / B200HNSW_K1_G) on a wave-built graph."""
import os, sys, time, numpy as np
sys.path.insert(0, '.')
import os
os.environ.setdefault('B200HNSW_E2E_CHUNKS', '1')  # kernel_ms = the traversal kernel alone
import hnsw_b200
from bench import synth_rows, draw_levels
N = int(sys.argv[1]); d = 128; M = 16; ef = 64
variants = sys.argv[2].split(',')
Bs = [int(x) for x in sys.argv[3].split(',')]
X = synth_rows(N, d, 42); lv = draw_levels(N, 0.25, 44)
g = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=hnsw_b200.EuclideanDistance, capacity=N)
g.add_arrays(np.arange(N, dtype=np.int64), X, lv, wave_max=4096)
print("built", g.Topography(), "build launches", g.last_stats["launches"], flush=True)
import torch
torch.cuda.cudart().cudaProfilerStart()
ref = {}
for v in variants:
    os.environ.pop("B200HNSW_K1_VARIANT", None); os.environ.pop("B200HNSW_K1_G", None)
    if v.startswith("G"): os.environ["B200HNSW_K1_G"] = v[1:]
    elif v != "0": os.environ["B200HNSW_K1_VARIANT"] = v
    for B in Bs:
        Q = synth_rows(B, d, 43)
        ms = []
        for it in range(6):
            rows, dists, cnt = g.search_rows(Q, 10, ef)
            ms.append(g.last_stats['kernel_ms'])
        if B not in ref: ref[B] = rows.copy()
        ok = np.array_equal(ref[B], rows)
        best = min(ms[1:])
        print(f"variant={v} B={B} kernel_ms best={best:.3f} med={np.median(ms[1:]):.3f} qps={B/best*1e3:.3e} same={ok}", flush=True)
