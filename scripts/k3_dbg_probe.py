"""Exploration: K3 tensor path timing with parts of the epilogue switched off (B200EXACT_DBG; results are then wrong)."""
import os, sys, time, numpy as np
sys.path.insert(0, '.')
import hnsw_b200
N, d, B, k, metric = [int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "1000000,128,10000,10,0").split(',')]
rng = np.random.default_rng(1)
X = rng.standard_normal((N, d), dtype=np.float32)
Q = rng.standard_normal((B, d), dtype=np.float32)
ix = hnsw_b200.NewExactIndex(hnsw_b200.EuclideanDistance if metric == 0 else hnsw_b200.CosineDistance, capacity=N)
ix.BatchAdd(list(range(N)), X)
for dbg in (None, None, 1, 3, 7, 2, None):
    if dbg is None:
        os.environ.pop("B200EXACT_DBG", None)
    else:
        os.environ["B200EXACT_DBG"] = str(dbg)
    best = 1e9
    for it in range(3):
        ix.search_rows(Q, k)
        best = min(best, ix.last_stats['kernel_ms'])
    print(f"dbg={dbg}: device {best:.3f} ms", flush=True)
