"""Exploration: which queries of test_exact_tensor_path_query_scales_and_deletes differ from the oracle, and how."""
import os, sys, numpy as np
sys.path.insert(0, '.')
import hnsw_b200
from oracle import binding as oracle
oracle.build()
os.environ["B200EXACT_PATH"] = "tensor"
rng = np.random.default_rng(16)
n, d, k, B = 40000, 128, 10, 96
X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
X[7] *= 4.0
X[11:200] *= 0.01
Q = rng.standard_normal((B, d)).astype(np.float32)
Q[:4] *= 1e5
Q[4:8] *= 1e4
Q[8:16] *= 1e-3
Q[16:20] = X[300:304]
for variant in ("full", "no_small_rows", "no_big_row", "plain_rows"):
    Xv = X.copy()
    if variant in ("no_small_rows", "plain_rows"):
        Xv[11:200] = rng.uniform(-1, 1, (189, d)).astype(np.float32)
    if variant in ("no_big_row", "plain_rows"):
        Xv[7] /= 4.0
    ix = hnsw_b200.NewExactIndex(hnsw_b200.EuclideanDistance)
    ix.BatchAdd(list(range(n)), Xv)
    rows, dists, counts = ix.search_rows(Q, k)
    st = ix.last_stats
    o_idx, o_d = oracle.exact_batch_search(Xv, Q, k, 0, n_threads=8)
    bad = [q for q in range(B) if not np.array_equal(rows[q], o_idx[q])]
    print(variant, "stats", st["launches"], st["expansions"], st["dist_evals"] / B, "bad queries", bad, flush=True)
    for q in bad[:6]:
        print("  q", q, "dev", rows[q].tolist(), "\n       ora", o_idx[q].tolist(), "\n       dev d", dists[q][:4], "ora d", o_d[q][:4], flush=True)
    ix.Close()
