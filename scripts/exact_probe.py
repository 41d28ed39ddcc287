"""Exploration: K3 exact search timing (tensor path) with a sampled oracle check."""
import os, sys, time, numpy as np
sys.path.insert(0, '.')
import hnsw_b200
from oracle import binding as oracle
oracle.build()
for spec in sys.argv[1:]:
    N, d, B, k, metric = [int(v) for v in spec.split(',')]
    rng = np.random.default_rng(1)
    X = rng.standard_normal((N, d), dtype=np.float32)
    Q = rng.standard_normal((B, d), dtype=np.float32)
    ix = hnsw_b200.NewExactIndex(hnsw_b200.EuclideanDistance if metric == 0 else hnsw_b200.CosineDistance, capacity=N)
    t = time.time(); ix.BatchAdd(list(range(N)), X); print(f"N={N} d={d}: add {time.time()-t:.2f}s", flush=True)
    for it in range(3):
        t = time.time(); rows, dists, cnt = ix.search_rows(Q, k); wall = time.time() - t
        st = ix.last_stats
        fl = 2.0 * B * N * d
        print(f"  B={B} k={k} metric={metric} it={it}: device {st['kernel_ms']:.1f} ms wall {wall*1e3:.1f} ms  {fl/st['kernel_ms']/1e9:.1f} TFLOP/s  "
              f"path={st['launches']} cands/query={st['dist_evals']/B:.1f} fallbacks={st['expansions']} qps={B/st['kernel_ms']*1e3:.0f}", flush=True)
    ns = min(B, 16)
    o_idx, o_d = oracle.exact_batch_search(X, Q[:ns], k, metric, n_threads=16)
    print("  oracle sample rows equal:", np.array_equal(rows[:ns], o_idx.astype(np.int64)), "dists bit-equal:", np.array_equal(dists[:ns].view(np.uint32), o_d.view(np.uint32)), flush=True)
    ix.Close()
