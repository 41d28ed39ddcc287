import sys, time, numpy as np
sys.path.insert(0, '.')
import hnsw_b200
N = int(sys.argv[1]); d = int(sys.argv[2]); M = int(sys.argv[3]); ef = int(sys.argv[4])
rng = np.random.default_rng(42)
X = rng.uniform(-1, 1, (N, d)).astype(np.float32)
g = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=hnsw_b200.EuclideanDistance, Rng=hnsw_b200.LevelRand(44), capacity=N)
step = max(N // 10, 1)
for a in range(0, N, step):
    b = min(N, a + step)
    t = time.time()
    g.BatchAdd([hnsw_b200.Node(i, X[i]) for i in range(a, b)])
    st = g.last_stats
    print(f"rows {a}-{b}: wall {time.time()-t:.2f}s kernel {st['kernel_ms']/1e3:.2f}s -> {(b-a)/(st['kernel_ms']/1e3):.0f} vec/s; per insert: E={st['dist_evals']/(b-a):.0f} evict={st['evict_evals']/(b-a):.0f} repl={st['replenish_evals']/(b-a):.0f} evictions={st['evictions']/(b-a):.1f}", flush=True)
print(g.Topography())
