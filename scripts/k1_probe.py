"""Exploration only (not the bench): K1 timing on an oracle-built graph, with a parity check."""
import sys, time, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from oracle import binding as oracle
from util import build_oracle_graph, flatten_oracle
import hnsw_b200
N = int(sys.argv[1]); d = int(sys.argv[2]); M = int(sys.argv[3]); ef = int(sys.argv[4])
Bs = [int(x) for x in sys.argv[5].split(',')]
rng = np.random.default_rng(42)
X = rng.uniform(-1, 1, (N, d)).astype(np.float32)
t = time.time(); g, _, _ = build_oracle_graph(oracle, X, M, 0.25, ef, 0); print("oracle build s", time.time() - t, g.topography(), flush=True)
dg = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=hnsw_b200.EuclideanDistance, capacity=N)
t = time.time(); dg.load_flat(*flatten_oracle(g)); print("load s", time.time() - t, flush=True)
for B in Bs:
    Q = np.random.default_rng(43).uniform(-1, 1, (B, d)).astype(np.float32)
    for it in range(4):
        rows, dists, cnt = dg.search_rows(Q, 10, ef)
        st = dg.last_stats
        byts = st['dist_evals'] * 4 * d + st['nbr_reads'] * 4
        print(f"B={B} it={it} kernel_ms={st['kernel_ms']:.3f} qps={B/st['kernel_ms']*1e3:.3e} E/q={st['dist_evals']/B:.1f} H/q={st['expansions']/B:.1f} GB/s={byts/st['kernel_ms']/1e6:.1f}", flush=True)
    if B <= 20000:
        t = time.time(); o_rows, o_d, o_cnt, o_ctr = g.batch_search(Q, 10, ef, n_threads=8); dt = time.time() - t
        print("oracle 8thr qps", B / dt, "parity rows", np.array_equal(o_rows, rows), "dists", np.array_equal(o_d.view(np.uint32), dists.view(np.uint32)), "E", o_ctr['dist_evals'] == st['dist_evals'], flush=True)
