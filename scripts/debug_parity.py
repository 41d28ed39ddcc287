import sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from oracle import binding as oracle
from util import build_oracle_graph, flatten_oracle
import hnsw_b200
n,d,M,ef,k,metric = [int(x) for x in sys.argv[1:7]]
rng = np.random.default_rng(42)
X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
g,_,_ = build_oracle_graph(oracle, X, M, 0.25, ef, metric)
dist = hnsw_b200.EuclideanDistance if metric == 0 else hnsw_b200.CosineDistance
dg = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=dist)
dg.load_flat(*flatten_oracle(g))
Q = np.random.default_rng(43).uniform(-1, 1, (257, d)).astype(np.float32)
o_rows, o_d, o_cnt, o_ctr = g.batch_search(Q, k, ef)
d_rows, d_d, d_cnt = dg.search_rows(Q, k, ef)
print("topo", g.topography(), "entries", g.entries(), dg.entries())
print("ctr", o_ctr, dg.last_stats)
bad = [i for i in range(len(Q)) if not (np.array_equal(o_rows[i], d_rows[i]) and np.array_equal(o_d[i].view(np.uint32), d_d[i].view(np.uint32)))]
print("bad queries", len(bad), bad[:10])
for i in bad[:3]:
    print(i, "oracle", o_rows[i], o_d[i], o_cnt[i]); print(i, "device", d_rows[i], d_d[i], d_cnt[i])
    # per-layer trace with oracle
    ent = g.entries()
    el = -1
    for l in range(g.n_layers-1, -1, -1):
        sp = ent[l] if el < 0 else el
        r, dd = g.layer_search(l, int(sp), 1 if l > 0 else k, ef, Q[i])
        print("  layer", l, "sp", sp, "->", r, dd)
        if l > 0 and len(r): el = int(r[0])
