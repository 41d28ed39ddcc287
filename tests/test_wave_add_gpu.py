"""K2b parity: device wave-batched Add (b200hnsw_add_batch_waves) vs the oracle's restatement of the SAME
specification (ora_graph_add_waves).  This is not a reference-parity test -- the reference inserts sequentially
(tests/test_add_parity_gpu.py covers that) -- it pins the parallel device schedule to a serial CPU restatement:
every neighbour list of every layer, topography, entries and all counters must be identical.
"""
import numpy as np
import pytest

from util import draw_levels

pytestmark = pytest.mark.gpu


def _assert_same_graph(g, dg):
    assert g.topography() == dg.Topography()
    assert list(g.entries()) == list(dg.entries())
    for l in range(g.n_layers):
        rows, ptr, nbr = g.dump_layer(l)
        dn, member = dg.neighbors(l)
        assert list(np.nonzero(member)[0]) == list(rows)
        for i, r in enumerate(rows):
            want = list(nbr[ptr[i]:ptr[i + 1]])
            got = [int(v) for v in dn[r] if v >= 0]
            assert got == want, (l, int(r), got, want)


@pytest.mark.parametrize("n,d,M,ef,metric,wave_max", [
    (3000, 128, 16, 64, 0, 4096),   # C2 parameters
    (2000, 96, 16, 20, 0, 64),      # C1/C4 parameters, small waves
    (1200, 768, 32, 64, 1, 256),    # C3: cosine, M=32
    (1500, 8, 6, 20, 0, 4096),      # low dimension: many evictions, several requests per target
    (600, 40, 48, 30, 1, 128),      # M > 32
    (400, 16, 8, 20, 0, 1),         # waves of one node
    (900, 32, 64, 128, 0, 512),     # MAX_M
])
def test_wave_add_matches_oracle(oracle, n, d, M, ef, metric, wave_max):
    import hnsw_b200
    rng = np.random.default_rng(11)
    X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    keys = rng.permutation(4 * n)[:n].astype(np.int64) - n
    levels = draw_levels(oracle, n, 0.25, 46)
    g = oracle.Graph(d, M, metric)
    g.add_waves(keys, X, levels, ef, wave_max, n_threads=4)
    dist = hnsw_b200.EuclideanDistance if metric == 0 else hnsw_b200.CosineDistance
    dg = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=dist)
    tot = {}
    for a, b in ((0, n // 3), (n // 3, n)):  # two calls: the second one starts from a non-empty graph
        dg.add_arrays(keys[a:b], X[a:b], levels[a:b], wave_max=wave_max)
        for kk, v in dg.last_stats.items():
            tot[kk] = tot.get(kk, 0) + v
    g2 = oracle.Graph(d, M, metric)
    g2.add_waves(keys[:n // 3], X[:n // 3], levels[:n // 3], ef, wave_max)
    g2.add_waves(keys[n // 3:], X[n // 3:], levels[n // 3:], ef, wave_max)
    _assert_same_graph(g2, dg)
    c = g2.counters()
    for kk in ("dist_evals", "expansions", "nbr_reads", "evict_evals", "replenish_evals", "evictions"):
        assert tot[kk] == c[kk], kk
    # searches on the wave-built graph agree bit for bit too
    Q = rng.uniform(-1, 1, (128, d)).astype(np.float32)
    o_rows, o_d, o_cnt, _ = g2.batch_search(Q, 10, ef)
    d_rows, d_d, d_cnt = dg.search_rows(Q, 10, ef)
    assert np.array_equal(o_rows, d_rows) and np.array_equal(o_d.view(np.uint32), d_d.view(np.uint32))
    assert len(g) == n


def test_wave_schedule_matches_oracle(oracle):
    # the wave partition is part of the specification: same formula on both sides (exercised through the graphs above);
    # here: the oracle's own schedule function on a few hand cases
    L = oracle.lib()
    import ctypes as C
    lv = np.array([0, 0, 3, 0, 0], dtype=np.int32)
    p = lv.ctypes.data_as(C.POINTER(C.c_int32))
    assert L.ora_wave_len(0, 0, p, 5, 4096) == 1          # first node alone
    assert L.ora_wave_len(100, 2, p, 5, 4096) == 2         # stops before the node that opens layers 2..3
    lv2 = np.array([3, 0, 0], dtype=np.int32)
    assert L.ora_wave_len(100, 2, lv2.ctypes.data_as(C.POINTER(C.c_int32)), 3, 4096) == 1
    assert L.ora_wave_len(100, 4, p, 5, 4096) == 5         # cap = 100/16 = 6
    assert L.ora_wave_len(100, 4, p, 5, 2) == 2
    assert L.ora_wave_len(10, 4, p, 5, 4096) == 1          # tiny graph: cap = max(1, 10/16)
