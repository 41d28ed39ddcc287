"""K2 parity: device Graph.Add / BatchAdd vs the oracle's restatement of graph.go:437-531.

Same vectors, keys, levels and parameters -> the whole graph must be identical: every neighbour list of every
layer (rows, in key order), topography, entries, and the E / eviction / replenish counters.
"""
import numpy as np
import pytest

from util import draw_levels

pytestmark = pytest.mark.gpu


def _build_both(oracle, X, keys, levels, M, ef, metric, chunks=(None,)):
    import hnsw_b200
    n, d = X.shape
    g = oracle.Graph(d, M, metric)
    g.add_many(keys, X, levels, ef)
    dist = hnsw_b200.EuclideanDistance if metric == 0 else hnsw_b200.CosineDistance
    dg = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=dist)
    totals = {"dist_evals": 0, "evict_evals": 0, "replenish_evals": 0, "evictions": 0, "expansions": 0}
    bounds = [0] + [c for c in chunks if c is not None] + [n]
    for a, b in zip(bounds, bounds[1:]):
        if b <= a:
            continue
        dg.BatchAdd([hnsw_b200.Node(int(keys[i]), X[i]) for i in range(a, b)], levels=levels[a:b])
        for kk in totals:
            totals[kk] += dg.last_stats[kk]
    return g, dg, totals


def _assert_same_graph(g, dg, M):
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


@pytest.mark.parametrize("n,d,M,ef,metric", [
    (600, 128, 16, 20, 0),    # C1 parameters
    (500, 96, 16, 64, 0),     # C2/C4 parameters
    (300, 768, 32, 64, 1),    # C3: cosine, M=32
    (400, 8, 6, 20, 0),       # low dimension: many evictions and replenishes
    (128, 1, 6, 20, 0),       # graph_test.go:86-133 shape: 128 points on a line, M=6
    (200, 40, 48, 30, 1),     # M > 32
    (260, 24, 64, 100, 0),    # MAX_M, ef > M
])
def test_add_matches_oracle(oracle, n, d, M, ef, metric):
    rng = np.random.default_rng(7)
    X = rng.uniform(-1, 1, (n, d)).astype(np.float32) if d > 1 else np.arange(n, dtype=np.float32).reshape(n, 1)
    keys = np.arange(n, dtype=np.int64)
    levels = draw_levels(oracle, n, 0.25 if d > 1 else 0.5, 44)
    g, dg, tot = _build_both(oracle, X, keys, levels, M, ef, metric)
    _assert_same_graph(g, dg, M)
    c = g.counters()
    assert tot["dist_evals"] == c["dist_evals"]
    assert tot["expansions"] == c["expansions"]
    assert tot["evict_evals"] == c["evict_evals"]
    assert tot["replenish_evals"] == c["replenish_evals"]
    assert tot["evictions"] == c["evictions"]


def test_add_in_several_calls_and_shuffled_keys(oracle):
    # keys not in insertion order: the sorted-by-key iteration order differs from row order
    rng = np.random.default_rng(8)
    n, d, M, ef = 500, 32, 8, 20
    X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    keys = rng.permutation(10 * n)[:n].astype(np.int64) - 2000  # negative keys too
    levels = draw_levels(oracle, n, 0.25, 45)
    g, dg, _ = _build_both(oracle, X, keys, levels, M, ef, 0, chunks=(1, 2, 50, 333))
    _assert_same_graph(g, dg, M)
    # and the searches agree afterwards
    Q = rng.uniform(-1, 1, (64, d)).astype(np.float32)
    o_rows, o_d, o_cnt, _ = g.batch_search(Q, 10, ef)
    d_rows, d_d, d_cnt = dg.search_rows(Q, 10, ef)
    assert np.array_equal(o_rows, d_rows) and np.array_equal(o_d.view(np.uint32), d_d.view(np.uint32))


def test_add_search_api_like_reference(oracle):
    # graph_test.go:86-133 through the public API: 128 1-D points, M=6, Ml=.5, ef=20, Euclidean
    import hnsw_b200
    g = hnsw_b200.Graph(M=6, Ml=0.5, EfSearch=20, Distance=hnsw_b200.EuclideanDistance, Rng=hnsw_b200.LevelRand(0))
    for i in range(128):
        g.Add(hnsw_b200.Node(i, [float(i)]))
    topo = g.Topography()
    assert topo[0] == 128 and g.Len() == 128 and all(a >= b for a, b in zip(topo, topo[1:]))
    nearest = g.Search([64.5], 4)
    assert len(nearest) == 4
    assert {nd.Key for nd in nearest} <= set(range(128))
    # graph_test.go:253-275
    g2 = hnsw_b200.NewGraph()
    g2.Add(hnsw_b200.Node(1, [1, 1]), hnsw_b200.Node(2, [0, 1]), hnsw_b200.Node(3, [1, -1]))
    out = g2.Search([0.5, 0.5], 1)
    assert len(out) == 1 and out[0].Key == 1 and list(out[0].Value) == [1, 1]
    with pytest.raises(ValueError, match="embedding dimension mismatch: 2 != 3"):
        g2.Add(hnsw_b200.Node(9, [1, 2, 3]))


def test_whole_graph_kat_on_device_under_the_reference_topography(oracle):
    """graph_test.go:86-133 on the device (see tests/test_oracle_kat.py for the CPU half): for level assignments with the
    reference's topography, exact sequential Add + Search([64.5], 4) through the public API return what the oracle
    returns, and the reference's list [64, 65, 62, 63] -- in its heap-array order -- is the most frequent answer."""
    import hnsw_b200
    topo = [128, 67, 28, 12, 6, 2, 1, 1]
    n = 128
    rng = np.random.default_rng(0)
    seen = {}
    for _ in range(24):
        lv = np.zeros(n, np.int32)
        members = np.arange(n)
        for l in range(1, len(topo)):
            members = rng.choice(members, topo[l], replace=False)
            lv[members] = l
        og = oracle.Graph(1, 6, oracle.EUCLIDEAN)
        for i in range(n):
            og.add(i, [float(i)], int(lv[i]), 20)
        want = [int(x) for x in og.batch_search(np.array([[64.5]], np.float32), 4, 20)[0][0]]
        g = hnsw_b200.Graph(M=6, Ml=0.5, EfSearch=20, Distance=hnsw_b200.EuclideanDistance)
        g.Add(*[hnsw_b200.Node(i, [float(i)]) for i in range(n)], levels=[int(x) for x in lv])
        assert g.Topography() == topo
        got = [nd.Key for nd in g.Search([64.5], 4)]
        assert got == want
        seen[tuple(got)] = seen.get(tuple(got), 0) + 1
        g.close()
    assert max(seen, key=seen.get) == (64, 65, 62, 63), seen
