"""K2s parity: the speculative windowed schedule of Graph.Add / BatchAdd (hnsw_b200/csrc/spec_insert.cu) builds the graph of
graph.go:437-531 bit for bit -- every neighbour list of every layer, entries, topography and the E / eviction / replenish
counters equal the oracle's, whatever the window.

The graphs here are SMALL on purpose: with a few hundred rows nearly every speculation reads a list an earlier commit of
its window wrote, so the reject-and-run-again path is what mostly runs; the 30 000-row case has both kinds.  Windows are
forced (the automatic choice is one node at a time below 2 000 rows).
"""
import numpy as np
import pytest

from util import draw_levels

pytestmark = pytest.mark.gpu


def _build(oracle, X, keys, levels, M, ef, metric, window, chunks=()):
    import hnsw_b200
    n, d = X.shape
    dist = hnsw_b200.EuclideanDistance if metric == 0 else hnsw_b200.CosineDistance
    dg = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=dist)
    dg.set_add_window(window)
    totals = {"dist_evals": 0, "evict_evals": 0, "replenish_evals": 0, "evictions": 0, "expansions": 0}
    bounds = [0] + list(chunks) + [n]
    for a, b in zip(bounds, bounds[1:]):
        if b > a:
            dg.add_arrays(keys[a:b], X[a:b], levels[a:b], wave_max=0)
            for kk in totals:
                totals[kk] += dg.last_stats[kk]
    return dg, totals


def _same_graph(g, dg):
    assert g.topography() == dg.Topography()
    assert list(g.entries()) == list(dg.entries())
    for l in range(g.n_layers):
        rows, ptr, nbr = g.dump_layer(l)
        dn, member = dg.neighbors(l)
        assert list(np.nonzero(member)[0]) == list(rows)
        want = np.full((len(rows), dn.shape[1]), -1, np.int32)
        for i in range(len(rows)):
            want[i, :ptr[i + 1] - ptr[i]] = nbr[ptr[i]:ptr[i + 1]]
        bad = np.nonzero((dn[rows] != want).any(axis=1))[0]
        assert len(bad) == 0, (l, int(rows[bad[0]]), dn[rows[bad[0]]], want[bad[0]])


@pytest.mark.parametrize("window", [2, 7, 32])
@pytest.mark.parametrize("n,d,M,ef,metric", [
    (600, 128, 16, 20, 0),    # C1 parameters
    (300, 768, 32, 64, 1),    # C3: cosine, M=32, 6 chunks of a row per lane
    (400, 8, 6, 20, 0),       # low dimension: many evictions and replenishes
    (128, 1, 6, 20, 0),       # graph_test.go:86-133 shape
    (260, 24, 64, 100, 0),    # MAX_M, ef > M
])
def test_windowed_add_matches_oracle(oracle, n, d, M, ef, metric, window):
    rng = np.random.default_rng(7)
    X = rng.uniform(-1, 1, (n, d)).astype(np.float32) if d > 1 else np.arange(n, dtype=np.float32).reshape(n, 1)
    keys = np.arange(n, dtype=np.int64)
    levels = draw_levels(oracle, n, 0.25 if d > 1 else 0.5, 44)
    g = oracle.Graph(d, M, metric)
    g.add_many(keys, X, levels, ef)
    dg, tot = _build(oracle, X, keys, levels, M, ef, metric, window)
    _same_graph(g, dg)
    c = g.counters()
    for name in ("dist_evals", "expansions", "evict_evals", "replenish_evals", "evictions"):
        assert tot[name] == c[name], name
    acc, redo, repaired = dg.add_window_stats()
    assert acc + redo == n            # every node went through the windowed kernel, one way or the other
    assert redo + repaired > 0        # and on a graph this small the run-again paths were exercised
    dg.close()


def test_windowed_add_in_several_calls_shuffled_keys_and_both_paths(oracle):
    """30 000 rows in five calls (an empty graph, a 1-node call, then growing ones), keys in random order: the window is
    accepted for most nodes by the end and rejected for most at the start."""
    rng = np.random.default_rng(9)
    n, d, M, ef = 30000, 32, 8, 20
    X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    keys = rng.permutation(10 * n)[:n].astype(np.int64) - 5000
    levels = draw_levels(oracle, n, 0.25, 45)
    g = oracle.Graph(d, M, 0)
    g.add_many(keys, X, levels, ef)
    dg, tot = _build(oracle, X, keys, levels, M, ef, 0, 12, chunks=(1, 2, 500, 9000))
    _same_graph(g, dg)
    c = g.counters()
    for name in ("dist_evals", "expansions", "evict_evals", "replenish_evals", "evictions"):
        assert tot[name] == c[name], name
    acc, redo, repaired = dg.add_window_stats()
    print("windowed add, 30 000 x 32, window 12: accepted", acc, "run again", redo, "repaired by workers", repaired)
    assert acc + redo == n - 2 and acc > redo      # the two one-node calls ran on the sequential kernel
    # the searches agree afterwards
    Q = rng.uniform(-1, 1, (256, d)).astype(np.float32)
    o_rows, o_d, o_cnt, _ = g.batch_search(Q, 10, ef)
    d_rows, d_d, d_cnt = dg.search_rows(Q, 10, ef)
    assert np.array_equal(o_rows, d_rows) and np.array_equal(o_d.view(np.uint32), d_d.view(np.uint32))
    dg.close()


def test_window_one_and_automatic_build_the_same_graph(oracle):
    """Device against device: the one-node-at-a-time kernel (window 1) and the automatic window on 40 000 rows."""
    rng = np.random.default_rng(10)
    n, d, M, ef = 40000, 16, 12, 32
    X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    keys = np.arange(n, dtype=np.int64)
    levels = draw_levels(oracle, n, 0.25, 46)
    a, ta = _build(oracle, X, keys, levels, M, ef, 1, 1)
    b, tb = _build(oracle, X, keys, levels, M, ef, 1, 0, chunks=(20000,))
    assert a.add_window_stats() == (0, 0, 0)
    # the first 2 048 rows go one at a time, everything after them through windows that grow with the graph
    assert sum(b.add_window_stats()[:2]) == n - 2048
    assert ta == tb
    assert a.Topography() == b.Topography() and list(a.entries()) == list(b.entries())
    for l in range(len(a.Topography())):
        na, ma = a.neighbors(l)
        nb, mb = b.neighbors(l)
        assert np.array_equal(ma, mb) and np.array_equal(na, nb)
    a.close()
    b.close()


def test_windowed_add_after_delete(oracle):
    """Deleted rows in the graph (dead elevators, graph.go:501-505): an Add that fails stops the batch in front of the
    failing node with the reference's error, exactly like the sequential kernel."""
    import hnsw_b200
    rng = np.random.default_rng(11)
    n, d, M, ef = 800, 16, 8, 20
    X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    keys = np.arange(n, dtype=np.int64)
    levels = draw_levels(oracle, n, 0.25, 44)
    g = oracle.Graph(d, M, 0)
    g.add_many(keys, X, levels, ef)
    dg, _ = _build(oracle, X, keys, levels, M, ef, 0, 5)
    victims = [int(k) for k in rng.choice(keys, 100, replace=False)]
    assert dg.BatchDelete(victims) == g.batch_delete(victims)
    m = 300
    Y = rng.uniform(-1, 1, (m, d)).astype(np.float32)
    nl = draw_levels(oracle, m, 0.25, 45) % 3
    went_in = 0
    for i in range(m):   # the oracle one node at a time: the first failure ends the batch
        try:
            g.add(10 ** 6 + i, Y[i], int(nl[i]), ef)
            went_in += 1
        except RuntimeError:
            break
    try:
        dg.add_arrays(np.arange(m, dtype=np.int64) + 10 ** 6, Y, nl, wave_max=0)
        failed = False
    except hnsw_b200.B200Error as e:
        assert "no nodes found in neighborhood search" in str(e)
        failed = True
    assert failed == (went_in < m)
    assert dg.Len() == n - 100 + went_in
    _same_graph(g, dg)
    dg.close()
