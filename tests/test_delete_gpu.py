"""N3 parity: device Graph.Delete / BatchDelete (graph.go:843-895, isolate graph.go:223-235, replenish
graph.go:172-219) vs the oracle, then Search and Add on the graph the deletions leave behind.

The reference's behaviour is kept, quirks included: a deleted node keeps its storage and its own list, a replenish
can link it again one-way, Search can walk through it and return it; a Search whose elevator is a deleted node
returns nothing; an Add in that situation fails with "no nodes found in neighborhood search".
"""
import numpy as np
import pytest

from test_add_parity_gpu import _assert_same_graph, _build_both
from test_search_parity_gpu import _assert_same
from util import draw_levels

pytestmark = pytest.mark.gpu


def _pair(oracle, n, d, M, ef, metric, seed=7):
    rng = np.random.default_rng(seed)
    X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    keys = np.arange(n, dtype=np.int64) * 3 + 1  # keys differ from rows
    levels = draw_levels(oracle, n, 0.25, 44)
    g, dg, _ = _build_both(oracle, X, keys, levels, M, ef, metric)
    return rng, X, keys, levels, g, dg


@pytest.mark.parametrize("n,d,M,ef,metric", [
    (900, 32, 8, 20, 0),
    (700, 128, 16, 64, 0),
    (400, 96, 16, 20, 1),
    (500, 8, 6, 20, 0),
])
def test_batch_delete_matches_oracle(oracle, n, d, M, ef, metric):
    rng, X, keys, levels, g, dg = _pair(oracle, n, d, M, ef, metric)
    victims = [int(k) for k in rng.choice(keys, n // 8, replace=False)]
    batch = victims[:10] + [10 ** 9, victims[0]] + victims[10:]  # unknown key, repeated key
    want = g.batch_delete(batch)
    got = dg.BatchDelete(batch)
    assert got == want
    assert sum(got) == len(victims) and got[10] is False and got[11] is False
    assert dg.Len() == len(g) == n - len(victims)
    _assert_same_graph(g, dg, M)
    for k in victims[:5]:
        assert dg.Lookup(k) == (None, False)
    # searches on what is left: empty answers (deleted elevator) and deleted rows among the results included
    Q = rng.uniform(-1, 1, (300, d)).astype(np.float32)
    _assert_same(oracle, g, dg, Q, 10, ef)
    _assert_same(oracle, g, dg, X[:64], 5, ef)
    # one at a time
    for k in victims[:3] + [int(keys[5]) if int(keys[5]) not in victims else int(keys[6])]:
        assert dg.Delete(k) == g.delete(k)
    _assert_same_graph(g, dg, M)
    _assert_same(oracle, g, dg, Q, 10, ef)


def test_add_after_delete_matches_oracle(oracle):
    import hnsw_b200
    n, d, M, ef = 800, 16, 8, 20
    rng, X, keys, levels, g, dg = _pair(oracle, n, d, M, ef, 0)
    victims = [int(k) for k in rng.choice(keys, 100, replace=False)]
    assert dg.BatchDelete(victims) == g.batch_delete(victims)
    new_levels = draw_levels(oracle, 200, 0.25, 45) % 3
    ok = bad = 0
    for i in range(200):
        key = 10 ** 6 + i
        v = rng.uniform(-1, 1, d).astype(np.float32)
        try:
            g.add(key, v, int(new_levels[i]), ef)
            oracle_ok = True
        except RuntimeError:
            oracle_ok = False
        try:
            dg.Add(hnsw_b200.Node(key, v), levels=[int(new_levels[i])])
            dev_ok = True
        except hnsw_b200.B200Error as e:
            assert "no nodes found in neighborhood search" in str(e)
            dev_ok = False
        assert dev_ok == oracle_ok, i
        ok += dev_ok
        bad += not dev_ok
    assert ok > 0
    assert dg.Len() == len(g)
    _assert_same_graph(g, dg, M)
    Q = rng.uniform(-1, 1, (200, d)).astype(np.float32)
    _assert_same(oracle, g, dg, Q, 10, ef)
    # a batch that fails in the middle keeps the nodes in front of the failing one, like the per-node loop
    if bad:
        assert dg.Len() == n - 100 + ok
    # replacing a key in place and re-using a deleted key: tests/test_dropin_gpu.py


def test_emptied_top_layers(oracle):
    import hnsw_b200
    n, d, M, ef = 600, 16, 8, 20
    rng, X, keys, levels, g, dg = _pair(oracle, n, d, M, ef, 0)
    top = int(levels.max())
    assert top >= 2
    victims = [int(keys[i]) for i in range(n) if levels[i] >= top - 1]  # empties the two highest layers
    assert dg.BatchDelete(victims) == g.batch_delete(victims)
    assert dg.Topography() == g.topography()
    assert dg.Topography()[top] == 0 and dg.Topography()[top - 1] == 0
    assert list(dg.entries()) == list(g.entries())
    _assert_same_graph(g, dg, M)
    Q = rng.uniform(-1, 1, (128, d)).astype(np.float32)
    _assert_same(oracle, g, dg, Q, 10, ef)
    # a node drawn below an emptied layer: the reference's behaviour there is not reproduced, both refuse
    v = rng.uniform(-1, 1, d).astype(np.float32)
    with pytest.raises(RuntimeError):
        g.add(10 ** 6, v, 0, ef)
    with pytest.raises(hnsw_b200.B200Error):
        dg.Add(hnsw_b200.Node(10 ** 6, v), levels=[0])
    # a node that reaches the emptied layers becomes their entry point
    g.add(10 ** 6 + 1, v, top, ef)
    dg.Add(hnsw_b200.Node(10 ** 6 + 1, v), levels=[top])
    assert dg.Topography() == g.topography() and list(dg.entries()) == list(g.entries())
    _assert_same_graph(g, dg, M)
    _assert_same(oracle, g, dg, Q, 10, ef)


def test_batch_delete_reference_cases():
    """batch_delete_test.go:10-105 through the device."""
    import hnsw_b200
    g = hnsw_b200.Graph(M=16, Ml=0.25, EfSearch=20, Distance=hnsw_b200.EuclideanDistance)
    for i in range(1, 11):
        g.Add(hnsw_b200.Node(i, [float(i), float(i), float(i)]))
    assert g.BatchDelete([1, 3, 5]) == [True, True, True]
    for k in (1, 3, 5):
        assert g.Lookup(k)[1] is False
    assert g.Len() == 7
    assert g.BatchDelete([11, 12, 13]) == [False, False, False]
    assert g.Len() == 7
    assert g.BatchDelete([2, 15, 4, 20]) == [True, False, True, False]
    assert g.Lookup(2)[1] is False and g.Lookup(4)[1] is False
    assert g.Len() == 5
    assert g.BatchDelete([]) == []
    assert g.BatchDelete([6, 7, 8, 9, 10]) == [True] * 5
    assert g.Len() == 0
    assert g.Search([1.0, 1.0, 1.0], 3) == []
    assert hnsw_b200.Graph().BatchDelete([1, 2]) == [False, False]  # graph.go:875-878: no layers yet


def test_batch_delete_reference_cases_cosine_and_three_batches():
    """batch_delete_test.go as the reference runs it: NewGraphWithConfig(16, 0.25, 20, CosineDistance) over the parallel
    vectors (i, i, i) -- every cosine distance is 0 up to rounding, i.e. ties everywhere -- first the five cases of
    TestBatchDelete (lines 10-105), then TestBatchDeleteConcurrency (107-158: 100 nodes, three batches of 20)."""
    import hnsw_b200
    g = hnsw_b200.NewGraphWithConfig(16, 0.25, 20, hnsw_b200.CosineDistance)
    for i in range(1, 11):
        g.Add(hnsw_b200.MakeNode(i, [float(i), float(i), float(i)]))
    assert g.Len() == 10
    assert g.BatchDelete([1, 3, 5]) == [True, True, True] and g.Len() == 7
    assert all(g.Lookup(k)[1] is False for k in (1, 3, 5)) and all(g.Lookup(k)[1] is True for k in (2, 4, 6, 7, 8, 9, 10))
    assert g.BatchDelete([11, 12, 13]) == [False, False, False] and g.Len() == 7
    assert g.BatchDelete([2, 15, 4, 20]) == [True, False, True, False] and g.Len() == 5
    assert all(g.Lookup(k)[1] is False for k in (2, 4)) and all(g.Lookup(k)[1] is True for k in (6, 7, 8, 9, 10))
    assert g.BatchDelete([]) == [] and g.Len() == 5
    assert g.BatchDelete([6, 7, 8, 9, 10]) == [True] * 5 and g.Len() == 0
    g = hnsw_b200.NewGraphWithConfig(16, 0.25, 20, hnsw_b200.CosineDistance)
    for i in range(1, 101):
        g.Add(hnsw_b200.MakeNode(i, [float(i), float(i), float(i)]))
    assert g.Len() == 100
    b1, b2, b3 = list(range(1, 21)), list(range(41, 61)), list(range(81, 101))
    assert g.BatchDelete(b1) == [True] * 20 and g.BatchDelete(b2) == [True] * 20 and g.BatchDelete(b3) == [True] * 20
    assert g.Len() == 40
    gone = set(b1) | set(b2) | set(b3)
    for i in range(1, 101):
        assert g.Lookup(i)[1] is (i not in gone)
