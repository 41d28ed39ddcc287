"""Graphs keyed by strings / floats (the reference's K is any cmp.Ordered): the device sees order-preserving int64 codes
(hnsw_b200/keys.py), so every list and every search must equal the oracle's on a graph keyed by the RANKS of the same
keys -- including after the codes were re-spaced on the device (b200hnsw_set_keys)."""
import random

import numpy as np
import pytest

from util import draw_levels

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("kind,gap", [("str", 1 << 32), ("str", 4), ("float", 1 << 32)])
def test_non_integer_keys_match_oracle(oracle, monkeypatch, kind, gap):
    import hnsw_b200
    from hnsw_b200 import keys as K
    monkeypatch.setattr(K, "GAP", gap)
    n, d, M, ef, k = 900, 32, 8, 32, 6
    rng = np.random.default_rng(17)
    X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    rnd = random.Random(3)
    if kind == "str":
        pool = list({"".join(rnd.choice("abcdefghij") for _ in range(rnd.randint(1, 7))) for _ in range(3 * n)})
    else:
        pool = list({rnd.uniform(-1000, 1000) for _ in range(3 * n)} | {0.0, -0.5, 1e-300})
    rnd.shuffle(pool)
    keys = pool[:n]
    rank = {key: r for r, key in enumerate(sorted(keys))}
    levels = draw_levels(oracle, n, 0.25, 46)
    og = oracle.Graph(d, M, oracle.EUCLIDEAN)
    og.add_many(np.asarray([rank[key] for key in keys], dtype=np.int64), X, levels, ef)
    g = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=hnsw_b200.EuclideanDistance)
    for a, b in ((0, 1), (1, 300), (300, 650), (650, n)):  # several batches: new keys land between stored ones
        g.BatchAdd([hnsw_b200.MakeNode(keys[i], X[i]) for i in range(a, b)], levels=levels[a:b])
    assert g.Len() == n and g.Topography() == og.topography()
    for l in range(og.n_layers):
        rows, ptr, nbr = og.dump_layer(l)
        dn, member = g.neighbors(l)
        assert list(np.nonzero(member)[0]) == list(rows)
        for i, r in enumerate(rows):
            assert [int(v) for v in dn[r] if v >= 0] == list(nbr[ptr[i]:ptr[i + 1]]), (l, int(r))
    Q = rng.uniform(-1, 1, (64, d)).astype(np.float32)
    o_rows, _, o_cnt, _ = og.batch_search(Q, k, ef)
    res = g.BatchSearch(list(Q), k)
    for i in range(len(Q)):
        assert [nd.Key for nd in res[i]] == [keys[r] for r in o_rows[i][:o_cnt[i]]]
        one = g.Search(Q[i], k)
        assert [nd.Key for nd in one] == [nd.Key for nd in res[i]]
    vec, ok = g.Lookup(keys[5])
    assert ok and np.array_equal(vec, X[5])
    assert g.Delete(keys[5]) and not g.Delete(keys[5]) and g.Len() == n - 1
    with pytest.raises(TypeError):
        g.BatchAdd([hnsw_b200.MakeNode(12345, X[0])])  # an int key in a str / float graph
