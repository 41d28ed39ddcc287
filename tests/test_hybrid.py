"""N4: the hybrid package's dispatch glue (hybrid/hybrid.go, hybrid/adapter.go) over the device indexes.
CPU part: host logic only (validation, MultiIndexAdapter's merge on stub indexes).  GPU part: tier switch and adapters
through the device."""
import numpy as np
import pytest


class _Stub:
    def __init__(self, keys, dists):
        self.keys, self.dists = keys, dists

    def Search(self, query, k):
        return self.keys[:k], np.asarray(self.dists[:k], dtype=np.float32)

    def Delete(self, key):
        return key in self.keys

    def BatchDelete(self, keys):
        return [k in self.keys for k in keys]

    def Len(self):
        return len(self.keys)


def _merge_spec(per_index, k):
    """adapter.go:257-308 restated: concatenate, insertion sort with strict >, first instance per key, first k."""
    allr = [(key, np.float32(d)) for keys, ds in per_index for key, d in zip(keys, ds)]
    for i in range(1, len(allr)):
        j = i
        while j > 0 and allr[j - 1][1] > allr[j][1]:
            allr[j - 1], allr[j] = allr[j], allr[j - 1]
            j -= 1
    seen, out = set(), []
    for key, d in allr:
        if key not in seen:
            seen.add(key)
            out.append((key, d))
    return out[:k]


def test_config_validation_and_defaults():
    from hnsw_b200 import hybrid
    cfg = hybrid.DefaultIndexConfig()
    assert (cfg.Type, cfg.ExactThreshold, cfg.M, cfg.Ml, cfg.EfSearch) == (hybrid.HybridIndexType, 1000, 16, 0.25, 20)
    assert hybrid.HybridIndex(hybrid.IndexConfig()).config.Type == hybrid.HybridIndexType  # zero value -> default
    for bad, msg in ((dict(M=0), "M must be greater than 0"), (dict(Ml=1.0), "Ml must be between 0 and 1"),
                     (dict(EfSearch=0), "EfSearch must be greater than 0")):
        c = hybrid.DefaultIndexConfig()
        for f, v in bad.items():
            setattr(c, f, v)
        with pytest.raises(ValueError, match=msg):
            hybrid.HybridIndex(c)
    c = hybrid.DefaultIndexConfig()
    c.Type = hybrid.LSHIndexType
    with pytest.raises(NotImplementedError):
        hybrid.HybridIndex(c)


def test_multi_index_merge_on_stubs():
    from hnsw_b200 import hybrid
    nan = float("nan")
    a = _Stub([1, 2, 3, 4], [0.1, 0.2, 0.2, 0.9])
    b = _Stub([3, 9, 2, 7], [0.05, 0.2, nan, 0.3])   # duplicate keys, a tie with index a, a NaN in the middle
    m = hybrid.MultiIndexAdapter([a, b])
    for k in (1, 3, 4, 10):
        keys, d = m.Search(None, k)
        want = _merge_spec([a.Search(None, k), b.Search(None, k)], k)
        assert keys == [w[0] for w in want]
        assert np.array_equal(np.asarray(d).view(np.uint32), np.asarray([w[1] for w in want], np.float32).view(np.uint32))
    assert m.Delete(3) is True and m.Delete(1) is False       # adapter.go:311-319: every index must hold the key
    assert m.BatchDelete([2, 9, 5]) == [True, False, False]
    assert m.Len() == 4 and hybrid.MultiIndexAdapter([]).Len() == 0


@pytest.mark.gpu
def test_hybrid_index_reference_case():
    """hybrid_test.go:10-75 through the device (100 vectors stay in the exact tier)."""
    from hnsw_b200 import hybrid
    rng = np.random.default_rng(3)
    idx = hybrid.HybridIndex(hybrid.DefaultIndexConfig())
    V = rng.random((100, 10), dtype=np.float32)
    for i in range(100):
        idx.Add(i, V[i])
    assert idx.Len() == 100 and idx.GetStats().ExactVectors == 100
    res = idx.Search(rng.random(10, dtype=np.float32), 5)
    assert 0 < len(res) <= 5
    for i in range(10):
        assert idx.Delete(i)
    assert idx.Len() == 90
    assert all(nd.Key >= 10 for nd in idx.Search(V[0], 5))
    idx.Close()


@pytest.mark.gpu
def test_tier_switch_matches_the_plain_indexes():
    import hnsw_b200
    from hnsw_b200 import hybrid
    rng = np.random.default_rng(5)
    n, d, thr = 400, 24, 150
    V = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    cfg = hybrid.DefaultIndexConfig()
    cfg.ExactThreshold, cfg.Distance = thr, hnsw_b200.EuclideanDistance
    idx = hybrid.HybridIndex(cfg)
    idx.hnswIndex.Rng = hnsw_b200.LevelRand(11)
    exact = hnsw_b200.ExactIndex(hnsw_b200.EuclideanDistance)
    graph = hnsw_b200.Graph(M=cfg.M, Ml=cfg.Ml, EfSearch=cfg.EfSearch, Distance=cfg.Distance, Rng=hnsw_b200.LevelRand(11))
    Q = rng.uniform(-1, 1, (20, d)).astype(np.float32)
    for i in range(thr):
        idx.Add(i, V[i])
        exact.Add(i, V[i])
    for q in Q:  # exact tier
        assert [nd.Key for nd in idx.Search(q, 7)] == [nd.Key for nd in exact.Search(q, 7)]
    idx.BatchAdd(list(range(thr, n)), list(V[thr:]))
    graph.BatchAdd([hnsw_b200.Node(i, V[i]) for i in range(thr, n)])
    st = idx.GetStats()
    assert (st.TotalVectors, st.ExactVectors, st.HNSWVectors) == (n, thr, n - thr)
    for q in Q:  # graph tier: only what was added after the switch is reachable (hybrid.go:379-395)
        got = [nd.Key for nd in idx.Search(q, 7)]
        assert got == [nd.Key for nd in graph.Search(q, 7)] and all(k >= thr for k in got)
    batched = idx.BatchSearch(Q, 7)
    assert [[nd.Key for nd in r] for r in batched] == [[nd.Key for nd in idx.Search(q, 7)] for q in Q]
    assert idx.BatchDelete([0, thr, 10 ** 6]) == [True, True, False] and idx.Len() == n - 2
    # the adapters over the same indexes and MultiIndexAdapter's merge
    m = hybrid.MultiIndexAdapter([hybrid.ExactAdapter(exact), hybrid.HNSWAdapter(graph, cfg.Distance)])
    for q in Q[:8]:
        per = [a.Search(q, 6) for a in m.indexes]
        for keys, ds in per:  # adapter distances are distance(query, stored vector) on the device
            for key, dist in zip(keys, ds):
                assert dist == hnsw_b200.EuclideanDistance(q, V[key])
        keys, ds = m.Search(q, 6)
        want = _merge_spec(per, 6)
        assert keys == [w[0] for w in want] and list(ds) == [w[1] for w in want]
    idx.Close()
