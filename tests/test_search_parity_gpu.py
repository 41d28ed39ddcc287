"""K1 parity: device Graph.Search / BatchSearch vs the oracle on identical graphs and queries.

Bit-exact bar: rows, their heap-array order, counts AND distances (the device reproduces the oracle's canonical
summation order), plus equal E/H counters ("device counters must equal them").
"""
import numpy as np
import pytest

from util import build_oracle_graph, flatten_oracle

pytestmark = pytest.mark.gpu


def _device_graph(oracle, g, M, ef, metric):
    import hnsw_b200
    dist = hnsw_b200.EuclideanDistance if metric == oracle.EUCLIDEAN else hnsw_b200.CosineDistance
    dg = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=dist)
    dg.load_flat(*flatten_oracle(g))
    return dg


def _assert_same(oracle, g, dg, Q, k, ef):
    g.reset_counters()
    o_rows, o_d, o_cnt, o_ctr = g.batch_search(Q, k, ef)
    d_rows, d_d, d_cnt = dg.search_rows(Q, k, ef)
    assert np.array_equal(o_cnt, d_cnt)
    assert np.array_equal(o_rows, d_rows)
    assert np.array_equal(o_d.view(np.uint32), d_d.view(np.uint32))  # bit-identical distances (NaN padding too)
    st = dg.last_stats
    assert st["dist_evals"] == o_ctr["dist_evals"]
    assert st["expansions"] == o_ctr["expansions"]
    assert st["nbr_reads"] == o_ctr["nbr_reads"]


@pytest.mark.parametrize("n,d,M,ef,k,metric", [
    (2000, 128, 16, 20, 10, 0),   # BASELINE config C1 shape (graph_benchmark_test.go), reduced N
    (2000, 128, 16, 64, 10, 0),   # C2 parameters
    (1500, 96, 16, 64, 10, 0),    # C4 dimension (24 chunks: partial warp)
    (800, 768, 32, 64, 10, 1),    # C3 shape: cosine, M=32, 6 chunks per lane
    (600, 1024, 16, 20, 5, 1),    # C5 dimension
    (500, 3, 6, 20, 4, 0),        # tiny dims like the reference's unit tests
    (300, 1, 6, 20, 4, 0),        # 1-D line graph (graph_test.go:86-133 shape)
    (400, 200, 48, 30, 64, 0),    # M > 32 (two list rounds), k > ef
    (700, 2048, 64, 1024, 256, 1),  # every limit of the C ABI at once: dim 2048, MAX_M, MAX_EF, MAX_K (cosine)
    (900, 1536, 64, 512, 200, 0),   # the same neighbourhood under L2
    (1200, 100, 64, 1024, 256, 0),  # dim not a multiple of 32, limits of M / ef / k
])
def test_batch_search_matches_oracle(oracle, n, d, M, ef, k, metric):
    rng = np.random.default_rng(42)
    X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    g, _, _ = build_oracle_graph(oracle, X, M, 0.25, ef, metric)
    dg = _device_graph(oracle, g, M, ef, metric)
    Q = np.random.default_rng(43).uniform(-1, 1, (257, d)).astype(np.float32)
    _assert_same(oracle, g, dg, Q, k, ef)
    # stored points as queries: exact-zero distances and ties
    _assert_same(oracle, g, dg, X[:64], k, ef)


def test_layer_search_kat_on_device(oracle):
    # graph_test.go:27-74 through the device: rows [4, 3] == Keys [5, 3]
    import hnsw_b200
    g = oracle.Graph(1, 6, oracle.EUCLIDEAN)
    for mk in range(6):
        g.put_node(mk, [float(mk)])
    for nb in (1, 2, 3):
        g.link(0, 0, nb)
    for nb in (4, 5):
        g.link(0, 3, nb)
    dg = hnsw_b200.Graph(M=6, Ml=0.5, EfSearch=4, Distance=hnsw_b200.EuclideanDistance)
    dg.load_flat(*flatten_oracle(g))
    rows, dists, cnt = dg.search_rows([[4.0]], 2, 4)
    assert cnt[0] == 2 and list(rows[0]) == [4, 3] and list(dists[0]) == [0.0, 1.0]


def test_entry_override_and_k_larger_than_graph(oracle):
    rng = np.random.default_rng(1)
    X = rng.uniform(-1, 1, (50, 16)).astype(np.float32)
    g, _, _ = build_oracle_graph(oracle, X, 8, 0.25, 20, 0)
    dg = _device_graph(oracle, g, 8, 20, 0)
    Q = rng.uniform(-1, 1, (16, 16)).astype(np.float32)
    _assert_same(oracle, g, dg, Q, 100, 20)  # k > N: fewer than k results
    # explicit entries (F5: the reference's entry is a random map element)
    ent = g.entries().copy()
    top = g.n_layers - 1
    members = [r for r in range(50) if g.level(r) >= top]
    ent[top] = members[-1]
    for q in Q:
        o_rows, o_d = g.search(q, 5, 20, entries=ent)
        d_rows, d_d, d_cnt = dg.search_rows(q, 5, 20, entries=ent)
        assert list(o_rows) == list(d_rows[0][:d_cnt[0]])


def test_distance_kats_on_device():
    # distance_test.go:9-31 through the device arithmetic
    import hnsw_b200
    assert hnsw_b200.EuclideanDistance([1, 2, 3], [4, 5, 6]) == np.float32(5.196152)
    assert abs(hnsw_b200.CosineDistance([1, 1, 1], [0.8, 0.8, 0.8])) <= 1e-6
    assert abs(hnsw_b200.CosineDistance([1, 0], [0, 1]) - 1) <= 1e-6
    assert abs(hnsw_b200.CosineDistance([1, 0], [1, 0])) <= 1e-6


def test_distance_bits_match_oracle(oracle):
    from hnsw_b200.distance import distance_batch
    rng = np.random.default_rng(9)
    for d in (1, 3, 96, 128, 130, 768, 1024, 1536):
        a = rng.standard_normal((64, d)).astype(np.float32)
        b = rng.standard_normal((64, d)).astype(np.float32)
        for metric in (0, 1):
            dev = distance_batch(metric, a, b)
            ref = np.array([oracle.distance(metric, a[i], b[i]) for i in range(64)], dtype=np.float32)
            assert np.array_equal(dev.view(np.uint32), ref.view(np.uint32)), (d, metric)


def test_empty_graph_and_errors():
    import hnsw_b200
    g = hnsw_b200.NewGraph()
    assert g.Search([0.5, 0.5], 1) is None  # graph.go:554-556
    with pytest.raises(ValueError, match="k must be greater than 0"):
        g.Search([0.5], 0)


def test_merge_kernel_matches_spec():
    # K4 vs its numpy specification (tests/merge_ref.py): ties -> lower shard, then list position; NaN last
    import torch
    from hnsw_b200 import sharded
    from merge_ref import merge_topk_ref
    rng = np.random.default_rng(3)
    for G, B, k in ((2, 50, 10), (8, 33, 10), (4, 20, 64), (3, 7, 1), (4, 29, 10), (2, 9, 64), (8, 17, 16), (16, 5, 8)):
        d = np.round(rng.uniform(0, 4, (G, B, k)), 1).astype(np.float32)  # coarse values: many ties
        d.sort(axis=2)
        d[rng.uniform(size=d.shape) < 0.02] = np.nan
        i = rng.integers(0, 1 << 40, (G, B, k)).astype(np.int64)
        c = rng.integers(0, k + 1, (G, B)).astype(np.int32)
        md, mi, mc = sharded.merge_topk_device(torch.from_numpy(d).cuda(), torch.from_numpy(i).cuda(),
                                               torch.from_numpy(c).cuda(), k)
        torch.cuda.synchronize()
        rd, ri, rc = merge_topk_ref(d, i, c, k)
        assert np.array_equal(mc.cpu().numpy(), rc)
        assert np.array_equal(mi.cpu().numpy(), ri)
        assert np.array_equal(md.cpu().numpy(), rd, equal_nan=True)


@pytest.mark.gpu
def test_packed_shard_merge_matches_spec():
    # the same K4 reading the search kernel's own output layout (rows | dists | counts per shard), ids = g*stride + row
    import torch
    from hnsw_b200 import sharded
    from merge_ref import merge_topk_ref
    rng = np.random.default_rng(4)
    for G, B, k, stride_ids in ((2, 50, 10, 1000), (8, 33, 10, 1 << 20), (4, 20, 64, 77), (1, 9, 3, 5)):
        d = np.round(rng.uniform(0, 4, (G, B, k)), 1).astype(np.float32)
        d.sort(axis=2)
        d[rng.uniform(size=d.shape) < 0.02] = np.nan
        rows = rng.integers(0, stride_ids, (G, B, k)).astype(np.int32)
        c = rng.integers(0, k + 1, (G, B)).astype(np.int32)
        pr = sharded.PackedResults(B, k, G, torch.device("cuda", 0))
        g = pr.gathered.view(G, pr.stride)
        g[:, :B * k] = torch.from_numpy(rows.reshape(G, -1)).cuda()
        g[:, B * k:2 * B * k] = torch.from_numpy(d.view(np.int32).reshape(G, -1)).cuda()
        g[:, 2 * B * k:] = torch.from_numpy(c).cuda()
        md, mi, mc = pr.merge(stride_ids)
        torch.cuda.synchronize()
        ids = rows.astype(np.int64) + (np.arange(G, dtype=np.int64) * stride_ids)[:, None, None]
        rd, ri, rc = merge_topk_ref(d, ids, c, k)
        assert np.array_equal(mc.cpu().numpy(), rc)
        assert np.array_equal(mi.cpu().numpy(), ri)
        assert np.array_equal(md.cpu().numpy(), rd, equal_nan=True)


def test_limits_are_refused_loudly(oracle):
    """Past the limits of the C ABI (include/b200hnsw.h: MAX_K 256, MAX_EF 1024, MAX_M 64, dim 2048) calls fail with a
    message; nothing is clamped silently."""
    import hnsw_b200
    from hnsw_b200 import _lib
    rng = np.random.default_rng(3)
    X = rng.uniform(-1, 1, (300, 16)).astype(np.float32)
    g, _, _ = build_oracle_graph(oracle, X, 8, 0.25, 20, 0)
    dg = _device_graph(oracle, g, 8, 20, 0)
    Q = X[:4]
    with pytest.raises(_lib.B200Error, match="k must be greater than 0 and <= 256"):
        dg.search_rows(Q, 257, 20)
    with pytest.raises(_lib.B200Error, match="EfSearch must be greater than 0 and <= 1024"):
        dg.search_rows(Q, 10, 1025)
    rows, _, cnt = dg.search_rows(Q, 256, 1024)              # the limits themselves are fine
    assert (cnt > 0).all() and (rows[:, 0] == np.arange(4)).all()
    big = hnsw_b200.Graph(M=65, Ml=0.25, EfSearch=20, Distance=hnsw_b200.EuclideanDistance)
    with pytest.raises(Exception):
        big.add_arrays([0, 1], np.zeros((2, 8), np.float32), levels=[0, 0])
    wide = hnsw_b200.Graph(M=8, Ml=0.25, EfSearch=20, Distance=hnsw_b200.EuclideanDistance)
    with pytest.raises(Exception):
        wide.add_arrays([0, 1], np.zeros((2, 2049), np.float32), levels=[0, 0])
