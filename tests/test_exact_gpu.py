"""K3 parity: device ExactIndex (hnsw-extensions/hybrid/exact.go) vs the oracle's restatement and a float64 brute force.

Bar: rows identical to the oracle (ties resolve to the lower row on both sides), distances bit-identical (canonical
order), counts = min(k, Len()).  Both device paths (fp32 scan; tensor-core candidates + fp32 re-evaluation) are held to
the same bar via B200EXACT_PATH.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _check(oracle, X, Q, k, metric, path, expect_no_fallback=True):
    import hnsw_b200
    os.environ["B200EXACT_PATH"] = path
    try:
        dist = hnsw_b200.EuclideanDistance if metric == 0 else hnsw_b200.CosineDistance
        ix = hnsw_b200.NewExactIndex(dist)
        half = len(X) // 2
        ix.BatchAdd(list(range(half)), X[:half])
        ix.BatchAdd(list(range(half, len(X))), X[half:])
        rows, dists, counts = ix.search_rows(Q, k)
        st = ix.last_stats
        # the path that was asked for is the path that answered (launches: 1 = scan, 2 = tensor), and on the tensor
        # path no query fell back to the scan (expansions = queries redone because their candidate list overflowed
        # or came up short): a filter bug that sent every query to the scan would otherwise stay green
        assert st["launches"] == {"scan": 1, "tensor": 2}[path], st
        if path == "tensor" and expect_no_fallback:
            assert st["expansions"] == 0, st
            assert k <= st["dist_evals"] / len(Q) < 4096, st   # candidates per query: a real filter, not everything
        o_idx, o_d = oracle.exact_batch_search(X, Q, k, metric, n_threads=8)
        kk = min(k, len(X))
        assert (counts == kk).all()
        assert np.array_equal(rows[:, :kk], o_idx.astype(np.int64))
        assert np.array_equal(dists[:, :kk].view(np.uint32), o_d.view(np.uint32))
        ix.Close()
    finally:
        os.environ.pop("B200EXACT_PATH", None)


@pytest.mark.parametrize("n,d,k,metric", [
    (1000, 128, 10, 0),     # the reference's ExactThreshold (hybrid.go:101) and benchmark dimension
    (5000, 96, 10, 0),      # several scan chunks
    (3000, 1024, 10, 1),    # C5 shape: cosine, dim 1024
    (700, 3, 5, 1),
    (2500, 40, 100, 0),     # large k: two merge levels
    (7, 16, 10, 0),         # k > N (exact.go:97-100)
])
def test_exact_scan_matches_oracle(oracle, n, d, k, metric):
    rng = np.random.default_rng(5)
    X = rng.standard_normal((n, d)).astype(np.float32)
    Q = rng.standard_normal((33, d)).astype(np.float32)
    _check(oracle, X, Q, k, metric, "scan")
    _check(oracle, X, X[:8], k, metric, "scan")   # stored points as queries: zero distances


def test_exact_api_like_reference(oracle):
    # hybrid_test.go:66-114 shape: add, search, delete; plus value checks the reference's tests do not make
    import hnsw_b200
    ix = hnsw_b200.NewExactIndex(hnsw_b200.EuclideanDistance)
    assert ix.Search([0.0, 0.0], 3) == [] and ix.Len() == 0
    pts = {10: [1, 0], 20: [0, 1], 30: [-1, 0], 40: [0, -1], 50: [2, 0]}
    for kx, v in pts.items():
        ix.Add(kx, v)
    assert ix.Len() == 5
    out = ix.Search([0, 0], 5)
    assert [nd.Key for nd in out] == [10, 20, 30, 40, 50]       # ties keep insertion order (exact.go:88-94)
    assert ix.Delete(20) and not ix.Delete(20) and ix.Len() == 4
    assert [nd.Key for nd in ix.Search([0, 0], 10)] == [10, 30, 40, 50]
    ix.Add(50, [0.1, 0])                                          # overwrite (exact.go:36)
    out = ix.Search([0, 0], 1)
    assert out[0].Key == 50 and list(out[0].Value) == [np.float32(0.1), 0] and ix.Len() == 4
    assert ix.BatchDelete([10, 11, 30]) == [True, False, True]
    with pytest.raises(ValueError, match="number of keys"):
        ix.BatchAdd([1, 2], [[0, 0]])
    ix.Close()


@pytest.mark.parametrize("n,d,k,metric,B", [
    (40000, 128, 10, 0, 300),
    (70000, 1024, 10, 1, 200),   # C5 shape
    (50000, 96, 32, 0, 130),
    (33000, 200, 10, 1, 257),    # dim not a multiple of 64
    (60000, 320, 10, 0, 150),    # Euclidean with the query tile NOT resident: the fold block travels in the main ring
])
def test_exact_tensor_path_matches_oracle(oracle, n, d, k, metric, B):
    rng = np.random.default_rng(6)
    X = rng.standard_normal((n, d)).astype(np.float32)
    if metric == 0:
        X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    Q = rng.standard_normal((B, d)).astype(np.float32)
    Q[:5] = X[100:105]  # exact hits
    _check(oracle, X, Q, k, metric, "tensor")


def test_exact_tensor_path_query_scales_and_deletes(oracle):
    """Euclidean tensor path: the per-row term comes out of the GEMM (fold block, exact_gemm.cu), which needs the query's
    fp16 scale within a window of the rows'.  Queries far outside it (x 1e4: moved into the window; x 1e5: given an infinite
    margin and answered by the scan path; x 1e-3: a smaller scale) and rows of very different norms must still
    give the oracle's rows and bits.  Deleted rows (the nearest ones among them) never come back."""
    import hnsw_b200
    os.environ["B200EXACT_PATH"] = "tensor"
    try:
        rng = np.random.default_rng(16)
        n, d, k, B = 40000, 128, 10, 96
        X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
        X[7] *= 4.0                                      # one row sets the scale of the fp16 copy
        X[11:200] *= 0.01
        Q = rng.standard_normal((B, d)).astype(np.float32)
        Q[:4] *= 1e5                                     # outside the window: scan path
        Q[4:8] *= 1e4                                    # moved into the window
        Q[8:16] *= 1e-3
        Q[16:20] = X[300:304]
        ix = hnsw_b200.NewExactIndex(hnsw_b200.EuclideanDistance)
        ix.BatchAdd(list(range(n)), X)
        rows, dists, counts = ix.search_rows(Q, k)
        st = ix.last_stats
        assert st["launches"] == 2, st
        assert 4 <= st["expansions"] <= 12, st          # the x 1e5 queries are answered by the scan, hardly any other
        o_idx, o_d = oracle.exact_batch_search(X, Q, k, 0, n_threads=8)
        assert np.array_equal(rows, o_idx.astype(np.int64))
        assert np.array_equal(dists.view(np.uint32), o_d.view(np.uint32))
        # delete the nearest rows of the ordinary queries and 500 others
        gone = np.unique(np.concatenate([rows[20:, :3].ravel(), rng.choice(n, 500, replace=False)]))
        for key in gone:
            assert ix.Delete(int(key))
        keep = np.ones(n, bool)
        keep[gone] = False
        ids = np.nonzero(keep)[0]
        rows2, dists2, counts2 = ix.search_rows(Q[16:], k)
        assert ix.last_stats["launches"] == 2
        o_idx2, o_d2 = oracle.exact_batch_search(X[keep], Q[16:], k, 0, n_threads=8)
        assert np.array_equal(rows2, ids[o_idx2].astype(np.int64))
        assert np.array_equal(dists2.view(np.uint32), o_d2.view(np.uint32))
        assert not np.isin(rows2, gone).any()
        ix.Close()
    finally:
        os.environ.pop("B200EXACT_PATH", None)
