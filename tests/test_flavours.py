"""The flavour divergence report (oracle/flavours.py): north_star allows ID divergence "only at float ties" and wants it
reported.  The kernels equal the oracle's CANONICAL summation order bit for bit; vek32's own orders (the scalar Go loop,
the AVX2 shape) are the oracle's SEQUENTIAL and SIMD flavours.  CPU only."""
import numpy as np

from util import draw_levels, flatten_oracle


def test_uniform_data_has_no_divergence_between_summation_orders(oracle):
    from oracle import flavours
    rng = np.random.default_rng(42)
    n, d, M, ef = 6000, 128, 16, 64
    X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    levels = draw_levels(oracle, n, 0.25, 44)
    keys = np.arange(n, dtype=np.int64)
    g = oracle.Graph(d, M, oracle.EUCLIDEAN, oracle.CANONICAL)
    g.add_many(keys, X, levels, ef)
    Q = rng.uniform(-1, 1, (1500, d)).astype(np.float32)
    rep = flavours.search_report(flatten_oracle(g), Q, 10, ef, oracle.EUCLIDEAN, M, d, n_threads=4)
    assert set(rep) == {"canonical_vs_sequential", "canonical_vs_simd", "sequential_vs_simd"}
    for pair, r in rep.items():
        assert r["non_tie"] == 0 and r["trace_truncated"] == 0, (pair, r)
        assert r["queries_with_id_difference"] == r["tie_class"], (pair, r)
        assert r["queries_with_id_difference"] <= 2, (pair, r)          # continuous 128-d data: ties are rare
        assert 0 < r["max_relative_distance_error"] < 1e-5, (pair, r)   # north_star's distance tolerance
    b = flavours.build_report(X[:2500], keys[:2500], levels[:2500], M, ef, oracle.EUCLIDEAN)
    for pair, r in b.items():
        assert r["topography_equal"] and r["entries_equal"]
        assert r["lists_different"] <= 0.002 * r["lists"], (pair, r)


def test_near_duplicates_diverge_only_at_ties(oracle):
    """150 distinct directions, 4000 rows that differ from them by a few ulp: the distances a walk compares are tied to
    within the rounding of a summation order, so the flavours DO return different rows -- and every such query has its
    divergence explained by a pair of rows the two flavours order differently, within 1e-5 relative."""
    from oracle import flavours
    rng = np.random.default_rng(5)
    n, d, M, ef = 4000, 48, 12, 32
    base = rng.uniform(-1, 1, (150, d)).astype(np.float32)
    X = (base[rng.integers(0, 150, n)] * (1 + rng.normal(0, 2e-7, (n, 1)))).astype(np.float32)
    levels = draw_levels(oracle, n, 0.25, 44)
    keys = np.arange(n, dtype=np.int64)
    for metric in (oracle.EUCLIDEAN, oracle.COSINE):
        g = oracle.Graph(d, M, metric, oracle.CANONICAL)
        g.add_many(keys, X, levels, ef)
        Q = rng.uniform(-1, 1, (600, d)).astype(np.float32)
        rep = flavours.search_report(flatten_oracle(g), Q, 10, ef, metric, M, d, n_threads=4)
        for pair, r in rep.items():
            assert r["queries_with_id_difference"] > 50, (pair, r)
            assert r["non_tie"] == 0, (pair, r)
            assert r["tie_class"] == r["queries_with_id_difference"], (pair, r)
            assert r["max_relative_gap_of_inverted_pairs"] < 1e-5, (pair, r)


def test_identical_traces_compare_equal(oracle):
    from oracle import flavours
    rng = np.random.default_rng(1)
    X = rng.uniform(-1, 1, (500, 16)).astype(np.float32)
    levels = draw_levels(oracle, 500, 0.25, 3)
    g = oracle.Graph(16, 8, oracle.EUCLIDEAN, oracle.CANONICAL)
    g.add_many(np.arange(500, dtype=np.int64), X, levels, 20)
    Q = rng.uniform(-1, 1, (50, 16)).astype(np.float32)
    a = g.batch_search_trace(Q, 5, 20)
    b = g.batch_search_trace(Q, 5, 20, n_threads=3)
    r = flavours.compare_search(a, b)
    assert r["queries_with_id_difference"] == 0 and r["max_relative_distance_error"] == 0.0
    rows, d, cnt, _ = g.batch_search(Q, 5, 20)
    assert np.array_equal(rows, a[0]) and np.array_equal(cnt, a[2])
    assert (a[5] > 0).all() and (a[3][np.arange(50), 0] >= 0).all()
