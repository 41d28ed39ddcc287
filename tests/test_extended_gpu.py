"""K1x parity: B200HNSW_MODE_EXTENDED (textbook HNSW layer search; NOT the reference's termination rule) on the device
vs its CPU statement, the oracle's ORA_MODE_EXTENDED.  Same bar as the reference mode: rows, order, counts, distance
bits and the E / H / neighbour-read counters.  The two modes are never mixed: the reference mode's results on the same
graph are checked to be untouched."""
import numpy as np
import pytest

from util import build_oracle_graph, flatten_oracle, same_float_bits, recall_at_k

pytestmark = pytest.mark.gpu

EXT = 1


def _device_graph(oracle, g, M, ef, metric):
    import hnsw_b200
    dist = hnsw_b200.EuclideanDistance if metric == oracle.EUCLIDEAN else hnsw_b200.CosineDistance
    dg = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=dist)
    dg.load_flat(*flatten_oracle(g))
    return dg


def _same(g, dg, Q, k, ef, mode):
    g.reset_counters()
    o_rows, o_d, o_cnt, o_ctr = g.batch_search(Q, k, ef, mode=mode)
    d_rows, d_d, d_cnt = dg.search_rows(Q, k, ef, mode=mode)
    assert np.array_equal(o_cnt, d_cnt)
    assert np.array_equal(o_rows, d_rows)
    assert same_float_bits(o_d, d_d)
    st = dg.last_stats
    assert (st["dist_evals"], st["expansions"], st["nbr_reads"]) == (
        o_ctr["dist_evals"], o_ctr["expansions"], o_ctr["nbr_reads"])
    return d_rows


@pytest.mark.parametrize("n,d,M,ef,k,metric", [
    (3000, 128, 16, 64, 10, 0),   # C2 parameters (G = 8, 4 chunks per lane)
    (3000, 128, 16, 16, 10, 0),   # small ef
    (2000, 96, 16, 64, 10, 0),    # C4 dimension
    (1200, 768, 32, 64, 10, 1),   # C3 shape: cosine, M = 32, a full warp per row
    (800, 1024, 16, 40, 5, 1),    # C5 dimension
    (900, 64, 8, 32, 3, 0),
    (700, 32, 8, 200, 100, 1),    # large ef and k: W spans several 32-entry chunks
    (500, 3, 6, 20, 4, 0),        # tiny dims like the reference's unit tests
    (400, 200, 48, 30, 64, 0),    # M > 32 (two list rounds), k > ef
    (300, 1, 6, 20, 4, 0),        # 1-D: many exact ties in distance -> the row tie-break decides
])
def test_extended_search_matches_oracle(oracle, n, d, M, ef, k, metric):
    rng = np.random.default_rng(42)
    X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    if d == 1:
        X = np.round(X * 20) / 20  # duplicates
    g, _, _ = build_oracle_graph(oracle, X, M, 0.25, ef, metric)
    dg = _device_graph(oracle, g, M, ef, metric)
    Q = np.random.default_rng(43).uniform(-1, 1, (203, d)).astype(np.float32)
    _same(g, dg, Q, k, ef, EXT)
    _same(g, dg, X[:40], k, ef, EXT)     # stored points: zero distances, ties
    _same(g, dg, Q, k, ef, 0)            # the reference mode on the same handle is what it was


def test_extended_mode_finds_more_than_the_reference_rule(oracle):
    """What the mode is for: on clustered data the reference's rule stops at the first non-improving hop."""
    rng = np.random.default_rng(7)
    n, d = 6000, 32
    cent = rng.uniform(-1, 1, (20, d)).astype(np.float32)
    X = (cent[rng.integers(0, 20, n)] + rng.normal(0, 0.1, (n, d))).astype(np.float32)
    Q = (cent[rng.integers(0, 20, 200)] + rng.normal(0, 0.1, (200, d))).astype(np.float32)
    g, _, _ = build_oracle_graph(oracle, X, 16, 0.25, 64, 0)
    dg = _device_graph(oracle, g, 16, 64, 0)
    d2 = ((Q[:, None, :] - X[None, :, :]) ** 2).sum(-1)
    truth = np.argsort(d2, axis=1)[:, :10]
    r_ref = _same(g, dg, Q, 10, 64, 0)
    r_ext = _same(g, dg, Q, 10, 64, EXT)
    assert recall_at_k(r_ext, truth) > recall_at_k(r_ref, truth) + 0.1
    # ascending order, no duplicates
    _, dd, cnt = dg.search_rows(Q, 10, 64, mode=EXT)
    assert (np.diff(dd, axis=1) >= 0).all()
    assert all(len(set(r)) == len(r) for r in r_ext)


def test_extended_mode_nan_and_deleted(oracle):
    import hnsw_b200
    rng = np.random.default_rng(9)
    n, d, M, ef = 1500, 24, 12, 48
    X = rng.standard_normal((n, d)).astype(np.float32)
    X[rng.choice(n, 100, replace=False)] = 0.0   # cosine: NaN distances never enter W
    g, keys, _ = build_oracle_graph(oracle, X, M, 0.25, ef, oracle.COSINE)
    dg = _device_graph(oracle, g, M, ef, oracle.COSINE)
    Q = rng.standard_normal((100, d)).astype(np.float32)
    Q[::9] = 0.0
    _same(g, dg, Q, 10, ef, EXT)
    # Delete: a traversal whose elevator is a deleted row answers with nothing (graph.go:573-593), in this mode too
    dead = [int(k) for k in rng.choice(n, 200, replace=False)]
    assert g.batch_delete(dead) == dg.BatchDelete(dead)
    _same(g, dg, Q, 10, ef, EXT)
    _same(g, dg, Q, 10, ef, 0)


def test_extended_visited_table_grows(oracle):
    """ef = 512 on a small dense graph visits far more nodes than the default table of the reference mode."""
    rng = np.random.default_rng(10)
    X = rng.uniform(-1, 1, (5000, 16)).astype(np.float32)
    g, _, _ = build_oracle_graph(oracle, X, 16, 0.25, 64, 0)
    dg = _device_graph(oracle, g, 16, 64, 0)
    Q = rng.uniform(-1, 1, (64, 16)).astype(np.float32)
    _same(g, dg, Q, 50, 512, EXT)
    _same(g, dg, Q, 10, 64, 0)


@pytest.mark.parametrize("n,d,M,ef,metric,wave_max", [
    (3000, 128, 16, 64, 0, 4096),
    (1500, 32, 8, 24, 1, 128),
    (900, 768, 32, 40, 1, 256),
])
def test_extended_wave_build_matches_oracle(oracle, n, d, M, ef, metric, wave_max):
    """The wave-batched Add with the textbook layer search inside (b200hnsw_set_build_mode; NOT the reference) against
    the oracle's add_waves in the same build mode: every list, the counters, and searches in both modes."""
    import hnsw_b200
    from hnsw_b200 import _lib
    from util import draw_levels
    rng = np.random.default_rng(31)
    X = rng.standard_normal((n, d)).astype(np.float32)
    keys = np.arange(n, dtype=np.int64)
    levels = draw_levels(oracle, n, 0.25, 46)
    g = oracle.Graph(d, M, metric)
    g.set_build_mode(oracle.MODE_EXTENDED)
    g.add_waves(keys, X, levels, ef, wave_max, n_threads=4)
    dist = hnsw_b200.EuclideanDistance if metric == 0 else hnsw_b200.CosineDistance
    dg = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=dist)
    dg.set_build_mode(_lib.MODE_EXTENDED)
    dg.add_arrays(keys, X, levels, wave_max=wave_max)
    st = dg.last_stats
    assert g.topography() == dg.Topography()
    for l in range(g.n_layers):
        rows, ptr, nbr = g.dump_layer(l)
        dn, member = dg.neighbors(l)
        for i, r in enumerate(rows):
            assert [int(v) for v in dn[r] if v >= 0] == list(nbr[ptr[i]:ptr[i + 1]]), (l, int(r))
    c = g.counters()
    for kk in ("dist_evals", "expansions", "nbr_reads", "evict_evals", "replenish_evals", "evictions"):
        assert st[kk] == c[kk], kk
    Q = rng.standard_normal((100, d)).astype(np.float32)
    _same(g, dg, Q, 10, ef, EXT)
    _same(g, dg, Q, 10, ef, 0)
    with pytest.raises(_lib.B200Error):   # the sequential Add is the reference's: no extended build there
        dg.add_arrays(np.array([n + 5], dtype=np.int64), X[:1], [0])
