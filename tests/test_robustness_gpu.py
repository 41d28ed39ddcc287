"""Edge cases the reference's own code anticipates (graph.go:65-66: NaN from cosine on zero embeddings; graph.go:435-436
"replaced"; a graph emptied by Delete) and the failure paths of the C ABI: the device must stay bit-identical to the
oracle through them, and a failed or retried call must leave the handle usable."""
import numpy as np
import pytest

from util import build_oracle_graph, flatten_oracle, same_float_bits

pytestmark = pytest.mark.gpu


def _device_graph(oracle, g, M, ef, metric):
    import hnsw_b200
    dist = hnsw_b200.EuclideanDistance if metric == oracle.EUCLIDEAN else hnsw_b200.CosineDistance
    dg = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=dist)
    dg.load_flat(*flatten_oracle(g))
    return dg


def _same(g, dg, Q, k, ef):
    g.reset_counters()
    o_rows, o_d, o_cnt, o_ctr = g.batch_search(Q, k, ef)
    d_rows, d_d, d_cnt = dg.search_rows(Q, k, ef)
    assert np.array_equal(o_cnt, d_cnt)
    assert np.array_equal(o_rows, d_rows)
    assert same_float_bits(o_d, d_d)
    st = dg.last_stats
    assert (st["dist_evals"], st["expansions"], st["nbr_reads"]) == (
        o_ctr["dist_evals"], o_ctr["expansions"], o_ctr["nbr_reads"])


@pytest.mark.parametrize("k,ef", [(10, 64), (1, 20), (40, 30)])
def test_cosine_with_zero_vectors_in_graph_and_queries(oracle, k, ef):
    """CosineDistance of a zero vector is NaN (0/0).  Go's heap never moves an element past a NaN, so after a Pop a NaN
    entry can sit at the root of the candidate queue and be expanded next: the kernel must keep the Go-exact queue
    for such graphs instead of its 'last improving row' shortcut."""
    rng = np.random.default_rng(11)
    n, d, M = 1500, 24, 12
    X = rng.standard_normal((n, d)).astype(np.float32)
    X[rng.choice(n, 150, replace=False)] = 0.0           # 10 % zero rows, spread over the layers
    g, _, _ = build_oracle_graph(oracle, X, M, 0.25, ef, oracle.COSINE)
    dg = _device_graph(oracle, g, M, ef, oracle.COSINE)
    Q = rng.standard_normal((200, d)).astype(np.float32)
    Q[::7] = 0.0                                          # zero queries: every distance NaN
    Q[3] = 1e-25                                          # |q|^2 underflows: mixed NaN / inf distances
    Q[5] = 3e19                                           # |q|^2 overflows
    _same(g, dg, Q, k, ef)
    _same(g, dg, X[:100], k, ef)


def test_zero_rows_added_on_the_device_keep_later_searches_exact(oracle):
    """The rows arrive through Add (both paths), not load: the device flags them as they land."""
    import hnsw_b200
    rng = np.random.default_rng(12)
    n, d, M, ef = 1200, 16, 8, 40
    X = rng.standard_normal((n, d)).astype(np.float32)
    X[400:440] = 0.0
    X[900] = np.inf
    from util import draw_levels
    levels = draw_levels(oracle, n, 0.25, 3)
    keys = np.arange(n, dtype=np.int64)
    for wave_max in (0, 64):
        og = oracle.Graph(d, M, oracle.COSINE)
        dg = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=hnsw_b200.CosineDistance)
        if wave_max:
            og.add_waves(keys, X, levels, ef, wave_max)
        else:
            og.add_many(keys, X, levels, ef)
        dg.add_arrays(keys, X, levels, wave_max=wave_max)
        assert og.topography() == dg.Topography()
        for l in range(og.n_layers):
            rows, ptr, nbr = og.dump_layer(l)
            dn, mem = dg.neighbors(l)
            for i, r in enumerate(rows):
                assert list(dn[r][dn[r] >= 0]) == list(nbr[ptr[i]:ptr[i + 1]]), (wave_max, l, r)
        Q = rng.standard_normal((64, d)).astype(np.float32)
        _same(og, dg, Q, 10, ef)


def test_visited_overflow_retries_and_leaves_the_handle_usable(oracle):
    """the larger table becomes the handle's hint only after a retry succeeded, and never a size no launch
    can have.  k = 200 needs > 128 visited nodes, more than half of the smallest table."""
    from hnsw_b200 import _lib
    rng = np.random.default_rng(13)
    X = rng.uniform(-1, 1, (3000, 32)).astype(np.float32)
    g, _, _ = build_oracle_graph(oracle, X, 16, 0.25, 64, 0)
    dg = _device_graph(oracle, g, 16, 64, 0)
    L = _lib.lib()
    _lib.check(L.b200hnsw_set_visited_capacity(dg._h, 256))
    assert L.b200hnsw_get_visited_capacity(dg._h) == 256
    Q = rng.uniform(-1, 1, (50, 32)).astype(np.float32)
    _same(g, dg, Q, 200, 256)                              # overflows 256, redone with 1024
    grown = L.b200hnsw_get_visited_capacity(dg._h)
    assert grown >= 1024 and grown & (grown - 1) == 0
    _same(g, dg, Q, 10, 64)                                # an ordinary search afterwards
    with pytest.raises(_lib.B200Error):
        _lib.check(L.b200hnsw_set_visited_capacity(dg._h, 1 << 20))   # more than a traversal can be given
    with pytest.raises(_lib.B200Error):
        _lib.check(L.b200hnsw_set_visited_capacity(dg._h, 300))       # not a power of two
    assert L.b200hnsw_get_visited_capacity(dg._h) == grown
    _same(g, dg, Q, 10, 64)


def test_dimension_is_kept_after_every_node_was_deleted(oracle):
    """Len() == 0 after deleting everything must not let a vector of another width through."""
    import hnsw_b200
    g = hnsw_b200.Graph(M=4, Ml=0.25, EfSearch=10, Distance=hnsw_b200.EuclideanDistance)
    g.Add(hnsw_b200.MakeNode(1, [1.0, 2.0, 3.0]), hnsw_b200.MakeNode(2, [2.0, 2.0, 3.0]), levels=[0, 0])
    assert g.BatchDelete([1, 2]) == [True, True] and g.Len() == 0
    with pytest.raises(ValueError, match="embedding dimension mismatch: 3 != 2"):
        g.Add(hnsw_b200.MakeNode(3, [1.0, 2.0]), levels=[0])
    with pytest.raises(ValueError, match="embedding dimension mismatch: 3 != 5"):
        g.add_arrays([4], np.zeros((1, 5), np.float32), levels=[0])


def test_failed_load_leaves_an_empty_graph(oracle):
    import hnsw_b200
    from hnsw_b200 import _lib
    rng = np.random.default_rng(14)
    X = rng.uniform(-1, 1, (300, 8)).astype(np.float32)
    g, _, _ = build_oracle_graph(oracle, X, 6, 0.25, 20, 0)
    dg = _device_graph(oracle, g, 6, 20, 0)
    flat = list(flatten_oracle(g))
    bad = [(r.copy(), p.copy(), b.copy()) for r, p, b in flat[3]]
    bad[0][2][0] = 10 ** 6                                   # a neighbour row that does not exist
    with pytest.raises(_lib.B200Error):
        dg.load_flat(flat[0], flat[1], flat[2], bad, flat[4])
    # rejected while validating: the old graph is untouched and still answers
    _same(g, dg, X[:10], 5, 20)


@pytest.mark.parametrize("kind", ["p2p", "nccl"])
def test_pipeline_search_all_raises_the_visited_capacity_and_reissues(oracle, kind):
    """The pipelines drive the device-pointer entry points, which report an overflowed visited set in d_counters[4]
    instead of retrying (include/b200hnsw.h): search_all reads it, raises the capacity and issues the batches again.
    One rank here (no process group); the 2-rank agreement on the gathered error word is in test_sharded_gpu.py."""
    import torch
    from hnsw_b200 import sharded
    rng = np.random.default_rng(21)
    n, d, k, ef, B = 3000, 32, 200, 256, 64
    X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    g, _, _ = build_oracle_graph(oracle, X, 16, 0.25, 64, 0)
    dg = _device_graph(oracle, g, 16, ef, 0)
    dg.visited_capacity = 256                                   # k = 200 visits more than that
    dev = torch.device("cuda:0")
    cls = sharded.P2PShardPipeline if kind == "p2p" else sharded.ShardPipeline
    pipe = cls(dg, None, B, k, n, dev)
    Qs = [rng.uniform(-1, 1, (B, d)).astype(np.float32) for _ in range(5)]
    qds = [torch.from_numpy(q).to(dev) for q in Qs]
    keep_i = torch.full((len(Qs), B, k), -9, dtype=torch.int64, device=dev)
    keep_c = torch.full((len(Qs), B), -9, dtype=torch.int32, device=dev)
    order = {}

    def after(stream, slot):
        i = order[slot].pop(0)
        md, mi, mc = pipe.merged(slot)
        keep_i[i].copy_(mi[:B], non_blocking=True)
        keep_c[i].copy_(mc[:B], non_blocking=True)

    def before(stream, slot):
        order.setdefault(slot, []).append(before.i % len(Qs))
        before.i += 1
    before.i = 0
    slots = pipe.search_all([q.data_ptr() for q in qds], before=before, after=after)
    assert len(slots) == len(Qs)
    assert before.i == 2 * len(Qs)                              # every batch was issued twice: 256 -> 1024 was enough
    assert dg.visited_capacity == 1024
    torch.cuda.synchronize(dev)
    for i, Q in enumerate(Qs):
        rows, dists, cnt = dg.search_rows(Q, k, ef)             # the host-pointer call (retries by itself)
        assert np.array_equal(cnt, keep_c[i].cpu().numpy())
        got = keep_i[i].cpu().numpy()
        for q in range(B):                                      # K4 returns the list ascending (stable), Search heap order
            o = np.argsort(dists[q, :cnt[q]], kind="stable")
            assert np.array_equal(rows[q, :cnt[q]][o], got[q, :cnt[q]])
    if hasattr(pipe, "close"):
        pipe.close()


def test_pageable_and_page_locked_host_buffers_give_the_same_lists(oracle):
    """b200hnsw_search_batch stages pageable buffers itself (chunked copies through the call context's page-locked
    buffer, results written there by the kernel and copied out); page-locked buffers are read by the copy engine and
    written by the kernel directly; B200HNSW_NO_HOST_STAGING=1 leaves pageable buffers to the driver.  Same lists, and
    the oracle's, on all three routes; a batch small enough to skip the staging (< 256 KB of queries) as well."""
    import ctypes as C
    import os
    import torch
    from hnsw_b200 import _lib
    rng = np.random.default_rng(31)
    n, d, k, ef = 6000, 128, 10, 48
    X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    g, _, _ = build_oracle_graph(oracle, X, 12, 0.25, 64, 0)
    dg = _device_graph(oracle, g, 12, ef, 0)
    for B in (4500, 300):
        Q = rng.uniform(-1, 1, (B, d)).astype(np.float32)
        o_rows, o_d, o_cnt, _ = g.batch_search(Q, k, ef)
        a = dg.search_rows(Q, k, ef)                                   # pageable numpy buffers: own staging
        os.environ["B200HNSW_NO_HOST_STAGING"] = "1"                   # read once per process: no effect if already read
        qp = torch.from_numpy(Q).pin_memory()
        rows = torch.empty((B, k), dtype=torch.int32).pin_memory()
        dists = torch.empty((B, k), dtype=torch.float32).pin_memory()
        cnt = torch.empty((B,), dtype=torch.int32).pin_memory()
        _lib.check(_lib.lib().b200hnsw_search_batch(
            dg._h, C.cast(qp.data_ptr(), _lib.f32p), B, k, ef, 0, _lib.MODE_REFERENCE, None,
            C.cast(rows.data_ptr(), _lib.i32p), C.cast(dists.data_ptr(), _lib.f32p), C.cast(cnt.data_ptr(), _lib.i32p),
            None))
        os.environ.pop("B200HNSW_NO_HOST_STAGING", None)
        for got in (a, (rows.numpy(), dists.numpy(), cnt.numpy())):
            assert np.array_equal(got[2], o_cnt)
            assert np.array_equal(got[0], o_rows)
            assert same_float_bits(got[1], o_d)
