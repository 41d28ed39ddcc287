"""Size-independent properties at BASELINE.json's FULL sizes (C2: 1 M x 128 L2, M = 16, EfSearch = 64, 10 k-query batches;
exact search over the same rows), where the oracle would need minutes: what must hold for ANY correct implementation of
the reference's data structures, checked on the device's own output.

  build   : every neighbour list is sorted by key, has no duplicate, no self link, at most M entries, and names only
            nodes that are members of the layer (graph.go:41-81, 516-520); layer sizes follow the drawn levels
  search  : the result array of the reference mode is a binary min-heap on distance (container/heap after Push /
            PopLast, graph.go:147-169) with distinct rows, and every distance is the canonical distance of (query,
            row) bit for bit (recomputed through b200hnsw_distance_batch); the extended mode is ascending with ties by
            row; a second identical call and the device-pointer entry point return the same bits (idempotence)
  exact   : ascending, ties by row, distances canonical; never worse than anything the graph search found
Smaller oracle-backed parity tests are in the other files; bench.py repeats the oracle comparison at full size on
every run.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N, D, M, EF, K, B = 1_000_000, 128, 16, 64, 10, 10_000


@pytest.fixture(scope="module")
def c2():
    import hnsw_b200
    from benchlib.data import draw_levels, gen
    X = gen("uniform", N, D, 42)
    levels = draw_levels(N, 0.25, 44)
    g = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=EF, Distance=hnsw_b200.EuclideanDistance, capacity=N)
    g.add_arrays(np.arange(N, dtype=np.int64), X, levels, wave_max=16384)
    Q = gen("uniform", B, D, 43)
    yield g, X, levels, Q
    g.close()


def test_full_size_graph_invariants(c2):
    g, X, levels, _ = c2
    topo = g.Topography()
    assert topo[0] == N
    for l, size in enumerate(topo):
        assert size == int((levels >= l).sum())                       # a node is a member of layers 0..level
    for l in range(len(topo)):
        nbr, member = g.neighbors(l)
        member = member.astype(bool)
        assert member.sum() == topo[l]
        sub = nbr[member]
        rows = np.nonzero(member)[0]
        valid = sub >= 0
        assert (valid.sum(axis=1) <= M).all()
        assert (valid[:, :-1] >= valid[:, 1:]).all()                   # -1 padding only at the end
        assert not (sub == rows[:, None]).any()                        # no self links
        assert member[sub[valid]].all()                                # neighbours are members of the layer
        nxt_ok = ~valid[:, 1:] | (sub[:, 1:] > sub[:, :-1])            # keys = rows here: strictly ascending
        assert nxt_ok.all()
        assert not (~member[:, None] & (nbr >= 0)).any()               # non-members hold no list on this layer


def _heap_ok(d, cnt):
    """dists[:cnt] is a binary min-heap."""
    idx = np.arange(1, d.shape[1])
    parent = (idx - 1) // 2
    child_valid = idx[None, :] < cnt[:, None]
    return bool((~child_valid | (d[:, parent] <= d[:, idx])).all())


def test_full_size_search_properties(c2):
    import torch
    import hnsw_b200
    from hnsw_b200 import _lib
    from hnsw_b200.distance import distance_batch
    g, X, _, Q = c2
    rows, dists, cnt = g.search_rows(Q, K, EF)
    assert (cnt == K).all()
    assert _heap_ok(dists, cnt)
    assert (np.sort(rows, axis=1)[:, 1:] != np.sort(rows, axis=1)[:, :-1]).all()      # distinct rows
    pick = np.random.default_rng(1).choice(B, 512, replace=False)
    a = np.repeat(Q[pick], K, axis=0)
    b = X[rows[pick].ravel()]
    canon = distance_batch(0, b, a).reshape(len(pick), K)                             # layerNode.search: distance(node, query)
    assert np.array_equal(canon.view(np.uint32), dists[pick].view(np.uint32))
    # idempotence and the device-pointer entry point
    rows2, dists2, cnt2 = g.search_rows(Q, K, EF)
    assert np.array_equal(rows, rows2) and np.array_equal(dists.view(np.uint32), dists2.view(np.uint32))
    dev = torch.device("cuda:0")
    qd = torch.from_numpy(Q).to(dev)
    r_d = torch.empty((B, K), dtype=torch.int32, device=dev)
    d_d = torch.empty((B, K), dtype=torch.float32, device=dev)
    c_d = torch.empty((B,), dtype=torch.int32, device=dev)
    ctr = torch.zeros(8, dtype=torch.int64, device=dev)
    g.search_device(qd.data_ptr(), B, K, r_d.data_ptr(), d_d.data_ptr(), c_d.data_ptr(), ctr.data_ptr(),
                    torch.cuda.current_stream(dev).cuda_stream)
    torch.cuda.synchronize(dev)
    assert int(ctr[4]) == 0
    assert np.array_equal(r_d.cpu().numpy(), rows) and np.array_equal(d_d.cpu().numpy().view(np.uint32), dists.view(np.uint32))
    # extended mode: ascending by (distance, row)
    er, ed, ec = g.search_rows(Q[:2000], K, EF, mode=_lib.MODE_EXTENDED)
    assert (ec == K).all()
    assert ((ed[:, 1:] > ed[:, :-1]) | ((ed[:, 1:] == ed[:, :-1]) & (er[:, 1:] > er[:, :-1]))).all()
    # the textbook termination hardly ever ends on a worse nearest neighbour than the reference rule's (no guarantee:
    # the two descents can reach layer 0 at different nodes)
    assert (ed[:, 0] <= dists[:2000].min(axis=1)).mean() > 0.99


def test_full_size_exact_properties(c2):
    import hnsw_b200
    from hnsw_b200.distance import distance_batch
    g, X, _, Q = c2
    ix = hnsw_b200.NewExactIndex(hnsw_b200.EuclideanDistance, capacity=N)
    ix.add_arrays(range(N), X)
    nq = 2000
    rows, dists, cnt = ix.search_rows(Q[:nq], K)
    assert ix.last_stats["launches"] == 2                                             # the tensor path answered
    assert (cnt == K).all()
    assert ((dists[:, 1:] > dists[:, :-1]) | ((dists[:, 1:] == dists[:, :-1]) & (rows[:, 1:] > rows[:, :-1]))).all()
    pick = np.random.default_rng(2).choice(nq, 256, replace=False)
    canon = distance_batch(0, X[rows[pick].ravel()], np.repeat(Q[pick], K, axis=0)).reshape(len(pick), K)
    assert np.array_equal(canon.view(np.uint32), dists[pick].view(np.uint32))
    # exact means exact: nothing the graph search found is closer than the exact k-th, and the exact nearest is at
    # least as near as the graph's
    gr, gd, gc = g.search_rows(Q[:nq], K, EF)
    assert (dists[:, 0] <= gd.min(axis=1)).all()
    # float64 brute force on a few queries: the same ids (no ties expected on continuous data)
    few = pick[:8]
    X64 = X.astype(np.float64)
    for q in few:
        dd = ((X64 - Q[q].astype(np.float64)) ** 2).sum(1)
        want = np.argsort(dd, kind="stable")[:K]
        assert np.array_equal(np.sort(want), np.sort(rows[q]))
    ix.Close()
