"""b200exact_search_batch_noalloc_device (include/b200hnsw.h): ExactIndex.Search with its workspace in a static device
array, for callers that cannot call cudaMalloc.  No counterpart in the reference; held to the existing entry point
b200exact_search_batch_device on the same inputs: rows and counts identical, distances bit-identical (both evaluate every
distance in canonical fp32), on calls made one after the other (the array keeps the previous call's contents).

Collected last on purpose: a failure here hides nothing behind it.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _same_float_bits(a, b):
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    na, nb = np.isnan(a), np.isnan(b)
    return bool(np.array_equal(na, nb) and np.array_equal(a.view(np.uint32)[~na], b.view(np.uint32)[~nb]))


def _both(torch, ix, Q, k):
    """(existing entry point, no-alloc entry point) results for the queries Q through the C-ABI, device pointers."""
    dev = torch.device("cuda:0")
    B = Q.shape[0]
    qd = torch.from_numpy(np.ascontiguousarray(Q, dtype=np.float32)).to(dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    outs = []
    for call in (ix.search_device, ix.search_device_noalloc):
        rows = torch.full((B, k), -7, dtype=torch.int64, device=dev)
        dists = torch.full((B, k), -7.0, dtype=torch.float32, device=dev)
        counts = torch.full((B,), -7, dtype=torch.int32, device=dev)
        call(qd.data_ptr(), B, k, rows.data_ptr(), dists.data_ptr(), counts.data_ptr(), st)
        torch.cuda.synchronize(dev)
        outs.append((rows.cpu().numpy(), dists.cpu().numpy(), counts.cpu().numpy()))
    return outs


def _assert_same(ref, got, what):
    assert np.array_equal(ref[2], got[2]), what + ": counts"
    assert np.array_equal(ref[0], got[0]), what + ": rows"
    assert _same_float_bits(ref[1], got[1]), what + ": distance bits"


def test_noalloc_matches_existing_entry_point_on_consecutive_calls():
    import torch
    import hnsw_b200
    rng = np.random.default_rng(11)
    n, d = 5000, 96                      # three scan chunks: one merge level
    X = rng.standard_normal((n, d)).astype(np.float32)
    ix = hnsw_b200.NewExactIndex(hnsw_b200.EuclideanDistance)
    ix.BatchAdd(list(range(n)), X)
    ix.search_device_noalloc(0, 0, 10, 0, 0, 0, 0)   # the set-up call: resolves the array's address, no work
    # call 1: a larger batch and k, fills a good part of the array
    Q1 = rng.standard_normal((257, d)).astype(np.float32)
    ref, got = _both(torch, ix, Q1, 100)
    assert (ref[2] == 100).all()
    _assert_same(ref, got, "call 1")
    # call 2, right behind it: smaller batch, other k -- whatever call 1 left in the array must not show
    Q2 = np.concatenate([X[:8], rng.standard_normal((25, d)).astype(np.float32)])
    ref, got = _both(torch, ix, Q2, 10)
    assert (ref[0][:8, 0] == np.arange(8)).all()      # stored points find themselves
    _assert_same(ref, got, "call 2")
    # call 3: after a Delete (tombstones are honoured), same queries
    for key in (0, 3, 4999):
        assert ix.Delete(key)
    ref, got = _both(torch, ix, Q2, 10)
    assert not np.isin(ref[0], [0, 3, 4999]).any()
    _assert_same(ref, got, "call 3")
    ix.Close()


def test_noalloc_sliced_batch_cosine_and_small_index():
    import torch
    import hnsw_b200
    rng = np.random.default_rng(12)
    # 49 chunks x 4000 queries: the partial lists of the whole batch do not fit 8 MiB, the scan runs in query slices
    n, d = 100_000, 32
    X = rng.standard_normal((n, d)).astype(np.float32)
    ix = hnsw_b200.NewExactIndex(hnsw_b200.CosineDistance)
    ix.BatchAdd(list(range(n)), X)
    Q = rng.standard_normal((4000, d)).astype(np.float32)
    ref, got = _both(torch, ix, Q, 10)
    _assert_same(ref, got, "sliced")
    ix.Close()
    # k > Len(): counts = Len(), the tail of every list is (-1, NaN) on both entry points
    ix = hnsw_b200.NewExactIndex(hnsw_b200.EuclideanDistance)
    ix.BatchAdd(list(range(7)), X[:7])
    ref, got = _both(torch, ix, Q[:5], 10)
    assert (ref[2] == 7).all() and (got[0][:, 7:] == -1).all()
    _assert_same(ref, got, "k > Len()")
    ix.Close()


def test_noalloc_rejects_sizes_past_the_array():
    import torch
    import hnsw_b200
    from hnsw_b200 import _lib
    rng = np.random.default_rng(13)
    X = rng.standard_normal((3000, 16)).astype(np.float32)
    ix = hnsw_b200.NewExactIndex(hnsw_b200.EuclideanDistance)
    ix.BatchAdd(list(range(3000)), X)
    dev = torch.device("cuda:0")
    dummy = torch.zeros(16, dtype=torch.float32, device=dev)   # never touched: the sizes are refused first
    B, k = 200_000, 10                                          # 16 MB of final keys alone
    with pytest.raises(_lib.B200Error) as ei:
        ix.search_device_noalloc(dummy.data_ptr(), B, k, dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), 0)
    assert ei.value.code == _lib.ERR_UNSUPPORTED
    assert "static array holds 8388608" in str(ei.value)
    # and the entry point still answers afterwards
    Q = rng.standard_normal((9, 16)).astype(np.float32)
    ref, got = _both(torch, ix, Q, 5)
    _assert_same(ref, got, "after the refusal")
    ix.Close()
