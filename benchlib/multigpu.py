"""The N > 1 layouts.  Sharded (north_star): every rank owns an independent sub-graph, the same batch goes to every
rank, per-shard top-k lists reach the merging rank over NVLink and the K4 kernel merges them.  Replicas:
the same graph on every GPU, queries split -- handled by bench.py's ordinary N = 1 path per rank.  Exact sharded
(BASELINE configs[4]): rows of the ExactIndex split over the ranks, K3 per shard + all-gather + K4."""
from __future__ import annotations

import sys
import time

import numpy as np

from .data import merge_spec


def _max_over_ranks(torch, dist, dev, x):
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _gather_np(torch, dist, dev, arr):
    """all_gather of equal-shape numpy arrays -> [world, ...] numpy (through the device: NCCL has no CPU tensors)."""
    t = torch.from_numpy(np.ascontiguousarray(arr)).to(dev)
    out = torch.empty((dist.get_world_size(),) + tuple(t.shape), dtype=t.dtype, device=dev)
    dist.all_gather_into_tensor(out.view(-1), t.view(-1))
    return out.cpu().numpy()


def make_pipe(sharded, g, dist, B, k, n, dev, depth, exchange):
    """p2p: K1 stores each rank's share of its results into that rank's receive buffer over NVLink (search and exchange
    are one kernel, flags instead of a collective); nccl: K1 + one NCCL all-to-all."""
    if exchange == "p2p":
        try:
            return sharded.P2PShardPipeline(g, dist, B, k, n, dev, depth=depth), "p2p"
        except RuntimeError as e:  # raised on every rank together; the collective variant needs no peer mapping
            print(f"bench.py: {e}; using the NCCL all-to-all exchange", file=sys.stderr)
    return sharded.ShardPipeline(g, dist, B, k, n, dev, depth=depth), "nccl"


def sharded_search(torch, dist, sharded, g, n, dim, B, k, Qh, steps, warmup, depth, exchange, rank, world, dev,
                   barrier):
    """Device-timed and end-to-end measurement of the sharded pipeline plus the parity of its merged lists.
    Qh: pinned torch tensor [nb, B, dim], identical on every rank.  Returns a dict (meaningful on rank 0)."""
    nb = Qh.shape[0]
    Qd = Qh.to(dev)
    counters = torch.zeros(8, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream(dev)
    pipe, exchange = make_pipe(sharded, g, dist, B, k, n, dev, depth, exchange)
    for i in range(warmup):
        pipe.search(Qd[i % nb].data_ptr(), counters.data_ptr())
    pipe.drain()
    barrier()
    counters.zero_()
    ka = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    kb = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for i in range(steps):
        pipe.search(Qd[(warmup + i) % nb].data_ptr(), counters.data_ptr(), timing=(ka[i], kb[i]))
    pipe.drain()  # the timed region ends when the last merge has finished
    ev1.record(stream)
    barrier()
    total_ms = _max_over_ranks(torch, dist, dev, ev0.elapsed_time(ev1))
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in zip(ka, kb)]))
    ctr = counters.cpu().numpy().copy()
    overflow = _max_over_ranks(torch, dist, dev, float(ctr[4] != 0))

    # ---- end to end: every rank uploads the batch from ITS pinned host memory in front of the traversal and returns the
    # merged lists of its share of the queries to pinned host memory behind the merge; nbuf batches in flight
    seg, nbuf = pipe.seg, pipe.nbuf
    qtmp = [torch.empty((B, dim), dtype=torch.float32, device=dev) for _ in range(nbuf)]
    h_m = [(torch.empty((seg, k), dtype=torch.int64).pin_memory(), torch.empty((seg, k), dtype=torch.float32).pin_memory(),
            torch.empty((seg,), dtype=torch.int32).pin_memory()) for _ in range(nbuf)]
    back = [torch.cuda.Event() for _ in range(nbuf)]
    queued = [False] * nbuf

    def e2e_step(i):
        slot = pipe.step % nbuf
        if queued[slot]:
            back[slot].synchronize()  # the host has this slot's previous results

        def before(st_, sl_):
            qtmp[sl_].copy_(Qh[i], non_blocking=True)

        def after(st_, sl_):
            md, mi, mc = pipe.merged(sl_)
            h_m[sl_][0].copy_(mi, non_blocking=True)
            h_m[sl_][1].copy_(md, non_blocking=True)
            h_m[sl_][2].copy_(mc, non_blocking=True)
            back[sl_].record(st_)

        pipe.search(qtmp[slot].data_ptr(), before=before, after=after)
        queued[slot] = True
        return slot

    def e2e_finish():
        for sl_ in range(nbuf):
            if queued[sl_]:
                back[sl_].synchronize()

    for i in range(max(warmup, 3)):
        e2e_step(i % nb)
    e2e_finish()
    barrier()
    t0 = time.perf_counter()
    for i in range(steps):
        e2e_step((warmup + i) % nb)
    e2e_finish()
    torch.cuda.synchronize(dev)
    e2e_s = _max_over_ranks(torch, dist, dev, time.perf_counter() - t0)

    # ---- parity of the merged lists (outside every timed region): one more step through the pipeline on a known batch;
    # every rank also answers the whole batch from its own shard through the host-pointer call, rank 0 applies the merge
    # SPECIFICATION (benchlib/data.py merge_spec) to those and compares with what the pipeline's ranks merged.
    last = (warmup + steps) % nb
    slot = e2e_step(last)
    e2e_finish()
    mi, md, mc = (x.numpy().copy() for x in h_m[slot])
    own_rows, own_d, own_c = g.search_rows(Qh[last].numpy(), k)
    own_ids = np.where(own_rows >= 0, own_rows.astype(np.int64) + rank * n, -1)
    all_i = _gather_np(torch, dist, dev, own_ids)
    all_d = _gather_np(torch, dist, dev, own_d)
    all_c = _gather_np(torch, dist, dev, own_c)
    got_i = _gather_np(torch, dist, dev, mi).reshape(world * seg, k)[:B]
    got_d = _gather_np(torch, dist, dev, md).reshape(world * seg, k)[:B]
    got_c = _gather_np(torch, dist, dev, mc).reshape(world * seg)[:B]
    parity = None
    if rank == 0:
        # shard lists arrive in the reference's heap-array order; the merge orders by (distance, shard, position)
        s_d, s_i, s_c = merge_spec(all_d, all_i, all_c, k)
        keep = np.arange(k)[None, :] < s_c[:, None]
        parity = {"checked_queries": int(B), "ranks": world,
                  "rows_equal": bool(np.array_equal(np.where(keep, got_i, -1), s_i)),
                  "dists_bit_equal": bool(np.array_equal(np.where(keep, got_d, 0).view(np.uint32),
                                                         np.where(keep, s_d, 0).view(np.uint32))),
                  "counts_equal": bool(np.array_equal(got_c, s_c)),
                  "against": "merge specification (numpy) applied to every rank's own b200hnsw_search_batch results for "
                             "the same batch, gathered outside the timed region"}
    if hasattr(pipe, "close"):
        pipe.close()  # collective: unmap the peers' receive buffers, then free mine
    return {"total_ms": total_ms, "event_kernel_ms": kernel_ms, "ctr": ctr, "overflow": bool(overflow),
            "e2e_s": e2e_s, "seg": seg, "nbuf": nbuf, "exchange": exchange, "parity": parity,
            "own_results": (own_rows, own_d, own_c), "last_batch": last}


def exact_sharded(torch, dist, sharded, hnsw_b200, X, metric, Qh, k, steps, warmup, rank, world, dev, barrier):
    """BASELINE configs[4]: the rows of one ExactIndex split round-robin over the ranks (global row = local * world +
    rank); per step K3 on the shard, ONE all-gather of the per-shard lists, K4 merge on every rank."""
    from .device import distance_of
    n_local, dim = X.shape
    B = Qh.shape[1]
    nb = Qh.shape[0]
    ix = hnsw_b200.ExactIndex(distance_of(hnsw_b200, metric), device=dev.index, capacity=n_local)
    t0 = time.perf_counter()
    ix.add_arrays(range(n_local), X)
    add_s = time.perf_counter() - t0
    Qd = Qh.to(dev)
    rows = torch.empty((B, k), dtype=torch.int64, device=dev)
    dd = torch.empty((B, k), dtype=torch.float32, device=dev)
    cnt = torch.empty((B,), dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev)

    def step(i):
        ix.search_device(Qd[i % nb].data_ptr(), B, k, rows.data_ptr(), dd.data_ptr(), cnt.data_ptr(), stream.cuda_stream)
        gids = torch.where(rows >= 0, rows * world + rank, rows)
        a, b, c = sharded.gather_shard_results(dist, dd, gids, cnt)
        return sharded.merge_topk_device(a, b, c, k, stream.cuda_stream)

    for i in range(max(3, warmup)):
        step(i)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(steps):
        m = step(warmup + i)
    e1.record(stream)
    barrier()
    ms = _max_over_ranks(torch, dist, dev, e0.elapsed_time(e1) / steps)
    # end to end on one batch: host queries in, merged lists out to the host
    hq = Qh[(warmup + steps - 1) % nb]
    qt = torch.empty((B, dim), dtype=torch.float32, device=dev)
    barrier()
    t0 = time.perf_counter()
    qt.copy_(hq, non_blocking=True)
    ix.search_device(qt.data_ptr(), B, k, rows.data_ptr(), dd.data_ptr(), cnt.data_ptr(), stream.cuda_stream)
    gids = torch.where(rows >= 0, rows * world + rank, rows)
    a, b, c = sharded.gather_shard_results(dist, dd, gids, cnt)
    md, mi, mc = sharded.merge_topk_device(a, b, c, k, stream.cuda_stream)
    h_i, h_d, h_c = mi.cpu().numpy(), md.cpu().numpy(), mc.cpu().numpy()
    e2e_s = _max_over_ranks(torch, dist, dev, time.perf_counter() - t0)
    st = None
    try:
        ix.search_rows(hq.numpy()[:256], k)
        st = ix.last_stats
    except Exception:  # noqa: BLE001
        pass
    # every rank must hold the same merged lists
    chk = _gather_np(torch, dist, dev, h_i)
    same_everywhere = bool((chk == chk[0:1]).all())
    ix.Close()
    return {"ms_per_step": ms, "e2e_s": e2e_s, "add_vec_per_s": n_local / add_s, "merged": (h_i, h_d, h_c),
            "ranks_agree": same_everywhere, "last_batch": (warmup + steps - 1) % nb, "stats": st}
