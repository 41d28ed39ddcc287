"""N>1 path on CPU: world_size 2 over gloo.  Each rank owns an independent sub-graph (built and searched by the
oracle here, since there is no GPU), the product's all-gather plumbing (hnsw_b200.sharded.gather_shard_results) moves
the per-shard lists, and the merged result must equal merge(oracle shard results) computed in one process
("merge(oracle shard results) == device merged results")."""
import os
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def _worker(rank, world, port, tmp):
    sys.path.insert(0, str(ROOT))
    sys.path.insert(0, str(ROOT / "tests"))
    import torch
    import torch.distributed as dist
    from hnsw_b200 import sharded
    from merge_ref import merge_topk_ref
    from oracle import binding as oracle
    from util import build_oracle_graph
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n, d, k, ef = 600, 32, 10, 20
    X = np.random.default_rng(100 + rank).uniform(-1, 1, (n, d)).astype(np.float32)
    Q = np.random.default_rng(7).uniform(-1, 1, (40, d)).astype(np.float32)  # the same on every rank ("broadcast")
    g, _, _ = build_oracle_graph(oracle, X, 8, 0.25, ef, 0, seed=50 + rank)
    rows, dd, cnt, _ = g.batch_search(Q, k, ef)
    ids = sharded.global_ids(torch.from_numpy(rows), rank, n)
    all_d, all_i, all_c = sharded.gather_shard_results(dist, torch.from_numpy(dd), ids, torch.from_numpy(cnt))
    m = merge_topk_ref(all_d.numpy(), all_i.numpy(), all_c.numpy(), k)
    # the packed layout the GPU path uses (one buffer, one collective) carries the same lists
    pr = sharded.PackedResults(len(Q), k, world, torch.device("cpu"))
    pr.rows.copy_(torch.from_numpy(rows))
    pr.dists.copy_(torch.from_numpy(dd))
    pr.counts.copy_(torch.from_numpy(cnt))
    pr.gather(dist)
    p_rows, p_d, p_c = pr.shard_views()
    assert np.array_equal(p_d.numpy(), all_d.numpy(), equal_nan=True) and np.array_equal(p_c.numpy(), all_c.numpy())
    gid = np.where(p_rows.numpy() >= 0, p_rows.numpy().astype(np.int64) + np.arange(world)[:, None, None] * n, -1)
    assert np.array_equal(gid, all_i.numpy())
    # the all-to-all layout: block s of `send` = my results for the queries rank s merges; after the exchange block g
    # of `recv` = shard g's results for MY queries
    sr = sharded.SegmentedResults(len(Q), k, world, rank, torch.device("cpu"))
    s_rows, s_d, s_c = sr.block_views(sr.send)
    for s_ in range(world):
        a, b = s_ * sr.seg, min(len(Q), (s_ + 1) * sr.seg)
        s_rows[s_, :b - a] = torch.from_numpy(rows[a:b])
        s_d[s_, :b - a] = torch.from_numpy(dd[a:b])
        s_c[s_, :b - a] = torch.from_numpy(cnt[a:b])
    sr.exchange(dist)
    r_rows, r_d, r_c = sr.block_views(sr.recv)
    nq = sr.q1 - sr.q0
    assert np.array_equal(r_d.numpy()[:, :nq], all_d.numpy()[:, sr.q0:sr.q1], equal_nan=True)
    assert np.array_equal(r_c.numpy()[:, :nq], all_c.numpy()[:, sr.q0:sr.q1])
    gid2 = np.where(r_rows.numpy() >= 0, r_rows.numpy().astype(np.int64) + np.arange(world)[:, None, None] * n, -1)
    assert np.array_equal(gid2[:, :nq], all_i.numpy()[:, sr.q0:sr.q1])
    np.savez(os.path.join(tmp, f"rank{rank}.npz"), d=m[0], i=m[1], c=m[2], rows=rows, dd=dd, cnt=cnt)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gather_and_merge(tmp_path):
    import torch.multiprocessing as mp
    sys.path.insert(0, str(ROOT / "tests"))
    from merge_ref import merge_topk_ref
    world, port = 2, 29517 + os.getpid() % 200
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r = [np.load(tmp_path / f"rank{i}.npz") for i in range(world)]
    # every rank ends with the same merged lists
    for key in ("d", "i", "c"):
        assert np.array_equal(r[0][key], r[1][key], equal_nan=True)
    # and they equal the merge of the per-shard oracle results done in one process
    all_d = np.stack([r[i]["dd"] for i in range(world)])
    all_i = np.stack([np.where(r[i]["rows"] >= 0, r[i]["rows"].astype(np.int64) + i * 600, -1) for i in range(world)])
    all_c = np.stack([r[i]["cnt"] for i in range(world)])
    m = merge_topk_ref(all_d, all_i, all_c, 10)
    assert np.array_equal(m[0], r[0]["d"], equal_nan=True) and np.array_equal(m[1], r[0]["i"])
    # merged distances ascend and ids come from both shards
    dd = r[0]["d"]
    assert (np.diff(dd, axis=1) >= 0).all()
    assert (r[0]["i"] >= 600).any() and (r[0]["i"] < 600).any()


def test_merge_ref_tie_rule():
    sys.path.insert(0, str(ROOT / "tests"))
    from merge_ref import merge_topk_ref
    d = np.array([[[1.0, 2.0, np.nan]], [[1.0, 1.5, 9.0]]], dtype=np.float32)  # G=2, B=1, k=3
    i = np.array([[[10, 11, 12]], [[20, 21, 22]]], dtype=np.int64)
    c = np.array([[3], [2]], dtype=np.int32)
    md, mi, mc = merge_topk_ref(d, i, c, 4)
    assert list(mi[0]) == [10, 20, 21, 11] and mc[0] == 4  # tie 1.0: shard 0 first; NaN last; 9.0 beyond count
