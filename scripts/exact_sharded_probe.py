"""Exploration: BASELINE config 5 (exact search, rows sharded over the GPUs of one box, K3 per shard + gather + K4).
Launch with torchrun; every rank holds N/world rows; prints the per-step time of rank 0 (max over ranks)."""
import os, sys, time
import numpy as np
sys.path.insert(0, '.')
import torch, torch.distributed as dist
import hnsw_b200
from hnsw_b200 import sharded
N, d, B, k, metric = [int(v) for v in sys.argv[1].split(',')]
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank); dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
n = N // world
X = np.random.default_rng(100 + rank).standard_normal((n, d), dtype=np.float32)
ix = hnsw_b200.NewExactIndex(hnsw_b200.EuclideanDistance if metric == 0 else hnsw_b200.CosineDistance, device=rank, capacity=n)
ix.add_arrays(range(n), X)
Qd = torch.from_numpy(np.random.default_rng(7).standard_normal((B, d), dtype=np.float32)).to(dev)
rows = torch.empty((B, k), dtype=torch.int64, device=dev); dd = torch.empty((B, k), dtype=torch.float32, device=dev)
cnt = torch.empty((B,), dtype=torch.int32, device=dev)
stream = torch.cuda.current_stream(dev)
def step():
    ix.search_device(Qd.data_ptr(), B, k, rows.data_ptr(), dd.data_ptr(), cnt.data_ptr(), stream.cuda_stream)
    gids = torch.where(rows >= 0, rows * world + rank, rows)
    a, b, c = sharded.gather_shard_results(dist, dd, gids, cnt)
    return sharded.merge_topk_device(a, b, c, k, stream.cuda_stream)
for _ in range(3): step()
dist.barrier(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
steps = 10
e0.record(stream)
for _ in range(steps): step()
e1.record(stream); dist.barrier(); torch.cuda.synchronize()
t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    ms = float(t.item())
    print(f"exact sharded: N={N} d={d} B={B} k={k} metric={metric} world={world}: {ms:.2f} ms/step, {B/ms*1e3:.0f} queries/s, "
          f"{2.0*N*B*d/ms/1e9:.0f} TFLOP/s aggregate", flush=True)
dist.barrier(); dist.destroy_process_group()
