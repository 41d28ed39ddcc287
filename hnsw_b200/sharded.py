"""Multi-GPU layout of north_star (no reference counterpart: the reference is one process): every rank owns an
independent sub-graph, queries are the same on all ranks, per-shard top-k lists are all-gathered (NCCL over NVLink on
GPUs; gloo in the CPU tests) and merged by the K4 kernel.  Tie rule of the merge: lower shard first, then position in
the shard's list -- the stable concatenate-then-sort of hnsw-extensions/hybrid/adapter.go:264-292.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib


def shard_of_row(row: int, world: int) -> int:
    """Round-robin partition of SURVEY 8(e): shards are i.i.d. samples of the data."""
    return row % world


def global_ids(rows, rank: int, rows_per_shard: int):
    """Shard-local rows -> global ids for contiguous per-rank id ranges (unused slots stay negative)."""
    import torch
    ids = rows.to(torch.int64)
    return torch.where(ids >= 0, ids + rank * rows_per_shard, ids)


def gather_shard_results(dist, dists, ids, counts):
    """All-gather of per-shard results: dists [B,k] f32, ids [B,k] i64, counts [B] i32 -> [G,B,k], [G,B,k], [G,B].
    One collective per array; B*k*12 bytes per rank (0.8 MB at B=10k, k=10): latency, not bandwidth, is the cost."""
    import torch
    world = dist.get_world_size()
    def gather(t):
        t = t.contiguous()
        out = torch.empty((world * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(out, t)  # concatenation along dim 0 (the layout gloo and NCCL both accept)
        return out.view((world,) + tuple(t.shape))
    return gather(dists), gather(ids), gather(counts)


def merge_topk_device(all_d, all_i, all_c, k: int, stream: int = 0):
    """K4 on CUDA tensors laid out as gather_shard_results returns them -> (dists [B,k], ids [B,k], counts [B])."""
    import torch
    G, B = all_c.shape
    dev = all_d.device
    if dev.type != "cuda":
        raise RuntimeError("merge_topk_device needs CUDA tensors; hnsw_b200 has no CPU fallback")
    m_d = torch.empty((B, k), dtype=torch.float32, device=dev)
    m_i = torch.empty((B, k), dtype=torch.int64, device=dev)
    m_c = torch.empty((B,), dtype=torch.int32, device=dev)
    _lib.check(_lib.lib().b200hnsw_merge_topk_device(
        dev.index or 0, G, B, k, C.c_void_p(all_d.data_ptr()), C.c_void_p(all_i.data_ptr()),
        C.c_void_p(all_c.data_ptr()), 0, C.c_void_p(m_d.data_ptr()), C.c_void_p(m_i.data_ptr()),
        C.c_void_p(m_c.data_ptr()), C.c_void_p(stream) if stream else None))
    return m_d, m_i, m_c
