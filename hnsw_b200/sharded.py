"""Multi-GPU layout of north_star (no reference counterpart: the reference is one process): every rank owns an
independent sub-graph, queries are the same on all ranks, per-shard top-k lists are all-gathered (NCCL over NVLink on
GPUs; gloo in the CPU tests) and merged by the K4 kernel.  Tie rule of the merge: lower shard first, then position in
the shard's list -- the stable concatenate-then-sort of hnsw-extensions/hybrid/adapter.go:264-292.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _lib
from .distance import metric_of as _metric_of

_MAX_PEERS = 16  # B200HNSW_MAX_PEERS


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


class PackedResults:
    """One contiguous int32 buffer per rank laid out as the search kernel's three outputs (rows [B,k] | dists [B,k] |
    counts [B]), so a step needs ONE all-gather and no repacking kernels; `merge` runs K4 on the gathered buffer and
    turns shard rows into global ids (shard * rows_per_shard + row) inside the kernel."""

    def __init__(self, B: int, k: int, world: int, device):
        import torch
        self.B, self.k, self.world = B, k, world
        self.stride = 2 * B * k + B
        self.local = torch.empty(self.stride, dtype=torch.int32, device=device)
        self.gathered = torch.empty(world * self.stride, dtype=torch.int32, device=device)
        self.rows = self.local[:B * k].view(B, k)
        self.dists = self.local[B * k:2 * B * k].view(torch.float32).view(B, k)
        self.counts = self.local[2 * B * k:]
        dev = self.local.device
        self.m_d = torch.empty((B, k), dtype=torch.float32, device=dev)
        self.m_i = torch.empty((B, k), dtype=torch.int64, device=dev)
        self.m_c = torch.empty((B,), dtype=torch.int32, device=dev)

    def gather(self, dist):
        dist.all_gather_into_tensor(self.gathered, self.local)
        return self.gathered

    def shard_views(self):
        """(rows [G,B,k] i32, dists [G,B,k] f32, counts [G,B] i32) views of the gathered buffer (for checks)."""
        import torch
        g = self.gathered.view(self.world, self.stride)
        B, k = self.B, self.k
        return (g[:, :B * k].reshape(self.world, B, k), g[:, B * k:2 * B * k].view(torch.float32).reshape(self.world, B, k),
                g[:, 2 * B * k:])

    def merge(self, rows_per_shard: int, stream: int = 0):
        if self.gathered.device.type != "cuda":
            raise RuntimeError("PackedResults.merge needs CUDA tensors; hnsw_b200 has no CPU fallback")
        _lib.check(_lib.lib().b200hnsw_merge_shards_device(
            self.gathered.device.index or 0, self.world, self.B, self.k, C.c_void_p(self.gathered.data_ptr()),
            self.stride, rows_per_shard, C.c_void_p(self.m_d.data_ptr()), C.c_void_p(self.m_i.data_ptr()),
            C.c_void_p(self.m_c.data_ptr()), C.c_void_p(stream) if stream else None))
        return self.m_d, self.m_i, self.m_c


class SegmentedResults:
    """Send/receive buffers of the all-to-all layout: rank r merges the queries [r*seg, (r+1)*seg) of every batch.
    `send` is what the search kernel writes (block s = this shard's results for the queries rank s merges: rows [seg,k] |
    dists [seg,k] | counts [seg]); after `exchange`, `recv` block g holds shard g's results for MY queries, which is the
    packed layout K4 reads.  Each rank moves (world-1)/world of one result set instead of receiving world-1 of them."""

    def __init__(self, B: int, k: int, world: int, rank: int, device):
        import torch
        self.B, self.k, self.world, self.rank = B, k, world, rank
        self.seg = (B + world - 1) // world
        self.stride = 2 * self.seg * k + self.seg
        self.send = torch.zeros(world * self.stride, dtype=torch.int32, device=device)  # counts of unused slots stay 0
        self.recv = torch.zeros(world * self.stride, dtype=torch.int32, device=device)
        self.q0 = min(B, rank * self.seg)
        self.q1 = min(B, (rank + 1) * self.seg)
        dev = self.send.device
        self.m_d = torch.empty((self.seg, k), dtype=torch.float32, device=dev)
        self.m_i = torch.empty((self.seg, k), dtype=torch.int64, device=dev)
        self.m_c = torch.empty((self.seg,), dtype=torch.int32, device=dev)

    def exchange(self, dist):
        if self.world == 1:
            self.recv.copy_(self.send)
        else:
            dist.all_to_all_single(self.recv, self.send)
        return self.recv

    def block_views(self, buf):
        """(rows [G,seg,k] i32, dists [G,seg,k] f32, counts [G,seg] i32) views of `send` or `recv` (for checks)."""
        import torch
        g = buf.view(self.world, self.stride)
        n = self.seg * self.k
        return (g[:, :n].reshape(self.world, self.seg, self.k),
                g[:, n:2 * n].view(torch.float32).reshape(self.world, self.seg, self.k), g[:, 2 * n:])

    def merge(self, rows_per_shard: int, stream: int = 0):
        """K4 over my query range: (dists [seg,k], global ids [seg,k], counts [seg]); rows beyond q1-q0 are padding."""
        if self.recv.device.type != "cuda":
            raise RuntimeError("SegmentedResults.merge needs CUDA tensors; hnsw_b200 has no CPU fallback")
        _lib.check(_lib.lib().b200hnsw_merge_shards_device(
            self.recv.device.index or 0, self.world, self.seg, self.k, C.c_void_p(self.recv.data_ptr()),
            self.stride, rows_per_shard, C.c_void_p(self.m_d.data_ptr()), C.c_void_p(self.m_i.data_ptr()),
            C.c_void_p(self.m_c.data_ptr()), C.c_void_p(stream) if stream else None))
        return self.m_d, self.m_i, self.m_c


class _CheckedRuns:
    """search_all: the visited-capacity retry of b200hnsw_search_batch for the pipelines, which drive the device-pointer
    entry points (those report a failed traversal in d_counters[4] instead of retrying, include/b200hnsw.h)."""

    def search_all(self, d_query_ptrs, before=None, after=None, max_rounds: int = 4):
        """Queue every batch of `d_query_ptrs` (device pointers), drain, then read the traversal error word on every
        rank.  If a traversal overflowed its visited set on ANY rank, every rank raises its graph's capacity (x4) and the
        whole list is issued again (traversals are deterministic: the same outputs are rewritten).  Collective when the
        pipeline has a process group.  Returns the slots of the batches, in order."""
        import torch
        if getattr(self, "_ctr", None) is None:
            self._ctr = torch.zeros(8, dtype=torch.int64, device=self.main.device)
        for _ in range(max_rounds):
            self._ctr.zero_()
            slots = [self.search(p, self._ctr.data_ptr(), before=before, after=after) for p in d_query_ptrs]
            self.drain()
            self.main.synchronize()
            err = self._ctr[4:5].to(torch.int32)
            if self.dist is not None and self.world > 1:
                codes = [torch.zeros_like(err) for _ in range(self.world)]
                self.dist.all_gather(codes, err)
                codes = [int(c.item()) for c in codes]
            else:
                codes = [int(err.item())]
            if not any(codes):
                return slots
            if any(c not in (0, 1) for c in codes):
                raise RuntimeError(f"a traversal failed with error {codes} (3: a query or row with a non-finite or extreme "
                                   f"norm; b200hnsw_search_batch handles those with the full candidate queue)")
            self.g.visited_capacity = self.g.visited_capacity * 4  # raises B200Error past what the device can give
        raise RuntimeError("search_all: traversals still overflow the visited set after raising it "
                           f"{max_rounds - 1} times")


class ShardPipeline(_CheckedRuns):
    """Sharded batched search as a pipeline.  `depth` traversal streams take the batches in turn (two traversal kernels
    in flight fill the SMs that one kernel's tail leaves idle: a 10k-query launch lasts as long as its slowest
    traversal), and the all-to-all + K4 merge of every batch run on ONE side stream (so the collectives are issued in
    the same order on every rank) under the traversals of the batches behind it.  2 * depth SegmentedResults buffers.
    `search` queues one batch and returns at once; `merged(slot)` are the tensors the merge of that slot writes for
    THIS rank's share of the queries (valid after `drain`, or after `done[slot]` has been waited on)."""

    def __init__(self, graph, dist, B: int, k: int, rows_per_shard: int, device, depth: int = 2):
        import torch
        self.g, self.dist, self.B, self.k, self.n = graph, dist, B, k, rows_per_shard
        self.world = dist.get_world_size() if dist is not None else 1
        self.rank = dist.get_rank() if dist is not None else 0
        self.depth = max(1, int(depth))
        self.nbuf = 2 * self.depth
        self.main = torch.cuda.current_stream(device)
        self.trav = [torch.cuda.Stream(device) for _ in range(self.depth)]
        # the exchange and the merge are short and on the critical path of slot reuse: their blocks go in front of the
        # traversal blocks that are waiting for an SM
        self.side = torch.cuda.Stream(device, priority=int(os.environ.get('B200HNSW_SIDE_PRIORITY', '-1')))
        self.buf = [SegmentedResults(B, k, self.world, self.rank, device) for _ in range(self.nbuf)]
        self.seg, self.q0, self.q1 = self.buf[0].seg, self.buf[0].q0, self.buf[0].q1
        self.searched = [torch.cuda.Event() for _ in range(self.nbuf)]
        self.done = [torch.cuda.Event() for _ in range(self.nbuf)]
        self.used = [False] * self.nbuf
        self.step = 0
        # optional per-stage timing (B200HNSW_PIPE_PROFILE=1): events around the exchange and the merge on the side stream
        self.profile = os.environ.get("B200HNSW_PIPE_PROFILE") == "1"
        self.prof = []

    def search(self, d_queries: int, d_counters: int = 0, timing=None, before=None, after=None) -> int:
        """Queue one batch (device pointer to [B, dim] f32); returns the slot its results will land in.
        timing = (start, end) CUDA events recorded around the traversal kernel on the stream it runs on;
        before(stream, slot) / after(stream, slot) queue the caller's own work in front of the traversal (e.g. the H2D
        copy of the queries) and behind the merge (e.g. the D2H copy of the merged lists)."""
        import torch
        slot = self.step % self.nbuf
        t = self.trav[self.step % self.depth]
        self.step += 1
        b = self.buf[slot]
        t.wait_stream(self.main)  # ordered behind whatever the caller queued on its stream (inputs, timing events)
        if self.used[slot]:
            t.wait_event(self.done[slot])  # the exchange + merge that last used this slot's buffers have finished
        if before is not None:
            with torch.cuda.stream(t):
                before(t, slot)
        if timing is not None:
            timing[0].record(t)
        self.g.search_segmented_device(d_queries, self.B, self.k, b.send.data_ptr(), b.seg, b.stride, d_counters,
                                       t.cuda_stream)
        if timing is not None:
            timing[1].record(t)
        self.searched[slot].record(t)
        self.side.wait_event(self.searched[slot])
        with torch.cuda.stream(self.side):
            if self.profile:
                e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
                e[0].record(self.side)
            b.exchange(self.dist)
            if self.profile:
                e[1].record(self.side)
            b.merge(self.n, self.side.cuda_stream)
            if self.profile:
                e[2].record(self.side)
                self.prof.append(e)
            if after is not None:
                after(self.side, slot)
        self.done[slot].record(self.side)
        self.used[slot] = True
        return slot

    def merged(self, slot: int):
        b = self.buf[slot]
        return b.m_d, b.m_i, b.m_c

    def drain(self):
        """Main stream waits for every queued merge."""
        for slot in range(self.nbuf):
            if self.used[slot]:
                self.main.wait_event(self.done[slot])

    def profile_summary(self):
        """Mean device time of the exchange and of the merge on the side stream (ms), over the steps profiled so far."""
        import torch
        torch.cuda.synchronize()
        if not self.prof:
            return None
        x = [e[0].elapsed_time(e[1]) for e in self.prof]
        m = [e[1].elapsed_time(e[2]) for e in self.prof]
        self.prof = []
        return {"exchange_ms": sum(x) / len(x), "merge_ms": sum(m) / len(m), "steps": len(x)}


class P2PShardPipeline(_CheckedRuns):
    """The sharded batched search with search and exchange FUSED: the traversal kernel writes every rank's share of its
    results straight into that rank's receive buffer (peer memory over NVLink / NVSwitch, mapped through CUDA IPC), so
    no collective and no extra kernel moves data.  What is left of the exchange is two flags per (slot, peer):

        ready[slot][src] = step + 1   set by src, in every peer, behind its traversal of `step`  (data has landed)
        freed[slot][dst] = step + 1   set by dst, in every peer, behind its merge of `step`      (slot may be reused)

    Per step i (slot = i mod nbuf), on one of `depth` traversal streams: wait freed[slot][*] >= i + 1 - nbuf, K1 with
    per-destination base pointers (b200hnsw_search_batch_scatter_device), signal ready[slot][me] = i + 1 everywhere;
    on the merge stream: wait ready[slot][*] >= i + 1, K4 over the local receive buffer, signal freed.  Same results as
    ShardPipeline (NCCL all-to-all); `merged(slot)`, `drain()`, `search(..., timing, before, after)` are the same."""

    def __init__(self, graph, dist, B: int, k: int, rows_per_shard: int, device, depth: int = 2):
        import torch
        self.g, self.dist, self.B, self.k, self.n = graph, dist, B, k, rows_per_shard
        self.world = dist.get_world_size() if dist is not None else 1
        self.rank = dist.get_rank() if dist is not None else 0
        if self.world > _MAX_PEERS:
            raise ValueError(f"at most {_MAX_PEERS} ranks")
        self.dev = torch.device(device)
        self.devi = self.dev.index or 0
        self.depth = max(1, int(depth))
        self.nbuf = int(os.environ.get("B200HNSW_PIPE_NBUF", 2 * self.depth))
        self.seg = (B + self.world - 1) // self.world
        self.stride = 2 * self.seg * k + self.seg          # int32 elements of one block (rows | dists | counts)
        self.q0, self.q1 = min(B, self.rank * self.seg), min(B, (self.rank + 1) * self.seg)
        # one peer-visible allocation per rank: nbuf receive buffers of world blocks, then the two flag arrays
        self.recv_elems = self.world * self.stride
        flag_off = self.nbuf * self.recv_elems
        self.flag_elems = self.nbuf * _MAX_PEERS
        total = (flag_off + 2 * self.flag_elems) * 4
        L = _lib.lib()
        mine, handle = C.c_void_p(), (C.c_uint8 * 64)()
        _lib.check(L.b200hnsw_peer_alloc(self.devi, total, C.byref(mine), handle))
        self._mine = mine.value
        handles = [None] * self.world
        if self.world > 1:
            dist.all_gather_object(handles, bytes(handle))
        else:
            handles[0] = bytes(handle)
        self._opened = []
        self.base = []
        problem = None
        try:
            if os.environ.get("B200HNSW_TEST_P2P_FAIL") == str(self.rank):
                raise RuntimeError("peer mapping refused (test hook)")
            for r in range(self.world):
                if r == self.rank:
                    self.base.append(self._mine)
                else:
                    ptr = C.c_void_p()
                    hb = (C.c_uint8 * 64).from_buffer_copy(handles[r])
                    _lib.check(L.b200hnsw_peer_open(self.devi, hb, C.byref(ptr)))
                    self._opened.append(ptr.value)
                    self.base.append(ptr.value)
        except Exception as e:  # e.g. no peer access between two devices
            problem = e
        if dist is not None and self.world > 1:
            # every rank learns whether EVERY rank mapped its peers, so that all of them give up together
            ok = torch.tensor([0 if problem else 1], dtype=torch.int32, device=device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if int(ok.item()) == 0 and problem is None:
                problem = RuntimeError("another rank could not map its peers' buffers")
        if problem is not None:
            for q in self._opened:
                L.b200hnsw_peer_close(C.c_void_p(q))
            if dist is not None and self.world > 1:
                dist.barrier()
            L.b200hnsw_peer_free(C.c_void_p(self._mine))
            self._mine = None
            raise RuntimeError(f"P2PShardPipeline: peer memory is not available ({problem})")
        self._flag_off = flag_off
        # host arrays of device pointers, one per slot: where my block lands in each rank, and the flags I set there
        self._seg_ptrs, self._ready_ptrs, self._freed_ptrs = [], [], []
        for slot in range(self.nbuf):
            VP = C.c_void_p * self.world
            self._seg_ptrs.append(VP(*[self.base[r] + 4 * (slot * self.recv_elems + self.rank * self.stride)
                                       for r in range(self.world)]))
            self._ready_ptrs.append(VP(*[self.base[r] + 4 * (flag_off + slot * _MAX_PEERS + self.rank)
                                         for r in range(self.world)]))
            self._freed_ptrs.append(VP(*[self.base[r] + 4 * (flag_off + self.flag_elems + slot * _MAX_PEERS + self.rank)
                                         for r in range(self.world)]))
        self.main = torch.cuda.current_stream(self.dev)
        self.trav = [torch.cuda.Stream(self.dev) for _ in range(self.depth)]
        self.side = torch.cuda.Stream(self.dev, priority=-1)
        self.m_d = [torch.empty((self.seg, k), dtype=torch.float32, device=self.dev) for _ in range(self.nbuf)]
        self.m_i = [torch.empty((self.seg, k), dtype=torch.int64, device=self.dev) for _ in range(self.nbuf)]
        self.m_c = [torch.empty((self.seg,), dtype=torch.int32, device=self.dev) for _ in range(self.nbuf)]
        self.searched = [torch.cuda.Event() for _ in range(self.nbuf)]
        self.done = [torch.cuda.Event() for _ in range(self.nbuf)]
        self.used = [False] * self.nbuf
        self.step = 0
        self.profile = False
        self.prof = []
        if dist is not None and self.world > 1:
            dist.barrier()  # every rank has opened every buffer before anyone writes

    # addresses inside my own allocation
    def _recv_ptr(self, slot):
        return self._mine + 4 * slot * self.recv_elems

    def _ready_flags(self, slot):
        return self._mine + 4 * (self._flag_off + slot * _MAX_PEERS)

    def _freed_flags(self, slot):
        return self._mine + 4 * (self._flag_off + self.flag_elems + slot * _MAX_PEERS)

    def search(self, d_queries: int, d_counters: int = 0, timing=None, before=None, after=None) -> int:
        import torch
        L = _lib.lib()
        i = self.step
        slot = i % self.nbuf
        t = self.trav[i % self.depth]
        self.step += 1
        t.wait_stream(self.main)
        if self.used[slot]:
            t.wait_event(self.done[slot])  # my own merge of the slot's previous step (its m_* tensors are rewritten)
        if i >= self.nbuf:  # every peer has merged what I wrote into this slot nbuf steps ago
            _lib.check(L.b200hnsw_peer_wait_device(self.devi, C.c_void_p(self._freed_flags(slot)), self.world,
                                                   i + 1 - self.nbuf, C.c_void_p(t.cuda_stream)))
        if before is not None:
            with torch.cuda.stream(t):
                before(t, slot)
        if timing is not None:
            timing[0].record(t)
        g = self.g
        _lib.check(L.b200hnsw_search_batch_scatter_device(
            g._h, C.c_void_p(d_queries), self.B, self.k, g.EfSearch, _metric_of(g.Distance), _lib.MODE_REFERENCE, None,
            self._seg_ptrs[slot], self.world, self.seg, C.c_void_p(d_counters) if d_counters else None,
            C.c_void_p(t.cuda_stream)))
        if timing is not None:
            timing[1].record(t)
        _lib.check(L.b200hnsw_peer_signal_device(self.devi, self._ready_ptrs[slot], self.world, i + 1,
                                                 C.c_void_p(t.cuda_stream)))
        self.searched[slot].record(t)
        self.side.wait_event(self.searched[slot])
        s = self.side
        _lib.check(L.b200hnsw_peer_wait_device(self.devi, C.c_void_p(self._ready_flags(slot)), self.world, i + 1,
                                               C.c_void_p(s.cuda_stream)))
        _lib.check(L.b200hnsw_merge_shards_device(
            self.devi, self.world, self.seg, self.k, C.c_void_p(self._recv_ptr(slot)), self.stride, self.n,
            C.c_void_p(self.m_d[slot].data_ptr()), C.c_void_p(self.m_i[slot].data_ptr()),
            C.c_void_p(self.m_c[slot].data_ptr()), C.c_void_p(s.cuda_stream)))
        _lib.check(L.b200hnsw_peer_signal_device(self.devi, self._freed_ptrs[slot], self.world, i + 1,
                                                 C.c_void_p(s.cuda_stream)))
        if after is not None:
            with torch.cuda.stream(s):
                after(s, slot)
        self.done[slot].record(s)
        self.used[slot] = True
        return slot

    def merged(self, slot: int):
        return self.m_d[slot], self.m_i[slot], self.m_c[slot]

    def drain(self):
        for slot in range(self.nbuf):
            if self.used[slot]:
                self.main.wait_event(self.done[slot])

    def profile_summary(self):
        return None

    def close(self):
        """Collective: every rank stops using the buffers, then unmaps and frees."""
        import torch
        if self._mine is None:
            return
        torch.cuda.synchronize(self.dev)
        if self.dist is not None and self.world > 1:
            self.dist.barrier()
        L = _lib.lib()
        for p in self._opened:
            L.b200hnsw_peer_close(C.c_void_p(p))
        self._opened = []
        if self.dist is not None and self.world > 1:
            self.dist.barrier()
        L.b200hnsw_peer_free(C.c_void_p(self._mine))
        self._mine = None
