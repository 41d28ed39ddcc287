"""Device-side measurements shared by the headline and the side workloads: build, device-resident search timing with
CUDA events, and the end-to-end host-pointer call (pinned or pageable buffers, one or several host callers).
Nothing here touches oracle/."""
from __future__ import annotations

import threading
import time

import numpy as np


def distance_of(hnsw_b200, metric):
    return hnsw_b200.EuclideanDistance if metric == "l2" else hnsw_b200.CosineDistance


def build_graph(hnsw_b200, X, levels, keys, m, ml, ef, metric, device, wave_max, build_mode=0):
    g = hnsw_b200.Graph(M=m, Ml=ml, EfSearch=ef, Distance=distance_of(hnsw_b200, metric), device=device,
                        capacity=len(X))
    if build_mode:
        g.set_build_mode(build_mode)  # NOT the reference: the wave Add runs the textbook layer search
    t0 = time.perf_counter()
    g.add_arrays(keys, X, levels, wave_max=wave_max)
    wall = time.perf_counter() - t0
    st = g.last_stats
    n = len(X)
    return g, {"vec_per_s": n / wall, "kernel_vec_per_s": n / (st["kernel_ms"] / 1e3), "wall_s": wall,
               "launches": st["launches"],
               "evals_per_insert": (st["dist_evals"] + st["evict_evals"] + st["replenish_evals"]) / n}


class ResidentSearch:
    """K steps of one BatchSearch each over query batches already in HBM, issued in turn on `inflight` streams (a
    10k-query launch lasts as long as its slowest traversal; the next launch fills the SMs its tail leaves idle).
    Every step is one K1 launch over one batch.  Timed with CUDA events on the launching streams."""

    def __init__(self, torch, g, dev, B, k, ef, mode, inflight):
        self.torch, self.g, self.dev, self.B, self.k, self.ef, self.mode = torch, g, dev, B, k, ef, mode
        self.depth = max(1, inflight)
        self.main = torch.cuda.current_stream(dev)
        self.streams = [torch.cuda.Stream(dev) for _ in range(self.depth)]
        self.outs = [(torch.empty((B, k), dtype=torch.int32, device=dev),
                      torch.empty((B, k), dtype=torch.float32, device=dev),
                      torch.empty((B,), dtype=torch.int32, device=dev)) for _ in range(self.depth)]
        self.counters = torch.zeros(8, dtype=torch.int64, device=dev)

    def step(self, i, qd, t0=None, t1=None):
        ts = self.streams[i % self.depth]
        r, d_, c = self.outs[i % self.depth]
        ts.wait_stream(self.main)
        if t0 is not None:
            t0.record(ts)
        self.g.search_device(qd.data_ptr(), self.B, self.k, r.data_ptr(), d_.data_ptr(), c.data_ptr(),
                             self.counters.data_ptr(), ts.cuda_stream, ef=self.ef, mode=self.mode)
        if t1 is not None:
            t1.record(ts)

    def join(self):
        for ts in self.streams:
            self.main.wait_stream(ts)

    def run(self, Qd, steps, warmup, barrier):
        """Qd [nb, B, dim] on the device.  Returns the timing record; `barrier()` = dist.barrier + synchronize."""
        torch = self.torch
        nb = Qd.shape[0]
        for i in range(warmup):
            self.step(i, Qd[i % nb])
        self.join()
        barrier()
        self.counters.zero_()
        ka = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        kb = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record(self.main)
        t_issue = time.perf_counter()
        for i in range(steps):
            self.step(i, Qd[(warmup + i) % nb], ka[i], kb[i])
        t_issue = (time.perf_counter() - t_issue) / steps * 1e3
        self.join()
        ev1.record(self.main)
        barrier()
        ctr = self.counters.cpu().numpy().copy()
        lo = self.outs[(steps - 1) % self.depth]
        return {"total_ms": ev0.elapsed_time(ev1), "kernel_ms": [a.elapsed_time(b) for a, b in zip(ka, kb)],
                "ctr": ctr, "host_issue_ms_per_step": t_issue, "last_batch": (warmup + steps - 1) % nb,
                "last": (lo[0].cpu().numpy(), lo[1].cpu().numpy(), lo[2].cpu().numpy())}

    def solo(self, Qd, reps=8):
        """The same launch with nothing else in flight (mean event time of reps - 2 launches)."""
        torch = self.torch
        nb = Qd.shape[0]
        sa = [torch.cuda.Event(enable_timing=True) for _ in range(reps)]
        sb = [torch.cuda.Event(enable_timing=True) for _ in range(reps)]
        r, d_, c = self.outs[0]
        for i in range(reps):
            sa[i].record(self.main)
            self.g.search_device(Qd[i % nb].data_ptr(), self.B, self.k, r.data_ptr(), d_.data_ptr(), c.data_ptr(), 0,
                                 self.main.cuda_stream, ef=self.ef, mode=self.mode)
            sb[i].record(self.main)
        torch.cuda.synchronize(self.dev)
        return float(np.mean([x.elapsed_time(y) for x, y in zip(sa, sb)][2:]))


def e2e_host_calls(torch, g, Qh, B, k, ef, mode, steps, warmup, callers, pinned, passes=3):
    """The drop-in call itself, b200hnsw_search_batch with HOST buffers (H2D copy, traversal, D2H all inside), made by
    `callers` host threads at once (the reference admits concurrent BatchSearch callers under its RWMutex read lock,
    graph.go:1048).  Qh: torch tensor [nb, B, dim] (pinned) or numpy array (pageable).  Returns (seconds for `steps`
    calls: median over `passes`, every pass, the last caller's outputs)."""
    nb = Qh.shape[0]

    def bufs():
        if pinned:
            return (torch.empty((B, k), dtype=torch.int32).pin_memory(),
                    torch.empty((B, k), dtype=torch.float32).pin_memory(),
                    torch.empty((B,), dtype=torch.int32).pin_memory())
        return (np.empty((B, k), np.int32), np.empty((B, k), np.float32), np.empty((B,), np.int32))

    def ptr(x):
        return x.data_ptr() if hasattr(x, "data_ptr") else x.ctypes.data

    outs = [bufs() for _ in range(callers)]
    gate = threading.Barrier(callers + 1)
    errors = []

    def call(t, i):
        r, d, c = outs[t]
        g.search_host_ptrs(ptr(Qh[i % nb]), B, k, ptr(r), ptr(d), ptr(c), ef=ef, mode=mode)

    def caller(t):
        try:
            gate.wait()
            # warm-up with every caller in flight at once: the library creates a call context (streams, scratch) the
            # first time a call finds every existing one busy, which must not happen inside the timed region
            for i in range(max(warmup, 3) + 3):
                call(t, t + i)
            for _ in range(passes):
                gate.wait()
                for i in range(t, steps, callers):
                    call(t, warmup + i)
                gate.wait()
        except BaseException as e:  # noqa: BLE001 - reported by the main thread
            errors.append(e)
            gate.abort()

    th = [threading.Thread(target=caller, args=(t,)) for t in range(callers)]
    for x in th:
        x.start()
    times = []
    try:
        gate.wait()
        for _ in range(passes):
            gate.wait()
            t0 = time.perf_counter()
            gate.wait()
            times.append(time.perf_counter() - t0)
    except threading.BrokenBarrierError:
        pass
    for x in th:
        x.join()
    if errors:
        raise errors[0]
    tl = (steps - 1) % callers
    last = tuple(np.array(x.numpy() if hasattr(x, "numpy") else x) for x in outs[tl])
    return float(np.median(times)), times, last
