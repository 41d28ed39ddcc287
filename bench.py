#!/usr/bin/env python
"""bench.py -- batched Graph.Search QPS on 1M x 128 L2 (BASELINE.json configs[1]) + build rate, on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun ... bench.py --gpus N ...   (one rank per GPU; RANK / LOCAL_RANK / WORLD_SIZE from the env)

A "step" is one BatchSearch of `--batch` synthetic queries (k=10, EfSearch=64) against the device-resident graph.
  value      : whole-job queries/s, queries already resident in HBM, CUDA events around the K steps, max over ranks
  e2e        : the same through the reference-facing C-ABI call with HOST buffers (H2D + kernel + D2H per step)
  roofline   : traversal kernel's algorithmic bytes (E*4d + neighbour ids + query + results) / its event time,
               against the measured HBM copy bandwidth in MEASURED_PEAKS.json
  cpu_baseline: the oracle (CPU port of the reference's Go algorithm) on all host cores, same graph, same queries
  --impl reference: only the CPU port is built and timed (no GPU code on that path)

Steps are issued in turn on --inflight traversal streams (default 3): every step is one K1 launch over one batch, and
the launches in flight fill the SMs that a single launch's tail leaves idle.  e2e: --e2e-callers host threads each
make the synchronous host-pointer call (the reference admits concurrent BatchSearch callers under its read lock).

Multi-GPU (north_star): every rank owns an independent sub-graph of --n rows (weak scaling), the query batch is
the same on all ranks; the traversal kernel stores each rank's share of the per-shard top-k lists straight into that
rank's receive buffer over NVLink (--exchange p2p; `nccl` = one all-to-all instead) and the K4 kernel merges them.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=1_000_000, help="rows per GPU (BASELINE configs[1]: 1M)")
    ap.add_argument("--dim", type=int, default=128)
    ap.add_argument("--batch", type=int, default=10_000)
    ap.add_argument("--m", type=int, default=16)
    ap.add_argument("--ef", type=int, default=64)
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--ml", type=float, default=0.25)
    ap.add_argument("--wave-max", type=int, default=16384,
                    help="nodes per construction wave (b200hnsw_add_batch_waves); 0 = exact sequential Graph.Add")
    ap.add_argument("--pool", type=int, default=16, help="distinct query batches cycled through the steps")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="target length of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-exact", action="store_true", help="skip the ExactIndex (K3) side measurement")
    ap.add_argument("--exact-steps", type=int, default=10)
    ap.add_argument("--inflight", type=int, default=3,
                    help="batches in flight in the device-resident measurement (traversal streams taking the steps in "
                         "turn); 1 = strictly one launch after the other")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"],
                    help="N > 1: how per-shard results reach the rank that merges them: p2p = the traversal kernel "
                         "stores them into peer memory itself (fused search + exchange); nccl = one all-to-all")
    ap.add_argument("--e2e-callers", type=int, default=4,
                    help="host threads making the synchronous host-pointer call concurrently in the e2e measurement "
                         "(the reference admits concurrent BatchSearch callers under its RWMutex read lock)")
    return ap.parse_args()


def synth_rows(n, dim, seed):
    # graph_benchmark_test.go:12-18: rand.Float32()*2-1, i.e. uniform [-1, 1)
    return np.random.default_rng(seed).uniform(-1.0, 1.0, (n, dim)).astype(np.float32)


def draw_levels(n, ml, seed):
    """graph.go:388-417 with a numpy stream: level draws for n consecutive insertions into an empty graph."""
    rng = np.random.default_rng(seed)
    go = rng.random((n, 32)) <= ml  # the loop continues while r <= Ml (graph.go:411 stops on r > Ml)
    run = np.where(go.all(axis=1), go.shape[1], go.argmin(axis=1))  # leading run of "continue" = level
    size0 = np.arange(n, dtype=np.float64)  # len(layer 0) seen by the i-th insertion
    with np.errstate(divide="ignore"):
        mx = np.floor(np.log(np.maximum(size0, 1.0)) / np.log(1.0 / ml) + 0.5).astype(np.int64) + 1  # maxLevel
    mx[0] = 1  # graph.go:395: no layers yet -> max = 1 (and maxLevel(ml, 0) == 1 as well)
    return np.minimum(run, mx).astype(np.int32)


class ClockSampler(threading.Thread):
    """Samples SM clock / power / throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index, period_s=0.002):
        super().__init__(daemon=True)
        self.index, self.period, self.rows, self._stop_ev = index, period_s, [], threading.Event()
        self.max_mhz, self.ok = None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception:
            self.ok = False

    def sample(self):
        nv = self.nv
        mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
        try:
            reasons = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:
            reasons = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        try:
            power = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
        except Exception:
            power = None
        self.rows.append((float(mhz), int(reasons), power))

    def run(self):
        if not self.ok:
            return
        errors = 0
        while not self._stop_ev.is_set() and errors < 100:
            try:
                self.sample()
            except Exception:
                errors += 1  # a transient NVML error must not end the sampling of the timed region
            time.sleep(self.period)

    def stop(self):
        self._stop_ev.set()
        self.join(timeout=2)
        if not self.ok or not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        nv = self.nv
        bits = {"hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
                "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
                "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4)}
        seen = set()
        for _, r, _ in self.rows:
            for name, bit in bits.items():
                if r & bit:
                    seen.add(name)
        pw = [p for _, _, p in self.rows if p is not None]
        return {"sm_mhz": float(np.median([m for m, _, _ in self.rows])), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(seen), "samples": len(self.rows), "power_w_max": max(pw) if pw else None}


def load_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def load_traffic():
    """dram bytes per launch of the traversal kernel from the committed ncu --set full capture, if any."""
    p = ROOT / "profiles" / "k1_traffic.json"
    if p.exists():
        try:
            return json.loads(p.read_text())
        except Exception:
            return None
    return None


def algorithmic_bytes(evals, nbr_reads, queries, dim, k):
    """E rows of 4*dim bytes + the neighbour ids read + the query + the (id, dist) results."""
    return evals * 4 * dim + nbr_reads * 4 + queries * (4 * dim + 8 * k + 4)


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args, rank, world):
    """The reference's own CPU path (its C port: oracle/), all host threads, same config.  No GPU code here."""
    if rank != 0:
        return 0
    from oracle import binding as oracle
    oracle.build()
    cores = os.cpu_count() or 1
    X = synth_rows(args.n, args.dim, 42)
    levels = draw_levels(args.n, args.ml, 44)
    keys = np.arange(args.n, dtype=np.int64)
    # Graph.Add exactly as the reference does it (serial, graph.go:445) on a bounded sample: the build rate reported
    ns = min(args.n, 20000)
    gs = oracle.Graph(args.dim, args.m, oracle.EUCLIDEAN, oracle.SIMD)
    t0 = time.perf_counter()
    gs.add_many(keys[:ns], X[:ns], levels[:ns], args.ef)
    serial_rate = ns / (time.perf_counter() - t0)
    del gs
    # the graph that is searched: the same one our arm builds (wave schedule, deterministic), built on all cores
    g = oracle.Graph(args.dim, args.m, oracle.EUCLIDEAN, oracle.SIMD)
    t0 = time.perf_counter()
    if args.wave_max > 0:
        g.add_waves(keys, X, levels, args.ef, args.wave_max, n_threads=cores)
    else:
        g.add_many(keys, X, levels, args.ef)
    build_s = time.perf_counter() - t0
    nb = min(args.warmup + args.steps, args.pool)
    Q = synth_rows(nb * args.batch, args.dim, 43).reshape(nb, args.batch, args.dim)
    for i in range(args.warmup):
        g.batch_search(Q[i % nb], args.k, args.ef, n_threads=cores)
    t0 = time.perf_counter()
    for i in range(args.steps):
        g.batch_search(Q[(args.warmup + i) % nb], args.k, args.ef, n_threads=cores)
    dt = time.perf_counter() - t0
    qps = args.steps * args.batch / dt
    line = {
        "impl": "reference", "metric": "batched Search QPS (Graph.BatchSearch, k=10)", "value": qps,
        "unit": "queries/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, 1),
        "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": cores, "kind": "port",
                         "sample": f"{args.steps} x {args.batch} queries, one query per thread over {cores} threads, "
                                   f"AVX2 distance; graph = the wave-scheduled construction our arm uses, built by the "
                                   f"port on {cores} threads in {build_s:.1f} s; the reference's serial Graph.Add runs "
                                   f"at {serial_rate:.0f} vec/s here (first {ns} rows)"},
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "build_vec_per_s": serial_rate, "gpu_launches": 0,
    }
    emit(line)
    return 0


def workload_config(args, world):
    return {"workload": f"{args.n}x{args.dim} uniform[-1,1) f32 rows per GPU, L2 (EuclideanDistance), M={args.m} "
                        f"Ml={args.ml} EfSearch={args.ef}, {args.batch}-query batch, k={args.k} "
                        f"(BASELINE.json configs[1])",
            "rows_per_gpu": args.n, "dim": args.dim, "M": args.m, "EfSearch": args.ef, "k": args.k,
            "batch": args.batch, "mode": "reference (bit-faithful layerNode.search termination)",
            "layout": "1 graph" if world == 1 else
            f"{world} independent sub-graphs (one per GPU), queries broadcast; rank r merges queries [r*B/{world}, "
            f"(r+1)*B/{world}): " + (
                "the traversal kernel stores each rank's share of its results into that rank's receive buffer over "
                "NVLink (CUDA IPC peer memory; ready/freed flags, no collective)" if args.exchange == "p2p" else
                "one NCCL all-to-all of the per-shard result blocks") +
            " + K4 top-k merge on a side stream under the traversals of the next steps (all inside the timed region)",
            "batches_in_flight": max(1, args.inflight),
            "pipelining": f"steps are issued in turn on {max(1, args.inflight)} traversal stream(s): each step is one K1 "
                          f"launch over one {args.batch}-query batch, at most {max(1, args.inflight)} of them in flight",
            "l2_policy": f"inputs larger than L2: {args.n * (args.dim * 4 + args.m * 4) / 1e6:.0f} MB of graph per GPU "
                         f"gathered at random; a pool of {args.pool} distinct query batches cycled through the steps"}


# ------------------------------------------------------------------------------------------------------ our arm
def run_ours(args, rank, local_rank, world):
    import ctypes as C

    import torch

    import hnsw_b200
    from hnsw_b200 import _lib, sharded

    _lib.lib()  # fail loudly if the CUDA extension is missing
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; hnsw_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    n, dim, B, k, ef = args.n, args.dim, args.batch, args.k, args.ef

    # ---- build this rank's sub-graph on the device -------------------------------------------------------------------
    X = synth_rows(n, dim, 42 + rank)
    levels = draw_levels(n, args.ml, 44 + rank)
    g = hnsw_b200.Graph(M=args.m, Ml=args.ml, EfSearch=ef, Distance=hnsw_b200.EuclideanDistance, device=local_rank,
                        capacity=n)
    keys = np.arange(n, dtype=np.int64) + rank * n  # global ids
    t0 = time.perf_counter()
    g.add_arrays(keys, X, levels, wave_max=args.wave_max)
    build_wall = time.perf_counter() - t0
    bst = g.last_stats
    build_kernel_s = bst["kernel_ms"] / 1e3
    seq_rate = None
    if rank == 0 and world == 1:  # the reference's exact sequential order on a bounded sample (K2, insert_kernel.cu)
        ns = min(n, 4000)
        gs = hnsw_b200.Graph(M=args.m, Ml=args.ml, EfSearch=ef, Distance=hnsw_b200.EuclideanDistance,
                             device=local_rank, capacity=ns)
        gs.add_arrays(keys[:ns], X[:ns], levels[:ns])
        seq_rate = ns / (gs.last_stats["kernel_ms"] / 1e3)
        gs.close()

    # ---- synthetic queries: a pool of distinct batches cycled through the steps, identical on every rank ----------
    nb = max(1, min(args.warmup + args.steps, args.pool))
    Qh = torch.from_numpy(synth_rows(nb * B, dim, 43).reshape(nb, B, dim)).pin_memory()
    Qd = Qh.to(dev)
    counters = torch.zeros(8, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream(dev)
    # `depth` batches in flight: a 10k-query launch lasts as long as its slowest traversal and leaves SMs idle towards
    # its end; the next batch's launch, on a second stream, fills them.  Every step is still one launch over one batch.
    depth = max(1, args.inflight)
    tstreams = [torch.cuda.Stream(dev) for _ in range(depth)]
    outs_d = [(torch.empty((B, k), dtype=torch.int32, device=dev), torch.empty((B, k), dtype=torch.float32, device=dev),
               torch.empty((B,), dtype=torch.int32, device=dev)) for _ in range(depth)]

    def search_step(i, qd, t0=None, t1=None):
        """step i on traversal stream i % depth (N = 1): b200hnsw_search_batch_device, one K1 launch"""
        ts = tstreams[i % depth]
        r, d_, c = outs_d[i % depth]
        ts.wait_stream(stream)
        if t0 is not None:
            t0.record(ts)
        g.search_device(qd.data_ptr(), B, k, r.data_ptr(), d_.data_ptr(), c.data_ptr(), counters.data_ptr(),
                        ts.cuda_stream)
        if t1 is not None:
            t1.record(ts)

    def join():
        for ts in tstreams:
            stream.wait_stream(ts)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # N > 1: the all-to-all + merge of a batch run on a side stream under the traversals of the batches behind it
    pipe = None
    if world > 1:
        # p2p: K1 stores each rank's share of its results into that rank's receive buffer over NVLink (search and
        # exchange are one kernel, flags instead of a collective); nccl: K1 + one NCCL all-to-all
        if args.exchange == "p2p":
            try:
                pipe = sharded.P2PShardPipeline(g, dist, B, k, n, dev, depth=depth)
            except RuntimeError as e:  # raised on every rank together; the collective variant needs no peer mapping
                print(f"bench.py: {e}; using the NCCL all-to-all exchange", file=sys.stderr)
                args.exchange = "nccl"
        if pipe is None:
            pipe = sharded.ShardPipeline(g, dist, B, k, n, dev, depth=depth)

    # NVML serialises its callers system-wide: rank 0 (whose samples go into the JSON line) samples every 2 ms, the
    # other ranks every 20 ms
    sampler = ClockSampler(local_rank, 0.002 if rank == 0 else 0.02)
    sampler.start()

    # ---- device-resident timing: W warm-up steps, then exactly K steps between events ---------------------------
    for i in range(args.warmup):
        if world > 1:
            pipe.search(Qd[i % nb].data_ptr(), counters.data_ptr())
        else:
            search_step(i, Qd[i % nb])
    if world > 1:
        pipe.drain()
    else:
        join()
    barrier()
    if pipe is not None:
        pipe.prof = []
    counters.zero_()
    ka = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    kb = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    t_issue = time.perf_counter()
    for i in range(args.steps):
        if world > 1:
            pipe.search(Qd[(args.warmup + i) % nb].data_ptr(), counters.data_ptr(), timing=(ka[i], kb[i]))
        else:
            search_step(i, Qd[(args.warmup + i) % nb], ka[i], kb[i])
    t_issue = (time.perf_counter() - t_issue) / args.steps * 1e3  # host time to queue one step (launch-bound check)
    if world > 1:
        pipe.drain()  # the timed region ends when the last merge has finished
    else:
        join()
    ev1.record(stream)
    barrier()
    pipe_prof = pipe.profile_summary() if pipe is not None and pipe.profile else None
    total_ms = ev0.elapsed_time(ev1)
    kernel_ms = [a.elapsed_time(b) for a, b in zip(ka, kb)]
    ctr = counters.cpu().numpy().copy()
    if int(ctr[4]) != 0:
        raise SystemExit("bench.py: a traversal overflowed the visited set in the device-pointer path")
    last = (args.warmup + args.steps - 1) % nb
    lo = outs_d[(args.steps - 1) % depth]
    last_rows, last_dists, last_counts = lo[0].cpu().numpy(), lo[1].cpu().numpy(), lo[2].cpu().numpy()
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    # the same launch alone on the device (nothing else in flight): the per-kernel figure earlier rounds reported
    solo_ms = None
    if world == 1:
        sa = [torch.cuda.Event(enable_timing=True) for _ in range(8)]
        sb = [torch.cuda.Event(enable_timing=True) for _ in range(8)]
        r, d_, c = outs_d[0]
        for i in range(8):
            sa[i].record(stream)
            g.search_device(Qd[i % nb].data_ptr(), B, k, r.data_ptr(), d_.data_ptr(), c.data_ptr(), 0,
                            stream.cuda_stream)
            sb[i].record(stream)
        torch.cuda.synchronize(dev)
        solo_ms = float(np.mean([x.elapsed_time(y) for x, y in zip(sa, sb)][2:]))

    # ---- end to end: host buffers in, host buffers out, every step ---------------------------------------------------
    # NVML queries take driver locks that the CUDA calls of the host threads below also need; the device-timed region
    # above was sampled every 2 ms, the end-to-end regions (synchronous calls, host in the loop) every 20 ms
    sampler.period = float(os.environ.get("B200_BENCH_E2E_SAMPLE_PERIOD", "0.02"))
    h_rows = torch.empty((B, k), dtype=torch.int32).pin_memory()
    h_dists = torch.empty((B, k), dtype=torch.float32).pin_memory()
    h_counts = torch.empty((B,), dtype=torch.int32).pin_memory()
    if world == 1:
        def e2e_step(i):  # the drop-in call itself: b200hnsw_search_batch(host pointers), one caller, one call at a time
            g.search_host_ptrs(Qh[i].data_ptr(), B, k, h_rows.data_ptr(), h_dists.data_ptr(), h_counts.data_ptr())

        def e2e_finish():
            pass
    else:
        # every rank receives the whole batch from ITS host (H2D in front of the traversal, on the traversal's stream),
        # searches its shard, and returns the merged lists of its share of the queries (B / world of them) to page-
        # locked host memory behind the merge (side stream).  The host takes a slot's results (waits for that copy)
        # before it queues the slot again, nbuf steps later: the serving loop a caller of ShardPipeline runs.
        seg = pipe.seg
        nbuf = pipe.nbuf
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

        def e2e_finish():
            for sl_ in range(nbuf):
                if queued[sl_]:
                    back[sl_].synchronize()

    for i in range(max(args.warmup, 3)):
        e2e_step(i % nb)
    e2e_finish()
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        e2e_step((args.warmup + i) % nb)
    e2e_finish()
    torch.cuda.synchronize(dev)
    e2e_s = time.perf_counter() - t0
    e2e_single_s = e2e_s
    # The reference serves concurrent BatchSearch callers under its read lock (graph.go:1048, sync.RWMutex); so does the
    # C ABI (a call context with its own streams per call in flight).  With `callers` host threads each making the same
    # synchronous b200hnsw_search_batch(host pointers) call, one call's H2D copy runs under another call's traversal
    # kernel.  Every step still copies its own queries in and its own results out inside the timed region.
    callers = max(1, args.e2e_callers) if world == 1 else 1
    E2E_PASSES = 3
    e2e_passes = []
    if callers > 1:
        outs = [(torch.empty((B, k), dtype=torch.int32).pin_memory(), torch.empty((B, k), dtype=torch.float32).pin_memory(),
                 torch.empty((B,), dtype=torch.int32).pin_memory()) for _ in range(callers)]
        gate = threading.Barrier(callers + 1)

        def caller(t):
            r, d, c = outs[t]
            # warm-up with all callers in flight at once: the library creates a call context (streams, scratch) the
            # first time a call finds every existing context busy, which must not happen inside the timed region
            gate.wait()
            for i in range(max(args.warmup, 3) + 3):
                g.search_host_ptrs(Qh[(t + i) % nb].data_ptr(), B, k, r.data_ptr(), d.data_ptr(), c.data_ptr())
            for _ in range(E2E_PASSES):
                gate.wait()
                for i in range(t, args.steps, callers):
                    g.search_host_ptrs(Qh[(args.warmup + i) % nb].data_ptr(), B, k, r.data_ptr(), d.data_ptr(),
                                       c.data_ptr())
                gate.wait()

        th = [threading.Thread(target=caller, args=(t,)) for t in range(callers)]
        for x in th:
            x.start()
        gate.wait()  # warm-up starts
        for _ in range(E2E_PASSES):
            gate.wait()  # everyone is ready: a timed pass of exactly K steps starts
            t0 = time.perf_counter()
            gate.wait()
            e2e_passes.append(time.perf_counter() - t0)
        for x in th:
            x.join()
        e2e_s = float(np.median(e2e_passes))  # every pass is reported (e2e.passes); the median is the value
        # the last step's results, as the last caller received them, must be the device-timed run's
        tl = (args.steps - 1) % callers
        if not (np.array_equal(outs[tl][0].numpy(), last_rows) and np.array_equal(outs[tl][2].numpy(), last_counts)):
            raise SystemExit("bench.py: concurrent host-pointer calls returned different rows than the device-pointer call")
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    clocks = sampler.stop()

    # ---- numbers ----------------------------------------------------------------------------------------------------
    units = args.steps * B * world  # (query, shard) traversals done by all ranks
    value = units / (total_ms / 1e3)
    e2e_value = units / e2e_s
    peak, peak_src = load_peaks()
    kern_avg_ms = float(np.mean(kernel_ms))
    alg_bytes = algorithmic_bytes(int(ctr[0]), int(ctr[2]), args.steps * B, dim, k) / args.steps
    # With several launches in flight the event time of ONE launch on its stream (kern_avg_ms) counts the time its
    # blocks wait behind, and share the SMs with, the other launches; the time the device spends per launch over the
    # timed region is total_ms / steps.  `achieved` uses the latter (launch durations that add up to the region); the
    # event time and the launch-alone figure are reported beside it.
    eff_ms = (total_ms / args.steps) if depth > 1 else kern_avg_ms
    achieved = alg_bytes / (eff_ms / 1e3) / 1e9
    traffic = load_traffic()
    line = {
        "metric": "batched Search QPS (Graph.BatchSearch, k=10)", "value": value, "unit": "queries/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
        "host_issue_ms_per_step": t_issue, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, world),
        "e2e": {"value": e2e_value, "unit": "queries/s", "h2d_bytes_per_step": B * dim * 4,
                "d2h_bytes_per_step": B * k * 8 + B * 4 + 40 if world == 1 else pipe.seg * (k * 12 + 4),
                "callers": callers, "single_caller_value": units / e2e_single_s,
                "passes": [units / x for x in e2e_passes] if e2e_passes else None,
                "call": (f"b200hnsw_search_batch (host pointers, synchronous) from {callers} host threads, median of "
                         f"{E2E_PASSES} passes of {args.steps} steps (all in `passes`); single_caller_value = one "
                         "thread, one call at a time") if world == 1 else
                        (f"ShardPipeline.search per rank with the H2D copy of the batch in front of the traversal and "
                         f"the D2H copy of the merged lists behind the merge, {pipe.nbuf} batches in flight")},
        "gpu_launches": args.steps * (1 if world == 1 else 2),  # K1 (+ K4 merge); the all-gather is NCCL's
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": (traffic or {}).get("dram_bytes_per_launch"),
                     "kernel": "b200::traverse_kernel (K1, traverse.cuh)", "kernel_ms": eff_ms,
                     "kernel_ms_definition": (f"timed region / launches ({depth} launches in flight on {depth} streams); "
                                              "event time of one launch on its own stream: event_kernel_ms") if depth > 1
                     else "CUDA events around each launch",
                     "event_kernel_ms": kern_avg_ms,
                     "algorithmic_bytes_per_launch": alg_bytes, "peak_source": peak_src,
                     "dist_evals_per_query": float(ctr[0]) / (args.steps * B),
                     "expansions_per_query": float(ctr[1]) / (args.steps * B),
                     # frac can pass 1: the bytes are ALGORITHMIC (E rows of 4*dim bytes per query) and `traffic` shows
                     # that ~90% of them are served by L2 (the traversals of a batch share the upper layers and the hub
                     # rows); the kernel's real bound is traversal latency x resident traversals.
                     # event_frac = one launch's bytes / its event time while sharing the SMs; solo = the launch alone.
                     "note": ("bytes are algorithmic (SURVEY 8d); frac > 1 means L2 reuse, not an error: see traffic "
                              "(DRAM bytes of one cold launch, ncu) and DESIGN.md 4/K1 (the kernel is bound by traversal "
                              "latency x resident traversals)"),
                     "launches_in_flight": depth,
                     "event_frac": alg_bytes / (kern_avg_ms / 1e3) / 1e9 / peak,
                     "solo_kernel_ms": solo_ms,
                     "solo_achieved": (alg_bytes / (solo_ms / 1e3) / 1e9) if solo_ms else None,
                     "solo_frac": (alg_bytes / (solo_ms / 1e3) / 1e9 / peak) if solo_ms else None},
        "clocks": clocks,
        "build": {"vec_per_s": n * world / build_wall, "kernel_vec_per_s": n / build_kernel_s, "wall_s": build_wall,
                  "semantics": (f"wave-batched Add, waves of <= {args.wave_max} nodes (b200hnsw_add_batch_waves; "
                                "bit-identical to the oracle's restatement of the same schedule, NOT the reference's "
                                "sequential graph)") if args.wave_max > 0 else
                  "exact sequential Graph.Add order (bit-identical to the oracle)",
                  "launches": bst["launches"],
                  "evals_per_insert": (bst["dist_evals"] + bst["evict_evals"] + bst["replenish_evals"]) / n,
                  "exact_sequential_vec_per_s": seq_rate},
    }

    if pipe_prof:
        line["pipeline_profile"] = pipe_prof
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line.update(cpu_baseline_and_parity(args, g, Qh, last, last_rows, last_dists, last_counts, torch, dev, X))
    if rank == 0 and world == 1 and not args.no_exact:
        line["exact"] = exact_measurement(args, X, Qh, Qd, torch, dev, stream, not args.no_cpu_baseline)
        # BASELINE.json words its metric "QPS @ recall@10 >= 0.95".  The reference's graph search cannot reach that on
        # this data; of the paths the reference has, its hybrid package's exact index is the one
        # that does, so that clause is answered by K3 on the same rows and queries, reported beside the headline.
        ex = line["exact"]
        line["at_recall_ge_0.95"] = {"path": "hybrid.ExactIndex.Search (K3, exact: recall@10 = 1.0 by construction, "
                                             "results bit-identical to the oracle)",
                                     "value": ex["value"], "unit": ex["unit"], "e2e": ex["e2e"]["value"],
                                     "recall_at_10": 1.0}
    if rank == 0:
        emit(line)
    if world > 1:
        if hasattr(pipe, "close"):
            pipe.close()  # collective: unmap the peers' receive buffers, then free mine
        dist.barrier()
        dist.destroy_process_group()
    return 0


def cpu_baseline_and_parity(args, g, Qh, last, last_rows, last_dists, last_counts, torch, dev, X):
    """Rank 0, N=1: time the CPU port on the SAME graph (read back from the device) and check the GPU's last batch
    against it.  The oracle is the checker and the reported baseline here, never the thing measured above."""
    from oracle import binding as oracle
    oracle.build()
    cores = os.cpu_count() or 1
    flat = g.dump_flat()
    B, k, ef = args.batch, args.k, args.ef
    out = {}
    # parity: canonical-order oracle vs the device results of the last timed step (bit-exact bar)
    oc = oracle.Graph(args.dim, args.m, oracle.EUCLIDEAN, oracle.CANONICAL)
    oc.load_flat(*flat)
    o_rows, o_d, o_cnt, o_ctr = oc.batch_search(Qh[last].numpy(), k, ef, n_threads=cores)
    out["parity"] = {"checked_queries": int(B), "rows_equal": bool(np.array_equal(o_rows, last_rows)),
                     "dists_bit_equal": bool(np.array_equal(o_d.view(np.uint32), last_dists.view(np.uint32))),
                     "counts_equal": bool(np.array_equal(o_cnt, last_counts))}
    del oc
    # recall@10 of the bit-faithful mode against float32 brute force on the GPU (measurement only)
    nq = min(1000, B)
    q = Qh[last][:nq].to(dev)
    xd = torch.from_numpy(X).to(dev)
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    d2 = (xd * xd).sum(1)[None, :] - 2.0 * (q @ xd.T)
    torch.backends.cuda.matmul.allow_tf32 = prev
    truth = d2.topk(k, dim=1, largest=False).indices.cpu().numpy()
    hits = sum(len(set(int(v) for v in last_rows[i] if v >= 0) & set(int(v) for v in truth[i])) for i in range(nq))
    out["recall_at_10"] = hits / (nq * k)
    del xd, d2
    # timing: AVX2-distance flavour, one query per thread on all cores, bounded sample
    os_ = oracle.Graph(args.dim, args.m, oracle.EUCLIDEAN, oracle.SIMD)
    os_.load_flat(*flat)
    nb = Qh.shape[0]
    os_.batch_search(Qh[0].numpy(), k, ef, n_threads=cores)
    done, t0 = 0, time.perf_counter()
    i = 0
    while True:
        os_.batch_search(Qh[i % nb].numpy(), k, ef, n_threads=cores)
        done += B
        i += 1
        el = time.perf_counter() - t0
        if el >= args.cpu_seconds or i >= 2000:
            break
    t1 = time.perf_counter()
    os_.batch_search(Qh[0].numpy()[: max(B // 10, 1)], k, ef, n_threads=1)
    single = max(B // 10, 1) / (time.perf_counter() - t1)
    out["cpu_baseline"] = {"value": done / el, "unit": "queries/s", "cores": cores, "kind": "port",
                           "sample": f"{done} queries ({i} batches of {B}) in {el:.1f} s, one query per thread, AVX2 "
                                     f"distance, same graph read back from the device",
                           "single_thread_value": single}
    return out


def exact_measurement(args, X, Qh, Qd, torch, dev, stream, with_cpu):
    """hybrid.ExactIndex.Search (exact.go:61-111) over the same rows and query batches: K3, the tensor path
    (exact_gemm.cu: fp16 tcgen05 scores + provably safe filter + canonical fp32 re-rank).  A side measurement carried
    in the same JSON line; the headline metric above is untouched."""
    import hnsw_b200
    n, dim, B, k = args.n, args.dim, args.batch, args.k
    nb = Qd.shape[0]
    x = hnsw_b200.ExactIndex(hnsw_b200.EuclideanDistance, device=dev.index, capacity=n)
    t0 = time.perf_counter()
    x.add_arrays(range(n), X)
    add_s = time.perf_counter() - t0
    rows = torch.empty((B, k), dtype=torch.int64, device=dev)
    dists = torch.empty((B, k), dtype=torch.float32, device=dev)
    counts = torch.empty((B,), dtype=torch.int32, device=dev)
    for i in range(3):
        x.search_device(Qd[i % nb].data_ptr(), B, k, rows.data_ptr(), dists.data_ptr(), counts.data_ptr(),
                        stream.cuda_stream)
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(args.exact_steps):
        x.search_device(Qd[(3 + i) % nb].data_ptr(), B, k, rows.data_ptr(), dists.data_ptr(), counts.data_ptr(),
                        stream.cuda_stream)
    e1.record(stream)
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / args.exact_steps
    last = (3 + args.exact_steps - 1) % nb
    # end to end through b200exact_search_batch (host buffers)
    x.search_rows(Qh[last].numpy(), k)  # untimed: first host-pointer call (staging buffers are allocated)
    e2e_t = []
    for _ in range(5):
        t0 = time.perf_counter()
        h_rows, h_d, h_c = x.search_rows(Qh[last].numpy(), k)
        e2e_t.append(time.perf_counter() - t0)
    e2e_s = float(np.median(e2e_t))
    st = x.last_stats
    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    peak = float(peaks.get("bf16_tflops_sustained", 1414.2))
    flops = 2.0 * n * B * dim
    out = {"metric": "ExactIndex.Search QPS (k=10)", "value": B / (ms / 1e3), "unit": "queries/s", "ms_per_step": ms,
           "steps": args.exact_steps, "dtype": "fp16 operands / fp32 accumulate scores, f32 canonical re-rank",
           "e2e": {"value": B / e2e_s, "unit": "queries/s", "h2d_bytes_per_step": B * dim * 4,
                   "d2h_bytes_per_step": B * k * 12 + B * 4},
           "roofline": {"bound": "tensor", "achieved": flops / (ms / 1e3) / 1e12, "peak": peak, "unit": "TFLOP/s",
                        "frac": flops / (ms / 1e3) / 1e12 / peak, "traffic": None,
                        "kernel": "b200::exact_gemm_kernel (K3 tensor path) + re-rank, whole call",
                        "algorithmic_flops_per_launch": flops,
                        "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained" if peaks else "fallback"},
           "path": {1: "scan", 2: "tensor"}.get(int(st["launches"]), str(st["launches"])),
           "candidates_per_query": st["dist_evals"] / B, "fallback_queries": int(st["expansions"]),
           "add_vec_per_s": n / add_s,
           "device_and_host_call_agree": bool(np.array_equal(h_rows, rows.cpu().numpy()))}
    if with_cpu:
        from oracle import binding as oracle
        cores = os.cpu_count() or 1
        nq = 64
        o_rows, o_d = oracle.exact_batch_search(X, Qh[last].numpy()[:nq], k, oracle.EUCLIDEAN, oracle.CANONICAL,
                                                n_threads=cores)[:2]
        out["parity"] = {"checked_queries": nq, "rows_equal": bool(np.array_equal(o_rows, h_rows[:nq])),
                         "dists_bit_equal": bool(np.array_equal(np.asarray(o_d).view(np.uint32),
                                                                h_d[:nq].view(np.uint32)))}
        nq = min(B, 16 * cores)
        t0 = time.perf_counter()
        oracle.exact_batch_search(X, Qh[last].numpy()[:nq], k, oracle.EUCLIDEAN, oracle.SIMD, n_threads=cores)
        el = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": nq / el, "unit": "queries/s", "cores": cores, "kind": "port",
                               "sample": f"{nq} queries, one per thread, AVX2 distance, in {el:.1f} s"}
    x.Close()
    return out


_REAL_STDOUT = None


def emit(line: dict):
    """The one JSON line goes to the real stdout; everything else (NCCL banners, warnings) was sent to stderr."""
    out = _REAL_STDOUT if _REAL_STDOUT is not None else sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    global _REAL_STDOUT
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)  # libraries that print to stdout (e.g. "NCCL version ...") must not break the one-line contract
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference(args, rank, world)
    return run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    sys.exit(main())
