"""The legs of bench.py that may execute oracle/ (the CPU restatement of the reference): the checker of the GPU's last
batch, the reported cpu_baseline, the flavour divergence sample and the `--impl reference` arm.  Never the thing
measured as `value`, never imported by hnsw_b200/."""
from __future__ import annotations

import os
import time

import numpy as np

from .data import draw_levels, gen


def _oracle():
    from oracle import binding as oracle
    oracle.build()
    return oracle


def same_bits(a, b):
    """float32 bit equality with every NaN counted as one value (sign / payload of a computed NaN is the platform's)."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    ua = np.where(np.isnan(a), np.uint32(0x7FC00000), a.view(np.uint32))
    ub = np.where(np.isnan(b), np.uint32(0x7FC00000), b.view(np.uint32))
    return bool(np.array_equal(ua, ub))


def metric_id(oracle, metric):
    return oracle.EUCLIDEAN if metric == "l2" else oracle.COSINE


class OracleView:
    """ONE oracle copy of a device graph (read back with dump_flat), searched under the canonical order for parity and
    under the AVX2-shaped order for the timed CPU baseline."""

    def __init__(self, flat, wl):
        self.oracle = _oracle()
        self.cores = os.cpu_count() or 1
        self.g = self.oracle.Graph(wl["dim"], wl["m"], metric_id(self.oracle, wl["metric"]), self.oracle.CANONICAL)
        self.g.load_flat(*flat)

    def parity(self, Q, k, ef, mode, rows, dists, counts):
        """Canonical-order oracle on the SAME graph vs the device's results for batch Q (bit-exact bar)."""
        self.g.set_flavor(self.oracle.CANONICAL)
        o_rows, o_d, o_cnt, o_ctr = self.g.batch_search(Q, k, ef, mode=mode, n_threads=self.cores)
        return {"checked_queries": int(len(Q)), "rows_equal": bool(np.array_equal(o_rows, rows)),
                "dists_bit_equal": same_bits(o_d, dists), "counts_equal": bool(np.array_equal(o_cnt, counts)),
                "against": "oracle (CPU restatement of graph.go:94-170 / 534-625; extended mode: its own statement in "
                           "oracle/ora_graph.c), canonical summation order, same graph read back from the device"}, o_ctr

    def baseline(self, Qpool, k, ef, mode, seconds):
        """The CPU port in the SAME mode on the same graph: AVX2-shaped distance, one query per thread over all cores
        (the generous analogue of N goroutines calling Search under the read lock), bounded sample; plus one thread,
        which is what the reference's BatchSearch really is (a serial loop, graph.go:1075)."""
        cores = self.cores
        self.g.set_flavor(self.oracle.SIMD)
        nb, B = Qpool.shape[0], Qpool.shape[1]
        self.g.batch_search(Qpool[0][: max(B // 4, 1)], k, ef, mode=mode, n_threads=cores)
        done, i, t0 = 0, 0, time.perf_counter()
        while True:
            self.g.batch_search(Qpool[i % nb], k, ef, mode=mode, n_threads=cores)
            done += B
            i += 1
            el = time.perf_counter() - t0
            if el >= seconds or i >= 2000:
                break
        ns = max(min(B // 10, int(2000 if mode == 0 else 300)), 1)
        t1 = time.perf_counter()
        self.g.batch_search(Qpool[0][:ns], k, ef, mode=mode, n_threads=1)
        single = ns / (time.perf_counter() - t1)
        self.g.set_flavor(self.oracle.CANONICAL)
        return {"value": done / el, "unit": "queries/s", "cores": cores, "kind": "port",
                "sample": f"{done} queries ({i} batches of {B}) in {el:.1f} s, one query per thread, AVX2 distance, same "
                          f"graph read back from the device, same search mode",
                "single_thread_value": single,
                "single_thread_note": "the reference's BatchSearch is a serial loop (graph.go:1075): this is its shape"}


def flavours_sample(wl, n_rows, n_queries, k, ef, ml, cores):
    """north_star: ID divergence only at float ties, and reported.  On the first n_rows rows of the workload: the graph
    built under the canonical order searched under vek's two orders (scalar Go loop, AVX2 shape), and the graph BUILT
    under each order (reference's sequential Graph.Add), compared list by list.  oracle/flavours.py."""
    oracle = _oracle()
    from oracle import flavours
    X = gen(wl["dist"], n_rows, wl["dim"], 42)
    Q = gen(wl["dist"], n_queries, wl["dim"], 43)
    levels = draw_levels(n_rows, ml, 44)
    keys = np.arange(n_rows, dtype=np.int64)
    met = metric_id(oracle, wl["metric"])
    t0 = time.perf_counter()
    g = oracle.Graph(wl["dim"], wl["m"], met, oracle.CANONICAL)
    g.add_many(keys, X, levels, ef)
    flat = (g.dump()[0], keys, levels, [g.dump_layer(l) for l in range(g.n_layers)], g.entries())
    del g
    srep = flavours.search_report(flat, Q, k, ef, met, wl["m"], wl["dim"], n_threads=cores)
    nb = min(n_rows, 10000)
    brep = flavours.build_report(X[:nb], keys[:nb], levels[:nb], wl["m"], ef, met)
    return {"sample": f"first {n_rows} rows of the workload (sequential Graph.Add order), {n_queries} queries; graphs "
                      f"built under each order on the first {nb} rows", "search": srep, "build": brep,
            "seconds": time.perf_counter() - t0,
            "reading": "tie_class = the two orders rank a pair of evaluated rows differently within 1e-5 relative; "
                       "non_tie must be 0"}


def recall_at_k(torch, dev, X, Q, rows, k):
    """recall@k of result rows against a float32 brute force on the GPU (measurement only; L2 or cosine by X's norm)."""
    xd = torch.from_numpy(X).to(dev)
    q = torch.from_numpy(np.ascontiguousarray(Q)).to(dev)
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    d2 = (xd * xd).sum(1)[None, :] - 2.0 * (q @ xd.T)
    torch.backends.cuda.matmul.allow_tf32 = prev
    truth = d2.topk(k, dim=1, largest=False).indices.cpu().numpy()
    del xd, d2
    hits = sum(len(set(int(v) for v in rows[i] if v >= 0) & set(int(v) for v in truth[i])) for i in range(len(Q)))
    return hits / (len(Q) * k)


def exact_cpu(X, Q, k, metric, cores, device_rows, device_dists):
    """ExactIndex.Search on the CPU port: parity on a sample, the fair (stable sort) variant timed on all cores and the
    reference's literal O(N^2) insertion sort (exact.go:88-94) timed on a few queries."""
    oracle = _oracle()
    met = metric_id(oracle, metric)
    nq = min(64, len(Q))
    o_rows, o_d = oracle.exact_batch_search(X, Q[:nq], k, met, oracle.CANONICAL, n_threads=cores)[:2]
    out = {"parity": {"checked_queries": nq, "rows_equal": bool(np.array_equal(o_rows, device_rows[:nq])),
                      "dists_bit_equal": same_bits(o_d, device_dists[:nq])}}
    nt = min(len(Q), 16 * cores)
    t0 = time.perf_counter()
    oracle.exact_batch_search(X, Q[:nt], k, met, oracle.SIMD, n_threads=cores)
    el = time.perf_counter() - t0
    out["cpu_baseline"] = {"value": nt / el, "unit": "queries/s", "cores": cores, "kind": "port",
                           "variant": "fair: all N distances + a stable sort with the same output as exact.go:88-94",
                           "sample": f"{nt} queries, one per thread, AVX2 distance, in {el:.1f} s"}
    nf = min(len(X), 20000)  # the literal insertion sort is O(N^2): timed on a slice of the rows and said so
    t0 = time.perf_counter()
    oracle.exact_batch_search(X[:nf], Q[:cores], k, met, oracle.SIMD, faithful=True, n_threads=cores)
    el = time.perf_counter() - t0
    out["cpu_baseline_faithful"] = {"value": cores / el, "unit": "queries/s", "cores": cores, "kind": "port",
                                    "variant": "faithful: exact.go:88-94's insertion sort of ALL items, O(N^2)",
                                    "sample": f"{cores} queries over the first {nf} rows only, in {el:.1f} s; the cost "
                                              f"grows with N^2, so at N = {len(X)} it is ~{(len(X) / nf) ** 2:.0f}x slower"}
    return out


def run_reference(args, wl, emit):
    """`--impl reference`: the reference's own CPU path for the workload (its C port, all host threads).  No GPU code."""
    oracle = _oracle()
    cores = os.cpu_count() or 1
    n, dim, B, k, ef, m = wl["n"], wl["dim"], wl["batch"], wl["k"], wl["ef"], wl["m"]
    mode = 1 if args.mode == "extended" else 0
    met = metric_id(oracle, wl["metric"])
    X = gen(wl["dist"], n, dim, 42)
    nb = max(1, min(args.warmup + args.steps, args.pool))
    Q = gen(wl["dist"], nb * B, dim, 43).reshape(nb, B, dim)
    if wl.get("exact"):
        # ExactIndex.Search, fair variant, a bounded number of queries per step
        per = max(1, min(B, 4 * cores))
        for i in range(args.warmup):
            oracle.exact_batch_search(X, Q[i % nb][:per], k, met, oracle.SIMD, n_threads=cores)
        t0 = time.perf_counter()
        for i in range(args.steps):
            oracle.exact_batch_search(X, Q[(args.warmup + i) % nb][:per], k, met, oracle.SIMD, n_threads=cores)
        dt = time.perf_counter() - t0
        qps = args.steps * per / dt
        sample = (f"{args.steps} steps of {per} queries (a bounded sample of the {B}-query batch), one query per thread "
                  f"over {cores} threads, AVX2 distance, stable-sort variant of exact.go:88-94")
        serial_rate = None
    else:
        levels = draw_levels(n, args.ml, 44)
        keys = np.arange(n, dtype=np.int64)
        # Graph.Add exactly as the reference does it (serial, graph.go:445) on a bounded sample: the build rate reported
        ns = min(n, 20000)
        gs = oracle.Graph(dim, m, met, oracle.SIMD)
        t0 = time.perf_counter()
        gs.add_many(keys[:ns], X[:ns], levels[:ns], ef)
        serial_rate = ns / (time.perf_counter() - t0)
        del gs
        # the graph that is searched: the same one our arm builds (wave schedule, deterministic), built on all cores
        g = oracle.Graph(dim, m, met, oracle.SIMD)
        t0 = time.perf_counter()
        if args.wave_max > 0:
            g.add_waves(keys, X, levels, ef, args.wave_max, n_threads=cores)
        else:
            g.add_many(keys, X, levels, ef)
        build_s = time.perf_counter() - t0
        for i in range(args.warmup):
            g.batch_search(Q[i % nb], k, ef, mode=mode, n_threads=cores)
        t0 = time.perf_counter()
        for i in range(args.steps):
            g.batch_search(Q[(args.warmup + i) % nb], k, ef, mode=mode, n_threads=cores)
        dt = time.perf_counter() - t0
        qps = args.steps * B / dt
        sample = (f"{args.steps} x {B} queries, one query per thread over {cores} threads, AVX2 distance; graph = the "
                  f"wave-scheduled construction our arm uses, built by the port on {cores} threads in {build_s:.1f} s; "
                  f"the reference's serial Graph.Add runs at {serial_rate:.0f} vec/s here (first {ns} rows)")
    return qps, dt, cores, sample, serial_rate
