"""BASELINE configs[1] at FULL size: the wave-batched construction of 1M x 128 on the device against the oracle's
restatement of the same schedule (ora_graph_add_waves, all host threads): every neighbour list of every layer, the
topography, the entries and the counters must be identical.  (tests/test_wave_add_gpu.py does this at sizes the oracle
finishes in seconds; this is the same check at 1M rows, ~20 s of CPU.)  One JSON line on stdout.

  python scripts/full_size_build_parity.py [N] [dim] [M] [ef] [wave_max]
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, '.')
import hnsw_b200  # noqa: E402
from bench import draw_levels, synth_rows  # noqa: E402
from oracle import binding as oracle  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 128
M = int(sys.argv[3]) if len(sys.argv) > 3 else 16
ef = int(sys.argv[4]) if len(sys.argv) > 4 else 64
WM = int(sys.argv[5]) if len(sys.argv) > 5 else 4096
oracle.build()
X = synth_rows(N, d, 42)
lv = draw_levels(N, 0.25, 44)
keys = np.arange(N, dtype=np.int64)
dg = hnsw_b200.Graph(M=M, Ml=0.25, EfSearch=ef, Distance=hnsw_b200.EuclideanDistance, capacity=N)
t0 = time.perf_counter()
dg.add_arrays(keys, X, lv, wave_max=WM)
dev_s = time.perf_counter() - t0
dst = dict(dg.last_stats)
cores = os.cpu_count() or 1
og = oracle.Graph(d, M, oracle.EUCLIDEAN)
t0 = time.perf_counter()
og.add_waves(keys, X, lv, ef, WM, n_threads=cores)
cpu_s = time.perf_counter() - t0
same_topo = og.topography() == dg.Topography()
same_entries = list(og.entries()) == list(dg.entries())
lists = mism = 0
for l in range(og.n_layers):
    rows, ptr, nbr = og.dump_layer(l)
    dn, member = dg.neighbors(l)
    if not np.array_equal(np.nonzero(member)[0], rows):
        mism += 1
        continue
    deg = np.diff(ptr)
    want = np.full((len(rows), dn.shape[1]), -1, dtype=np.int64)
    col = np.arange(len(nbr)) - np.repeat(ptr[:-1], deg)
    want[np.repeat(np.arange(len(rows)), deg), col] = nbr
    got = dn[rows].astype(np.int64)
    bad = (want != got).any(axis=1)
    mism += int(bad.sum())
    lists += len(rows)
c = og.counters()
ctr_equal = all(int(dst[k]) == int(c[k]) for k in ("dist_evals", "expansions", "nbr_reads", "evict_evals",
                                                   "replenish_evals", "evictions"))
print(json.dumps({"workload": f"{N}x{d} L2 M={M} ef={ef}, waves <= {WM}", "lists_compared": lists,
                  "lists_different": mism, "topography_equal": same_topo, "entries_equal": same_entries,
                  "counters_equal": ctr_equal, "topography": dg.Topography(),
                  "device_wall_s": dev_s, "device_kernel_s": dst["kernel_ms"] / 1e3,
                  "oracle_s": cpu_s, "oracle_threads": cores,
                  "evals_per_insert": (dst["dist_evals"] + dst["evict_evals"] + dst["replenish_evals"]) / N}))
