"""Full-size flavour divergence report (oracle/flavours.py) for a BASELINE config; CPU only, test infrastructure.
usage: flavour_report.py c2|c3 [N] [build_rows]  -> one JSON line.  The searched graph is built by the wave schedule
(what bench.py builds); the per-flavour builds use the same schedule on the first build_rows rows."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import binding as ob, flavours
from benchlib.data import WORKLOADS, gen, draw_levels
ob.build()
wl = dict(WORKLOADS[sys.argv[1]])
N = int(sys.argv[2]) if len(sys.argv) > 2 else wl["n"]
nb = int(sys.argv[3]) if len(sys.argv) > 3 else min(N, 200_000)
cores = os.cpu_count() or 1
met = ob.EUCLIDEAN if wl["metric"] == "l2" else ob.COSINE
X = gen(wl["dist"], N, wl["dim"], 42); Q = gen(wl["dist"], wl["batch"], wl["dim"], 43)
lv = draw_levels(N, 0.25, 44); keys = np.arange(N, dtype=np.int64)
t0 = time.time()
g = ob.Graph(wl["dim"], wl["m"], met, ob.CANONICAL)
g.add_waves(keys, X, lv, wl["ef"], 16384, n_threads=cores)
flat = (X, keys, lv, [g.dump_layer(l) for l in range(g.n_layers)], g.entries())
del g
out = {"workload": f"{sys.argv[1]}: {N} x {wl['dim']} {wl['dist']} {wl['metric']} M={wl['m']} ef={wl['ef']}, {wl['batch']} queries, k={wl['k']}",
       "graph": "wave schedule (waves <= 16384) under the canonical order", "build_s": time.time() - t0}
t0 = time.time()
out["search"] = flavours.search_report(flat, Q, wl["k"], wl["ef"], met, wl["m"], wl["dim"], n_threads=cores)
out["search_s"] = time.time() - t0
t0 = time.time()
out["build"] = flavours.build_report(X[:nb], keys[:nb], lv[:nb], wl["m"], wl["ef"], met, wave_max=16384, n_threads=cores)
out["build_rows"] = nb
out["build_report_s"] = time.time() - t0
print(json.dumps(out))
