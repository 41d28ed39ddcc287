"""How far is the wave-batched construction (b200hnsw_add_batch_waves, the throughput path) from the graph the
reference's sequential Graph.Add builds?  CPU only, with the oracle (both schedules are restated there and the device
equals each bit for bit): same rows, same levels; list overlap and search agreement.
usage: wave_vs_reference_order.py N [wave_max] [dist]  -> one JSON line"""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import binding as ob  # test infrastructure: a CPU study, not a product path
from benchlib.data import gen, draw_levels
ob.build()
N = int(sys.argv[1]); wave_max = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
dist = sys.argv[3] if len(sys.argv) > 3 else "uniform"
d, M, ef, k, nq = 128, 16, 64, 10, 10000
cores = os.cpu_count() or 1
X = gen(dist, N, d, 42); Q = gen(dist, nq, d, 43); lv = draw_levels(N, 0.25, 44); keys = np.arange(N, dtype=np.int64)
t = time.time(); gs = ob.Graph(d, M, ob.EUCLIDEAN, ob.SIMD); gs.add_many(keys, X, lv, ef); t_seq = time.time() - t
t = time.time(); gw = ob.Graph(d, M, ob.EUCLIDEAN, ob.SIMD); gw.add_waves(keys, X, lv, ef, wave_max, n_threads=cores); t_wav = time.time() - t
out = {"workload": f"{N} x {d} {dist} L2 M={M} ef={ef}, waves <= {wave_max}", "serial_add_s": t_seq,
       "serial_add_vec_per_s": N / t_seq, "wave_add_s_cpu": t_wav, "cores": cores,
       "topography_equal": gs.topography() == gw.topography(), "layers": {}}
tot = same = links = common = 0
for l in range(gs.n_layers):
    ra, pa, na = gs.dump_layer(l); rb, pb, nb = gw.dump_layer(l)
    assert np.array_equal(ra, rb)
    ls = lc = ll = 0
    for i in range(len(ra)):
        a = na[pa[i]:pa[i + 1]]; b = nb[pb[i]:pb[i + 1]]
        ls += int(len(a) == len(b) and np.array_equal(a, b))
        lc += len(np.intersect1d(a, b, assume_unique=True)); ll += max(len(a), len(b))
    out["layers"][str(l)] = {"lists": int(len(ra)), "identical": ls, "link_overlap": lc / max(ll, 1)}
    tot += len(ra); same += ls; links += ll; common += lc
out["lists"] = tot; out["lists_identical_frac"] = same / tot; out["link_overlap_frac"] = common / links
ra_, da_, ca_, ctr_a = gs.batch_search(Q, k, ef, n_threads=cores)
rb_, db_, cb_, ctr_b = gw.batch_search(Q, k, ef, n_threads=cores)
ident = float(np.mean([np.array_equal(ra_[i], rb_[i]) for i in range(nq)]))
agree = float(np.mean([len(set(ra_[i][:ca_[i]]) & set(rb_[i][:cb_[i]])) / k for i in range(nq)]))
top1 = float(np.mean(ra_[:, 0] == rb_[:, 0]))
out["search_reference_mode"] = {"queries": nq, "identical_result_lists_frac": ident, "top10_agreement": agree,
                                "top1_agreement": top1, "evals_per_query_serial": ctr_a["dist_evals"] / nq,
                                "evals_per_query_wave": ctr_b["dist_evals"] / nq}
re_, _, ce_, _ = gs.batch_search(Q, k, ef, mode=1, n_threads=cores)
rf_, _, cf_, _ = gw.batch_search(Q, k, ef, mode=1, n_threads=cores)
out["search_extended_mode"] = {"top10_agreement": float(np.mean([len(set(re_[i][:ce_[i]]) & set(rf_[i][:cf_[i]])) / k
                                                                 for i in range(nq)]))}
nt = 500
d2 = (X * X).sum(1)[None, :] - 2.0 * (Q[:nt] @ X.T)
truth = np.argpartition(d2, k, axis=1)[:, :k]
rec = lambda R: float(np.mean([len(set(R[i][:k]) & set(truth[i])) / k for i in range(nt)]))
out["recall_at_10"] = {"serial_graph_reference_mode": rec(ra_), "wave_graph_reference_mode": rec(rb_),
                       "serial_graph_extended_mode": rec(re_), "wave_graph_extended_mode": rec(rf_)}
print(json.dumps(out))
