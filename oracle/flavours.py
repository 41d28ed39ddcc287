"""Flavour divergence report (TEST INFRASTRUCTURE, like everything under oracle/).

north_star: "Any ID divergence is allowed only at float ties and must be reported."  The kernels are bit-exact to the
oracle's CANONICAL summation order; the reference's own order lives in github.com/viterin/vek v0.4.2 (distance.go:16,22;
absent from the tree, platform dependent: AVX2/FMA assembly on amd64, a scalar Go loop elsewhere).  The oracle restates
both shapes (SEQUENTIAL = vek's pure-Go loop, SIMD = the shape of its AVX2 path), so the bridge from "equals our oracle"
to "equals the Go reference" is: how often do results differ BETWEEN flavours, and is every difference a float tie?

Search (same graph, same queries, flavour A vs flavour B).  Every decision layerNode.search takes (graph.go:147, 150,
heap sifts) is a `<` between the distances of two rows it has evaluated.  So two flavours walk the same path for as
long as they ORDER the rows evaluated so far the same way; the first step at which their evaluation traces differ has,
among the rows evaluated before it, a pair (x, y) ordered differently by the two flavours -- a float tie at the
flavours' resolution.  Per differing query the report finds that point, collects the inverted pairs and their relative
gap: `tie_class` = every inverted pair is within `tol` (1e-5 relative, north_star's distance tolerance), `non_tie` =
anything else (must be 0: it would mean a divergence that is not explained by summation order).

Build (same rows, same levels, graph built under each flavour): neighbour lists that differ, per layer and in total.
"""
from __future__ import annotations

import numpy as np

from . import binding as ob

NAMES = {ob.CANONICAL: "canonical", ob.SEQUENTIAL: "sequential", ob.SIMD: "simd"}


def _inverted_pairs(da: np.ndarray, db: np.ndarray):
    """Pairs (i, j) of evaluated rows that flavour A and flavour B order differently (strict `<` on one side, not on
    the other).  Returns (count, max relative gap over them) -- gap measured as |d_i - d_j| / max(|d_i|, |d_j|) under
    either flavour, whichever is larger."""
    n = len(da)
    if n < 2:
        return 0, 0.0
    # O(n^2) on a few hundred evaluations per query; exact and simple
    la = da[:, None] < da[None, :]
    lb = db[:, None] < db[None, :]
    inv = la != lb
    # NaN distances compare false on both sides under both flavours unless only one flavour produced the NaN
    cnt = int(np.triu(inv | inv.T, 1).sum())
    if cnt == 0:
        return 0, 0.0
    ii, jj = np.nonzero(np.triu(inv | inv.T, 1))
    with np.errstate(invalid="ignore", divide="ignore"):
        ga = np.abs(da[ii] - da[jj]) / np.maximum(np.maximum(np.abs(da[ii]), np.abs(da[jj])), 1e-30)
        gb = np.abs(db[ii] - db[jj]) / np.maximum(np.maximum(np.abs(db[ii]), np.abs(db[jj])), 1e-30)
    g = np.nanmax(np.maximum(ga, gb)) if len(ii) else 0.0
    return cnt, float(g)


def compare_search(res_a, res_b, tol=1e-5):
    """res_* = Graph.batch_search_trace(...) of two flavours on the same graph and queries."""
    rows_a, d_a, cnt_a, tr_a, td_a, tl_a = res_a
    rows_b, d_b, cnt_b, tr_b, td_b, tl_b = res_b
    nq, cap = tr_a.shape
    differing = tie_class = non_tie = truncated = 0
    set_diff_queries = 0
    max_gap = 0.0
    max_rel_err = 0.0
    for q in range(nq):
        la, lb = int(min(tl_a[q], cap)), int(min(tl_b[q], cap))
        m = min(la, lb)
        same_prefix = tr_a[q, :m] == tr_b[q, :m]
        t = m if same_prefix.all() else int(np.argmin(same_prefix))  # first step at which the walks differ
        if t > 0:
            with np.errstate(invalid="ignore", divide="ignore"):
                rel = np.abs(td_a[q, :t] - td_b[q, :t]) / np.maximum(np.abs(td_a[q, :t]), 1e-30)
            rel = rel[np.isfinite(rel)]
            if len(rel):
                max_rel_err = max(max_rel_err, float(rel.max()))
        ids_equal = cnt_a[q] == cnt_b[q] and np.array_equal(rows_a[q], rows_b[q])
        walks_equal = t == m and tl_a[q] == tl_b[q]
        if ids_equal and walks_equal:
            continue
        if ids_equal:
            continue  # the walks parted and met again: no ID divergence to report
        differing += 1
        if set(rows_a[q][:cnt_a[q]].tolist()) != set(rows_b[q][:cnt_b[q]].tolist()):
            set_diff_queries += 1
        if max(tl_a[q], tl_b[q]) > cap and t >= cap:
            truncated += 1
            continue
        # the divergence shows either in the walk (a different row evaluated at step t) or, with identical walks, in the
        # result queue's layout: both are decided by comparisons among the rows evaluated up to that point
        upto = t if not walks_equal else m
        n_inv, gap = _inverted_pairs(td_a[q, :upto], td_b[q, :upto])
        max_gap = max(max_gap, gap)
        if n_inv > 0 and gap <= tol:
            tie_class += 1
        else:
            non_tie += 1
    return {"queries": int(nq), "queries_with_id_difference": int(differing),
            "of_which_different_id_sets": int(set_diff_queries), "tie_class": int(tie_class),
            "non_tie": int(non_tie), "trace_truncated": int(truncated),
            "max_relative_gap_of_inverted_pairs": max_gap, "max_relative_distance_error": max_rel_err, "tol": tol}


def search_report(flat, queries, k, ef, metric, M, dim, n_threads=1, mode=ob.MODE_REFERENCE, trace_cap=2048, tol=1e-5,
                  flavours=(ob.CANONICAL, ob.SEQUENTIAL, ob.SIMD)):
    """flat = (vectors, keys, levels, layer_lists, entries): ONE graph, searched under every flavour."""
    res = {}
    for f in flavours:
        g = ob.Graph(dim, M, metric, f)
        g.load_flat(*flat)
        res[f] = g.batch_search_trace(queries, k, ef, mode=mode, n_threads=n_threads, trace_cap=trace_cap)
        del g
    out = {}
    fl = list(flavours)
    for i in range(len(fl)):
        for j in range(i + 1, len(fl)):
            out[f"{NAMES[fl[i]]}_vs_{NAMES[fl[j]]}"] = compare_search(res[fl[i]], res[fl[j]], tol)
    return out


def build_report(X, keys, levels, M, ef, metric, wave_max=0, n_threads=1,
                 flavours=(ob.CANONICAL, ob.SEQUENTIAL, ob.SIMD)):
    """The same rows and levels inserted under every flavour (sequential Graph.Add order unless wave_max > 0):
    neighbour lists that differ between the built graphs."""
    graphs = {}
    for f in flavours:
        g = ob.Graph(X.shape[1], M, metric, f)
        if wave_max > 0:
            g.add_waves(keys, X, levels, ef, wave_max, n_threads=n_threads)
        else:
            g.add_many(keys, X, levels, ef)
        graphs[f] = [g.dump_layer(l) for l in range(g.n_layers)], g.topography(), list(g.entries())
        del g
    out = {}
    fl = list(flavours)
    for i in range(len(fl)):
        for j in range(i + 1, len(fl)):
            la, ta, ea = graphs[fl[i]]
            lb, tb, eb = graphs[fl[j]]
            lists = diff = links = link_diff = 0
            for (ra, pa, na), (rb, pb, nb) in zip(la, lb):
                assert np.array_equal(ra, rb)
                lists += len(ra)
                da, db = np.diff(pa), np.diff(pb)
                for r in range(len(ra)):
                    a, b = na[pa[r]:pa[r + 1]], nb[pb[r]:pb[r + 1]]
                    links += len(a)
                    if len(a) != len(b) or not np.array_equal(a, b):
                        diff += 1
                        link_diff += len(set(a.tolist()) ^ set(b.tolist()))
            out[f"{NAMES[fl[i]]}_vs_{NAMES[fl[j]]}"] = {
                "lists": int(lists), "lists_different": int(diff), "links": int(links),
                "links_in_symmetric_difference": int(link_diff), "topography_equal": ta == tb,
                "entries_equal": ea == eb}
    return out
