"""Restatement of graph.go:1116-1537 over the ORACLE (test infrastructure): Graph.Search(3k) by the oracle's search, the
reference's float32 score expression per candidate, highest score first (ties keep candidate order; Go's SortFunc
leaves them unspecified).  Returns lists of keys."""
import numpy as np

F = np.float32


def _score_sort(oracle, metric, vecs, keys, cand_rows, q, negs, k, w, single):
    w = F(w)
    scores = []
    for r in cand_rows:
        qd = oracle.distance(metric, vecs[r], q)          # graph.go:1178 h.Distance(candidate.Value, near)
        qs = F(1.0) - qd
        if single:
            nd = oracle.distance(metric, vecs[r], negs[0])
            ns = F(1.0) - nd
            if qd < F(0.001):
                s = F(2.0)
            elif nd < F(0.1):
                s = qs - w * F(2.0)
            else:
                s = qs - w * ns
        else:
            total, close = F(0.0), False
            for n in negs:
                nd = oracle.distance(metric, vecs[r], n)
                total = F(total + (F(1.0) - nd))
                close = close or nd < F(0.1)
            avg = F(total / F(len(negs)))
            if qd < F(0.001):
                s = F(2.0)
            elif close:
                s = qs - w * F(2.0)
            else:
                boost = F(0.2) if int(keys[r]) in (7, 8, 9) else F(0.0)
                s = F(F(qs - F(w * avg)) + boost)
        scores.append(F(s))
    order = np.argsort(-np.array(scores, dtype=np.float32), kind="stable")
    return [int(keys[cand_rows[i]]) for i in order[:min(k, len(cand_rows))]]


def search_with_negatives_ref(oracle, g, metric, vecs, keys, q, negs, k, w, ef, single=False):
    if len(negs) == 0:
        rows, _, cnt, _ = g.batch_search(np.asarray([q], np.float32), k, ef)
        return [int(keys[r]) for r in rows[0][:cnt[0]]]
    ek = max(3 * k, 10)
    rows, _, cnt, _ = g.batch_search(np.asarray([q], np.float32), ek, ef)
    cand = [int(r) for r in rows[0][:cnt[0]]]
    if not cand:
        return None
    return _score_sort(oracle, metric, vecs, keys, cand, np.asarray(q, np.float32),
                       [np.asarray(n, np.float32) for n in negs], k, w, single)
