"""Specification of the K4 merge in numpy (test infrastructure): per query, the G*k candidates are ordered by
(distance, shard, position in the shard's list); NaN distances last; slots beyond a shard's count are ignored."""
import numpy as np


def merge_topk_ref(all_d, all_i, all_c, k):
    G, B, kk = all_d.shape
    out_d = np.full((B, k), np.nan, dtype=np.float32)
    out_i = np.full((B, k), -1, dtype=np.int64)
    out_c = np.zeros(B, dtype=np.int32)
    for q in range(B):
        items = []
        for g in range(G):
            for p in range(min(int(all_c[g, q]), kk)):
                d = all_d[g, q, p]
                items.append((np.isnan(d), d if not np.isnan(d) else 0.0, g, p))
        items.sort()
        n = min(k, len(items))
        for t in range(n):
            _, _, g, p = items[t]
            out_d[q, t] = all_d[g, q, p]
            out_i[q, t] = all_i[g, q, p]
        out_c[q] = min(k, int(all_c[:, q].clip(max=kk).sum()))
    return out_d, out_i, out_c
