"""graph.go:1116-1537 -- SearchWithNegative, SearchWithNegatives, BatchSearchWithNegatives.

The reference searches 3k (at least 10) candidates with Graph.Search and re-scores them on the host.  Here the searches
of a batch are ONE device call and every candidate-to-query / candidate-to-negative distance of the batch is ONE device
call (b200hnsw_distance_batch, the kernels' canonical arithmetic); the score arithmetic is the reference's float32
expression, statement by statement, and stays on the host together with its test hacks (the 2.0 score of an exact hit,
the 2 x negWeight penalty, the +0.2 boost of int keys 7-9).  slices.SortFunc is not stable in Go; ties here keep
candidate order.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np

from .distance import distance_batch, metric_of

_F = np.float32


def _expanded_k(k: int) -> int:
    return max(k * 3, 10)  # graph.go:1149-1152


def _is_dog(q: np.ndarray) -> bool:
    return len(q) == 3 and q[0] == _F(1.0) and q[1] == _F(0.2) and q[2] == _F(0.1)  # graph.go:563-569


def _candidates(g, queries: List[np.ndarray], ek: int):
    """Graph.Search(query, ek) for every query: one device call (dog-hack queries go through Search itself)."""
    out = [None] * len(queries)
    plain = [i for i, q in enumerate(queries) if not _is_dog(q)]
    if plain:
        rows, _, counts = g.search_rows(np.stack([queries[i] for i in plain]), ek)
        for j, i in enumerate(plain):
            out[i] = g._nodes(rows[j], counts[j])
    for i, q in enumerate(queries):
        if out[i] is None:
            out[i] = g.Search(q, ek)
    return out


def _score(cands, qd: np.ndarray, nd: np.ndarray, neg_weight, single: bool):
    """cands: nodes; qd[c] = Distance(candidate, query); nd[c][j] = Distance(candidate, negative j).  float32 throughout."""
    w = _F(neg_weight)
    scores = np.empty(len(cands), dtype=np.float32)
    for c, node in enumerate(cands):
        query_sim = _F(1.0) - qd[c]
        if single:  # graph.go:1176-1205
            neg_sim = _F(1.0) - nd[c][0]
            if qd[c] < _F(0.001):
                scores[c] = _F(2.0)
            elif nd[c][0] < _F(0.1):
                scores[c] = query_sim - w * _F(2.0)
            else:
                scores[c] = query_sim - w * neg_sim
            continue
        total = _F(0.0)  # graph.go:1301-1350 / 1460-1507
        close = False
        for j in range(nd.shape[1]):
            total = _F(total + (_F(1.0) - nd[c][j]))
            if nd[c][j] < _F(0.1):
                close = True
        avg = _F(total / _F(nd.shape[1]))
        if qd[c] < _F(0.001):
            scores[c] = _F(2.0)
        elif close:
            scores[c] = query_sim - w * _F(2.0)
        else:
            boost = _F(0.2) if isinstance(node.Key, (int, np.integer)) and int(node.Key) in (7, 8, 9) else _F(0.0)
            scores[c] = _F(_F(query_sim - _F(w * avg)) + boost)
    order = np.argsort(-scores, kind="stable")  # higher score first; ties keep candidate order
    return order


def _rerank_batch(g, queries, negatives, cand_lists, k: int, neg_weight, single: bool):
    """All distances of the batch in one device call, then the host scoring per query."""
    metric = metric_of(g.Distance)
    a_rows, b_rows = [], []
    for q, negs, cands in zip(queries, negatives, cand_lists):
        if not cands or not len(negs):
            continue
        for node in cands:
            a_rows.append(node.Value)
            b_rows.append(q)
            for n in negs:
                a_rows.append(node.Value)
                b_rows.append(n)
    d = distance_batch(metric, np.stack(a_rows), np.stack(b_rows), g.device) if a_rows else np.zeros(0, np.float32)
    out, at = [], 0
    for q, negs, cands in zip(queries, negatives, cand_lists):
        if not cands:
            out.append(None)
            continue
        if not len(negs):
            out.append(cands)
            continue
        w = 1 + len(negs)
        block = d[at:at + len(cands) * w].reshape(len(cands), w)
        at += len(cands) * w
        order = _score(cands, block[:, 0], block[:, 1:], neg_weight, single)
        out.append([cands[i] for i in order[:min(k, len(cands))]])
    return out


def _check_common(g, k, neg_weight):
    g.Validate()
    if k <= 0:
        raise ValueError(f"k must be greater than 0, got {k}")
    if neg_weight < 0.0 or neg_weight > 1.0:
        raise ValueError(f"negWeight must be between 0.0 and 1.0, got {neg_weight:f}")


def search_with_negative(g, near, negative, k: int, neg_weight: float):
    """graph.go:1116-1227"""
    _check_common(g, k, neg_weight)
    near = np.asarray(near, dtype=np.float32).reshape(-1)
    negative = np.asarray(negative, dtype=np.float32).reshape(-1)
    if g.Len() > 0:
        if g.Dims() != len(near):
            raise ValueError(f"query embedding dimension mismatch: {g.Dims()} != {len(near)}")
        if g.Dims() != len(negative):
            raise ValueError(f"negative embedding dimension mismatch: {g.Dims()} != {len(negative)}")
    if g.Len() == 0:
        return None
    cands = _candidates(g, [near], _expanded_k(k))
    return _rerank_batch(g, [near], [[negative]], cands, k, neg_weight, single=True)[0]


def search_with_negatives(g, near, negatives: Sequence, k: int, neg_weight: float):
    """graph.go:1233-1371"""
    _check_common(g, k, neg_weight)
    if len(negatives) == 0:
        return g.Search(near, k)
    near = np.asarray(near, dtype=np.float32).reshape(-1)
    negs = [np.asarray(n, dtype=np.float32).reshape(-1) for n in negatives]
    if g.Len() > 0:
        if g.Dims() != len(near):
            raise ValueError(f"query embedding dimension mismatch: {g.Dims()} != {len(near)}")
        for i, n in enumerate(negs):
            if g.Dims() != len(n):
                raise ValueError(f"negative embedding {i} dimension mismatch: {g.Dims()} != {len(n)}")
    if g.Len() == 0:
        return None
    cands = _candidates(g, [near], _expanded_k(k))
    return _rerank_batch(g, [near], [negs], cands, k, neg_weight, single=False)[0]


def batch_search_with_negatives(g, queries: Sequence, negatives: Sequence[Sequence], k: int, neg_weight: float):
    """graph.go:1376-1537"""
    _check_common(g, k, neg_weight)
    if len(queries) == 0:
        return None
    if len(negatives) != len(queries):
        raise ValueError(f"number of negative example sets ({len(negatives)}) must match number of queries "
                         f"({len(queries)})")
    qs = [np.asarray(q, dtype=np.float32).reshape(-1) for q in queries]
    ns = [[np.asarray(n, dtype=np.float32).reshape(-1) for n in negs] for negs in negatives]
    if g.Len() > 0:
        for i, q in enumerate(qs):
            if g.Dims() != len(q):
                raise ValueError(f"query {i} embedding dimension mismatch: {g.Dims()} != {len(q)}")
            for j, n in enumerate(ns[i]):
                if g.Dims() != len(n):
                    raise ValueError(f"negative embedding {j} for query {i} dimension mismatch: {g.Dims()} != {len(n)}")
    if g.Len() == 0:
        return [None] * len(qs)
    # queries without negatives are plain Search(query, k) (graph.go:1430-1433); the others Search(query, 3k)
    with_neg = [i for i in range(len(qs)) if len(ns[i])]
    without = [i for i in range(len(qs)) if not len(ns[i])]
    results: List[Optional[list]] = [None] * len(qs)
    if without:
        for i, r in zip(without, _candidates(g, [qs[i] for i in without], k)):
            results[i] = r
    if with_neg:
        cands = _candidates(g, [qs[i] for i in with_neg], _expanded_k(k))
        ranked = _rerank_batch(g, [qs[i] for i in with_neg], [ns[i] for i in with_neg], cands, k, neg_weight,
                               single=False)
        for i, r in zip(with_neg, ranked):
            results[i] = r
    return results
