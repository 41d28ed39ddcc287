"""BASELINE.json's configs as bench workloads, and the synthetic data SURVEY.md 8(d) specifies for each."""
from __future__ import annotations

import numpy as np

# n is per GPU unless total=True (the configs that shard a fixed dataset over the GPUs of the box)
WORKLOADS = {
    "c2": dict(n=1_000_000, dim=128, metric="l2", m=16, ef=64, k=10, batch=10_000, dist="uniform", total=False,
               baseline="BASELINE.json configs[1]: 1M x 128 synthetic L2, M=16 EfSearch=64, 10k-query batch, k=10"),
    "c2c": dict(n=1_000_000, dim=128, metric="l2", m=16, ef=64, k=10, batch=10_000, dist="clustered", total=False,
                baseline="configs[1] shape on SURVEY 8(d)'s clustered variant: 100 Gaussian centroids, sigma 0.1 "
                         "(hnsw-extensions/hybrid/benchmark_test.go:64-87)"),
    "c3": dict(n=1_000_000, dim=768, metric="cosine", m=32, ef=64, k=10, batch=10_000, dist="normal_unit", total=False,
               baseline="BASELINE.json configs[2]: 1M x 768 synthetic cosine embeddings, M=32, batched Add + Search"),
    "c4": dict(n=10_000_000, dim=96, metric="l2", m=16, ef=64, k=10, batch=10_000, dist="uniform", total=True,
               baseline="BASELINE.json configs[3]: 10M x 96 L2 sharded sub-graphs across the GPUs, top-k merge over "
                        "NVLink"),
    "c5": dict(n=1_000_000, dim=1024, metric="cosine", m=0, ef=0, k=10, batch=10_000, dist="normal_unit", total=True,
               exact=True,
               baseline="BASELINE.json configs[4]: hybrid exact search 1M x 1024 cosine brute-force top-k as a tcgen05 "
                        "distance GEMM, rows sharded over the GPUs"),
}

_CENTROID_SEED = 4242


def synth_rows(n, dim, seed):
    # graph_benchmark_test.go:12-18: rand.Float32()*2-1, i.e. uniform [-1, 1)
    return np.random.default_rng(seed).uniform(-1.0, 1.0, (n, dim)).astype(np.float32)


def gen(dist, n, dim, seed):
    """n x dim float32 rows (or queries) of the named distribution; rows and queries of a workload share the
    distribution and differ in the seed."""
    rng = np.random.default_rng(seed)
    if dist == "uniform":
        return rng.uniform(-1.0, 1.0, (n, dim)).astype(np.float32)
    if dist == "clustered":  # hybrid/benchmark_test.go:64-87: Gaussian blobs around 100 uniform centroids
        cent = np.random.default_rng(_CENTROID_SEED).uniform(-1.0, 1.0, (100, dim)).astype(np.float32)
        out = rng.standard_normal((n, dim), dtype=np.float32)
        out *= np.float32(0.1)
        out += cent[rng.integers(0, 100, n)]
        return out
    if dist == "normal_unit":  # embeddings: N(0, 1) then L2-normalised
        out = rng.standard_normal((n, dim), dtype=np.float32)
        out /= np.linalg.norm(out, axis=1, keepdims=True).astype(np.float32)
        return out
    raise ValueError(dist)


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


def algorithmic_bytes(evals, nbr_reads, queries, dim, k):
    """E rows of 4*dim bytes + the neighbour ids read + the query + the (id, dist) results."""
    return evals * 4 * dim + nbr_reads * 4 + queries * (4 * dim + 8 * k + 4)


def merge_spec(all_d, all_i, all_c, k):
    """The K4 merge as a specification (numpy): per query the candidates of all shards ordered by (distance, shard,
    position in the shard's list), NaN last, slots beyond a shard's count ignored.  all_d/all_i [G, B, kk], all_c [G, B]."""
    G, B, kk = all_d.shape
    pos = np.arange(kk)[None, None, :]
    valid = pos < np.minimum(all_c, kk)[:, :, None]
    d = np.where(valid, all_d, np.inf).astype(np.float64)
    nan = np.isnan(d)
    d = np.where(nan, np.inf, d)
    # lexicographic key: (invalid, nan, distance, shard, position); stable argsort over the flattened [G*kk] axis
    flat_d = d.transpose(1, 0, 2).reshape(B, G * kk)
    flat_bad = (~valid | nan).transpose(1, 0, 2).reshape(B, G * kk)
    flat_inv = (~valid).transpose(1, 0, 2).reshape(B, G * kk)
    order = np.lexsort((np.broadcast_to(np.arange(G * kk), (B, G * kk)), flat_d, flat_bad, flat_inv), axis=1)[:, :k]
    fi = all_i.transpose(1, 0, 2).reshape(B, G * kk)
    fd = all_d.transpose(1, 0, 2).reshape(B, G * kk)
    out_c = np.minimum(k, np.minimum(all_c, kk).sum(axis=0)).astype(np.int32)
    keep = np.arange(k)[None, :] < out_c[:, None]
    out_i = np.where(keep, np.take_along_axis(fi, order, 1), -1).astype(np.int64)
    out_d = np.where(keep, np.take_along_axis(fd, order, 1), np.nan).astype(np.float32)
    return out_d, out_i, out_c
