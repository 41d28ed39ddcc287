"""Shared helpers for the parity tests: build graphs with the oracle, flatten them for the device."""
import numpy as np


def has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def draw_levels(oracle, n, ml, seed):
    """Level sequence with the reference's rule (graph.go:388-417) from a numpy stream."""
    rng = np.random.default_rng(seed)
    levels = np.empty(n, dtype=np.int32)
    n_layers, size0 = 0, 0
    for i in range(n):
        lvl, _ = oracle.random_level(ml, n_layers, size0, rng.random(64))
        levels[i] = lvl
        n_layers = max(n_layers, lvl + 1)
        size0 += 1
    return levels


def build_oracle_graph(oracle, X, M, ml, ef, metric, seed=44, keys=None, flavor=0):
    n, d = X.shape
    g = oracle.Graph(d, M, metric, flavor)
    levels = draw_levels(oracle, n, ml, seed)
    keys = np.arange(n, dtype=np.int64) if keys is None else np.asarray(keys, dtype=np.int64)
    g.add_many(keys, X, levels, ef)
    return g, keys, levels


def flatten_oracle(g):
    """Oracle graph -> (vectors, keys, levels, layer_lists, entries) as hnsw_b200.Graph.load_flat expects."""
    vec, keys, levels = g.dump()
    layer_lists = [g.dump_layer(l) for l in range(g.n_layers)]
    return vec, keys, levels, layer_lists, g.entries()


def recall_at_k(found_rows, truth_rows):
    k = truth_rows.shape[1]
    hits = 0
    for f, t in zip(found_rows, truth_rows):
        hits += len(set(int(x) for x in f if x >= 0) & set(int(x) for x in t))
    return hits / (k * len(truth_rows))


def same_float_bits(a, b):
    """Bit equality of two float32 arrays with every NaN counted as the same value: the sign and payload of a computed
    NaN are the platform's (x86 0/0 = 0xFFC00000, CUDA 0x7FFFFFFF, arm64 0x7FC00000), not the algorithm's."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    if a.shape != b.shape:
        return False
    ua = np.where(np.isnan(a), np.uint32(0x7FC00000), a.view(np.uint32))
    ub = np.where(np.isnan(b), np.uint32(0x7FC00000), b.view(np.uint32))
    return bool(np.array_equal(ua, ub))
