"""ctypes binding of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
The product package hnsw_b200 never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB = None

EUCLIDEAN, COSINE = 0, 1
CANONICAL, SEQUENTIAL, SIMD = 0, 1, 2
MODE_REFERENCE, MODE_EXTENDED = 0, 1

f32p = C.POINTER(C.c_float)
i32p = C.POINTER(C.c_int32)
u8p = C.POINTER(C.c_uint8)
i64p = C.POINTER(C.c_int64)
f64p = C.POINTER(C.c_double)


class Counters(C.Structure):
    _fields_ = [(n, C.c_int64) for n in
                ("dist_evals", "expansions", "nbr_reads", "evict_evals", "replenish_evals", "evictions")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


class _Heap(C.Structure):
    _fields_ = [("dist", f32p), ("id", i32p), ("len", C.c_int), ("cap", C.c_int)]


def build(force: bool = False) -> Path:
    so = _HERE / "liboracle.so"
    srcs = [_HERE / n for n in ("hnsw_oracle.c", "ora_graph.c", "ora_encode_exact.c", "hnsw_oracle.h")]
    if force or not so.exists() or any(s.stat().st_mtime > so.stat().st_mtime for s in srcs):
        subprocess.run(["make", "-C", str(_HERE), "-B", "liboracle.so"], check=True, capture_output=True)
    return so


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    so = _HERE / "liboracle.so"
    if not so.exists():
        build()
    L = C.CDLL(str(so))
    L.ora_distance.restype = C.c_float
    L.ora_distance.argtypes = [C.c_int, C.c_int, f32p, f32p, C.c_int]
    L.ora_max_level.restype = C.c_int
    L.ora_max_level.argtypes = [C.c_double, C.c_int64]
    L.ora_random_level.restype = C.c_int
    L.ora_random_level.argtypes = [C.c_double, C.c_int, C.c_int64, f64p, C.POINTER(C.c_int)]
    L.ora_graph_new.restype = C.c_void_p
    L.ora_graph_new.argtypes = [C.c_int] * 4
    L.ora_graph_free.argtypes = [C.c_void_p]
    for name, res in (("ora_graph_len", C.c_int64), ("ora_graph_layers", C.c_int), ("ora_graph_M", C.c_int),
                      ("ora_graph_metric", C.c_int), ("ora_graph_dim", C.c_int)):
        getattr(L, name).restype = res
        getattr(L, name).argtypes = [C.c_void_p]
    L.ora_graph_layer_size.restype = C.c_int64
    L.ora_graph_layer_size.argtypes = [C.c_void_p, C.c_int]
    L.ora_graph_entry.restype = C.c_int32
    L.ora_graph_entry.argtypes = [C.c_void_p, C.c_int]
    L.ora_graph_set_entry.argtypes = [C.c_void_p, C.c_int, C.c_int32]
    L.ora_graph_level.restype = C.c_int
    L.ora_graph_level.argtypes = [C.c_void_p, C.c_int32]
    L.ora_graph_key.restype = C.c_int64
    L.ora_graph_key.argtypes = [C.c_void_p, C.c_int32]
    L.ora_graph_vector.restype = f32p
    L.ora_graph_vector.argtypes = [C.c_void_p, C.c_int32]
    L.ora_graph_neighbors.restype = C.c_int
    L.ora_graph_neighbors.argtypes = [C.c_void_p, C.c_int, C.c_int32, i32p]
    L.ora_graph_counters.argtypes = [C.c_void_p, C.POINTER(Counters)]
    L.ora_graph_reset_counters.argtypes = [C.c_void_p]
    L.ora_graph_layer_rows.restype = C.c_int64
    L.ora_graph_layer_rows.argtypes = [C.c_void_p, C.c_int]
    L.ora_graph_delete.restype = C.c_int
    L.ora_graph_delete.argtypes = [C.c_void_p, C.c_int64]
    L.ora_graph_delete_row.restype = C.c_int
    L.ora_graph_delete_row.argtypes = [C.c_void_p, C.c_int32]
    L.ora_graph_is_dead.restype = C.c_int
    L.ora_graph_is_dead.argtypes = [C.c_void_p, C.c_int32]
    L.ora_graph_dead_count.restype = C.c_int64
    L.ora_graph_dead_count.argtypes = [C.c_void_p]
    L.ora_graph_set_build_mode.restype = None
    L.ora_graph_set_build_mode.argtypes = [C.c_void_p, C.c_int]
    L.ora_graph_set_flavor.restype = None
    L.ora_graph_set_flavor.argtypes = [C.c_void_p, C.c_int]
    L.ora_graph_add.restype = C.c_int32
    L.ora_graph_add.argtypes = [C.c_void_p, C.c_int64, f32p, C.c_int, C.c_int]
    L.ora_graph_put_node.restype = C.c_int32
    L.ora_graph_put_node.argtypes = [C.c_void_p, C.c_int64, f32p, C.c_int]
    L.ora_graph_link.restype = C.c_int
    L.ora_graph_link.argtypes = [C.c_void_p, C.c_int, C.c_int32, C.c_int32]
    L.ora_layer_search.restype = C.c_int
    L.ora_layer_search.argtypes = [C.c_void_p, C.c_int, C.c_int32, C.c_int, C.c_int, f32p, C.c_int, i32p, f32p]
    L.ora_graph_search.restype = C.c_int
    L.ora_graph_search.argtypes = [C.c_void_p, f32p, C.c_int, C.c_int, C.c_int, i32p, i32p, f32p]
    L.ora_graph_batch_search.argtypes = [C.c_void_p, f32p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, i32p, f32p,
                                         i32p, C.POINTER(Counters)]
    L.ora_graph_batch_search_trace.argtypes = [C.c_void_p, f32p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, i32p,
                                               f32p, i32p, C.POINTER(Counters), i32p, f32p, i64p, C.c_int64]
    L.ora_put_varint.restype = C.c_int
    L.ora_put_varint.argtypes = [u8p, C.c_int64]
    L.ora_get_varint.restype = C.c_int
    L.ora_get_varint.argtypes = [u8p, C.c_size_t, C.POINTER(C.c_int64)]
    L.ora_graph_export.restype = C.c_int64
    L.ora_graph_export.argtypes = [C.c_void_p, C.c_double, C.c_int, u8p, C.c_int64]
    L.ora_graph_import.restype = C.c_void_p
    L.ora_graph_import.argtypes = [u8p, C.c_int64, C.c_int, f64p, C.POINTER(C.c_int)]
    L.ora_exact_search.restype = C.c_int
    L.ora_exact_search.argtypes = [f32p, C.c_int64, C.c_int, C.c_int, C.c_int, f32p, C.c_int, C.c_int, i32p, f32p]
    L.ora_exact_batch_search.argtypes = [f32p, C.c_int64, C.c_int, C.c_int, C.c_int, f32p, C.c_int64, C.c_int,
                                         C.c_int, C.c_int, i32p, f32p]
    L.ora_graph_dump_layer.restype = C.c_int64
    L.ora_graph_dump_layer.argtypes = [C.c_void_p, C.c_int, i32p, i64p, i32p]
    L.ora_graph_dump_rows.argtypes = [C.c_void_p, f32p, i64p, i32p]
    L.ora_graph_add_many.restype = C.c_int64
    L.ora_graph_add_many.argtypes = [C.c_void_p, i64p, f32p, i32p, C.c_int64, C.c_int]
    L.ora_graph_load_flat.restype = C.c_int
    L.ora_graph_load_flat.argtypes = [C.c_void_p, C.c_int64, f32p, i64p, i32p, C.c_int, i32p, i64p, i32p, i64p, i32p]
    L.ora_wave_len.restype = C.c_int64
    L.ora_wave_len.argtypes = [C.c_int64, C.c_int, i32p, C.c_int64, C.c_int64]
    L.ora_graph_add_waves.restype = C.c_int64
    L.ora_graph_add_waves.argtypes = [C.c_void_p, i64p, f32p, i32p, C.c_int64, C.c_int, C.c_int64, C.c_int]
    L.ora_heap_init.argtypes = [C.POINTER(_Heap), C.c_int]
    L.ora_heap_free.argtypes = [C.POINTER(_Heap)]
    L.ora_heap_push.argtypes = [C.POINTER(_Heap), C.c_float, C.c_int32]
    L.ora_heap_pop.argtypes = [C.POINTER(_Heap), f32p, i32p]
    L.ora_heap_poplast.argtypes = [C.POINTER(_Heap), f32p, i32p]
    _LIB = L
    return L


def _f32(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(f32p)


def distance(metric, a, b, flavor=CANONICAL) -> np.float32:
    a, pa = _f32(a)
    b, pb = _f32(b)
    assert a.shape == b.shape and a.ndim == 1
    return np.float32(lib().ora_distance(metric, flavor, pa, pb, a.shape[0]))


def max_level(ml: float, n: int) -> int:
    return lib().ora_max_level(ml, n)


def random_level(ml, n_layers, layer0_size, uniforms):
    u = np.ascontiguousarray(uniforms, dtype=np.float64)
    used = C.c_int(0)
    lvl = lib().ora_random_level(ml, n_layers, layer0_size, u.ctypes.data_as(f64p), C.byref(used))
    return lvl, used.value


class Heap:
    """heap/heap.go Heap[searchCandidate] (dist, id)."""

    def __init__(self, cap=8):
        self._h = _Heap()
        lib().ora_heap_init(C.byref(self._h), cap)

    def __del__(self):
        try:
            lib().ora_heap_free(C.byref(self._h))
        except Exception:
            pass

    def __len__(self):
        return self._h.len

    def push(self, d, i):
        lib().ora_heap_push(C.byref(self._h), float(d), int(i))

    def pop(self):
        d, i = C.c_float(), C.c_int32()
        lib().ora_heap_pop(C.byref(self._h), C.byref(d), C.byref(i))
        return np.float32(d.value), i.value

    def pop_last(self):
        d, i = C.c_float(), C.c_int32()
        lib().ora_heap_poplast(C.byref(self._h), C.byref(d), C.byref(i))
        return np.float32(d.value), i.value

    def slice(self):
        n = self._h.len
        return [(np.float32(self._h.dist[t]), self._h.id[t]) for t in range(n)]


class Graph:
    """Row-based restatement of hnsw.Graph[int] (graph.go:305-332)."""

    def __init__(self, dim, M, metric, flavor=CANONICAL, _handle=None):
        self.L = lib()
        self.h = _handle if _handle is not None else self.L.ora_graph_new(dim, M, metric, flavor)
        self.dim, self.M, self.metric, self.flavor = dim, M, metric, flavor

    def __del__(self):
        try:
            self.L.ora_graph_free(self.h)
        except Exception:
            pass

    def __len__(self):
        return int(self.L.ora_graph_len(self.h))

    @property
    def n_layers(self):
        return self.L.ora_graph_layers(self.h)

    def topography(self):
        """analyzer.go Topography: nodes per layer."""
        return [int(self.L.ora_graph_layer_size(self.h, l)) for l in range(self.n_layers)]

    def entries(self):
        return np.array([self.L.ora_graph_entry(self.h, l) for l in range(self.n_layers)], dtype=np.int32)

    def set_entry(self, layer, row):
        self.L.ora_graph_set_entry(self.h, layer, row)

    def add(self, key, vec, level, ef):
        v, pv = _f32(vec)
        assert v.shape == (self.dim,)
        row = self.L.ora_graph_add(self.h, int(key), pv, int(level), int(ef))
        if row < 0:
            raise RuntimeError(f"ora_graph_add failed: {row}")
        return row

    def delete(self, key) -> bool:
        """Graph.Delete (graph.go:843-866)."""
        return bool(self.L.ora_graph_delete(self.h, int(key)))

    def batch_delete(self, keys):
        """Graph.BatchDelete (graph.go:871-895): the same loop under one lock."""
        return [self.delete(k) for k in keys]

    def is_dead(self, row) -> bool:
        return self.L.ora_graph_is_dead(self.h, int(row)) == 1

    def set_build_mode(self, mode):
        self.L.ora_graph_set_build_mode(self.h, int(mode))

    def set_flavor(self, flavor):
        """Summation order used from now on (one loaded graph searched under several orders)."""
        self.L.ora_graph_set_flavor(self.h, int(flavor))
        self.flavor = int(flavor)

    def add_many(self, keys, vecs, levels, ef):
        vecs = np.ascontiguousarray(vecs, dtype=np.float32)
        keys = np.ascontiguousarray(keys, dtype=np.int64)
        levels = np.ascontiguousarray(levels, dtype=np.int32)
        done = self.L.ora_graph_add_many(self.h, keys.ctypes.data_as(i64p), vecs.ctypes.data_as(f32p),
                                         levels.ctypes.data_as(i32p), len(keys), int(ef))
        if done != len(keys):
            raise RuntimeError(f"ora_graph_add failed at node {done}")

    def add_waves(self, keys, vecs, levels, ef, wave_max=4096, n_threads=1):
        """Wave-batched Add (NOT the reference's algorithm; restates the device library's batched construction)."""
        vecs = np.ascontiguousarray(vecs, dtype=np.float32)
        keys = np.ascontiguousarray(keys, dtype=np.int64)
        levels = np.ascontiguousarray(levels, dtype=np.int32)
        done = self.L.ora_graph_add_waves(self.h, keys.ctypes.data_as(i64p), vecs.ctypes.data_as(f32p),
                                          levels.ctypes.data_as(i32p), len(keys), int(ef), int(wave_max),
                                          int(n_threads))
        if done != len(keys):
            raise RuntimeError(f"ora_graph_add_waves failed at node {done}")

    def load_flat(self, vectors, keys, levels, layer_lists, entries=None):
        """Fill an empty graph from (vectors, keys, levels, [(rows, ptr, nbr) per layer], entries)."""
        vectors = np.ascontiguousarray(vectors, dtype=np.float32)
        keys = np.ascontiguousarray(keys, dtype=np.int64)
        levels = np.ascontiguousarray(levels, dtype=np.int32)
        nl = len(layer_lists)
        sizes = np.array([len(r) for r, _, _ in layer_lists], dtype=np.int64)
        rows = np.ascontiguousarray(np.concatenate([np.asarray(r, np.int32) for r, _, _ in layer_lists]))
        ptr = np.ascontiguousarray(np.concatenate([np.asarray(p, np.int64) for _, p, _ in layer_lists]))
        nbr = np.ascontiguousarray(np.concatenate([np.asarray(b, np.int32) for _, _, b in layer_lists] +
                                                  [np.zeros(1, np.int32)]))
        ent = None if entries is None else np.ascontiguousarray(entries, dtype=np.int32)
        rc = self.L.ora_graph_load_flat(self.h, len(keys), vectors.ctypes.data_as(f32p), keys.ctypes.data_as(i64p),
                                        levels.ctypes.data_as(i32p), nl,
                                        ent.ctypes.data_as(i32p) if ent is not None else None,
                                        sizes.ctypes.data_as(i64p), rows.ctypes.data_as(i32p),
                                        ptr.ctypes.data_as(i64p), nbr.ctypes.data_as(i32p))
        if rc:
            raise RuntimeError(f"ora_graph_load_flat failed: {rc}")

    def dump_layer(self, layer):
        """CSR of one layer over its member rows: (rows, ptr, nbr)."""
        n = int(self.L.ora_graph_layer_rows(self.h, layer))  # deleted rows keep their lists
        nn = self.L.ora_graph_dump_layer(self.h, layer, None, None, None)
        rows = np.empty(n, dtype=np.int32)
        ptr = np.empty(n + 1, dtype=np.int64)
        nbr = np.empty(max(nn, 1), dtype=np.int32)
        self.L.ora_graph_dump_layer(self.h, layer, rows.ctypes.data_as(i32p), ptr.ctypes.data_as(i64p),
                                    nbr.ctypes.data_as(i32p))
        return rows, ptr, nbr[:nn]

    def put_node(self, key, vec, level=0):
        v, pv = _f32(vec)
        return self.L.ora_graph_put_node(self.h, int(key), pv, int(level))

    def link(self, layer, a, b):
        assert self.L.ora_graph_link(self.h, layer, a, b) == 0

    def level(self, row):
        return self.L.ora_graph_level(self.h, row)

    def key(self, row):
        return int(self.L.ora_graph_key(self.h, row))

    def vector(self, row):
        p = self.L.ora_graph_vector(self.h, row)
        return np.ctypeslib.as_array(p, shape=(self.dim,)).copy()

    def neighbors(self, layer, row):
        out = np.empty(self.M + 1, dtype=np.int32)
        d = self.L.ora_graph_neighbors(self.h, layer, row, out.ctypes.data_as(i32p))
        return None if d < 0 else out[:d].copy()

    def counters(self):
        c = Counters()
        self.L.ora_graph_counters(self.h, C.byref(c))
        return c.as_dict()

    def reset_counters(self):
        self.L.ora_graph_reset_counters(self.h)

    def layer_search(self, layer, entry, k, ef, target, mode=MODE_REFERENCE):
        t, pt = _f32(target)
        rows = np.empty(k, dtype=np.int32)
        d = np.empty(k, dtype=np.float32)
        n = self.L.ora_layer_search(self.h, layer, entry, k, ef, pt, mode, rows.ctypes.data_as(i32p),
                                    d.ctypes.data_as(f32p))
        return rows[:n].copy(), d[:n].copy()

    def search(self, query, k, ef, mode=MODE_REFERENCE, entries=None):
        q, pq = _f32(query)
        rows = np.empty(k, dtype=np.int32)
        d = np.empty(k, dtype=np.float32)
        pe = None
        if entries is not None:
            e = np.ascontiguousarray(entries, dtype=np.int32)
            pe = e.ctypes.data_as(i32p)
        n = self.L.ora_graph_search(self.h, pq, k, ef, mode, pe, rows.ctypes.data_as(i32p), d.ctypes.data_as(f32p))
        return rows[:n].copy(), d[:n].copy()

    def batch_search(self, queries, k, ef, mode=MODE_REFERENCE, n_threads=1):
        q = np.ascontiguousarray(queries, dtype=np.float32)
        nq = q.shape[0]
        rows = np.full((nq, k), -1, dtype=np.int32)
        d = np.full((nq, k), np.nan, dtype=np.float32)
        cnt = np.zeros(nq, dtype=np.int32)
        c = Counters()
        self.L.ora_graph_batch_search(self.h, q.ctypes.data_as(f32p), nq, k, ef, mode, n_threads,
                                      rows.ctypes.data_as(i32p), d.ctypes.data_as(f32p), cnt.ctypes.data_as(i32p),
                                      C.byref(c))
        return rows, d, cnt, c.as_dict()

    def batch_search_trace(self, queries, k, ef, mode=MODE_REFERENCE, n_threads=1, trace_cap=1024):
        """batch_search plus, per query, the rows it evaluated in evaluation order with this flavour's distances:
        (rows, dists, counts, trace_rows[nq, cap], trace_dists[nq, cap], trace_len[nq])."""
        q = np.ascontiguousarray(queries, dtype=np.float32)
        nq = q.shape[0]
        rows = np.full((nq, k), -1, dtype=np.int32)
        d = np.full((nq, k), np.nan, dtype=np.float32)
        cnt = np.zeros(nq, dtype=np.int32)
        tr = np.full((nq, trace_cap), -1, dtype=np.int32)
        td = np.full((nq, trace_cap), np.nan, dtype=np.float32)
        tl = np.zeros(nq, dtype=np.int64)
        c = Counters()
        self.L.ora_graph_batch_search_trace(self.h, q.ctypes.data_as(f32p), nq, k, ef, mode, n_threads,
                                            rows.ctypes.data_as(i32p), d.ctypes.data_as(f32p),
                                            cnt.ctypes.data_as(i32p), C.byref(c), tr.ctypes.data_as(i32p),
                                            td.ctypes.data_as(f32p), tl.ctypes.data_as(i64p), trace_cap)
        return rows, d, cnt, tr, td, tl

    def export(self, ml, ef) -> bytes:
        n = self.L.ora_graph_export(self.h, ml, ef, None, 0)
        buf = (C.c_uint8 * n)()
        self.L.ora_graph_export(self.h, ml, ef, buf, n)
        return bytes(buf)

    @classmethod
    def import_(cls, data: bytes, flavor=CANONICAL):
        L = lib()
        buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
        ml, ef = C.c_double(), C.c_int()
        h = L.ora_graph_import(buf, len(data), flavor, C.byref(ml), C.byref(ef))
        if not h:
            raise ValueError("import failed")
        g = cls(L.ora_graph_dim(h), L.ora_graph_M(h), L.ora_graph_metric(h), flavor, _handle=h)
        return g, ml.value, ef.value

    def dump(self):
        """(vectors[N,d], keys[N], levels[N], lists per layer as {layer: (rows[N_l], adj[N_l,M], deg[N_l])})."""
        n = int(self.L.ora_graph_len(self.h)) if self.n_layers else 0
        # rows are dense 0..n-1 when nothing was deleted
        vec = np.zeros((n, self.dim), np.float32)
        keys = np.zeros(n, dtype=np.int64)
        levels = np.zeros(n, dtype=np.int32)
        if n:
            self.L.ora_graph_dump_rows(self.h, vec.ctypes.data_as(f32p), keys.ctypes.data_as(i64p),
                                       levels.ctypes.data_as(i32p))
        return vec, keys, levels


def exact_search(vecs, query, k, metric, flavor=CANONICAL, faithful=False):
    v = np.ascontiguousarray(vecs, dtype=np.float32)
    q, pq = _f32(query)
    idx = np.empty(k, dtype=np.int32)
    d = np.empty(k, dtype=np.float32)
    n = lib().ora_exact_search(v.ctypes.data_as(f32p), v.shape[0], v.shape[1] if v.ndim == 2 else 0, metric, flavor, pq,
                               k, int(faithful), idx.ctypes.data_as(i32p), d.ctypes.data_as(f32p))
    return idx[:n].copy(), d[:n].copy()


def exact_batch_search(vecs, queries, k, metric, flavor=CANONICAL, faithful=False, n_threads=1):
    v = np.ascontiguousarray(vecs, dtype=np.float32)
    q = np.ascontiguousarray(queries, dtype=np.float32)
    nq = q.shape[0]
    kk = min(k, v.shape[0])
    idx = np.full((nq, k), -1, dtype=np.int32)
    d = np.full((nq, k), np.nan, dtype=np.float32)
    lib().ora_exact_batch_search(v.ctypes.data_as(f32p), v.shape[0], v.shape[1], metric, flavor,
                                 q.ctypes.data_as(f32p), nq, k, int(faithful), n_threads,
                                 idx.ctypes.data_as(i32p), d.ctypes.data_as(f32p))
    return idx[:, :kk], d[:, :kk]


def put_varint(v: int) -> bytes:
    buf = (C.c_uint8 * 10)()
    n = lib().ora_put_varint(buf, v)
    return bytes(buf[:n])


def get_varint(data: bytes):
    buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
    v = C.c_int64()
    n = lib().ora_get_varint(buf, len(data), C.byref(v))
    return v.value, n
