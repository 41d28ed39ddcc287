"""hnsw-extensions/hybrid/exact.go:13-160 -- ExactIndex[K]: brute-force exact search, on the device.

Host-side mirror over the C ABI (b200exact_* in include/b200hnsw.h).  The host keeps what a cgo host would keep: the
key <-> row mapping and a copy of each vector for the returned Node values.  Go's map iteration order (random) is
replaced by insertion (row) order wherever exact.go iterates the map, so ties resolve to the earlier-added key.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Sequence

import numpy as np

from . import _lib
from .distance import metric_of
from .graph import Node


class ExactIndex:
    def __init__(self, distance, device: int = 0, capacity: int = 0):
        self.distance = distance
        self._metric = metric_of(distance)  # only the two registered metrics have kernels; anything else raises
        self.device, self._capacity = device, capacity
        self._x = None
        self._dim = 0
        self._row_of = {}
        self._key_of: List = []
        self._vecs: List[np.ndarray] = []
        self.last_stats = None

    def _ensure(self, dim):
        if self._x is None:
            x = C.c_void_p()
            _lib.check(_lib.lib().b200exact_create(dim, self._metric, self.device, self._capacity, C.byref(x)))
            self._x, self._dim = x, dim

    # exact.go:27-38
    def Add(self, key, vector) -> None:
        self.BatchAdd([key], [vector])

    # exact.go:41-58
    def BatchAdd(self, keys: Sequence, vectors: Sequence) -> None:
        if len(keys) != len(vectors):
            raise ValueError(f"number of keys ({len(keys)}) does not match number of vectors ({len(vectors)})")
        if len(keys) == 0:
            return
        V = np.ascontiguousarray(np.stack([np.asarray(v, dtype=np.float32).reshape(-1) for v in vectors]))
        self._ensure(V.shape[1])
        if V.shape[1] != self._dim:
            raise ValueError(f"embedding dimension mismatch: {self._dim} != {V.shape[1]}")
        # map assignment overwrites an existing key (exact.go:36, 54): the old row is tombstoned; a key repeated inside
        # the batch keeps its last vector
        last = {}
        for i, k in enumerate(keys):
            last[k] = i
        keep = sorted(last.values())
        stale = [self._row_of[k] for k in last if k in self._row_of]
        if stale:
            self._remove_rows(stale)
        first = C.c_int64()
        Vk = np.ascontiguousarray(V[keep])
        _lib.check(_lib.lib().b200exact_add_batch(self._x, Vk.ctypes.data_as(_lib.f32p), len(keep), C.byref(first)))
        for j, i in enumerate(keep):
            row = first.value + j
            self._row_of[keys[i]] = row
            self._key_of.append(keys[i])
            self._vecs.append(Vk[j].copy())

    def add_arrays(self, keys, V) -> None:
        """Array-level BatchAdd for keys not seen before (what a cgo host calls after its own map bookkeeping)."""
        V = np.ascontiguousarray(V, dtype=np.float32)
        self._ensure(V.shape[1])
        if V.shape[1] != self._dim:
            raise ValueError(f"embedding dimension mismatch: {self._dim} != {V.shape[1]}")
        first = C.c_int64()
        _lib.check(_lib.lib().b200exact_add_batch(self._x, V.ctypes.data_as(_lib.f32p), V.shape[0], C.byref(first)))
        for j, key in enumerate(keys):
            self._row_of[key] = first.value + j
        self._key_of.extend(keys)
        self._vecs.extend(V)

    def search_device(self, d_queries: int, B: int, k: int, d_rows: int, d_dists: int, d_counts: int,
                      stream: int = 0) -> None:
        """Device-pointer search (rows int64 [B,k], dists f32 [B,k], counts i32 [B]) on `stream`.  The scan path is
        asynchronous; the tensor path synchronises the stream once inside the call (candidate-overflow check)."""
        _lib.check(_lib.lib().b200exact_search_batch_device(self._x, C.c_void_p(d_queries), B, k, C.c_void_p(d_rows),
                                                            C.c_void_p(d_dists), C.c_void_p(d_counts),
                                                            C.c_void_p(stream)))

    def search_device_noalloc(self, d_queries: int, B: int, k: int, d_rows: int, d_dists: int, d_counts: int,
                              stream: int = 0) -> None:
        """NOT in the reference (added for callers that cannot call cudaMalloc): the scan path with its workspace in a
        static 8 MiB device array; no allocation inside the call.  Fails past the array's size; calls that overlap on
        the GPU race (include/b200hnsw.h: b200exact_search_batch_noalloc_device).  B = 0 is the set-up call."""
        _lib.check(_lib.lib().b200exact_search_batch_noalloc_device(self._x, C.c_void_p(d_queries), B, k,
                                                                    C.c_void_p(d_rows), C.c_void_p(d_dists),
                                                                    C.c_void_p(d_counts), C.c_void_p(stream)))

    def _remove_rows(self, rows):
        r = np.ascontiguousarray(rows, dtype=np.int64)
        _lib.check(_lib.lib().b200exact_remove(self._x, r.ctypes.data_as(_lib.i64p), len(r)))

    # exact.go:61-111
    def search_rows(self, queries, k: int):
        """(rows[B,k] int64, dists[B,k], counts[B]) ascending by distance, ties in row order."""
        q = np.ascontiguousarray(queries, dtype=np.float32)
        if q.ndim == 1:
            q = q.reshape(1, -1)
        B = q.shape[0]
        rows = np.full((B, k), -1, dtype=np.int64)
        dists = np.full((B, k), np.nan, dtype=np.float32)
        counts = np.zeros(B, dtype=np.int32)
        if self._x is None:
            return rows, dists, counts
        if q.shape[1] != self._dim:
            raise ValueError(f"embedding dimension mismatch: {self._dim} != {q.shape[1]}")
        st = _lib.Stats()
        _lib.check(_lib.lib().b200exact_search_batch(self._x, q.ctypes.data_as(_lib.f32p), B, k,
                                                     rows.ctypes.data_as(_lib.i64p), dists.ctypes.data_as(_lib.f32p),
                                                     counts.ctypes.data_as(_lib.i32p), C.byref(st)))
        self.last_stats = st.as_dict()
        return rows, dists, counts

    def Search(self, query, k: int) -> List[Node]:
        if self.Len() == 0:
            return []
        rows, _, counts = self.search_rows(query, k)
        return [Node(self._key_of[r], self._vecs[r]) for r in rows[0][:counts[0]]]

    # exact.go:114-145
    def Delete(self, key) -> bool:
        return self.BatchDelete([key])[0]

    def BatchDelete(self, keys) -> List[bool]:
        out, rows = [], []
        for k in keys:
            r = self._row_of.pop(k, None)
            out.append(r is not None)
            if r is not None:
                rows.append(r)
        if rows:
            self._remove_rows(rows)
        return out

    # exact.go:148-153
    def Len(self) -> int:
        return 0 if self._x is None else int(_lib.lib().b200exact_len(self._x))

    # exact.go:156-160
    def Close(self) -> None:
        if self._x is not None:
            _lib.lib().b200exact_destroy(self._x)
            self._x = None
        self._row_of, self._key_of, self._vecs = {}, [], []

    def __del__(self):
        try:
            self.Close()
        except Exception:
            pass


def NewExactIndex(distance, **kw) -> ExactIndex:
    """exact.go:20-25"""
    return ExactIndex(distance, **kw)
