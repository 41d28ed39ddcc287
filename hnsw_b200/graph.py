"""Host-side mirror of the reference's Graph[K] API (graph.go), routing the hot path through the C ABI.

Same names, argument meaning and error texts as the Go methods so the parity tests read like the reference's:
NewGraph / NewGraphWithConfig (graph.go:340-366), Graph.Add / BatchAdd (437-531, 942-1042), Search / BatchSearch /
ParallelSearch (534-625, 1047-1110, 631-826), Len / Dims / Lookup / Validate (419-429, 829-838, 898-937),
Export / Import (encode.go:133-262).  The host keeps what the reference's host would keep under cgo: validation,
the key <-> row mapping, the level RNG (graph.go:388-417), a host copy of the vectors for returning Node values,
and the test hacks of graph.go:563-569, 595-619.  Everything data-parallel runs on the device.
"""
from __future__ import annotations

import ctypes as C
import math
import threading
import time
from dataclasses import dataclass
from typing import Iterable, List, Optional, Sequence

import numpy as np

from . import _lib
from .distance import CosineDistance, EuclideanDistance, distanceFuncs, distanceFuncToName, metric_of
from .keys import KeyCodec


@dataclass
class Node:
    """graph.go:20-27"""
    Key: int
    Value: np.ndarray

    def __eq__(self, other):
        return isinstance(other, Node) and self.Key == other.Key and np.array_equal(self.Value, other.Value)


def MakeNode(key, vec) -> Node:
    return Node(key, np.asarray(vec, dtype=np.float32))


class _RWLock:
    """Graph.mu (graph.go:305-332 holds a sync.RWMutex: Add / BatchAdd / Delete / BatchDelete take it for writing, Search /
    BatchSearch / Lookup for reading).  The write side is re-entrant for its owner: BatchAdd replaces an existing key
    through BatchDelete (the reference's Add deadlocks there)."""

    def __init__(self):
        self._c = threading.Condition()
        self._readers = 0
        self._writer = None
        self._depth = 0

    def r_acquire(self) -> bool:
        me = threading.get_ident()
        with self._c:
            if self._writer == me:
                return False  # nested in my own write section
            while self._writer is not None:
                self._c.wait()
            self._readers += 1
            return True

    def r_release(self):
        with self._c:
            self._readers -= 1
            if self._readers == 0:
                self._c.notify_all()

    def w_acquire(self):
        me = threading.get_ident()
        with self._c:
            if self._writer == me:
                self._depth += 1
                return
            while self._writer is not None or self._readers:
                self._c.wait()
            self._writer, self._depth = me, 1

    def w_release(self):
        with self._c:
            self._depth -= 1
            if self._depth == 0:
                self._writer = None
                self._c.notify_all()


def _writes(fn):
    def wrapped(self, *a, **kw):
        self._mu.w_acquire()
        try:
            return fn(self, *a, **kw)
        finally:
            self._mu.w_release()
    wrapped.__name__, wrapped.__doc__ = fn.__name__, fn.__doc__
    return wrapped


def _reads(fn):
    def wrapped(self, *a, **kw):
        mine = self._mu.r_acquire()
        try:
            return fn(self, *a, **kw)
        finally:
            if mine:
                self._mu.r_release()
    wrapped.__name__, wrapped.__doc__ = fn.__name__, fn.__doc__
    return wrapped


class LevelRand:
    """What Graph.Rng has to be: anything with Float64() (the one method randomLevel, graph.go:388-417, calls on its
    *rand.Rand).  This one is a numpy PCG64 stream, NOT Go's math/rand: a seed gives a reproducible stream here, not the
    levels the Go reference draws from the same seed (graph_test.go:86-133 therefore cannot be replayed value for value;
    tests pass explicit levels wherever they compare with the oracle)."""

    def __init__(self, seed: Optional[int] = None):
        self._g = np.random.default_rng(seed if seed is not None else time.time_ns())

    def Float64(self) -> float:
        return float(self._g.random())


def maxLevel(ml: float, numNodes: int) -> int:
    """graph.go:370-385"""
    if ml == 0:
        raise ValueError("ml must be greater than 0")
    if numNodes == 0:
        return 1
    l = math.log(float(numNodes)) / math.log(1 / ml)
    # Go math.Round: half away from zero
    return int(math.floor(l + 0.5) if l >= 0 else math.ceil(l - 0.5)) + 1


class Graph:
    """graph.go:305-332.  Public mutable fields: Distance, Rng, M, Ml, EfSearch."""

    def __init__(self, M=16, Ml=0.25, EfSearch=20, Distance=CosineDistance, Rng=None, device: int = 0,
                 capacity: int = 0):
        self.M, self.Ml, self.EfSearch, self.Distance = M, Ml, EfSearch, Distance
        self.Rng = Rng if Rng is not None else LevelRand()
        self.device = device
        self._capacity = capacity
        self._h = None          # b200hnsw_t*
        self._mu = _RWLock()    # Graph.mu
        self._dim = 0
        self._max_m = 0
        self._keys: List[int] = []          # row -> key
        self._row_of = {}                   # key -> row (live keys only)
        self._retired = set()               # keys removed by Delete
        self._vecs = np.zeros((0, 0), np.float32)  # host copy, grows by doubling
        self._codec = KeyCodec()            # K (cmp.Ordered) -> the int64 the device orders neighbour lists by
        self.last_stats = None

    # ---- lifetime -------------------------------------------------------------------------------------------
    def _ensure_device(self, dim: int):
        if self._h is not None:
            return
        cfg = _lib.Config(dim=dim, max_m=min(max(self.M, 1), _lib.MAX_M), device=self.device, reserved=0,
                          capacity=self._capacity)
        h = C.c_void_p()
        _lib.check(_lib.lib().b200hnsw_create(C.byref(cfg), C.byref(h)))
        self._h, self._dim, self._max_m = h, dim, cfg.max_m
        if getattr(self, "_build_mode", 0):
            _lib.check(_lib.lib().b200hnsw_set_build_mode(self._h, self._build_mode))
        if getattr(self, "_add_window", 0):
            _lib.check(_lib.lib().b200hnsw_set_add_window(self._h, self._add_window))
        self._vecs = np.zeros((max(self._capacity, 1024), dim), np.float32)

    def close(self):
        if self._h is not None:
            _lib.lib().b200hnsw_destroy(self._h)
            self._h = None
        self._keys, self._row_of, self._dim = [], {}, 0
        self._codec = KeyCodec()
        self._retired = set()
        self._vecs = np.zeros((0, 0), np.float32)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- graph.go:916-937 ---------------------------------------------------------------------------------------
    def Validate(self):
        if self.M <= 0:
            raise ValueError(f"M must be greater than 0, got {self.M}")
        if self.Ml <= 0 or self.Ml >= 1:
            raise ValueError(f"Ml must be between 0 and 1 (exclusive), got {self.Ml:f}")
        if self.EfSearch <= 0:
            raise ValueError(f"EfSearch must be greater than 0, got {self.EfSearch}")
        if self.Distance is None:
            raise ValueError("Distance function must be set")

    def Len(self) -> int:
        return 0 if self._h is None else int(_lib.lib().b200hnsw_len(self._h))

    def Dims(self) -> int:
        return self._dim if self.Len() > 0 else 0

    @_reads
    def Lookup(self, key):
        row = self._row_of.get(key)
        if row is None:
            return None, False
        return self._vecs[row], True

    def Topography(self) -> List[int]:
        """analyzer.go: nodes per layer."""
        if self._h is None:
            return []
        L = _lib.lib()
        return [int(L.b200hnsw_layer_size(self._h, l)) for l in range(L.b200hnsw_layers(self._h))]

    # ---- graph.go:388-417 -----------------------------------------------------------------------------------------
    def randomLevel(self) -> int:
        mx = 1
        if self._h is not None and _lib.lib().b200hnsw_layers(self._h) > 0:
            if self.Ml == 0:
                raise ValueError("(*Graph).Ml must be greater than 0")
            mx = maxLevel(self.Ml, self.Len())
        for level in range(mx):
            if self.Rng is None:
                self.Rng = LevelRand()
            if self.Rng.Float64() > self.Ml:
                return level
        return mx

    # ---- Add / BatchAdd ---------------------------------------------------------------------------------------------
    def _store_host(self, keys: Sequence[int], vecs: np.ndarray):
        n0 = len(self._keys)
        need = n0 + len(keys)
        if need > self._vecs.shape[0]:
            grown = np.zeros((max(need, 2 * self._vecs.shape[0]), self._dim), np.float32)
            grown[:n0] = self._vecs[:n0]
            self._vecs = grown
        self._vecs[n0:need] = vecs
        keys = list(keys)
        self._row_of.update(zip(keys, range(n0, need)))
        self._keys.extend(keys)

    def _draw_levels(self, n: int) -> np.ndarray:
        """randomLevel() for n consecutive insertions: the bound maxLevel(Ml, len(layer0)) grows as nodes land."""
        out = np.empty(n, dtype=np.int32)
        have_layers = self._h is not None and _lib.lib().b200hnsw_layers(self._h) > 0
        size0 = self.Len()
        for i in range(n):
            mx = maxLevel(self.Ml, size0) if have_layers else 1
            lvl = mx
            for level in range(mx):
                if self.Rng.Float64() > self.Ml:
                    lvl = level
                    break
            out[i] = lvl
            have_layers = True
            size0 += 1
        return out

    @_writes
    def BatchAdd(self, nodes: Iterable[Node], levels: Optional[Sequence[int]] = None):
        """graph.go:942-1042.  `levels` (testing hook) replaces the Rng draws."""
        self.Validate()
        nodes = list(nodes)
        if not nodes:
            return
        metric = metric_of(self.Distance)
        vecs = [np.asarray(nd.Value, dtype=np.float32).reshape(-1) for nd in nodes]
        # graph.go:450-455 checks against Dims() (0 for an empty graph).  The device slab keeps its row width once it
        # exists, so a graph whose nodes were all deleted still takes vectors of that width only.
        dim0 = self._dim if self._h is not None else len(vecs[0])
        for v in vecs:
            if len(v) != dim0:  # graph.go:450-455
                raise ValueError(f"embedding dimension mismatch: {dim0} != {len(v)}")
        self._ensure_device(dim0)
        if self.M > self._max_m:
            raise ValueError(f"M={self.M} exceeds the device list stride {self._max_m} fixed at the first Add")
        lv_all = None if levels is None else list(levels)
        if lv_all is not None and len(lv_all) != len(nodes):
            raise ValueError("levels must have one entry per node")

        def norm(k):
            self._codec.check(k)  # int, float, str or bytes, the same kind for the whole graph (Go: the type parameter K)
            if isinstance(k, (int, np.integer)):
                return int(k)
            if isinstance(k, (float, np.floating)):
                return float(k)
            return k

        def flush(lo, hi):
            if hi > lo:
                self._add_keys([norm(nd.Key) for nd in nodes[lo:hi]], np.stack(vecs[lo:hi]),
                               None if lv_all is None else lv_all[lo:hi], metric)

        start, seen = 0, set()
        for i, nd in enumerate(nodes):
            k = norm(nd.Key)
            if k in self._row_of or k in seen:
                # "If another node with the same ID exists, it is replaced" (graph.go:435-436).  The reference deletes the
                # old node in the middle of the insertion (graph.go:1016-1024: isolate on every layer) and then fails its
                # own invariant -- Len() did not grow -- so BatchAdd returns "node not added" with the replacement done and
                # the rest of the batch untouched (graph.go:1035-1037); Add deadlocks (graph.go:438 + 844).  Here: the
                # nodes in front go in, the old node is deleted, the new one inserted under a new row, and the
                # reference's error is raised.  The delete runs in FRONT of the insertion, not inside it, so the link
                # structure is Delete + Add's, which is what the doc comment promises.
                flush(start, i)
                self.BatchDelete([k])
                self._retired.discard(k)
                flush(i, i + 1)
                if self.REPLACE_RAISES:
                    raise ValueError("node not added")
                start = i + 1
                seen = set()
                continue
            seen.add(k)
        flush(start, len(nodes))

    # graph.go:1035-1037: the reference's BatchAdd reports "node not added" after replacing an existing key; False makes
    # the replacement silent and lets the batch continue
    REPLACE_RAISES = True

    def _add_keys(self, keys, V, levels, metric):
        """Insert nodes whose keys are new to the graph (or were deleted: the key then names the new row; lists that
        still hold the deleted row under that key are overwritten when the new row reaches them, as the Go map would)."""
        for k in keys:
            self._retired.discard(k)
        if self._codec.kind == "int":
            self._add_arrays(keys, V, levels, metric)
            return
        # other ordered key types: the device sees order-preserving int64 codes (hnsw_b200/keys.py)
        known = getattr(self._codec, "_code_of", {})
        fresh = [k for k in keys if self._codec.kind != "ordered" or k not in known]
        fresh_codes, respaced = self._codec.assign(fresh)
        by_key = dict(zip(fresh, (int(c) for c in fresh_codes)))
        # a key that was deleted and comes back keeps its code (the codec never forgets a stored key)
        codes = np.asarray([by_key[k] if k in by_key else (respaced or known)[k] for k in keys], dtype=np.int64)
        try:
            if respaced:  # a gap between codes ran out: every stored row gets its new code
                allc = np.asarray([respaced[k] for k in self._keys], dtype=np.int64)
                _lib.check(_lib.lib().b200hnsw_set_keys(self._h, 0, len(allc), allc.ctypes.data_as(_lib.i64p)))
            self._add_arrays(codes, V, levels, metric, user_keys=keys)
        except BaseException:
            self._codec.forget_new([k for k in keys if k not in self._row_of])  # nodes in front of a failing one are in
            raise

    def _add_arrays(self, keys, V, levels, metric, wave_max=0, user_keys=None):
        """keys: the int64 order keys the device gets; user_keys: what the host mirror stores for them (the same unless
        K is not an integer type)."""
        lv = np.ascontiguousarray(levels, dtype=np.int32) if levels is not None else self._draw_levels(len(keys))
        V = np.ascontiguousarray(V, dtype=np.float32)
        K = np.ascontiguousarray(keys, dtype=np.int64)
        L = _lib.lib()
        rows0 = int(L.b200hnsw_rows(self._h))
        n0 = len(self._keys)
        t0 = time.perf_counter()
        # the host mirror (keys, key -> row, vectors: what Search hands back as Node.Value) is filled while the device
        # call runs; the call does not hold the interpreter lock
        host_keys = list(user_keys) if user_keys is not None else K.tolist()
        mirror_err = []

        def fill_mirror():
            try:
                self._store_host(host_keys, V)
            except BaseException as e:  # a thread does not propagate exceptions: hand it to the caller
                mirror_err.append(e)

        mirror = threading.Thread(target=fill_mirror)
        mirror.start()
        total = None
        try:
            off, stalled = 0, False
            while True:
                st = _lib.Stats()
                try:
                    if wave_max and wave_max > 0:  # batched construction in waves (not the reference's sequential order)
                        _lib.check(L.b200hnsw_add_batch_waves(
                            self._h, V[off:].ctypes.data_as(_lib.f32p), K[off:].ctypes.data_as(_lib.i64p),
                            lv[off:].ctypes.data_as(_lib.i32p), len(K) - off, self.M, self.EfSearch, metric,
                            int(wave_max), C.byref(st)))
                    else:
                        _lib.check(L.b200hnsw_add_batch(
                            self._h, V[off:].ctypes.data_as(_lib.f32p), K[off:].ctypes.data_as(_lib.i64p),
                            lv[off:].ctypes.data_as(_lib.i32p), len(K) - off, self.M, self.EfSearch, metric,
                            C.byref(st)))
                except _lib.B200Error as e:
                    done = int(L.b200hnsw_rows(self._h)) - rows0
                    # rows that can yield NaN distances entered the graph with this batch: the handle now runs the full
                    # candidate queue; the nodes that were not inserted go in again (include/b200hnsw.h)
                    if e.code == _lib.ERR_STATE and "add the rest again" in str(e) and (done > off or not stalled):
                        stalled = done == off  # no progress is fine once: the handle learned about the rows only now
                        off = done
                        continue
                    raise
                d = st.as_dict()
                total = d if total is None else {k: total[k] + d[k] for k in d}
                break
        except BaseException:
            mirror.join()
            done = int(L.b200hnsw_rows(self._h)) - rows0  # nodes in front of the failing one are in
            for key in self._keys[n0 + done:]:
                self._row_of.pop(key, None)
            del self._keys[n0 + done:]
            raise
        self.last_stats = total
        self.last_stats["call_s"] = time.perf_counter() - t0  # wall time of the C call (staging + kernels)
        mirror.join()
        if mirror_err:
            raise mirror_err[0]

    @_writes
    def add_arrays(self, keys, vectors, levels: Optional[Sequence[int]] = None, wave_max: int = 0):
        """BatchAdd for array inputs (keys[n] int64, vectors[n, dim] float32): same semantics, no per-node objects.
        Keys must be new and distinct.  wave_max > 0 selects the wave-batched construction
        (b200hnsw_add_batch_waves): parallel, deterministic, NOT the reference's sequential graph."""
        self.Validate()
        vectors = np.asarray(vectors, dtype=np.float32)
        if vectors.ndim != 2 or len(keys) != vectors.shape[0]:
            raise ValueError("keys and vectors must have the same length")
        if len(keys) == 0:
            return
        dim0 = self._dim if self._h is not None else vectors.shape[1]  # see BatchAdd
        if vectors.shape[1] != dim0:
            raise ValueError(f"embedding dimension mismatch: {dim0} != {vectors.shape[1]}")
        metric = metric_of(self.Distance)
        if self._codec.kind not in (None, "int"):
            raise TypeError("add_arrays takes int64 keys; this graph's key type is " + self._codec.kind)
        self._codec.kind = "int"
        self._ensure_device(dim0)
        if self.M > self._max_m:
            raise ValueError(f"M={self.M} exceeds the device list stride {self._max_m} fixed at the first Add")
        self._add_arrays(keys, vectors, levels, metric, wave_max)

    # ---- Delete / BatchDelete -------------------------------------------------------------------------------------
    @_writes
    def BatchDelete(self, keys) -> List[bool]:
        """graph.go:871-895.  The reference's quirks are kept (see b200hnsw_delete_batch in include/b200hnsw.h): a
        deleted node can stay linked one-way and Search can still return it; Lookup and Len stop seeing it."""
        keys = list(keys)
        if self._h is None or _lib.lib().b200hnsw_layers(self._h) == 0 or not keys:
            return [False] * len(keys)
        def row_of(k):
            try:
                return self._row_of.get(k, -1)
            except TypeError:  # unhashable: cannot be a key
                return -1
        rows = np.array([row_of(k) for k in keys], dtype=np.int32)
        out = np.zeros(len(keys), dtype=np.uint8)
        st = _lib.Stats()
        _lib.check(_lib.lib().b200hnsw_delete_batch(self._h, rows.ctypes.data_as(_lib.i32p), len(keys), self.M,
                                                    out.ctypes.data_as(_lib.u8p), C.byref(st)))
        self.last_stats = st.as_dict()
        for k, ok in zip(keys, out):
            if ok:
                del self._row_of[k]
                self._retired.add(k)
        return [bool(x) for x in out]

    @_writes
    def Delete(self, key) -> bool:
        """graph.go:843-866."""
        return self.BatchDelete([key])[0]

    @property
    def visited_capacity(self) -> int:
        """Entries of the per-traversal visited set the device-pointer entry points use (b200hnsw_get_visited_capacity)."""
        if self._h is None:
            raise RuntimeError("the graph has no device handle yet (nothing was added or loaded)")
        return int(_lib.lib().b200hnsw_get_visited_capacity(self._h))

    @visited_capacity.setter
    def visited_capacity(self, entries: int):
        if self._h is None:
            raise RuntimeError("the graph has no device handle yet (nothing was added or loaded)")
        _lib.check(_lib.lib().b200hnsw_set_visited_capacity(self._h, int(entries)))

    def set_build_mode(self, mode: int):
        """NOT the reference: the layer search the wave-batched Add runs (_lib.MODE_EXTENDED = textbook search)."""
        self._build_mode = int(mode)
        if self._h is not None:
            _lib.check(_lib.lib().b200hnsw_set_build_mode(self._h, self._build_mode))

    def set_add_window(self, window: int):
        """The SCHEDULE of Add / BatchAdd, never the graph they build (include/b200hnsw.h: b200hnsw_set_add_window):
        0 = automatic, 1 = one node at a time on one CTA, >= 2 = that many insertions speculated in parallel and committed
        in input order."""
        self._add_window = int(window)
        if self._h is not None:
            _lib.check(_lib.lib().b200hnsw_set_add_window(self._h, self._add_window))

    def add_window_stats(self):
        """(speculations accepted, insertions run again by the committer, speculations run again by their workers) since
        the handle was created."""
        if self._h is None:
            return 0, 0, 0
        a = np.zeros(3, np.int64)
        p = a.ctypes.data_as(_lib.i64p)
        _lib.check(_lib.lib().b200hnsw_get_add_window_stats(
            self._h, p, C.cast(C.addressof(p.contents) + 8, _lib.i64p), C.cast(C.addressof(p.contents) + 16, _lib.i64p)))
        return int(a[0]), int(a[1]), int(a[2])

    @_writes
    def Add(self, *nodes: Node, levels: Optional[Sequence[int]] = None):
        """graph.go:437-531 (same body as BatchAdd; the reference's Add deadlocks on a duplicate key, F9)."""
        self.BatchAdd(nodes, levels=levels)

    # ---- load a graph built elsewhere (Import, or the oracle in tests) ----------------------------------------------
    @_writes
    def load_flat(self, vectors, keys, levels, layer_lists, entries=None):
        """layer_lists[l] = (rows[N_l], ptr[N_l+1], nbr[...]) CSR of layer l.  Replaces the content."""
        vectors = np.ascontiguousarray(vectors, dtype=np.float32)
        n, dim = vectors.shape
        self._ensure_device(dim)
        user_keys = list(keys)
        codec = KeyCodec()
        for k in user_keys[:1]:
            codec.check(k)
        if codec.kind in (None, "int"):
            codec.kind = "int"
            user_keys = [int(k) for k in user_keys]
            keys = np.ascontiguousarray(user_keys, dtype=np.int64)
        else:  # float / str / bytes keys: the device orders rows by order-preserving codes (hnsw_b200/keys.py)
            if codec.kind == "float":
                user_keys = [float(k) for k in user_keys]
            codes, _ = codec.assign(user_keys)
            keys = np.ascontiguousarray(codes, dtype=np.int64)
        levels = np.ascontiguousarray(levels, dtype=np.int32)
        nl = len(layer_lists)
        sizes = np.array([len(r) for r, _, _ in layer_lists], dtype=np.int64)
        rows = np.ascontiguousarray(np.concatenate([np.asarray(r, np.int32) for r, _, _ in layer_lists])
                                    if nl else np.zeros(0, np.int32))
        ptr = np.ascontiguousarray(np.concatenate([np.asarray(p, np.int64) for _, p, _ in layer_lists])
                                   if nl else np.zeros(0, np.int64))
        nbr = np.ascontiguousarray(np.concatenate([np.asarray(b, np.int32) for _, _, b in layer_lists])
                                   if nl else np.zeros(0, np.int32))
        ent = None
        if entries is not None:
            ent = np.ascontiguousarray(entries, dtype=np.int32)
        _lib.check(_lib.lib().b200hnsw_load(
            self._h, n, vectors.ctypes.data_as(_lib.f32p), keys.ctypes.data_as(_lib.i64p),
            levels.ctypes.data_as(_lib.i32p), nl, ent.ctypes.data_as(_lib.i32p) if ent is not None else None,
            sizes.ctypes.data_as(_lib.i64p), rows.ctypes.data_as(_lib.i32p), ptr.ctypes.data_as(_lib.i64p),
            nbr.ctypes.data_as(_lib.i32p)))
        self._codec = codec
        self._keys = user_keys
        self._row_of = {k: i for i, k in enumerate(self._keys)}
        self._retired = set()
        self._vecs = vectors.copy()

    def entries(self) -> np.ndarray:
        L = _lib.lib()
        return np.array([L.b200hnsw_entry(self._h, l) for l in range(L.b200hnsw_layers(self._h))], dtype=np.int32)

    def set_entry(self, layer: int, row: int):
        _lib.check(_lib.lib().b200hnsw_set_entry(self._h, layer, row))

    def neighbors(self, layer: int):
        """(nbr[N][max_m] rows padded with -1, member[N]) for a layer -- used by Export / Analyzer / parity tests."""
        n = len(self._keys)
        nbr = np.empty((n, self._max_m), dtype=np.int32)
        mem = np.empty(n, dtype=np.uint8)
        _lib.check(_lib.lib().b200hnsw_get_neighbors(self._h, layer, 0, n, nbr.ctypes.data_as(_lib.i32p),
                                                     mem.ctypes.data_as(_lib.u8p)))
        return nbr, mem.astype(bool)

    def dump_flat(self):
        """Inverse of load_flat: (vectors, keys, levels, layer_lists, entries) read back from the device."""
        L = _lib.lib()
        n = len(self._keys)
        vec = np.empty((n, self._dim), dtype=np.float32)
        keys = np.empty(n, dtype=np.int64)
        levels = np.empty(n, dtype=np.int32)
        _lib.check(L.b200hnsw_get_rows(self._h, 0, n, vec.ctypes.data_as(_lib.f32p), keys.ctypes.data_as(_lib.i64p),
                                       levels.ctypes.data_as(_lib.i32p)))
        layer_lists = []
        for l in range(L.b200hnsw_layers(self._h)):
            nbr, member = self.neighbors(l)
            rows = np.nonzero(member)[0].astype(np.int32)
            sub = nbr[rows]
            deg = (sub >= 0).sum(axis=1)
            ptr = np.concatenate([[0], np.cumsum(deg)]).astype(np.int64)
            layer_lists.append((rows, ptr, sub[sub >= 0].astype(np.int32)))
        return vec, keys, levels, layer_lists, self.entries()

    # ---- Search / BatchSearch -----------------------------------------------------------------------------------------
    @_reads
    def search_rows(self, queries, k: int, ef: Optional[int] = None, mode: int = _lib.MODE_REFERENCE, entries=None):
        """Raw device call: (rows[B,k], dists[B,k], counts[B]) in the reference's heap-array order."""
        q = np.ascontiguousarray(queries, dtype=np.float32)
        if q.ndim == 1:
            q = q.reshape(1, -1)
        B = q.shape[0]
        rows = np.empty((B, k), dtype=np.int32)
        dists = np.empty((B, k), dtype=np.float32)
        counts = np.empty(B, dtype=np.int32)
        st = _lib.Stats()
        ent = None if entries is None else np.ascontiguousarray(entries, dtype=np.int32)
        _lib.check(_lib.lib().b200hnsw_search_batch(
            self._h, q.ctypes.data_as(_lib.f32p), B, k, ef if ef is not None else self.EfSearch,
            metric_of(self.Distance), mode, ent.ctypes.data_as(_lib.i32p) if ent is not None else None,
            rows.ctypes.data_as(_lib.i32p), dists.ctypes.data_as(_lib.f32p), counts.ctypes.data_as(_lib.i32p),
            C.byref(st)))
        self.last_stats = st.as_dict()
        return rows, dists, counts

    def search_device(self, d_queries: int, B: int, k: int, d_rows: int, d_dists: int, d_counts: int,
                      d_counters: int = 0, stream: int = 0, ef: Optional[int] = None,
                      mode: int = _lib.MODE_REFERENCE):
        """Device-pointer variant (b200hnsw_search_batch_device): inputs already resident in HBM, asynchronous on
        `stream` (a cudaStream_t as int).  d_counters: int64[8] or 0."""
        _lib.check(_lib.lib().b200hnsw_search_batch_device(
            self._h, C.c_void_p(d_queries), B, k, ef if ef is not None else self.EfSearch, metric_of(self.Distance),
            mode, None, C.c_void_p(d_rows), C.c_void_p(d_dists), C.c_void_p(d_counts),
            C.c_void_p(d_counters) if d_counters else None, C.c_void_p(stream) if stream else None))

    def search_segmented_device(self, d_queries: int, B: int, k: int, d_packed: int, seg_len: int, seg_stride: int,
                                d_counters: int = 0, stream: int = 0, ef: Optional[int] = None,
                                mode: int = _lib.MODE_REFERENCE):
        """b200hnsw_search_batch_segmented_device: results of queries [s*seg_len, (s+1)*seg_len) go to the block at
        d_packed + s*seg_stride (int32 elements) as rows | dists | counts -- the send buffer of the all-to-all."""
        _lib.check(_lib.lib().b200hnsw_search_batch_segmented_device(
            self._h, C.c_void_p(d_queries), B, k, ef if ef is not None else self.EfSearch, metric_of(self.Distance),
            mode, None, C.c_void_p(d_packed), seg_len, seg_stride,
            C.c_void_p(d_counters) if d_counters else None, C.c_void_p(stream) if stream else None))

    def search_host_ptrs(self, q_ptr: int, B: int, k: int, rows_ptr: int, dists_ptr: int, counts_ptr: int,
                         ef: Optional[int] = None, mode: int = _lib.MODE_REFERENCE):
        """b200hnsw_search_batch on caller-owned HOST buffers given as addresses (e.g. pinned torch tensors): the
        call a cgo host makes.  H2D copy, traversal kernel and D2H copy all happen inside; returns the stats."""
        st = _lib.Stats()
        _lib.check(_lib.lib().b200hnsw_search_batch(
            self._h, C.cast(q_ptr, _lib.f32p), B, k, ef if ef is not None else self.EfSearch,
            metric_of(self.Distance), mode, None, C.cast(rows_ptr, _lib.i32p), C.cast(dists_ptr, _lib.f32p),
            C.cast(counts_ptr, _lib.i32p), C.byref(st)))
        return st.as_dict()

    def _nodes(self, rows, count) -> List[Node]:
        return [Node(self._keys[r], self._vecs[r]) for r in rows[:count]]

    @_reads
    def Search(self, near, k: int) -> List[Node]:
        """graph.go:534-625"""
        self.Validate()
        if k <= 0:
            raise ValueError(f"k must be greater than 0, got {k}")
        near = np.asarray(near, dtype=np.float32).reshape(-1)
        if self.Len() > 0 and self.Dims() != len(near):
            raise ValueError(f"embedding dimension mismatch: {self.Dims()} != {len(near)}")
        if self._h is None or _lib.lib().b200hnsw_layers(self._h) == 0:
            return None  # graph.go:554-556: (nil, nil)
        if self.Len() == 0:
            return []    # every node deleted: the layers are still there, their entry points are nil
        ef = self.EfSearch
        # graph.go:563-569: the reference's hard-wired "dog query" test hack doubles efSearch
        is_dog = len(near) == 3 and near[0] == np.float32(1.0) and near[1] == np.float32(0.2) and \
            near[2] == np.float32(0.1)
        if is_dog:
            ef *= 2
        rows, _, counts = self.search_rows(near, k, ef)
        out = self._nodes(rows[0], counts[0])
        if is_dog and len(out) == 3 and all(nd.Key != 3 for nd in out) and 3 in self._row_of:
            out[2] = Node(3, self._vecs[self._row_of[3]])  # graph.go:595-619
        return out

    @_reads
    def BatchSearch(self, queries, k: int) -> List[List[Node]]:
        """graph.go:1047-1110 (no dog hack in the reference's batch loop)."""
        self.Validate()
        if k <= 0:
            raise ValueError(f"k must be greater than 0, got {k}")
        qs = [np.asarray(q, dtype=np.float32).reshape(-1) for q in queries]
        if self.Len() > 0:
            for i, q in enumerate(qs):
                if self.Dims() != len(q):
                    raise ValueError(f"embedding dimension mismatch for query {i}: {self.Dims()} != {len(q)}")
        if self.Len() == 0:
            return [None] * len(qs)
        if not qs:
            return []
        rows, _, counts = self.search_rows(np.stack(qs), k)
        return [self._nodes(rows[i], counts[i]) for i in range(len(qs))]

    # ---- graph.go:1116-1537 (host re-ranking over device searches and device distances: negatives.py) -----------------
    @_reads
    def SearchWithNegative(self, near, negative, k: int, negWeight: float):
        from .negatives import search_with_negative
        return search_with_negative(self, near, negative, k, negWeight)

    @_reads
    def SearchWithNegatives(self, near, negatives, k: int, negWeight: float):
        from .negatives import search_with_negatives
        return search_with_negatives(self, near, negatives, k, negWeight)

    @_reads
    def BatchSearchWithNegatives(self, queries, negatives, k: int, negWeight: float):
        from .negatives import batch_search_with_negatives
        return batch_search_with_negatives(self, queries, negatives, k, negWeight)

    # ---- encode.go:133-262 -----------------------------------------------------------------------------------------
    @_reads
    def Export(self, w) -> None:
        from .encode import export_graph
        export_graph(self, w)

    @_writes
    def Import(self, r, key_type=None) -> None:
        """key_type (int, float, str, bytes) plays Go's type parameter K when the graph has no keys yet."""
        from .encode import import_graph
        import_graph(self, r, key_type)

    @_reads
    def ParallelSearch(self, near, k: int, numWorkers: int = 0) -> List[Node]:
        """graph.go:631-826.  The reference's goroutine fan-out consumes results in channel-arrival order
        (non-deterministic); the deterministic branch (graph.go:791-809) is Search, which is what runs here."""
        return self.Search(near, k)


def NewGraph(**kw) -> Graph:
    """graph.go:340-349: M=16, Ml=0.25, CosineDistance, EfSearch=20."""
    return Graph(M=16, Ml=0.25, EfSearch=20, Distance=CosineDistance, **kw)


def NewGraphWithConfig(m: int, ml: float, efSearch: int, distance, **kw) -> Graph:
    """graph.go:352-366: validates and returns the error."""
    g = Graph(M=m, Ml=ml, EfSearch=efSearch, Distance=distance, **kw)
    g.Validate()
    return g
