"""hybrid -- host mirror of the dispatch glue of hnsw-extensions/hybrid over the device indexes.

What is here is what sits either side of the hot path: `HybridIndex` (hybrid/hybrid.go:124-470) choosing between the
exact index (K3) and the graph (K1/K2) by `TotalVectors <= ExactThreshold`, and the `SearchableIndex` adapters with
`MultiIndexAdapter`'s merge (hybrid/adapter.go:10-354).  The reference's behaviour is kept, including that a
HybridIndexType index stops seeing its first ExactThreshold vectors once it switches to the graph (they were only ever
added to the exact index, hybrid.go:247-262 / 379-395).  LSH (hybrid/lsh.go) and the partitioner
(hybrid/partitioner.go) are separate algorithms outside the path: the reference computes a partition and LSH
candidates in HybridIndexType and throws both away (hybrid.go:253-254, 386-388), so leaving them out changes no
result; LSHIndexType itself is refused.  No CPU fallback: every search and distance below runs on the device.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np

from .distance import CosineDistance, distance_batch, metric_of
from .exact import ExactIndex
from .graph import Graph, Node

# hybrid.go:62-76
ExactIndexType, HNSWIndexType, LSHIndexType, HybridIndexType = 0, 1, 2, 3


@dataclass
class IndexConfig:
    """hybrid.go:84-105"""
    Type: int = ExactIndexType
    ExactThreshold: int = 0
    M: int = 0
    Ml: float = 0.0
    EfSearch: int = 0
    Distance: Optional[Callable] = None
    NumHashTables: int = 0
    NumHashBits: int = 0
    NumPartitions: int = 0
    PartitionSize: int = 0


def DefaultIndexConfig() -> IndexConfig:
    """hybrid.go:108-122"""
    return IndexConfig(Type=HybridIndexType, ExactThreshold=1000, M=16, Ml=0.25, EfSearch=20, Distance=CosineDistance,
                       NumHashTables=4, NumHashBits=8, NumPartitions=10, PartitionSize=10000)


@dataclass
class IndexStats:
    """hybrid.go:147-153"""
    TotalVectors: int = 0
    ExactVectors: int = 0
    HNSWVectors: int = 0
    NumPartitions: int = 0
    AvgPartitionSize: float = 0.0


class HybridIndex:
    """hybrid.go:124-470"""

    def __init__(self, config: Optional[IndexConfig] = None, device: int = 0):
        config = config if config is not None else IndexConfig()
        if config.Type == 0:  # hybrid.go:157-159: the zero value (which is also ExactIndexType) means "default"
            config = DefaultIndexConfig()
        if config.M <= 0:
            raise ValueError("M must be greater than 0")
        if config.Ml <= 0 or config.Ml >= 1:
            raise ValueError("Ml must be between 0 and 1")
        if config.EfSearch <= 0:
            raise ValueError("EfSearch must be greater than 0")
        self.config = config
        self.stats = IndexStats()
        self.vectors = {}
        self.exactIndex: Optional[ExactIndex] = None
        self.hnswIndex: Optional[Graph] = None
        if config.Type == LSHIndexType:
            raise NotImplementedError("LSHIndexType: hybrid/lsh.go is a separate algorithm outside the device path")
        if config.Type in (ExactIndexType, HybridIndexType):
            self.exactIndex = ExactIndex(config.Distance, device=device)
        if config.Type in (HNSWIndexType, HybridIndexType):
            self.hnswIndex = Graph(M=config.M, Ml=config.Ml, EfSearch=config.EfSearch, Distance=config.Distance,
                                   device=device)
            self.hnswIndex.Validate()
        if config.Type not in (ExactIndexType, HNSWIndexType, HybridIndexType):
            raise ValueError(f"unknown index type: {config.Type}")

    # hybrid.go:233-281
    def Add(self, key, vector) -> None:
        self.vectors[key] = vector
        self.stats.TotalVectors += 1
        t = self.config.Type
        if t == ExactIndexType:
            return self.exactIndex.Add(key, vector)
        if t == HNSWIndexType:
            return self.hnswIndex.Add(Node(key, np.asarray(vector, dtype=np.float32)))
        if self.stats.TotalVectors <= self.config.ExactThreshold:
            self.stats.ExactVectors += 1
            return self.exactIndex.Add(key, vector)
        self.stats.HNSWVectors += 1
        return self.hnswIndex.Add(Node(key, np.asarray(vector, dtype=np.float32)))

    # hybrid.go:284-355
    def BatchAdd(self, keys: Sequence, vectors: Sequence) -> None:
        if len(keys) != len(vectors):
            raise ValueError(f"number of keys ({len(keys)}) does not match number of vectors ({len(vectors)})")
        for k, v in zip(keys, vectors):
            self.vectors[k] = v
        self.stats.TotalVectors += len(keys)
        t = self.config.Type
        if t == ExactIndexType:
            return self.exactIndex.BatchAdd(keys, vectors)
        nodes = [Node(k, np.asarray(v, dtype=np.float32)) for k, v in zip(keys, vectors)]
        if t == HNSWIndexType:
            return self.hnswIndex.BatchAdd(nodes)  # the reference's loop of Add: same graph, one device call
        if self.stats.TotalVectors <= self.config.ExactThreshold:
            self.stats.ExactVectors += len(keys)
            return self.exactIndex.BatchAdd(keys, vectors)
        self.hnswIndex.BatchAdd(nodes)
        self.stats.HNSWVectors += len(keys)

    # hybrid.go:358-403
    def Search(self, query, k: int) -> List[Node]:
        t = self.config.Type
        if t == ExactIndexType:
            return self.exactIndex.Search(query, k)
        if t == HNSWIndexType:
            return self.hnswIndex.Search(query, k)
        if self.stats.TotalVectors <= self.config.ExactThreshold:
            return self.exactIndex.Search(query, k)
        return self.hnswIndex.Search(query, k)  # both larger tiers end in hnswIndex.Search (hybrid.go:383-395)

    def _route(self, n_total: int):
        t = self.config.Type
        if t == ExactIndexType or (t == HybridIndexType and n_total <= self.config.ExactThreshold):
            return self.exactIndex
        return self.hnswIndex

    def BatchSearch(self, queries, k: int) -> List[List[Node]]:
        """Not in the reference (its callers loop over Search): the same dispatch with one device call per batch."""
        idx = self._route(self.stats.TotalVectors)
        if isinstance(idx, Graph):
            return idx.BatchSearch(queries, k)
        if idx.Len() == 0:
            return [[] for _ in range(len(queries))]
        rows, _, counts = idx.search_rows(queries, k)
        return [[Node(idx._key_of[r], idx._vecs[r]) for r in rows[i][:counts[i]]] for i in range(len(counts))]

    # hybrid.go:406-437
    def Delete(self, key) -> bool:
        return self.BatchDelete([key])[0]

    # hybrid.go:440-477
    def BatchDelete(self, keys) -> List[bool]:
        keys = list(keys)
        t = self.config.Type
        if t == ExactIndexType:
            results = self.exactIndex.BatchDelete(keys)
        elif t == HNSWIndexType:
            results = self.hnswIndex.BatchDelete(keys)
        else:
            e = self.exactIndex.BatchDelete(keys)
            h = self.hnswIndex.BatchDelete(keys)
            results = [a or b for a, b in zip(e, h)]
        for k, ok in zip(keys, results):
            if ok:
                self.stats.TotalVectors -= 1
                self.vectors.pop(k, None)
        return results

    def Len(self) -> int:
        return self.stats.TotalVectors

    def Close(self) -> None:
        if self.exactIndex is not None:
            self.exactIndex.Close()
        if self.hnswIndex is not None:
            self.hnswIndex.close()
        self.vectors = None

    def GetStats(self) -> IndexStats:
        return self.stats


def NewHybridIndex(config: Optional[IndexConfig] = None, **kw) -> HybridIndex:
    return HybridIndex(config, **kw)


# ---- SearchableIndex adapters -- hybrid/adapter.go --------------------------------------------------------------------
def _distances(distance, query, nodes, device=0) -> np.ndarray:
    """distance(query, node.Value) for every node (adapter.go:58-66, 133-139), one device call."""
    if not nodes:
        return np.zeros(0, np.float32)
    q = np.asarray(query, dtype=np.float32).reshape(1, -1)
    V = np.stack([np.asarray(nd.Value, dtype=np.float32) for nd in nodes])
    return distance_batch(metric_of(distance), np.repeat(q, len(nodes), axis=0), V, device)


class HNSWAdapter:
    """adapter.go:10-96"""

    def __init__(self, graph: Graph, distance):
        self.graph, self.distance = graph, distance

    def Add(self, key, vector):
        return self.graph.Add(Node(key, np.asarray(vector, dtype=np.float32)))

    def BatchAdd(self, keys, vectors) -> list:
        if len(keys) != len(vectors):
            return [ValueError(f"number of keys ({len(keys)}) does not match number of vectors ({len(vectors)})")]
        errors = []
        for k, v in zip(keys, vectors):  # adapter.go:41-48: one Add, one error slot per key
            try:
                self.Add(k, v)
                errors.append(None)
            except Exception as e:  # noqa: BLE001 -- the reference returns the error value per key
                errors.append(e)
        return errors

    def Search(self, query, k: int) -> Tuple[list, np.ndarray]:
        try:
            nodes = self.graph.Search(query, k) or []
        except ValueError:
            return None, None  # adapter.go:53-56
        d = _distances(self.distance, query, nodes, self.graph.device) if self.distance is not None else \
            np.zeros(len(nodes), np.float32)
        return [nd.Key for nd in nodes], d

    def Delete(self, key) -> bool:
        return self.graph.Delete(key)

    def BatchDelete(self, keys) -> List[bool]:
        return self.graph.BatchDelete(keys)

    def Len(self) -> int:
        return self.graph.Len()

    def Close(self) -> None:
        pass  # adapter.go:93-95


class ExactAdapter:
    """adapter.go:98-170"""

    def __init__(self, index: ExactIndex):
        self.index = index

    def Add(self, key, vector):
        return self.index.Add(key, vector)

    def BatchAdd(self, keys, vectors) -> list:
        if len(keys) != len(vectors):
            return [ValueError(f"number of keys ({len(keys)}) does not match number of vectors ({len(vectors)})")]
        try:
            self.index.BatchAdd(keys, vectors)
            return [None] * len(keys)
        except Exception as e:  # noqa: BLE001
            return [e] * len(keys)

    def Search(self, query, k: int) -> Tuple[list, np.ndarray]:
        try:
            nodes = self.index.Search(query, k)
        except ValueError:
            return None, None
        d = _distances(self.index.distance, query, nodes, self.index.device)
        return [nd.Key for nd in nodes], d

    def Delete(self, key) -> bool:
        return self.index.Delete(key)

    def BatchDelete(self, keys) -> List[bool]:
        return self.index.BatchDelete(keys)

    def Len(self) -> int:
        return self.index.Len()

    def Close(self) -> None:
        self.index.Close()


@dataclass
class MultiIndexAdapter:
    """adapter.go:205-354"""
    indexes: list = field(default_factory=list)

    def Add(self, key, vector) -> None:
        for idx in self.indexes:
            idx.Add(key, vector)

    def BatchAdd(self, keys, vectors) -> list:
        if len(keys) != len(vectors):
            return [ValueError(f"number of keys ({len(keys)}) does not match number of vectors ({len(vectors)})")]
        errors = [None] * len(keys)
        for idx in self.indexes:
            for i, e in enumerate(idx.BatchAdd(keys, vectors)):
                if e is not None and errors[i] is None:
                    errors[i] = e
        return errors

    def Search(self, query, k: int) -> Tuple[list, np.ndarray]:
        """adapter.go:257-308: concatenate in index order, stable insertion sort on distance (strict >, so NaN and equal
        distances stay where they are), keep the first instance of each key, cut at k."""
        keys: list = []
        dists: List[np.float32] = []
        for idx in self.indexes:
            ks, ds = idx.Search(query, k)
            if ks is None:
                continue
            keys.extend(ks)
            dists.extend(np.float32(x) for x in ds)
        order = list(range(len(keys)))
        for i in range(1, len(order)):  # the reference's insertion sort, literally (NaN-safe where sorted() is not)
            j = i
            while j > 0 and dists[order[j - 1]] > dists[order[j]]:
                order[j - 1], order[j] = order[j], order[j - 1]
                j -= 1
        seen, out_k, out_d = set(), [], []
        for i in order:
            if keys[i] not in seen:
                seen.add(keys[i])
                out_k.append(keys[i])
                out_d.append(dists[i])
        return out_k[:k], np.asarray(out_d[:k], dtype=np.float32)

    def Delete(self, key) -> bool:
        ok = True
        for idx in self.indexes:
            if not idx.Delete(key):
                ok = False
        return ok

    def BatchDelete(self, keys) -> List[bool]:
        results = [True] * len(keys)
        for idx in self.indexes:
            for i, s in enumerate(idx.BatchDelete(keys)):
                if not s:
                    results[i] = False
        return results

    def Len(self) -> int:
        return self.indexes[0].Len() if self.indexes else 0

    def Close(self) -> None:
        for idx in self.indexes:
            idx.Close()
