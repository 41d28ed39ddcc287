"""distance.go:11-46 -- DistanceFunc, CosineDistance, EuclideanDistance and the name registry.

The two registered functions are sentinels that select the device metric; evaluating one directly runs the device
kernel on a single pair (b200hnsw_distance_batch), so even the distance KATs go through the CUDA arithmetic.
Any other callable cannot be used with the device path (north_star: no CPU fallback) and is rejected with an error.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib


def _device_distance(metric: int, a, b, device: int = 0) -> np.float32:
    a = np.ascontiguousarray(a, dtype=np.float32).reshape(1, -1)
    b = np.ascontiguousarray(b, dtype=np.float32).reshape(1, -1)
    if a.shape != b.shape:
        raise ValueError(f"embedding dimension mismatch: {a.shape[1]} != {b.shape[1]}")
    out = np.empty(1, dtype=np.float32)
    _lib.check(_lib.lib().b200hnsw_distance_batch(device, metric, a.ctypes.data_as(_lib.f32p),
                                                  b.ctypes.data_as(_lib.f32p), 1, a.shape[1],
                                                  out.ctypes.data_as(_lib.f32p)))
    return out[0]


def CosineDistance(a, b) -> np.float32:
    """distance.go:15-17: 1 - cosine similarity."""
    return _device_distance(_lib.COSINE, a, b)


def EuclideanDistance(a, b) -> np.float32:
    """distance.go:20-23: sqrt(sum((a-b)^2))."""
    return _device_distance(_lib.EUCLIDEAN, a, b)


def distance_batch(metric: int, a, b, device: int = 0) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    assert a.shape == b.shape and a.ndim == 2
    out = np.empty(a.shape[0], dtype=np.float32)
    _lib.check(_lib.lib().b200hnsw_distance_batch(device, metric, a.ctypes.data_as(_lib.f32p),
                                                  b.ctypes.data_as(_lib.f32p), a.shape[0], a.shape[1],
                                                  out.ctypes.data_as(_lib.f32p)))
    return out


# distance.go:25-28
distanceFuncs = {"euclidean": EuclideanDistance, "cosine": CosineDistance}
_METRIC_OF_NAME = {"euclidean": _lib.EUCLIDEAN, "cosine": _lib.COSINE}


def distanceFuncToName(fn):
    """distance.go:30-39: reverse lookup by function identity."""
    for name, f in distanceFuncs.items():
        if f is fn:
            return name, True
    return "", False


def RegisterDistanceFunc(name: str, fn) -> None:
    """distance.go:44-46.  A registered custom function can be exported/imported by name, but only the two built-in
    metrics have device kernels; using another one in Add/Search raises."""
    distanceFuncs[name] = fn


def metric_of(fn) -> int:
    name, ok = distanceFuncToName(fn)
    if not ok or name not in _METRIC_OF_NAME or distanceFuncs[name] not in (EuclideanDistance, CosineDistance):
        raise ValueError(f"distance function {fn!r} has no device kernel (only EuclideanDistance and CosineDistance); "
                         f"hnsw_b200 has no CPU fallback")
    return _METRIC_OF_NAME[name]
