"""ctypes loader for the C-ABI library (hnsw_b200/lib/libb200hnsw.so, declared in include/b200hnsw.h).

There is no CPU fallback: if the library is missing, or no CUDA device is present when a compute call is made,
the call raises.  Nothing here imports the oracle.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_PKG = Path(__file__).resolve().parent
# B200HNSW_LIB: another build of the same CUDA library (kernel tuning experiments); never a CPU path
LIB_PATH = Path(os.environ["B200HNSW_LIB"]) if os.environ.get("B200HNSW_LIB") else _PKG / "lib" / "libb200hnsw.so"

OK = 0
ERR_BAD_ARG, ERR_OOM, ERR_CUDA, ERR_UNSUPPORTED, ERR_VISITED_OVERFLOW, ERR_STATE = -1, -2, -3, -4, -5, -6
EUCLIDEAN, COSINE = 0, 1
MODE_REFERENCE, MODE_EXTENDED = 0, 1
MAX_LAYERS, MAX_M, MAX_K, MAX_EF = 64, 64, 256, 1024

f32p = C.POINTER(C.c_float)
i32p = C.POINTER(C.c_int32)
i64p = C.POINTER(C.c_int64)
u8p = C.POINTER(C.c_uint8)


class Config(C.Structure):
    _fields_ = [("dim", C.c_int32), ("max_m", C.c_int32), ("device", C.c_int32), ("reserved", C.c_int32),
                ("capacity", C.c_int64)]


class Stats(C.Structure):
    _fields_ = [("dist_evals", C.c_int64), ("expansions", C.c_int64), ("nbr_reads", C.c_int64),
                ("evict_evals", C.c_int64), ("replenish_evals", C.c_int64), ("evictions", C.c_int64),
                ("kernel_ms", C.c_float), ("launches", C.c_int32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class B200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"b200hnsw error {code}: {msg}")
        self.code = code


# every symbol include/b200hnsw.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "b200hnsw_abi_version": (C.c_int, []),
    "b200hnsw_last_error": (C.c_char_p, []),
    "b200hnsw_create": (C.c_int, [C.POINTER(Config), C.POINTER(C.c_void_p)]),
    "b200hnsw_destroy": (None, [C.c_void_p]),
    "b200hnsw_reserve": (C.c_int, [C.c_void_p, C.c_int64]),
    "b200hnsw_len": (C.c_int64, [C.c_void_p]),
    "b200hnsw_layers": (C.c_int32, [C.c_void_p]),
    "b200hnsw_layer_size": (C.c_int64, [C.c_void_p, C.c_int32]),
    "b200hnsw_entry": (C.c_int32, [C.c_void_p, C.c_int32]),
    "b200hnsw_set_entry": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    "b200hnsw_load": (C.c_int, [C.c_void_p, C.c_int64, f32p, i64p, i32p, C.c_int32, i32p, i64p, i32p, i64p, i32p]),
    "b200hnsw_get_rows": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, f32p, i64p, i32p]),
    "b200hnsw_set_keys": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, i64p]),
    "b200hnsw_get_neighbors": (C.c_int, [C.c_void_p, C.c_int32, C.c_int64, C.c_int64, i32p, u8p]),
    "b200hnsw_search_batch": (C.c_int, [C.c_void_p, f32p, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, i32p,
                                        i32p, f32p, i32p, C.POINTER(Stats)]),
    "b200hnsw_search_batch_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                               C.c_int32, i32p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                               C.c_void_p]),
    "b200hnsw_add_batch": (C.c_int, [C.c_void_p, f32p, i64p, i32p, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                     C.POINTER(Stats)]),
    "b200hnsw_add_batch_waves": (C.c_int, [C.c_void_p, f32p, i64p, i32p, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                           C.c_int64, C.POINTER(Stats)]),
    "b200hnsw_delete_batch": (C.c_int, [C.c_void_p, i32p, C.c_int64, C.c_int32, u8p, C.POINTER(Stats)]),
    "b200hnsw_is_deleted": (C.c_int32, [C.c_void_p, C.c_int32]),
    "b200hnsw_rows": (C.c_int64, [C.c_void_p]),
    "b200hnsw_set_build_mode": (C.c_int, [C.c_void_p, C.c_int32]),
    "b200hnsw_get_build_mode": (C.c_int32, [C.c_void_p]),
    "b200hnsw_set_add_window": (C.c_int, [C.c_void_p, C.c_int32]),
    "b200hnsw_get_add_window_stats": (C.c_int, [C.c_void_p, i64p, i64p, i64p]),
    "b200hnsw_get_visited_capacity": (C.c_int32, [C.c_void_p]),
    "b200hnsw_set_visited_capacity": (C.c_int, [C.c_void_p, C.c_int32]),
    "b200hnsw_merge_topk_device": (C.c_int, [C.c_int32, C.c_int32, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200hnsw_search_batch_segmented_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32,
                                                         C.c_int32, C.c_int32, i32p, C.c_void_p, C.c_int64, C.c_int64,
                                                         C.c_void_p, C.c_void_p]),
    "b200hnsw_merge_shards_device": (C.c_int, [C.c_int32, C.c_int32, C.c_int64, C.c_int32, C.c_void_p, C.c_int64,
                                               C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200hnsw_search_batch_scatter_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32,
                                                       C.c_int32, C.c_int32, i32p, C.POINTER(C.c_void_p), C.c_int32,
                                                       C.c_int64, C.c_void_p, C.c_void_p]),
    "b200hnsw_peer_alloc": (C.c_int, [C.c_int32, C.c_int64, C.POINTER(C.c_void_p), u8p]),
    "b200hnsw_peer_open": (C.c_int, [C.c_int32, u8p, C.POINTER(C.c_void_p)]),
    "b200hnsw_peer_close": (C.c_int, [C.c_void_p]),
    "b200hnsw_peer_free": (C.c_int, [C.c_void_p]),
    "b200hnsw_peer_signal_device": (C.c_int, [C.c_int32, C.POINTER(C.c_void_p), C.c_int32, C.c_int32, C.c_void_p]),
    "b200hnsw_peer_wait_device": (C.c_int, [C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "b200exact_create": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.POINTER(C.c_void_p)]),
    "b200exact_destroy": (None, [C.c_void_p]),
    "b200exact_len": (C.c_int64, [C.c_void_p]),
    "b200exact_add_batch": (C.c_int, [C.c_void_p, f32p, C.c_int64, i64p]),
    "b200exact_remove": (C.c_int, [C.c_void_p, i64p, C.c_int64]),
    "b200exact_search_batch": (C.c_int, [C.c_void_p, f32p, C.c_int64, C.c_int32, i64p, f32p, i32p, C.POINTER(Stats)]),
    "b200exact_search_batch_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p,
                                                C.c_void_p, C.c_void_p]),
    "b200exact_search_batch_noalloc_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p,
                                                        C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200hnsw_distance_batch": (C.c_int, [C.c_int32, C.c_int32, f32p, f32p, C.c_int64, C.c_int32, f32p]),
}

_LIB = None


def lib():
    """Load the CUDA library; raise if it is not built (no fallback)."""
    global _LIB
    if _LIB is None:
        if not LIB_PATH.exists():
            raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              f"(nvcc, sm_100a). hnsw_b200 has no CPU fallback.")
        L = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)  # AttributeError if the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


def check(rc):
    if rc != 0:
        raise B200Error(rc, lib().b200hnsw_last_error().decode("utf-8", "replace"))
