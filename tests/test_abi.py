"""The C-ABI library loads on a machine without a GPU and exports every function include/b200hnsw.h declares (no
compute calls here); the ctypes table of the host mirror covers the same set."""
import ctypes as C
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def _declared():
    text = (ROOT / "include" / "b200hnsw.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)  # comments name functions too
    names = re.findall(r"\b(b200(?:hnsw|exact)_\w+)\s*\(", text)
    return sorted(set(names))


def test_header_declares_the_documented_entry_points():
    names = _declared()
    for must in ("b200hnsw_create", "b200hnsw_search_batch", "b200hnsw_search_batch_device", "b200hnsw_add_batch",
                 "b200hnsw_add_batch_waves", "b200hnsw_delete_batch", "b200hnsw_load", "b200hnsw_merge_shards_device",
                 "b200hnsw_search_batch_scatter_device", "b200hnsw_peer_alloc", "b200exact_search_batch"):
        assert must in names
    assert len(names) >= 38


def test_library_exports_every_declared_symbol():
    from hnsw_b200 import _lib
    L = C.CDLL(str(_lib.LIB_PATH))
    missing = [n for n in _declared() if not hasattr(L, n)]
    assert not missing, f"declared in include/b200hnsw.h but not exported: {missing}"


def test_ctypes_table_matches_the_header():
    from hnsw_b200 import _lib
    assert sorted(_lib.SYMBOLS) == _declared()
    L = _lib.lib()  # resolves and types every symbol
    text = (ROOT / "include" / "b200hnsw.h").read_text()
    ver = int(re.search(r"#define\s+B200HNSW_ABI_VERSION\s+(\d+)", text).group(1))
    assert L.b200hnsw_abi_version() == ver
    assert L.b200hnsw_last_error() == b""
    assert int(re.search(r"#define\s+B200HNSW_MAX_PEERS\s+(\d+)", text).group(1)) == 16


def test_compute_call_without_a_device_fails_loudly():
    import torch
    if torch.cuda.is_available():
        return  # on the GPU box the GPU tests cover the calls
    from hnsw_b200 import _lib
    L = _lib.lib()
    cfg = _lib.Config(dim=8, max_m=4, device=0, reserved=0, capacity=0)
    h = C.c_void_p()
    rc = L.b200hnsw_create(C.byref(cfg), C.byref(h))
    assert rc == -3 and h.value is None  # B200HNSW_ERR_CUDA: no CPU fallback
    assert b"" != L.b200hnsw_last_error()


def test_add_window_api_without_a_device():
    """b200hnsw_set_add_window / get_add_window_stats (the schedule of the reference-order Add, include/b200hnsw.h): argument
    checks and the host mirror's behaviour before a handle exists need no GPU."""
    import hnsw_b200
    from hnsw_b200 import _lib
    L = _lib.lib()
    assert L.b200hnsw_set_add_window(None, 8) == _lib.ERR_BAD_ARG
    assert b"set_add_window" in L.b200hnsw_last_error()
    assert L.b200hnsw_get_add_window_stats(None, None, None, None) == _lib.ERR_BAD_ARG
    g = hnsw_b200.Graph()
    g.set_add_window(8)                      # remembered, applied when the first Add creates the handle
    assert g.add_window_stats() == (0, 0, 0)
