"""tests/golden/reference_kats.json: the known-answer vectors of the reference's own tests, checked against
the oracle on the CPU and against the device arithmetic / kernels through the C ABI on the GPU."""
import json
import math
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
KATS = json.loads((ROOT / "tests" / "golden" / "reference_kats.json").read_text())


@pytest.fixture(scope="module")
def oracle():
    from oracle import binding
    binding.build()
    return binding


def test_golden_distances_oracle(oracle):
    for flavor in (oracle.CANONICAL, oracle.SEQUENTIAL, oracle.SIMD):
        for c in KATS["euclidean"]:
            got = oracle.distance(oracle.EUCLIDEAN, c["a"], c["b"], flavor)
            assert np.float32(got) == np.float32(c["expect_f32"]), c["ref"]
        for c in KATS["cosine"]:
            assert abs(oracle.distance(oracle.COSINE, c["a"], c["b"], flavor) - c["expect"]) <= c["delta"], c["ref"]


def test_golden_max_level_and_wire_bytes(oracle):
    for c in KATS["max_level"]:
        assert oracle.max_level(c["ml"], c["n"]) == c["expect"], c["ref"]
    from hnsw_b200 import encode
    for c in KATS["varint"]:
        assert oracle.put_varint(c["value"]).hex() == c["bytes_hex"] and len(oracle.put_varint(c["value"])) == c["n_bytes"]
        assert encode.put_varint(c["value"]).hex() == c["bytes_hex"], c["ref"]
        v, n = oracle.get_varint(bytes.fromhex(c["bytes_hex"]) + bytes(c["trailing_untouched"]))
        assert (v, n) == (c["value"], c["n_bytes"])  # does not read past the varint
    for c in KATS["string"]:
        assert encode.put_string(c["value"]).hex() == c["bytes_hex"] and len(encode.put_string(c["value"])) == c["n_bytes"]


def test_golden_layer_search_oracle(oracle):
    for c in KATS["layer_search"]:
        g = oracle.Graph(c["dim"], 6, oracle.EUCLIDEAN)
        for nd in c["nodes"]:
            assert g.put_node(nd["map_key"], nd["value"]) == nd["map_key"]
        for a, b in c["links"]:
            g.link(0, a, b)
        rows, _ = g.layer_search(0, c["entry"], c["k"], c["ef"], c["target"])
        assert [c["nodes"][r]["key"] for r in rows] == c["expect_keys"], c["ref"]


@pytest.mark.gpu
def test_golden_distances_device():
    from hnsw_b200 import EuclideanDistance, CosineDistance  # evaluated on the device (b200hnsw_distance_batch)
    for c in KATS["euclidean"]:
        assert np.float32(EuclideanDistance(c["a"], c["b"])) == np.float32(c["expect_f32"]), c["ref"]
    for c in KATS["cosine"]:
        assert abs(float(CosineDistance(c["a"], c["b"])) - c["expect"]) <= c["delta"], c["ref"]


@pytest.mark.gpu
def test_golden_layer_search_device():
    # graph_test.go:27-74 through the C ABI: the six nodes loaded as a one-layer graph, Search = layer search from
    # the entry; the device returns rows [4, 3] -> Keys [5, 3]
    import hnsw_b200
    for c in KATS["layer_search"]:
        n = len(c["nodes"])
        vec = np.array([nd["value"] for nd in c["nodes"]], dtype=np.float32)
        keys = np.array([nd["map_key"] for nd in c["nodes"]], dtype=np.int64)
        adj = {i: [] for i in range(n)}
        for a, b in c["links"]:
            adj[a].append(b)
        ptr = np.zeros(n + 1, dtype=np.int64)
        nbr = []
        for i in range(n):
            nbr += sorted(adj[i])
            ptr[i + 1] = len(nbr)
        g = hnsw_b200.Graph(M=6, Ml=0.25, EfSearch=c["ef"], Distance=hnsw_b200.EuclideanDistance)
        g.load_flat(vec, keys, np.zeros(n, dtype=np.int32), [(np.arange(n, dtype=np.int32), ptr, np.array(nbr, np.int32))],
                    entries=[c["entry"]])
        rows, dists, cnt = g.search_rows(np.array([c["target"]], dtype=np.float32), c["k"], c["ef"])
        assert [c["nodes"][r]["key"] for r in rows[0][:cnt[0]]] == c["expect_keys"], c["ref"]
