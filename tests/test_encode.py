"""encode.go v1 wire format: byte-level KATs on CPU (encode_test.go:11-50) and, on the GPU, Export/Import of the
device-resident graph against the oracle's restatement (encode_test.go:120-181)."""
import io

import numpy as np
import pytest

from util import build_oracle_graph, flatten_oracle


def test_varint_kat():
    # encode_test.go:11-33: 1337 encodes to 2 bytes and decoding does not over-read
    from hnsw_b200.encode import put_varint, read_varint
    b = put_varint(1337)
    assert len(b) == 2
    r = io.BytesIO(b + b"\xff\xff")
    assert read_varint(r) == 1337 and r.tell() == 2
    for v in (0, 1, -1, 63, 64, -64, -65, 2 ** 31, -2 ** 31, 2 ** 62, -2 ** 63, 2 ** 63 - 1):
        assert read_varint(io.BytesIO(put_varint(v))) == v


def test_varint_matches_oracle(oracle):
    from hnsw_b200.encode import put_varint
    for v in (0, 1, -1, 1337, -1337, 300, 2 ** 40 + 17, -2 ** 40):
        assert put_varint(v) == oracle.put_varint(v)


def test_string_kat():
    # encode_test.go:35-50: "hello" -> 6 bytes
    from hnsw_b200.encode import put_string, read_string
    b = put_string("hello")
    assert len(b) == 6 and read_string(io.BytesIO(b)) == "hello"


@pytest.mark.gpu
def test_export_import_roundtrip_on_device(oracle, tmp_path):
    import hnsw_b200
    rng = np.random.default_rng(12)
    n, d = 400, 24
    X = rng.uniform(-1, 1, (n, d)).astype(np.float32)
    keys = rng.permutation(5 * n)[:n].astype(np.int64) - 700
    g, _, _ = build_oracle_graph(oracle, X, 8, 0.25, 20, 0, keys=keys)
    dg = hnsw_b200.Graph(M=8, Ml=0.25, EfSearch=20, Distance=hnsw_b200.EuclideanDistance)
    dg.load_flat(*flatten_oracle(g))
    buf = io.BytesIO()
    dg.Export(buf)
    assert buf.getvalue() == g.export(0.25, 20)  # same bytes as the oracle's Export (ascending key order)
    # encode_test.go:120-158: Import into a fresh graph -> equal Len / topography / parameters / Search output
    g2 = hnsw_b200.NewGraph()
    g2.Import(io.BytesIO(buf.getvalue()))
    assert (g2.M, g2.Ml, g2.EfSearch) == (8, 0.25, 20) and g2.Distance is hnsw_b200.EuclideanDistance
    assert g2.Len() == dg.Len() and g2.Topography() == dg.Topography()
    Q = rng.uniform(-1, 1, (32, d)).astype(np.float32)
    r1, d1, c1 = dg.search_rows(Q, 10)
    r2, d2, c2 = g2.search_rows(Q, 10)
    k1 = [[dg._keys[r] for r in row[:c]] for row, c in zip(r1, c1)]
    k2 = [[g2._keys[r] for r in row[:c]] for row, c in zip(r2, c2)]
    assert k1 == k2 and np.array_equal(d1.view(np.uint32), d2.view(np.uint32))
    out = io.BytesIO()
    g2.Export(out)
    assert out.getvalue() == buf.getvalue()
    # the oracle reads our bytes too
    g3, ml, ef = oracle.Graph.import_(buf.getvalue())
    assert g3.topography() == g.topography() and (ml, ef) == (0.25, 20)
    # encode_test.go:160-181: SavedGraph
    path = str(tmp_path / "graph.bin")
    sg = hnsw_b200.LoadSavedGraph(path)
    assert sg.Len() == 0
    for i in range(50):
        sg.Add(hnsw_b200.Node(i, X[i]))
    sg.Save()
    sg2 = hnsw_b200.LoadSavedGraph(path)
    assert sg2.Len() == 50
    a = [nd.Key for nd in sg.Search(X[3], 5)]
    b = [nd.Key for nd in sg2.Search(X[3], 5)]
    assert a == b and len(a) == 5
    with pytest.raises(ValueError, match="incompatible encoding version"):
        hnsw_b200.NewGraph().Import(io.BytesIO(b"\x04" + buf.getvalue()[1:]))
