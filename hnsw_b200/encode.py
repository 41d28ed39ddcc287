"""encode.go:16-327 -- the v1 wire format (Export / Import / SavedGraph) for int keys, between byte streams and the
device-resident graph.

Format (encode.go:128-174): varint(version=1), varint(M), float64 LE (Ml), varint(EfSearch), string(distance name:
varint length + bytes), varint(nLayers); per layer varint(nNodes); per node: key (int -> binary.PutVarint, i.e.
zig-zag LEB128), vector (varint length + float32 LE), varint(nNeighbors), neighbour keys.  Go writes nodes and
neighbours in map order (random); here they are written in ascending key order (the order the oracle uses too, so the
two produce identical bytes).  Import accepts any order.
"""
from __future__ import annotations

import io
import os
import struct
import tempfile
from typing import BinaryIO

import numpy as np

from .distance import distanceFuncs, distanceFuncToName

encodingVersion = 1  # encode.go:128


def put_varint(v: int) -> bytes:
    """encoding/binary.PutVarint: zig-zag then base-128 little endian (encode_test.go:11-33: 1337 -> 2 bytes)."""
    ux = ((v << 1) ^ (v >> 63)) & 0xFFFFFFFFFFFFFFFF
    out = bytearray()
    while ux >= 0x80:
        out.append((ux & 0x7F) | 0x80)
        ux >>= 7
    out.append(ux)
    return bytes(out)


def read_varint(r: BinaryIO) -> int:
    """encoding/binary.ReadVarint (reads byte by byte, never past the value: encode_test.go:25-32)."""
    ux, shift = 0, 0
    for i in range(10):
        b = r.read(1)
        if len(b) != 1:
            raise EOFError("unexpected EOF")
        byte = b[0]
        if byte < 0x80:
            if i == 9 and byte > 1:
                raise OverflowError("varint overflows a 64-bit integer")
            ux |= byte << shift
            x = ux >> 1
            return ~x if ux & 1 else x
        ux |= (byte & 0x7F) << shift
        shift += 7
    raise OverflowError("varint overflows a 64-bit integer")


def put_string(s: str) -> bytes:
    """encode.go:77-88 (encode_test.go:35-50: "hello" -> 6 bytes)."""
    b = s.encode("utf-8")
    return put_varint(len(b)) + b


def read_string(r: BinaryIO) -> str:
    n = read_varint(r)
    b = r.read(n)
    if len(b) != n:
        raise EOFError("unexpected EOF")
    return b.decode("utf-8")


def _read_exact(r, n):
    b = r.read(n)
    if len(b) != n:
        raise EOFError("unexpected EOF")
    return b


# ---- keys on the wire (encode.go:66-126): K = int -> binary.PutVarint (zig-zag); K = string -> varint length + bytes;
#      any other fixed-size K -> binary.Write little endian.  The Python host has int (Go int), float (Go float64) and
#      str / bytes (Go string) keys.
def _key_kind(g, key_type=None):
    if key_type is not None:
        return {int: "int", float: "float", str: "str", bytes: "bytes"}[key_type]
    kind = getattr(getattr(g, "_codec", None), "kind", None)
    if kind in (None, "int"):
        return "int"
    if kind == "float":
        return "float"
    sample = next((k for k in g._keys), "")
    return "bytes" if isinstance(sample, bytes) else "str"


def put_key(kind: str, key) -> bytes:
    if kind == "int":
        return put_varint(int(key))
    if kind == "float":
        return struct.pack("<d", float(key))
    b = key if isinstance(key, bytes) else str(key).encode("utf-8")
    return put_varint(len(b)) + b


def read_key(kind: str, r: BinaryIO):
    if kind == "int":
        return read_varint(r)
    if kind == "float":
        return struct.unpack("<d", _read_exact(r, 8))[0]
    n = read_varint(r)
    b = _read_exact(r, n)
    return b if kind == "bytes" else b.decode("utf-8")


def export_graph(g, w: BinaryIO) -> None:
    """Graph.Export (encode.go:133-174): reads the lists back from the device.  Like the reference it writes what
    layer.nodes holds: after a Delete that is the live nodes only, whose neighbour lists can still name deleted keys
    (isolate leaves one-way links and replenish re-links isolated nodes, graph.go:172-235); Import resolves those to nil
    neighbours, which Search skips (encode.go:253-257, graph.go:140-143)."""
    name, ok = distanceFuncToName(g.Distance)
    if not ok:
        raise ValueError(f"distance function {g.Distance!r} must be registered with RegisterDistanceFunc")
    w.write(put_varint(encodingVersion) + put_varint(g.M) + struct.pack("<d", float(g.Ml)) +
            put_varint(g.EfSearch) + put_string(name))
    if g._h is None or not g._keys:
        w.write(put_varint(0))
        return
    kind = _key_kind(g)
    vec, codes, levels, layer_lists, _ = g.dump_flat()
    w.write(put_varint(len(layer_lists)))
    dim_prefix = put_varint(vec.shape[1])
    user_keys = g._keys                      # row -> the caller's key (the device orders rows by `codes`)
    key_bytes = [put_key(kind, k) for k in user_keys]
    live = np.zeros(len(user_keys), dtype=bool)
    live[np.fromiter(g._row_of.values(), dtype=np.int64, count=len(g._row_of))] = True
    for rows, ptr, nbr in layer_lists:
        member = live[rows]
        w.write(put_varint(int(member.sum())))
        order = np.argsort(codes[rows], kind="stable")
        for i in order:
            if not member[i]:
                continue  # delete(layer.nodes, key): the node is no longer in the map Export iterates
            r = int(rows[i])
            nb = nbr[ptr[i]:ptr[i + 1]]  # already in ascending key order on the device
            w.write(key_bytes[r] + dim_prefix + vec[r].astype("<f4").tobytes() + put_varint(len(nb)) +
                    b"".join(key_bytes[int(t)] for t in nb))


def import_graph(g, r: BinaryIO, key_type=None) -> None:
    """Graph.Import (encode.go:180-262): parameters come from the stream, the content replaces the graph's.  key_type
    (int, float, str, bytes) plays Go's type parameter K; default: the graph's current key type, else int."""
    kind = _key_kind(g, key_type)
    version = read_varint(r)
    m = read_varint(r)
    ml = struct.unpack("<d", _read_exact(r, 8))[0]
    ef = read_varint(r)
    dist = read_string(r)
    if dist not in distanceFuncs:
        raise ValueError(f'unknown distance function "{dist}"')  # encode.go:194-197
    if version != encodingVersion:
        raise ValueError(f"incompatible encoding version: {version}")  # encode.go:202-204
    g.M, g.Ml, g.EfSearch, g.Distance = m, ml, ef, distanceFuncs[dist]
    n_layers = read_varint(r)
    row_of, keys, vecs, levels = {}, [], [], []
    layers = []
    for l in range(n_layers):
        n_nodes = read_varint(r)
        members = []
        for _ in range(n_nodes):
            key = read_key(kind, r)
            ln = read_varint(r)
            v = np.frombuffer(_read_exact(r, 4 * ln), dtype="<f4")
            nn = read_varint(r)
            nbk = [read_key(kind, r) for _ in range(nn)]
            if l == 0:
                if key in row_of:
                    continue  # a Go map keeps one node per key
                row_of[key] = len(keys)
                keys.append(key)
                vecs.append(v)
                levels.append(0)
            row = row_of.get(key)
            if row is None:
                continue  # a node of an upper layer that layer 0 does not hold cannot be reached by the device layout
            levels[row] = max(levels[row], l)
            members.append((row, nbk))
        layers.append(members)
    if not keys:
        g.close()
        return
    layer_lists = []
    for l, members in enumerate(layers):
        present = {row for row, _ in members}
        rows, ptr, nbr = [], [0], []
        for row, nbk in members:
            rows.append(row)
            for kx in nbk:  # neighbours absent from the layer resolve to nil in Go (encode.go:253-257) and are skipped
                t = row_of.get(kx)
                if t is not None and t in present and t != row:
                    nbr.append(t)
            ptr.append(len(nbr))
        layer_lists.append((np.asarray(rows, np.int32), np.asarray(ptr, np.int64), np.asarray(nbr, np.int32)))
    # the device layout holds a node of level L in every layer 0..L
    for l, (rows, _, _) in enumerate(layer_lists):
        have = set(int(x) for x in rows)
        for row, lv in enumerate(levels):
            if lv >= l and row not in have:
                raise ValueError(f"key {keys[row]} is in layer {lv} but not in layer {l}: not a graph Export wrote")
    g.close()
    g.load_flat(np.stack(vecs).astype(np.float32), keys, np.asarray(levels, np.int32), layer_lists)


class SavedGraph:
    """encode.go:268-327: a graph that persists to a file upon Save()."""

    def __init__(self, graph, path: str):
        self.Graph, self.Path = graph, path

    def __getattr__(self, name):
        return getattr(self.Graph, name)

    def Save(self) -> None:
        """Temporary file in the same directory + atomic rename (renameio.TempFile / CloseAtomicallyReplace)."""
        d = os.path.dirname(os.path.abspath(self.Path))
        fd, tmp = tempfile.mkstemp(prefix=".tmp-", dir=d)
        try:
            with os.fdopen(fd, "wb") as f:
                w = io.BufferedWriter(f)
                export_graph(self.Graph, w)
                w.flush()
                os.fsync(f.fileno())
            os.replace(tmp, self.Path)
        except BaseException:
            try:
                os.unlink(tmp)
            except OSError:
                pass
            raise


def LoadSavedGraph(path: str, **kw) -> SavedGraph:
    """encode.go:283-304: a missing or empty file gives the equivalent of NewGraph."""
    from .graph import NewGraph
    g = NewGraph(**kw)
    if os.path.exists(path) and os.path.getsize(path) > 0:
        with open(path, "rb") as f:
            import_graph(g, io.BufferedReader(f))
    else:
        open(path, "ab").close()  # os.O_CREATE
    return SavedGraph(g, path)
