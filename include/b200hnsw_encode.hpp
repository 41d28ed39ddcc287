// b200hnsw_encode.hpp -- encode.go:16-262 for the C++ host (include/b200hnsw_host.hpp): the v1 wire format between byte
// streams and the device-resident graph.
//
// Format (encode.go:128-174): varint(version=1), varint(M), float64 LE (Ml), varint(EfSearch), string(distance name:
// varint length + bytes), varint(nLayers); per layer varint(nNodes); per node: key (int -> binary.PutVarint, zig-zag
// LEB128), vector (varint length + float32 LE), varint(nNeighbors), neighbour keys.  Go writes nodes and neighbours in
// map order (random); here they are written in ascending key order, as the oracle does, so the two produce identical
// bytes.  Import accepts any order.
#pragma once
#include <algorithm>
#include <istream>
#include <ostream>

#include "b200hnsw_host.hpp"

namespace hnsw {

constexpr int encodingVersion = 1;  // encode.go:128

namespace enc {
// encoding/binary.PutVarint (encode_test.go:11-33: 1337 -> 2 bytes)
inline void putVarint(std::string &out, int64_t v) {
  uint64_t ux = ((uint64_t)v << 1) ^ (uint64_t)(v >> 63);
  while (ux >= 0x80) {
    out.push_back((char)((ux & 0x7f) | 0x80));
    ux >>= 7;
  }
  out.push_back((char)ux);
}
// encoding/binary.ReadVarint: byte by byte, never past the value (encode_test.go:25-32)
inline int64_t readVarint(std::istream &r) {
  uint64_t ux = 0;
  int shift = 0;
  for (int i = 0; i < 10; i++) {
    int c = r.get();
    if (c == std::char_traits<char>::eof()) throw Error("unexpected EOF");
    uint8_t b = (uint8_t)c;
    if (b < 0x80) {
      if (i == 9 && b > 1) throw Error("binary: varint overflows a 64-bit integer");
      ux |= (uint64_t)b << shift;
      int64_t x = (int64_t)(ux >> 1);
      return (ux & 1) ? ~x : x;
    }
    ux |= (uint64_t)(b & 0x7f) << shift;
    shift += 7;
  }
  throw Error("binary: varint overflows a 64-bit integer");
}
// encode.go:77-88 (encode_test.go:35-50: "hello" -> 6 bytes)
inline void putString(std::string &out, const std::string &s) {
  putVarint(out, (int64_t)s.size());
  out += s;
}
inline void readExact(std::istream &r, char *dst, size_t n) {
  r.read(dst, (std::streamsize)n);
  if ((size_t)r.gcount() != n) throw Error("unexpected EOF");
}
inline std::string readString(std::istream &r) {
  int64_t n = readVarint(r);
  if (n < 0) throw Error("negative string length");
  std::string s((size_t)n, '\0');
  readExact(r, s.data(), (size_t)n);
  return s;
}
inline void putF32LE(std::string &out, float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  for (int i = 0; i < 4; i++) out.push_back((char)((u >> (8 * i)) & 0xff));
}
inline void putF64LE(std::string &out, double f) {
  uint64_t u;
  std::memcpy(&u, &f, 8);
  for (int i = 0; i < 8; i++) out.push_back((char)((u >> (8 * i)) & 0xff));
}
}  // namespace enc

// Graph.Export (encode.go:133-174): the lists are read back from the device
template <class K>
void Export(const Graph<K> &g, std::ostream &w) {
  auto name = distanceFuncToName(g.Distance);
  if (!name.second) throw Error("distance function must be registered with RegisterDistanceFunc");
  std::string out;
  enc::putVarint(out, encodingVersion);
  enc::putVarint(out, g.M);
  enc::putF64LE(out, g.Ml);
  enc::putVarint(out, g.EfSearch);
  enc::putString(out, name.first);
  const b200hnsw_t *h = g.Handle();
  const int64_t rows = h ? b200hnsw_rows(h) : 0;
  // After a Delete the reference writes what layer.nodes holds: the live nodes only, whose neighbour lists can still
  // name deleted keys (isolate leaves one-way links, replenish re-links isolated nodes, graph.go:172-235); Import
  // resolves those to nil neighbours, which Search skips (encode.go:253-257, graph.go:140-143).
  if (rows == 0) {
    enc::putVarint(out, 0);
    w.write(out.data(), (std::streamsize)out.size());
    return;
  }
  const int dim = g.Dims(), nl = b200hnsw_layers(h);
  std::vector<float> vec((size_t)rows * dim);
  std::vector<int64_t> keys((size_t)rows);
  std::vector<int32_t> levels((size_t)rows);
  detail::check(b200hnsw_get_rows(h, 0, rows, vec.data(), keys.data(), levels.data()));
  const int stride = g.ListStride();  // neighbour slots per list on the device (max_m fixed at the first Add)
  std::vector<int32_t> nbr((size_t)rows * stride);
  std::vector<uint8_t> member((size_t)rows);
  enc::putVarint(out, nl);
  std::vector<int32_t> order;
  for (int l = 0; l < nl; l++) {
    detail::check(b200hnsw_get_neighbors(h, l, 0, rows, nbr.data(), member.data()));
    order.clear();
    for (int64_t r = 0; r < rows; r++)
      if (member[(size_t)r] && b200hnsw_is_deleted(h, (int32_t)r) == 0) order.push_back((int32_t)r);
    std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return keys[a] < keys[b]; });
    enc::putVarint(out, (int64_t)order.size());
    for (int32_t r : order) {
      enc::putVarint(out, keys[(size_t)r]);
      enc::putVarint(out, dim);
      for (int i = 0; i < dim; i++) enc::putF32LE(out, vec[(size_t)r * dim + i]);
      const int32_t *list = &nbr[(size_t)r * stride];
      int deg = 0;
      while (deg < stride && list[deg] >= 0) deg++;
      enc::putVarint(out, deg);
      for (int i = 0; i < deg; i++) enc::putVarint(out, keys[(size_t)list[i]]);  // ascending key order on the device
    }
    w.write(out.data(), (std::streamsize)out.size());
    out.clear();
  }
}

// Graph.Import (encode.go:180-262): parameters come from the stream, the content replaces the graph's
template <class K>
void Import(Graph<K> &g, std::istream &r) {
  const int64_t version = enc::readVarint(r);
  const int64_t m = enc::readVarint(r);
  char b8[8];
  enc::readExact(r, b8, 8);
  uint64_t u = 0;
  for (int i = 0; i < 8; i++) u |= (uint64_t)(uint8_t)b8[i] << (8 * i);
  double ml;
  std::memcpy(&ml, &u, 8);
  const int64_t ef = enc::readVarint(r);
  const std::string dist = enc::readString(r);
  auto it = distanceFuncs().find(dist);
  if (it == distanceFuncs().end()) throw Error("unknown distance function \"" + dist + "\"");  // encode.go:194-197
  if (version != encodingVersion)
    throw Error(detail::fmt("incompatible encoding version: %lld", (long long)version));     // encode.go:202-204
  g.M = (int)m, g.Ml = ml, g.EfSearch = (int)ef, g.Distance = it->second;
  const int64_t nLayers = enc::readVarint(r);
  std::unordered_map<int64_t, int32_t> rowOf;
  std::vector<int64_t> keys;
  std::vector<Vector> vecs;
  std::vector<int32_t> levels;
  struct Member {
    int32_t row;
    std::vector<int64_t> nbk;
  };
  std::vector<std::vector<Member>> layers((size_t)std::max<int64_t>(nLayers, 0));
  for (int64_t l = 0; l < nLayers; l++) {
    const int64_t nNodes = enc::readVarint(r);
    for (int64_t i = 0; i < nNodes; i++) {
      const int64_t key = enc::readVarint(r);
      const int64_t ln = enc::readVarint(r);
      Vector v((size_t)ln);
      for (int64_t j = 0; j < ln; j++) {
        char b4[4];
        enc::readExact(r, b4, 4);
        uint32_t w = 0;
        for (int t = 0; t < 4; t++) w |= (uint32_t)(uint8_t)b4[t] << (8 * t);
        std::memcpy(&v[(size_t)j], &w, 4);
      }
      const int64_t nn = enc::readVarint(r);
      std::vector<int64_t> nbk((size_t)nn);
      for (auto &x : nbk) x = enc::readVarint(r);
      if (l == 0) {
        if (rowOf.count(key)) continue;  // a Go map keeps one node per key
        rowOf[key] = (int32_t)keys.size();
        keys.push_back(key);
        vecs.push_back(std::move(v));
        levels.push_back(0);
      }
      auto f = rowOf.find(key);
      if (f == rowOf.end()) continue;  // an upper-layer node that layer 0 does not hold: unreachable in the device layout
      levels[(size_t)f->second] = std::max(levels[(size_t)f->second], (int32_t)l);
      layers[(size_t)l].push_back(Member{f->second, std::move(nbk)});
    }
  }
  g.LoadFlatBegin();
  if (keys.empty()) return;
  const int64_t n = (int64_t)keys.size();
  const int dim = (int)vecs[0].size();
  std::vector<float> flat((size_t)n * dim);
  for (int64_t i = 0; i < n; i++) {
    if ((int)vecs[(size_t)i].size() != dim) throw Error("vectors of different lengths in one stream");
    std::memcpy(&flat[(size_t)i * dim], vecs[(size_t)i].data(), sizeof(float) * dim);
  }
  std::vector<int64_t> sizes;
  std::vector<int32_t> lrows, lnbr, entries((size_t)nLayers, -1);
  std::vector<int64_t> lptr;
  for (int64_t l = 0; l < nLayers; l++) {
    auto &mem = layers[(size_t)l];
    std::unordered_set<int32_t> present;
    for (auto &mm : mem) present.insert(mm.row);
    for (int64_t row = 0; row < n; row++)  // the device layout holds a node of level L in every layer 0..L
      if (levels[(size_t)row] >= l && !present.count((int32_t)row))
        throw Error(detail::fmt("key %lld is in layer %d but not in layer %lld: not a graph Export wrote",
                                (long long)keys[(size_t)row], levels[(size_t)row], (long long)l));
    sizes.push_back((int64_t)mem.size());
    lptr.push_back(0);
    int64_t cnt = 0;
    for (auto &mm : mem) {
      lrows.push_back(mm.row);
      for (int64_t kx : mm.nbk) {  // neighbours absent from the layer resolve to nil in Go (encode.go:253-257): skipped
        auto t = rowOf.find(kx);
        if (t != rowOf.end() && present.count(t->second) && t->second != mm.row) {
          lnbr.push_back(t->second);
          cnt++;
        }
      }
      lptr.push_back(cnt);
    }
  }
  g.LoadFlat(n, dim, flat, keys, levels, (int)nLayers, entries, sizes, lrows, lptr, lnbr, vecs);
}

}  // namespace hnsw
