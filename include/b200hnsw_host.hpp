// b200hnsw_host.hpp -- the reference's Go API surface (graph.go, distance.go) as a header-only C++17 host over the C ABI
// of include/b200hnsw.h.  The reference's host is Go; this image has no Go toolchain, so this is the compiled-language
// host a maintainer can read next to graph.go: same names, same argument meaning, the same error texts (thrown as
// hnsw::Error instead of returned), the same public mutable fields.  Everything data-parallel is a C-ABI call; what stays
// here is what would stay in Go under cgo: validation, the key <-> row map, the host copy of the vectors
// (results alias it, graph.go:592), the level RNG (graph.go:388-417) and the test hacks of graph.go:563-569, 595-619.
// There is no CPU fallback: without the library and a CUDA device every compute call throws.
//
//   hnsw::Graph<int> g = hnsw::NewGraph<int>();           // graph.go:340-349
//   g.Add(hnsw::MakeNode(1, {1, 1, 1}), ...);             // graph.go:437-531
//   auto nodes = g.Search({0.5f, 0.5f, 0.5f}, 1);         // graph.go:534-625
#pragma once
#include <cmath>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <initializer_list>
#include <map>
#include <memory>
#include <mutex>
#include <optional>
#include <random>
#include <shared_mutex>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <unordered_map>
#include <unordered_set>
#include <utility>
#include <vector>

#include "b200hnsw.h"

namespace hnsw {

// a Go `error` with the reference's text
struct Error : std::runtime_error {
  using std::runtime_error::runtime_error;
};

namespace detail {
inline std::string fmt(const char *f, ...) __attribute__((format(printf, 1, 2)));
inline std::string fmt(const char *f, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, f);
  vsnprintf(buf, sizeof buf, f, ap);
  va_end(ap);
  return buf;
}
inline void check(int rc) {
  if (rc != 0) throw Error(std::string("b200hnsw: ") + b200hnsw_last_error());
}
}  // namespace detail

using Vector = std::vector<float>;  // graph.go:17

// graph.go:20-27
template <class K>
struct Node {
  K Key;
  Vector Value;
};
template <class K>
Node<K> MakeNode(K key, Vector vec) {
  return Node<K>{key, std::move(vec)};
}

// ---- distance.go:11-46 ------------------------------------------------------------------------------------------------
// The two registered functions select the device metric; called directly they evaluate ONE pair on the device
// (b200hnsw_distance_batch), so even the distance known-answer tests run the kernels' arithmetic.
using DistanceFunc = float (*)(const Vector &, const Vector &);

namespace detail {
inline float device_distance(int metric, const Vector &a, const Vector &b) {
  if (a.size() != b.size())
    throw Error(fmt("embedding dimension mismatch: %zu != %zu", a.size(), b.size()));
  float out = 0.f;
  check(b200hnsw_distance_batch(0, metric, a.data(), b.data(), 1, (int32_t)a.size(), &out));
  return out;
}
}  // namespace detail

inline float CosineDistance(const Vector &a, const Vector &b) {  // distance.go:15-17
  return detail::device_distance(B200HNSW_COSINE, a, b);
}
inline float EuclideanDistance(const Vector &a, const Vector &b) {  // distance.go:20-23
  return detail::device_distance(B200HNSW_EUCLIDEAN, a, b);
}

// distance.go:25-28
inline std::map<std::string, DistanceFunc> &distanceFuncs() {
  static std::map<std::string, DistanceFunc> m{{"euclidean", EuclideanDistance}, {"cosine", CosineDistance}};
  return m;
}
// distance.go:30-39: reverse lookup by function identity
inline std::pair<std::string, bool> distanceFuncToName(DistanceFunc fn) {
  for (auto &kv : distanceFuncs())
    if (kv.second == fn) return {kv.first, true};
  return {"", false};
}
// distance.go:44-46
inline void RegisterDistanceFunc(const std::string &name, DistanceFunc fn) { distanceFuncs()[name] = fn; }

namespace detail {
// only the two built-in metrics have kernels; anything else is an error (north_star: no CPU fallback)
inline int metric_of(DistanceFunc fn) {
  if (fn == EuclideanDistance) return B200HNSW_EUCLIDEAN;
  if (fn == CosineDistance) return B200HNSW_COSINE;
  throw Error("distance function has no device kernel (only EuclideanDistance and CosineDistance); no CPU fallback");
}
}  // namespace detail

// graph.go:370-385
inline int maxLevel(double ml, int numNodes) {
  if (ml == 0) throw Error("ml must be greater than 0");
  if (numNodes == 0) return 1;
  double l = std::log((double)numNodes);
  l /= std::log(1 / ml);
  return (int)std::round(l) + 1;  // Go math.Round: half away from zero, as std::round
}

// *rand.Rand stand-in: anything with Float64().  Go's own math/rand stream is not reproducible outside Go.
struct Rand {
  std::function<double()> Float64;
  explicit Rand(uint64_t seed) {
    auto eng = std::make_shared<std::mt19937_64>(seed);
    Float64 = [eng]() { return std::uniform_real_distribution<double>(0.0, 1.0)(*eng); };
  }
  explicit Rand(std::function<double()> f) : Float64(std::move(f)) {}
};

// ---- graph.go:305-332: Graph[K].  Public mutable fields Distance, Rng, M, Ml, EfSearch ---------------------------------
template <class K>
class Graph {
  static_assert(std::is_integral<K>::value,
                "the device orders neighbour lists by an int64 image of K (cmp.Ordered); integer keys map to themselves, "
                "other ordered key types need an order-preserving host map (INTEGRATION.md)");

 public:
  DistanceFunc Distance = CosineDistance;
  std::optional<Rand> Rng;
  int M = 16;
  double Ml = 0.25;
  int EfSearch = 20;
  int Device = 0;  // CUDA device ordinal of the mirror (no counterpart in the reference)
  bool ReplaceRaises = true;  // graph.go:1035-1037: BatchAdd of an existing key replaces it AND returns "node not added"

  Graph() = default;
  Graph(const Graph &) = delete;
  Graph &operator=(const Graph &) = delete;
  Graph(Graph &&o) noexcept { *this = std::move(o); }
  Graph &operator=(Graph &&o) noexcept {
    if (this != &o) {
      close();
      Distance = o.Distance, Rng = std::move(o.Rng), M = o.M, Ml = o.Ml, EfSearch = o.EfSearch, Device = o.Device;
      h_ = o.h_, dim_ = o.dim_, max_m_ = o.max_m_, add_window_ = o.add_window_;
      keys_ = std::move(o.keys_), row_of_ = std::move(o.row_of_), retired_ = std::move(o.retired_);
      vecs_ = std::move(o.vecs_);
      o.h_ = nullptr;
    }
    return *this;
  }
  ~Graph() { close(); }

  // graph.go:916-937
  void Validate() const {
    if (M <= 0) throw Error(detail::fmt("M must be greater than 0, got %d", M));
    if (Ml <= 0 || Ml >= 1) throw Error(detail::fmt("Ml must be between 0 and 1 (exclusive), got %f", Ml));
    if (EfSearch <= 0) throw Error(detail::fmt("EfSearch must be greater than 0, got %d", EfSearch));
    if (Distance == nullptr) throw Error("Distance function must be set");
  }

  // graph.go:829-838
  int Len() const { return h_ ? (int)b200hnsw_len(h_) : 0; }
  // graph.go:419-429
  int Dims() const { return Len() > 0 ? dim_ : 0; }
  // graph.go:898-912
  std::pair<Vector, bool> Lookup(K key) const {
    std::shared_lock<std::shared_mutex> lk(mu_);
    auto it = row_of_.find((int64_t)key);
    if (it == row_of_.end()) return {Vector{}, false};
    return {vecs_[(size_t)it->second], true};
  }
  // analyzer.go: nodes per layer
  std::vector<int> Topography() const {
    std::vector<int> t;
    if (h_)
      for (int l = 0, n = b200hnsw_layers(h_); l < n; l++) t.push_back((int)b200hnsw_layer_size(h_, l));
    return t;
  }
  const b200hnsw_t *Handle() const { return h_; }  // for parity harnesses (levels, neighbour lists)
  int ListStride() const { return max_m_; }         // neighbour slots per list on the device

  // Import (b200hnsw_encode.hpp): drop the content, then install a flattened graph (rows, CSR lists per layer)
  void LoadFlatBegin() {
    close();
    dim_ = max_m_ = 0;
    keys_.clear(), row_of_.clear(), retired_.clear(), vecs_.clear();
  }
  void LoadFlat(int64_t n, int dim, const std::vector<float> &flat, const std::vector<int64_t> &keys,
                const std::vector<int32_t> &levels, int nLayers, const std::vector<int32_t> &entries,
                const std::vector<int64_t> &layerSizes, const std::vector<int32_t> &layerRows,
                const std::vector<int64_t> &layerPtr, const std::vector<int32_t> &layerNbr,
                const std::vector<Vector> &vecs) {
    b200hnsw_config cfg{};
    cfg.dim = dim, cfg.max_m = M < B200HNSW_MAX_M ? M : B200HNSW_MAX_M, cfg.device = Device, cfg.capacity = n;
    detail::check(b200hnsw_create(&cfg, &h_));
    if (add_window_) detail::check(b200hnsw_set_add_window(h_, add_window_));
    dim_ = dim, max_m_ = cfg.max_m;
    detail::check(b200hnsw_load(h_, n, flat.data(), keys.data(), levels.data(), nLayers, entries.data(),
                                layerSizes.data(), layerRows.data(), layerPtr.data(), layerNbr.data()));
    keys_ = keys;
    vecs_ = vecs;
    for (int64_t i = 0; i < n; i++) row_of_[keys[(size_t)i]] = (int32_t)i;
  }

  // graph.go:388-417
  int randomLevel() {
    int max = 1;
    if (h_ && b200hnsw_layers(h_) > 0) {
      if (Ml == 0) throw Error("(*Graph).Ml must be greater than 0");
      max = maxLevel(Ml, Len());
    }
    for (int level = 0; level < max; level++) {
      if (!Rng) Rng.emplace((uint64_t)std::random_device{}());  // graph.go:407-409: time-seeded default
      if (Rng->Float64() > Ml) return level;
    }
    return max;
  }

  // graph.go:437-531.  The reference's Add deadlocks on a key that is already present; here, as in
  // BatchAdd below, re-using a key is refused.
  template <class... Ns>
  void Add(const Ns &...nodes) {
    BatchAdd(std::vector<Node<K>>{nodes...});
  }

  // graph.go:942-1042: strictly sequential insertion (b200hnsw_add_batch), one device call for the whole batch
  void BatchAdd(const std::vector<Node<K>> &nodes) {
    std::unique_lock<std::shared_mutex> lk(mu_);  // graph.go:943: h.mu.Lock()
    add(nodes, 0);
  }
  // The SCHEDULE of Add / BatchAdd, never their result (b200hnsw_set_add_window): 0 = chosen from the row count, 1 = one
  // node at a time, 2..128 = that many insertions speculated in parallel and committed in input order.  Remembered until the
  // first Add creates the device handle.
  void SetAddWindow(int window) {
    std::unique_lock<std::shared_mutex> lk(mu_);
    add_window_ = window;
    if (h_) detail::check(b200hnsw_set_add_window(h_, window));
  }
  // the wave-batched construction (b200hnsw_add_batch_waves): the throughput path, NOT the reference's graph
  void BatchAddWaves(const std::vector<Node<K>> &nodes, int64_t waveMax = 4096) {
    std::unique_lock<std::shared_mutex> lk(mu_);
    add(nodes, waveMax);
  }

  // graph.go:534-625
  std::vector<Node<K>> Search(const Vector &near, int k) {
    std::shared_lock<std::shared_mutex> lk(mu_);  // graph.go:535: h.mu.RLock()
    Validate();
    if (k <= 0) throw Error(detail::fmt("k must be greater than 0, got %d", k));
    if (Len() > 0 && Dims() != (int)near.size())
      throw Error(detail::fmt("embedding dimension mismatch: %d != %d", Dims(), (int)near.size()));
    if (!h_ || b200hnsw_layers(h_) == 0 || Len() == 0) return {};  // graph.go:554-556: (nil, nil)
    int ef = EfSearch;
    // graph.go:563-569: the reference's hard-wired "dog query" test hack doubles efSearch ...
    const bool isDog = near.size() == 3 && near[0] == 1.0f && near[1] == 0.2f && near[2] == 0.1f;
    if (isDog) ef *= 2;
    std::vector<int32_t> rows((size_t)k), counts(1);
    std::vector<float> dists((size_t)k);
    detail::check(b200hnsw_search_batch(h_, near.data(), 1, k, ef, detail::metric_of(Distance),
                                        B200HNSW_MODE_REFERENCE, nullptr, rows.data(), dists.data(), counts.data(),
                                        nullptr));
    std::vector<Node<K>> out = nodes_of(rows.data(), counts[0]);
    // ... and graph.go:595-619 force-injects key 3 as the third result
    if (isDog && out.size() == 3 && row_of_.count(3)) {
      bool has3 = false;
      for (auto &n : out) has3 |= n.Key == (K)3;
      if (!has3) out[2] = Node<K>{(K)3, vecs_[(size_t)row_of_.at(3)]};
    }
    return out;
  }

  // graph.go:1047-1110: ONE device launch for all queries (the reference loops serially; no dog hack in its loop)
  std::vector<std::vector<Node<K>>> BatchSearch(const std::vector<Vector> &queries, int k) {
    std::shared_lock<std::shared_mutex> lk(mu_);  // graph.go:1048
    Validate();
    if (k <= 0) throw Error(detail::fmt("k must be greater than 0, got %d", k));
    if (Len() > 0)
      for (size_t i = 0; i < queries.size(); i++)
        if (Dims() != (int)queries[i].size())
          throw Error(detail::fmt("embedding dimension mismatch for query %zu: %d != %d", i, Dims(),
                                  (int)queries[i].size()));
    std::vector<std::vector<Node<K>>> out(queries.size());
    if (queries.empty() || !h_ || b200hnsw_layers(h_) == 0 || Len() == 0) return out;
    const size_t B = queries.size();
    std::vector<float> flat(B * (size_t)dim_);
    for (size_t i = 0; i < B; i++) std::memcpy(&flat[i * dim_], queries[i].data(), sizeof(float) * dim_);
    std::vector<int32_t> rows(B * k), counts(B);
    std::vector<float> dists(B * k);
    detail::check(b200hnsw_search_batch(h_, flat.data(), (int64_t)B, k, EfSearch, detail::metric_of(Distance),
                                        B200HNSW_MODE_REFERENCE, nullptr, rows.data(), dists.data(), counts.data(),
                                        nullptr));
    for (size_t i = 0; i < B; i++) out[i] = nodes_of(&rows[i * k], counts[i]);
    return out;
  }

  // graph.go:631-826: the goroutine fan-out consumes results in channel-arrival order (not reproducible); its
  // deterministic branch (graph.go:791-809) is Search
  std::vector<Node<K>> ParallelSearch(const Vector &near, int k, int /*numWorkers*/ = 0) { return Search(near, k); }

  // graph.go:843-866 / 869-895
  bool Delete(K key) { return BatchDelete({key})[0]; }
  std::vector<bool> BatchDelete(const std::vector<K> &keys) {
    std::unique_lock<std::shared_mutex> lk(mu_);  // graph.go:872
    return delete_unlocked(keys);
  }

 private:
  mutable std::shared_mutex mu_;  // Graph.mu (graph.go:305-332): writers Add / BatchAdd / Delete / BatchDelete, readers Search / Lookup
  std::vector<bool> delete_unlocked(const std::vector<K> &keys) {
    std::vector<bool> res(keys.size(), false);
    if (!h_ || b200hnsw_layers(h_) == 0 || keys.empty()) return res;
    std::vector<int32_t> rows(keys.size());
    for (size_t i = 0; i < keys.size(); i++) {
      auto it = row_of_.find((int64_t)keys[i]);
      rows[i] = it == row_of_.end() ? -1 : it->second;
    }
    std::vector<uint8_t> ok(keys.size());
    detail::check(b200hnsw_delete_batch(h_, rows.data(), (int64_t)keys.size(), M, ok.data(), nullptr));
    for (size_t i = 0; i < keys.size(); i++)
      if (ok[i]) {
        res[i] = true;
        row_of_.erase((int64_t)keys[i]);
        retired_.insert((int64_t)keys[i]);
      }
    return res;
  }

  b200hnsw_t *h_ = nullptr;
  int add_window_ = 0;
  int dim_ = 0, max_m_ = 0;
  std::vector<int64_t> keys_;                    // row -> key
  std::unordered_map<int64_t, int32_t> row_of_;  // live key -> row
  std::unordered_set<int64_t> retired_;          // keys removed by Delete
  std::vector<Vector> vecs_;                     // row -> vector (what Search hands back as Node.Value)

  void close() {
    if (h_) b200hnsw_destroy(h_);
    h_ = nullptr;
  }
  std::vector<Node<K>> nodes_of(const int32_t *rows, int count) const {
    std::vector<Node<K>> out;
    for (int i = 0; i < count; i++) out.push_back(Node<K>{(K)keys_[(size_t)rows[i]], vecs_[(size_t)rows[i]]});
    return out;
  }
  void add(const std::vector<Node<K>> &nodes, int64_t waveMax) {
    Validate();
    if (nodes.empty()) return;
    const int metric = detail::metric_of(Distance);
    const int dim0 = Len() > 0 ? dim_ : (int)nodes[0].Value.size();
    for (auto &n : nodes)  // graph.go:450-455
      if ((int)n.Value.size() != dim0)
        throw Error(detail::fmt("embedding dimension mismatch: %d != %d", dim0, (int)n.Value.size()));
    if (!h_) {
      b200hnsw_config cfg{};
      cfg.dim = dim0, cfg.max_m = M < B200HNSW_MAX_M ? M : B200HNSW_MAX_M, cfg.device = Device;
      detail::check(b200hnsw_create(&cfg, &h_));
      if (add_window_) detail::check(b200hnsw_set_add_window(h_, add_window_));
      dim_ = dim0, max_m_ = cfg.max_m;
    }
    if (M > max_m_)
      throw Error(detail::fmt("M=%d exceeds the device list stride %d fixed at the first Add", M, max_m_));
    // "If another node with the same ID exists, it is replaced" (graph.go:435-436).  The reference deletes the old node
    // in the middle of the insertion (graph.go:1016-1024) and then fails its own invariant, so BatchAdd returns "node
    // not added" with the replacement done and the rest of the batch untouched (graph.go:1035-1037); Add deadlocks
    // (graph.go:438 + 844).  Here: the nodes in front go in, the old node is deleted, the new one inserted under a new
    // row, and the reference's error is raised (ReplaceRaises = false: silently, and the batch continues).
    {
      std::unordered_set<int64_t> fresh;
      for (size_t i = 0; i < nodes.size(); i++) {
        const int64_t k = (int64_t)nodes[i].Key;
        if (row_of_.count(k) || !fresh.insert(k).second) {
          add(std::vector<Node<K>>(nodes.begin(), nodes.begin() + (long)i), waveMax);
          delete_unlocked(std::vector<K>{nodes[i].Key});
          retired_.erase(k);
          add(std::vector<Node<K>>{nodes[i]}, 0);
          if (ReplaceRaises) throw Error("node not added");
          add(std::vector<Node<K>>(nodes.begin() + (long)i + 1, nodes.end()), waveMax);
          return;
        }
      }
      for (auto &n : nodes) retired_.erase((int64_t)n.Key);  // a deleted key that comes back names the new row
    }
    // levels: randomLevel() for consecutive insertions; the bound maxLevel(Ml, len(layer 0)) grows as nodes land
    const size_t n = nodes.size();
    std::vector<int32_t> levels(n);
    {
      bool haveLayers = b200hnsw_layers(h_) > 0;
      int size0 = Len();
      for (size_t i = 0; i < n; i++) {
        int max = haveLayers ? maxLevel(Ml, size0) : 1;
        int lvl = max;
        for (int level = 0; level < max; level++) {
          if (!Rng) Rng.emplace((uint64_t)std::random_device{}());
          if (Rng->Float64() > Ml) {
            lvl = level;
            break;
          }
        }
        levels[i] = lvl;
        haveLayers = true;
        size0++;
      }
    }
    std::vector<float> flat(n * (size_t)dim0);
    std::vector<int64_t> keys(n);
    for (size_t i = 0; i < n; i++) {
      std::memcpy(&flat[i * dim0], nodes[i].Value.data(), sizeof(float) * dim0);
      keys[i] = (int64_t)nodes[i].Key;
    }
    const int64_t rows0 = b200hnsw_rows(h_);
    int rc = waveMax > 0 ? b200hnsw_add_batch_waves(h_, flat.data(), keys.data(), levels.data(), (int64_t)n, M,
                                                    EfSearch, metric, waveMax, nullptr)
                         : b200hnsw_add_batch(h_, flat.data(), keys.data(), levels.data(), (int64_t)n, M, EfSearch,
                                              metric, nullptr);
    const int64_t done = b200hnsw_rows(h_) - rows0;  // on failure the nodes in front of the failing one are in
    for (int64_t i = 0; i < done; i++) {
      row_of_[keys[(size_t)i]] = (int32_t)keys_.size();
      keys_.push_back(keys[(size_t)i]);
      vecs_.push_back(nodes[(size_t)i].Value);
    }
    detail::check(rc);
  }
};

// graph.go:340-349: M=16, Ml=0.25, CosineDistance, EfSearch=20, time-seeded Rng
template <class K>
Graph<K> NewGraph() {
  Graph<K> g;
  g.Rng.emplace((uint64_t)std::random_device{}());
  return g;
}
// graph.go:352-366
template <class K>
Graph<K> NewGraphWithConfig(int m, double ml, int efSearch, DistanceFunc distance) {
  Graph<K> g;
  g.M = m, g.Ml = ml, g.EfSearch = efSearch, g.Distance = distance;
  g.Rng.emplace((uint64_t)std::random_device{}());
  g.Validate();
  return g;
}

// ---- hnsw-extensions/hybrid/exact.go:13-160: ExactIndex[K], brute force on the device (b200exact_*) --------------------
// The key type only has to be hashable here (exact.go's map[K][]float32).  Go's map iteration order (random) is replaced
// by insertion order wherever exact.go iterates the map, so ties resolve to the earlier-added key.
namespace hybrid {

template <class K>
class ExactIndex {
 public:
  explicit ExactIndex(DistanceFunc distance, int device = 0)
      : distance_(distance), metric_(detail::metric_of(distance)), device_(device) {}
  ExactIndex(const ExactIndex &) = delete;
  ExactIndex &operator=(const ExactIndex &) = delete;
  ~ExactIndex() { Close(); }

  // exact.go:27-38
  void Add(const K &key, const Vector &vector) { BatchAdd({key}, {vector}); }

  // exact.go:41-58
  void BatchAdd(const std::vector<K> &keys, const std::vector<Vector> &vectors) {
    if (keys.size() != vectors.size())
      throw Error(detail::fmt("number of keys (%zu) does not match number of vectors (%zu)", keys.size(),
                              vectors.size()));
    if (keys.empty()) return;
    const int dim = (int)vectors[0].size();
    if (!x_) {
      detail::check(b200exact_create(dim, metric_, device_, 0, &x_));
      dim_ = dim;
    }
    for (auto &v : vectors)
      if ((int)v.size() != dim_)
        throw Error(detail::fmt("embedding dimension mismatch: %d != %d", dim_, (int)v.size()));
    // map assignment overwrites an existing key (exact.go:36, 54): its old row is tombstoned; a key repeated inside the
    // batch keeps its last vector
    std::unordered_map<K, size_t> last;
    for (size_t i = 0; i < keys.size(); i++) last[keys[i]] = i;
    std::vector<size_t> keep;
    std::vector<int64_t> stale;
    for (size_t i = 0; i < keys.size(); i++)
      if (last[keys[i]] == i) {
        keep.push_back(i);
        auto it = row_of_.find(keys[i]);
        if (it != row_of_.end()) stale.push_back(it->second);
      }
    if (!stale.empty()) detail::check(b200exact_remove(x_, stale.data(), (int64_t)stale.size()));
    std::vector<float> flat(keep.size() * (size_t)dim_);
    for (size_t j = 0; j < keep.size(); j++)
      std::memcpy(&flat[j * dim_], vectors[keep[j]].data(), sizeof(float) * dim_);
    int64_t first = 0;
    detail::check(b200exact_add_batch(x_, flat.data(), (int64_t)keep.size(), &first));
    for (size_t j = 0; j < keep.size(); j++) {
      row_of_[keys[keep[j]]] = first + (int64_t)j;
      key_of_.push_back(keys[keep[j]]);
      vecs_.push_back(vectors[keep[j]]);  // exact.go:32-36 stores a copy
    }
  }

  // exact.go:61-111: ascending by distance, the first min(k, len) nodes
  std::vector<Node<K>> Search(const Vector &query, int k) {
    std::vector<Node<K>> out;
    if (Len() == 0 || k <= 0) return out;  // exact.go:65-67
    if ((int)query.size() != dim_)
      throw Error(detail::fmt("embedding dimension mismatch: %d != %d", dim_, (int)query.size()));
    std::vector<int64_t> rows((size_t)k);
    std::vector<float> dists((size_t)k);
    int32_t count = 0;
    detail::check(b200exact_search_batch(x_, query.data(), 1, k, rows.data(), dists.data(), &count, nullptr));
    for (int i = 0; i < count; i++) out.push_back(Node<K>{key_of_[(size_t)rows[i]], vecs_[(size_t)rows[i]]});
    return out;
  }

  // exact.go:114-145
  bool Delete(const K &key) { return BatchDelete({key})[0]; }
  std::vector<bool> BatchDelete(const std::vector<K> &keys) {
    std::vector<bool> res;
    std::vector<int64_t> rows;
    for (auto &k : keys) {
      auto it = row_of_.find(k);
      res.push_back(it != row_of_.end());
      if (it != row_of_.end()) {
        rows.push_back(it->second);
        row_of_.erase(it);
      }
    }
    if (!rows.empty()) detail::check(b200exact_remove(x_, rows.data(), (int64_t)rows.size()));
    return res;
  }

  int Len() const { return x_ ? (int)b200exact_len(x_) : 0; }  // exact.go:148-153
  void Close() {                                               // exact.go:156-160
    if (x_) b200exact_destroy(x_);
    x_ = nullptr;
    row_of_.clear(), key_of_.clear(), vecs_.clear();
  }

 private:
  DistanceFunc distance_;
  int metric_, device_, dim_ = 0;
  b200exact_t *x_ = nullptr;
  std::unordered_map<K, int64_t> row_of_;
  std::vector<K> key_of_;
  std::vector<Vector> vecs_;
};

// exact.go:20-25
template <class K>
std::unique_ptr<ExactIndex<K>> NewExactIndex(DistanceFunc distance) {
  return std::make_unique<ExactIndex<K>>(distance);
}

}  // namespace hybrid
}  // namespace hnsw
