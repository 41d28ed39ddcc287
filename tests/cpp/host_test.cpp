// host_test.cpp -- the C++ host mirror (include/b200hnsw_host.hpp) exercised the way the reference's own tests exercise
// the Go API, plus parity against the CPU oracle (oracle/hnsw_oracle.h: test infrastructure, linked here only).
//   host_test --cpu : what needs no device -- validation texts (graph_test.go:415-459), maxLevel (graph_test.go:14-25),
//                     the loud failure of a compute call without a CUDA device
//   host_test --gpu : distance KATs (distance_test.go:9-31), TestGraph_DefaultCosine (graph_test.go:253-275), and a random
//                     graph built through Graph.Add / BatchAddWaves whose Search / BatchSearch keys, in order, equal the
//                     oracle's on the same rows and levels
#include <cstdio>
#include <cstdlib>
#include <atomic>
#include <string>
#include <thread>
#include <vector>

#include <sstream>

#include "b200hnsw_encode.hpp"
#include "b200hnsw_host.hpp"
#include "hnsw_oracle.h"

static int failures = 0;
#define CHECK(cond)                                                           \
  do {                                                                        \
    if (!(cond)) {                                                            \
      std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond);             \
      failures++;                                                             \
    }                                                                         \
  } while (0)

template <class F>
static std::string error_of(F &&f) {
  try {
    f();
  } catch (const hnsw::Error &e) {
    return e.what();
  }
  return "";
}
static bool contains(const std::string &s, const char *sub) { return s.find(sub) != std::string::npos; }

static void cpu_tests() {
  using namespace hnsw;
  // graph_test.go:14-25
  CHECK(maxLevel(0.5, 10) == 4);
  CHECK(maxLevel(0.5, 1000) == 11);
  CHECK(contains(error_of([] { maxLevel(0, 10); }), "ml must be greater than 0"));
  // graph_test.go:415-459
  CHECK(error_of([] { NewGraphWithConfig<int>(16, 0.25, 20, CosineDistance); }).empty());
  CHECK(contains(error_of([] { NewGraphWithConfig<int>(0, 0.25, 20, CosineDistance); }), "M must be greater than 0"));
  CHECK(contains(error_of([] { NewGraphWithConfig<int>(16, 0, 20, CosineDistance); }), "Ml must be between 0 and 1"));
  CHECK(contains(error_of([] { NewGraphWithConfig<int>(16, 1.5, 20, CosineDistance); }), "Ml must be between 0 and 1"));
  CHECK(contains(error_of([] { NewGraphWithConfig<int>(16, 0.25, 0, CosineDistance); }),
                 "EfSearch must be greater than 0"));
  CHECK(contains(error_of([] { NewGraphWithConfig<int>(16, 0.25, 20, nullptr); }), "Distance function must be set"));
  {
    auto g = NewGraph<int>();
    CHECK(contains(error_of([&] { g.Search({1, 2, 3}, 0); }), "k must be greater than 0"));
    CHECK(contains(error_of([&] { g.ParallelSearch({1, 2, 3}, -1, 4); }), "k must be greater than 0"));
    CHECK(g.Search({1, 2, 3}, 1).empty());  // graph.go:554-556: empty graph -> (nil, nil)
    CHECK(g.Len() == 0 && g.Dims() == 0 && !g.Lookup(1).second);
    CHECK(g.M == 16 && g.Ml == 0.25 && g.EfSearch == 20 && g.Distance == CosineDistance);  // graph.go:340-349
  }
  {  // encode_test.go:11-50
    std::string b;
    enc::putVarint(b, 1337);
    CHECK(b.size() == 2);
    std::istringstream is(b);
    CHECK(enc::readVarint(is) == 1337);
    std::string s;
    enc::putString(s, "hello");
    CHECK(s.size() == 6);
    std::istringstream is2(s);
    CHECK(enc::readString(is2) == "hello");
    for (int64_t v : {0LL, -1LL, 1LL, -64LL, 63LL, 64LL, (long long)INT64_MAX, (long long)INT64_MIN}) {
      std::string t;
      enc::putVarint(t, v);
      uint8_t ref[16];
      CHECK(ora_put_varint(ref, v) == (int)t.size() && std::memcmp(ref, t.data(), t.size()) == 0);
      std::istringstream is3(t);
      CHECK(enc::readVarint(is3) == v);
    }
  }
  // distance.go:25-46
  CHECK(distanceFuncToName(EuclideanDistance).first == "euclidean" && distanceFuncToName(CosineDistance).second);
  CHECK(!distanceFuncToName(nullptr).second);
}

static void no_device_test() {
  using namespace hnsw;
  auto g = NewGraph<int>();
  std::string e = error_of([&] { g.Add(MakeNode(1, {1, 1})); });
  CHECK(contains(e, "b200hnsw:"));  // B200HNSW_ERR_CUDA surfaced as an error: no CPU fallback
  CHECK(g.Len() == 0);
}

static void gpu_tests() {
  using namespace hnsw;
  // distance_test.go:9-31
  CHECK(EuclideanDistance({1, 2, 3}, {4, 5, 6}) == 5.196152f);
  CHECK(std::fabs(CosineDistance({1, 1, 1}, {0.8f, 0.8f, 0.8f})) < 1e-6f);
  CHECK(std::fabs(CosineDistance({1, 0}, {0, 1}) - 1.f) < 1e-6f);
  CHECK(std::fabs(CosineDistance({1, 0}, {1, 0})) < 1e-6f);
  {  // graph_test.go:253-275
    auto g = NewGraph<int>();
    g.Add(MakeNode(1, {1, 1}), MakeNode(2, {0, 1}), MakeNode(3, {1, -1}));
    auto nb = g.Search({0.5f, 0.5f}, 1);
    CHECK(nb.size() == 1 && nb[0].Key == 1 && nb[0].Value == Vector({1, 1}));
    CHECK(g.Len() == 3 && g.Dims() == 2 && g.Lookup(2).second && g.Lookup(2).first == Vector({0, 1}));
    CHECK(contains(error_of([&] { g.Search({0.5f}, 1); }), "embedding dimension mismatch: 2 != 1"));
    CHECK(contains(error_of([&] { g.Add(MakeNode(9, {1, 2, 3})); }), "embedding dimension mismatch: 2 != 3"));
  }
  // random graph: host Graph.Add (exact sequential order) and BatchAddWaves vs the oracle on the same rows and levels
  for (int waves = 0; waves < 2; waves++) {
    const int n = waves ? 3000 : 1200, d = 24, M = 8, ef = 32, k = 5, nq = 100;
    std::mt19937 rng(7 + waves);
    std::uniform_real_distribution<float> U(-1.f, 1.f);
    std::vector<Node<int64_t>> nodes;
    for (int i = 0; i < n; i++) {
      Vector v(d);
      for (auto &x : v) x = U(rng);
      nodes.push_back(MakeNode<int64_t>(1000 - 3 * i, v));  // descending keys: list order != insertion order
    }
    auto g = NewGraphWithConfig<int64_t>(M, 0.25, ef, EuclideanDistance);
    g.Rng.emplace((uint64_t)(99 + waves));
    if (!waves) g.SetAddWindow(6);  // the speculative schedule of the reference-order Add: the same graph (spec_insert.cu)
    if (waves) g.BatchAddWaves(nodes, 256);
    else g.BatchAdd(nodes);
    CHECK(g.Len() == n);
    // the levels the host drew, read back from the device
    std::vector<float> vec((size_t)n * d);
    std::vector<int64_t> keys(n);
    std::vector<int32_t> levels(n);
    CHECK(b200hnsw_get_rows(g.Handle(), 0, n, vec.data(), keys.data(), levels.data()) == 0);
    ora_graph *og = ora_graph_new(d, M, ORA_EUCLIDEAN, ORA_FLAVOR_CANONICAL);
    if (waves) ora_graph_add_waves(og, keys.data(), vec.data(), levels.data(), n, ef, 256, 2);
    else ora_graph_add_many(og, keys.data(), vec.data(), levels.data(), n, ef);
    auto topo = g.Topography();
    CHECK((int)topo.size() == ora_graph_layers(og));
    for (size_t l = 0; l < topo.size(); l++) CHECK(topo[l] == (int)ora_graph_layer_size(og, (int)l));
    std::vector<Vector> qs;
    for (int i = 0; i < nq; i++) {
      Vector q(d);
      for (auto &x : q) x = U(rng);
      qs.push_back(q);
    }
    auto batch = g.BatchSearch(qs, k);
    int mismatched = 0;
    for (int i = 0; i < nq; i++) {
      int32_t rows[16];
      float dist[16];
      int cnt = ora_graph_search(og, qs[i].data(), k, ef, ORA_MODE_REFERENCE, nullptr, rows, dist);
      auto one = g.Search(qs[i], k);
      bool ok = (int)one.size() == cnt && (int)batch[i].size() == cnt;
      for (int j = 0; ok && j < cnt; j++)
        ok = one[j].Key == ora_graph_key(og, rows[j]) && batch[i][j].Key == one[j].Key &&
             one[j].Value == Vector(ora_graph_vector(og, rows[j]), ora_graph_vector(og, rows[j]) + d);
      mismatched += !ok;
    }
    CHECK(mismatched == 0);
    {  // Export (encode.go:133-174) is byte-identical to the oracle's; Import of that stream searches identically
      std::ostringstream os;
      Export(g, os);
      const std::string mine = os.str();
      std::vector<uint8_t> buf(mine.size() + 1024);
      int64_t olen = ora_graph_export(og, 0.25, ef, buf.data(), (int64_t)buf.size());
      CHECK(olen == (int64_t)mine.size() && std::memcmp(buf.data(), mine.data(), mine.size()) == 0);
      auto g2 = NewGraph<int64_t>();
      std::istringstream is(mine);
      Import(g2, is);
      CHECK(g2.M == M && g2.Ml == 0.25 && g2.EfSearch == ef && g2.Distance == EuclideanDistance && g2.Len() == n);
      CHECK(g2.Topography() == g.Topography());
      int diff = 0;
      for (int i = 0; i < 20; i++) {
        auto a = g.Search(qs[i], k), b = g2.Search(qs[i], k);
        bool ok = a.size() == b.size();
        for (size_t j = 0; ok && j < a.size(); j++) ok = a[j].Key == b[j].Key && a[j].Value == b[j].Value;
        diff += !ok;
      }
      CHECK(diff == 0);
      std::istringstream bad(std::string("\x04\x10"));  // version 2
      CHECK(!error_of([&] { auto g3 = NewGraph<int64_t>(); Import(g3, bad); }).empty());
    }
    // Delete through the host: the key leaves Lookup / Len, a second Delete reports false (graph.go:843-866)
    CHECK(g.Delete(nodes[5].Key) && !g.Delete(nodes[5].Key) && g.Len() == n - 1 && !g.Lookup(nodes[5].Key).second);
    CHECK(ora_graph_delete(og, nodes[5].Key) == 1);
    auto after = g.Search(qs[0], k);
    int32_t rows[16];
    float dist[16];
    int cnt = ora_graph_search(og, qs[0].data(), k, ef, ORA_MODE_REFERENCE, nullptr, rows, dist);
    CHECK((int)after.size() == cnt);
    for (int j = 0; j < cnt && j < (int)after.size(); j++) CHECK(after[j].Key == ora_graph_key(og, rows[j]));
    // a key that exists is replaced and the reference's own error comes back (graph.go:435-436, 1016-1037)
    {
      auto repl = nodes[7];
      for (auto &v : repl.Value) v = -v;
      CHECK(contains(error_of([&] { g.Add(repl); }), "node not added"));
      CHECK(g.Len() == n - 1 && g.Lookup(repl.Key).second && g.Lookup(repl.Key).first == repl.Value);
      // Export after Delete writes the live nodes only; Import of that stream answers like the exporter's live part
      std::ostringstream os2;
      Export(g, os2);
      std::istringstream is2(os2.str());
      auto g4 = NewGraph<int64_t>();
      Import(g4, is2);
      CHECK(g4.Len() == g.Len() && !g4.Lookup(nodes[5].Key).second && g4.Lookup(repl.Key).first == repl.Value);
    }
    ora_graph_free(og);
  }
}

// hybrid.ExactIndex (exact.go:13-160) through the C++ host against the oracle's restatement of ExactIndex.Search
static void gpu_exact_tests() {
  using namespace hnsw;
  for (int metric = 0; metric < 2; metric++) {
    const int n = 5000, d = 48, k = 7, nq = 60;
    std::mt19937 rng(21 + metric);
    std::uniform_real_distribution<float> U(-1.f, 1.f);
    std::vector<std::string> keys;
    std::vector<Vector> vecs;
    std::vector<float> flat;
    for (int i = 0; i < n; i++) {
      Vector v(d);
      for (auto &x : v) x = U(rng);
      keys.push_back("k" + std::to_string(i));  // string keys: the exact index only hashes them
      vecs.push_back(v);
      flat.insert(flat.end(), v.begin(), v.end());
    }
    auto x = hybrid::NewExactIndex<std::string>(metric ? CosineDistance : EuclideanDistance);
    CHECK(x->Search(vecs[0], 3).empty());  // exact.go:65-67
    x->BatchAdd(keys, vecs);
    CHECK(x->Len() == n);
    int bad = 0;
    for (int i = 0; i < nq; i++) {
      Vector q(d);
      for (auto &v : q) v = U(rng);
      int32_t idx[16];
      float dist[16];
      int cnt = ora_exact_search(flat.data(), n, d, metric, ORA_FLAVOR_CANONICAL, q.data(), k, 0, idx, dist);
      auto got = x->Search(q, k);
      bool ok = (int)got.size() == cnt;
      for (int j = 0; ok && j < cnt; j++) ok = got[j].Key == keys[(size_t)idx[j]] && got[j].Value == vecs[(size_t)idx[j]];
      bad += !ok;
    }
    CHECK(bad == 0);
    // overwrite (exact.go:36) and delete (exact.go:114-126)
    x->Add("k3", vecs[10]);
    CHECK(x->Len() == n);
    auto hit = x->Search(vecs[10], 2);
    CHECK(hit.size() == 2 && hit[0].Key == "k10" && hit[1].Key == "k3");  // equal distance: the earlier row first
    CHECK(x->Delete("k10") && !x->Delete("k10") && x->Len() == n - 1);
    hit = x->Search(vecs[10], 1);
    CHECK(hit.size() == 1 && hit[0].Key == "k3");
    CHECK(contains(error_of([&] { x->BatchAdd({"a", "b"}, {vecs[0]}); }), "number of keys (2) does not match"));
    x->Close();
    CHECK(x->Len() == 0);
  }
}

// TestThreadSafety (graph_test.go): goroutines mixing Add / Search / Delete / Lookup on one graph.  Graph.mu is a
// std::shared_mutex in the host mirror; an Add may legitimately fail with the reference's own "no nodes found" text when
// a concurrent Delete emptied the neighbourhood (the reference test only logs those), everything else must hold.
static void gpu_thread_safety_test() {
  using namespace hnsw;
  auto g = NewGraph<int64_t>();
  g.Rng = Rand(5);
  const int d = 8, threads = 8, ops = 40;
  for (int i = 0; i < 32; i++) {
    Vector v(d);
    for (int j = 0; j < d; j++) v[j] = (float)((i * 7 + j * 3) % 11) / 11.f + 0.01f;
    g.Add(MakeNode<int64_t>(i, v));
  }
  std::atomic<int> hard{0}, soft{0};
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; t++)
    pool.emplace_back([&, t] {
      std::mt19937 rng(100 + t);
      std::uniform_real_distribution<float> U(0.f, 1.f);
      for (int i = 0; i < ops; i++) {
        Vector v(d);
        for (auto &x : v) x = U(rng) + 0.01f;
        const int64_t key = 1000 + t * ops + i;
        try {
          switch (i % 4) {
            case 0: g.Add(MakeNode<int64_t>(key, v)); break;
            case 1: {
              auto r = g.Search(v, 3);
              if (r.size() > 3) hard++;
            } break;
            case 2: g.Delete(key - 2); break;
            default: g.Lookup(key - 3);
          }
        } catch (const Error &e) {
          if (i % 4 == 0) soft++; else hard++;
        }
      }
    });
  for (auto &th : pool) th.join();
  CHECK(hard == 0);
  CHECK(g.Len() >= 1 && g.Len() <= 32 + threads * ops / 4);
  auto r = g.Search(Vector(d, 0.5f), 3);  // may be empty: a deleted elevator ends the descent (graph.go:573-593)
  CHECK(r.size() <= 3);
  std::printf("thread safety: %d add(s) refused with the reference's own error, len %d\n", (int)soft, g.Len());
}

int main(int argc, char **argv) {
  const std::string mode = argc > 1 ? argv[1] : "--cpu";
  cpu_tests();
  if (mode == "--gpu") {
    gpu_tests();
    gpu_exact_tests();
    gpu_thread_safety_test();
  }
  else if (mode == "--cpu-nodevice") no_device_test();
  std::printf("%s: %d failure(s)\n", mode.c_str(), failures);
  return failures ? 1 : 0;
}
