// search_kernel.cu -- launcher of K1 (traverse.cuh) for Graph.Search / Graph.BatchSearch (graph.go:534-625, 1047-1110)
// and for the search half of the batched Add (wave_insert.cu).  Bound: HBM gather (E rows of 4*dim bytes per
// traversal; E is counted per launch in d_counters).
#include <cstdlib>
#include <mutex>

#include "index.hpp"
#include "traverse.cuh"

namespace b200 {

template <int G, int CH, int U, int MINB = 1, int TW = TRAV_WARPS, bool NANG = false>
static int launch_t(b200hnsw *h, const GraphView &g, TraverseParams &p, cudaStream_t s) {
  auto kern = traverse_kernel<G, CH, U, MINB, TW, NANG>;
  constexpr int NG = 32 / G;
  size_t smem = traverse_group_bytes(p) * TW * NG;
  if (smem > h->smem_optin - 512) {
    set_error("search: k=%d ef=%d visited=%d need %zu B of shared memory per block (> %zu)", p.k, p.ef, p.vis_cap,
              smem, h->smem_optin);
    return B200HNSW_ERR_UNSUPPORTED;
  }
  {
    // the opt-in limit of a kernel is per device and shared by every call in flight: it is only ever raised, under a
    // lock, so a concurrent call that needs less can never lower it between another call's test and its launch
    static std::mutex mu;
    static size_t raised[64] = {};
    const int dv = h->cfg.device & 63;
    std::lock_guard<std::mutex> lk(mu);
    if (smem > raised[dv]) {
      B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      raised[dv] = smem;
    }
  }
  int per_sm = 0;
  B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, TW * 32, smem));
  if (per_sm < 1) per_sm = 1;
  long long blocks_needed = (p.B + TW * NG - 1) / (TW * NG);
  long long resident = (long long)per_sm * h->sm_count;
  int grid = (int)(blocks_needed < resident ? blocks_needed : resident);
  if (grid < 1) grid = 1;
  kern<<<grid, TW * 32, smem, s>>>(g, p);
  B200_CUDA(cudaGetLastError());
  return 0;
}

static int round_pow2(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

// group width: the fewest lanes that keep a row within 4 chunks per lane (overridable for tuning)
static int pick_group(int row_chunks) {
  const char *env = getenv("B200HNSW_K1_G");
  if (env) {
    int v = atoi(env);
    if (v == 8 || v == 16 || v == 32) return v;
  }
  if (row_chunks <= 32) return 8;
  if (row_chunks <= 64) return 16;
  return 32;
}

int max_visited_cap(const b200hnsw *h, int k, int ef) {
  // one traversal per warp (G = 32), TRAV_WARPS warps per block, full-length queues
  const size_t fixed = ((size_t)((ef + 2) & ~1) + (size_t)((k + 1) & ~1)) * 8 + (size_t)((h->M + 3) & ~3) * 4;
  int cap = 256;
  while ((fixed + (size_t)cap * 2 * 4) * TRAV_WARPS <= h->smem_optin - 512) cap *= 2;
  return cap;
}

// the dispatch table: group width x chunks per lane x rows per batch (x the NaN-guard instantiation, traverse.cuh)
template <bool NANG>
static int traverse_dispatch(b200hnsw *h, const GraphView &g, TraverseParams &p, int G, int ch, cudaStream_t s) {
#define K1(G_, CH_, U_) \
  return launch_t<G_, CH_, U_, ((CH_) >= 12 ? 2 : ((CH_) * (U_) <= 8 ? 5 : 4)), TRAV_WARPS, NANG>(h, g, p, s)
  if (G == 8) {
    // G = 8 kernels load rows without bounds tests: CH must equal row_chunks / 8 exactly (gdist.cuh)
    if (ch == 1) K1(8, 1, 8);
    if (ch == 2) K1(8, 2, 4);
    if (ch == 3) K1(8, 3, 2);
    // dim 128: two-warp blocks, 10 per SM (measured with three launches in flight: 73.9M queries/s against 72.9M for
    // four-warp blocks and 70.1M for one-warp blocks, profiles/r4j_k1_block_shape.log)
    if (ch == 4) return launch_t<8, 4, 2, 10, 2, NANG>(h, g, p, s);
    G = 16;
    ch = (g.row_chunks + 15) / 16;
  }
  if (G == 16) {
    if (ch <= 1) K1(16, 1, 8);
    if (ch <= 2) K1(16, 2, 4);
    if (ch <= 4) K1(16, 4, 4);
    G = 32;
    ch = (g.row_chunks + 31) / 32;
  }
  if (ch <= 1) K1(32, 1, 8);
  if (ch <= 2) K1(32, 2, 4);
  if (ch <= 4) K1(32, 4, 4);
  if (ch <= 6) K1(32, 6, 2);
  if (ch <= 8) K1(32, 8, 2);
  if (ch <= 12) K1(32, 12, 1);
  if (ch <= 16) K1(32, 16, 1);
#undef K1
  set_error("search: dim %d > 2048 not supported", g.dim);
  return B200HNSW_ERR_UNSUPPORTED;
}

int launch_traverse(b200hnsw *h, TraverseParams &p, int vis_cap, cudaStream_t s) {
  GraphView g = h->view();
  if (p.n_layers == 0) p.n_layers = h->n_layers;
  int kr = p.k > p.k_upper ? p.k : p.k_upper;
  // The candidate queue is only maintained until the result queue is full (traverse.cuh): at most kr - 1 rows were
  // evaluated before the step that fills it and at most M in that step, and PopLast keeps it at ef: min(ef, kr-1+M)
  // entries, +1 for the push in front of PopLast.
  // With NaN distances possible (rows flagged at Add time, or a retry after the kernel reported error 3) the queue is
  // kept for the whole layer and needs its full ef + 1 entries.
  {
    int need = kr - 1 + g.M;
    if (need > p.ef || p.heap_c_full || (h->row_flags.load() & 1)) {
      need = p.ef;
      p.heap_c_full = 1;
    }
    p.heap_c = (need + 2) & ~1;
  }
  p.row_flags = h->d_row_flags;
  p.heap_r = (kr + 1) & ~1;
  p.fresh_cap = (g.M + 3) & ~3;
  p.vis_cap = round_pow2(vis_cap < 256 ? 256 : vis_cap);
  p.off_r = p.heap_c * 8;
  p.off_fresh = (p.heap_c + p.heap_r) * 8;
  p.off_vis = p.off_fresh + p.fresh_cap * 4;
  p.group_bytes = (int)traverse_group_bytes(p);
  int G = pick_group(g.row_chunks);
  // a large visited table (retry after overflow) only fits with one traversal per warp
  if (traverse_group_bytes(p) * TRAV_WARPS * (32 / G) > h->smem_optin - 512) G = 32;
  int ch = (g.row_chunks + G - 1) / G;
#ifdef B200_K1_TUNING  // tuning builds only (scripts/k1_variants.py): eighteen more instantiations for dim 128
  if (const char *v = getenv("B200HNSW_K1_VARIANT")) {  // tuning experiments (dim 128 only)
    int id = atoi(v);
    if (g.row_chunks == 32) {
      if (id == 1) return launch_t<8, 4, 4, 5>(h, g, p, s);
      if (id == 2) return launch_t<8, 4, 8, 1>(h, g, p, s);
      if (id == 3) return launch_t<16, 2, 8, 1>(h, g, p, s);
      if (id == 4) return launch_t<16, 2, 4, 4>(h, g, p, s);
      if (id == 5) return launch_t<32, 1, 16, 1>(h, g, p, s);
      if (id == 6) return launch_t<16, 2, 8, 3>(h, g, p, s);
      if (id == 7) return launch_t<8, 4, 2, 5>(h, g, p, s);
      if (id == 8) return launch_t<8, 4, 1, 6>(h, g, p, s);
      if (id == 9) return launch_t<8, 4, 2, 6>(h, g, p, s);
      if (id == 10) return launch_t<16, 2, 2, 6>(h, g, p, s);
      if (id == 11) return launch_t<16, 2, 4, 6>(h, g, p, s);
      if (id == 12) return launch_t<8, 4, 1, 8>(h, g, p, s);
      if (id == 13) return launch_t<16, 2, 2, 8>(h, g, p, s);
      if (id == 14) return launch_t<16, 2, 4, 8>(h, g, p, s);
      if (id == 15) return launch_t<8, 4, 2, 6>(h, g, p, s);
      if (id == 16) return launch_t<16, 2, 2, 6>(h, g, p, s);
      if (id == 17) return launch_t<8, 4, 2, 21, 1>(h, g, p, s);  // one-warp blocks: 21 warps per SM instead of 20
      if (id == 18) return launch_t<8, 4, 2, 10, 2>(h, g, p, s);
    }
  }
#endif
  return p.heap_c_full ? traverse_dispatch<true>(h, g, p, G, ch, s) : traverse_dispatch<false>(h, g, p, G, ch, s);
}

int launch_search(b200hnsw *h, const float *d_queries, int64_t B, int k, int ef, int metric, int mode,
                  const int32_t *entries, int32_t *d_rows, float *d_dists, int32_t *d_counts,
                  unsigned long long *d_counters, int *d_error, int vis_cap, cudaStream_t s, int64_t seg_len,
                  int64_t seg_stride, int32_t *const *seg_ptrs, int n_seg_ptrs, int full_queue) {
  if (mode != B200HNSW_MODE_REFERENCE && mode != B200HNSW_MODE_EXTENDED) {
    set_error("search: unknown mode %d", mode);
    return B200HNSW_ERR_BAD_ARG;
  }
  TraverseParams p{};
  p.queries = d_queries;
  p.q_stride = h->cfg.dim;
  p.B = B;
  p.k = k;
  p.k_upper = 1;  // graph.go:579
  p.ef = ef;
  p.metric = metric;
  for (int l = 0; l < B200HNSW_MAX_LAYERS; l++)
    p.entries[l] = l < h->n_layers ? (entries ? entries[l] : h->entry[l]) : -1;
  p.out_rows = d_rows;
  p.out_dists = d_dists;
  p.out_counts = d_counts;
  p.out_seg_len = seg_len;
  p.out_seg_stride = seg_stride;
  p.n_seg_ptr = n_seg_ptrs;
  for (int i = 0; i < n_seg_ptrs && i < B200HNSW_MAX_PEERS; i++) p.seg_ptr[i] = seg_ptrs[i];
  p.counters = d_counters;
  p.error = d_error;
  p.dead = h->n_dead ? h->d.dead : nullptr;
  p.heap_c_full = full_queue;
  p.n_layers = h->live_layers();  // layers Delete emptied at the top are passed through without an elevator
  if (mode == B200HNSW_MODE_EXTENDED) {
    // textbook layer search (traverse_ext.cuh): greedy k = ef = 1 descent on the upper layers, W of max(ef, k) on layer 0
    const int efk = ef > k ? ef : k;
    int want = ext_visited_default(efk, h->M);
    const int mx = max_visited_cap_ext(h, efk);
    if (want > mx) want = mx;
    // the host-pointer call passes its own (retry-grown) size; device-pointer calls pass the reference mode's hint, which
    // is smaller than `want` unless an extended retry raised vis_hint_ext
    const int hint = h->vis_hint_ext.load();
    if (hint > want) want = hint;
    return launch_traverse_ext(h, p, 1, vis_cap > want ? vis_cap : want, s);
  }
  return launch_traverse(h, p, vis_cap, s);
}

}  // namespace b200
