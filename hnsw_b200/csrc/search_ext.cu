// search_ext.cu -- launcher of K1x (traverse_ext.cuh), the EXTENDED search mode: textbook HNSW layer search, one warp
// per traversal.  Not the reference's algorithm (include/b200hnsw.h, B200HNSW_MODE_EXTENDED); its specification and
// CPU restatement are oracle/ora_graph.c's ORA_MODE_EXTENDED.  Bound: HBM gather.
#include <mutex>

#include "index.hpp"
#include "traverse_ext.cuh"

namespace b200 {

template <int G, int CH, int U>
static int launch_x(b200hnsw *h, const GraphView &g, TraverseParams &p, int ef_upper, int efk_max, cudaStream_t s) {
  constexpr int TW = 1;  // one-warp blocks: the block scheduler packs as many traversals per SM as shared memory allows
  auto kern = traverse_ext_kernel<G, CH, U, TW>;
  const ExtLayout lay = ext_layout(efk_max, g.M, p.vis_cap);
  const size_t smem = (size_t)lay.warp_bytes * TW;
  if (smem > h->smem_optin - 1024) {
    set_error("extended search: k=%d ef=%d visited=%d need %zu B of shared memory per traversal (> %zu)", p.k, p.ef,
              p.vis_cap, smem, h->smem_optin - 1024);
    return B200HNSW_ERR_UNSUPPORTED;
  }
  {
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
  const long long blocks_needed = (p.B + TW - 1) / TW;
  const long long resident = (long long)per_sm * h->sm_count;
  int grid = (int)(blocks_needed < resident ? blocks_needed : resident);
  if (grid < 1) grid = 1;
  kern<<<grid, TW * 32, smem, s>>>(g, p, ef_upper, efk_max);
  B200_CUDA(cudaGetLastError());
  return 0;
}

static int pow2_ge(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

// visited-table entries an extended traversal is given when the handle's hint is smaller: a search with W of efk
// entries expands about efk..2 efk nodes of up to M neighbours, of which roughly a third to a half are new
int ext_visited_default(int efk, int M) {
  long long est = (long long)(efk + M) * M;  // nodes visited, generous
  if (est < 256) est = 256;
  if (est > (1 << 20)) est = 1 << 20;
  return pow2_ge((int)(2 * est));
}

int max_visited_cap_ext(const b200hnsw *h, int efk_max) {
  int cap = 256;
  while ((size_t)ext_layout(efk_max, h->M, cap * 2).warp_bytes <= h->smem_optin - 1024) cap *= 2;
  return cap;
}

int launch_traverse_ext(b200hnsw *h, TraverseParams &p, int ef_upper, int vis_cap, cudaStream_t s) {
  GraphView g = h->view();
  if (p.n_layers == 0) p.n_layers = h->n_layers;
  const int e0 = p.ef > p.k ? p.ef : p.k, eU = ef_upper > p.k_upper ? ef_upper : p.k_upper;
  const int efk_max = e0 > eU ? e0 : eU;
  p.vis_cap = pow2_ge(vis_cap < 256 ? 256 : vis_cap);
  const int rc = g.row_chunks;
#define K1X(G_, CH_, U_) return launch_x<G_, CH_, U_>(h, g, p, ef_upper, efk_max, s)
  // G = 8 kernels load rows without bounds tests: CH must equal row_chunks / 8 exactly (gdist.cuh)
  if (rc == 8) K1X(8, 1, 8);
  if (rc == 16) K1X(8, 2, 4);
  if (rc == 24) K1X(8, 3, 4);
  if (rc == 32) K1X(8, 4, 4);
  if (rc <= 64) K1X(16, 4, 4);
  if (rc <= 128) K1X(32, 4, 4);
  if (rc <= 192) K1X(32, 6, 2);
  if (rc <= 256) K1X(32, 8, 2);
  if (rc <= 384) K1X(32, 12, 1);
  if (rc <= 512) K1X(32, 16, 1);
#undef K1X
  set_error("search: dim %d > 2048 not supported", g.dim);
  return B200HNSW_ERR_UNSUPPORTED;
}

}  // namespace b200
