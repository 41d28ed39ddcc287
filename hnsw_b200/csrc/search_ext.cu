// search_ext.cu -- launcher of K1x (traverse_ext.cuh), the EXTENDED search mode: textbook HNSW layer search, one warp
// per traversal.  Not the reference's algorithm (include/b200hnsw.h, B200HNSW_MODE_EXTENDED); its specification and
// CPU restatement are oracle/ora_graph.c's ORA_MODE_EXTENDED.  Bound: HBM gather.
#include "index.hpp"
#include "traverse_ext.cuh"

namespace b200 {

int launch_ext_e2(b200hnsw *h, const GraphView &g, TraverseParams &p, int ef_upper, cudaStream_t s);
int launch_ext_e4(b200hnsw *h, const GraphView &g, TraverseParams &p, int ef_upper, cudaStream_t s);
int launch_ext_e8(b200hnsw *h, const GraphView &g, TraverseParams &p, int ef_upper, cudaStream_t s);
int launch_ext_e32(b200hnsw *h, const GraphView &g, TraverseParams &p, int ef_upper, cudaStream_t s);

static int pow2_ge(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

// Visited-table entries an extended traversal is given by default.  Measured (1M x 128, M = 16): a walk with W of efk
// entries visits ~21 efk nodes at efk = 16, ~13 efk at 64, ~8 efk at 256; the table may fill to 3/4.  A walk that
// outgrows it fails loudly and the host-pointer call retries with 4x.
int ext_visited_default(int efk, int M) {
  long long est = (10LL * efk + 512) * (M > 16 ? (M + 15) / 16 : 1);
  if (est > (1 << 20)) est = 1 << 20;
  return pow2_ge((int)(est * 4 / 3 + 1));
}

int max_visited_cap_ext(const b200hnsw *h, int efk_max) {
  (void)efk_max;  // W lives in registers: shared memory holds the visited table and the fresh list only
  int cap = 256;
  while ((size_t)ext_layout(h->M, cap * 2).warp_bytes <= h->smem_optin - 1024) cap *= 2;
  return cap;
}

int launch_traverse_ext(b200hnsw *h, TraverseParams &p, int ef_upper, int vis_cap, cudaStream_t s) {
  GraphView g = h->view();
  if (p.n_layers == 0) p.n_layers = h->n_layers;
  const int e0 = p.ef > p.k ? p.ef : p.k, eU = ef_upper > p.k_upper ? ef_upper : p.k_upper;
  const int efk_max = e0 > eU ? e0 : eU;
  p.vis_cap = pow2_ge(vis_cap < 256 ? 256 : vis_cap);
  if (efk_max <= 64) return launch_ext_e2(h, g, p, ef_upper, s);
  if (efk_max <= 128) return launch_ext_e4(h, g, p, ef_upper, s);
  if (efk_max <= 256) return launch_ext_e8(h, g, p, ef_upper, s);
  if (efk_max <= 1024) return launch_ext_e32(h, g, p, ef_upper, s);
  set_error("extended search: max(ef, k) = %d > 1024 not supported", efk_max);
  return B200HNSW_ERR_UNSUPPORTED;
}

}  // namespace b200
