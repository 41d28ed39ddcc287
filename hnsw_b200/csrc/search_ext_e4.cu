// K1x kernels whose result set W has up to 4 entries per lane (max(ef, k) <= 128): see search_ext_impl.cuh
#include "search_ext_impl.cuh"
namespace b200 {
int launch_ext_e4(b200hnsw *h, const GraphView &g, TraverseParams &p, int ef_upper, cudaStream_t s) {
  return launch_ext_epl<4>(h, g, p, ef_upper, s);
}
}  // namespace b200
