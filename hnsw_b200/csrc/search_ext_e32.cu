// K1x kernels whose result set W has up to 32 entries per lane (max(ef, k) <= 1024): see search_ext_impl.cuh
#include "search_ext_impl.cuh"
namespace b200 {
int launch_ext_e32(b200hnsw *h, const GraphView &g, TraverseParams &p, int ef_upper, cudaStream_t s) {
  return launch_ext_epl<32>(h, g, p, ef_upper, s);
}
}  // namespace b200
