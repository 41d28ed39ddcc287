// search_ext_impl.cuh -- the launch code of K1x (traverse_ext.cuh), instantiated once per W size (entries per lane) in
// search_ext_e2.cu / _e8.cu / _e32.cu so the three sets of kernels compile in parallel.
#pragma once
#include <mutex>

#include "index.hpp"
#include "traverse_ext.cuh"

namespace b200 {

// resident one-warp blocks per SM the register file should allow: rows in flight + query + W + ~44 working registers
template <int CH, int U, int EPL>
struct ExtMinBlocks {
  static constexpr int regs = (4 * CH * U + 4 * CH + 4 * EPL + 52 + 7) & ~7;
  static constexpr int raw = 65536 / 32 / regs;
  static constexpr int value = raw > 32 ? 32 : (raw < 1 ? 1 : raw);
};

template <int G, int CH, int U, int EPL>
static int launch_x(b200hnsw *h, const GraphView &g, TraverseParams &p, int ef_upper, cudaStream_t s) {
  auto kern = traverse_ext_kernel<G, CH, U, EPL, ExtMinBlocks<CH, U, EPL>::value>;
  const ExtLayout lay = ext_layout(g.M, p.vis_cap);
  const size_t smem = (size_t)lay.warp_bytes;
  if (smem > h->smem_optin - 1024) {
    set_error("extended search: visited=%d need %zu B of shared memory per traversal (> %zu)", p.vis_cap, smem,
              h->smem_optin - 1024);
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
  B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 32, smem));
  if (per_sm < 1) per_sm = 1;
  const long long resident = (long long)per_sm * h->sm_count;
  int grid = (int)(p.B < resident ? p.B : resident);  // persistent: one traversal at a time per warp
  if (grid < 1) grid = 1;
  kern<<<grid, 32, smem, s>>>(g, p, ef_upper);
  B200_CUDA(cudaGetLastError());
  return 0;
}

template <int EPL>
static int launch_ext_epl(b200hnsw *h, const GraphView &g, TraverseParams &p, int ef_upper, cudaStream_t s) {
  const int rc = g.row_chunks;
#define K1X(G_, CH_, U_) return launch_x<G_, CH_, U_, EPL>(h, g, p, ef_upper, s)
  // G = 8 kernels load rows without bounds tests: CH must equal row_chunks / 8 exactly (gdist.cuh)
  if (rc == 8) K1X(8, 1, 4);
  if (rc == 16) K1X(8, 2, 4);
  if (rc == 24) K1X(8, 3, 2);
  if (rc == 32) K1X(8, 4, 2);
  if (rc <= 64) K1X(16, 4, 2);
  if (rc <= 128) K1X(32, 4, 2);
  if (rc <= 192) K1X(32, 6, 2);
  if (rc <= 256) K1X(32, 8, 1);
  if (rc <= 384) K1X(32, 12, 1);
  if (rc <= 512) K1X(32, 16, 1);
#undef K1X
  set_error("search: dim %d > 2048 not supported", g.dim);
  return B200HNSW_ERR_UNSUPPORTED;
}

}  // namespace b200
