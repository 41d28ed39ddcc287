// exact_gemm.cu -- K3 tensor path (placeholder until the tcgen05 kernel lands): reports "unsupported" so the
// caller uses the exact fp32 scan.
#include "exact_index.hpp"
namespace b200 {
int exact_candidates_tensor(b200exact *, const float *, int64_t, int, int32_t *, int32_t *, int, cudaStream_t) {
  return B200HNSW_ERR_UNSUPPORTED;
}
int exact_after_add(b200exact *, int64_t, int64_t, cudaStream_t) { return 0; }
}  // namespace b200
