// exact_index.hpp -- host object behind b200exact_t (hybrid.ExactIndex, hnsw-extensions/hybrid/exact.go:13-25).
#pragma once
#include <cuda_fp16.h>

#include "gdist.cuh"
#include "index.hpp"

struct b200exact {
  int dim = 0, metric = 0, device = 0;
  int row_floats = 0;   // dim padded to a multiple of 32 floats
  int kpad = 0;         // dim padded to a multiple of 64 (fp16 K extent of the tensor path)
  int64_t n = 0, cap = 0, live = 0;
  float *d_vec = nullptr;            // [cap][row_floats] fp32, zero padded (what results are computed from)
  uint8_t *d_alive = nullptr;        // [cap] 0 = deleted (exact.go:114-145)
  __half *d_vech = nullptr;          // [cap][kpad] fp16 copy of the rows scaled by 2^sx_exp, for the tensor path
  float *d_aux = nullptr;            // [cap] L2: |x|^2, cosine: 1/|x|; +inf / 0 for deleted rows
  __half *d_vext = nullptr;          // [cap][16] L2 only: -|x 2^sx|^2 2^-11 in three fp16 pieces (extra K columns of the
                                     // tensor path: the per-row term of the score comes out of the GEMM itself)
  int sx_exp = 0;                    // power-of-two scale of d_vech: chosen so that every component fits fp16
  bool have_scale = false;
  std::vector<uint8_t> h_alive;
  int sm_count = 148;
  size_t smem_optin = 0;
  cudaStream_t stream = nullptr;
  void *d_scratch = nullptr;
  size_t scratch_bytes = 0;
  void *tmap_x = nullptr;            // CUtensorMap of d_vech, then of d_vext (exact_gemm.cu), rebuilt when storage moves
  int64_t tmap_rows = -1;
  uint32_t *d_maxn2 = nullptr;       // [0] bits of max |x|^2, [1] bits of min |x|^2 > 0, [2] bits of max |x_i| over rows ever added
  int64_t aux_dead_cap = -1;         // capacity up to which d_aux of unused rows holds the dead value
  void *d_gemm_ws = nullptr;         // query-side workspace of the tensor path
  size_t gemm_ws_bytes = 0;
  std::mutex mu;                     // searches share the scratch buffer: one at a time per index
  // recorded on the caller's stream behind the last kernel of a search; the next search (possibly on another stream)
  // waits for it before it touches d_scratch / d_gemm_ws, so two device-pointer calls in flight never share them
  cudaEvent_t scratch_ev = nullptr;
  int reserve(size_t bytes);
  int ensure_capacity(int64_t rows);
};

namespace b200 {
// exact_gemm.cu: fills cand[q][0..cand_cap) / cand_cnt[q] with rows that provably contain the k nearest of every
// query; returns B200HNSW_ERR_UNSUPPORTED when the tensor path cannot be used (caller falls back to the scan).
int exact_candidates_tensor(b200exact *x, const float *d_queries, int64_t B, int k, int32_t *d_cand,
                            int32_t *d_cand_cnt, int cand_cap, cudaStream_t s);
int exact_after_add(b200exact *x, int64_t first, int64_t count, cudaStream_t s);  // fp16 copy + norms of new rows (synchronises the stream)
}  // namespace b200
