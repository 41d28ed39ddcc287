// index.hpp -- host-side object behind b200hnsw_t: owns the device-resident graph (vector slab + fixed-stride
// neighbour lists), a pool of per-call contexts (stream, events, scratch) and the error plumbing.
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/b200hnsw.h"
#include "layer_search.cuh"

namespace b200 {

void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what);

#define B200_CUDA(call)                                   \
  do {                                                    \
    cudaError_t e__ = (call);                             \
    if (e__ != cudaSuccess) return cuda_fail(e__, #call); \
  } while (0)

// per-call execution context: search calls may come from several host threads at once
struct CallCtx {
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  void *d_scratch = nullptr;   // device staging (queries, outputs, counters)
  size_t scratch_bytes = 0;
  void *h_pinned = nullptr;    // pinned host staging
  size_t pinned_bytes = 0;
  static constexpr int MAX_SUB = 8;
  cudaStream_t sub[MAX_SUB] = {};   // chunk streams of a host-pointer search (copy / kernel overlap)
  cudaEvent_t sub_ev[MAX_SUB] = {};
  int n_sub = 0;
  int ensure_sub(int n);
  int reserve_device(size_t bytes);
  int reserve_pinned(size_t bytes);
};

struct DeviceGraph {
  float *vec = nullptr;         // [cap][row_floats]
  int64_t *keys = nullptr;      // [cap]
  int32_t *adj0 = nullptr;      // [cap][M]
  int32_t *upper_off = nullptr; // [cap]
  int32_t *adjU = nullptr;      // [capU][M]
  int64_t cap = 0, capU = 0;
  uint8_t *dead = nullptr;      // [cap] 1 = removed from its layers by Delete (graph.go:843-866); row and lists stay
  int32_t *inbox = nullptr;     // 4 x inbox_n: request counts/segments and eviction heads/owners per list slot (wave Add)
  int64_t inbox_n = 0;
};

}  // namespace b200

struct b200hnsw {
  b200hnsw_config cfg{};
  int row_floats = 0;  // dim padded to a multiple of 32 floats (128 B)
  int M = 0;           // list stride (= cfg.max_m)
  b200::DeviceGraph d;
  int64_t n = 0, nU = 0;
  int32_t n_layers = 0;
  int32_t entry[B200HNSW_MAX_LAYERS];
  int64_t layer_size[B200HNSW_MAX_LAYERS];
  // host mirrors of the small per-row metadata (the host decides slots; the device never allocates)
  std::vector<int64_t> h_keys;
  std::vector<int32_t> h_level;
  std::vector<int32_t> h_upper_off;
  std::vector<uint8_t> h_dead;
  int64_t n_dead = 0;
  int *d_err_word = nullptr;  // error flag for device-pointer calls that pass no counters
  int vis_hint = 512;  // visited-table entries per traversal; raised (sticky) when a search overflows it
  int sm_count = 148;
  size_t smem_optin = 0;
  std::mutex pool_mu;
  std::vector<b200::CallCtx *> pool;

  b200::CallCtx *acquire();
  void release(b200::CallCtx *c);
  int ensure_capacity(int64_t rows, int64_t upper_slots, cudaStream_t s);
  int ensure_inboxes(cudaStream_t s);
  b200::GraphView view() const;
  // layers that still have a member: Delete can empty the top ones, which Search then passes through without an
  // elevator (graph.go:571-583), i.e. exactly as if they were not there
  int32_t live_layers() const {
    int32_t l = n_layers;
    while (l > 0 && layer_size[l - 1] == 0) l--;
    return l;
  }
};

namespace b200 {
// launchers implemented in the .cu files
int launch_search(b200hnsw *h, const float *d_queries, int64_t B, int k, int ef, int metric, int mode,
                  const int32_t *entries, int32_t *d_rows, float *d_dists, int32_t *d_counts,
                  unsigned long long *d_counters, int *d_error, int vis_cap, cudaStream_t s, int64_t seg_len = 0,
                  int64_t seg_stride = 0, int32_t *const *seg_ptrs = nullptr, int n_seg_ptrs = 0);
struct TraverseParams;
int launch_traverse(b200hnsw *h, TraverseParams &p, int vis_cap, cudaStream_t s);
}  // namespace b200
