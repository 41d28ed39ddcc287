// index.hpp -- host-side object behind b200hnsw_t: owns the device-resident graph (vector slab + fixed-stride
// neighbour lists), a pool of per-call contexts (stream, events, scratch) and the error plumbing.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
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
  // visited-table entries per traversal; raised (sticky) only after a retry with the larger table has succeeded
  std::atomic<int> vis_hint{512};
  std::atomic<int> vis_hint_ext{0};  // the same for the extended mode (its walks visit ~10x the nodes; 0 = sized from ef)
  // bit 0: some stored row has |x|^2 = 0, non-finite or outside [1e-18, 1e18] (NaN distances possible: traverse.cuh).
  // d_row_flags is the device word the upload paths OR into (flag_rows); row_flags its host mirror, refreshed at the end
  // of every mutating call.
  // the speculative windowed Add (spec_insert.cu): 0 = window chosen from the row count, 1 = the sequential kernel,
  // >= 2 = that many nodes in flight (b200hnsw_set_add_window); d_spec = its device scratch
  int add_window = 0;
  void *d_spec = nullptr;
  size_t spec_bytes = 0;
  unsigned long long spec_accepted_redone[3] = {0, 0, 0};  // + repaired by workers
  int build_mode = 0;  // B200HNSW_MODE_*: the layer search the wave Add runs (b200hnsw_set_build_mode)
  int *d_row_flags = nullptr;
  std::atomic<int> row_flags{0};
  int flag_rows(int64_t first, int64_t count, cudaStream_t s);   // queue the scan of rows [first, first + count)
  int refresh_row_flags(cudaStream_t s);                         // synchronises s
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
                  int64_t seg_stride = 0, int32_t *const *seg_ptrs = nullptr, int n_seg_ptrs = 0, int full_queue = 0);
// largest visited table (entries, power of two) a traversal can be given on this device
int max_visited_cap(const b200hnsw *h, int k, int ef);
struct TraverseParams;
int launch_traverse(b200hnsw *h, TraverseParams &p, int vis_cap, cudaStream_t s);
// extended mode (traverse_ext.cuh): ef_upper = ef of the layers above 0 (1 for Search, EfSearch for Add)
int launch_traverse_ext(b200hnsw *h, TraverseParams &p, int ef_upper, int vis_cap, cudaStream_t s);
int ext_visited_default(int efk, int M);
int max_visited_cap_ext(const b200hnsw *h, int efk_max);
}  // namespace b200
