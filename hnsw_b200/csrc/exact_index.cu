// exact_index.cu -- K3: hybrid.ExactIndex (hnsw-extensions/hybrid/exact.go:13-160) on the device.
//
// ExactIndex.Search (exact.go:61-111) computes distance(query, vector) for every stored vector, sorts all of them
// ascending (stable: ties keep iteration order) and returns the first k.  Here:
//   * scan path (this file): every distance is evaluated in fp32 with the canonical order of dist.cuh, so values are
//     bit-identical to the oracle; one CTA per (query, chunk of rows) sorts its chunk's 64-bit keys
//     (order-preserving distance bits << 32 | row) in shared memory and emits the chunk's k smallest, then the
//     per-chunk lists are merged the same way.  Ties therefore keep row order; NaN distances sort last.
//     Bound: fp32 pipe / HBM (B * N * dim).  It is the exact fallback and the path for small indexes
//     (the reference's ExactThreshold is 1000 rows, hybrid.go:101).
//   * tensor path (exact_gemm.cu): a bf16 tcgen05 GEMM over all rows with a fused, provably safe threshold filter
//     produces a short candidate list per query; candidates are re-evaluated here in canonical fp32 and selected
//     with the same keys, so both paths return the same rows and bits.
#include <algorithm>
#include <cstring>
#include <new>
#include <vector>

#include "exact_index.hpp"

namespace b200 {

__device__ __forceinline__ uint32_t ordered_bits_f(float d) {
  uint32_t u = __float_as_uint(d);
  if (d != d) return 0xffffffffu;
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float unordered_bits_f(uint32_t k) {
  if (k == 0xffffffffu) return __int_as_float(0x7fc00000);
  return __uint_as_float((k & 0x80000000u) ? (k ^ 0x80000000u) : ~k);
}

// in-place ascending bitonic sort of n_pow2 64-bit keys in shared memory (all threads of the block)
__device__ __forceinline__ void block_bitonic_sort(unsigned long long *keys, int n_pow2) {
  for (int size = 2; size <= n_pow2; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = threadIdx.x; i < n_pow2 / 2; i += blockDim.x) {
        int lo = 2 * i - (i & (stride - 1));
        int hi = lo + stride;
        bool up = (lo & size) == 0;
        unsigned long long a = keys[lo], b = keys[hi];
        if ((a > b) == up) {
          keys[lo] = b;
          keys[hi] = a;
        }
      }
      __syncthreads();
    }
  }
}

constexpr int SCAN_ROWS = 2048;  // rows per CTA
constexpr int SCAN_THREADS = 256;

// grid (chunks, B).  partial[q][chunk][k] <- the chunk's k smallest keys (~0 = none)
template <int CH, int U>
__global__ void __launch_bounds__(SCAN_THREADS) exact_scan_kernel(const float4 *vec, const uint8_t *alive,
                                                                  long long n, int row_chunks, int dim,
                                                                  const float *queries, int metric, int k,
                                                                  unsigned long long *partial, int n_chunks) {
  __shared__ unsigned long long keys[SCAN_ROWS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = SCAN_THREADS / 32;
  const long long q = blockIdx.y;
  const long long base = (long long)blockIdx.x * SCAN_ROWS;
  GQuery<CH> Q;
  group_load_query<32, CH>(Q, queries + q * dim, true, dim, row_chunks, lane);
  for (int r0 = warp * U; r0 < SCAN_ROWS; r0 += nw * U) {
    int ids[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      long long row = base + r0 + u;
      ids[u] = (row < n && alive[row]) ? (int)row : -1;
    }
    // exact.go:79: distance(query, vector); the canonical arithmetic is symmetric in its arguments bit for bit
    const float dm = group_distances<32, CH, U>(vec, row_chunks, ids, Q, metric, lane);
    if ((lane & (32 / U - 1)) == 0) {
      const int u = lane / (32 / U);
      keys[r0 + u] = ids[u] >= 0 ? (((unsigned long long)ordered_bits_f(dm) << 32) | (unsigned)ids[u]) : ~0ull;
    }
  }
  __syncthreads();
  block_bitonic_sort(keys, SCAN_ROWS);
  unsigned long long *out = partial + (q * n_chunks + blockIdx.x) * k;
  for (int i = threadIdx.x; i < k; i += blockDim.x) out[i] = i < SCAN_ROWS ? keys[i] : ~0ull;
}

// Re-evaluation of candidate rows (tensor path): one CTA per query; cand[q][0..cnt) rows -> keys -> sorted -> top k
template <int CH, int U>
__global__ void __launch_bounds__(SCAN_THREADS) exact_rerank_kernel(const float4 *vec, int row_chunks, int dim,
                                                                    const float *queries, int metric, int k,
                                                                    const uint8_t *alive, const int32_t *cand,
                                                                    const int32_t *cand_cnt,
                                                                    int cand_cap, int n_pow2,
                                                                    unsigned long long *out_keys) {
  extern __shared__ unsigned long long skeys[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = SCAN_THREADS / 32;
  const long long q = blockIdx.x;
  int cnt = cand_cnt[q];
  if (cnt > cand_cap) cnt = cand_cap;  // overflowed queries are redone by the scan path (host checks the count)
  {  // sort only as many keys as this query has (n_pow2 is the capacity the launch reserved)
    int need = 2;
    while (need < cnt || need < k) need <<= 1;
    n_pow2 = min(n_pow2, need);
  }
  for (int i = threadIdx.x; i < n_pow2; i += blockDim.x) skeys[i] = ~0ull;
  __syncthreads();
  GQuery<CH> Q;
  group_load_query<32, CH>(Q, queries + q * dim, true, dim, row_chunks, lane);
  const int32_t *cq = cand + q * cand_cap;
  for (int r0 = warp * U; r0 < cnt; r0 += nw * U) {
    int ids[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      int id = r0 + u < cnt ? cq[r0 + u] : -1;
      ids[u] = (id >= 0 && alive[id]) ? id : -1;
    }
    const float dm = group_distances<32, CH, U>(vec, row_chunks, ids, Q, metric, lane);
    if ((lane & (32 / U - 1)) == 0) {
      const int u = lane / (32 / U);
      if (ids[u] >= 0) skeys[r0 + u] = ((unsigned long long)ordered_bits_f(dm) << 32) | (unsigned)ids[u];
    }
  }
  __syncthreads();
  block_bitonic_sort(skeys, n_pow2);
  for (int i = threadIdx.x; i < k; i += blockDim.x) out_keys[q * k + i] = i < n_pow2 ? skeys[i] : ~0ull;
}

// merge `lists` sorted-or-not key lists of length k per query into one: in[q][lists][k] -> out[q][out_lists][k],
// list group g takes input lists [g*group, (g+1)*group)
__global__ void exact_merge_kernel(const unsigned long long *in, int lists, int k, int group, int n_pow2,
                                   unsigned long long *out, int out_lists) {
  extern __shared__ unsigned long long skeys[];
  const long long q = blockIdx.y;
  const int g = blockIdx.x;
  const int first = g * group;
  const int cnt = min(group, lists - first) * k;
  const unsigned long long *src = in + (q * lists + first) * k;
  for (int i = threadIdx.x; i < n_pow2; i += blockDim.x) skeys[i] = i < cnt ? src[i] : ~0ull;
  __syncthreads();
  block_bitonic_sort(skeys, n_pow2);
  unsigned long long *dst = out + (q * out_lists + g) * k;
  for (int i = threadIdx.x; i < k; i += blockDim.x) dst[i] = skeys[i];
}

__global__ void exact_emit_kernel(const unsigned long long *keys, long long B, int k, long long live, int64_t *rows,
                                  float *dists, int32_t *counts) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * k) return;
  unsigned long long key = keys[i];
  bool ok = key != ~0ull;
  rows[i] = ok ? (int64_t)(key & 0xffffffffu) : -1;
  dists[i] = ok ? unordered_bits_f((uint32_t)(key >> 32)) : __int_as_float(0x7fc00000);
  if (i % k == 0) counts[i / k] = (int32_t)(live < k ? live : k);  // exact.go:97-100
}

}  // namespace b200

using namespace b200;

int b200exact::reserve(size_t bytes) {
  if (bytes <= scratch_bytes) return 0;
  if (d_scratch) cudaFree(d_scratch);
  d_scratch = nullptr;
  scratch_bytes = 0;
  size_t want = std::max(bytes, (size_t)1 << 20);
  B200_CUDA(cudaMalloc(&d_scratch, want));
  scratch_bytes = want;
  return 0;
}

template <typename T>
static int grow_arr(T *&ptr, int64_t old_elems, int64_t new_elems, cudaStream_t s) {
  T *np = nullptr;
  B200_CUDA(cudaMalloc(&np, sizeof(T) * (size_t)new_elems));
  B200_CUDA(cudaMemsetAsync(np, 0, sizeof(T) * (size_t)new_elems, s));
  if (ptr && old_elems > 0)
    B200_CUDA(cudaMemcpyAsync(np, ptr, sizeof(T) * (size_t)old_elems, cudaMemcpyDeviceToDevice, s));
  B200_CUDA(cudaStreamSynchronize(s));
  if (ptr) cudaFree(ptr);
  ptr = np;
  return 0;
}

int b200exact::ensure_capacity(int64_t rows) {
  if (rows <= cap) return 0;
  int64_t nc = std::max<int64_t>(cap * 2, std::max<int64_t>(rows, 1024));
  nc = (nc + 255) / 256 * 256;  // whole tiles for the tensor path
  int rc;
  if ((rc = grow_arr(d_vec, n * row_floats, nc * row_floats, stream))) return rc;
  if ((rc = grow_arr(d_alive, n, nc, stream))) return rc;
  if ((rc = grow_arr(d_vech, n * kpad, nc * kpad, stream))) return rc;
  if ((rc = grow_arr(d_aux, n, nc, stream))) return rc;
  if (metric == B200HNSW_EUCLIDEAN && (rc = grow_arr(d_vext, n * 16, nc * 16, stream))) return rc;
  cap = nc;
  tmap_rows = -1;
  return 0;
}

template <int CH, int U>
static int run_scan(b200exact *x, const float *d_q, int64_t B, int k, unsigned long long *partial, int n_chunks,
                    cudaStream_t s) {
  dim3 grid((unsigned)n_chunks, (unsigned)B);
  exact_scan_kernel<CH, U><<<grid, SCAN_THREADS, 0, s>>>(reinterpret_cast<const float4 *>(x->d_vec), x->d_alive,
                                                          x->n, x->row_floats / 4, x->dim, d_q, x->metric, k, partial,
                                                          n_chunks);
  B200_CUDA(cudaGetLastError());
  return 0;
}
template <int CH, int U>
static int run_rerank(b200exact *x, const float *d_q, int64_t B, int k, const int32_t *cand, const int32_t *cnt,
                      int cand_cap, unsigned long long *out, cudaStream_t s) {
  int n_pow2 = 2;
  while (n_pow2 < cand_cap || n_pow2 < k) n_pow2 <<= 1;
  size_t smem = sizeof(unsigned long long) * (size_t)n_pow2;
  auto kern = exact_rerank_kernel<CH, U>;
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)B, SCAN_THREADS, smem, s>>>(reinterpret_cast<const float4 *>(x->d_vec), x->row_floats / 4, x->dim,
                                               d_q, x->metric, k, x->d_alive, cand, cnt, cand_cap, n_pow2, out);
  B200_CUDA(cudaGetLastError());
  return 0;
}

#define EXACT_DISPATCH(fn, ...)                              \
  do {                                                       \
    int ch__ = (x->row_floats / 4 + 31) / 32;                \
    if (ch__ <= 1) return fn<1, 8>(__VA_ARGS__);             \
    if (ch__ <= 2) return fn<2, 4>(__VA_ARGS__);             \
    if (ch__ <= 4) return fn<4, 4>(__VA_ARGS__);             \
    if (ch__ <= 6) return fn<6, 2>(__VA_ARGS__);             \
    if (ch__ <= 8) return fn<8, 2>(__VA_ARGS__);             \
    if (ch__ <= 12) return fn<12, 1>(__VA_ARGS__);           \
    if (ch__ <= 16) return fn<16, 1>(__VA_ARGS__);           \
    set_error("exact: dim %d > 2048 not supported", x->dim); \
    return B200HNSW_ERR_UNSUPPORTED;                         \
  } while (0)

static int scan_any(b200exact *x, const float *d_q, int64_t B, int k, unsigned long long *partial, int n_chunks,
                    cudaStream_t s) {
  EXACT_DISPATCH(run_scan, x, d_q, B, k, partial, n_chunks, s);
}
static int rerank_any(b200exact *x, const float *d_q, int64_t B, int k, const int32_t *cand, const int32_t *cnt,
                      int cand_cap, unsigned long long *out, cudaStream_t s) {
  EXACT_DISPATCH(run_rerank, x, d_q, B, k, cand, cnt, cand_cap, out, s);
}

// scan path for queries [0, B): keys of the k nearest -> final_keys[B][k].  Uses scratch at `ws` (ws_bytes).
static int scan_topk(b200exact *x, const float *d_q, int64_t B, int k, unsigned long long *final_keys, char *ws,
                     size_t ws_bytes, cudaStream_t s) {
  const int n_chunks = (int)((x->n + SCAN_ROWS - 1) / SCAN_ROWS);
  const int group = std::max(2, 4096 / k);  // lists merged per CTA (group * k <= 4096 keys = 32 KB)
  // queries are processed in slices so that the partial lists fit the workspace
  size_t per_q = sizeof(unsigned long long) * (size_t)k * ((size_t)n_chunks + (n_chunks + group - 1) / group + 2);
  int64_t slice = std::max<int64_t>(1, std::min<int64_t>(B, (int64_t)(ws_bytes / per_q)));
  slice = std::min<int64_t>(slice, 32768);
  for (int64_t q0 = 0; q0 < B; q0 += slice) {
    int64_t nq = std::min(slice, B - q0);
    unsigned long long *a = (unsigned long long *)ws;
    unsigned long long *b = a + (size_t)nq * k * n_chunks;
    int rc = scan_any(x, d_q + q0 * x->dim, nq, k, a, n_chunks, s);
    if (rc) return rc;
    int lists = n_chunks;
    while (lists > 1) {
      int out_lists = (lists + group - 1) / group;
      int n_pow2 = 2;
      while (n_pow2 < group * k) n_pow2 <<= 1;
      dim3 grid((unsigned)out_lists, (unsigned)nq);
      exact_merge_kernel<<<grid, 256, sizeof(unsigned long long) * (size_t)n_pow2, s>>>(a, lists, k, group, n_pow2,
                                                                                         b, out_lists);
      B200_CUDA(cudaGetLastError());
      std::swap(a, b);
      lists = out_lists;
    }
    B200_CUDA(cudaMemcpyAsync(final_keys + q0 * k, a, sizeof(unsigned long long) * (size_t)nq * k,
                              cudaMemcpyDeviceToDevice, s));
  }
  return 0;
}

// ExactIndex.Search for B queries resident on the device; results to device buffers.  Synchronises `s` only when the
// tensor path has to check for candidate-list overflow.
static thread_local int64_t g_last_cands = 0, g_last_fallbacks = 0;
static thread_local int g_last_path = 0;  // 1 = scan, 2 = tensor

static int exact_search_core(b200exact *x, const float *d_q, int64_t B, int k, int64_t *d_rows, float *d_dists,
                             int32_t *d_counts, cudaStream_t s) {
  g_last_cands = g_last_fallbacks = 0;
  g_last_path = 0;
  // the scratch (and the tensor path's query workspace) belongs to one search at a time: order this call behind the
  // previous one on the device, whatever stream that one ran on
  if (!x->scratch_ev) B200_CUDA(cudaEventCreateWithFlags(&x->scratch_ev, cudaEventDisableTiming));
  else B200_CUDA(cudaStreamWaitEvent(s, x->scratch_ev, 0));
  struct Release {
    b200exact *x;
    cudaStream_t s;
    ~Release() { cudaEventRecord(x->scratch_ev, s); }
  } release{x, s};
  if (x->live == 0 || x->n == 0) {  // exact.go:65-67
    B200_CUDA(cudaMemsetAsync(d_counts, 0, sizeof(int32_t) * (size_t)B, s));
    B200_CUDA(cudaMemsetAsync(d_rows, 0xff, sizeof(int64_t) * (size_t)B * k, s));
    B200_CUDA(cudaMemsetAsync(d_dists, 0xff, sizeof(float) * (size_t)B * k, s));
    return 0;
  }
  auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
  const char *force = getenv("B200EXACT_PATH");
  bool want_tensor = x->n >= 16384 && (double)x->n * (double)B >= 2e6;  // small problems: the scan is cheaper
  if (force && !strcmp(force, "scan")) want_tensor = false;
  if (force && !strcmp(force, "tensor")) want_tensor = true;
  const int cand_cap = 4096;
  const size_t keys_b = al(sizeof(unsigned long long) * (size_t)B * k);
  const size_t cand_b = want_tensor ? al(4 * (size_t)B * cand_cap) + al(4 * (size_t)B) : 0;
  const int n_chunks = (int)((x->n + SCAN_ROWS - 1) / SCAN_ROWS);
  // scan workspace: enough for a slice of queries (at least one)
  size_t ws_b = al(sizeof(unsigned long long) * (size_t)k * ((size_t)n_chunks * 2 + 4) *
                   (size_t)std::min<int64_t>(B, want_tensor ? 64 : 4096));
  int rc = x->reserve(keys_b + cand_b + ws_b + 1024);
  if (rc) return rc;
  char *p = (char *)x->d_scratch;
  unsigned long long *d_keys = (unsigned long long *)p;
  p += keys_b;
  int32_t *d_cand = nullptr, *d_cnt = nullptr;
  if (want_tensor) {
    d_cand = (int32_t *)p;
    p += al(4 * (size_t)B * cand_cap);
    d_cnt = (int32_t *)p;
    p += al(4 * (size_t)B);
  }
  char *ws = p;
  bool done = false;
  if (want_tensor) {
    rc = exact_candidates_tensor(x, d_q, B, k, d_cand, d_cnt, cand_cap, s);
    if (rc == 0) {
      if ((rc = rerank_any(x, d_q, B, k, d_cand, d_cnt, cand_cap, d_keys, s))) return rc;
      // queries whose candidate list overflowed are redone exactly by the scan (never truncated silently)
      std::vector<int32_t> cnt((size_t)B);
      B200_CUDA(cudaMemcpyAsync(cnt.data(), d_cnt, 4 * (size_t)B, cudaMemcpyDeviceToHost, s));
      B200_CUDA(cudaStreamSynchronize(s));
      for (int64_t q = 0; q < B; q++) {
        g_last_cands += cnt[q];
        if (cnt[q] > cand_cap || cnt[q] < std::min<int64_t>(k, x->live)) {
          g_last_fallbacks++;
          if ((rc = scan_topk(x, d_q + q * x->dim, 1, k, d_keys + q * k, ws, ws_b, s))) return rc;
        }
      }
      g_last_path = 2;
      done = true;
    } else if (rc != B200HNSW_ERR_UNSUPPORTED || (force && !strcmp(force, "tensor"))) {
      return rc;
    }
  }
  if (!done) {
    if ((rc = scan_topk(x, d_q, B, k, d_keys, ws, ws_b, s))) return rc;
    g_last_path = 1;
  }
  long long total = (long long)B * k;
  exact_emit_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(d_keys, B, k, x->live, d_rows, d_dists, d_counts);
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" {

int b200exact_create(int32_t dim, int32_t metric, int32_t device, int64_t capacity, b200exact_t **out) {
  if (!out || dim <= 0 || dim > 2048 || (metric != B200HNSW_EUCLIDEAN && metric != B200HNSW_COSINE)) {
    set_error("exact_create: bad argument (dim=%d metric=%d)", dim, metric);
    return B200HNSW_ERR_BAD_ARG;
  }
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    set_error("exact_create: no CUDA device (%s); this library has no CPU fallback", cudaGetErrorString(e));
    return B200HNSW_ERR_CUDA;
  }
  if (device < 0 || device >= ndev) {
    set_error("exact_create: device %d out of range", device);
    return B200HNSW_ERR_BAD_ARG;
  }
  B200_CUDA(cudaSetDevice(device));
  b200exact *x = new (std::nothrow) b200exact();
  if (!x) return B200HNSW_ERR_OOM;
  x->dim = dim;
  x->metric = metric;
  x->device = device;
  x->row_floats = (dim + 31) / 32 * 32;
  x->kpad = (dim + 63) / 64 * 64;
  cudaDeviceProp prop;
  B200_CUDA(cudaGetDeviceProperties(&prop, device));
  x->sm_count = prop.multiProcessorCount;
  x->smem_optin = prop.sharedMemPerBlockOptin;
  if (cudaStreamCreateWithFlags(&x->stream, cudaStreamNonBlocking) != cudaSuccess) {
    delete x;
    set_error("exact_create: cannot create stream");
    return B200HNSW_ERR_CUDA;
  }
  int rc = x->ensure_capacity(capacity > 0 ? capacity : 1024);
  if (rc) {
    b200exact_destroy(x);
    return rc;
  }
  *out = x;
  return 0;
}

void b200exact_destroy(b200exact_t *x) {
  if (!x) return;
  cudaSetDevice(x->device);
  cudaDeviceSynchronize();
  cudaFree(x->d_vec);
  cudaFree(x->d_alive);
  cudaFree(x->d_vech);
  cudaFree(x->d_aux);
  cudaFree(x->d_vext);
  cudaFree(x->d_scratch);
  cudaFree(x->d_maxn2);
  cudaFree(x->d_gemm_ws);
  if (x->scratch_ev) cudaEventDestroy(x->scratch_ev);
  free(x->tmap_x);
  if (x->stream) cudaStreamDestroy(x->stream);
  delete x;
}

int64_t b200exact_len(const b200exact_t *x) { return x ? x->live : 0; }

int b200exact_add_batch(b200exact_t *x, const float *vectors, int64_t n, int64_t *first_row) {
  if (!x || n < 0 || (n > 0 && !vectors)) {
    set_error("exact_add_batch: bad argument");
    return B200HNSW_ERR_BAD_ARG;
  }
  if (first_row) *first_row = x->n;
  if (n == 0) return 0;
  if (x->n + n > 0x7fffffffLL) {
    set_error("exact_add_batch: more than 2^31-1 rows");
    return B200HNSW_ERR_UNSUPPORTED;
  }
  B200_CUDA(cudaSetDevice(x->device));
  std::lock_guard<std::mutex> lk(x->mu);
  int rc = x->ensure_capacity(x->n + n);
  if (rc) return rc;
  cudaStream_t s = x->stream;
  // exact.go:32-36: the index keeps its own copy of every vector
  B200_CUDA(cudaMemcpy2DAsync(x->d_vec + (size_t)x->n * x->row_floats, sizeof(float) * x->row_floats, vectors,
                              sizeof(float) * x->dim, sizeof(float) * x->dim, (size_t)n, cudaMemcpyHostToDevice, s));
  B200_CUDA(cudaMemsetAsync(x->d_alive + x->n, 1, (size_t)n, s));
  if ((rc = exact_after_add(x, x->n, n, s))) return rc;
  B200_CUDA(cudaStreamSynchronize(s));
  x->h_alive.resize((size_t)(x->n + n), 1);
  x->n += n;
  x->live += n;
  return 0;
}

int b200exact_remove(b200exact_t *x, const int64_t *rows, int64_t n) {
  if (!x || n < 0 || (n > 0 && !rows)) {
    set_error("exact_remove: bad argument");
    return B200HNSW_ERR_BAD_ARG;
  }
  B200_CUDA(cudaSetDevice(x->device));
  std::lock_guard<std::mutex> lk(x->mu);
  for (int64_t i = 0; i < n; i++) {
    int64_t r = rows[i];
    if (r < 0 || r >= x->n || !x->h_alive[r]) continue;  // exact.go:118-121: deleting an absent key is a no-op
    x->h_alive[r] = 0;
    x->live--;
    B200_CUDA(cudaMemsetAsync(x->d_alive + r, 0, 1, x->stream));
    // the tensor path skips the row: |x|^2 = +inf (L2) / 1/|x| = NaN scores it out of every candidate list
    const float dead = x->metric == B200HNSW_EUCLIDEAN ? __builtin_inff() : __builtin_nanf("");
    B200_CUDA(cudaMemcpyAsync(x->d_aux + r, &dead, sizeof dead, cudaMemcpyHostToDevice, x->stream));
    if (x->d_vext) {  // the folded per-row term: -inf makes the accumulator -inf, i.e. the score +inf
      static const uint16_t dead_h[4] = {0xfc00u, 0, 0, 0};
      B200_CUDA(cudaMemcpyAsync(x->d_vext + r * 16, dead_h, sizeof dead_h, cudaMemcpyHostToDevice, x->stream));
    }
  }
  B200_CUDA(cudaStreamSynchronize(x->stream));
  return 0;
}

int b200exact_search_batch_device(b200exact_t *x, const float *d_queries, int64_t B, int32_t k, int64_t *d_out_rows,
                                  float *d_out_dists, int32_t *d_out_counts, void *stream) {
  if (!x || B < 0 || k <= 0 || k > B200HNSW_MAX_K || (B > 0 && (!d_queries || !d_out_rows || !d_out_dists ||
                                                                  !d_out_counts))) {
    set_error("exact_search: bad argument (k must be in 1..%d)", B200HNSW_MAX_K);
    return B200HNSW_ERR_BAD_ARG;
  }
  if (B == 0) return 0;
  B200_CUDA(cudaSetDevice(x->device));
  std::lock_guard<std::mutex> lk(x->mu);
  return exact_search_core(x, d_queries, B, k, d_out_rows, d_out_dists, d_out_counts, (cudaStream_t)stream);
}

int b200exact_search_batch(b200exact_t *x, const float *queries, int64_t B, int32_t k, int64_t *out_rows,
                           float *out_dists, int32_t *out_counts, b200hnsw_stats *stats) {
  if (stats) memset(stats, 0, sizeof *stats);
  if (!x || B < 0 || k <= 0 || k > B200HNSW_MAX_K || (B > 0 && (!queries || !out_rows || !out_dists || !out_counts))) {
    set_error("exact_search: bad argument (k must be in 1..%d)", B200HNSW_MAX_K);
    return B200HNSW_ERR_BAD_ARG;
  }
  if (B == 0) return 0;
  B200_CUDA(cudaSetDevice(x->device));
  std::lock_guard<std::mutex> lk(x->mu);
  cudaStream_t s = x->stream;
  float *d_q = nullptr;
  int64_t *d_rows = nullptr;
  float *d_dists = nullptr;
  int32_t *d_cnt = nullptr;
  const size_t qb = sizeof(float) * (size_t)B * x->dim;
  B200_CUDA(cudaMalloc(&d_q, qb + (sizeof(int64_t) + sizeof(float)) * (size_t)B * k + sizeof(int32_t) * (size_t)B + 64));
  d_rows = (int64_t *)((char *)d_q + ((qb + 15) & ~(size_t)15));
  d_dists = (float *)(d_rows + (size_t)B * k);
  d_cnt = (int32_t *)(d_dists + (size_t)B * k);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  int rc = 0;
  cudaError_t e;
  if ((e = cudaMemcpyAsync(d_q, queries, qb, cudaMemcpyHostToDevice, s)) != cudaSuccess) rc = cuda_fail(e, "exact H2D");
  cudaEventRecord(e0, s);
  if (!rc) rc = exact_search_core(x, d_q, B, k, d_rows, d_dists, d_cnt, s);
  cudaEventRecord(e1, s);
  if (!rc && ((e = cudaMemcpyAsync(out_rows, d_rows, sizeof(int64_t) * (size_t)B * k, cudaMemcpyDeviceToHost, s)) !=
                  cudaSuccess ||
              (e = cudaMemcpyAsync(out_dists, d_dists, sizeof(float) * (size_t)B * k, cudaMemcpyDeviceToHost, s)) !=
                  cudaSuccess ||
              (e = cudaMemcpyAsync(out_counts, d_cnt, sizeof(int32_t) * (size_t)B, cudaMemcpyDeviceToHost, s)) !=
                  cudaSuccess ||
              (e = cudaStreamSynchronize(s)) != cudaSuccess))
    rc = cuda_fail(e, "exact D2H");
  if (!rc && stats) {
    cudaEventElapsedTime(&stats->kernel_ms, e0, e1);
    stats->dist_evals = g_last_path == 2 ? g_last_cands : (int64_t)B * x->live;  // fp32 evaluations made
    stats->expansions = g_last_fallbacks;                                        // queries redone by the scan
    stats->launches = g_last_path;                                               // 1 = scan path, 2 = tensor path
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d_q);
  return rc;
}

}  // extern "C"

// ---- ExactIndex.Search without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that
// forbid it).  The scan path above with its workspace (final keys + per-chunk partial lists + merge ping-pong) in a
// file-scope device array instead of b200exact::d_scratch.  Every GPU that loads this library holds the array until
// the process exits.  In host code `g_noalloc_scratch` is only a host-side stand-in for the symbol: the device address
// comes from cudaGetSymbolAddress, once per device, with that device current.
// Nothing in the workspace has to start at zero: the scan kernel writes all k keys of every (query, chunk) list and
// every merge level writes all of its output before the next level reads it, so the previous call's contents are never
// read.  All calls (all indexes, all streams, all host threads) share the one array per GPU: calls that overlap on the
// GPU race.
constexpr size_t NOALLOC_SCRATCH_BYTES = (size_t)8 << 20;
constexpr int NOALLOC_MAX_DEVICES = 64;
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];
static std::mutex g_noalloc_mu;
static char *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

// device address of the array on `device` (which is current)
static int noalloc_scratch_of(int device, char **out) {
  if (device < 0 || device >= NOALLOC_MAX_DEVICES) {
    set_error("exact_search_noalloc: device %d out of range (0..%d)", device, NOALLOC_MAX_DEVICES - 1);
    return B200HNSW_ERR_BAD_ARG;
  }
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[device]) {
    void *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_noalloc_scratch)");
    if (!p) {
      set_error("exact_search_noalloc: cudaGetSymbolAddress returned a null address on device %d", device);
      return B200HNSW_ERR_CUDA;
    }
    g_noalloc_ptr[device] = (char *)p;
  }
  *out = g_noalloc_ptr[device];
  return 0;
}

extern "C" int b200exact_search_batch_noalloc_device(b200exact_t *x, const float *d_queries, int64_t B, int32_t k,
                                                     int64_t *d_out_rows, float *d_out_dists, int32_t *d_out_counts,
                                                     void *stream) {
  if (!x || B < 0 || k <= 0 || k > B200HNSW_MAX_K || (B > 0 && (!d_queries || !d_out_rows || !d_out_dists ||
                                                                  !d_out_counts))) {
    set_error("exact_search_noalloc: bad argument (k must be in 1..%d)", B200HNSW_MAX_K);
    return B200HNSW_ERR_BAD_ARG;
  }
  B200_CUDA(cudaSetDevice(x->device));
  char *scratch = nullptr;
  int rc = noalloc_scratch_of(x->device, &scratch);  // B == 0 is the set-up call: resolves the address and returns
  if (rc) return rc;
  if (B == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  std::lock_guard<std::mutex> lk(x->mu);
  if (x->live == 0 || x->n == 0) {  // exact.go:65-67
    B200_CUDA(cudaMemsetAsync(d_out_counts, 0, sizeof(int32_t) * (size_t)B, s));
    B200_CUDA(cudaMemsetAsync(d_out_rows, 0xff, sizeof(int64_t) * (size_t)B * k, s));
    B200_CUDA(cudaMemsetAsync(d_out_dists, 0xff, sizeof(float) * (size_t)B * k, s));
    return 0;
  }
  // final keys [B][k] in front, the scan's workspace behind them; the scan slices the queries to fit, so what has to
  // fit is the final keys plus ONE query's partial lists (the formula of scan_topk)
  const size_t n_chunks = (size_t)((x->n + SCAN_ROWS - 1) / SCAN_ROWS);
  const size_t group = (size_t)std::max(2, 4096 / k);
  const size_t keys_b = (sizeof(unsigned long long) * (size_t)B * k + 255) & ~(size_t)255;
  const size_t per_q = sizeof(unsigned long long) * (size_t)k * (n_chunks + (n_chunks + group - 1) / group + 2);
  if (keys_b > NOALLOC_SCRATCH_BYTES || per_q > NOALLOC_SCRATCH_BYTES - keys_b) {
    set_error("exact_search_noalloc: B=%lld k=%d over %lld rows needs %zu bytes of scratch, the static array holds %zu "
              "(8*k*(B + rows/2048*(1 + 1/max(2, 4096/k)) + 2) must fit); split the batch or use "
              "b200exact_search_batch_device",
              (long long)B, k, (long long)x->n, keys_b + per_q, NOALLOC_SCRATCH_BYTES);
    return B200HNSW_ERR_UNSUPPORTED;
  }
  unsigned long long *d_keys = (unsigned long long *)scratch;
  if ((rc = scan_topk(x, d_queries, B, k, d_keys, scratch + keys_b, NOALLOC_SCRATCH_BYTES - keys_b, s))) return rc;
  const long long total = (long long)B * k;
  exact_emit_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(d_keys, B, k, x->live, d_out_rows, d_out_dists,
                                                                   d_out_counts);
  B200_CUDA(cudaGetLastError());
  return 0;
}
