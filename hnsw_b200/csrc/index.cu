// index.cu -- C ABI (include/b200hnsw.h): lifetime, device-resident layout, load / read-back, search entry points.
#include <chrono>
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <new>

#include <sched.h>

#include "index.hpp"

namespace b200 {

static thread_local std::string g_err;

void set_error(const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
}

int cuda_fail(cudaError_t e, const char *what) {
  set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
  return e == cudaErrorMemoryAllocation ? B200HNSW_ERR_OOM : B200HNSW_ERR_CUDA;
}

int CallCtx::reserve_device(size_t bytes) {
  if (bytes <= scratch_bytes) return 0;
  if (d_scratch) cudaFree(d_scratch);
  d_scratch = nullptr;
  scratch_bytes = 0;
  size_t want = std::max(bytes, (size_t)1 << 20);
  B200_CUDA(cudaMalloc(&d_scratch, want));
  scratch_bytes = want;
  return 0;
}
int CallCtx::ensure_sub(int n) {
  while (n_sub < n && n_sub < MAX_SUB) {
    B200_CUDA(cudaStreamCreateWithFlags(&sub[n_sub], cudaStreamNonBlocking));
    B200_CUDA(cudaEventCreateWithFlags(&sub_ev[n_sub], cudaEventDisableTiming));
    n_sub++;
  }
  return 0;
}
int CallCtx::reserve_pinned(size_t bytes) {
  if (bytes <= pinned_bytes) return 0;
  if (h_pinned) cudaFreeHost(h_pinned);
  h_pinned = nullptr;
  pinned_bytes = 0;
  size_t want = std::max(bytes, (size_t)1 << 20);
  B200_CUDA(cudaMallocHost(&h_pinned, want));
  pinned_bytes = want;
  return 0;
}

template <typename T>
static int grow(T *&ptr, int64_t old_elems, int64_t new_elems, cudaStream_t s, int fill_byte = -1) {
  T *np = nullptr;
  B200_CUDA(cudaMalloc(&np, sizeof(T) * (size_t)new_elems));
  if (fill_byte >= 0) B200_CUDA(cudaMemsetAsync(np, fill_byte, sizeof(T) * (size_t)new_elems, s));
  if (ptr && old_elems > 0)
    B200_CUDA(cudaMemcpyAsync(np, ptr, sizeof(T) * (size_t)old_elems, cudaMemcpyDeviceToDevice, s));
  B200_CUDA(cudaStreamSynchronize(s));
  if (ptr) cudaFree(ptr);
  ptr = np;
  return 0;
}

}  // namespace b200

using namespace b200;

b200::CallCtx *b200hnsw::acquire() {
  {
    std::lock_guard<std::mutex> lk(pool_mu);
    if (!pool.empty()) {
      CallCtx *c = pool.back();
      pool.pop_back();
      return c;
    }
  }
  CallCtx *c = new (std::nothrow) CallCtx();
  if (!c) return nullptr;
  if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess) {
    delete c;
    return nullptr;
  }
  return c;
}
void b200hnsw::release(CallCtx *c) {
  std::lock_guard<std::mutex> lk(pool_mu);
  pool.push_back(c);
}

b200::GraphView b200hnsw::view() const {
  GraphView g;
  g.vec = reinterpret_cast<const float4 *>(d.vec);
  g.adj0 = d.adj0;
  g.adjU = d.adjU;
  g.upper_off = d.upper_off;
  g.row_chunks = row_floats / 4;
  g.M = M;
  g.dim = cfg.dim;
  return g;
}

int b200hnsw::ensure_capacity(int64_t rows, int64_t upper_slots, cudaStream_t s) {
  if (rows > d.cap) {
    int64_t nc = std::max<int64_t>(d.cap * 2, std::max<int64_t>(rows, 1024));
    int rc;
    // vectors are zero-filled so the padding lanes of every row are exact zeros
    if ((rc = grow(d.vec, n * row_floats, nc * row_floats, s, 0))) return rc;
    if ((rc = grow(d.keys, n, nc, s))) return rc;
    if ((rc = grow(d.adj0, n * M, nc * M, s, 0xff))) return rc;
    if ((rc = grow(d.upper_off, n, nc, s, 0xff))) return rc;
    if ((rc = grow(d.dead, n, nc, s, 0))) return rc;
    d.cap = nc;
  }
  if (upper_slots > d.capU) {
    int64_t nc = std::max<int64_t>(d.capU * 2, std::max<int64_t>(upper_slots, 1024));
    int rc;
    if ((rc = grow(d.adjU, nU * M, nc * M, s, 0xff))) return rc;
    d.capU = nc;
  }
  return 0;
}

int b200hnsw::ensure_inboxes(cudaStream_t s) {
  const int64_t want = d.cap + d.capU;
  if (d.inbox && d.inbox_n == want) return 0;  // inboxes are left empty by every wave
  if (d.inbox) cudaFree(d.inbox);
  d.inbox = nullptr;
  d.inbox_n = 0;
  B200_CUDA(cudaMalloc(&d.inbox, sizeof(int32_t) * 4 * (size_t)want));
  // request counts = 0 (their segment bases need no initial value); eviction heads = -1 (empty list), owners =
  // 0x7f7f7f7f (larger than any record id)
  B200_CUDA(cudaMemsetAsync(d.inbox, 0, sizeof(int32_t) * 2 * (size_t)want, s));
  B200_CUDA(cudaMemsetAsync(d.inbox + 2 * want, 0xff, sizeof(int32_t) * (size_t)want, s));
  B200_CUDA(cudaMemsetAsync(d.inbox + 3 * want, 0x7f, sizeof(int32_t) * (size_t)want, s));
  d.inbox_n = want;
  return 0;
}

// Wait for a stream the way a caller thread among many should: cudaStreamSynchronize spins on its core until the
// stream is done, so more waiting callers than cores (8 ranks x 4 callers on a 32-core box) only take cycles from
// each other and from the threads that still have calls to queue.  Poll instead, and hand the core over between polls
// once the first few (the common case of a short wait) have come back "not ready".
static cudaError_t wait_stream(cudaStream_t s) {
  for (int spins = 0;; spins++) {
    cudaError_t e = cudaStreamQuery(s);
    if (e != cudaErrorNotReady) return e;
    if (spins >= 64) sched_yield();
  }
}

extern "C" {

int b200hnsw_abi_version(void) { return B200HNSW_ABI_VERSION; }
const char *b200hnsw_last_error(void) { return g_err.c_str(); }

int b200hnsw_create(const b200hnsw_config *cfg, b200hnsw_t **out) {
  if (!cfg || !out) {
    set_error("create: null argument");
    return B200HNSW_ERR_BAD_ARG;
  }
  *out = nullptr;
  if (cfg->dim <= 0 || cfg->max_m <= 0 || cfg->max_m > B200HNSW_MAX_M || cfg->reserved != 0) {
    set_error("create: dim=%d max_m=%d out of range (max_m <= %d)", cfg->dim, cfg->max_m, B200HNSW_MAX_M);
    return B200HNSW_ERR_BAD_ARG;
  }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    set_error("create: no CUDA device (%s); this library has no CPU fallback", cudaGetErrorString(e));
    return B200HNSW_ERR_CUDA;
  }
  if (cfg->device < 0 || cfg->device >= ndev) {
    set_error("create: device %d out of range (%d devices)", cfg->device, ndev);
    return B200HNSW_ERR_BAD_ARG;
  }
  B200_CUDA(cudaSetDevice(cfg->device));
  b200hnsw *h = new (std::nothrow) b200hnsw();
  if (!h) return B200HNSW_ERR_OOM;
  h->cfg = *cfg;
  h->row_floats = (cfg->dim + 31) / 32 * 32;
  h->M = cfg->max_m;
  for (int l = 0; l < B200HNSW_MAX_LAYERS; l++) {
    h->entry[l] = -1;
    h->layer_size[l] = 0;
  }
  cudaDeviceProp prop;
  B200_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
  h->sm_count = prop.multiProcessorCount;
  h->smem_optin = prop.sharedMemPerBlockOptin;
  CallCtx *c = h->acquire();
  if (!c) {
    delete h;
    set_error("create: cannot create stream");
    return B200HNSW_ERR_CUDA;
  }
  int rc = h->ensure_capacity(cfg->capacity > 0 ? cfg->capacity : 1024, 1024, c->stream);
  if (!rc && cudaMalloc(&h->d_err_word, 64) != cudaSuccess) rc = B200HNSW_ERR_OOM;
  if (!rc) {
    cudaMemset(h->d_err_word, 0, 64);
    h->d_row_flags = h->d_err_word + 8;
  }
  h->release(c);
  if (rc) {
    b200hnsw_destroy(h);
    return rc;
  }
  *out = h;
  return 0;
}

void b200hnsw_destroy(b200hnsw_t *h) {
  if (!h) return;
  cudaSetDevice(h->cfg.device);
  cudaDeviceSynchronize();
  for (CallCtx *c : h->pool) {
    for (int i = 0; i < c->n_sub; i++) {
      cudaEventDestroy(c->sub_ev[i]);
      cudaStreamDestroy(c->sub[i]);
    }
    if (c->d_scratch) cudaFree(c->d_scratch);
    if (c->h_pinned) cudaFreeHost(c->h_pinned);
    cudaEventDestroy(c->ev0);
    cudaEventDestroy(c->ev1);
    cudaStreamDestroy(c->stream);
    delete c;
  }
  cudaFree(h->d.vec);
  cudaFree(h->d.keys);
  cudaFree(h->d.adj0);
  cudaFree(h->d.upper_off);
  cudaFree(h->d.dead);
  cudaFree(h->d.adjU);
  cudaFree(h->d.inbox);
  cudaFree(h->d_err_word);
  if (h->d_spec) cudaFree(h->d_spec);
  delete h;
}

int b200hnsw_reserve(b200hnsw_t *h, int64_t rows) {
  if (!h || rows < 0) return B200HNSW_ERR_BAD_ARG;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  CallCtx *c = h->acquire();
  if (!c) return B200HNSW_ERR_CUDA;
  int rc = h->ensure_capacity(rows, rows / 2 + 1024, c->stream);
  h->release(c);
  return rc;
}

int64_t b200hnsw_len(const b200hnsw_t *h) { return h && h->n_layers ? h->layer_size[0] : 0; }
int64_t b200hnsw_rows(const b200hnsw_t *h) { return h ? h->n : 0; }
int32_t b200hnsw_layers(const b200hnsw_t *h) { return h ? h->n_layers : 0; }
int64_t b200hnsw_layer_size(const b200hnsw_t *h, int32_t l) {
  return (h && l >= 0 && l < h->n_layers) ? h->layer_size[l] : 0;
}
int32_t b200hnsw_entry(const b200hnsw_t *h, int32_t l) { return (h && l >= 0 && l < h->n_layers) ? h->entry[l] : -1; }
int b200hnsw_set_entry(b200hnsw_t *h, int32_t l, int32_t row) {
  if (!h || l < 0 || l >= h->n_layers || row < 0 || row >= h->n || h->h_level[row] < l) {
    set_error("set_entry: row %d is not a member of layer %d", row, l);
    return B200HNSW_ERR_BAD_ARG;
  }
  h->entry[l] = row;
  return 0;
}

int b200hnsw_load(b200hnsw_t *h, int64_t n, const float *vectors, const int64_t *keys, const int32_t *levels,
                  int32_t n_layers, const int32_t *entries, const int64_t *layer_sizes, const int32_t *layer_rows,
                  const int64_t *layer_ptr, const int32_t *layer_nbr) {
  if (!h || n < 0 || n_layers < 0 || n_layers > B200HNSW_MAX_LAYERS || (n > 0 && (!vectors || !keys || !levels))) {
    set_error("load: bad argument");
    return B200HNSW_ERR_BAD_ARG;
  }
  if (n > 0 && (n_layers == 0 || !layer_sizes || !layer_rows || !layer_ptr || layer_sizes[0] != n)) {
    set_error("load: layer 0 must list every row");
    return B200HNSW_ERR_BAD_ARG;
  }
  B200_CUDA(cudaSetDevice(h->cfg.device));
  const int M = h->M, dim = h->cfg.dim, rf = h->row_floats;
  // slots for upper layers
  std::vector<int32_t> upper_off((size_t)n, -1);
  int64_t nU = 0;
  for (int64_t r = 0; r < n; r++) {
    if (levels[r] < 0 || levels[r] >= n_layers) {
      set_error("load: level %d of row %lld outside 0..%d", levels[r], (long long)r, n_layers - 1);
      return B200HNSW_ERR_BAD_ARG;
    }
    if (levels[r] > 0) {
      upper_off[r] = (int32_t)nU;
      nU += levels[r];
    }
  }
  std::vector<int32_t> adj0((size_t)n * M, -1), adjU((size_t)nU * M, -1);
  int64_t roff = 0, poff = 0, noff = 0;
  std::vector<int32_t> tmp;
  for (int l = 0; l < n_layers; l++) {
    int64_t nl = layer_sizes[l];
    for (int64_t i = 0; i < nl; i++) {
      int32_t row = layer_rows[roff + i];
      if (row < 0 || row >= n || levels[row] < l) {
        set_error("load: row %d listed in layer %d but its level is lower", row, l);
        return B200HNSW_ERR_BAD_ARG;
      }
      int64_t b = layer_ptr[poff + i], e = layer_ptr[poff + i + 1];
      if (e - b > M || e < b) {
        set_error("load: row %d has %lld neighbours in layer %d (max_m=%d)", row, (long long)(e - b), l, M);
        return B200HNSW_ERR_BAD_ARG;
      }
      tmp.assign(layer_nbr + noff + b, layer_nbr + noff + e);
      for (int32_t t : tmp)
        if (t < 0 || t >= n || levels[t] < l) {
          set_error("load: neighbour %d of row %d is not a member of layer %d", t, row, l);
          return B200HNSW_ERR_BAD_ARG;
        }
      std::sort(tmp.begin(), tmp.end(), [&](int32_t a, int32_t c) { return keys[a] < keys[c]; });
      int32_t *dst = l == 0 ? &adj0[(size_t)row * M] : &adjU[((size_t)upper_off[row] + (l - 1)) * M];
      std::copy(tmp.begin(), tmp.end(), dst);
    }
    noff += layer_ptr[poff + nl];
    roff += nl;
    poff += nl + 1;
  }
  CallCtx *c = h->acquire();
  if (!c) return B200HNSW_ERR_CUDA;
  // Load replaces the content: from here on the old graph is gone.  Whatever fails below leaves a consistent EMPTY
  // graph (host mirrors and device agree), never a half-reset handle.
  h->n = 0;
  h->nU = 0;
  h->n_layers = 0;
  h->n_dead = 0;
  h->h_keys.clear();
  h->h_level.clear();
  h->h_dead.clear();
  h->h_upper_off.clear();
  for (int l = 0; l < B200HNSW_MAX_LAYERS; l++) {
    h->entry[l] = -1;
    h->layer_size[l] = 0;
  }
  int rc = h->ensure_capacity(n, nU, c->stream);
  if (rc) {
    h->release(c);
    return rc;
  }
  cudaStream_t s = c->stream;
  auto fail = [&](int code) {
    h->release(c);
    return code;
  };
  if (cudaMemsetAsync(h->d_row_flags, 0, sizeof(int), s) != cudaSuccess) return fail(cuda_fail(cudaGetLastError(), "load"));
  h->row_flags.store(0);
  // wipe the whole reserved area: later Adds rely on untouched rows being zero vectors / empty lists
  if (cudaMemsetAsync(h->d.vec, 0, sizeof(float) * (size_t)h->d.cap * rf, s) != cudaSuccess ||
      cudaMemsetAsync(h->d.adj0, 0xff, sizeof(int32_t) * (size_t)h->d.cap * M, s) != cudaSuccess ||
      cudaMemsetAsync(h->d.upper_off, 0xff, sizeof(int32_t) * (size_t)h->d.cap, s) != cudaSuccess ||
      cudaMemsetAsync(h->d.adjU, 0xff, sizeof(int32_t) * (size_t)h->d.capU * M, s) != cudaSuccess)
    return fail(cuda_fail(cudaGetLastError(), "load wipe"));
  if (n > 0) {
    if (
        cudaMemcpy2DAsync(h->d.vec, sizeof(float) * rf, vectors, sizeof(float) * dim, sizeof(float) * dim, (size_t)n,
                          cudaMemcpyHostToDevice, s) != cudaSuccess ||
        cudaMemcpyAsync(h->d.keys, keys, sizeof(int64_t) * (size_t)n, cudaMemcpyHostToDevice, s) != cudaSuccess ||
        cudaMemcpyAsync(h->d.adj0, adj0.data(), sizeof(int32_t) * adj0.size(), cudaMemcpyHostToDevice, s) !=
            cudaSuccess ||
        cudaMemcpyAsync(h->d.upper_off, upper_off.data(), sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, s) !=
            cudaSuccess ||
        (nU > 0 && cudaMemcpyAsync(h->d.adjU, adjU.data(), sizeof(int32_t) * adjU.size(), cudaMemcpyHostToDevice,
                                   s) != cudaSuccess))
      return fail(cuda_fail(cudaGetLastError(), "load copies"));
  }
  if (h->d.dead && cudaMemsetAsync(h->d.dead, 0, (size_t)h->d.cap, s) != cudaSuccess)
    return fail(cuda_fail(cudaGetLastError(), "load: clearing the deleted flags"));
  if ((rc = h->flag_rows(0, n, s)) || (rc = h->refresh_row_flags(s))) return fail(rc);  // synchronises s
  h->n = n;
  h->nU = nU;
  h->n_layers = n_layers;
  h->h_keys.assign(keys, keys + n);
  h->h_level.assign(levels, levels + n);
  h->h_dead.assign((size_t)n, 0);
  h->n_dead = 0;
  h->h_upper_off = upper_off;
  roff = 0;
  for (int l = 0; l < B200HNSW_MAX_LAYERS; l++) {
    h->entry[l] = -1;
    h->layer_size[l] = 0;
  }
  for (int l = 0; l < n_layers; l++) {
    h->layer_size[l] = layer_sizes[l];
    int32_t en = entries ? entries[l] : -1;
    if (en < 0 && layer_sizes[l] > 0) {
      // "any node of the layer" (graph.go:246-258): the member with the smallest row = first inserted
      en = layer_rows[roff];
      for (int64_t i = 1; i < layer_sizes[l]; i++) en = std::min(en, layer_rows[roff + i]);
    }
    h->entry[l] = en;
    roff += layer_sizes[l];
  }
  h->release(c);
  return 0;
}

int b200hnsw_get_rows(const b200hnsw_t *hc, int64_t first, int64_t count, float *vectors, int64_t *keys,
                      int32_t *levels) {
  b200hnsw *h = const_cast<b200hnsw *>(hc);
  if (!h || first < 0 || count < 0 || first + count > h->n) {
    set_error("get_rows: range [%lld, +%lld) outside 0..%lld", (long long)first, (long long)count,
              h ? (long long)h->n : 0LL);
    return B200HNSW_ERR_BAD_ARG;
  }
  B200_CUDA(cudaSetDevice(h->cfg.device));
  if (vectors && count > 0)
    B200_CUDA(cudaMemcpy2D(vectors, sizeof(float) * h->cfg.dim, h->d.vec + (size_t)first * h->row_floats,
                           sizeof(float) * h->row_floats, sizeof(float) * h->cfg.dim, (size_t)count,
                           cudaMemcpyDeviceToHost));
  if (keys) std::copy(h->h_keys.begin() + first, h->h_keys.begin() + first + count, keys);
  if (levels) std::copy(h->h_level.begin() + first, h->h_level.begin() + first + count, levels);
  return 0;
}

int b200hnsw_set_keys(b200hnsw_t *h, int64_t first, int64_t count, const int64_t *keys) {
  if (!h || first < 0 || count < 0 || first + count > h->n || (count > 0 && !keys)) {
    set_error("set_keys: range [%lld, +%lld) outside 0..%lld", (long long)first, (long long)count,
              h ? (long long)h->n : 0LL);
    return B200HNSW_ERR_BAD_ARG;
  }
  if (count == 0) return 0;
  // the new keys must order the rows exactly as the old ones do, or the lists (kept sorted by key) would no longer be
  // in the reference's expansion order: checked on the range given (strictly, as keys are unique)
  {
    std::vector<int64_t> idx((size_t)count);
    for (int64_t i = 0; i < count; i++) idx[(size_t)i] = i;
    std::sort(idx.begin(), idx.end(), [&](int64_t a, int64_t b) { return h->h_keys[first + a] < h->h_keys[first + b]; });
    for (int64_t i = 1; i < count; i++)
      if (!(keys[idx[(size_t)i - 1]] < keys[idx[(size_t)i]])) {
        set_error("set_keys: the new keys do not preserve the order of rows %lld and %lld",
                  (long long)(first + idx[(size_t)i - 1]), (long long)(first + idx[(size_t)i]));
        return B200HNSW_ERR_BAD_ARG;
      }
  }
  B200_CUDA(cudaSetDevice(h->cfg.device));
  B200_CUDA(cudaMemcpy(h->d.keys + first, keys, sizeof(int64_t) * (size_t)count, cudaMemcpyHostToDevice));
  std::copy(keys, keys + count, h->h_keys.begin() + first);
  return 0;
}

int b200hnsw_get_neighbors(const b200hnsw_t *hc, int32_t layer, int64_t first, int64_t count, int32_t *nbr,
                           uint8_t *member) {
  b200hnsw *h = const_cast<b200hnsw *>(hc);
  if (!h || layer < 0 || layer >= h->n_layers || first < 0 || count < 0 || first + count > h->n || !nbr) {
    set_error("get_neighbors: bad argument");
    return B200HNSW_ERR_BAD_ARG;
  }
  B200_CUDA(cudaSetDevice(h->cfg.device));
  const int M = h->M;
  if (layer == 0) {
    if (count > 0)
      B200_CUDA(cudaMemcpy(nbr, h->d.adj0 + (size_t)first * M, sizeof(int32_t) * (size_t)count * M,
                           cudaMemcpyDeviceToHost));
    if (member) memset(member, 1, (size_t)count);
    return 0;
  }
  std::vector<int32_t> all((size_t)h->nU * M);
  if (h->nU > 0)
    B200_CUDA(cudaMemcpy(all.data(), h->d.adjU, sizeof(int32_t) * all.size(), cudaMemcpyDeviceToHost));
  for (int64_t i = 0; i < count; i++) {
    int64_t row = first + i;
    bool mem = h->h_level[row] >= layer;
    if (member) member[i] = mem;
    if (mem)
      std::copy(&all[((size_t)h->h_upper_off[row] + (layer - 1)) * M],
                &all[((size_t)h->h_upper_off[row] + (layer - 1)) * M] + M, nbr + (size_t)i * M);
    else
      std::fill(nbr + (size_t)i * M, nbr + (size_t)(i + 1) * M, -1);
  }
  return 0;
}

int b200hnsw_set_build_mode(b200hnsw_t *h, int32_t mode) {
  if (!h || (mode != B200HNSW_MODE_REFERENCE && mode != B200HNSW_MODE_EXTENDED)) {
    set_error("set_build_mode: unknown mode %d", mode);
    return B200HNSW_ERR_BAD_ARG;
  }
  h->build_mode = mode;
  return 0;
}
int32_t b200hnsw_get_build_mode(const b200hnsw_t *h) { return h ? h->build_mode : 0; }

int b200hnsw_set_add_window(b200hnsw_t *h, int32_t window) {
  if (!h || window < 0 || window > 128) {
    set_error("set_add_window: window must be 0 (automatic), 1 (sequential) or 2..128, got %d", window);
    return B200HNSW_ERR_BAD_ARG;
  }
  h->add_window = window;
  return 0;
}
int b200hnsw_get_add_window_stats(const b200hnsw_t *h, int64_t *accepted, int64_t *redone, int64_t *repaired) {
  if (!h) return B200HNSW_ERR_BAD_ARG;
  if (accepted) *accepted = (int64_t)h->spec_accepted_redone[0];
  if (redone) *redone = (int64_t)h->spec_accepted_redone[1];
  if (repaired) *repaired = (int64_t)h->spec_accepted_redone[2];
  return 0;
}

int32_t b200hnsw_get_visited_capacity(const b200hnsw_t *h) { return h ? h->vis_hint.load() : 0; }
int b200hnsw_set_visited_capacity(b200hnsw_t *h, int32_t entries) {
  if (!h || entries < 256 || (entries & (entries - 1))) {
    set_error("set_visited_capacity: %d is not a power of two >= 256", entries);
    return B200HNSW_ERR_BAD_ARG;
  }
  const int mx = max_visited_cap(h, 1, 1);
  if (entries > mx) {
    set_error("set_visited_capacity: %d exceeds what the device can give a traversal (%d)", entries, mx);
    return B200HNSW_ERR_UNSUPPORTED;
  }
  h->vis_hint.store(entries);
  return 0;
}

static int check_search_args(b200hnsw_t *h, int64_t B, int32_t k, int32_t ef, int32_t metric) {
  if (!h || B < 0) {
    set_error("search: bad handle or batch size");
    return B200HNSW_ERR_BAD_ARG;
  }
  if (k <= 0 || k > B200HNSW_MAX_K) {  // graph.go:542-544: "k must be greater than 0"
    set_error("k must be greater than 0 and <= %d, got %d", B200HNSW_MAX_K, k);
    return B200HNSW_ERR_BAD_ARG;
  }
  if (ef <= 0 || ef > B200HNSW_MAX_EF) {  // graph.go:928-930
    set_error("EfSearch must be greater than 0 and <= %d, got %d", B200HNSW_MAX_EF, ef);
    return B200HNSW_ERR_BAD_ARG;
  }
  if (metric != B200HNSW_EUCLIDEAN && metric != B200HNSW_COSINE) {
    set_error("unknown metric %d", metric);
    return B200HNSW_ERR_BAD_ARG;
  }
  return 0;
}

int b200hnsw_search_batch_device(b200hnsw_t *h, const float *d_queries, int64_t B, int32_t k, int32_t ef,
                                 int32_t metric, int32_t mode, const int32_t *entries, int32_t *d_out_rows,
                                 float *d_out_dists, int32_t *d_out_counts, int64_t *d_counters, void *stream) {
  int rc = check_search_args(h, B, k, ef, metric);
  if (rc) return rc;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  cudaStream_t s = (cudaStream_t)stream;
  if (B == 0) return 0;
  if (h->live_layers() == 0) {  // graph.go:554-556: empty graph -> no results
    B200_CUDA(cudaMemsetAsync(d_out_counts, 0, sizeof(int32_t) * (size_t)B, s));
    B200_CUDA(cudaMemsetAsync(d_out_rows, 0xff, sizeof(int32_t) * (size_t)B * k, s));
    B200_CUDA(cudaMemsetAsync(d_out_dists, 0xff, sizeof(float) * (size_t)B * k, s));
    return 0;
  }
  // d_counters is int64[8]: evals, expansions, nbr_reads, overflowed queries, error flag, 3 spare
  int *d_err = d_counters ? reinterpret_cast<int *>(d_counters + 4) : h->d_err_word;
  return launch_search(h, d_queries, B, k, ef, metric, mode, entries, d_out_rows, d_out_dists, d_out_counts,
                       reinterpret_cast<unsigned long long *>(d_counters), d_err, h->vis_hint.load(), s);
}

int b200hnsw_search_batch_segmented_device(b200hnsw_t *h, const float *d_queries, int64_t B, int32_t k, int32_t ef,
                                           int32_t metric, int32_t mode, const int32_t *entries, int32_t *d_packed,
                                           int64_t seg_len, int64_t seg_stride, int64_t *d_counters, void *stream) {
  int rc = check_search_args(h, B, k, ef, metric);
  if (rc) return rc;
  if (seg_len <= 0 || seg_stride < 2 * seg_len * k + seg_len || (B > 0 && !d_packed)) {
    set_error("search: bad segment layout (seg_len=%lld seg_stride=%lld, need >= %lld)", (long long)seg_len,
              (long long)seg_stride, (long long)(2 * seg_len * k + seg_len));
    return B200HNSW_ERR_BAD_ARG;
  }
  B200_CUDA(cudaSetDevice(h->cfg.device));
  cudaStream_t s = (cudaStream_t)stream;
  if (B == 0) return 0;
  const int64_t nseg = (B + seg_len - 1) / seg_len;
  if (h->live_layers() == 0) {  // graph.go:554-556: empty graph -> no results
    for (int64_t g = 0; g < nseg; g++) {
      int32_t *base = d_packed + g * seg_stride;
      B200_CUDA(cudaMemsetAsync(base, 0xff, sizeof(int32_t) * 2 * (size_t)seg_len * k, s));
      B200_CUDA(cudaMemsetAsync(base + 2 * seg_len * k, 0, sizeof(int32_t) * (size_t)seg_len, s));
    }
    return 0;
  }
  int *d_err = d_counters ? reinterpret_cast<int *>(d_counters + 4) : h->d_err_word;
  return launch_search(h, d_queries, B, k, ef, metric, mode, entries, d_packed, nullptr, nullptr,
                       reinterpret_cast<unsigned long long *>(d_counters), d_err, h->vis_hint.load(), s, seg_len, seg_stride);
}

int b200hnsw_search_batch_scatter_device(b200hnsw_t *h, const float *d_queries, int64_t B, int32_t k, int32_t ef,
                                         int32_t metric, int32_t mode, const int32_t *entries,
                                         int32_t *const *seg_ptrs, int32_t n_seg, int64_t seg_len,
                                         int64_t *d_counters, void *stream) {
  int rc = check_search_args(h, B, k, ef, metric);
  if (rc) return rc;
  if (seg_len <= 0 || n_seg <= 0 || n_seg > B200HNSW_MAX_PEERS || !seg_ptrs || (int64_t)n_seg * seg_len < B) {
    set_error("search: bad scatter layout (seg_len=%lld n_seg=%d for B=%lld; n_seg <= %d)", (long long)seg_len, n_seg,
              (long long)B, B200HNSW_MAX_PEERS);
    return B200HNSW_ERR_BAD_ARG;
  }
  for (int i = 0; i < n_seg; i++)
    if (!seg_ptrs[i]) {
      set_error("search: scatter pointer %d is null", i);
      return B200HNSW_ERR_BAD_ARG;
    }
  B200_CUDA(cudaSetDevice(h->cfg.device));
  cudaStream_t s = (cudaStream_t)stream;
  if (B == 0) return 0;
  if (h->live_layers() == 0) {  // graph.go:554-556: empty graph -> no results
    for (int g = 0; g < n_seg; g++) {
      B200_CUDA(cudaMemsetAsync(seg_ptrs[g], 0xff, sizeof(int32_t) * 2 * (size_t)seg_len * k, s));
      B200_CUDA(cudaMemsetAsync(seg_ptrs[g] + 2 * seg_len * k, 0, sizeof(int32_t) * (size_t)seg_len, s));
    }
    return 0;
  }
  int *d_err = d_counters ? reinterpret_cast<int *>(d_counters + 4) : h->d_err_word;
  // out_rows doubles as the "Search sink" marker of the kernel: any non-null value
  return launch_search(h, d_queries, B, k, ef, metric, mode, entries, seg_ptrs[0], nullptr, nullptr,
                       reinterpret_cast<unsigned long long *>(d_counters), d_err, h->vis_hint.load(), s, seg_len, 0, seg_ptrs,
                       n_seg);
}

// device alias of a page-locked, mapped host buffer (cudaHostAlloc / cudaHostRegister under UVA), or nullptr
static void *mapped_host(const void *p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return a.type == cudaMemoryTypeHost ? a.devicePointer : nullptr;
}

int b200hnsw_search_batch(b200hnsw_t *h, const float *queries, int64_t B, int32_t k, int32_t ef, int32_t metric,
                          int32_t mode, const int32_t *entries, int32_t *out_rows, float *out_dists,
                          int32_t *out_counts, b200hnsw_stats *stats) {
  int rc = check_search_args(h, B, k, ef, metric);
  if (rc) return rc;
  if (B > 0 && (!queries || !out_rows || !out_dists || !out_counts)) {
    set_error("search: null buffer");
    return B200HNSW_ERR_BAD_ARG;
  }
  if (stats) memset(stats, 0, sizeof *stats);
  if (B == 0) return 0;
  if (h->live_layers() == 0) {
    std::fill(out_counts, out_counts + B, 0);
    std::fill(out_rows, out_rows + B * k, -1);
    std::fill(out_dists, out_dists + B * k, NAN);
    return 0;
  }
  B200_CUDA(cudaSetDevice(h->cfg.device));
  static const bool trace_slow = getenv("B200HNSW_TRACE_SLOW") != nullptr;  // diagnostics: report calls over 5 ms
  const auto w0 = std::chrono::steady_clock::now();
  auto since = [&]() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - w0).count(); };
  double w_acq = 0, w_queued = 0;
  CallCtx *c = h->acquire();
  if (!c) return B200HNSW_ERR_CUDA;
  w_acq = since();
  const int dim = h->cfg.dim;
  size_t qb = sizeof(float) * (size_t)B * dim, rb = sizeof(int32_t) * (size_t)B * k, db = sizeof(float) * (size_t)B * k,
         cb = sizeof(int32_t) * (size_t)B;
  auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
  size_t total = al(qb) + al(rb) + al(db) + al(cb) + 256;
  auto fail = [&](int code) {
    h->release(c);
    return code;
  };
  if ((rc = c->reserve_device(total))) return fail(rc);
  char *p = (char *)c->d_scratch;
  float *d_q = (float *)p;
  p += al(qb);
  int32_t *d_rows = (int32_t *)p;
  p += al(rb);
  float *d_dists = (float *)p;
  p += al(db);
  int32_t *d_cnt = (int32_t *)p;
  p += al(cb);
  unsigned long long *d_ctr = (unsigned long long *)p;  // [4] + error word
  int *d_err = (int *)(d_ctr + 4);
  cudaStream_t s = c->stream;
  cudaError_t e;
  // Page-locked result buffers are written by the kernel itself (80 B of posted PCIe writes per query at the end of
  // its traversal): no device staging, no D2H copies behind the kernel.
  bool direct_out = false;
  // Pageable buffers (what a cgo host passes from make([]float32, ...)): cudaMemcpyAsync from or to pageable memory goes
  // through the driver's own staging, synchronously, at ~9 GB/s here.  The call stages them itself instead, through the
  // context's page-locked buffer: queries are copied there a chunk at a time by this thread (each chunk's H2D copy and
  // traversal run while the next chunk is being copied), the kernel writes its results there, and they are copied out
  // after the wait.  B200HNSW_NO_HOST_STAGING=1 keeps the driver's path.
  static const bool own_staging = getenv("B200HNSW_NO_HOST_STAGING") == nullptr;
  const bool q_pageable = own_staging && mapped_host(queries) == nullptr && qb >= (256u << 10);
  bool out_staged = false;
  char *h_stage_q = nullptr, *h_stage_o = nullptr;
  if (!getenv("B200HNSW_NO_DIRECT_OUT")) {
    void *zr = mapped_host(out_rows), *zd = mapped_host(out_dists), *zc = mapped_host(out_counts);
    if (zr && zd && zc) {
      d_rows = (int32_t *)zr;
      d_dists = (float *)zd;
      d_cnt = (int32_t *)zc;
      direct_out = true;
    } else if (own_staging && !zr && !zd && !zc) {
      out_staged = true;
    }
  }
  if (q_pageable || out_staged) {
    if ((rc = c->reserve_pinned(256 + (q_pageable ? al(qb) : 0) + (out_staged ? al(rb) + al(db) + al(cb) : 0))))
      return fail(rc);
    char *hp = (char *)c->h_pinned + 256;  // the first 64 bytes are the counters
    if (q_pageable) {
      h_stage_q = hp;
      hp += al(qb);
    }
    if (out_staged) {
      h_stage_o = hp;
      void *zo = mapped_host(h_stage_o);
      if (zo) {  // the kernel's result stores go straight into the page-locked staging
        d_rows = (int32_t *)zo;
        d_dists = (float *)((char *)zo + al(rb));
        d_cnt = (int32_t *)((char *)zo + al(rb) + al(db));
        direct_out = true;
      } else {
        out_staged = false;
      }
    }
  }
  // The batch is cut in chunks that run on their own streams: the H2D copy of chunk i+1 and the D2H copy of chunk
  // i-1 overlap the traversal kernel of chunk i, and the kernels of different chunks share the SMs.
  int nchunk = B >= 1024 ? 2 : 1;  // measured (profiles/r3h_e2e_callers.log): 2 beats 4 alone and with callers in flight
  if (q_pageable && B >= 4096) nchunk = 4;  // the host copy of chunk i+1 runs under the H2D copy and traversal of chunk i
  if (const char *v = getenv("B200HNSW_E2E_CHUNKS")) {
    int t = atoi(v);
    if (t >= 1 && t <= CallCtx::MAX_SUB) nchunk = t;
  }
  if ((rc = c->ensure_sub(nchunk)) || (rc = c->reserve_pinned(256))) return fail(rc);
  // counters come back into page-locked memory: a copy into pageable memory would go through the driver's staging
  // path, which serialises against the other calls in flight
  volatile unsigned long long *ctr = (volatile unsigned long long *)c->h_pinned;
  for (int i = 0; i < 5; i++) ctr[i] = 0;
  int launches = 0;
  bool staged = false;
  // Page-locked QUERIES could be read by the kernel itself as well (each traversal fetching its 4*dim bytes over PCIe
  // when it starts).  Measured and rejected as the default (profiles/r5j_direct_in_ab.log): one caller 29.4 M vs 30.2 M
  // queries/s, four callers 54 M vs 71 M -- a traversal that waits 2 us for its query holds its slot on the SM, and
  // the copy engine moves the same bytes without occupying one.  B200HNSW_DIRECT_IN=1 selects it for experiments.
  const float *q_src = d_q;
  {
    const char *v = getenv("B200HNSW_DIRECT_IN");
    if (v && v[0] == '1') {
      if (void *zq = mapped_host(queries)) {
        q_src = (const float *)zq;
        staged = true;
      }
    }
  }
  const bool direct_in = q_src != d_q;
  if (nchunk == 1 && !staged) {  // one chunk: copy first, so that kernel_ms (ev0..ev1) is the traversal kernel alone
    if (q_pageable) memcpy(h_stage_q, queries, qb);
    if ((e = cudaMemcpyAsync(d_q, q_pageable ? (const float *)h_stage_q : queries, qb, cudaMemcpyHostToDevice, s)) != cudaSuccess)
      return fail(cuda_fail(e, "search staging"));
    staged = true;
  }
  // A layer search that visits more nodes than the shared-memory visited table holds fails loudly in the kernel;
  // the whole batch is then redone with a table 4x the size, so results are always exact.  The larger size becomes
  // the handle's hint only AFTER a retry with it has succeeded, and never exceeds what the device can give a
  // traversal: past that the call fails with B200HNSW_ERR_VISITED_OVERFLOW and leaves the handle as it was.
  // Error 3 (NaN distances met with the short candidate queue, traverse.cuh) is retried once with the full queue.
  int vis0 = h->vis_hint.load(), vis_max = max_visited_cap(h, k, ef);
  if (mode == B200HNSW_MODE_EXTENDED) {  // the extended walk visits ~10x the nodes: its own default and its own ceiling
    const int efk = ef > k ? ef : k;
    vis_max = max_visited_cap_ext(h, efk);
    vis0 = ext_visited_default(efk, h->M);
    if (h->vis_hint_ext.load() > vis0) vis0 = h->vis_hint_ext.load();
  }
  int vis = vis0 < vis_max ? vis0 : vis_max, full_queue = 0;
  for (;;) {
    if ((e = cudaMemsetAsync(d_ctr, 0, 64, s)) != cudaSuccess || (e = cudaEventRecord(c->ev0, s)) != cudaSuccess)
      return fail(cuda_fail(e, "search staging"));
    for (int ci = 0; ci < nchunk; ci++) {
      const int64_t q0 = B * ci / nchunk, q1 = B * (ci + 1) / nchunk, nq = q1 - q0;
      if (nq <= 0) continue;
      cudaStream_t sc = nchunk == 1 ? s : c->sub[ci];
      if (nchunk > 1 && (e = cudaStreamWaitEvent(sc, c->ev0, 0)) != cudaSuccess)
        return fail(cuda_fail(e, "search chunk wait"));
      if (!staged) {
        const float *src = queries + q0 * dim;
        if (q_pageable) {
          float *st = (float *)h_stage_q + q0 * dim;
          memcpy(st, src, sizeof(float) * (size_t)nq * dim);
          src = st;
        }
        if ((e = cudaMemcpyAsync(d_q + q0 * dim, src, sizeof(float) * (size_t)nq * dim, cudaMemcpyHostToDevice, sc)) !=
            cudaSuccess)
          return fail(cuda_fail(e, "search staging"));
      }
      rc = launch_search(h, q_src + q0 * dim, nq, k, ef, metric, mode, entries, d_rows + q0 * k, d_dists + q0 * k,
                         d_cnt + q0, d_ctr, d_err, vis, sc, 0, 0, nullptr, 0, full_queue);
      if (rc) return fail(rc);
      launches++;
      if (!direct_out &&
          ((e = cudaMemcpyAsync(out_rows + q0 * k, d_rows + q0 * k, sizeof(int32_t) * (size_t)nq * k,
                               cudaMemcpyDeviceToHost, sc)) != cudaSuccess ||
          (e = cudaMemcpyAsync(out_dists + q0 * k, d_dists + q0 * k, sizeof(float) * (size_t)nq * k,
                               cudaMemcpyDeviceToHost, sc)) != cudaSuccess ||
          (e = cudaMemcpyAsync(out_counts + q0, d_cnt + q0, sizeof(int32_t) * (size_t)nq, cudaMemcpyDeviceToHost,
                               sc)) != cudaSuccess))
        return fail(cuda_fail(e, "search copy-back"));
      if (nchunk > 1 && ((e = cudaEventRecord(c->sub_ev[ci], sc)) != cudaSuccess ||
                         (e = cudaStreamWaitEvent(s, c->sub_ev[ci], 0)) != cudaSuccess))
        return fail(cuda_fail(e, "search chunk join"));
    }
    staged = true;
    w_queued = since();
    if ((e = cudaEventRecord(c->ev1, s)) != cudaSuccess ||
        (e = cudaMemcpyAsync((void *)ctr, d_ctr, 40, cudaMemcpyDeviceToHost, s)) != cudaSuccess ||
        (e = wait_stream(s)) != cudaSuccess)
      return fail(cuda_fail(e, "search kernel"));
    const int err = (int)(ctr[4] & 0xffffffffu);
    if (err == 0) break;
    if (err == 3 && !full_queue) {  // the kernel writes the most recent code: an overflow may hide behind it, see below
      full_queue = 1;
      continue;
    }
    if (vis >= vis_max) {
      set_error("search: %llu queries overflowed the visited set (a layer search visited > %d nodes, the largest table "
                "this device can give a traversal)", ctr[3], vis / 2);
      return fail(B200HNSW_ERR_VISITED_OVERFLOW);
    }
    vis = vis * 4 < vis_max ? vis * 4 : vis_max;
  }
  if (out_staged) {
    memcpy(out_rows, h_stage_o, rb);
    memcpy(out_dists, h_stage_o + al(rb), db);
    memcpy(out_counts, h_stage_o + al(rb) + al(db), cb);
  }
  if (vis > vis0) {  // sticky, but only now that a search with this table has succeeded; concurrent callers keep the largest
    std::atomic<int> &hint = mode == B200HNSW_MODE_EXTENDED ? h->vis_hint_ext : h->vis_hint;
    int cur = hint.load();
    while (vis > cur && !hint.compare_exchange_weak(cur, vis)) {
    }
  }
  if (stats) {
    stats->dist_evals = (int64_t)ctr[0];
    stats->expansions = (int64_t)ctr[1];
    stats->nbr_reads = (int64_t)ctr[2];
    stats->launches = launches;
    cudaEventElapsedTime(&stats->kernel_ms, c->ev0, c->ev1);
  }
  if (trace_slow && since() > 5.0) {
    float kms = 0.f;
    cudaEventElapsedTime(&kms, c->ev0, c->ev1);
    fprintf(stderr, "b200hnsw_search_batch slow call: total %.2f ms (context %.2f, queued at %.2f), device ev0..ev1 %.2f ms, "
                    "B=%lld chunks=%d visited table %d direct_out=%d direct_in=%d\n", since(), w_acq, w_queued, kms,
            (long long)B, nchunk, h->vis_hint.load(), (int)direct_out, (int)direct_in);
  }
  h->release(c);
  return 0;
}

}  // extern "C"
