// inserter.cuh -- the block-wide machinery of Graph.Add / Graph.Delete shared by the sequential kernels (insert_kernel.cu)
// and the speculative windowed Add (spec_insert.cu): addNeighbor (graph.go:41-81), replenish (graph.go:172-219), isolate
// (graph.go:223-235) executed by one 512-thread CTA on one node at a time.
//
// SPEC = false: lists are read and written in place (the sequential kernels).
// SPEC = true : the CTA works on a node SPECULATIVELY: every list it reads is logged (rlog), every list it writes goes to a
//               private overlay (ovl, indexed by a shared-memory hash of list ids) instead of the graph, and list reads
//               bypass L1 (another CTA commits to the graph while this one runs).  spec_insert.cu decides afterwards whether
//               the logged reads were still current and copies the overlay into the graph in input order.
#pragma once
#include "index.hpp"
#include "listops.cuh"

namespace b200 {

constexpr int INS_WARPS = 16;
constexpr int INS_THREADS = INS_WARPS * 32;

struct InsertParams {
  long long first_row;
  int n;                   // nodes in this batch
  const int32_t *levels;   // [n] device
  int m, ef, metric;
  int n_layers;
  int entries[B200HNSW_MAX_LAYERS];
  int *state_out;                 // [1 + MAX_LAYERS]: n_layers, entries
  unsigned long long *counters;   // [8]: evals, expansions, nbr_reads, evict_evals, replenish_evals, evictions
  int *error;
  int heap_c, heap_r, vis_cap;    // search scratch (uint2 slots, uint2 slots, u32 entries)
  int cand_cap, rep_cap;          // replenish candidates (<= M*M) and its visited table (power of two)
  const uint8_t *dead;            // [rows] deleted flags, null while nothing was deleted
  const int *row_flags;           // device word: bit 0 = some stored row can yield NaN distances (index.hpp)
  // Delete (delete_kernel): rows to isolate, in call order
  const int32_t *del_rows;
  const int32_t *del_levels;
  uint8_t *dead_set;              // deleted flags, written by the delete kernel
};

struct MutGraph {
  float4 *vec;
  int64_t *keys;
  int32_t *adj0;
  int32_t *adjU;
  int32_t *upper_off;
  int32_t row_chunks, M, dim;
  __device__ __forceinline__ int32_t *list(int layer, int row) const {
    if (layer == 0) return adj0 + (size_t)row * M;
    return adjU + ((size_t)upper_off[row] + (layer - 1)) * M;
  }
};

// shared-memory control block
struct InsCtl {
  int n_layers;
  int elevator;
  int cnt;        // neighbourhood size of the current layer search, -1 on overflow
  int cur;        // block_layer_search: node to expand next, -1 = the layer ends
  int fresh_n;    // block_layer_search: unvisited neighbours of this expansion, -1 on visited overflow
  int vcount;     // block_layer_search: entries in the visited table
  int own_len;    // entries in the new node's own list on the layer being linked (own / ownk)
  int own_dirty;  // the new node itself was evicted from a neighbour (its own list lost the back-link and was replenished,
                  // graph.go:73-79): the shared-memory copy is dead, the rest of the layer goes through add_neighbor
  int wsum[17];   // replenish: candidates found per warp (block-wide compaction)
  int flag;       // overflow / continue flag of the current block-wide step
  int tmp_len;    // entries in tmp list
  int worst;      // evicted neighbour
  int wl_len;     // degree of the node being replenished
  int ncand;      // replenish candidates
  int need_heap;  // replenish needs the exact heap (NaN present or more than one pop)
  int abort_;
  int n_rd, n_ov, ov_fail;  // SPEC: lists read (log entries), lists written (overlay entries), overlay overflow
  unsigned long long ent_mask;  // SPEC: layers that were empty and take this node as their entry
  int entries[B200HNSW_MAX_LAYERS];
  int cnts[B200HNSW_MAX_LAYERS];  // neighbourhood sizes of the node being inserted, -1 = the layer was empty
};

// ---- block-wide state -------------------------------------------------------------------------------------------
constexpr int OV_MAX = 1024;   // overlay lists per speculated node (33 per layer the node joins at M neighbours)
constexpr int OV_HASH = 2048;  // shared-memory hash slots over them

#ifdef B200_K2_TIMING  // exploration build only: cycles per phase as thread 0 sees them, printed by the kernel at its end
#define K2_T(i) { if (tid == 0) { const long long t__ = clock64(); tph[i] += t__ - tlast; tlast = t__; } }
#else
#define K2_T(i)
#endif

template <int CPL, int U, bool SPEC = false>
struct Inserter {
#ifdef B200_K2_TIMING
  long long tph[16], tlast;
#endif
  MutGraph g;
  InsertParams p;
  InsCtl *ctl;
  int *nbh;        // [MAX_LAYERS][M] neighbourhood rows per layer (heap-array order)
  int *tmp;        // [M+1] list being edited
  float *tmpd;     // [M+1] eviction-scan distances
  int *wl;         // [M+1] list of the node being replenished
  int *cand;       // [cand_cap] replenish candidates in push order
  float *candd;    // [cand_cap]
  uint32_t *rep;   // [rep_cap] replenish: hash of the rows met (neighbours of neighbours, and the excluded ones)
  int *repi;       // [rep_cap] replenish: position of the first occurrence of that row in the scan, -1 = excluded
  int *own;        // [M + 1] the new node's own list on the layer being linked, sorted by key
  long long *ownk; // [M + 1] their keys
  long long *nbk;  // [M] keys of the layer's neighbourhood (nbh)
  GoHeap C, R;
  Visited V;
  int tid, lane, warp;
  unsigned long long c_evict, c_repl, c_evictions;  // thread-local counters (summed at the end)
  // SPEC only
  uint32_t *ovk;   // shared [OV_HASH]: list id, ~0 = empty
  int *ovx;        // shared [OV_HASH]: overlay entry of that id
  int32_t *ovl;    // global [OV_MAX + 1][M]: the overlay lists (entry OV_MAX takes the writes after an overflow)
  uint32_t *ovid;  // global [OV_MAX]: list id of every overlay entry, in first-write order
  uint32_t *rlog;  // global [rl_cap]: ids of the lists read
  int rl_cap;

  // one 32-bit name per list: the row on layer 0, the upper slot (top bit set) above
  __device__ __forceinline__ uint32_t list_id(int layer, int row) const {
    return layer == 0 ? (uint32_t)row : 0x80000000u | (uint32_t)(g.upper_off[row] + (layer - 1));
  }
  __device__ __forceinline__ int32_t *base_list(uint32_t id) const {
    return (id & 0x80000000u) ? g.adjU + (size_t)(id & 0x7fffffffu) * g.M : g.adj0 + (size_t)id * g.M;
  }
  __device__ __forceinline__ int ov_find(uint32_t id) const {
    uint32_t h = (id * 2654435761u) >> 16;
    for (;;) {
      h &= OV_HASH - 1;
      const uint32_t k = ovk[h];
      if (k == id) return ovx[h];
      if (k == ~0u) return -1;
      h++;
    }
  }
  // The list a warp reads (warp-uniform arguments).  SPEC: logged, and the overlay's copy if this node wrote it before.
  __device__ __forceinline__ const int32_t *rlist(int layer, int row) {
    if constexpr (!SPEC) {
      return g.list(layer, row);
    } else {
      const uint32_t id = list_id(layer, row);
      if (lane == 0) {
        const int k = atomicAdd(&ctl->n_rd, 1);
        if (k < rl_cap) rlog[k] = id;
      }
      const int x = ov_find(id);
      return x >= 0 ? ovl + (size_t)x * g.M : base_list(id);
    }
  }
  // The list a warp writes (warp-uniform arguments; every store rewrites the whole stride, so nothing is copied in).
  __device__ __forceinline__ int32_t *wlist(int layer, int row) {
    if constexpr (!SPEC) {
      return g.list(layer, row);
    } else {
      const uint32_t id = list_id(layer, row);
      int x = ov_find(id);
      if (x < 0) {
        if (lane == 0) {
          x = atomicAdd(&ctl->n_ov, 1);
          if (x >= OV_MAX) {
            ctl->ov_fail = 1;
            x = OV_MAX;
          } else {
            ovid[x] = id;
            uint32_t h = (id * 2654435761u) >> 16;
            for (;;) {  // two warps may add different ids at once (add_neighbor), never the same one
              h &= OV_HASH - 1;
              if (atomicCAS(&ovk[h], ~0u, id) == ~0u) break;
              h++;
            }
            ovx[h] = x;
          }
        }
        x = __shfl_sync(FULL, x, 0);
      }
      return ovl + (size_t)x * g.M;
    }
  }
  static __device__ __forceinline__ int32_t ldl(const int32_t *p) {
    if constexpr (SPEC) return __ldcg(p);
    else return *p;
  }

  // distance(row = a, query = b) for a set of rows against one "query" row, warps striding over the items
  __device__ __forceinline__ void block_distances_q(const int *ids, int n, const QueryRegs<CPL> &Q, int metric,
                                                    float *out) {
    for (int base = warp * U; base < n; base += INS_WARPS * U) {
      float4 rows[U][CPL];
#pragma unroll
      for (int u = 0; u < U; u++)
        if (base + u < n) load_row<CPL>(rows[u], g.vec + (size_t)ids[base + u] * g.row_chunks, g.row_chunks, lane);
#pragma unroll
      for (int u = 0; u < U; u++)
        if (base + u < n) {
          float d = finish_distance<CPL>(rows[u], Q, metric);
          if (lane == 0) out[base + u] = d;
        }
    }
  }
  __device__ __forceinline__ void block_distances(const int *ids, int n, int qrow, int metric, float *out) {
    QueryRegs<CPL> Q;
    load_query<CPL>(Q, reinterpret_cast<const float *>(g.vec + (size_t)qrow * g.row_chunks), g.dim, g.row_chunks,
                    lane);
    block_distances_q(ids, n, Q, metric, out);
  }

  // layerNode.search (graph.go:94-170) for the node being inserted (k = m), BLOCK-wide: warp 0 reads the popped node's
  // list and filters it through the visited set, every warp evaluates rows of the expansion (all of them in flight at
  // once: one memory round trip per expansion instead of one per U rows), thread 0 replays the order-dependent part
  // (graph.go:146-166) in key order on the Go-exact heaps.  Every warp holds the query in registers (Q).
  // Returns the result count (R holds the heap array, valid for every thread) or -1 on visited overflow.
  __device__ int block_layer_search(int layer, int entry, const QueryRegs<CPL> &Q, SearchCounters &sc) {
    const int M = g.M, k = p.m, ef = p.ef, metric = p.metric;
    int *fresh = cand;       // [<= M] unvisited neighbours of the expansion, key order
    float *freshd = candd;
    __syncthreads();  // cand / candd / ctl are free
    if (warp == 0) {
      V.clear(lane);
      if (lane == 0) fresh[0] = entry;
    }
    __syncthreads();
    block_distances_q(fresh, 1, Q, metric, freshd);  // graph.go:107-108
    __syncthreads();
    // Once the result queue is full it stays full, and from then on the candidate queue is never looked at again: the
    // loop goes on only when a row improved on result.Min(), that row is strictly below everything evaluated on the
    // layer, so candidates.Push sifted it to the root and the next Pop returns exactly the last improving row (the
    // argument of traverse.cuh, K1).  NaN distances break it (Go's up() never moves an element past a NaN): graphs that
    // hold rows able to produce them keep the queue for the whole layer.
    const bool keepc = p.row_flags != nullptr && (__ldcg(p.row_flags) & 1);
    bool cfree = false;  // thread 0
    if (tid == 0) {
      // graph.go:109-127: candidates.Push(entry); result.Push(entry); visited[entry]; the first Pop takes it back
      C.len = 0;
      R.len = 0;
      sc.evals++;
      C.push(freshd[0], entry);
      R.push(freshd[0], entry);
      V.insert((uint32_t)entry);
      ctl->cur = (int)C.pop().y;
      ctl->vcount = 1;
      ctl->cnt = 1;
    }
    for (;;) {
      __syncthreads();
      const int cur = ctl->cur;
      if (cur < 0) break;
      if (warp == 0) {
        const int32_t *list = rlist(layer, cur);
        int nf = 0, vc = ctl->vcount, nbr = 0;
        bool ovf = false;
        for (int base = 0; base < M; base += 32) {  // graph.go:137-144: ascending key order, unvisited only
          int nb = -1;
          if (base + lane < M) nb = ldl(list + base + lane);
          const bool valid = nb >= 0;
          const unsigned vm = __ballot_sync(FULL, valid);
          if (vm == 0) break;
          nbr += __popc(vm);
          const bool fr = valid && V.insert((uint32_t)nb);
          const unsigned fm = __ballot_sync(FULL, fr);
          if (fr) {
            fresh[nf + __popc(fm & ((1u << lane) - 1u))] = nb;
            const float4 *rowp = g.vec + (size_t)nb * g.row_chunks;
            for (int c = 0; c < g.row_chunks; c += 8) prefetch_l2(rowp + c);
            if (layer == 0) prefetch_l2(g.adj0 + (size_t)nb * M);  // one of these rows is the next to be expanded
          }
          nf += __popc(fm);
          vc += __popc(fm);
          if (vc * 2 > (int)V.mask) {  // load factor > 1/2: fail loudly
            ovf = true;
            break;
          }
        }
        if (lane == 0) {
          ctl->fresh_n = ovf ? -1 : nf;
          ctl->vcount = vc;
          sc.expansions++;
          sc.nbr_reads += nbr;
        }
      }
      __syncthreads();
      const int nf = ctl->fresh_n;
      if (nf < 0) return -1;
      if (nf > 0) {
        block_distances_q(fresh, nf, Q, metric, freshd);
        __syncthreads();
      }
      if (tid == 0) {
        bool improved = false;
        int imp_id = -1;
        for (int i = 0; i < nf; i++) {  // graph.go:146-159, key order
          const float d = freshd[i];
          const int id = fresh[i];
          if (R.len > 0 && d < R.min_dist()) {  // graph.go:147
            improved = true;
            imp_id = id;
          }
          if (R.len < k) {
            R.push(d, id);
          } else if (d < R.max_dist()) {  // result.Max() is the last slot
            R.pop_last();
            R.push(d, id);
          }
          if (!cfree) {
            C.push(d, id);
            if (C.len > ef) C.pop_last();  // graph.go:157-159
          }
        }
        sc.evals += nf;
        int next;
        if (cfree) next = improved ? imp_id : -1;            // what candidates.Pop() would return (see above)
        else if (!improved && R.len >= k) next = -1;         // graph.go:164-166
        else next = C.len > 0 ? (int)C.pop().y : -1;         // graph.go:125-127
        if (R.len >= k && !keepc) cfree = true;
        ctl->cur = next;
        ctl->cnt = R.len;
      }
    }
    return ctl->cnt;
  }

  // layerNode.replenish(m) -- graph.go:172-219; always CosineDistance.  Block-wide; ends with a barrier.
  // have_wl: wl[0 .. ctl->wl_len) already holds w's list (the caller has just taken a back-link out of it): no reload.
  __device__ void replenish(int layer, int w, bool have_wl = false) {
    const int m = p.m, M = g.M;
    __syncthreads();  // every thread has finished reading the control block of the previous step
    for (int i = tid; i < p.rep_cap; i += INS_THREADS) {
      rep[i] = ~0u;
      repi[i] = 0x7fffffff;
    }
    if (warp == 0) {
      int deg = have_wl ? ctl->wl_len : warp_load_list<SPEC>(rlist(layer, w), M, wl, lane);
      if (lane == 0) {
        ctl->wl_len = deg;
        ctl->flag = deg < m;  // graph.go:173-175
        ctl->ncand = 0;
      }
    }
    __syncthreads();
    K2_T(4)
    if (!ctl->flag) return;
    // all neighbours' lists are fetched at once (independent gathers), one list per warp, into shared memory
    int *nl = reinterpret_cast<int *>(candd);  // [deg][M], free until the distances are written
    {
      const int deg = ctl->wl_len;
      for (int i = warp; i < deg; i += INS_WARPS) {
        const int32_t *ln = rlist(layer, wl[i]);
        for (int j = lane; j < M; j += 32) nl[i * M + j] = ldl(ln + j);
      }
    }
    __syncthreads();
    K2_T(5)
    {
      // graph.go:183-210: the candidates are the neighbours of neighbours outside {w} U neighbours(w), in the order a
      // scan of the lists (ascending key at both levels) FIRST meets them -- the heap's tie rule depends on that order.
      // Block-wide: every row met goes into a hash with the smallest scan position it was seen at (-1 for the excluded
      // rows); a position is a candidate iff it holds that smallest position; a prefix sum compacts them in scan order.
      const int deg = ctl->wl_len;
      const uint32_t rmask = (uint32_t)p.rep_cap - 1;
      auto note = [&](uint32_t v, int pos) {
        uint32_t h = (v * 2654435761u) >> 7;
        for (;;) {
          h &= rmask;
          const uint32_t old = atomicCAS(&rep[h], ~0u, v);
          if (old == ~0u || old == v) break;
          h++;
        }
        atomicMin(&repi[h], pos);
      };
      if (tid == 0) note((uint32_t)w, -1);
      if (tid >= 32 && tid - 32 < deg) note((uint32_t)wl[tid - 32], -1);
      const int total = deg * M;
      for (int e = tid; e < total; e += INS_THREADS) {
        const int v = nl[e];
        if (v >= 0) note((uint32_t)v, e);
      }
      __syncthreads();
      int base_n = 0;
      for (int c0 = 0; c0 < total; c0 += INS_THREADS) {
        const int e = c0 + tid;
        int v = -1;
        bool is = false;
        if (e < total) {
          v = nl[e];
          if (v >= 0) {
            uint32_t h = ((uint32_t)v * 2654435761u) >> 7;
            for (;;) {
              h &= rmask;
              if (rep[h] == (uint32_t)v) break;
              h++;
            }
            is = repi[h] == e;
          }
        }
        const unsigned bm = __ballot_sync(FULL, is);
        if (lane == 0) ctl->wsum[warp] = __popc(bm);
        __syncthreads();
        int off = base_n, tot = 0;
        for (int ww = 0; ww < INS_WARPS; ww++) {
          const int c = ctl->wsum[ww];
          if (ww < warp) off += c;
          tot += c;
        }
        if (is) cand[off + __popc(bm & ((1u << lane) - 1u))] = v;
        base_n += tot;
        __syncthreads();
      }
      if (tid == 0) ctl->ncand = base_n;
    }
    __syncthreads();
    K2_T(6)
    const int nc = ctl->ncand;
    if (nc == 0) return;
    block_distances(cand, nc, w, 1 /*cosine*/, candd);  // CosineDistance(candidate.Value, n.Value)
    if (tid == 0) c_repl += nc;
    __syncthreads();
    K2_T(7)
    if (warp == 0) {
      // graph.go:213-218: pop best-first until the list is full.  The first pop of a min-heap filled by pushes is
      // the first-pushed global minimum; that shortcut is exact when one pop suffices and no distance is NaN.
      int deg = ctl->wl_len;
      const int need = m - deg;
      bool nan = false;
      for (int i = lane; i < nc; i += 32) nan = nan || (candd[i] != candd[i]);
      nan = __any_sync(FULL, nan);
      if (need == 1 && !nan) {
        float bd = __int_as_float(0x7f800000);
        int bi = 0x7fffffff;
        for (int i = lane; i < nc; i += 32) {
          float d = candd[i];
          if (d < bd || (d == bd && i < bi)) {
            bd = d;
            bi = i;
          }
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
          float od = __shfl_xor_sync(FULL, bd, off);
          int oi = __shfl_xor_sync(FULL, bi, off);
          if (od < bd || (od == bd && oi < bi)) {
            bd = od;
            bi = oi;
          }
        }
        deg = warp_sorted_insert(wl, deg, cand[bi], g.keys, lane);
      } else {
        // exact heap emulation (lane 0) in the search heap C's storage is too small in general: use candd/cand
        // in place as a binary heap built by successive pushes.
        if (lane == 0) {
          // build: push i-th element into heap of size i (arrays themselves are the heap storage)
          for (int i = 1; i < nc; i++) {
            float d = candd[i];
            int id_ = cand[i];
            int j = i;
            while (j > 0) {
              int pi = (j - 1) >> 1;
              if (!(d < candd[pi])) break;
              candd[j] = candd[pi];
              cand[j] = cand[pi];
              j = pi;
            }
            candd[j] = d;
            cand[j] = id_;
          }
        }
        __syncwarp();
        int hn = nc;
        while (hn > 0 && deg < m) {
          int best = cand[0];
          __syncwarp();
          if (lane == 0) {  // container/heap.Pop
            int n = --hn;
            float md = candd[n];
            int mid = cand[n];
            int i = 0;
            for (;;) {
              int j1 = 2 * i + 1;
              if (j1 >= n) break;
              int j = j1;
              if (j1 + 1 < n && candd[j1 + 1] < candd[j1]) j = j1 + 1;
              if (!(candd[j] < md)) break;
              candd[i] = candd[j];
              cand[i] = cand[j];
              i = j;
            }
            if (n > 0) {
              candd[i] = md;
              cand[i] = mid;
            }
          }
          hn = __shfl_sync(FULL, hn, 0);
          __syncwarp();
          deg = warp_sorted_insert(wl, deg, best, g.keys, lane);  // addNeighbor(best): len <= m, no overflow
        }
      }
      warp_store_list(wlist(layer, w), M, wl, deg, lane);
    }
    __syncthreads();
    K2_T(8)
  }

  __device__ __forceinline__ void prefetch_neighbourhood(int layer, int n) {
    const int M = g.M;
    const int32_t *ln = g.list(layer, n);
    for (int j = lane; j < M; j += 32) {
      const int e = ldl(ln + j);
      if (e < 0) continue;
      const float4 *rowp = g.vec + (size_t)e * g.row_chunks;
      for (int c = 0; c < g.row_chunks; c += 8) prefetch_l2(rowp + c);
      prefetch_l2(g.keys + e);
      if (layer == 0) prefetch_l2(g.adj0 + (size_t)e * M);
    }
  }

  // layerNode.addNeighbor(newNode, m, dist) on node n -- graph.go:41-81.  Block-wide; ends with a barrier.
  // next_n: the node the caller's loop will call this for next (-1: none).  A spare warp requests its list, and the rows,
  // keys and lists of its neighbours, into L2 now: a hint only (never logged, never consulted), it turns the dependent
  // DRAM round trips of the next call into L2 hits.
  __device__ void add_neighbor(int layer, int n, int x, int metric, int next_n = -1) {
    const int m = p.m, M = g.M;
    __syncthreads();  // every thread has finished reading the control block of the previous step
    if (warp == INS_WARPS - 1 && next_n >= 0) prefetch_neighbourhood(layer, next_n);
    if (warp == 0) {
      int deg = warp_load_list<SPEC>(rlist(layer, n), M, tmp, lane);
      int nd = warp_sorted_insert(tmp, deg, x, g.keys, lane);  // graph.go:50
      bool over = nd > m;                                      // graph.go:51-53
      if (!over && nd != deg) warp_store_list(wlist(layer, n), M, tmp, nd, lane);
      if (lane == 0) {
        ctl->flag = over;
        ctl->tmp_len = nd;
      }
    }
    __syncthreads();
    K2_T(0)
    if (!ctl->flag) return;
    const int len = ctl->tmp_len;  // m + 1
    // graph.go:60-71: dist(neighbor.Value, n.Value) for every neighbour
    block_distances(tmp, len, n, metric, tmpd);
    __syncthreads();
    K2_T(1)
    if (warp == 0) {
      // graph.go:60-71 over the entries in key order (the stand-in for Go's map order): `d > worst || worst == nil`
      // keeps the FIRST maximum; a NaN never wins unless it is the first entry, and then nothing replaces it.
      const float d0 = tmpd[0];
      float bd = __int_as_float(0xff800000);  // -Inf
      int bi = 0x7fffffff;
      for (int i = lane; i < len; i += 32) {
        float d = tmpd[i];
        if (d != d) d = __int_as_float(0xff800000);
        if (d > bd || (d == bd && i < bi)) {
          bd = d;
          bi = i;
        }
      }
#pragma unroll
      for (int off = 16; off >= 1; off >>= 1) {
        const float od = __shfl_xor_sync(FULL, bd, off);
        const int oi = __shfl_xor_sync(FULL, bi, off);
        if (od > bd || (od == bd && oi < bi)) {
          bd = od;
          bi = oi;
        }
      }
      const int wpos = (d0 != d0) ? 0 : bi;
      const int worst = tmp[wpos];
      __syncwarp();
      int moved[3];
      int cm = 0;
      for (int i = wpos + lane; i + 1 < len; i += 32) moved[cm++] = tmp[i + 1];  // delete(n.neighbors, worst.Key)
      __syncwarp();
      cm = 0;
      for (int i = wpos + lane; i + 1 < len; i += 32) tmp[i] = moved[cm++];
      if (lane == 0) {
        c_evict += len;
        if (worst == x) ctl->own_dirty = 1;  // x is being linked by the caller's loop: its shared-memory list is stale now
        ctl->worst = worst;
        ctl->tmp_len = len - 1;
        c_evictions++;
      }
    }
    __syncthreads();
    K2_T(2)
    const int worst = ctl->worst;
    if (warp == 0) warp_store_list(wlist(layer, n), M, tmp, ctl->tmp_len, lane);
    if (warp == 1) {  // graph.go:75-78: delete the back-link
      int dw = warp_load_list<SPEC>(rlist(layer, worst), M, wl, lane);
      int nd = warp_list_remove(wl, dw, n, lane);
      if (nd != dw) warp_store_list(wlist(layer, worst), M, wl, nd, lane);
      if (lane == 0) ctl->wl_len = nd;  // replenish continues on this copy
    }
    __syncthreads();
    K2_T(3)
    replenish(layer, worst, true);  // graph.go:79
  }

  // The new node's side of graph.go:516-520 (newNode.addNeighbor(node)): its list on this layer starts empty and, as long
  // as the node is not itself evicted from a neighbour (own_dirty), only this loop writes it and it cannot overflow (the
  // neighbourhood has at most m members): it is kept in shared memory, sorted by key, and stored after every insertion --
  // replenish may read it in between through another node's list.
  __device__ void own_begin(int layer, const int *nbrs, int cnt) {
    __syncthreads();
    for (int i = tid; i < cnt; i += INS_THREADS) nbk[i] = g.keys[nbrs[i]];
    if (warp == INS_WARPS - 1 && cnt > 0) prefetch_neighbourhood(layer, nbrs[0]);
    if (tid == 0) {
      ctl->own_len = 0;
      ctl->own_dirty = 0;
    }
  }
  __device__ void add_own(int layer, int row, int x, long long kx) {
    if (warp != 0) return;
    const int M = g.M;
    const int len = ctl->own_len;
    int pos = 0, eqpos = -1;
    bool present = false;
    for (int base = 0; base < len; base += 32) {
      const int i = base + lane;
      bool lt = false, eq = false;
      if (i < len) {
        present = present || own[i] == x;
        lt = ownk[i] < kx;
        eq = ownk[i] == kx && own[i] != x;  // another row under the same key (listops.cuh: warp_sorted_insert)
      }
      pos += __popc(__ballot_sync(FULL, lt));
      const unsigned em = __ballot_sync(FULL, eq);
      if (em && eqpos < 0) eqpos = base + __ffs(em) - 1;
    }
    if (__any_sync(FULL, present)) return;
    int nlen = len;
    if (eqpos >= 0) {
      if (lane == 0) own[eqpos] = x;
    } else {
      int mv[3];
      long long mk[3];
      int cm = 0;
      for (int i = pos + lane; i < len; i += 32) {
        mv[cm] = own[i];
        mk[cm++] = ownk[i];
      }
      __syncwarp();
      cm = 0;
      for (int i = pos + lane; i < len; i += 32) {
        own[i + 1] = mv[cm];
        ownk[i + 1] = mk[cm++];
      }
      if (lane == 0) {
        own[pos] = x;
        ownk[pos] = kx;
        ctl->own_len = len + 1;
      }
      nlen = len + 1;
    }
    __syncwarp();
    warp_store_list(wlist(layer, row), M, own, nlen, lane);
  }

  // layerNode.isolate(m) -- graph.go:223-235: every neighbour drops its link to n and is replenished at once, in
  // ascending key order (the stand-in for Go's map order).  n's own list is left as it is, as in the reference.
  __device__ void isolate(int layer, int n) {
    const int M = g.M;
    __syncthreads();
    if (warp == 0) {
      int deg = warp_load_list(g.list(layer, n), M, nbh, lane);
      if (lane == 0) ctl->cnt = deg;
    }
    __syncthreads();
    const int deg = ctl->cnt;
    for (int i = 0; i < deg; i++) {
      const int nb = nbh[i];
      if (warp == 1) {  // graph.go:232: delete(neighbor.neighbors, n.Key)
        int32_t *l = g.list(layer, nb);
        int d = warp_load_list(l, M, wl, lane);
        int nd = warp_list_remove(wl, d, n, lane);
        if (nd != d) warp_store_list(l, M, wl, nd, lane);
      }
      replenish(layer, nb);  // graph.go:233 (starts with a barrier)
    }
  }
};

static __host__ __device__ inline size_t al16(size_t v) { return (v + 15) & ~(size_t)15; }

struct InsLayout {
  size_t ctl, heapC, heapR, vis, nbh, tmp, tmpd, wl, cand, candd, rep, repi, own, ownk, nbk, total;
  __host__ __device__ InsLayout(const InsertParams &p, int M) {
    size_t o = 0;
    ctl = o;   o += al16(sizeof(InsCtl));
    heapC = o; o += al16(sizeof(uint2) * p.heap_c);
    heapR = o; o += al16(sizeof(uint2) * p.heap_r);
    vis = o;   o += al16(sizeof(uint32_t) * p.vis_cap);
    nbh = o;   o += al16(sizeof(int) * (size_t)B200HNSW_MAX_LAYERS * M);
    tmp = o;   o += al16(sizeof(int) * (M + 1));
    tmpd = o;  o += al16(sizeof(float) * (M + 1));
    wl = o;    o += al16(sizeof(int) * (M + 1));
    cand = o;  o += al16(sizeof(int) * p.cand_cap);
    candd = o; o += al16(sizeof(float) * p.cand_cap);
    rep = o;   o += al16(sizeof(uint32_t) * p.rep_cap);
    repi = o;  o += al16(sizeof(int) * p.rep_cap);
    own = o;   o += al16(sizeof(int) * (M + 1));
    ownk = o;  o += al16(sizeof(long long) * (M + 1));
    nbk = o;   o += al16(sizeof(long long) * M);
    total = o;
  }
};


// ---- host side, shared by the launchers ------------------------------------------------------------------------------
inline int pow2_at_least(int v) {
  int r = 1;
  while (r < v) r <<= 1;
  return r;
}
inline MutGraph mut_graph_of(b200hnsw *h) {
  MutGraph g;
  g.vec = reinterpret_cast<float4 *>(h->d.vec);
  g.keys = h->d.keys;
  g.adj0 = h->d.adj0;
  g.adjU = h->d.adjU;
  g.upper_off = h->d.upper_off;
  g.row_chunks = h->row_floats / 4;
  g.M = h->M;
  g.dim = h->cfg.dim;
  return g;
}
inline void size_insert_scratch(InsertParams &p, const MutGraph &g) {
  p.heap_c = (p.ef + 2) & ~1;
  p.heap_r = (p.m + 2) & ~1;
  p.vis_cap = 1024;
  p.cand_cap = g.M * g.M + 4;  // also holds the [deg][M] staging of the neighbours' lists
  p.rep_cap = pow2_at_least(2 * (g.M * g.M + g.M + 2));
}
int spec_window_auto(int64_t rows);
int launch_spec_insert(b200hnsw *h, InsertParams &p, int depth, int *done, int *err, int32_t *n_layers,
                       int32_t *entries, unsigned long long *accepted_redone, cudaStream_t s);

}  // namespace b200
