// wave_insert.cu -- K2b: batched Graph.Add in WAVES (north_star: "Batched Add evaluates candidates in parallel in the
// same kernels and applies the same neighbor-selection heuristic").  NOT the reference's sequential order -- that is
// insert_kernel.cu -- but a deterministic parallel schedule of the same steps, restated on the CPU by
// oracle/ora_graph.c (ora_graph_add_waves) and checked bit for bit against it:
//   schedule : wave_len() below == ora_wave_len()
//   phase S  : K1 (traverse.cuh) with k = M on every layer, one traversal per new node, against the graph as it was
//              before the wave (graph.go:475-508)
//   phase A  : link_kernel     -- the new node's own lists = its results (graph.go:519); requests queued per target
//   phase B  : targets_kernel  -- per (layer, target), one warp: insert the new rows in input order, on overflow
//              evict the farthest (graph.go:50-73)
//   phase C  : backlink_kernel -- per evicted node: delete the evictors from its list (graph.go:75-78)
//   phase D  : replenish_kernel-- per evicted node short of M links: neighbours of neighbours scored with
//              CosineDistance (graph.go:172-219), staged
//   phase E  : scatter_kernel  -- staged lists replace the old ones
// Requests are grouped per target without sorting: link_kernel counts them (atomicAdd per list slot), group_kernel
// gives every target a contiguous segment of the request arena and drops each request into its place; the warp holding
// a target's first-counted request owns the target.  Evicted rows are grouped per row the same way they were before
// (atomicExch linked lists + atomicMin owner election).  No result depends on the scheduling (see phase B).
// Bound: HBM gather (search + eviction scans + replenish rows), counted in the same counters as the oracle's.
#include <algorithm>
#include <cstring>
#include <vector>

#include "index.hpp"
#include "listops.cuh"
#include "traverse.cuh"

namespace b200 {

constexpr int WV_WARPS = 4;  // warps per block in the commit kernels

struct WaveParams {
  // graph
  float4 *vec;
  int64_t *keys;
  int32_t *adj0, *adjU, *upper_off;
  int32_t row_chunks, M, dim;
  long long cap;  // list slot of (row, 0) = row; of (row, l > 0) = cap + upper_off[row] + l - 1
  // inboxes, one entry per list slot
  int32_t *in_cnt, *in_base, *ev_head, *ev_owner;  // requests per slot / their arena segment; eviction lists
  // this wave
  int first_row, w, m, metric;
  const int32_t *levels;      // [w]
  const long long *slot_off;  // [w]
  long long n_slots;
  int32_t *slot_cnt;    // [n_slots]
  int32_t *slot_rows;   // [n_slots][M]
  float *slot_dists;    // [n_slots][M]
  int32_t *slot_xrow;   // [n_slots] new row of the slot
  int32_t *slot_layer;  // [n_slots]
  int32_t *req_pos;     // [n_slots * M] position of the request inside its target's segment
  int32_t *req_arena;   // [n_slots * M] request ids grouped by target
  int32_t *ev_rec;      // [n_slots * M][3]: layer, n, worst
  int32_t *ev_next;     // [n_slots * M]
  int32_t *rep_rec;     // [n_slots * M][2]: layer, row
  int32_t *staged;      // [n_slots * M][M + 1]: new list, then its degree
  int *wave_ctr;        // [0] evictions recorded, [1] replenish work items, [2] waves committed, [3] arena fill,
                        // [4] NaN seen
  unsigned long long *counters;  // [8]: evals, expansions, nbr_reads, evict_evals, replenish_evals, evictions
  int *error;
  int rep_cap;          // replenish visited-table entries per warp (power of two)

  __device__ __forceinline__ int32_t *list(int layer, int row) const {
    if (layer == 0) return adj0 + (size_t)row * M;
    return adjU + ((size_t)upper_off[row] + (layer - 1)) * M;
  }
  __device__ __forceinline__ long long lslot(int layer, int row) const {
    return layer == 0 ? (long long)row : cap + upper_off[row] + (layer - 1);
  }
};

// ---- phase A + request queueing: one warp per new node -------------------------------------------------------------
__global__ void __launch_bounds__(WV_WARPS * 32) link_kernel(WaveParams p) {
  const int lane = threadIdx.x & 31;
  const int j = blockIdx.x * WV_WARPS + (threadIdx.x >> 5);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    p.wave_ctr[0] = 0;
    p.wave_ctr[1] = 0;
    p.wave_ctr[3] = 0;
  }
  if (*p.error || j >= p.w) return;
  const int x = p.first_row + j;
  const int M = p.M;
  for (int l = 0; l <= p.levels[j]; l++) {
    const long long slot = p.slot_off[j] + l;
    if (lane == 0) {
      p.slot_xrow[slot] = x;
      p.slot_layer[slot] = l;
    }
    const int cnt = p.slot_cnt[slot];
    if (cnt <= 0) continue;  // a layer this node opens: nothing to link to
    const int32_t *rows = p.slot_rows + slot * M;
    int32_t *lx = p.list(l, x);
    for (int t = lane; t < cnt; t += 32) {
      const int n = rows[t];
      const float dn = p.slot_dists[slot * M + t];
      if (dn != dn) p.wave_ctr[4] = 1;  // phase B then takes the literal, order-dependent path
      const long long kn = p.keys[n];
      int rank = 0;
      for (int u = 0; u < cnt; u++) rank += p.keys[rows[u]] < kn;  // keys are unique
      lx[rank] = n;  // graph.go:519: newNode.addNeighbor(n), lists kept in key order
      const int r = (int)(slot * M + t);  // requests are ordered by (node, layer, position)
      const long long s = p.lslot(l, n);
      p.req_pos[r] = atomicAdd(&p.in_cnt[s], 1);
    }
  }
}

// ---- requests grouped by target: one thread per request, two passes (segment allocation, placement) ------------------
__global__ void group_kernel(WaveParams p, int pass) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (*p.error || r >= p.n_slots * p.M) return;
  const long long slot = r / p.M;
  if ((int)(r - slot * p.M) >= p.slot_cnt[slot]) return;
  const long long s = p.lslot(p.slot_layer[slot], p.slot_rows[r]);
  const int pos = p.req_pos[r];
  if (pass == 0) {
    if (pos == 0) p.in_base[s] = atomicAdd(&p.wave_ctr[3], p.in_cnt[s]);
  } else {
    p.req_arena[p.in_base[s] + pos] = (int)r;
  }
}


// distances of ids[0..n) (shared memory) to the row held in Q, one warp, U rows in flight; out[i] by lane 0
template <int CH, int U>
__device__ __forceinline__ void warp_distances(const WaveParams &p, const int *ids, int n, const GQuery<CH> &Q,
                                               int metric, float *out, int lane) {
  for (int b0 = 0; b0 < n; b0 += U) {
    int r[U];
#pragma unroll
    for (int u = 0; u < U; u++) r[u] = b0 + u < n ? ids[b0 + u] : -1;
    const float dm = group_distances<32, CH, U>(p.vec, p.row_chunks, r, Q, metric, lane);
    // row u sits in lanes [u*32/U, (u+1)*32/U)
    if ((lane & (32 / U - 1)) == 0) {
      const int u = lane / (32 / U);
      if (b0 + u < n) out[b0 + u] = dm;
    }
  }
  __syncwarp();
}

// ---- phase B: one warp per (layer, target) that received requests ------------------------------------------------
// graph.go:50-73 applied to target n for every request in input order: insert x; once the list holds m + 1 rows, drop
// the one farthest from n (`d > worst || worst == nil` scanning in ascending key order, i.e. among equal maxima the
// smallest key).  Two ways to get there:
//   * target_sequential: literally that, request by request (any distances, NaN included);
//   * target_select: when no distance involved is NaN the rule is a total order (farther first, then smaller key), and
//     "insert, then drop the maximum" repeated over any sequence leaves exactly the m smallest of list U requests:
//     the outcome does not depend on the order, so requests are streamed as they come off the inbox.  The distance
//     of a new row x to n is the one its own search already produced (slot_dists; both metrics are symmetric bit for
//     bit in the canonical order), so only the rows already in the list are evaluated, once per target.
//   evict_evals still counts what the reference would evaluate (m + 1 per eviction).
struct TargetSmem {
  int *kid;          // [M + 4] list / kept rows
  float *kd;         // [M + 4] their distance to the target
  long long *kkey;   // [M + 4] their keys
  int *ebuf;         // [64] rows evicted, not yet recorded
  __host__ __device__ static size_t bytes(int M) { return (size_t)(M + 4) * 16 + 64 * 4; }
  __device__ TargetSmem(unsigned char *base, int M) {
    kkey = reinterpret_cast<long long *>(base);
    kid = reinterpret_cast<int *>(kkey + (M + 4));
    kd = reinterpret_cast<float *>(kid + (M + 4));
    ebuf = reinterpret_cast<int *>(kd + (M + 4));
  }
};

// rows ev[0..cnt) were dropped by target (l, n): one record each, grouped per dropped row for phase C
__device__ __forceinline__ void record_evictions(const WaveParams &p, int l, int n, const int *ev, int cnt, int lane) {
  int e0 = 0;
  if (lane == 0) e0 = atomicAdd(&p.wave_ctr[0], cnt);
  e0 = __shfl_sync(FULL, e0, 0);
  for (int j = lane; j < cnt; j += 32) {
    const int e = e0 + j, worst = ev[j];
    p.ev_rec[3 * e] = l;
    p.ev_rec[3 * e + 1] = n;
    p.ev_rec[3 * e + 2] = worst;
    if (worst < p.first_row) {  // rows of this wave keep their own link and are not replenished (see oracle)
      const long long sw = p.lslot(l, worst);
      p.ev_next[e] = atomicExch(&p.ev_head[sw], e);
      atomicMin(&p.ev_owner[sw], e);
    }
  }
  __syncwarp();
}

template <int CH, int U>
__device__ void target_sequential(const WaveParams &p, int l, int n, long long s, const TargetSmem &T,
                                  unsigned long long &c_evict, unsigned long long &c_evictions, int lane) {
  const int M = p.M, m = p.m;
  int *tmp = T.kid;
  float *td = T.kd;
  int32_t *ln = p.list(l, n);
  int deg = warp_load_list(ln, M, tmp, lane);
  const int r = p.in_cnt[s];
  const int32_t *seg = p.req_arena + p.in_base[s];
  GQuery<CH> Q;
  bool have_q = false;
  int last = -1;
  for (int it = 0; it < r; it++) {  // requests in ascending order (= input order of the new nodes)
    int best = 0x7fffffff;
    for (int i = lane; i < r; i += 32) {
      const int r2 = seg[i];
      if (r2 > last && r2 < best) best = r2;
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) best = min(best, __shfl_xor_sync(FULL, best, off));
    last = best;
    const int x = p.slot_xrow[best / M];
    deg = warp_sorted_insert(tmp, deg, x, p.keys, lane);  // graph.go:50
    if (deg <= m) continue;                               // graph.go:51-53
    if (!have_q) {
      group_load_query<32, CH>(Q, reinterpret_cast<const float *>(p.vec + (size_t)n * p.row_chunks), true, p.dim,
                               p.row_chunks, lane);
      have_q = true;
    }
    warp_distances<CH, U>(p, tmp, deg, Q, p.metric, td, lane);  // graph.go:60-63: dist(neighbor.Value, n.Value)
    c_evict += deg;
    int wpos = 0;
    if (lane == 0) {  // graph.go:64-71 in ascending key order: `d > worst || worst == nil`
      float wd = td[0];
      for (int i = 1; i < deg; i++)
        if (td[i] > wd) {
          wd = td[i];
          wpos = i;
        }
    }
    wpos = __shfl_sync(FULL, wpos, 0);
    const int worst = tmp[wpos];
    __syncwarp();
    deg = warp_list_remove(tmp, deg, worst, lane);  // graph.go:73
    c_evictions++;
    if (lane == 0) T.ebuf[0] = worst;
    __syncwarp();
    record_evictions(p, l, n, T.ebuf, 1, lane);
  }
  warp_store_list(ln, M, tmp, deg, lane);
}

// position of the row to drop next among T.k*[0..deg): farthest, then smallest key (uniform over the warp)
__device__ __forceinline__ int kept_worst(const TargetSmem &T, int deg, int lane) {
  float wd = 0.f;
  long long wk = 0;
  int wp = -1;
  for (int i = lane; i < deg; i += 32) {
    const float d = T.kd[i];
    const long long k = T.kkey[i];
    if (wp < 0 || d > wd || (d == wd && k < wk)) {
      wd = d;
      wk = k;
      wp = i;
    }
  }
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const float od = __shfl_xor_sync(FULL, wd, off);
    const long long ok = __shfl_xor_sync(FULL, wk, off);
    const int op = __shfl_xor_sync(FULL, wp, off);
    if (op >= 0 && (wp < 0 || od > wd || (od == wd && ok < wk))) {
      wd = od;
      wk = ok;
      wp = op;
    }
  }
  return wp;
}

// T.kid[0..deg) with keys T.kkey -> the list in key order, -1 padded
__device__ __forceinline__ void store_by_key(const TargetSmem &T, int32_t *ln, int deg, int M, int lane) {
  int mine[2], rank[2];
  int cnt = 0;
  for (int j = lane; j < deg; j += 32) {
    const long long kj = T.kkey[j];
    int rk = 0;
    for (int u = 0; u < deg; u++) rk += T.kkey[u] < kj;  // keys are unique
    mine[cnt] = T.kid[j];
    rank[cnt++] = rk;
  }
  __syncwarp();
  for (int c = 0; c < cnt; c++) ln[rank[c]] = mine[c];
  for (int j = deg + lane; j < M; j += 32) ln[j] = -1;
}

template <int CH, int U>
__device__ void target_select(const WaveParams &p, int l, int n, long long s, const TargetSmem &T,
                              unsigned long long &c_evict, unsigned long long &c_evictions, int lane) {
  const int M = p.M, m = p.m;
  int32_t *ln = p.list(l, n);
  const int deg0 = warp_load_list(ln, M, T.kid, lane);
  const int r = p.in_cnt[s];
  const int32_t *seg = p.req_arena + p.in_base[s];
  for (int j = lane; j < deg0; j += 32) T.kkey[j] = p.keys[T.kid[j]];
  if (deg0 + r <= m) {  // graph.go:51-53: room for everybody, no distance needed
    for (int i = lane; i < r; i += 32) {
      const int x = p.slot_xrow[seg[i] / M];
      T.kid[deg0 + i] = x;
      T.kkey[deg0 + i] = p.keys[x];
    }
    __syncwarp();
    store_by_key(T, ln, deg0 + r, M, lane);
    return;
  }
  if (deg0 > 0) {  // the rows that were in the list before the wave: their distance to n, once
    GQuery<CH> Q;
    group_load_query<32, CH>(Q, reinterpret_cast<const float *>(p.vec + (size_t)n * p.row_chunks), true, p.dim,
                             p.row_chunks, lane);
    warp_distances<CH, U>(p, T.kid, deg0, Q, p.metric, T.kd, lane);
    bool nan = false;
    for (int j = lane; j < deg0; j += 32) nan = nan || (T.kd[j] != T.kd[j]);
    if (__any_sync(FULL, nan)) {  // not a total order: nothing has been recorded yet, do it the literal way
      target_sequential<CH, U>(p, l, n, s, T, c_evict, c_evictions, lane);
      return;
    }
  }
  int deg = deg0, wpos = -1, nev = 0;
  for (int b0 = 0; b0 < r; b0 += 32) {
    bool pending = b0 + lane < r;
    int x = -1;
    float dx = 0.f;
    long long kx = 0;
    if (pending) {
      const int r2 = seg[b0 + lane];
      x = p.slot_xrow[r2 / M];
      dx = p.slot_dists[r2];  // d(x, n) from x's own search == dist(x.Value, n.Value), graph.go:60-63
      kx = p.keys[x];
    }
    for (;;) {
      unsigned mask;
      if (deg < m) {
        mask = __ballot_sync(FULL, pending);  // room left: anybody
      } else {
        if (wpos < 0) wpos = kept_worst(T, deg, lane);
        const float wd = T.kd[wpos];
        const long long wk = T.kkey[wpos];
        mask = __ballot_sync(FULL, pending && !(dx > wd || (dx == wd && kx < wk)));  // beats the row to drop
      }
      if (!mask) break;
      const int i = __ffs(mask) - 1;
      const int xi = __shfl_sync(FULL, x, i);
      const float di = __shfl_sync(FULL, dx, i);
      const long long ki = __shfl_sync(FULL, kx, i);
      const int at = deg < m ? deg : wpos;
      __syncwarp();
      if (lane == 0) {
        if (deg >= m) T.ebuf[nev] = T.kid[at];
        T.kid[at] = xi;
        T.kd[at] = di;
        T.kkey[at] = ki;
      }
      if (deg < m) deg++;
      else nev++;
      wpos = -1;
      if (lane == i) pending = false;
      __syncwarp();
      if (nev >= 32) {
        record_evictions(p, l, n, T.ebuf, nev, lane);
        c_evictions += nev;
        c_evict += (unsigned long long)nev * (m + 1);  // what graph.go:60-63 evaluates per eviction
        nev = 0;
      }
    }
    // whoever is still pending is farther than every kept row: dropped on arrival
    const unsigned rej = __ballot_sync(FULL, pending);
    if (pending) T.ebuf[nev + __popc(rej & ((1u << lane) - 1u))] = x;
    nev += __popc(rej);
    __syncwarp();
    if (nev >= 32) {
      record_evictions(p, l, n, T.ebuf, nev, lane);
      c_evictions += nev;
      c_evict += (unsigned long long)nev * (m + 1);
      nev = 0;
    }
  }
  if (nev > 0) {
    record_evictions(p, l, n, T.ebuf, nev, lane);
    c_evictions += nev;
    c_evict += (unsigned long long)nev * (m + 1);
  }
  store_by_key(T, ln, deg, M, lane);
}

template <int CH, int U>
__global__ void __launch_bounds__(WV_WARPS * 32) targets_kernel(WaveParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int M = p.M;
  const TargetSmem T(smem + TargetSmem::bytes(M) * warp, M);
  const long long slot = (long long)blockIdx.x * WV_WARPS + warp;
  if (*p.error || slot >= p.n_slots) return;
  const int cnt = p.slot_cnt[slot];
  const int l = p.slot_layer[slot];
  const bool any_nan = p.wave_ctr[4] != 0;  // some search of this wave returned a NaN distance (link_kernel)
  unsigned long long c_evict = 0, c_evictions = 0;
  for (int t = 0; t < cnt; t++) {
    const long long r = slot * M + t;
    if (p.req_pos[r] != 0) continue;  // another warp owns this target (warp-uniform)
    const int n = p.slot_rows[r];
    const long long s = p.lslot(l, n);
    if (any_nan) target_sequential<CH, U>(p, l, n, s, T, c_evict, c_evictions, lane);
    else target_select<CH, U>(p, l, n, s, T, c_evict, c_evictions, lane);
    __syncwarp();
    if (lane == 0) p.in_cnt[s] = 0;
    __syncwarp();
  }
  if (lane == 0) {
    if (c_evict) atomicAdd(&p.counters[3], c_evict);
    if (c_evictions) atomicAdd(&p.counters[5], c_evictions);
  }
}

// ---- phase C: one warp per evicted node: drop its links to the evictors ------------------------------------------
__global__ void __launch_bounds__(WV_WARPS * 32) backlink_kernel(WaveParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int M = p.M;
  int *tmp = reinterpret_cast<int *>(smem) + (size_t)warp * (M + 4);
  if (*p.error) return;
  const int nev = p.wave_ctr[0];
  for (int e = blockIdx.x * WV_WARPS + warp; e < nev; e += gridDim.x * WV_WARPS) {
    const int l = p.ev_rec[3 * e], worst = p.ev_rec[3 * e + 2];
    if (worst >= p.first_row) continue;
    const long long sw = p.lslot(l, worst);
    if (p.ev_owner[sw] != e) continue;
    int32_t *lw = p.list(l, worst);
    int deg = warp_load_list(lw, M, tmp, lane);
    for (int e2 = p.ev_head[sw]; e2 >= 0; e2 = p.ev_next[e2])
      deg = warp_list_remove(tmp, deg, p.ev_rec[3 * e2 + 1], lane);  // graph.go:75-78
    warp_store_list(lw, M, tmp, deg, lane);
    if (lane == 0) {
      p.ev_head[sw] = -1;
      p.ev_owner[sw] = 0x7fffffff;
      if (deg < p.m) {  // graph.go:173-175
        const int i = atomicAdd(&p.wave_ctr[1], 1);
        p.rep_rec[2 * i] = l;
        p.rep_rec[2 * i + 1] = worst;
      }
    }
    __syncwarp();
  }
}


struct RepLayout {
  size_t wl, cand, candd, tab, total;
  __host__ __device__ RepLayout(int M, int rep_cap) {
    size_t o = 0;
    wl = o;    o += (size_t)(M + 4) * 4;
    cand = o;  o += (size_t)(M * M + 4) * 4;
    candd = o; o += (size_t)(M * M + 4) * 4;
    tab = o;   o += (size_t)rep_cap * 4;
    total = (o + 15) & ~(size_t)15;
  }
};

// ---- phase D: one warp per evicted node short of M links (layerNode.replenish, graph.go:172-219) --------------------
template <int CH, int U>
__global__ void replenish_kernel(WaveParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int M = p.M, m = p.m;
  const RepLayout lay(M, p.rep_cap);
  unsigned char *base = smem + lay.total * warp;
  int *wl = reinterpret_cast<int *>(base + lay.wl);
  int *cand = reinterpret_cast<int *>(base + lay.cand);
  float *candd = reinterpret_cast<float *>(base + lay.candd);
  int *nl = reinterpret_cast<int *>(candd);  // [deg][M] staging of the neighbours' lists, dead before candd is written
  uint32_t *tab = reinterpret_cast<uint32_t *>(base + lay.tab);
  if (*p.error) return;
  const int nrep = p.wave_ctr[1];
  unsigned long long c_repl = 0;
  for (int idx = blockIdx.x * nw + warp; idx < nrep; idx += gridDim.x * nw) {
    const int l = p.rep_rec[2 * idx], w = p.rep_rec[2 * idx + 1];
    int deg = warp_load_list(p.list(l, w), M, wl, lane);
    for (int i = lane; i < p.rep_cap; i += 32) tab[i] = ~0u;
    for (int i = 0; i < deg; i++) {  // independent gathers of the neighbours' lists
      const int32_t *ln = p.list(l, wl[i]);
      for (int j = lane; j < M; j += 32) nl[i * M + j] = ln[j];
    }
    __syncwarp();
    Visited RV{tab, (uint32_t)p.rep_cap - 1, 0};
    if (lane == 0) RV.insert((uint32_t)w);  // graph.go:183-190: visited = {n} U neighbours(n)
    for (int i = lane; i < deg; i += 32) RV.insert((uint32_t)wl[i]);
    __syncwarp();
    int nc = 0;  // graph.go:193-210: neighbours of neighbours, ascending key order at both levels
    for (int i = 0; i < deg; i++) {
      const int *ln = nl + i * M;
      for (int b0 = 0; b0 < M; b0 += 32) {
        int v = -1;
        if (b0 + lane < M) v = ln[b0 + lane];
        const unsigned vm = __ballot_sync(FULL, v >= 0);
        if (!vm) break;
        const bool fresh = v >= 0 && RV.insert((uint32_t)v);
        const unsigned fm = __ballot_sync(FULL, fresh);
        if (fresh) cand[nc + __popc(fm & ((1u << lane) - 1))] = v;
        nc += __popc(fm);
        if (vm != FULL) break;
      }
    }
    __syncwarp();
    if (nc > 0) {
      GQuery<CH> Q;
      group_load_query<32, CH>(Q, reinterpret_cast<const float *>(p.vec + (size_t)w * p.row_chunks), true, p.dim,
                               p.row_chunks, lane);
      warp_distances<CH, U>(p, cand, nc, Q, 1 /*CosineDistance, graph.go:204*/, candd, lane);
      c_repl += nc;
      // graph.go:213-218: pop best-first until the list is full.  The first pop of a heap filled by pushes is the
      // first-pushed global minimum: exact shortcut when one pop suffices and no distance is NaN.
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
        deg = warp_sorted_insert(wl, deg, cand[bi], p.keys, lane);
      } else {
        if (lane == 0) {  // container/heap pushes, the arrays themselves are the heap storage
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
          const int best = cand[0];
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
          deg = warp_sorted_insert(wl, deg, best, p.keys, lane);
        }
      }
    }
    int32_t *st = p.staged + (size_t)idx * (M + 1);
    for (int i = lane; i < M; i += 32) st[i] = i < deg ? wl[i] : -1;
    if (lane == 0) st[M] = deg;
    __syncwarp();
  }
  if (lane == 0 && c_repl) atomicAdd(&p.counters[4], c_repl);
}

// ---- phase E: staged lists replace the old ones ---------------------------------------------------------------------
__global__ void __launch_bounds__(WV_WARPS * 32) scatter_kernel(WaveParams p) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (*p.error) return;
  const int nrep = p.wave_ctr[1];
  for (int idx = blockIdx.x * WV_WARPS + warp; idx < nrep; idx += gridDim.x * WV_WARPS) {
    const int l = p.rep_rec[2 * idx], w = p.rep_rec[2 * idx + 1];
    int32_t *lw = p.list(l, w);
    const int32_t *st = p.staged + (size_t)idx * (p.M + 1);
    for (int i = lane; i < p.M; i += 32) lw[i] = st[i];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    p.wave_ctr[2]++;
    p.wave_ctr[4] = 0;
  }
}


// == ora_wave_len (oracle/ora_graph.c): nodes of the next wave
static int64_t wave_len(int64_t cur_rows, int n_layers, const int32_t *levels, int64_t remaining, int64_t wave_max) {
  if (remaining <= 0) return 0;
  if (n_layers == 0) return 1;
  int64_t cap = cur_rows / 16;
  if (cap < 1) cap = 1;
  if (cap > wave_max) cap = wave_max;
  int64_t w = 0;
  while (w < remaining && w < cap) {
    if (levels[w] >= n_layers) {
      if (w == 0) w = 1;
      break;
    }
    w++;
  }
  return w;
}

static int pow2_ge(int v) {
  int r = 1;
  while (r < v) r <<= 1;
  return r;
}

template <int CH, int U>
static int launch_commit(b200hnsw *h, WaveParams &p, cudaStream_t s) {
  const int M = p.M;
  const int blocks_nodes = (p.w + WV_WARPS - 1) / WV_WARPS;
  const int blocks_slots = (int)((p.n_slots + WV_WARPS - 1) / WV_WARPS);
  link_kernel<<<blocks_nodes, WV_WARPS * 32, 0, s>>>(p);
  const int blocks_req = (int)((p.n_slots * M + 255) / 256);
  group_kernel<<<blocks_req, 256, 0, s>>>(p, 0);
  group_kernel<<<blocks_req, 256, 0, s>>>(p, 1);
  targets_kernel<CH, U><<<blocks_slots, WV_WARPS * 32, TargetSmem::bytes(M) * WV_WARPS, s>>>(p);
  // evictions <= requests; the grids below stride over the device-side counts
  long long max_items = p.n_slots * M;
  int grid_e = (int)std::min<long long>((max_items + WV_WARPS - 1) / WV_WARPS, (long long)h->sm_count * 8);
  backlink_kernel<<<grid_e, WV_WARPS * 32, (size_t)WV_WARPS * (M + 4) * 4, s>>>(p);
  RepLayout lay(M, p.rep_cap);
  int rw = (int)std::min<size_t>(WV_WARPS, (h->smem_optin - 1024) / lay.total);
  if (rw < 1) {
    set_error("add: M=%d needs %zu B of shared memory per replenish warp (> %zu)", M, lay.total, h->smem_optin);
    return B200HNSW_ERR_UNSUPPORTED;
  }
  auto rk = replenish_kernel<CH, U>;
  B200_CUDA(cudaFuncSetAttribute(rk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(lay.total * rw)));
  int grid_r = (int)std::min<long long>((max_items + rw - 1) / rw, (long long)h->sm_count * 8);
  rk<<<grid_r, rw * 32, lay.total * rw, s>>>(p);
  scatter_kernel<<<grid_e, WV_WARPS * 32, 0, s>>>(p);
  B200_CUDA(cudaGetLastError());
  return 0;
}

static int launch_commit_any(b200hnsw *h, WaveParams &p, cudaStream_t s) {
  int ch = (p.row_chunks + 31) / 32;
  if (ch <= 1) return launch_commit<1, 8>(h, p, s);
  if (ch <= 2) return launch_commit<2, 4>(h, p, s);
  if (ch <= 4) return launch_commit<4, 4>(h, p, s);
  if (ch <= 6) return launch_commit<6, 2>(h, p, s);
  if (ch <= 8) return launch_commit<8, 2>(h, p, s);
  if (ch <= 12) return launch_commit<12, 1>(h, p, s);
  if (ch <= 16) return launch_commit<16, 1>(h, p, s);
  set_error("add: dim %d > 2048 not supported", p.dim);
  return B200HNSW_ERR_UNSUPPORTED;
}

}  // namespace b200

using namespace b200;

extern "C" int b200hnsw_add_batch_waves(b200hnsw_t *h, const float *vectors, const int64_t *keys,
                                        const int32_t *levels, int64_t n, int32_t m, int32_t ef, int32_t metric,
                                        int64_t wave_max, b200hnsw_stats *stats) {
  if (stats) memset(stats, 0, sizeof *stats);
  if (!h || n < 0 || (n > 0 && (!vectors || !keys || !levels)) || wave_max < 1) {
    set_error("add_waves: bad argument");
    return B200HNSW_ERR_BAD_ARG;
  }
  if (m <= 0 || m > h->M) {
    set_error("M must be greater than 0 and <= max_m=%d, got %d", h->M, m);
    return B200HNSW_ERR_BAD_ARG;
  }
  if (ef <= 0 || ef > B200HNSW_MAX_EF) {
    set_error("EfSearch must be greater than 0 and <= %d, got %d", B200HNSW_MAX_EF, ef);
    return B200HNSW_ERR_BAD_ARG;
  }
  if (metric != B200HNSW_EUCLIDEAN && metric != B200HNSW_COSINE) {
    set_error("unknown metric %d", metric);
    return B200HNSW_ERR_BAD_ARG;
  }
  if (h->n_dead) {
    set_error("add_batch_waves: the graph holds deleted rows; Add after Delete goes through b200hnsw_add_batch");
    return B200HNSW_ERR_UNSUPPORTED;
  }
  if (n == 0) return 0;
  if (h->n + n > 0x7fffffffLL) {
    set_error("add: more than 2^31-1 rows");
    return B200HNSW_ERR_UNSUPPORTED;
  }
  int64_t add_slots = 0;
  for (int64_t i = 0; i < n; i++) {
    if (levels[i] < 0 || levels[i] >= B200HNSW_MAX_LAYERS) {
      set_error("invalid level: %d", levels[i]);
      return B200HNSW_ERR_BAD_ARG;
    }
    add_slots += levels[i];
  }
  B200_CUDA(cudaSetDevice(h->cfg.device));
  CallCtx *c = h->acquire();
  if (!c) return B200HNSW_ERR_CUDA;
  cudaStream_t s = c->stream;
  auto fail = [&](int code) {
    h->release(c);
    return code;
  };
  int rc = h->ensure_capacity(h->n + n, h->nU + add_slots, s);
  if (rc) return fail(rc);
  if ((rc = h->ensure_inboxes(s))) return fail(rc);
  const int dim = h->cfg.dim, rf = h->row_floats, M = h->M;

  // ---- host-side schedule: waves, slots, layers and entries are all functions of `levels` ---------------------
  struct Wave {
    int64_t start, w, n_slots;
    int n_layers;
    int32_t entries[B200HNSW_MAX_LAYERS];
    bool kernels;
  };
  std::vector<Wave> waves;
  std::vector<int32_t> upper_off((size_t)n, -1);
  std::vector<long long> slot_rel((size_t)n, 0);
  int64_t nU = h->nU, max_slots = 1, max_w = 1;
  {
    int nl = h->n_layers;
    int32_t ent[B200HNSW_MAX_LAYERS];
    memcpy(ent, h->entry, sizeof ent);
    int64_t i = 0;
    while (i < n) {
      Wave wv;
      wv.start = i;
      wv.w = wave_len(h->n + i, nl, levels + i, n - i, wave_max);
      wv.n_layers = nl;
      memcpy(wv.entries, ent, sizeof ent);
      wv.kernels = nl > 0;
      int64_t ns = 0;
      for (int64_t j = 0; j < wv.w; j++) {
        int L = levels[i + j];
        if (L > 0) {
          upper_off[i + j] = (int32_t)nU;
          nU += L;
        }
        slot_rel[i + j] = ns;
        ns += L + 1;
        for (int l = nl; l <= L; l++) ent[l] = (int32_t)(h->n + i + j);  // opens the layer: becomes its entry
        if (L + 1 > nl) nl = L + 1;
      }
      wv.n_slots = ns;
      max_slots = std::max(max_slots, ns);
      max_w = std::max(max_w, wv.w);
      waves.push_back(wv);
      i += wv.w;
    }
  }
  // ---- device scratch ------------------------------------------------------------------------------------------------
  auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
  const size_t items = (size_t)max_slots * M;
  size_t need = al(4 * (size_t)n) + al(8 * (size_t)n) + al(4 * (size_t)max_slots) * 3 + al(4 * items) * 5 +
                al(12 * items) + al(8 * items) + al(4 * items * (M + 1)) + 1024;
  if ((rc = c->reserve_device(need))) return fail(rc);
  char *sp = (char *)c->d_scratch;
  auto carve = [&](size_t bytes) {
    char *r = sp;
    sp += al(bytes);
    return r;
  };
  int32_t *d_levels = (int32_t *)carve(4 * (size_t)n);
  long long *d_slot_rel = (long long *)carve(8 * (size_t)n);
  WaveParams p{};
  p.slot_cnt = (int32_t *)carve(4 * (size_t)max_slots);
  p.slot_xrow = (int32_t *)carve(4 * (size_t)max_slots);
  p.slot_layer = (int32_t *)carve(4 * (size_t)max_slots);
  p.slot_rows = (int32_t *)carve(4 * items);
  p.slot_dists = (float *)carve(4 * items);
  p.req_pos = (int32_t *)carve(4 * items);
  p.req_arena = (int32_t *)carve(4 * items);
  p.ev_next = (int32_t *)carve(4 * items);
  p.ev_rec = (int32_t *)carve(12 * items);
  p.rep_rec = (int32_t *)carve(8 * items);
  p.staged = (int32_t *)carve(4 * items * (M + 1));
  char *ctl = carve(256);
  unsigned long long *d_ctr = (unsigned long long *)ctl;
  int *d_err = (int *)(ctl + 64);
  int *d_wave_ctr = (int *)(ctl + 128);
  cudaError_t e;
  // The vectors go up in chunks on a copy stream, one chunk ahead of the waves that need them: the host stages chunk
  // j+1 (pageable memory goes through the driver's bounce buffers) while the device runs the waves of chunk j.  Chunks
  // start small, because the first waves are (a wave never exceeds 1/16 of the rows already present).
  if ((rc = c->ensure_sub(2))) return fail(rc);
  cudaStream_t cs = c->sub[0];
  std::vector<int64_t> cut{0};
  {
    const int64_t big = std::max<int64_t>(4096, ((int64_t)32 << 20) / ((int64_t)sizeof(float) * dim));
    int64_t len = std::min<int64_t>(big, 8192);
    while (cut.back() < n) {
      cut.push_back(std::min<int64_t>(n, cut.back() + len));
      len = std::min<int64_t>(big, len * 2);
    }
  }
  const int n_chunks = (int)cut.size() - 1;
  auto upload = [&](int j) -> cudaError_t {
    cudaError_t r = cudaMemcpy2DAsync(h->d.vec + (size_t)(h->n + cut[j]) * rf, sizeof(float) * rf,
                                      vectors + (size_t)cut[j] * dim, sizeof(float) * dim, sizeof(float) * dim,
                                      (size_t)(cut[j + 1] - cut[j]), cudaMemcpyHostToDevice, cs);
    // rows that can yield NaN distances are flagged as they land (traverse.cuh reads the device word)
    if (r == cudaSuccess && h->flag_rows(h->n + cut[j], cut[j + 1] - cut[j], cs)) r = cudaGetLastError();
    return r != cudaSuccess ? r : cudaEventRecord(c->sub_ev[j & 1], cs);
  };
  // the slab may have just been grown on s (ensure_capacity): the copy stream starts behind that
  if ((e = cudaEventRecord(c->ev0, s)) != cudaSuccess || (e = cudaStreamWaitEvent(cs, c->ev0, 0)) != cudaSuccess ||
      (e = upload(0)) != cudaSuccess)
    return fail(cuda_fail(e, "add_waves staging"));
  if ((e = cudaMemcpyAsync(h->d.keys + h->n, keys, sizeof(int64_t) * (size_t)n, cudaMemcpyHostToDevice, s)) !=
          cudaSuccess ||
      (e = cudaMemcpyAsync(h->d.upper_off + h->n, upper_off.data(), sizeof(int32_t) * (size_t)n,
                           cudaMemcpyHostToDevice, s)) != cudaSuccess ||
      (e = cudaMemcpyAsync(d_levels, levels, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, s)) != cudaSuccess ||
      (e = cudaMemcpyAsync(d_slot_rel, slot_rel.data(), sizeof(long long) * (size_t)n, cudaMemcpyHostToDevice, s)) !=
          cudaSuccess ||
      (e = cudaMemsetAsync(ctl, 0, 256, s)) != cudaSuccess || (e = cudaEventRecord(c->ev0, s)) != cudaSuccess)
    return fail(cuda_fail(e, "add_waves staging"));
  p.vec = reinterpret_cast<float4 *>(h->d.vec);
  p.keys = h->d.keys;
  p.adj0 = h->d.adj0;
  p.adjU = h->d.adjU;
  p.upper_off = h->d.upper_off;
  p.row_chunks = rf / 4;
  p.M = M;
  p.dim = dim;
  p.cap = h->d.cap;
  p.in_cnt = h->d.inbox;
  p.in_base = h->d.inbox + h->d.inbox_n;
  p.ev_head = h->d.inbox + 2 * h->d.inbox_n;
  p.ev_owner = h->d.inbox + 3 * h->d.inbox_n;
  p.m = m;
  p.metric = metric;
  p.wave_ctr = d_wave_ctr;
  p.counters = d_ctr;
  p.error = d_err;
  p.rep_cap = pow2_ge(2 * (M * M + M + 2));
  int launches = 0, kernel_waves = 0;
  const int n_layers_before = h->n_layers;
  size_t wi = 0;
  for (int j = 0; j < n_chunks; j++) {
   if (j + 1 < n_chunks && (e = upload(j + 1)) != cudaSuccess) return fail(cuda_fail(e, "add_waves staging"));
   if ((e = cudaStreamWaitEvent(s, c->sub_ev[j & 1], 0)) != cudaSuccess)  // rows below cut[j + 1] are in HBM
     return fail(cuda_fail(e, "add_waves staging"));
   for (; wi < waves.size() && waves[wi].start + waves[wi].w <= cut[j + 1]; wi++) {
    const Wave &wv = waves[wi];
    if (!wv.kernels) continue;
    // phase S: K1 with the Add sink; it must see exactly the layers that exist before the wave
    h->n_layers = wv.n_layers;
    TraverseParams t{};
    t.queries = h->d.vec + (size_t)(h->n + wv.start) * rf;
    t.q_stride = rf;
    t.B = wv.w;
    t.k = m;
    t.k_upper = m;  // graph.go:500: search(M, ...) on every layer
    t.ef = ef;
    t.metric = metric;
    for (int l = 0; l < B200HNSW_MAX_LAYERS; l++) t.entries[l] = l < wv.n_layers ? wv.entries[l] : -1;
    t.levels = d_levels + wv.start;
    t.slot_off = d_slot_rel + wv.start;
    t.slot_cnt = p.slot_cnt;
    t.slot_rows = p.slot_rows;
    t.slot_dists = p.slot_dists;
    t.sink_stride = M;
    t.counters = d_ctr;
    t.error = d_err;
    if ((e = cudaMemsetAsync(p.slot_cnt, 0, 4 * (size_t)wv.n_slots, s)) != cudaSuccess) {
      h->n_layers = n_layers_before;
      return fail(cuda_fail(e, "add_waves memset"));
    }
    if (h->build_mode == B200HNSW_MODE_EXTENDED) {
      // NOT the reference: the textbook layer search (traverse_ext.cuh) with k = M, ef = EfSearch on every layer
      const int efk = ef > m ? ef : m;
      int vis = ext_visited_default(efk, M);
      const int mx = max_visited_cap_ext(h, efk);
      rc = launch_traverse_ext(h, t, ef, vis < mx ? vis : mx, s);
    } else {
      rc = launch_traverse(h, t, 1024, s);
    }
    h->n_layers = n_layers_before;
    if (rc) return fail(rc);
    p.first_row = (int)(h->n + wv.start);
    p.w = (int)wv.w;
    p.levels = d_levels + wv.start;
    p.slot_off = d_slot_rel + wv.start;
    p.n_slots = wv.n_slots;
    if ((rc = launch_commit_any(h, p, s))) return fail(rc);
    launches += 8;
    kernel_waves++;
   }
  }
  unsigned long long ctr[8];
  int err = 0, wave_ctr[4] = {0, 0, 0, 0};
  if ((e = cudaEventRecord(c->ev1, s)) != cudaSuccess ||
      (e = cudaMemcpyAsync(ctr, d_ctr, sizeof ctr, cudaMemcpyDeviceToHost, s)) != cudaSuccess ||
      (e = cudaMemcpyAsync(&err, d_err, sizeof err, cudaMemcpyDeviceToHost, s)) != cudaSuccess ||
      (e = cudaMemcpyAsync(wave_ctr, d_wave_ctr, sizeof wave_ctr, cudaMemcpyDeviceToHost, s)) != cudaSuccess ||
      (e = cudaStreamSynchronize(s)) != cudaSuccess)
    return fail(cuda_fail(e, "add_waves kernels"));
  if ((rc = h->refresh_row_flags(s))) return fail(rc);
  // ---- commit the host mirrors for the waves that completed -----------------------------------------------------------
  int done_kernel_waves = wave_ctr[2];
  int64_t done_nodes = 0;
  int kw = 0;
  for (const Wave &wv : waves) {
    if (wv.kernels) {
      if (kw >= done_kernel_waves) break;
      kw++;
    }
    for (int64_t j = 0; j < wv.w; j++) {
      int64_t i = wv.start + j;
      int L = levels[i];
      h->h_keys.push_back(keys[i]);
      h->h_level.push_back(L);
      h->h_dead.push_back(0);
      h->h_upper_off.push_back(upper_off[i]);
      for (int l = 0; l <= L; l++) {
        if (h->layer_size[l] == 0) h->entry[l] = (int32_t)(h->n + i);
        h->layer_size[l]++;
      }
      if (L + 1 > h->n_layers) h->n_layers = L + 1;
      if (L > 0) h->nU = upper_off[i] + L;
    }
    done_nodes = wv.start + wv.w;
  }
  h->n += done_nodes;
  if (stats) {
    stats->dist_evals = (int64_t)ctr[0];
    stats->expansions = (int64_t)ctr[1];
    stats->nbr_reads = (int64_t)ctr[2];
    stats->evict_evals = (int64_t)ctr[3];
    stats->replenish_evals = (int64_t)ctr[4];
    stats->evictions = (int64_t)ctr[5];
    stats->launches = launches;
    cudaEventElapsedTime(&stats->kernel_ms, c->ev0, c->ev1);
  }
  h->release(c);
  if (err == 3) {
    // rows that can yield NaN distances (zero norm under cosine, non-finite components) entered the graph with this
    // batch and a later wave met them with the short candidate queue: the handle now knows (row_flags), so the caller
    // adds the remaining nodes again and they run with the full queue
    set_error("add_waves: rows with zero or non-finite norm entered the graph; %lld of %lld nodes were inserted, add the "
              "rest again", (long long)done_nodes, (long long)n);
    return B200HNSW_ERR_STATE;
  }
  if (err || done_nodes != n) {
    set_error("add_waves: a layer search overflowed the visited set; %lld of %lld nodes were inserted",
              (long long)done_nodes, (long long)n);
    return B200HNSW_ERR_VISITED_OVERFLOW;
  }
  return 0;
}
