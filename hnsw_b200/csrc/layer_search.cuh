// layer_search.cuh -- layerNode.search (graph.go:94-170) for one query, executed by one warp.
//
// Data-parallel parts (lanes cooperate): reading the neighbour list, the visited test, the gather of the
// unvisited neighbours' vectors (one 16 B load per lane per 512 B of row, up to U rows in flight) and the
// distance reduction.  Order-dependent parts (queue updates, the `improved` flag) are replayed serially in
// ascending-key order by lane 0 on shared-memory heaps (goheap.cuh), exactly as Go does.
#pragma once
#include "dist.cuh"
#include "goheap.cuh"

namespace b200 {

// Device-resident graph (see DESIGN.md "HBM layout").  Lists are fixed-stride, sorted by key, padded with -1.
struct GraphView {
  const float4 *vec;         // [cap][row_chunks] float4, rows 128 B aligned, zero padded
  const int32_t *adj0;       // [cap][M]           layer-0 lists
  const int32_t *adjU;       // [capU][M]          upper-layer lists, slot = upper_off[row] + layer - 1
  const int32_t *upper_off;  // [cap]              first upper slot of a row, -1 if level 0
  int32_t row_chunks;        // padded_dim / 4
  int32_t M;                 // list stride
  int32_t dim;
};

// Exact visited set: open addressing over 32-bit row ids; lives in shared memory (or global scratch when the
// host asks for more capacity than fits).  Empty slot = 0xffffffff.
struct Visited {
  uint32_t *tab;
  uint32_t mask;  // capacity - 1 (power of two)
  int count;

  __device__ __forceinline__ void clear(int lane) {
    uint4 e = make_uint4(~0u, ~0u, ~0u, ~0u);
    uint4 *t4 = reinterpret_cast<uint4 *>(tab);
    for (uint32_t i = lane; i < (mask + 1) / 4; i += 32) t4[i] = e;
    count = 0;
    __syncwarp();
  }
  // true if `id` was not yet in the set (and is now)
  __device__ __forceinline__ bool insert(uint32_t id) {
    uint32_t h = (id * 2654435761u) >> 7;
    for (;;) {
      h &= mask;
      uint32_t old = atomicCAS(&tab[h], ~0u, id);
      if (old == ~0u) return true;
      if (old == id) return false;
      h++;
    }
  }
};

struct SearchCounters {
  unsigned long long evals, expansions, nbr_reads;
};

// Returns the length of the result heap or -1 on visited-set overflow.
// C (capacity ef+1) and R (capacity k) are shared-memory heaps owned by this warp.  ONLY LANE 0 touches them
// (the other lanes' C.len / R.len are meaningless); the popped node travels to the other lanes by shuffle, and
// the caller must __syncwarp() before other lanes read R's array (this function ends with one).
template <int CPL, int U>
__device__ __forceinline__ int layer_search_ref(const GraphView &g, int layer, int entry, int k, int ef, int metric,
                                                const QueryRegs<CPL> &Q, GoHeap &C, GoHeap &R, Visited &V,
                                                SearchCounters &ctr, int lane) {
  C.len = 0;
  R.len = 0;
  V.clear(lane);
  int next_cur = -1;  // lane 0: node to expand next, -1 = stop
  // graph.go:109-123: candidates.Push(entry); result.Push(candidates.Min()); visited[entry] = true
  {
    float4 r[CPL];
    load_row<CPL>(r, g.vec + (size_t)entry * g.row_chunks, g.row_chunks, lane);
    float d0 = finish_distance<CPL>(r, Q, metric);
    ctr.evals++;
    if (lane == 0) {
      C.push(d0, entry);
      R.push(d0, entry);
      V.insert((uint32_t)entry);
      next_cur = (int)C.pop().y;  // graph.go:125-127, first iteration
    }
    V.count = 1;
    __syncwarp();
  }
  const int M = g.M;
  for (;;) {
    int cur = __shfl_sync(FULL, next_cur, 0);
    if (cur < 0) break;
    bool improved = false;  // lane 0 only
    const int32_t *list;
    if (layer == 0) {
      list = g.adj0 + (size_t)cur * M;
    } else {
      int off = g.upper_off[cur];
      list = g.adjU + ((size_t)off + (layer - 1)) * M;
    }
    ctr.expansions++;
    // graph.go:137-144: neighbours in ascending key order (stored that way), skipping visited ones.
    for (int base = 0; base < M; base += 32) {
      int nb = -1;
      if (base + lane < M) nb = list[base + lane];
      bool valid = nb >= 0;
      unsigned vmask = __ballot_sync(FULL, valid);
      if (vmask == 0) break;
      ctr.nbr_reads += __popc(vmask);
      bool fresh = valid && V.insert((uint32_t)nb);
      unsigned fmask = __ballot_sync(FULL, fresh);
      V.count += __popc(fmask);
      if (V.count * 2 > (int)V.mask) return -1;  // load factor > 1/2: scratch too small, fail loudly
      while (fmask) {
        // take the next <= U unvisited neighbours, issue all their row loads, then reduce
        int ids[U];
        float4 rows[U][CPL];
        int cnt = 0;
#pragma unroll
        for (int u = 0; u < U; u++) {
          ids[u] = -1;
          if (fmask) {
            int src = __ffs(fmask) - 1;
            fmask &= fmask - 1;
            ids[u] = __shfl_sync(FULL, nb, src);
            cnt++;
          }
        }
#pragma unroll
        for (int u = 0; u < U; u++)
          if (ids[u] >= 0) load_row<CPL>(rows[u], g.vec + (size_t)ids[u] * g.row_chunks, g.row_chunks, lane);
        float dist[U];
#pragma unroll
        for (int u = 0; u < U; u++)
          if (ids[u] >= 0) dist[u] = finish_distance<CPL>(rows[u], Q, metric);
        ctr.evals += cnt;
        // graph.go:146-159, serially in key order, by lane 0
        if (lane == 0) {
#pragma unroll
          for (int u = 0; u < U; u++) {
            if (ids[u] < 0) continue;
            float d = dist[u];
            improved = improved || (R.len > 0 && d < R.min_dist());  // graph.go:147
            if (R.len < k) {
              R.push(d, ids[u]);
            } else if (d < R.max_dist()) {  // result.Max() is the last slot
              R.pop_last();
              R.push(d, ids[u]);
            }
            C.push(d, ids[u]);
            if (C.len > ef) C.pop_last();  // graph.go:157-159
          }
        }
      }
    }
    if (lane == 0) {
      if (!improved && R.len >= k) next_cur = -1;              // graph.go:164-166
      else next_cur = C.len > 0 ? (int)C.pop().y : -1;         // graph.go:125-127
    }
  }
  int rlen = __shfl_sync(FULL, R.len, 0);
  __syncwarp();
  return rlen;
}

}  // namespace b200
