// layer_search.cuh -- what every kernel that walks the graph shares: the device-resident view of it, the exact visited set
// and the counters of a layer search.  The searches themselves: traverse.cuh (K1, groups of lanes per query),
// traverse_ext.cuh (K1x) and inserter.cuh: block_layer_search (K2 / K2s, one CTA per inserted node) -- each a restatement
// of layerNode.search (graph.go:94-170) with the order-dependent part replayed serially in ascending-key order on the
// Go-exact heaps (goheap.cuh).
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

}  // namespace b200
