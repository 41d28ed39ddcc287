// goheap.cuh -- the reference's heap.Heap[searchCandidate] (heap/heap.go:42-95) over Go's container/heap,
// as an array in shared memory.  Layout-exact: Graph.Search returns the raw heap array (graph.go:169),
// Max() is the LAST ARRAY SLOT (heap.go:89-91) and PopLast() == Remove(len-1) is a truncation (heap.go:73-75),
// so both membership and order of the result depend on reproducing Go's sift sequences exactly.
//
// One lane (lane 0) of the owning warp executes these; other lanes read the array only after a __syncwarp().
// (Redundant execution by all 32 lanes is NOT safe: lanes are not guaranteed to stay in lockstep.)
// Elements are (dist, id) packed in a uint2.  `less` is strict '<' on dist (graph.go:88-90).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

struct GoHeap {
  uint2 *a;
  int len;

  __device__ __forceinline__ float dist(int i) const { return __uint_as_float(a[i].x); }
  __device__ __forceinline__ int id(int i) const { return (int)a[i].y; }
  __device__ __forceinline__ float min_dist() const { return __uint_as_float(a[0].x); }       // Heap.Min
  __device__ __forceinline__ float max_dist() const { return __uint_as_float(a[len - 1].x); }  // Heap.Max (!)

  // container/heap.Push: append, up(len-1).  The moving element is held in registers and parents are shifted
  // down, which leaves the same array as Go's pairwise swaps.
  __device__ __forceinline__ void push(float d, int id_) {
    int j = len++;
    while (j > 0) {
      int i = (j - 1) >> 1;
      uint2 p = a[i];
      if (!(d < __uint_as_float(p.x))) break;
      a[j] = p;
      j = i;
    }
    a[j] = make_uint2(__float_as_uint(d), (unsigned)id_);
  }

  // container/heap.Pop: swap(0, n-1); down(0, n-1); remove last.  Returns the minimum.
  __device__ __forceinline__ uint2 pop() {
    int n = --len;
    uint2 root = a[0];
    if (n > 0) {
      uint2 mv = a[n];
      float md = __uint_as_float(mv.x);
      int i = 0;
      for (;;) {
        int j1 = 2 * i + 1;
        if (j1 >= n) break;
        int j = j1;
        uint2 c = a[j1];
        int j2 = j1 + 1;
        if (j2 < n) {
          uint2 c2 = a[j2];
          if (__uint_as_float(c2.x) < __uint_as_float(c.x)) {
            j = j2;
            c = c2;
          }
        }
        if (!(__uint_as_float(c.x) < md)) break;
        a[i] = c;
        i = j;
      }
      a[i] = mv;
    }
    return root;
  }

  // Heap.PopLast: Remove(len-1) -> i == n, no swap, no sift: truncation.
  __device__ __forceinline__ void pop_last() { --len; }
};

}  // namespace b200
