// listops.cuh -- warp-level operations on neighbour lists (fixed stride M, sorted by key, -1 padded): the device
// counterpart of Go's map insert / delete on layerNode.neighbors (graph.go:50, 73-78).  Shared by the sequential
// insert kernel (insert_kernel.cu) and the wave-batched one (wave_insert.cu).
#pragma once
#include "dist.cuh"

namespace b200 {

// ---- warp-level list helpers (lists: stride M, sorted by key, -1 padded) -----------------------------------------

// load a list into shared memory; returns its degree (uniform over the warp)
// CG: read past L1 (the speculative Add: another CTA may have rewritten the list since this SM last saw it)
template <bool CG = false>
__device__ __forceinline__ int warp_load_list(const int32_t *lp, int M, int *dst, int lane) {
  int deg = 0;
  for (int base = 0; base < M; base += 32) {
    int v = -1;
    if (base + lane < M) v = CG ? __ldcg(lp + base + lane) : lp[base + lane];
    unsigned mk = __ballot_sync(FULL, v >= 0);
    if constexpr (CG) {
      // a list caught while its owner rewrites it can show a gap; only the entries in front of it are used (the read is
      // thrown away by the commit test anyway, it just must not index anything with -1)
      const int lead = __ffs(~mk) - 1;  // -1 when all 32 are valid
      if (lead >= 0) {
        mk = (1u << lead) - 1u;
        if (lane >= lead) v = -1;
      }
    }
    if (base + lane < M) dst[base + lane] = v;
    deg += __popc(mk);
    if (mk != FULL) break;
  }
  __syncwarp();
  return deg;
}

// write `len` entries from shared memory, pad the rest of the stride with -1
__device__ __forceinline__ void warp_store_list(int32_t *lp, int M, const int *src, int len, int lane) {
  for (int i = lane; i < M; i += 32) lp[i] = i < len ? src[i] : -1;
}

// sorted insert of x into the shared-memory list l[0..deg): returns new degree (unchanged if x is present: Go map
// overwrite, graph.go:50).  Executed by one warp; the result is in l[].
__device__ __forceinline__ int warp_sorted_insert(int *l, int deg, int x, const int64_t *keys, int lane) {
  long long kx = keys[x];
  int pos = 0, eqpos = -1;
  bool present = false;
  for (int base = 0; base < deg; base += 32) {
    int i = base + lane;
    int v = i < deg ? l[i] : -1;
    bool lt = false, eq = false;
    if (v >= 0) {
      present = present || (v == x);
      const long long kv = keys[v];
      lt = kv < kx;
      eq = kv == kx && v != x;  // another row under the same key: the stale row of a key that was deleted and added again
    }
    pos += __popc(__ballot_sync(FULL, lt));
    const unsigned em = __ballot_sync(FULL, eq);
    if (em && eqpos < 0) eqpos = base + __ffs(em) - 1;
  }
  if (__any_sync(FULL, present)) return deg;
  if (eqpos >= 0) {  // n.neighbors[key] = newNode overwrites the entry the map holds under that key (graph.go:50)
    __syncwarp();
    if (lane == 0) l[eqpos] = x;
    __syncwarp();
    return deg;
  }
  // shift right from the end (single pass, each lane moves its element after all have read)
  int moved[2];
  int cntm = 0;
  for (int i = pos + lane; i < deg; i += 32) moved[cntm++] = l[i];
  __syncwarp();
  cntm = 0;
  for (int i = pos + lane; i < deg; i += 32) l[i + 1] = moved[cntm++];
  if (lane == 0) l[pos] = x;
  __syncwarp();
  return deg + 1;
}

// remove x from l[0..deg) if present; returns new degree
__device__ __forceinline__ int warp_list_remove(int *l, int deg, int x, int lane) {
  int pos = -1;
  for (int base = 0; base < deg; base += 32) {
    int i = base + lane;
    bool hit = i < deg && l[i] == x;
    unsigned mk = __ballot_sync(FULL, hit);
    if (mk) {
      pos = base + __ffs(mk) - 1;
      break;
    }
  }
  if (pos < 0) return deg;
  int moved[2];
  int cntm = 0;
  for (int i = pos + 1 + lane; i < deg; i += 32) moved[cntm++] = l[i];
  __syncwarp();
  cntm = 0;
  for (int i = pos + 1 + lane; i < deg; i += 32) l[i - 1] = moved[cntm++];
  __syncwarp();
  return deg - 1;
}

}  // namespace b200
