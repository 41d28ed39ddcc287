// traverse_ext.cuh -- K1x: the EXTENDED search mode (B200HNSW_MODE_EXTENDED).  NOT the reference's layerNode.search:
// textbook HNSW layer search (Malkov & Yashunin, Alg. 2) for callers who need recall the reference's
// first-non-improving-hop termination (graph.go:147, 164-166) cannot give.  The specification is the one written out in
// oracle/ora_graph.c (ORA_MODE_EXTENDED) and the kernel is held to it bit for bit:
//
//   key(d, row): distance ascending, ties by row; -0 counts as +0; a NaN distance never enters W (entry excepted)
//   W = the efk = max(ef, k) smallest keys among the nodes evaluated so far, each with an `expanded` flag
//   repeat: smallest unexpanded element of W -> mark, evaluate its unvisited neighbours, merge them into W
//   stop when every element of W is expanded; result = first min(k, |W|) elements, ascending
//
// Because W is "the efk smallest of a set", the order in which the rows of one expansion are merged does not matter:
// nothing here is serial except the walk itself.  ONE WARP PER TRAVERSAL:
//   * the popped node's list is read by the lanes (one id each), filtered through an exact visited set (open
//     addressing in shared memory, atomicCAS) and compacted;
//   * the fresh rows are evaluated 32/G * U at a time: every group of G lanes takes U rows (16 B loads, every row of
//     the round in flight before the first reduction; canonical-order arithmetic of gdist.cuh, so distances are the
//     bits the oracle computes);
//   * W is a sorted array of 64-bit keys (order-preserving distance bits << 32 | row, bit 31 = expanded) in shared
//     memory; a new key is ranked by all lanes at once (ballot + popc) and the tail shifted by all lanes at once;
//   * the neighbour list of every row that enters W on layer 0 is prefetched into L2: it is the list a later
//     expansion reads.
// Bound: HBM gather (E rows of 4*dim bytes per traversal; E is ~10x the reference mode's and rows are shared far less
// between the traversals of a batch).  Upper layers run the same search with k = ef = 1 (greedy descent) for Search, and
// with k = M, ef = EfSearch for the search half of the extended wave Add.
#pragma once
#include "gdist.cuh"
#include "layer_search.cuh"
#include "traverse.cuh"

namespace b200 {

// order-preserving image of a distance: ascending unsigned order == ascending float order (oracle: ord_bits)
__device__ __forceinline__ uint32_t ext_ord_bits(float d) {
  if (d != d) return 0x7fc00000u | 0x80000000u;
  d = __fadd_rn(d, 0.0f);  // -0 -> +0
  uint32_t u = __float_as_uint(d);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ext_ord_float(uint32_t o) {
  return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

constexpr unsigned long long EXT_FLAG = 1ull << 31;            // expanded
constexpr unsigned long long EXT_MASK = ~EXT_FLAG;             // compare keys with the flag removed

struct ExtLayout {  // per-warp shared memory, in bytes
  int w_off, fresh_off, vis_off, warp_bytes;
};
__host__ __device__ inline ExtLayout ext_layout(int efk_max, int M, int vis_cap) {
  ExtLayout L;
  L.w_off = 0;
  int o = ((efk_max + 1 + 1) & ~1) * 8;
  L.fresh_off = o;
  o += ((M + 3) & ~3) * 4;
  o = (o + 15) & ~15;
  L.vis_off = o;
  o += vis_cap * 4;
  L.warp_bytes = o;
  return L;
}

template <int G, int CH, int U, int TW>
__global__ void __launch_bounds__(TW * 32) traverse_ext_kernel(GraphView g, TraverseParams p, int ef_upper,
                                                               int efk_max) {
  constexpr int NG = 32 / G;
  constexpr int RPR = NG * U;  // rows per round
  static_assert(U <= G, "a group returns row u in lane u * (G / U)");
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int gl = lane & (G - 1);
  const int grp = lane / G;
  const ExtLayout lay = ext_layout(efk_max, g.M, p.vis_cap);
  unsigned char *base = smem + (size_t)lay.warp_bytes * warp;
  unsigned long long *W = reinterpret_cast<unsigned long long *>(base + lay.w_off);
  int *fresh = reinterpret_cast<int *>(base + lay.fresh_off);
  uint32_t *vtab = reinterpret_cast<uint32_t *>(base + lay.vis_off);
  const uint32_t vmask = (uint32_t)p.vis_cap - 1;
  const int M = g.M;
  const bool holder = (gl % (G / U)) == 0;  // this lane receives the distance of row gl / (G / U) of its group
  const int my_u = gl / (G / U);

  unsigned long long c_evals = 0, c_exp = 0, c_nbr = 0, c_over = 0;  // warp-uniform; lane 0 reports
  const long long n_warps = (long long)gridDim.x * TW;
  for (long long q = (long long)blockIdx.x * TW + warp; q < p.B; q += n_warps) {
    GQuery<CH> Q;
    group_load_query<G, CH>(Q, p.queries + q * p.q_stride, true, g.dim, g.row_chunks, gl);
    const int lvl = p.levels ? p.levels[q] : -1;
    int elevator = -1;
    unsigned long long elev_key = 0;
    bool failed = false, empty_out = false;
    int wn = 0;
    for (int layer = p.n_layers - 1; layer >= 0 && !failed; layer--) {
      const int kcur = layer > 0 ? p.k_upper : p.k;
      const int efl = layer > 0 ? ef_upper : p.ef;
      const int efk = efl > kcur ? efl : kcur;
      int cur = elevator >= 0 ? elevator : p.entries[layer];
      wn = 0;
      if (elevator >= 0 && p.dead && p.dead[elevator]) {
        // layer.nodes[*elevator] is nil once the key was deleted: nothing is found on this layer or below
        empty_out = true;
        break;
      }
      if (cur < 0) continue;  // a layer without an entry (emptied by Delete): passed through
      // ---- visited set, W = {entry} ---------------------------------------------------------------------------------
      {
        uint4 e4 = make_uint4(~0u, ~0u, ~0u, ~0u);
        uint4 *t4 = reinterpret_cast<uint4 *>(vtab);
        for (uint32_t i = lane; i < (vmask + 1) / 4; i += 32) t4[i] = e4;
      }
      __syncwarp();
      if (elevator >= 0) {
        // the elevator's distance to this query was computed on the layer above (same row, same arithmetic)
        if (lane == 0) W[0] = elev_key & EXT_MASK;
      } else {
        int ids[U];
#pragma unroll
        for (int u = 0; u < U; u++) ids[u] = (u == 0 && grp == 0) ? cur : -1;
        const float d = group_distances<G, CH, U>(g.vec, g.row_chunks, ids, Q, p.metric, gl);
        const float d0 = __shfl_sync(FULL, d, 0);
        if (lane == 0) W[0] = ((unsigned long long)ext_ord_bits(d0) << 32) | (unsigned)cur;
      }
      c_evals++;  // the oracle evaluates the entry on every layer
      if (lane == 0) vtab[(((uint32_t)cur * 2654435761u) >> 7) & vmask] = (uint32_t)cur;
      wn = 1;
      int vcount = 1, cursor = 0;
      __syncwarp();
      // ---- the walk -----------------------------------------------------------------------------------------------------
      for (;;) {
        int ci = -1;
        for (int b0 = cursor & ~31; b0 < wn; b0 += 32) {
          const int i = b0 + lane;
          const bool un = i < wn && i >= cursor && !(W[i] & EXT_FLAG);
          const unsigned m = __ballot_sync(FULL, un);
          if (m) {
            ci = b0 + __ffs(m) - 1;
            break;
          }
        }
        if (ci < 0) break;
        cur = (int)(W[ci] & 0x7fffffffull);
        __syncwarp();
        if (lane == 0) W[ci] |= EXT_FLAG;
        cursor = ci + 1;
        c_exp++;
        const int32_t *list = layer == 0 ? g.adj0 + (size_t)cur * M
                                         : g.adjU + ((size_t)g.upper_off[cur] + (layer - 1)) * M;
        int nfresh = 0;
        for (int b0 = 0; b0 < M; b0 += 32) {
          int nb = -1;
          if (b0 + lane < M) nb = list[b0 + lane];
          const bool valid = nb >= 0;
          const unsigned gv = __ballot_sync(FULL, valid);
          bool fr = false;
          if (valid) {
            uint32_t h = ((uint32_t)nb * 2654435761u) >> 7;
            for (;;) {
              h &= vmask;
              uint32_t old = atomicCAS(&vtab[h], ~0u, (uint32_t)nb);
              if (old == ~0u) {
                fr = true;
                break;
              }
              if (old == (uint32_t)nb) break;
              h++;
            }
          }
          const unsigned gf = __ballot_sync(FULL, fr);
          if (fr) {
            fresh[nfresh + __popc(gf & ((1u << lane) - 1u))] = nb;
            const float4 *row = g.vec + (size_t)nb * g.row_chunks;
            for (int c = 0; c < g.row_chunks; c += 8) prefetch_l2(row + c);
          }
          nfresh += __popc(gf);
          c_nbr += __popc(gv);
          const int in_range = M - b0 < 32 ? M - b0 : 32;
          if (__popc(gv) != in_range) break;  // lists are -1 padded at the end
        }
        vcount += nfresh;
        if (vcount * 2 > (int)vmask) {  // load factor > 1/2: fail loudly, never approximate
          failed = true;
          break;
        }
        __syncwarp();
        c_evals += nfresh;
        for (int r0 = 0; r0 < nfresh; r0 += RPR) {
          int ids[U];
#pragma unroll
          for (int u = 0; u < U; u++) {
            const int j = r0 + grp * U + u;
            ids[u] = j < nfresh ? fresh[j] : -1;
          }
          const float d = group_distances<G, CH, U>(g.vec, g.row_chunks, ids, Q, p.metric, gl);
          int my_id = -1;
#pragma unroll
          for (int u = 0; u < U; u++)
            if (u == my_u) my_id = ids[u];
          unsigned long long mykey = ~0ull;
          bool want = false;
          if (holder && my_id >= 0 && d == d) {
            mykey = ((unsigned long long)ext_ord_bits(d) << 32) | (unsigned)my_id;
            want = wn < efk || mykey < (W[wn - 1] & EXT_MASK);
          }
          unsigned m = __ballot_sync(FULL, want);
          while (m) {
            const int src = __ffs(m) - 1;
            m &= m - 1;
            const unsigned long long key = __shfl_sync(FULL, mykey, src);
            if (wn == efk) {
              if (!(key < (W[wn - 1] & EXT_MASK))) continue;  // W's worst moved since the ballot
              wn--;
              __syncwarp();
            }
            int pos = 0;
            for (int b0 = 0; b0 < wn; b0 += 32) {
              const int i = b0 + lane;
              const bool lt = i < wn && (W[i] & EXT_MASK) < key;
              pos += __popc(__ballot_sync(FULL, lt));
            }
            for (int b0 = (wn - 1) & ~31; b0 >= (pos & ~31) && wn > 0; b0 -= 32) {  // shift [pos, wn) up, top chunk first
              const int i = b0 + lane;
              const bool mv = i >= pos && i < wn;
              unsigned long long v = 0;
              if (mv) v = W[i];
              __syncwarp();
              if (mv) W[i + 1] = v;
              __syncwarp();
            }
            if (lane == 0) {
              W[pos] = key;
              if (layer == 0) prefetch_l2(g.adj0 + (size_t)(key & 0x7fffffffull) * M);
            }
            wn++;
            if (pos < cursor) cursor = pos;
            __syncwarp();
          }
        }
      }
      if (failed) break;
      // ---- end of the layer -------------------------------------------------------------------------------------------
      const int rlen = wn < kcur ? wn : kcur;
      if (!p.out_rows && layer <= lvl) {  // Add sink (graph.go:510): this layer takes the new node
        const long long slot = p.slot_off[q] + layer;
        for (int i = lane; i < rlen; i += 32) {
          const unsigned long long kv = W[i];
          p.slot_rows[slot * p.sink_stride + i] = (int)(kv & 0x7fffffffull);
          p.slot_dists[slot * p.sink_stride + i] = ext_ord_float((uint32_t)(kv >> 32));
        }
        if (lane == 0) p.slot_cnt[slot] = rlen;
      }
      elev_key = W[0];
      elevator = (int)(elev_key & 0x7fffffffull);
      __syncwarp();
    }
    // ---- results -------------------------------------------------------------------------------------------------------
    if (failed) {
      if (lane == 0) {
        *p.error = 1;
        c_over++;
        if (p.out_rows) *search_out(p, q).count = -1;
        else p.slot_cnt[p.slot_off[q]] = -1;
      }
    } else if (p.out_rows) {
      const int rlen = empty_out ? 0 : (wn < p.k ? wn : p.k);
      const OutPtrs o = search_out(p, q);
      for (int i = lane; i < p.k; i += 32) {
        const bool ok = i < rlen;
        const unsigned long long kv = ok ? W[i] : 0;
        o.rows[i] = ok ? (int)(kv & 0x7fffffffull) : -1;
        o.dists[i] = ok ? ext_ord_float((uint32_t)(kv >> 32)) : __int_as_float(0x7fc00000);
      }
      if (lane == 0) *o.count = rlen;
    } else if (empty_out && lane == 0) {
      *p.error = 2;  // Add: "no nodes found in neighborhood search" (graph.go:501-505)
    }
    __syncwarp();
  }
  if (p.counters && lane == 0) {
    if (c_evals) atomicAdd(&p.counters[0], c_evals);
    if (c_exp) atomicAdd(&p.counters[1], c_exp);
    if (c_nbr) atomicAdd(&p.counters[2], c_nbr);
    if (c_over) atomicAdd(&p.counters[3], c_over);
  }
}

}  // namespace b200
