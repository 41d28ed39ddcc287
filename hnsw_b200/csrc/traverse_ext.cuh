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
//   * W is a sorted array of 64-bit keys (order-preserving distance bits << 32 | row, bit 31 = expanded) held in
//     REGISTERS, EPL entries per lane (position p lives in lane p % 32, slot p / 32): a new key is ranked by all lanes
//     at once (compare + ballot + popc per slot) and the tail moves up by one with a shuffle per slot -- no shared
//     memory traffic and no serial section (v1 kept W in shared memory: 40 % of its instructions were this insertion);
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
  int fresh_off, vis_off, warp_bytes;
};
__host__ __device__ inline ExtLayout ext_layout(int M, int vis_cap) {
  ExtLayout L;
  L.fresh_off = 0;
  int o = ((M + 3) & ~3) * 4;
  o = (o + 15) & ~15;
  L.vis_off = o;
  o += vis_cap * 4;
  L.warp_bytes = o;
  return L;
}

// W[idx] for a warp-uniform idx: slot idx / 32 of lane idx % 32.  The slot is picked by a binary tree of selects over
// statically indexed registers (an equality chain `if (e == slot) v = w[e]` is folded back into an indexed load by the
// compiler, which sends the whole array to local memory).
template <int EPL>
__device__ __forceinline__ unsigned long long ext_w_at(const unsigned long long (&w)[EPL], int idx) {
  static_assert((EPL & (EPL - 1)) == 0, "EPL is a power of two");
  unsigned long long t[EPL];
#pragma unroll
  for (int e = 0; e < EPL; e++) t[e] = w[e];
  const int slot = idx >> 5;
#pragma unroll
  for (int stride = 1; stride < EPL; stride <<= 1) {
#pragma unroll
    for (int i = 0; i + stride < EPL; i += 2 * stride) t[i] = (slot & stride) ? t[i + stride] : t[i];
  }
  return __shfl_sync(FULL, t[0], idx & 31);
}

template <int G, int CH, int U, int EPL, int MINB>
__global__ void __launch_bounds__(32, MINB) traverse_ext_kernel(GraphView g, TraverseParams p, int ef_upper) {
  constexpr int NG = 32 / G;
  constexpr int RPR = NG * U;  // rows per round
  static_assert(U <= G, "a group returns row u in lane u * (G / U)");
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31;
  const int gl = lane & (G - 1);
  const int grp = lane / G;
  const ExtLayout lay = ext_layout(g.M, p.vis_cap);
  int *fresh = reinterpret_cast<int *>(smem + lay.fresh_off);
  uint32_t *vtab = reinterpret_cast<uint32_t *>(smem + lay.vis_off);
  const uint32_t vmask = (uint32_t)p.vis_cap - 1;
  const int vlimit = (p.vis_cap >> 2) * 3;  // load factor 3/4: past it the layer search fails loudly
  const int M = g.M;
  const bool holder = (gl % (G / U)) == 0;  // this lane receives the distance of row gl / (G / U) of its group
  const int my_u = gl / (G / U);

  unsigned long long c_evals = 0, c_exp = 0, c_nbr = 0, c_over = 0;  // warp-uniform; lane 0 reports
  for (long long q = blockIdx.x; q < p.B; q += gridDim.x) {
    GQuery<CH> Q;
    group_load_query<G, CH>(Q, p.queries + q * p.q_stride, true, g.dim, g.row_chunks, gl);
    const int lvl = p.levels ? p.levels[q] : -1;
    int elevator = -1;
    unsigned long long elev_key = 0;
    bool failed = false, empty_out = false;
    int wn = 0;
    unsigned long long w[EPL];
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
#pragma unroll
      for (int e = 0; e < EPL; e++) w[e] = ~0ull;
      unsigned long long k0;
      if (elevator >= 0) {
        // the elevator's distance to this query was computed on the layer above (same row, same arithmetic)
        k0 = elev_key & EXT_MASK;
      } else {
        int ids[U];
#pragma unroll
        for (int u = 0; u < U; u++) ids[u] = (u == 0 && grp == 0) ? cur : -1;
        const float d = group_distances<G, CH, U>(g.vec, g.row_chunks, ids, Q, p.metric, gl);
        const float d0 = __shfl_sync(FULL, d, 0);
        k0 = ((unsigned long long)ext_ord_bits(d0) << 32) | (unsigned)cur;
      }
      if (lane == 0) w[0] = k0;
      c_evals++;  // the oracle evaluates the entry on every layer
      __syncwarp();
      if (lane == 0) vtab[(((uint32_t)cur * 2654435761u) >> 7) & vmask] = (uint32_t)cur;
      wn = 1;
      int vcount = 1;
      __syncwarp();
      // ---- the walk -----------------------------------------------------------------------------------------------------
      for (;;) {
        int ci = -1;
#pragma unroll
        for (int e = 0; e < EPL; e++) {
          if (ci < 0 && e * 32 < wn) {
            const bool un = e * 32 + lane < wn && !(w[e] & EXT_FLAG);
            const unsigned m = __ballot_sync(FULL, un);
            if (m) ci = e * 32 + __ffs(m) - 1;
          }
        }
        if (ci < 0) break;
        cur = (int)(ext_w_at<EPL>(w, ci) & 0x7fffffffull);
#pragma unroll
        for (int e = 0; e < EPL; e++)  // every slot is rewritten (see ext_w_at on why not `if (e == slot)`)
          w[e] |= (unsigned long long)(lane == (ci & 31) && ((ci >> 5) ^ e) == 0) << 31;
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
        if (vcount > vlimit) {  // fail loudly, never approximate: the host retries with a larger table
          failed = true;
          break;
        }
        __syncwarp();
        c_evals += nfresh;
        unsigned long long worst = wn == efk ? (ext_w_at<EPL>(w, wn - 1) & EXT_MASK) : ~0ull;
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
            want = mykey < worst;
          }
          unsigned m = __ballot_sync(FULL, want);
          while (m) {
            const int src = __ffs(m) - 1;
            m &= m - 1;
            const unsigned long long key = __shfl_sync(FULL, mykey, src);
            if (!(key < worst)) continue;  // W's worst moved since the ballot
            int pos = 0;
#pragma unroll
            for (int e = 0; e < EPL; e++)
              if (e * 32 < wn) pos += __popc(__ballot_sync(FULL, (w[e] & EXT_MASK) < key));
            // [pos, wn) moves up by one, top slot first; the element pushed past efk - 1 falls off
#pragma unroll
            for (int e = EPL - 1; e >= 0; e--) {
              if (e * 32 + 31 >= pos && e * 32 <= wn) {
                unsigned long long up = __shfl_up_sync(FULL, w[e], 1);
                if (e > 0) {
                  const unsigned long long carry = __shfl_sync(FULL, w[e - 1], 31);
                  if (lane == 0) up = carry;
                }
                const int pp = e * 32 + lane;
                if (pp > pos) w[e] = up;
                else if (pp == pos) w[e] = key;
                if (pp >= efk) w[e] = ~0ull;
              }
            }
            if (wn < efk) wn++;
            if (wn == efk) worst = ext_w_at<EPL>(w, wn - 1) & EXT_MASK;
            if (lane == 0 && layer == 0) prefetch_l2(g.adj0 + (size_t)(key & 0x7fffffffull) * M);
          }
        }
      }
      if (failed) break;
      // ---- end of the layer -------------------------------------------------------------------------------------------
      const int rlen = wn < kcur ? wn : kcur;
      if (!p.out_rows && layer <= lvl) {  // Add sink (graph.go:510): this layer takes the new node
        const long long slot = p.slot_off[q] + layer;
#pragma unroll
        for (int e = 0; e < EPL; e++) {
          const int i = e * 32 + lane;
          if (i < rlen) {
            p.slot_rows[slot * p.sink_stride + i] = (int)(w[e] & 0x7fffffffull);
            p.slot_dists[slot * p.sink_stride + i] = ext_ord_float((uint32_t)(w[e] >> 32));
          }
        }
        if (lane == 0) p.slot_cnt[slot] = rlen;
      }
      elev_key = __shfl_sync(FULL, w[0], 0);
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
#pragma unroll
      for (int e = 0; e < EPL; e++) {
        const int i = e * 32 + lane;
        if (i < p.k) {
          const bool ok = i < rlen;
          o.rows[i] = ok ? (int)(w[e] & 0x7fffffffull) : -1;
          o.dists[i] = ok ? ext_ord_float((uint32_t)(w[e] >> 32)) : __int_as_float(0x7fc00000);
        }
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
