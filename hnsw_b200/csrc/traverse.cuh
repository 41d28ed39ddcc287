// traverse.cuh -- K1: the layer descent of Graph.Search / BatchSearch (graph.go:571-593, 1075-1106) and of the search
// half of Graph.Add (graph.go:475-508) as a STEP MACHINE.
//
// A warp serves 32/G traversals at once, one per group of G lanes.  Every iteration of the warp-uniform loop lets each
// group do one step of its own traversal, whatever layer it is in:
//     INIT   : evaluate the layer's entry point (graph.go:107-123)
//     EXPAND : read the popped node's neighbour list (ascending key order, graph.go:137-138), drop the visited ones,
//              evaluate the rest U rows at a time (16 B loads, all U rows in flight before the first reduction)
// then the group's leader lane replays the order-dependent part (graph.go:146-166) on the Go-exact shared-memory
// heaps (goheap.cuh).  The leaders of all groups run that serial code together, so its cost is shared by 32/G
// traversals, and a group never waits for another group's layer to end.
#pragma once
#include "gdist.cuh"
#include "goheap.cuh"
#include "layer_search.cuh"

namespace b200 {

struct TraverseParams {
  const float *queries;   // query t starts at queries + t * q_stride
  long long q_stride;
  long long B;
  int k;         // result size on layer 0
  int k_upper;   // result size on layers > 0: 1 for Search (graph.go:579), M for Add (graph.go:500)
  int ef, metric, n_layers;
  int entries[B200HNSW_MAX_LAYERS];
  // Search sink: layer-0 results in heap-array order
  int32_t *out_rows;    // [B][k]
  float *out_dists;     // [B][k]
  int32_t *out_counts;  // [B]
  // segmented Search sink (out_seg_len > 0): queries are cut in runs of out_seg_len; run s writes, at
  // out_rows + s * out_seg_stride (int32 elements): rows [seg][k] | dists f32 [seg][k] | counts [seg] -- one
  // self-contained block per destination rank of the multi-GPU all-to-all
  long long out_seg_len, out_seg_stride;
  // scattered segmented sink (n_seg_ptr > 0): block s starts at seg_ptr[s] instead of out_rows + s * out_seg_stride --
  // the pointers may name PEER memory (another GPU's receive buffer, mapped over NVLink), so the traversal kernel
  // delivers each rank's share of the results itself and no collective follows it
  int n_seg_ptr;
  int32_t *seg_ptr[B200HNSW_MAX_PEERS];
  // Add sink (out_rows == nullptr): results of every layer <= levels[t] go to slot slot_off[t] + layer
  const int32_t *levels;     // [B]
  const long long *slot_off; // [B]
  int32_t *slot_cnt;         // [slots]
  int32_t *slot_rows;        // [slots][sink_stride]
  float *slot_dists;         // [slots][sink_stride]
  int sink_stride;
  unsigned long long *counters;  // [4]: evals, expansions, nbr_reads, overflowed traversals (may be null)
  int *error;                    // set to 1 on visited overflow
  int vis_cap;                   // visited-table entries per group (power of two)
  int heap_c, heap_r, fresh_cap; // per-group slots: uint2, uint2, int
  // byte layout of one group's shared memory, computed once by the launcher (the kernel re-derives these addresses
  // wherever it runs out of registers, so they should cost one constant load, not a chain of multiplies)
  int group_bytes, off_r, off_fresh, off_vis;
  const uint8_t *dead;           // [rows] rows Delete removed from layer.nodes; null while nothing was deleted
  // NaN distances (cosine against a zero-norm row, non-finite components) break the argument that lets the kernel stop
  // maintaining the candidate queue once the result queue is full (Go's up()/down() never move an element past a NaN).
  // *row_flags bit 0 = the slab holds a row whose |x|^2 is 0, non-finite or outside [1e-18, 1e18]; such graphs, and
  // queries whose |q|^2 is outside [1e-20, 1e20] under cosine, keep the full Go-exact queue for the whole layer.
  const int *row_flags;
  int heap_c_full;               // the candidate queue has room for ef + 1 entries (else min(ef, k-1+M) + 1)
};

struct OutPtrs {
  int32_t *rows;
  float *dists;
  int32_t *count;
};
__device__ __forceinline__ OutPtrs search_out(const TraverseParams &p, long long q) {
  if (p.out_seg_len == 0) return {p.out_rows + q * p.k, p.out_dists + q * p.k, p.out_counts + q};
  const long long s = q / p.out_seg_len, ql = q - s * p.out_seg_len;
  int32_t *base = p.n_seg_ptr ? p.seg_ptr[s] : p.out_rows + s * p.out_seg_stride;
  return {base + ql * p.k, reinterpret_cast<float *>(base + p.out_seg_len * p.k) + ql * p.k,
          base + 2 * p.out_seg_len * p.k + ql};
}

__host__ __device__ inline size_t traverse_group_bytes(const TraverseParams &p) {
  return (size_t)(p.heap_c + p.heap_r) * 8 + (size_t)p.fresh_cap * 4 + (size_t)p.vis_cap * 4;
}

constexpr int TRAV_WARPS = 4;

#ifdef B200_K1_TIMING  // exploration build only (scripts/): cycles per phase of warp 0 of block 0, printed at the end
#define K1_T(i) { long long t__ = clock64(); tphase[i] += t__ - tlast; tlast = t__; }
#else
#define K1_T(i)
#endif

enum { ST_NEED = 0, ST_INIT = 1, ST_EXPAND = 2, ST_IDLE = 3 };

// NANG (NaN guard): the instantiation for graphs that hold rows able to produce NaN distances, and for the retry of a
// batch with such a query: the Go-exact candidate queue is kept for the whole layer (full ef + 1 capacity).  The
// NANG = false kernels -- the ones every ordinary search runs -- carry no per-step cost for it (measured: the guard
// inside the hot loop cost 7 %, profiles/r5g_k1_guard_ab.log): they refuse to run on a flagged graph and fail a risky
// query loudly (error 3), and the host runs the batch again through the NANG = true kernels.
template <int G, int CH, int U, int MINB, int TW = TRAV_WARPS, bool NANG = false>
__global__ void __launch_bounds__(TW * 32, MINB) traverse_kernel(GraphView g, TraverseParams p) {
  constexpr int NG = 32 / G;
  constexpr unsigned GBITS = G == 32 ? 0xffffffffu : ((1u << G) - 1u);
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int gl = lane & (G - 1);
  const int grp = lane / G;
  const int gshift = grp * G;
  const int leader_lane = gshift;
  const bool leader = gl == 0;
  unsigned char *base = smem + (unsigned)(p.group_bytes * (warp * NG + grp));
  GoHeap C{reinterpret_cast<uint2 *>(base), 0};
  GoHeap R{reinterpret_cast<uint2 *>(base + p.off_r), 0};
  int *fresh = reinterpret_cast<int *>(base + p.off_fresh);
  uint32_t *vtab = reinterpret_cast<uint32_t *>(base + p.off_vis);
  const uint32_t vmask = (uint32_t)p.vis_cap - 1;
  const int M = g.M;

  const long long n_groups = (long long)gridDim.x * TW * NG;
  long long q = ((long long)blockIdx.x * TW + warp) * NG + grp - n_groups;  // advanced before first use
  int st = ST_NEED, layer = 0, cur = -1, elevator = -1, kcur = 1, vcount = 0, lvl = -1;
  GQuery<CH> Q;
#pragma unroll
  for (int t = 0; t < CH; t++) Q.q[t] = make_float4(0.f, 0.f, 0.f, 0.f);
  Q.qq = 0.f;
  unsigned long long c_evals = 0, c_exp = 0, c_nbr = 0, c_over = 0;  // meaningful in leader lanes

  float best_d = 0.f;  // leader: the one-element result queue of a k = 1 layer lives in registers
  int best_id = -1, r1 = 0;
  int imp_id = -1;      // leader: last row that improved on result.Min() in this step
  bool cfree = false;   // leader: the candidate queue is no longer needed on this layer (see the decision below)
  constexpr bool keepc = NANG;  // never drop the candidate queue (see the decision below)
  if constexpr (!NANG) {
    if (p.row_flags != nullptr && (__ldg(p.row_flags) & 1)) {  // rows that can yield NaN entered the graph: not my job
      if (threadIdx.x == 0) *p.error = 3;
      return;
    }
  }

#ifdef B200_K1_TIMING
  long long tphase[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tlast = clock64();
  int tsteps = 0;
#endif
  for (;;) {
    K1_T(7)
    // ---- fetch the next traversal for groups that finished one --------------------------------------------------
    const bool need = st == ST_NEED;
    if (__any_sync(FULL, need)) {
      if (need) {
        q += n_groups;
        if (q >= p.B) st = ST_IDLE;
      }
      const bool on = need && st != ST_IDLE;
      GQuery<CH> NQ;
      group_load_query<G, CH>(NQ, p.queries + (on ? q : 0) * p.q_stride, on, g.dim, g.row_chunks, gl);
      if (on) {
        Q = NQ;
        layer = p.n_layers - 1;
        elevator = -1;
        lvl = p.levels ? p.levels[q] : -1;
        st = ST_INIT;
        if constexpr (!NANG) {
          // a query whose |q|^2 is 0, non-finite or out of range under cosine can meet NaN distances: fail it loudly
          if (p.metric != 0 && !(NQ.qq >= 1e-20f && NQ.qq <= 1e20f)) {
            if (leader) {
              *p.error = 3;
              c_over++;
              if (p.out_rows) *search_out(p, q).count = -1;
              else p.slot_cnt[p.slot_off[q]] = -1;
            }
            st = ST_NEED;
          }
        }
      }
    }
    if (__all_sync(FULL, st == ST_IDLE)) break;
    const bool init = st == ST_INIT, expand = st == ST_EXPAND;
    bool improved = false;  // leader
    int nfresh = 0;
    // ---- INIT (top layer, or a layer reached without an elevator): the entry point is the only row to evaluate ----
    if (init) {
      cur = elevator >= 0 ? elevator : p.entries[layer];  // graph.go:572-575 / 492-498 (constant bank, indexed)
      kcur = layer > 0 ? p.k_upper : p.k;
      C.len = 0;
      R.len = 0;
      r1 = 0;
      cfree = false;
      uint4 e4 = make_uint4(~0u, ~0u, ~0u, ~0u);
      uint4 *t4 = reinterpret_cast<uint4 *>(vtab);
      for (uint32_t i = gl; i < (vmask + 1) / 4; i += G) t4[i] = e4;
      vcount = 0;
      if (cur >= 0) {
        if (leader) fresh[0] = cur;
        nfresh = 1;
      }
    }
    __syncwarp();
    K1_T(0)
    // ---- EXPAND: neighbour list of `cur`, ascending key order, unvisited only (graph.go:137-144) ----------------
    if (__any_sync(FULL, expand)) {
      const int32_t *list = nullptr;
      if (expand) {
        if (layer == 0) list = g.adj0 + (size_t)cur * M;
        else list = g.adjU + ((size_t)g.upper_off[cur] + (layer - 1)) * M;
        if (leader) c_exp++;
      }
      bool more = expand;
      int nb_next = -1;
      if (more && gl < M) nb_next = list[gl];
      for (int b0 = 0; b0 < M; b0 += G) {
        const int nb = nb_next;
        nb_next = -1;  // the next G entries are requested before these are looked at
        if (more && b0 + G + gl < M) nb_next = list[b0 + G + gl];
        const bool valid = nb >= 0;
        const unsigned gv = (__ballot_sync(FULL, valid) >> gshift) & GBITS;
        bool fr = false;
        if (valid) {  // exact visited set: open addressing, empty = 0xffffffff
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
        const unsigned gf = (__ballot_sync(FULL, fr) >> gshift) & GBITS;
        if (fr) {
          fresh[nfresh + __popc(gf & ((1u << gl) - 1u))] = nb;
          // every fresh row is requested from HBM now (into L2), so the U-at-a-time loads below find it on chip
          // instead of paying one DRAM latency per batch
          const float4 *row = g.vec + (size_t)nb * g.row_chunks;
          for (int c = 0; c < g.row_chunks; c += 8) prefetch_l2(row + c);
        }
        nfresh += __popc(gf);
        if (leader) c_nbr += __popc(gv);
        if (gv != GBITS) more = false;  // lists are -1 padded at the end
        if (!__any_sync(FULL, more)) break;
      }
      vcount += nfresh;
    }
    __syncwarp();
    K1_T(1)
    const bool overflow = vcount * 2 > (int)vmask;  // load factor > 1/2: fail loudly, never approximate
    if (overflow) nfresh = 0;
    // ---- evaluate the fresh rows, U at a time; the leader replays graph.go:146-159 in key order -------------------
    // Software pipeline: the loads of batch b+1 are issued right after the multiply-adds of batch b, so they fly during
    // b's reduction and the leader's serial replay.
    int nid[U];
    float4 rr[U][CH];
#pragma unroll
    for (int u = 0; u < U; u++) nid[u] = u < nfresh ? fresh[u] : -1;
    group_load_rows<G, CH, U>(rr, g.vec, g.row_chunks, nid, gl);
#ifdef B200_K1_TIMING
    if (__any_sync(FULL, nfresh > 0)) {  // wait for the first batch explicitly so that it is timed on its own
      float tw = rr[0][0].x;
      asm volatile("" ::"f"(tw));
      tsteps++;
    }
    K1_T(2)
#endif
    for (int b0 = 0; __any_sync(FULL, b0 < nfresh); b0 += U) {
      int ids[U];
#pragma unroll
      for (int u = 0; u < U; u++) ids[u] = nid[u];
      GPartial<U> P;
      group_partials<G, CH, U>(P, rr, Q, p.metric);
#ifdef B200_K1_TIMING
      asm volatile("" ::"f"(P.a[0]));
      K1_T(3)
#endif
#pragma unroll
      for (int u = 0; u < U; u++) nid[u] = b0 + U + u < nfresh ? fresh[b0 + U + u] : -1;
      group_load_rows<G, CH, U>(rr, g.vec, g.row_chunks, nid, gl);
      const float dmine = group_finish<G, CH, U>(P, Q, p.metric, gl);
      float d[U];
#pragma unroll
      for (int u = 0; u < U; u++) d[u] = __shfl_sync(FULL, dmine, leader_lane + u * (G / U));
#ifdef B200_K1_TIMING
      asm volatile("" ::"f"(d[0]));
      K1_T(5)
#endif
      if (leader) {
#pragma unroll
        for (int u = 0; u < U; u++) {
          if (ids[u] < 0) continue;
          if (init) {
            // graph.go:109-127: candidates.Push(entry); result.Push(entry); visited[entry]; the first Pop takes the
            // entry back out, so the candidate queue stays empty and the entry is the first node expanded
            if (kcur == 1) {
              best_d = d[u];
              best_id = ids[u];
              r1 = 1;
            } else {
              R.push(d[u], ids[u]);
            }
            uint32_t h = ((uint32_t)ids[u] * 2654435761u) >> 7;
            vtab[h & vmask] = (uint32_t)ids[u];
            continue;
          }
          const float du = d[u];
          if (kcur == 1) {
            // a full one-element queue: `improved` (graph.go:147) and the replacement test (graph.go:150,
            // result.Max() is that element) are the same comparison
            if (du < best_d) {
              improved = true;
              imp_id = ids[u];
              best_d = du;
              best_id = ids[u];
            }
          } else {
            if (R.len > 0 && du < R.min_dist()) {  // graph.go:147
              improved = true;
              imp_id = ids[u];
            }
            if (R.len < kcur) {
              R.push(du, ids[u]);
            } else if (du < R.max_dist()) {  // result.Max() is the last array slot
              R.pop_last();
              R.push(du, ids[u]);
            }
          }
          if (!cfree) {
            C.push(du, ids[u]);
            if (C.len > p.ef) C.pop_last();  // graph.go:157-159
          }
        }
      }
    }
    K1_T(6)
    if (leader) c_evals += nfresh;  // E: one per evaluated row
    if (init && nfresh > 0) vcount = 1;
    // ---- leader decides: next node, end of layer ---------------------------------------------------------------------
    int next = -1;  // leader: row to expand next; -1 = the layer ends
    if (leader && (init || expand) && !overflow) {
      const int rl = kcur == 1 ? r1 : R.len;
      if (init) {
        if (nfresh > 0) next = cur;                                 // graph.go:125-127, first iteration
      } else if (cfree) {
        if (improved) next = imp_id;                                // what candidates.Pop() would return, see below
      } else if (!(!improved && rl >= kcur) && C.len > 0) {        // graph.go:164-166, 125
        next = (int)C.pop().y;
      }
      // Once the result queue is full it stays full (PopLast + Push), and from then on the candidate queue is never
      // looked at again: the loop goes on only when some row improved on result.Min() (graph.go:164-166), and that
      // row is strictly below everything evaluated on this layer, so candidates.Push sifted it to the root, where
      // neither a later Push of the same step (strict <) nor PopLast (ef >= 1 keeps slot 0) can move it -- the next
      // Pop (graph.go:126) returns exactly the last improving row.  So the queue is not maintained any more.
      // With NaN distances possible the argument fails (a NaN can sit at the root after a Pop): keep the queue.
      if (rl >= kcur && !keepc) cfree = true;
    }
    next = __shfl_sync(FULL, next, leader_lane);
    const int rlen = __shfl_sync(FULL, kcur == 1 ? r1 : R.len, leader_lane);
    const int r0_id = __shfl_sync(FULL, best_id, leader_lane);    // k = 1 layers: the result itself
    const float r0_d = __shfl_sync(FULL, best_d, leader_lane);
    __syncwarp();  // the leader's heap writes are visible to its group
    K1_T(4)
    bool chain = false;
    float elev_d = 0.f;
    if (init || expand) {
      if (overflow) {
        if (leader) {
          *p.error = 1;
          c_over++;
          if (p.out_rows) *search_out(p, q).count = -1;
          else p.slot_cnt[p.slot_off[q]] = -1;
        }
        st = ST_NEED;
      } else if (next >= 0) {
        cur = next;
        st = ST_EXPAND;
      } else {
        // ---- end of this layer: results are the raw heap array (graph.go:169) ----------------------------------
        const bool one = kcur == 1;
        if (p.out_rows) {
          if (layer == 0) {  // graph.go:588-593
            const OutPtrs o = search_out(p, q);
            for (int i = gl; i < p.k; i += G) {
              const bool ok = i < rlen;
              o.rows[i] = ok ? (one ? r0_id : R.id(i)) : -1;
              o.dists[i] = ok ? (one ? r0_d : R.dist(i)) : __int_as_float(0x7fc00000);
            }
            if (leader) *o.count = rlen;
          }
        } else if (layer <= lvl) {  // graph.go:510: this layer takes the new node
          const long long slot = p.slot_off[q] + layer;
          for (int i = gl; i < rlen; i += G) {
            p.slot_rows[slot * p.sink_stride + i] = one ? r0_id : R.id(i);
            p.slot_dists[slot * p.sink_stride + i] = one ? r0_d : R.dist(i);
          }
          if (leader) p.slot_cnt[slot] = rlen;
        }
        if (rlen > 0) {  // graph.go:584 / 508: the nearest result is the next layer's entry point
          elevator = one ? r0_id : R.id(0);
          elev_d = one ? r0_d : R.dist(0);
        }
        if (layer == 0) {
          st = ST_NEED;
        } else {
          layer--;
          chain = rlen > 0;
          st = chain ? ST_EXPAND : ST_INIT;
          if (chain && p.dead && p.dead[elevator]) {
            // layer.nodes[*elevator] is nil once the key was deleted (graph.go:573-575): the search from a nil node
            // returns nothing on this and, the elevator never changing again, on every layer below -- Search
            // answers with an empty list (graph.go:580-593).
            chain = false;
            st = ST_NEED;
            if (p.out_rows) {
              const OutPtrs o = search_out(p, q);
              for (int i = gl; i < p.k; i += G) {
                o.rows[i] = -1;
                o.dists[i] = __int_as_float(0x7fc00000);
              }
              if (leader) *o.count = 0;
            } else if (leader) {
              *p.error = 2;  // Add: "no nodes found in neighborhood search" (graph.go:501-505)
            }
          }
        }
      }
    }
    __syncwarp();  // every lane has read R before the leader reuses it
    // ---- the next layer starts at the elevator.  Its distance to the query was just computed in the layer above
    //      (same row, same query, same arithmetic), so graph.go:107-127 needs no new evaluation: the queues are
    //      {entry} and {} after the first Pop, and the entry is the node expanded next. --------------------------------
    if (chain) {
      cur = elevator;
      kcur = layer > 0 ? p.k_upper : p.k;
      uint4 e4 = make_uint4(~0u, ~0u, ~0u, ~0u);
      uint4 *t4 = reinterpret_cast<uint4 *>(vtab);
      for (uint32_t i = gl; i < (vmask + 1) / 4; i += G) t4[i] = e4;
      vcount = 1;
      if (leader) {
        c_evals++;  // the reference evaluates the entry again; counted so E equals the oracle's
        C.len = 0;
        cfree = kcur == 1 && !keepc;  // the result queue {entry} is already full
        if (kcur == 1) {
          best_d = elev_d;
          best_id = elevator;
          r1 = 1;
        } else {
          R.len = 0;
          R.push(elev_d, elevator);
        }
      }
    }
    __syncwarp();
    if (chain && leader) {
      uint32_t h = ((uint32_t)elevator * 2654435761u) >> 7;
      vtab[h & vmask] = (uint32_t)elevator;
    }
    __syncwarp();
  }
#ifdef B200_K1_TIMING
  if (blockIdx.x == 0 && threadIdx.x == 0)
    printf("K1 timing (cycles, warp 0): init %lld list+visited %lld first-rows %lld wait+math %lld reduce %lld leader %lld "
           "decide %lld tail %lld steps %d\n", tphase[0], tphase[1], tphase[2], tphase[3], tphase[5], tphase[6],
           tphase[4], tphase[7], tsteps);
#endif
  if (p.counters && leader) {
    if (c_evals) atomicAdd(&p.counters[0], c_evals);
    if (c_exp) atomicAdd(&p.counters[1], c_exp);
    if (c_nbr) atomicAdd(&p.counters[2], c_nbr);
    if (c_over) atomicAdd(&p.counters[3], c_over);
  }
}

}  // namespace b200
