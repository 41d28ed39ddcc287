// insert_kernel.cu -- K2: Graph.Add / Graph.BatchAdd (graph.go:437-531, 942-1042) with the reference's exact
// sequential semantics, one node at a time: nodes are inserted one after the other, each seeing every link of its
// predecessors; K2d: Graph.Delete / BatchDelete; and the host entry points of both (b200hnsw_add_batch also drives the
// speculative windowed schedule of spec_insert.cu, which builds the same graph).
//
// One persistent CTA walks the batch in input order.  Per node and per layer (top -> 0), all in inserter.cuh:
//   block_layer_search : layerNode.search(k=M, ef) from the elevator, the block evaluating an expansion's rows together
//   add_neighbor / add_own : for every result in heap-array order, addNeighbor both ways (graph.go:516-520):
//              sorted insert; on overflow the M+1 distances of the eviction scan (graph.go:60-71) across the warps,
//              the worst is dropped with its back-link, and replenish (graph.go:172-219) scores the
//              neighbours-of-neighbours with CosineDistance, all warps gathering rows in parallel.
// Order-dependent decisions (scan order, tie rule `d > worst || worst == nil`, heap pop order) are taken by one
// thread / one warp over the values the warps produced, so the result equals the oracle's bit for bit.
// Bound: latency of dependent gathers (~290 round trips per insertion), not bandwidth.
#include <algorithm>
#include <cstring>
#include <vector>

#include "inserter.cuh"

namespace b200 {

template <int CPL, int U>
__global__ void __launch_bounds__(INS_THREADS) insert_kernel(MutGraph g, InsertParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const InsLayout lay(p, g.M);
  Inserter<CPL, U> I;
  I.g = g;
  I.p = p;
  I.ctl = reinterpret_cast<InsCtl *>(smem + lay.ctl);
  I.nbh = reinterpret_cast<int *>(smem + lay.nbh);
  I.tmp = reinterpret_cast<int *>(smem + lay.tmp);
  I.tmpd = reinterpret_cast<float *>(smem + lay.tmpd);
  I.wl = reinterpret_cast<int *>(smem + lay.wl);
  I.cand = reinterpret_cast<int *>(smem + lay.cand);
  I.candd = reinterpret_cast<float *>(smem + lay.candd);
  I.rep = reinterpret_cast<uint32_t *>(smem + lay.rep);
  I.repi = reinterpret_cast<int *>(smem + lay.repi);
  I.own = reinterpret_cast<int *>(smem + lay.own);
  I.ownk = reinterpret_cast<long long *>(smem + lay.ownk);
  I.nbk = reinterpret_cast<long long *>(smem + lay.nbk);
  I.C = GoHeap{reinterpret_cast<uint2 *>(smem + lay.heapC), 0};
  I.R = GoHeap{reinterpret_cast<uint2 *>(smem + lay.heapR), 0};
  I.V = Visited{reinterpret_cast<uint32_t *>(smem + lay.vis), (uint32_t)p.vis_cap - 1, 0};
  I.tid = threadIdx.x;
  I.lane = threadIdx.x & 31;
  I.warp = threadIdx.x >> 5;
  I.c_evict = I.c_repl = I.c_evictions = 0;
  InsCtl *ctl = I.ctl;
  const int tid = I.tid, lane = I.lane, warp = I.warp;
  if (tid == 0) {
    ctl->n_layers = p.n_layers;
    ctl->abort_ = 0;
    for (int l = 0; l < B200HNSW_MAX_LAYERS; l++) ctl->entries[l] = p.entries[l];
  }
  __syncthreads();
  SearchCounters sc{0, 0, 0};
  bool aborted = false;
#ifdef B200_K2_TIMING
  for (int i = 0; i < 16; i++) I.tph[i] = 0;
  I.tlast = clock64();
#endif
  for (int j = 0; j < p.n && !aborted; j++) {
    const int row = (int)(p.first_row + j);
    const int L = p.levels[j];                          // randomLevel() drawn by the host (graph.go:457)
    const int old_layers = ctl->n_layers;
    const int top = (old_layers > L + 1 ? old_layers : L + 1) - 1;  // graph.go:462-464
    // Every layer search first, then the linking.  A layer's search reads that layer's lists only and linking at
    // layer l writes layer-l lists only, so the neighbourhoods equal those of the reference's interleaved loop
    // (graph.go:475-522); a failed Add (deleted elevator) then leaves the graph untouched.
    {
      QueryRegs<CPL> Q;  // every warp holds the new node's vector: the block evaluates an expansion's rows together
      load_query<CPL>(Q, reinterpret_cast<const float *>(g.vec + (size_t)row * g.row_chunks), g.dim, g.row_chunks,
                      lane);
      int elevator = -1, bad = 0;
      for (int l = top; l >= 0; l--) {                  // graph.go:475
        int cnt = -1;
        if (l < old_layers && ctl->entries[l] >= 0) {   // else graph.go:485-488: an empty layer takes the node
          int sp = elevator >= 0 ? elevator : ctl->entries[l];  // graph.go:492-498
          if (elevator >= 0 && p.dead && p.dead[elevator]) {
            cnt = 0;  // layer.nodes[*elevator] == nil: the search returns nothing (graph.go:101-103)
          } else {
            cnt = I.block_layer_search(l, sp, Q, sc);   // graph.go:500
            for (int i = tid; i < cnt; i += INS_THREADS) I.nbh[l * g.M + i] = I.R.id(i);
            if (cnt > 0) elevator = I.R.id(0);          // graph.go:508
          }
          if (cnt <= 0) {  // visited overflow (< 0) or "no nodes found in neighborhood search" (graph.go:501-505)
            bad = cnt < 0 ? 1 : 2;
            break;
          }
        }
        if (tid == 0) ctl->cnts[l] = cnt;
      }
      if (tid == 0) ctl->flag = bad;
    }
    __syncthreads();
#ifdef B200_K2_TIMING
    if (tid == 0) { const long long t__ = clock64(); I.tph[9] += t__ - I.tlast; I.tlast = t__; }
#endif
    if (ctl->flag) {
      aborted = true;
      if (tid == 0) {
        *p.error = ctl->flag;
        p.state_out[1 + B200HNSW_MAX_LAYERS] = j;  // nodes before j are in
      }
      break;
    }
    for (int l = top; l >= 0; l--) {
      const int cnt = ctl->cnts[l];
      if (cnt < 0) {
        if (tid == 0) ctl->entries[l] = row;
        continue;
      }
      if (L >= l) {  // graph.go:510-521
        I.own_begin(l, I.nbh + l * g.M, cnt);
        for (int t = 0; t < cnt; t++) {
          const int nb = I.nbh[l * g.M + t];
          I.add_neighbor(l, nb, row, p.metric, t + 1 < cnt ? I.nbh[l * g.M + t + 1] : -1);
          if (ctl->own_dirty) I.add_neighbor(l, row, nb, p.metric);  // the general path (inserter.cuh: own_dirty)
          else I.add_own(l, row, nb, I.nbk[t]);
        }
      }
    }
    __syncthreads();
    if (tid == 0 && L + 1 > old_layers) ctl->n_layers = L + 1;
    __syncthreads();
  }
#ifdef B200_K2_TIMING
  if (tid == 0)
    printf("K2 cycles per node: search %lld | addNeighbor: load+insert %lld scan-dist %lld decide %lld store+backlink %lld | "
           "replenish: load %lld nbr-lists %lld candidates %lld cand-dist %lld select+store %lld\n", I.tph[9] / p.n,
           I.tph[0] / p.n, I.tph[1] / p.n, I.tph[2] / p.n, I.tph[3] / p.n, I.tph[4] / p.n, I.tph[5] / p.n, I.tph[6] / p.n,
           I.tph[7] / p.n, I.tph[8] / p.n);
#endif
  if (tid == 0) {
    p.state_out[0] = ctl->n_layers;
    for (int l = 0; l < B200HNSW_MAX_LAYERS; l++) p.state_out[1 + l] = ctl->entries[l];
    atomicAdd(&p.counters[3], I.c_evict);
    atomicAdd(&p.counters[4], I.c_repl);
    atomicAdd(&p.counters[5], I.c_evictions);
  }
  if (tid == 0) {
    atomicAdd(&p.counters[0], sc.evals);
    atomicAdd(&p.counters[1], sc.expansions);
    atomicAdd(&p.counters[2], sc.nbr_reads);
  }
}

// K2d: Graph.Delete / Graph.BatchDelete (graph.go:843-895).  One CTA walks the rows in call order; per row, base layer
// first (the order of h.layers), isolate() on every layer the row is a member of.  The host has already taken the
// rows out of its layer bookkeeping and set their deleted flags.  Bound: latency of dependent gathers.
template <int CPL, int U>
__global__ void __launch_bounds__(INS_THREADS) delete_kernel(MutGraph g, InsertParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const InsLayout lay(p, g.M);
  Inserter<CPL, U> I;
  I.g = g;
  I.p = p;
  I.ctl = reinterpret_cast<InsCtl *>(smem + lay.ctl);
  I.nbh = reinterpret_cast<int *>(smem + lay.nbh);
  I.tmp = reinterpret_cast<int *>(smem + lay.tmp);
  I.tmpd = reinterpret_cast<float *>(smem + lay.tmpd);
  I.wl = reinterpret_cast<int *>(smem + lay.wl);
  I.cand = reinterpret_cast<int *>(smem + lay.cand);
  I.candd = reinterpret_cast<float *>(smem + lay.candd);
  I.rep = reinterpret_cast<uint32_t *>(smem + lay.rep);
  I.repi = reinterpret_cast<int *>(smem + lay.repi);
  I.own = reinterpret_cast<int *>(smem + lay.own);
  I.ownk = reinterpret_cast<long long *>(smem + lay.ownk);
  I.nbk = reinterpret_cast<long long *>(smem + lay.nbk);
  I.C = GoHeap{reinterpret_cast<uint2 *>(smem + lay.heapC), 0};
  I.R = GoHeap{reinterpret_cast<uint2 *>(smem + lay.heapR), 0};
  I.V = Visited{reinterpret_cast<uint32_t *>(smem + lay.vis), (uint32_t)p.vis_cap - 1, 0};
  I.tid = threadIdx.x;
  I.lane = threadIdx.x & 31;
  I.warp = threadIdx.x >> 5;
  I.c_evict = I.c_repl = I.c_evictions = 0;
  for (int j = 0; j < p.n; j++) {
    const int row = p.del_rows[j];
    const int L = p.del_levels[j];
    if (I.tid == 0) p.dead_set[row] = 1;  // read by Search / Add when they look an elevator up
    for (int l = 0; l <= L; l++) I.isolate(l, row);
  }
  if (I.tid == 0) atomicAdd(&p.counters[4], I.c_repl);
}

template <int CPL, int U>
static int launch_insert_t(b200hnsw *h, const MutGraph &g, InsertParams &p, cudaStream_t s) {
  auto kern = p.del_rows ? delete_kernel<CPL, U> : insert_kernel<CPL, U>;
  InsLayout lay(p, g.M);
  if (lay.total > h->smem_optin) {
    set_error("add: M=%d ef=%d need %zu B of shared memory (> %zu)", p.m, p.ef, lay.total, h->smem_optin);
    return B200HNSW_ERR_UNSUPPORTED;
  }
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lay.total));
  kern<<<1, INS_THREADS, lay.total, s>>>(g, p);
  B200_CUDA(cudaGetLastError());
  return 0;
}

static int launch_insert(b200hnsw *h, InsertParams &p, cudaStream_t s) {
  MutGraph g = mut_graph_of(h);
  size_insert_scratch(p, g);
  int cpl = (g.row_chunks + 31) / 32;
  switch (cpl) {
    case 1: return launch_insert_t<1, 4>(h, g, p, s);
    case 2: return launch_insert_t<2, 4>(h, g, p, s);
    case 3: return launch_insert_t<3, 2>(h, g, p, s);
    case 4: return launch_insert_t<4, 2>(h, g, p, s);
    case 5:
    case 6: return launch_insert_t<6, 2>(h, g, p, s);
    case 7:
    case 8: return launch_insert_t<8, 1>(h, g, p, s);
    case 9:
    case 10:
    case 11:
    case 12: return launch_insert_t<12, 1>(h, g, p, s);
    case 13:
    case 14:
    case 15:
    case 16: return launch_insert_t<16, 1>(h, g, p, s);
    default:
      set_error("add: dim %d > 2048 not supported", g.dim);
      return B200HNSW_ERR_UNSUPPORTED;
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200hnsw_add_batch(b200hnsw_t *h, const float *vectors, const int64_t *keys, const int32_t *levels,
                                  int64_t n, int32_t m, int32_t ef, int32_t metric, b200hnsw_stats *stats) {
  if (stats) memset(stats, 0, sizeof *stats);
  if (!h || n < 0 || (n > 0 && (!vectors || !keys || !levels))) {
    set_error("add: bad argument");
    return B200HNSW_ERR_BAD_ARG;
  }
  if (m <= 0 || m > h->M) {  // graph.go:918-920
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
  if (n == 0) return 0;
  if (h->build_mode != B200HNSW_MODE_REFERENCE) {
    set_error("add: the extended build mode exists for the wave-batched Add only (b200hnsw_add_batch_waves); the "
              "sequential Add is the reference's and uses the reference's layer search");
    return B200HNSW_ERR_UNSUPPORTED;
  }
  if (h->n + n > 0x7fffffffLL) {
    set_error("add: more than 2^31-1 rows");
    return B200HNSW_ERR_UNSUPPORTED;
  }
  int64_t add_slots = 0;
  for (int64_t i = 0; i < n; i++) {
    if (levels[i] < 0 || levels[i] >= B200HNSW_MAX_LAYERS) {  // graph.go:466-468 "invalid level"
      set_error("invalid level: %d", levels[i]);
      return B200HNSW_ERR_BAD_ARG;
    }
    add_slots += levels[i];
  }
  if (h->n_dead) {
    // A layer emptied by Delete above a node's drawn level: the reference parks the node there without the layers
    // in between (graph.go:485-488), after which every Search loses its elevator.  Not reproduced: refused.
    int top_empty = -1;  // highest empty layer below n_layers
    for (int l = h->n_layers - 1; l >= 0; l--)
      if (h->layer_size[l] == 0) {
        top_empty = l;
        break;
      }
    for (int64_t i = 0; i < n && top_empty >= 0; i++) {
      if (levels[i] < top_empty) {
        set_error("add: layer %d was emptied by Delete and node %lld of the batch draws level %d below it; the "
                  "reference's behaviour here (graph.go:485-488) is not reproduced", top_empty, (long long)i,
                  levels[i]);
        return B200HNSW_ERR_UNSUPPORTED;
      }
      top_empty = -1;  // this node fills every empty layer up to its level, which is >= top_empty
    }
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
  const int dim = h->cfg.dim, rf = h->row_floats;
  std::vector<int32_t> upper_off((size_t)n, -1);
  int64_t nU = h->nU;
  for (int64_t i = 0; i < n; i++)
    if (levels[i] > 0) {
      upper_off[i] = (int32_t)nU;
      nU += levels[i];
    }
  const int64_t CH = 1 << 16;  // insertions per launch
  size_t scratch = sizeof(int32_t) * (size_t)std::min(n, CH) + 256 + 64 * 8 + sizeof(int) * (2 + B200HNSW_MAX_LAYERS);
  if ((rc = c->reserve_device(scratch + 1024))) return fail(rc);
  char *sp = (char *)c->d_scratch;
  int32_t *d_levels = (int32_t *)sp;
  sp += (sizeof(int32_t) * (size_t)std::min(n, CH) + 255) & ~(size_t)255;
  unsigned long long *d_ctr = (unsigned long long *)sp;
  sp += 64;
  int *d_err = (int *)sp;
  sp += 64;
  int *d_state = (int *)sp;
  cudaError_t e;
  if ((e = cudaMemcpy2DAsync(h->d.vec + (size_t)h->n * rf, sizeof(float) * rf, vectors, sizeof(float) * dim,
                             sizeof(float) * dim, (size_t)n, cudaMemcpyHostToDevice, s)) != cudaSuccess ||
      (e = cudaMemcpyAsync(h->d.keys + h->n, keys, sizeof(int64_t) * (size_t)n, cudaMemcpyHostToDevice, s)) !=
          cudaSuccess ||
      (e = cudaMemcpyAsync(h->d.upper_off + h->n, upper_off.data(), sizeof(int32_t) * (size_t)n,
                           cudaMemcpyHostToDevice, s)) != cudaSuccess ||
      (e = cudaMemsetAsync(d_ctr, 0, 128, s)) != cudaSuccess)
    return fail(cuda_fail(e, "add staging"));
  if ((rc = h->flag_rows(h->n, n, s))) return fail(rc);
  if ((e = cudaEventRecord(c->ev0, s)) != cudaSuccess) return fail(cuda_fail(e, "event"));
  int launches = 0;
  int n_layers = h->n_layers;
  int32_t entries[B200HNSW_MAX_LAYERS];
  memcpy(entries, h->entry, sizeof entries);
  auto commit = [&](int64_t count) {  // host mirrors of the first `count` nodes of the batch
    int64_t used = h->nU;
    for (int64_t i = 0; i < count; i++) {
      h->h_keys.push_back(keys[i]);
      h->h_level.push_back(levels[i]);
      h->h_dead.push_back(0);
      h->h_upper_off.push_back(upper_off[i]);
      for (int l = 0; l <= levels[i]; l++) h->layer_size[l]++;
      used += levels[i];
    }
    h->n += count;
    h->nU = used;
    h->n_layers = n_layers;
    memcpy(h->entry, entries, sizeof entries);
  };
  // `count` nodes starting at batch index `first` through the sequential kernel; *bad = nodes of them that went in
  auto run_sequential = [&](int64_t first, int64_t count, const int32_t *d_lev, int *err, int64_t *bad) -> int {
    InsertParams p{};
    p.first_row = h->n + first;
    p.n = (int)count;
    p.levels = d_lev;
    p.m = m;
    p.ef = ef;
    p.metric = metric;
    p.n_layers = n_layers;
    memcpy(p.entries, entries, sizeof entries);
    p.state_out = d_state;
    p.counters = d_ctr;
    p.error = d_err;
    p.dead = h->n_dead ? h->d.dead : nullptr;
    p.row_flags = h->d_row_flags;
    int rc2 = launch_insert(h, p, s);
    if (rc2) return rc2;
    launches++;
    int state[2 + B200HNSW_MAX_LAYERS];
    *err = 0;
    cudaError_t e2;
    if ((e2 = cudaMemcpyAsync(state, d_state, sizeof state, cudaMemcpyDeviceToHost, s)) != cudaSuccess ||
        (e2 = cudaMemcpyAsync(err, d_err, sizeof(int), cudaMemcpyDeviceToHost, s)) != cudaSuccess ||
        (e2 = cudaStreamSynchronize(s)) != cudaSuccess)
      return cuda_fail(e2, "add kernel");
    n_layers = state[0];
    memcpy(entries, state + 1, sizeof entries);
    *bad = *err ? state[1 + B200HNSW_MAX_LAYERS] : count;
    return 0;
  };
  auto report = [&](int err, int64_t bad) {  // the kernel stopped in front of node `bad`: its predecessors stay
    commit(bad);
    if (err == 9) {
      set_error("add: internal wait timed out in the windowed Add (node %lld of the batch)", (long long)bad);
      return fail(B200HNSW_ERR_CUDA);
    }
    set_error(err == 1 ? "add: a layer search overflowed the visited set (node %lld of the batch)"
                       : "no nodes found in neighborhood search (node %lld of the batch)", (long long)bad);
    return fail(err == 1 ? B200HNSW_ERR_VISITED_OVERFLOW : B200HNSW_ERR_STATE);
  };
  for (int64_t off = 0, cnt = 0; off < n; off += cnt) {
    // launches grow with the graph while it is small, so that the window is chosen anew as the rows double
    cnt = std::min(std::min(CH, n - off), std::max<int64_t>(2048, h->n + off));
    if ((e = cudaMemcpyAsync(d_levels, levels + off, sizeof(int32_t) * (size_t)cnt, cudaMemcpyHostToDevice, s)) !=
            cudaSuccess ||
        (e = cudaMemsetAsync(d_err, 0, sizeof(int), s)) != cudaSuccess)
      return fail(cuda_fail(e, "levels copy"));
    // The same graph either way (spec_insert.cu): a window of insertions speculated in parallel and committed in input
    // order where that pays (enough rows for two neighbourhoods to be apart), else one CTA one node at a time.
    int window = h->add_window ? h->add_window : spec_window_auto(h->n + off);
    int64_t pos = 0;
    while (pos < cnt) {
      int err = 0;
      if (window >= 2 && cnt - pos >= 2) {
        InsertParams p{};
        p.first_row = h->n + off + pos;
        p.n = (int)(cnt - pos);
        p.levels = d_levels + pos;
        p.m = m;
        p.ef = ef;
        p.metric = metric;
        p.n_layers = n_layers;
        memcpy(p.entries, entries, sizeof entries);
        p.counters = d_ctr;
        p.dead = h->n_dead ? h->d.dead : nullptr;
        p.row_flags = h->d_row_flags;
        int done = 0;
        rc = launch_spec_insert(h, p, window, &done, &err, &n_layers, entries, h->spec_accepted_redone, s);
        if (rc < 0) return fail(rc);
        if (rc > 0) {
          window = 1;  // does not fit this device / this dim: sequential
          continue;
        }
        launches++;
        pos += done;
        if (err == 8) {  // this node writes more lists than the overlay holds: the sequential kernel takes it
          int64_t bad = 0;
          if ((rc = run_sequential(off + pos, 1, d_levels + pos, &err, &bad))) return fail(rc);
          if (err) return report(err, off + pos);
          pos += 1;
        } else if (err) {
          return report(err, off + pos);
        }
      } else {
        int64_t bad = 0;
        if ((rc = run_sequential(off + pos, cnt - pos, d_levels + pos, &err, &bad))) return fail(rc);
        if (err) return report(err, off + pos + bad);
        pos = cnt;
      }
    }
  }
  unsigned long long ctr[8];
  if ((e = cudaEventRecord(c->ev1, s)) != cudaSuccess ||
      (e = cudaMemcpyAsync(ctr, d_ctr, sizeof ctr, cudaMemcpyDeviceToHost, s)) != cudaSuccess ||
      (e = cudaStreamSynchronize(s)) != cudaSuccess)
    return fail(cuda_fail(e, "add counters"));
  if ((rc = h->refresh_row_flags(s))) return fail(rc);
  commit(n);
  if (stats) {
    stats->dist_evals = (int64_t)ctr[0];
    stats->evict_evals = (int64_t)ctr[3];
    stats->replenish_evals = (int64_t)ctr[4];
    stats->evictions = (int64_t)ctr[5];
    stats->expansions = (int64_t)ctr[1];
    stats->nbr_reads = (int64_t)ctr[2];
    stats->launches = launches;
    cudaEventElapsedTime(&stats->kernel_ms, c->ev0, c->ev1);
  }
  h->release(c);
  return 0;
}

// Graph.Delete / Graph.BatchDelete -- graph.go:843-895.  rows[i] is the row the caller's key map holds for the i-th
// key (-1 for a key it does not know); out_deleted[i] mirrors the reference's []bool.
extern "C" int b200hnsw_delete_batch(b200hnsw_t *h, const int32_t *rows, int64_t n, int32_t m, uint8_t *out_deleted,
                                     b200hnsw_stats *stats) {
  if (stats) memset(stats, 0, sizeof *stats);
  if (!h || n < 0 || (n > 0 && !rows)) {
    set_error("delete: bad argument");
    return B200HNSW_ERR_BAD_ARG;
  }
  if (m <= 0 || m > h->M) {  // graph.go:918-920
    set_error("M must be greater than 0 and <= max_m=%d, got %d", h->M, m);
    return B200HNSW_ERR_BAD_ARG;
  }
  if (out_deleted) memset(out_deleted, 0, (size_t)n);
  if (n == 0 || h->n_layers == 0) return 0;  // graph.go:847-849
  // The new host state (deleted flags, layer sizes, entry points, the caller's []bool) is built in temporaries and
  // committed only after the delete kernel has finished: a failure on the way leaves host and device agreeing.
  std::vector<int32_t> del, lev;
  std::vector<uint8_t> t_dead(h->h_dead);
  std::vector<uint8_t> t_out((size_t)n, 0);
  int64_t t_size[B200HNSW_MAX_LAYERS];
  int32_t t_entry[B200HNSW_MAX_LAYERS];
  memcpy(t_size, h->layer_size, sizeof t_size);
  memcpy(t_entry, h->entry, sizeof t_entry);
  for (int64_t i = 0; i < n; i++) {
    const int32_t r = rows[i];
    if (r < 0 || r >= h->n || t_dead[(size_t)r]) continue;  // unknown key, or deleted earlier (in this batch too)
    t_dead[(size_t)r] = 1;
    t_out[(size_t)i] = 1;
    del.push_back(r);
    lev.push_back(h->h_level[(size_t)r]);
    for (int l = 0; l <= h->h_level[(size_t)r]; l++) {  // delete(layer.nodes, key)
      t_size[l]--;
      if (t_entry[l] == r) {  // layer.entry(): the lowest live member row takes over (documented stand-in)
        int32_t en = -1;
        if (t_size[l] > 0)
          for (int64_t x = 0; x < h->n; x++)
            if (!t_dead[(size_t)x] && h->h_level[(size_t)x] >= l) {
              en = (int32_t)x;
              break;
            }
        t_entry[l] = en;
      }
    }
  }
  if (del.empty()) return 0;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  CallCtx *c = h->acquire();
  if (!c) return B200HNSW_ERR_CUDA;
  cudaStream_t s = c->stream;
  auto fail = [&](int code) {
    h->release(c);
    return code;
  };
  const size_t nd = del.size();
  const size_t lev_off = (sizeof(int32_t) * nd + 255) & ~(size_t)255;
  int rc = c->reserve_device(2 * lev_off + nd + 512);
  if (rc) return fail(rc);
  char *sp = (char *)c->d_scratch;
  int32_t *d_rows = (int32_t *)sp;
  int32_t *d_lev = (int32_t *)(sp + lev_off);
  unsigned long long *d_ctr = (unsigned long long *)(sp + 2 * lev_off);
  cudaError_t e;
  if ((e = cudaMemcpyAsync(d_rows, del.data(), sizeof(int32_t) * nd, cudaMemcpyHostToDevice, s)) != cudaSuccess ||
      (e = cudaMemcpyAsync(d_lev, lev.data(), sizeof(int32_t) * nd, cudaMemcpyHostToDevice, s)) != cudaSuccess ||
      (e = cudaMemsetAsync(d_ctr, 0, 128, s)) != cudaSuccess)
    return fail(cuda_fail(e, "delete staging"));
  if ((e = cudaEventRecord(c->ev0, s)) != cudaSuccess) return fail(cuda_fail(e, "event"));
  InsertParams p{};
  p.n = (int)nd;
  p.m = m;
  p.ef = 1;
  p.del_rows = d_rows;
  p.del_levels = d_lev;
  p.dead_set = h->d.dead;
  p.counters = d_ctr;
  if ((rc = launch_insert(h, p, s))) return fail(rc);
  unsigned long long ctr[8];
  if ((e = cudaEventRecord(c->ev1, s)) != cudaSuccess ||
      (e = cudaMemcpyAsync(ctr, d_ctr, sizeof ctr, cudaMemcpyDeviceToHost, s)) != cudaSuccess ||
      (e = cudaStreamSynchronize(s)) != cudaSuccess)
    return fail(cuda_fail(e, "delete kernel"));
  if (stats) {
    stats->replenish_evals = (int64_t)ctr[4];
    stats->launches = 1;
    cudaEventElapsedTime(&stats->kernel_ms, c->ev0, c->ev1);
  }
  h->h_dead.swap(t_dead);
  h->n_dead += (int64_t)nd;
  memcpy(h->layer_size, t_size, sizeof t_size);
  memcpy(h->entry, t_entry, sizeof t_entry);
  if (out_deleted) memcpy(out_deleted, t_out.data(), (size_t)n);
  h->release(c);
  return 0;
}

extern "C" int32_t b200hnsw_is_deleted(const b200hnsw_t *h, int32_t row) {
  return (h && row >= 0 && row < h->n) ? (int32_t)h->h_dead[(size_t)row] : -1;
}
