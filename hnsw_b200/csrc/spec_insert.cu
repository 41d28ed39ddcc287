// spec_insert.cu -- K2s: Graph.Add / Graph.BatchAdd (graph.go:437-531, 942-1042) in the REFERENCE'S ORDER, with the work of
// a window of pending nodes speculated in parallel and committed one after the other.  The graph it leaves is the one the
// sequential kernel (insert_kernel.cu) leaves, bit for bit; only the schedule differs.
//
// Why it is exact.  An insertion is a function of (a) the layer state (number of layers, entry points) and (b) the
// neighbour lists it READS (layer searches, addNeighbor's load, the back-link, replenish's neighbours-of-neighbours); rows
// and keys are immutable during the call.  A worker CTA runs the whole insertion of node j -- the same device code as the
// sequential kernel (inserter.cuh, SPEC = true) -- against the graph as it is while its predecessors are still being
// committed, but writes nothing into it: every list it would write goes to a private overlay, every list it reads is
// logged.  The committer CTA takes the nodes in input order: the speculation of node j, started when e0 nodes were
// committed, is ACCEPTED iff no list it read was written by a commit in [e0, j) and the layer state has not changed since;
// then its overlay is exactly what the sequential kernel would have stored, and it is copied into the graph.  Otherwise
// the committer runs node j itself, now as the head of the order, where nothing can be stale.  `last_writer[list]`
// (commit index + 1, zeroed per launch) makes the test one load per logged read.
//
// Memory ordering.  The committer stores the lists, fences, then publishes `done`; a worker reads `done` (-> e0), fences,
// and reads lists past L1 (ld.global.cg, the overlay and log likewise on the committer's side).  A list the committer is
// rewriting while a worker reads it may be seen torn -- every entry is still a valid row, and that read is rejected by
// the test above because its writer's index is >= e0.
//
// Repair.  A worker's next ticket is a window away, so while its node waits for its turn it keeps the speculation
// current: the same test against `last_writer`, and when a read has gone stale it takes the slot back (atomicCAS on the
// slot's ready word: j + 1 = ready, 0 = being run again, -(j + 1) = taken by the committer) and runs the node again with a
// fresh e0 -- off the committer's path.  The committer's own test stays the one that counts.
//
// Progress.  Tickets are handed out in input order; a worker holding node j waits until j < done + depth, the committer
// waits for the slot of node `done`, whose worker needs nothing from anyone.  Every wait polls an abort word and gives
// up after a few seconds (error 9) instead of hanging the device; no wait crosses a process or a GPU.
// Bound: the dependency between insertions -- a node that reads what one of its R*T nearest predecessors writes (R =
// commit rate, T = one insertion's latency) cannot be right before that one is in (~29 k/s at 1 M rows
// for today's T, 16-17 k measured).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "inserter.cuh"

namespace b200 {

struct SpecShared {
  int ticket;    // next node (index in this launch) to hand to a worker
  int done;      // nodes committed
  int abort_;    // everybody leaves
  int err;       // 1 visited overflow, 2 "no nodes found", 8 overlay / read-log overflow at the head, 9 wait timed out
  int err_node;
  int sver;      // bumped after every change of the layer state below
  int n_layers;
  int entries[B200HNSW_MAX_LAYERS];
  unsigned long long n_accepted, n_redone;
  unsigned long long n_repaired;  // speculations a worker ran again itself while its node waited for its turn
  unsigned long long cyc[4];      // committer cycles: waiting for the next node, checking it, running it again, committing
};

static_assert(B200HNSW_MAX_LAYERS <= 64, "ent_mask holds one bit per layer");

struct SpecSlot {
  int ready;     // node index + 1: speculation complete; -(index + 1): the committer has taken it; 0: its worker is
                 // running it again
  int status;    // 0, or the error the insertion would raise (1, 2), or 8
  int e0;        // nodes committed when the speculation started
  int sver0;     // layer-state version it used
  int n_rd, n_ov;
  int new_layers;               // L + 1 if the node opens new layers, else 0
  int pad_;
  unsigned long long ent_mask;  // layers (< B200HNSW_MAX_LAYERS = 64) that were empty and take the node as entry
  unsigned long long ctr[6];
};

struct SpecParams {
  SpecShared *S;
  SpecSlot *slots;   // [depth]
  uint32_t *rlogs;   // [depth][rl_cap]
  int32_t *ovls;     // [depth][(OV_MAX + 1) * M]
  uint32_t *ovids;   // [depth][OV_MAX]
  int *lastw0;       // [rows]   commit index + 1 of the last write to a layer-0 list in this launch
  int *lastwU;       // [upper slots]
  int depth, rl_cap;
};

__device__ __forceinline__ int ldv(const int *p) { return *reinterpret_cast<const volatile int *>(p); }
__device__ __forceinline__ void stv(int *p, int v) { *reinterpret_cast<volatile int *>(p) = v; }
constexpr long long SPEC_TIMEOUT = 12000000000LL;  // cycles (~6 s): a wait that long means a bug, not load

__device__ __forceinline__ void spec_give_up(SpecShared *S) {
  atomicCAS(&S->err, 0, 9);
  stv(&S->abort_, 1);
}

// The whole insertion of node j (graph.go:445-528) against the graph as it is now, into slot `si`: overlay + read log +
// header.  Block-wide; ends with a barrier.  Same steps as insert_kernel's loop body.
template <int CPL, int U>
__device__ void spec_node(Inserter<CPL, U, true> &I, const InsertParams &p, const SpecParams &sp, int j, int si,
                          int e0, int ready_val) {
  SpecShared *S = sp.S;
  InsCtl *ctl = I.ctl;
  const int tid = I.tid, lane = I.lane, warp = I.warp;
  const int M = I.g.M;
  __shared__ int sh_sver0;
  if (tid == 0) {
    sh_sver0 = ldv(&S->sver);  // the version first: a state newer than its version only makes the check fail
    __threadfence();
    ctl->n_layers = ldv(&S->n_layers);
    for (int l = 0; l < B200HNSW_MAX_LAYERS; l++) ctl->entries[l] = ldv(&S->entries[l]);
    ctl->n_rd = 0;
    ctl->n_ov = 0;
    ctl->ov_fail = 0;
    ctl->ent_mask = 0;
  }
  for (int i = tid; i < OV_HASH; i += INS_THREADS) I.ovk[i] = ~0u;
  I.ovl = sp.ovls + (size_t)si * (OV_MAX + 1) * M;
  I.ovid = sp.ovids + (size_t)si * OV_MAX;
  I.rlog = sp.rlogs + (size_t)si * sp.rl_cap;
  I.rl_cap = sp.rl_cap;
  I.c_evict = I.c_repl = I.c_evictions = 0;
  SearchCounters sc{0, 0, 0};
  __syncthreads();
  const int row = (int)(p.first_row + j);
  const int L = p.levels[j];
  const int old_layers = ctl->n_layers;
  const int top = (old_layers > L + 1 ? old_layers : L + 1) - 1;  // graph.go:462-464
  {
    QueryRegs<CPL> Q;
    load_query<CPL>(Q, reinterpret_cast<const float *>(I.g.vec + (size_t)row * I.g.row_chunks), I.g.dim,
                    I.g.row_chunks, lane);
    int elevator = -1, bad = 0;
    for (int l = top; l >= 0; l--) {                  // graph.go:475
      int cnt = -1;
      if (l < old_layers && ctl->entries[l] >= 0) {   // else graph.go:485-488: an empty layer takes the node
        int spn = elevator >= 0 ? elevator : ctl->entries[l];  // graph.go:492-498
        if (elevator >= 0 && p.dead && p.dead[elevator]) {
          cnt = 0;  // layer.nodes[*elevator] == nil (graph.go:101-103)
        } else {
          cnt = I.block_layer_search(l, spn, Q, sc);  // graph.go:500; the lists it reads are logged
          for (int i = tid; i < cnt; i += INS_THREADS) I.nbh[l * M + i] = I.R.id(i);
          if (cnt > 0) elevator = I.R.id(0);          // graph.go:508
        }
        if (cnt <= 0) {
          bad = cnt < 0 ? 1 : 2;
          break;
        }
      }
      if (tid == 0) ctl->cnts[l] = cnt;
    }
    if (tid == 0) ctl->flag = bad;
  }
  __syncthreads();
  int status = ctl->flag;
  if (!status) {
    for (int l = top; l >= 0; l--) {
      const int cnt = ctl->cnts[l];
      if (cnt < 0) {
        if (tid == 0) ctl->ent_mask |= 1ull << l;
        continue;
      }
      if (L >= l) {  // graph.go:510-521
        I.own_begin(l, I.nbh + l * M, cnt);
        for (int t = 0; t < cnt; t++) {
          const int nb = I.nbh[l * M + t];
          I.add_neighbor(l, nb, row, p.metric, t + 1 < cnt ? I.nbh[l * M + t + 1] : -1);
          if (ctl->own_dirty) I.add_neighbor(l, row, nb, p.metric);  // the general path (inserter.cuh: own_dirty)
          else I.add_own(l, row, nb, I.nbk[t]);
        }
      }
    }
    __syncthreads();
    if (ctl->ov_fail || ctl->n_rd > I.rl_cap) status = 8;
  }
  __threadfence();  // every thread's overlay stores and log entries are out before the header says so
  __syncthreads();
  if (tid == 0) {
    SpecSlot *sl = sp.slots + si;
    sl->status = status;
    sl->e0 = e0;
    sl->sver0 = sh_sver0;
    sl->n_rd = ctl->n_rd < I.rl_cap ? ctl->n_rd : I.rl_cap;
    sl->n_ov = ctl->n_ov < OV_MAX ? ctl->n_ov : OV_MAX;
    sl->new_layers = L + 1 > old_layers ? L + 1 : 0;
    sl->ent_mask = ctl->ent_mask;
    sl->ctr[0] = sc.evals;
    sl->ctr[1] = sc.expansions;
    sl->ctr[2] = sc.nbr_reads;
    sl->ctr[3] = I.c_evict;
    sl->ctr[4] = I.c_repl;
    sl->ctr[5] = I.c_evictions;
    __threadfence();
    stv(&sl->ready, ready_val);
  }
  __syncthreads();
}

// the committer's copy, out of line: it runs one node in twenty or so, and a second inlined body of the whole insertion
// doubles the compile time for nothing
template <int CPL, int U>
__device__ __noinline__ void spec_node_again(Inserter<CPL, U, true> &I, const InsertParams &p, const SpecParams &sp, int j,
                                             int si, int e0, int ready_val) {
  spec_node<CPL, U>(I, p, sp, j, si, e0, ready_val);
}

template <int CPL, int U>
__global__ void __launch_bounds__(INS_THREADS) spec_insert_kernel(MutGraph g, InsertParams p, SpecParams sp) {
  extern __shared__ __align__(16) unsigned char smem[];
  const InsLayout lay(p, g.M);
  Inserter<CPL, U, true> I;
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
  I.ovk = reinterpret_cast<uint32_t *>(smem + lay.total);
  I.ovx = reinterpret_cast<int *>(smem + lay.total + sizeof(uint32_t) * OV_HASH);
  I.tid = threadIdx.x;
  I.lane = threadIdx.x & 31;
  I.warp = threadIdx.x >> 5;
  const int tid = I.tid, lane = I.lane, warp = I.warp;
  SpecShared *S = sp.S;
  __shared__ int sh_j, sh_e0, sh_ok, sh_nrd;

  if (blockIdx.x != 0) {
    // ---- worker: speculate the nodes the ticket counter hands out --------------------------------------------------
    int j = 0, my_e0 = 0;
    bool repair = false;  // block-uniform: run node j again (its speculation went stale while it waited)
    for (;;) {
      if (!repair) {
        if (tid == 0) {
          int jt = atomicAdd(&S->ticket, 1);
          int e0 = 0;
          if (jt < p.n) {
            const long long t0 = clock64();
            for (;;) {
              if (ldv(&S->abort_)) {
                jt = p.n;
                break;
              }
              const int d = ldv(&S->done);
              if (jt < d + sp.depth) {  // slot jt % depth is free again and the window reaches jt
                e0 = d;
                break;
              }
              __nanosleep(256);
              if (clock64() - t0 > SPEC_TIMEOUT) {
                spec_give_up(S);
                jt = p.n;
                break;
              }
            }
          }
          __threadfence();
          sh_j = jt;
          sh_e0 = e0;
        }
        __syncthreads();
        j = sh_j;
        my_e0 = sh_e0;
        if (j >= p.n) break;
      }
      const int si = j % sp.depth;
      SpecSlot *sl = sp.slots + si;
      spec_node<CPL, U>(I, p, sp, j, si, my_e0, j + 1);  // ends with a barrier: the sh_ words may be rewritten
      if (repair && tid == 0) atomicAdd(&S->n_repaired, 1ull);
      repair = false;
      // This worker's next ticket is a window away, so until node j is taken it has nothing better to do than to keep the
      // speculation current: if a commit has meanwhile written a list it read, run it again now, off the committer's
      // path.  (The committer's own test stays the one that counts.)
      for (;;) {
        if (tid == 0) sh_ok = !ldv(&S->abort_) && ldv(&S->done) <= j && ldv(&sl->ready) == j + 1;
        __syncthreads();
        const int waiting = sh_ok;
        __syncthreads();
        if (!waiting) break;
        const int nrd = I.ctl->n_rd < I.rl_cap ? I.ctl->n_rd : I.rl_cap;
        int stale = 0;
        for (int r = tid; r < nrd; r += INS_THREADS) {
          const uint32_t id = __ldcg(I.rlog + r);
          const int lw = __ldcg((id & 0x80000000u) ? sp.lastwU + (id & 0x7fffffffu) : sp.lastw0 + id);
          stale |= lw > my_e0;
        }
        stale = __syncthreads_or(stale);
        if (!stale) {
          __nanosleep(2000);
          continue;
        }
        if (tid == 0) {
          const int mine = atomicCAS(&sl->ready, j + 1, 0) == j + 1;  // the committer may have taken it just now
          sh_ok = mine;
          if (mine) {
            sh_e0 = ldv(&S->done);
            __threadfence();
          }
        }
        __syncthreads();
        const int mine = sh_ok;
        const int e0n = sh_e0;
        __syncthreads();
        if (mine) {
          my_e0 = e0n;
          repair = true;
        }
        break;
      }
    }
    return;
  }

  // ---- committer: input order -------------------------------------------------------------------------------------
  unsigned long long acc[6] = {0, 0, 0, 0, 0, 0}, n_acc = 0, n_redo = 0;
  long long cyc[4] = {0, 0, 0, 0}, tl = clock64();
#define SPEC_T(i) { const long long t__ = clock64(); cyc[i] += t__ - tl; tl = t__; }
  for (int j = 0; j < p.n; j++) {
    const int si = j % sp.depth;
    SpecSlot *sl = sp.slots + si;
    if (tid == 0) {
      int ok = -1;
      const long long t0 = clock64();
      for (;;) {
        if (ldv(&sl->ready) == j + 1 && atomicCAS(&sl->ready, j + 1, -(j + 1)) == j + 1) {  // taken: its worker lets go
          __threadfence();
          ok = ldv(&sl->status) == 0 && ldv(&sl->sver0) == ldv(&S->sver);
          break;
        }
        if (ldv(&S->abort_)) break;
        __nanosleep(64);
        if (clock64() - t0 > SPEC_TIMEOUT) {
          spec_give_up(S);
          break;
        }
      }
      sh_ok = ok;
      sh_e0 = ok > 0 ? ldv(&sl->e0) : 0;
      sh_nrd = ok > 0 ? ldv(&sl->n_rd) : 0;
    }
    __syncthreads();
    SPEC_T(0)
    if (sh_ok < 0) break;
    int redo = !sh_ok;
    if (!redo) {
      // was any list this speculation read written by a commit it could not have seen?
      const uint32_t *rl = sp.rlogs + (size_t)si * sp.rl_cap;
      const int e0 = sh_e0, nrd = sh_nrd;
      int stale = 0;
      for (int r = tid; r < nrd; r += INS_THREADS) {
        const uint32_t id = __ldcg(rl + r);
        const int lw = (id & 0x80000000u) ? sp.lastwU[id & 0x7fffffffu] : sp.lastw0[id];
        stale |= lw > e0;
      }
      redo = __syncthreads_or(stale);
    }
    __syncthreads();
    SPEC_T(1)
    if (redo) {
      spec_node_again<CPL, U>(I, p, sp, j, si, j, -(j + 1));  // head of the order: nothing it reads can be stale
      n_redo++;
      const int status = ldv(&sl->status);
      if (status != 0) {  // the insertion itself fails (graph.go:501-505, visited overflow) or does not fit the overlay
        if (tid == 0) {
          S->err = status;
          S->err_node = j;
          __threadfence();
          stv(&S->abort_, 1);
        }
        break;
      }
    } else {
      n_acc++;
    }
    SPEC_T(2)
    // ---- commit: the overlay becomes the graph ----------------------------------------------------------------------
    const int nov = ldv(&sl->n_ov);
    const int32_t *ovl = sp.ovls + (size_t)si * (OV_MAX + 1) * g.M;
    const uint32_t *ovid = sp.ovids + (size_t)si * OV_MAX;
    for (int e = warp; e < nov; e += INS_WARPS) {
      const uint32_t id = __ldcg(ovid + e);
      int32_t *dst = I.base_list(id);
      for (int i = lane; i < g.M; i += 32) dst[i] = __ldcg(ovl + (size_t)e * g.M + i);
      if (lane == 0) {
        if (id & 0x80000000u) sp.lastwU[id & 0x7fffffffu] = j + 1;
        else sp.lastw0[id] = j + 1;
      }
    }
    if (tid == 0) {
      const unsigned long long em = *reinterpret_cast<const volatile unsigned long long *>(&sl->ent_mask);
      const int nl = ldv(&sl->new_layers);
      if (em || nl) {
        const int row = (int)(p.first_row + j);
        for (int l = 0; l < B200HNSW_MAX_LAYERS; l++)
          if ((em >> l) & 1ull) stv(&S->entries[l], row);
        if (nl) stv(&S->n_layers, nl);
        __threadfence();
        stv(&S->sver, ldv(&S->sver) + 1);  // after the state: a reader that pairs new state with the old version is rejected
      }
      for (int i = 0; i < 6; i++) acc[i] += *reinterpret_cast<const volatile unsigned long long *>(&sl->ctr[i]);
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) stv(&S->done, j + 1);
    SPEC_T(3)
  }
  if (tid == 0) {
    atomicAdd(&p.counters[0], acc[0]);
    atomicAdd(&p.counters[1], acc[1]);
    atomicAdd(&p.counters[2], acc[2]);
    atomicAdd(&p.counters[3], acc[3]);
    atomicAdd(&p.counters[4], acc[4]);
    atomicAdd(&p.counters[5], acc[5]);
    S->n_accepted = n_acc;
    S->n_redone = n_redo;
    for (int i = 0; i < 4; i++) S->cyc[i] = (unsigned long long)cyc[i];
  }
}

template <int CPL, int U>
static int launch_spec_t(b200hnsw *h, const MutGraph &g, InsertParams &p, const SpecParams &sp, cudaStream_t s) {
  auto kern = spec_insert_kernel<CPL, U>;
  InsLayout lay(p, g.M);
  const size_t total = lay.total + (sizeof(uint32_t) + sizeof(int)) * OV_HASH;
  if (total > h->smem_optin) return 1;  // the caller runs the sequential kernel instead
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)total));
  kern<<<1 + sp.depth, INS_THREADS, total, s>>>(g, p, sp);
  B200_CUDA(cudaGetLastError());
  return 0;
}

// Window depth (= worker CTAs).  Two insertions collide when one reads a list the other writes (~200 lists read, ~55
// written per node: probability ~1e4 / rows for a given pair).  A node that collides with one of its R*T nearest
// predecessors (R = commit rate, T = one insertion's latency) has to be run again after that predecessor is in, which
// stalls the committer: R ~ 1 / (T * sqrt(1e4 / 2 rows)) is the ceiling, ~29 k/s at 1 M rows.  The window only has to be
// deep enough to keep that many insertions in flight; deeper ones cost nothing but idle workers, because a waiting
// worker repairs its own speculation.  Measured (profiles/r8o_k2s_*.log, r8n_k2s_1m.log): 5-9 k rows: 4 -> 1.3x the
// one-node-at-a-time kernel; 20-36 k: 16 -> 2.1x; 100 k: 32 -> 3.0x; 1 M: 32 -> 5.2x (48: the same).
int spec_window_auto(int64_t rows) {
  if (rows < 2000) return 1;
  if (rows < 10000) return 4;
  if (rows < 50000) return 16;
  return 32;
}

// One launch: nodes [0, p.n) of the batch slice described by p, window `depth`.  On return *done nodes are in the graph
// and n_layers / entries are the layer state after them; *err is 0, or the error of node *done (1, 2 as the sequential
// kernel reports them, 8 = that node needs the sequential kernel, 9 = internal wait timed out).
// Returns 1 (nothing launched) when the kernel does not fit this device's shared memory.
int launch_spec_insert(b200hnsw *h, InsertParams &p, int depth, int *done, int *err, int32_t *n_layers,
                       int32_t *entries, unsigned long long *accepted_redone, cudaStream_t s) {
  MutGraph g = mut_graph_of(h);
  size_insert_scratch(p, g);
  if (depth > h->sm_count - 1) depth = h->sm_count - 1;
  if (depth > p.n) depth = p.n;
  const int rl_cap = 16384;
  auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
  const size_t o_slots = al(sizeof(SpecShared));
  const size_t o_rlogs = o_slots + al(sizeof(SpecSlot) * depth);
  const size_t o_ovids = o_rlogs + al(sizeof(uint32_t) * (size_t)depth * rl_cap);
  const size_t o_ovls = o_ovids + al(sizeof(uint32_t) * (size_t)depth * OV_MAX);
  const size_t o_lw0 = o_ovls + al(sizeof(int32_t) * (size_t)depth * (OV_MAX + 1) * g.M);
  const size_t o_lwU = o_lw0 + al(sizeof(int) * (size_t)h->d.cap);
  const size_t total = o_lwU + al(sizeof(int) * (size_t)(h->d.capU > 0 ? h->d.capU : 1));
  if (total > h->spec_bytes) {
    if (h->d_spec) cudaFree(h->d_spec);
    h->d_spec = nullptr;
    h->spec_bytes = 0;
    B200_CUDA(cudaMalloc(&h->d_spec, total));
    h->spec_bytes = total;
  }
  char *base = (char *)h->d_spec;
  SpecParams sp{};
  sp.S = (SpecShared *)base;
  sp.slots = (SpecSlot *)(base + o_slots);
  sp.rlogs = (uint32_t *)(base + o_rlogs);
  sp.ovids = (uint32_t *)(base + o_ovids);
  sp.ovls = (int32_t *)(base + o_ovls);
  sp.lastw0 = (int *)(base + o_lw0);
  sp.lastwU = (int *)(base + o_lwU);
  sp.depth = depth;
  sp.rl_cap = rl_cap;
  SpecShared S0{};
  S0.n_layers = p.n_layers;
  memcpy(S0.entries, p.entries, sizeof S0.entries);
  B200_CUDA(cudaMemsetAsync(base + o_slots, 0, o_rlogs - o_slots, s));
  B200_CUDA(cudaMemsetAsync(base + o_lw0, 0, total - o_lw0, s));
  B200_CUDA(cudaMemcpyAsync(sp.S, &S0, sizeof S0, cudaMemcpyHostToDevice, s));
  int rc;
  const int cpl = (g.row_chunks + 31) / 32;
  switch (cpl) {
    case 1: rc = launch_spec_t<1, 4>(h, g, p, sp, s); break;
    case 2: rc = launch_spec_t<2, 4>(h, g, p, sp, s); break;
    case 3:
    case 4: rc = launch_spec_t<4, 2>(h, g, p, sp, s); break;
    case 5:
    case 6:
    case 7:
    case 8: rc = launch_spec_t<8, 1>(h, g, p, sp, s); break;
    default: rc = 1;  // wider rows: the sequential kernel
  }
  if (rc) return rc;
  SpecShared S1;
  B200_CUDA(cudaMemcpyAsync(&S1, sp.S, sizeof S1, cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  static const bool trace = getenv("B200HNSW_TRACE_ADD") != nullptr;
  if (trace)
    fprintf(stderr, "[b200hnsw] windowed add: first_row %lld n %d window %d -> done %d err %d (node %d) layers %d -> %d "
            "accepted %llu run again %llu (repaired by workers %llu)\n", p.first_row, p.n, depth, S1.done, S1.err,
            S1.err_node, p.n_layers, S1.n_layers, S1.n_accepted, S1.n_redone, S1.n_repaired);
  if (trace)
    fprintf(stderr, "[b200hnsw]   committer cycles per node: wait %.0f check %.0f run-again %.0f commit %.0f\n",
            (double)S1.cyc[0] / p.n, (double)S1.cyc[1] / p.n, (double)S1.cyc[2] / p.n, (double)S1.cyc[3] / p.n);
  *done = S1.done;
  *err = S1.err;
  *n_layers = S1.n_layers;
  memcpy(entries, S1.entries, sizeof S1.entries);
  if (accepted_redone) {
    accepted_redone[0] += S1.n_accepted;
    accepted_redone[1] += S1.n_redone;
    accepted_redone[2] += S1.n_repaired;
  }
  return 0;
}

}  // namespace b200
