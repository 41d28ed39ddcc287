// search_kernel.cu -- K1: batched Graph.Search / Graph.BatchSearch (graph.go:571-593, 1075-1106).
// One warp per query; queries are striped over a grid sized to fill the SMs.  Bound: HBM gather
// (E rows of 4*dim bytes per query, E counted per launch in d_counters).
#include "index.hpp"

namespace b200 {

struct SearchParams {
  const float *queries;  // [B][dim] device
  long long B;
  int k, ef, metric, mode, n_layers;
  int entries[B200HNSW_MAX_LAYERS];
  int32_t *out_rows;   // [B][k]
  float *out_dists;    // [B][k]
  int32_t *out_counts; // [B]
  unsigned long long *counters;  // [4]: evals, expansions, nbr_reads, overflowed queries (may be null)
  int *error;                    // set to 1 on visited overflow
  int vis_cap;                   // visited-table entries per warp (power of two)
  int heap_c, heap_r;            // uint2 slots reserved per warp for the two queues
};

constexpr int WARPS_PER_BLOCK = 4;

template <int CPL, int U>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32) search_kernel(GraphView g, SearchParams p) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const size_t per_warp = (size_t)(p.heap_c + p.heap_r) * sizeof(uint2) + (size_t)p.vis_cap * sizeof(uint32_t);
  unsigned char *base = smem + per_warp * warp;
  GoHeap C{reinterpret_cast<uint2 *>(base), 0};
  GoHeap R{reinterpret_cast<uint2 *>(base) + p.heap_c, 0};
  Visited V{reinterpret_cast<uint32_t *>(base + (size_t)(p.heap_c + p.heap_r) * sizeof(uint2)),
            (uint32_t)p.vis_cap - 1, 0};
  SearchCounters ctr{0, 0, 0};
  unsigned long long overflow = 0;
  const long long stride = (long long)gridDim.x * WARPS_PER_BLOCK;
  for (long long q = (long long)blockIdx.x * WARPS_PER_BLOCK + warp; q < p.B; q += stride) {
    QueryRegs<CPL> Q;
    load_query<CPL>(Q, p.queries + q * g.dim, g.dim, g.row_chunks, lane);
    int elevator = -1;
    int n = 0;
    for (int layer = p.n_layers - 1; layer >= 0; layer--) {
      int sp = p.entries[layer];          // graph.go:572
      if (elevator >= 0) sp = elevator;   // graph.go:573-575
      if (sp < 0) continue;               // empty layer: search(nil) returns nil (graph.go:101-103, 580-583)
      if (layer > 0) {
        n = layer_search_ref<CPL, U>(g, layer, sp, 1, p.ef, p.metric, Q, C, R, V, ctr, lane);
        if (n < 0) break;
        if (n > 0) elevator = R.id(0);    // graph.go:584 (array published by the __syncwarp in layer_search_ref)
        __syncwarp();                     // all lanes have read R before lane 0 reuses it
        n = 0;
        continue;
      }
      n = layer_search_ref<CPL, U>(g, 0, sp, p.k, p.ef, p.metric, Q, C, R, V, ctr, lane);
    }
    if (n < 0) {
      overflow++;
      n = 0;
      if (lane == 0) {
        *p.error = 1;
        p.out_counts[q] = -1;
      }
    } else if (lane == 0) {
      p.out_counts[q] = n;
    }
    // graph.go:588-593: results in heap-array order
    for (int i = lane; i < p.k; i += 32) {
      bool ok = i < n;
      p.out_rows[q * p.k + i] = ok ? R.id(i) : -1;
      p.out_dists[q * p.k + i] = ok ? R.dist(i) : __int_as_float(0x7fc00000);
    }
    __syncwarp();
  }
  if (p.counters && lane == 0) {
    atomicAdd(&p.counters[0], ctr.evals);
    atomicAdd(&p.counters[1], ctr.expansions);
    atomicAdd(&p.counters[2], ctr.nbr_reads);
    if (overflow) atomicAdd(&p.counters[3], overflow);
  }
}

template <int CPL, int U>
static int launch_t(b200hnsw *h, const GraphView &g, SearchParams &p, cudaStream_t s) {
  auto kern = search_kernel<CPL, U>;
  size_t per_warp = (size_t)(p.heap_c + p.heap_r) * sizeof(uint2) + (size_t)p.vis_cap * sizeof(uint32_t);
  size_t smem = per_warp * WARPS_PER_BLOCK;
  if (smem > h->smem_optin) {
    set_error("search: k=%d ef=%d need %zu B of shared memory per block (> %zu)", p.k, p.ef, smem, h->smem_optin);
    return B200HNSW_ERR_UNSUPPORTED;
  }
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, WARPS_PER_BLOCK * 32, smem));
  if (per_sm < 1) per_sm = 1;
  long long blocks_needed = (p.B + WARPS_PER_BLOCK - 1) / WARPS_PER_BLOCK;
  long long resident = (long long)per_sm * h->sm_count;
  int grid = (int)(blocks_needed < resident ? blocks_needed : resident);
  if (grid < 1) grid = 1;
  kern<<<grid, WARPS_PER_BLOCK * 32, smem, s>>>(g, p);
  B200_CUDA(cudaGetLastError());
  return 0;
}

static int round_pow2(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

int launch_search(b200hnsw *h, const float *d_queries, int64_t B, int k, int ef, int metric, int mode,
                  const int32_t *entries, int32_t *d_rows, float *d_dists, int32_t *d_counts,
                  unsigned long long *d_counters, int *d_error, cudaStream_t s) {
  if (mode != B200HNSW_MODE_REFERENCE) {
    set_error("search: mode %d not implemented yet", mode);
    return B200HNSW_ERR_UNSUPPORTED;
  }
  SearchParams p{};
  p.queries = d_queries;
  p.B = B;
  p.k = k;
  p.ef = ef;
  p.metric = metric;
  p.mode = mode;
  p.n_layers = h->n_layers;
  for (int l = 0; l < B200HNSW_MAX_LAYERS; l++) p.entries[l] = l < h->n_layers ? (entries ? entries[l] : h->entry[l]) : -1;
  p.out_rows = d_rows;
  p.out_dists = d_dists;
  p.out_counts = d_counts;
  p.counters = d_counters;
  p.error = d_error;
  p.heap_c = (ef + 1 + 1) & ~1;  // keep 16 B alignment of what follows
  p.heap_r = (k + 1) & ~1;
  // the reference's rule stops at the first non-improving hop, so a layer search visits a few dozen nodes;
  // 1024 entries (load factor <= 1/2 => 512 visits) is ample and overflow is reported, never ignored
  p.vis_cap = round_pow2(1024);
  GraphView g = h->view();
  int cpl = (g.row_chunks + 31) / 32;
  switch (cpl) {
    case 1: return launch_t<1, 8>(h, g, p, s);
    case 2: return launch_t<2, 8>(h, g, p, s);
    case 3: return launch_t<3, 4>(h, g, p, s);
    case 4: return launch_t<4, 4>(h, g, p, s);
    case 5:
    case 6: return launch_t<6, 2>(h, g, p, s);
    case 7:
    case 8: return launch_t<8, 2>(h, g, p, s);
    case 9:
    case 10:
    case 11:
    case 12: return launch_t<12, 1>(h, g, p, s);
    case 13:
    case 14:
    case 15:
    case 16: return launch_t<16, 1>(h, g, p, s);
    default:
      set_error("search: dim %d > 2048 not supported", g.dim);
      return B200HNSW_ERR_UNSUPPORTED;
  }
}

}  // namespace b200
