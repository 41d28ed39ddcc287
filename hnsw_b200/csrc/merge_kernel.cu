// merge_kernel.cu -- K4: merge G per-shard result lists per query into one top-k (multi-GPU search and exact search).
// No reference counterpart (the reference is single-process); the tie rule mirrors the stable
// concatenate-then-sort of hnsw-extensions/hybrid/adapter.go:264-292: equal distances keep shard order, then the
// position inside the shard's list.  One CTA per query: the G*k candidates are bitonic-sorted in shared memory on
// a 64-bit key (order-preserving distance bits << 32 | shard*k + position); NaN distances sort last.
// Bound: launch latency / NVLink all-gather latency (G*k*12 bytes per query).
#include "index.hpp"

namespace b200 {

__device__ __forceinline__ uint32_t ordered_bits(float d) {
  uint32_t u = __float_as_uint(d);
  if (d != d) return 0xffffffffu;                      // NaN last
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);   // total order of the finite/inf floats
}

// PACKED = false: dists [G][B][k], ids int64 [G][B][k], counts [G][B].
// PACKED = true : `packed` holds, per shard g at packed + g * shard_stride (int32 elements), what one K1 launch wrote:
//                 rows int32 [B][k] | dists f32 [B][k] | counts int32 [B]; the global id of a row is g * id_stride + row.
//                 This is the buffer an all-gather of the per-rank result buffers produces: no repacking in between.
template <bool PACKED>
__global__ void merge_topk_kernel(int G, long long B, int k, const float *__restrict__ dists,
                                  const int64_t *__restrict__ ids, const int32_t *__restrict__ counts,
                                  const int32_t *__restrict__ packed, long long shard_stride, long long id_stride,
                                  int n_pow2, float *__restrict__ out_d, int64_t *__restrict__ out_id,
                                  int32_t *__restrict__ out_cnt) {
  extern __shared__ unsigned long long keys[];
  const long long q = blockIdx.x;
  const int total = G * k;
  for (int i = threadIdx.x; i < n_pow2; i += blockDim.x) {
    unsigned long long key = ~0ull;
    if (i < total) {
      int g = i / k, pos = i - g * k;
      int cnt;
      float d = 0.f;
      if constexpr (PACKED) {
        const int32_t *sh = packed + g * shard_stride;
        cnt = sh[2 * B * k + q];
        if (pos < cnt) d = __int_as_float(sh[B * k + q * k + pos]);
      } else {
        cnt = counts[(long long)g * B + q];
        if (pos < cnt) d = dists[((long long)g * B + q) * k + pos];
      }
      if (pos < cnt) {
        key = ((unsigned long long)ordered_bits(d) << 32) | (unsigned)i;
      }
    }
    keys[i] = key;
  }
  __syncthreads();
  for (int size = 2; size <= n_pow2; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = threadIdx.x; i < n_pow2 / 2; i += blockDim.x) {
        int lo = 2 * i - (i & (stride - 1));
        int hi = lo + stride;
        bool up = (lo & size) == 0;
        unsigned long long a = keys[lo], b = keys[hi];
        if ((a > b) == up) {
          keys[lo] = b;
          keys[hi] = a;
        }
      }
      __syncthreads();
    }
  }
  int valid = 0;
  for (int i = threadIdx.x; i < k; i += blockDim.x) {
    unsigned long long key = keys[i];
    bool ok = key != ~0ull;
    int src = (int)(key & 0xffffffffu);
    int g = ok ? src / k : 0, pos = ok ? src - g * k : 0;
    float d = __int_as_float(0x7fc00000);
    long long id = -1;
    if (ok) {
      if constexpr (PACKED) {
        const int32_t *sh = packed + g * shard_stride;
        d = __int_as_float(sh[B * k + q * k + pos]);
        id = g * id_stride + sh[q * k + pos];
      } else {
        const long long off = ((long long)g * B + q) * k + pos;
        d = dists[off];
        id = ids[off];
      }
    }
    out_d[q * k + i] = d;
    out_id[q * k + i] = id;
    valid += ok;
  }
  // count = min(k, sum of counts)
  if (threadIdx.x == 0) {
    long long s = 0;
    for (int g = 0; g < G; g++) {
      int c = PACKED ? packed[g * shard_stride + 2 * B * k + q] : counts[(long long)g * B + q];
      s += c > 0 ? c : 0;  // a failed traversal reports -1
    }
    out_cnt[q] = (int)(s < k ? s : k);
  }
  (void)valid;
}

// The same merge with ONE WARP per query for G*k <= 128: the keys live in registers (NPL per lane, element j*32+lane),
// compare-exchange steps with stride >= 32 are between a lane's own registers, the others are shuffles.  No shared
// memory, no block barriers, 8 queries per block: on the multi-GPU path this kernel runs on the side stream between
// the traversal kernels' blocks, so it has to be small (the one-block-per-query kernel took 27 us of a 150 us step).
template <bool PACKED, int NPL>
__global__ void __launch_bounds__(256) merge_topk_warp_kernel(int G, long long B, int k, const float *__restrict__ dists,
                                                               const int64_t *__restrict__ ids,
                                                               const int32_t *__restrict__ counts,
                                                               const int32_t *__restrict__ packed, long long shard_stride,
                                                               long long id_stride, float *__restrict__ out_d,
                                                               int64_t *__restrict__ out_id,
                                                               int32_t *__restrict__ out_cnt) {
  const int lane = threadIdx.x & 31;
  const long long q = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (q >= B) return;
  const int total = G * k;
  unsigned long long key[NPL];
  int csum = 0;
#pragma unroll
  for (int j = 0; j < NPL; j++) {
    const int i = j * 32 + lane;
    key[j] = ~0ull;
    if (i < total) {
      const int g = i / k, pos = i - g * k;
      int cnt;
      float d = 0.f;
      if constexpr (PACKED) {
        const int32_t *sh = packed + g * shard_stride;
        cnt = sh[2 * B * k + q];
        if (pos < cnt) d = __int_as_float(sh[B * k + q * k + pos]);
      } else {
        cnt = counts[(long long)g * B + q];
        if (pos < cnt) d = dists[((long long)g * B + q) * k + pos];
      }
      if (pos < cnt) key[j] = ((unsigned long long)ordered_bits(d) << 32) | (unsigned)i;
      if (pos == 0 && cnt > 0) csum += cnt;  // one element per shard carries its count (a failed traversal reports -1)
    }
  }
  constexpr int N = NPL * 32;
#pragma unroll
  for (int size = 2; size <= N; size <<= 1) {
#pragma unroll
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      if (stride >= 32) {
        constexpr int dummy = 0;
        (void)dummy;
        const int js = stride >> 5;
#pragma unroll
        for (int j = 0; j < NPL; j++) {
          if ((j & js) == 0) {  // j is the lower element of the pair (j, j + js)
            const int i = j * 32 + lane;
            const bool up = (i & size) == 0 || size == N;
            const unsigned long long a = key[j], b = key[j | js];
            const bool sw = (a > b) == up;
            key[j] = sw ? b : a;
            key[j | js] = sw ? a : b;
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < NPL; j++) {
          const int i = j * 32 + lane;
          const bool up = (i & size) == 0 || size == N;
          const bool lower = (lane & stride) == 0;
          const unsigned long long a = key[j];
          const unsigned long long b = __shfl_xor_sync(0xffffffffu, a, stride);
          const bool want_min = lower == up;
          key[j] = want_min ? (a < b ? a : b) : (a > b ? a : b);
        }
      }
    }
  }
  // ascending: element i < k is the i-th result
#pragma unroll
  for (int j = 0; j < NPL; j++) {
    const int i = j * 32 + lane;
    if (i < k) {
      const unsigned long long kk = key[j];
      const bool ok = kk != ~0ull;
      const int src = (int)(kk & 0xffffffffu);
      const int g = ok ? src / k : 0, pos = ok ? src - g * k : 0;
      float d = __int_as_float(0x7fc00000);
      long long id = -1;
      if (ok) {
        if constexpr (PACKED) {
          const int32_t *sh = packed + g * shard_stride;
          d = __int_as_float(sh[B * k + q * k + pos]);
          id = g * id_stride + sh[q * k + pos];
        } else {
          const long long off = ((long long)g * B + q) * k + pos;
          d = dists[off];
          id = ids[off];
        }
      }
      out_d[q * k + i] = d;
      out_id[q * k + i] = id;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) csum += __shfl_xor_sync(0xffffffffu, csum, o);
  if (lane == 0) out_cnt[q] = csum < k ? csum : k;
}

}  // namespace b200

using namespace b200;

static int launch_merge(int32_t device, int32_t G, int64_t B, int32_t k, const float *d_dists, const int64_t *d_ids,
                        const int32_t *d_counts, const int32_t *d_packed, int64_t shard_stride, int64_t id_stride,
                        float *d_out_dists, int64_t *d_out_ids, int32_t *d_out_counts, void *stream) {
  if (B == 0) return 0;
  B200_CUDA(cudaSetDevice(device));
  if (G * k <= 128 && k <= 128) {  // one warp per query, keys in registers
    const int npl = G * k <= 32 ? 1 : (G * k <= 64 ? 2 : 4);
    const unsigned blocks = (unsigned)((B + 7) / 8);
    cudaStream_t st = (cudaStream_t)stream;
#define WARP_MERGE(P, N)                                                                                              \
  merge_topk_warp_kernel<P, N><<<blocks, 256, 0, st>>>(G, B, k, d_dists, d_ids, d_counts, d_packed, shard_stride,    \
                                                       id_stride, d_out_dists, d_out_ids, d_out_counts)
    if (d_packed) {
      if (npl == 1) WARP_MERGE(true, 1);
      else if (npl == 2) WARP_MERGE(true, 2);
      else WARP_MERGE(true, 4);
    } else {
      if (npl == 1) WARP_MERGE(false, 1);
      else if (npl == 2) WARP_MERGE(false, 2);
      else WARP_MERGE(false, 4);
    }
#undef WARP_MERGE
    B200_CUDA(cudaGetLastError());
    return 0;
  }
  int n_pow2 = 2;
  while (n_pow2 < G * k) n_pow2 <<= 1;
  int threads = n_pow2 / 2 < 32 ? 32 : (n_pow2 / 2 > 256 ? 256 : n_pow2 / 2);
  size_t smem = sizeof(unsigned long long) * (size_t)n_pow2;
  auto kern = d_packed ? merge_topk_kernel<true> : merge_topk_kernel<false>;
  if (smem > 48 * 1024) B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<(unsigned)B, threads, smem, (cudaStream_t)stream>>>(G, B, k, d_dists, d_ids, d_counts, d_packed, shard_stride,
                                                             id_stride, n_pow2, d_out_dists, d_out_ids, d_out_counts);
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200hnsw_merge_topk_device(int32_t device, int32_t G, int64_t B, int32_t k, const float *d_dists,
                                          const int64_t *d_ids, const int32_t *d_counts, int32_t sorted_in,
                                          float *d_out_dists, int64_t *d_out_ids, int32_t *d_out_counts,
                                          void *stream) {
  (void)sorted_in;  // both input orders go through the same sort
  if (G <= 0 || B < 0 || k <= 0 || k > B200HNSW_MAX_K || (long long)G * k > 8192 ||
      (B > 0 && (!d_dists || !d_ids || !d_counts || !d_out_dists || !d_out_ids || !d_out_counts))) {
    set_error("merge_topk: bad argument (G=%d k=%d; G*k <= 8192)", G, k);
    return B200HNSW_ERR_BAD_ARG;
  }
  return launch_merge(device, G, B, k, d_dists, d_ids, d_counts, nullptr, 0, 0, d_out_dists, d_out_ids, d_out_counts,
                      stream);
}

extern "C" int b200hnsw_merge_shards_device(int32_t device, int32_t G, int64_t B, int32_t k, const int32_t *d_packed,
                                            int64_t shard_stride, int64_t id_stride, float *d_out_dists,
                                            int64_t *d_out_ids, int32_t *d_out_counts, void *stream) {
  if (G <= 0 || B < 0 || k <= 0 || k > B200HNSW_MAX_K || (long long)G * k > 8192 ||
      shard_stride < 2 * B * k + B || (B > 0 && (!d_packed || !d_out_dists || !d_out_ids || !d_out_counts))) {
    set_error("merge_shards: bad argument (G=%d k=%d; G*k <= 8192; shard_stride >= 2*B*k + B)", G, k);
    return B200HNSW_ERR_BAD_ARG;
  }
  return launch_merge(device, G, B, k, nullptr, nullptr, nullptr, d_packed, shard_stride, id_stride, d_out_dists,
                      d_out_ids, d_out_counts, stream);
}
