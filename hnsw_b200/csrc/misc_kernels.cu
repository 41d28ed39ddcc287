// misc_kernels.cu -- b200hnsw_distance_batch: EuclideanDistance / CosineDistance (distance.go:15-23) on pairs,
// one warp per pair, canonical order (dist.cuh).
#include <cuda.h>

#include <cstdlib>
#include <cstring>
#include <vector>

#include "index.hpp"

namespace b200 {

__global__ void distance_pairs_kernel(const float4 *a, const float4 *b, long long n, int nchunks, int metric,
                                      float *out) {
  long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  if (w >= n) return;
  float d = warp_distance_generic(a + w * nchunks, b + w * nchunks, nchunks, metric, lane);
  if (lane == 0) out[w] = d;
}

// One warp per row: |x|^2 (any summation order will do: the test is a coarse range check).  Rows whose squared norm is
// zero, non-finite or outside [1e-18, 1e18] can produce NaN distances (distance.go:15-17: 0/0 under cosine; inf - inf).
__global__ void flag_rows_kernel(const float4 *vec, long long first, long long count, int nchunks, int *flags) {
  long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= count) return;
  const float4 *row = vec + (first + w) * nchunks;
  float s = 0.f;
  for (int c = lane; c < nchunks; c += 32) {
    float4 v = row[c];
    s += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  for (int off = 16; off >= 1; off >>= 1) s += __shfl_xor_sync(FULL, s, off);
  if (lane == 0 && !(s >= 1e-18f && s <= 1e18f)) atomicOr(flags, 1);
}

// ---- flags of the fused search + exchange (sharded.py P2PShardPipeline) -------------------------------------------------
struct PeerFlagPtrs {
  int32_t *p[B200HNSW_MAX_PEERS];
};
// everything this GPU wrote before (the traversal kernel's peer stores) is made visible system-wide, then the flags
__global__ void peer_signal_kernel(PeerFlagPtrs f, int n, int value) {
  __threadfence_system();
  if ((int)threadIdx.x < n)
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(f.p[threadIdx.x]), "r"(value) : "memory");
}
__global__ void peer_wait_kernel(const int32_t *flags, int n, int value) {
  if ((int)threadIdx.x < n) {
    int v;
    do {
      asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(flags + threadIdx.x) : "memory");
      if (v < value) __nanosleep(200);
    } while (v < value);
  }
  __syncthreads();
  __threadfence_system();
}

}  // namespace b200

using namespace b200;

int b200hnsw::flag_rows(int64_t first, int64_t count, cudaStream_t s) {
  if (count <= 0) return 0;
  const long long blocks = (count * 32 + 255) / 256;
  flag_rows_kernel<<<(unsigned)blocks, 256, 0, s>>>(reinterpret_cast<const float4 *>(d.vec), first, count,
                                                     row_floats / 4, d_row_flags);
  B200_CUDA(cudaGetLastError());
  return 0;
}
int b200hnsw::refresh_row_flags(cudaStream_t s) {
  int v = 0;
  B200_CUDA(cudaMemcpyAsync(&v, d_row_flags, sizeof v, cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  row_flags.store(v);
  return 0;
}

extern "C" int b200hnsw_peer_alloc(int32_t device, int64_t bytes, void **d_ptr, uint8_t handle[64]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  if (bytes <= 0 || !d_ptr || !handle) {
    set_error("peer_alloc: bad argument");
    return B200HNSW_ERR_BAD_ARG;
  }
  B200_CUDA(cudaSetDevice(device));
  void *p = nullptr;
  B200_CUDA(cudaMalloc(&p, (size_t)bytes));
  cudaError_t e = cudaMemset(p, 0, (size_t)bytes);
  cudaIpcMemHandle_t hd;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&hd, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return cuda_fail(e, "peer_alloc");
  }
  memcpy(handle, &hd, 64);
  *d_ptr = p;
  return 0;
}
extern "C" int b200hnsw_peer_open(int32_t device, const uint8_t handle[64], void **d_ptr) {
  if (!d_ptr || !handle) {
    set_error("peer_open: bad argument");
    return B200HNSW_ERR_BAD_ARG;
  }
  B200_CUDA(cudaSetDevice(device));
  cudaIpcMemHandle_t hd;
  memcpy(&hd, handle, 64);
  B200_CUDA(cudaIpcOpenMemHandle(d_ptr, hd, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
extern "C" int b200hnsw_peer_close(void *d_ptr) {
  if (d_ptr) B200_CUDA(cudaIpcCloseMemHandle(d_ptr));
  return 0;
}
extern "C" int b200hnsw_peer_free(void *d_ptr) {
  if (d_ptr) B200_CUDA(cudaFree(d_ptr));
  return 0;
}
// Stream memory operations (cuStreamBatchMemOp): the flag writes and waits are executed by the stream's front end, not by
// a kernel, so they need no SM slot.  That matters here: with three traversal launches in flight a one-block flag kernel
// queues behind every block of the launches issued before it.  B200HNSW_PEER_FLAGS=kernel selects the kernels.
typedef CUresult (*BatchMemOpFn)(CUstream, unsigned int, CUstreamBatchMemOpParams *, unsigned int);
static BatchMemOpFn batch_memop() {
  static BatchMemOpFn fn = []() -> BatchMemOpFn {
    const char *v = getenv("B200HNSW_PEER_FLAGS");
    if (v && strcmp(v, "kernel") == 0) return nullptr;
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuStreamBatchMemOp", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      return nullptr;
    }
    return (BatchMemOpFn)p;
  }();
  return fn;
}

extern "C" int b200hnsw_peer_signal_device(int32_t device, int32_t *const *flag_ptrs, int32_t n, int32_t value,
                                           void *stream) {
  if (n <= 0 || n > B200HNSW_MAX_PEERS || !flag_ptrs) {
    set_error("peer_signal: bad argument (n=%d, at most %d)", n, B200HNSW_MAX_PEERS);
    return B200HNSW_ERR_BAD_ARG;
  }
  B200_CUDA(cudaSetDevice(device));
  if (BatchMemOpFn op = batch_memop()) {
    CUstreamBatchMemOpParams ops[B200HNSW_MAX_PEERS];
    memset(ops, 0, sizeof ops);
    for (int i = 0; i < n; i++) {
      ops[i].writeValue.operation = CU_STREAM_MEM_OP_WRITE_VALUE_32;
      ops[i].writeValue.address = (CUdeviceptr)(uintptr_t)flag_ptrs[i];
      ops[i].writeValue.value = (cuuint32_t)value;
      ops[i].writeValue.flags = CU_STREAM_WRITE_VALUE_DEFAULT;  // preceded by a system-wide memory fence
    }
    CUresult r = op((CUstream)stream, (unsigned)n, ops, 0);
    if (r != CUDA_SUCCESS) {
      set_error("peer_signal: cuStreamBatchMemOp failed (%d)", (int)r);
      return B200HNSW_ERR_CUDA;
    }
    return 0;
  }
  PeerFlagPtrs f{};
  for (int i = 0; i < n; i++) f.p[i] = flag_ptrs[i];
  peer_signal_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(f, n, value);
  B200_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int b200hnsw_peer_wait_device(int32_t device, const int32_t *d_flags, int32_t n, int32_t value,
                                         void *stream) {
  if (n <= 0 || n > 32 || !d_flags) {
    set_error("peer_wait: bad argument (n=%d, at most 32)", n);
    return B200HNSW_ERR_BAD_ARG;
  }
  B200_CUDA(cudaSetDevice(device));
  if (BatchMemOpFn op = batch_memop()) {
    CUstreamBatchMemOpParams ops[32];
    memset(ops, 0, sizeof ops);
    for (int i = 0; i < n; i++) {
      ops[i].waitValue.operation = CU_STREAM_MEM_OP_WAIT_VALUE_32;
      ops[i].waitValue.address = (CUdeviceptr)(uintptr_t)(d_flags + i);
      ops[i].waitValue.value = (cuuint32_t)value;
      ops[i].waitValue.flags = CU_STREAM_WAIT_VALUE_GEQ;
    }
    CUresult r = op((CUstream)stream, (unsigned)n, ops, 0);
    if (r != CUDA_SUCCESS) {
      set_error("peer_wait: cuStreamBatchMemOp failed (%d)", (int)r);
      return B200HNSW_ERR_CUDA;
    }
    return 0;
  }
  peer_wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(d_flags, n, value);
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200hnsw_distance_batch(int32_t device, int32_t metric, const float *a, const float *b, int64_t n,
                                       int32_t dim, float *out) {
  if (n < 0 || dim <= 0 || (n > 0 && (!a || !b || !out)) || (metric != 0 && metric != 1)) {
    set_error("distance_batch: bad argument");
    return B200HNSW_ERR_BAD_ARG;
  }
  if (n == 0) return 0;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    set_error("distance_batch: no CUDA device (%s); no CPU fallback", cudaGetErrorString(e));
    return B200HNSW_ERR_CUDA;
  }
  B200_CUDA(cudaSetDevice(device));
  int rf = (dim + 31) / 32 * 32;
  float *da = nullptr, *db = nullptr, *dout = nullptr;
  B200_CUDA(cudaMalloc(&da, sizeof(float) * (size_t)n * rf));
  B200_CUDA(cudaMalloc(&db, sizeof(float) * (size_t)n * rf));
  B200_CUDA(cudaMalloc(&dout, sizeof(float) * (size_t)n));
  int rc = 0;
  auto ck = [&](cudaError_t err, const char *what) {
    if (err != cudaSuccess && !rc) rc = cuda_fail(err, what);
  };
  ck(cudaMemset(da, 0, sizeof(float) * (size_t)n * rf), "memset");
  ck(cudaMemset(db, 0, sizeof(float) * (size_t)n * rf), "memset");
  ck(cudaMemcpy2D(da, sizeof(float) * rf, a, sizeof(float) * dim, sizeof(float) * dim, (size_t)n,
                  cudaMemcpyHostToDevice), "copy a");
  ck(cudaMemcpy2D(db, sizeof(float) * rf, b, sizeof(float) * dim, sizeof(float) * dim, (size_t)n,
                  cudaMemcpyHostToDevice), "copy b");
  if (!rc) {
    int threads = 128;
    long long blocks = (n * 32 + threads - 1) / threads;
    distance_pairs_kernel<<<(unsigned)blocks, threads>>>((const float4 *)da, (const float4 *)db, n, rf / 4, metric,
                                                         dout);
    ck(cudaGetLastError(), "distance kernel");
    ck(cudaMemcpy(out, dout, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost), "copy out");
  }
  cudaFree(da);
  cudaFree(db);
  cudaFree(dout);
  return rc;
}
