// misc_kernels.cu -- b200hnsw_distance_batch: EuclideanDistance / CosineDistance (distance.go:15-23) on pairs,
// one warp per pair, canonical order (dist.cuh).
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

}  // namespace b200

using namespace b200;

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
