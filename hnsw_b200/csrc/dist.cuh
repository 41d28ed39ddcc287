// dist.cuh -- warp-cooperative EuclideanDistance / CosineDistance (reference: distance.go:15-23).
//
// Canonical summation order (shared with oracle/hnsw_oracle.c so results are bit-identical):
//   the padded row is cut in float4 chunks; chunk c belongs to lane (c mod 32); each lane runs ONE fused
//   multiply-add chain over its elements in index order; the 32 partial sums are combined by an xor butterfly
//   with offsets 16, 8, 4, 2, 1.  Rows are zero-padded to a multiple of 32 floats (128 B), which adds exact
//   zeros to the chains and therefore changes nothing.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ float butterfly(float v) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) v = __fadd_rn(v, __shfl_xor_sync(FULL, v, off));
  return v;
}

// read-only, streaming 16 B load (vectors are immutable once written; rows are rarely re-used by the same SM)
__device__ __forceinline__ float4 ldg_f4(const float4 *p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}

// Query held in registers: lane L keeps chunks L, L+32, ... (CPL of them; chunks >= nchunks are zero).
template <int CPL>
struct QueryRegs {
  float4 q[CPL];
  float qq;  // |q|^2 in canonical order (cosine)
};

template <int CPL>
__device__ __forceinline__ void load_query(QueryRegs<CPL> &Q, const float *query, int dim, int nchunks, int lane) {
  // the caller's query is [dim] floats, unpadded and possibly unaligned for float4: scalar loads, once
  float acc = 0.f;
#pragma unroll
  for (int c = 0; c < CPL; c++) {
    int chunk = lane + 32 * c;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (chunk < nchunks) {
#pragma unroll
      for (int e = 0; e < 4; e++) {
        int i = chunk * 4 + e;
        if (i < dim) v[e] = query[i];
      }
    }
    Q.q[c] = make_float4(v[0], v[1], v[2], v[3]);
#pragma unroll
    for (int e = 0; e < 4; e++) acc = __fmaf_rn(v[e], v[e], acc);
  }
  Q.qq = butterfly(acc);
}

// issue the loads of one stored row (lane's chunks); nchunks = padded_dim/4
template <int CPL>
__device__ __forceinline__ void load_row(float4 (&r)[CPL], const float4 *row, int nchunks, int lane) {
#pragma unroll
  for (int c = 0; c < CPL; c++) {
    int chunk = lane + 32 * c;
    r[c] = (chunk < nchunks) ? ldg_f4(row + chunk) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// distance(row, query): row plays `a` (neighbor.Value), the query plays `b` (target)  -- graph.go:146
template <int CPL>
__device__ __forceinline__ float finish_distance(const float4 (&r)[CPL], const QueryRegs<CPL> &Q, int metric) {
  if (metric == 0) {
    float acc = 0.f;
#pragma unroll
    for (int c = 0; c < CPL; c++) {
      float dx = __fsub_rn(r[c].x, Q.q[c].x), dy = __fsub_rn(r[c].y, Q.q[c].y);
      float dz = __fsub_rn(r[c].z, Q.q[c].z), dw = __fsub_rn(r[c].w, Q.q[c].w);
      acc = __fmaf_rn(dx, dx, acc);
      acc = __fmaf_rn(dy, dy, acc);
      acc = __fmaf_rn(dz, dz, acc);
      acc = __fmaf_rn(dw, dw, acc);
    }
    return __fsqrt_rn(butterfly(acc));
  }
  float dot = 0.f, na = 0.f;
#pragma unroll
  for (int c = 0; c < CPL; c++) {
    dot = __fmaf_rn(r[c].x, Q.q[c].x, dot);
    dot = __fmaf_rn(r[c].y, Q.q[c].y, dot);
    dot = __fmaf_rn(r[c].z, Q.q[c].z, dot);
    dot = __fmaf_rn(r[c].w, Q.q[c].w, dot);
    na = __fmaf_rn(r[c].x, r[c].x, na);
    na = __fmaf_rn(r[c].y, r[c].y, na);
    na = __fmaf_rn(r[c].z, r[c].z, na);
    na = __fmaf_rn(r[c].w, r[c].w, na);
  }
  dot = butterfly(dot);
  na = butterfly(na);
  // 1 - dot / sqrt(|a|^2 * |b|^2)
  return __fsub_rn(1.0f, __fdiv_rn(dot, __fsqrt_rn(__fmul_rn(na, Q.qq))));
}

// generic-dimension variant (any number of chunks): lanes stride over chunks, both operands in memory
__device__ __forceinline__ float warp_distance_generic(const float4 *a, const float4 *b, int nchunks, int metric,
                                                       int lane) {
  if (metric == 0) {
    float acc = 0.f;
    for (int chunk = lane; chunk < nchunks; chunk += 32) {
      float4 x = a[chunk], y = b[chunk];
      float dx = __fsub_rn(x.x, y.x), dy = __fsub_rn(x.y, y.y), dz = __fsub_rn(x.z, y.z), dw = __fsub_rn(x.w, y.w);
      acc = __fmaf_rn(dx, dx, acc);
      acc = __fmaf_rn(dy, dy, acc);
      acc = __fmaf_rn(dz, dz, acc);
      acc = __fmaf_rn(dw, dw, acc);
    }
    return __fsqrt_rn(butterfly(acc));
  }
  float dot = 0.f, na = 0.f, nb = 0.f;
  for (int chunk = lane; chunk < nchunks; chunk += 32) {
    float4 x = a[chunk], y = b[chunk];
    dot = __fmaf_rn(x.x, y.x, dot);
    dot = __fmaf_rn(x.y, y.y, dot);
    dot = __fmaf_rn(x.z, y.z, dot);
    dot = __fmaf_rn(x.w, y.w, dot);
    na = __fmaf_rn(x.x, x.x, na);
    na = __fmaf_rn(x.y, x.y, na);
    na = __fmaf_rn(x.z, x.z, na);
    na = __fmaf_rn(x.w, x.w, na);
    nb = __fmaf_rn(y.x, y.x, nb);
    nb = __fmaf_rn(y.y, y.y, nb);
    nb = __fmaf_rn(y.z, y.z, nb);
    nb = __fmaf_rn(y.w, y.w, nb);
  }
  dot = butterfly(dot);
  na = butterfly(na);
  nb = butterfly(nb);
  return __fsub_rn(1.0f, __fdiv_rn(dot, __fsqrt_rn(__fmul_rn(na, nb))));
}

}  // namespace b200
