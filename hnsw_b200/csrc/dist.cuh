// dist.cuh -- warp-cooperative EuclideanDistance / CosineDistance (reference: distance.go:15-23).
//
// Canonical summation order (shared with oracle/hnsw_oracle.c so results are bit-identical):
//   the padded row is cut in float4 chunks; chunk c belongs to lane (c mod 32); each lane runs TWO fused
//   multiply-add chains over its chunks in index order -- the even one takes elements x then z of every chunk, the
//   odd one y then w -- and adds them (even + odd) at the end; the 32 partial sums are then combined by an xor
//   butterfly with offsets 16, 8, 4, 2, 1.  Two chains per lane is the shape of Blackwell's packed fp32 pipe:
//   FADD2/FFMA2 (add/fma.rn.f32x2) advance both chains with one instruction, each half a plain IEEE rn operation,
//   so a float4 costs 4 issue slots instead of 8.  Rows are zero-padded to a multiple of 32 floats (128 B), which
//   adds exact zeros to the chains and therefore changes nothing.
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

// ---- packed fp32 pairs (sm_100: FADD2 / FFMA2) ------------------------------------------------------------------------
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
// even chain + odd chain
__device__ __forceinline__ float sum2(f32x2 v) {
  float lo, hi;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
  return __fadd_rn(lo, hi);
}
// one float4 of squared differences / products onto a chain pair
__device__ __forceinline__ f32x2 acc_l2(const float4 &r, const float4 &q, f32x2 acc) {
  f32x2 d0 = sub2(pk2(r.x, r.y), pk2(q.x, q.y));
  f32x2 d1 = sub2(pk2(r.z, r.w), pk2(q.z, q.w));
  acc = fma2(d0, d0, acc);
  return fma2(d1, d1, acc);
}
__device__ __forceinline__ f32x2 acc_dot(const float4 &a, const float4 &b, f32x2 acc) {
  acc = fma2(pk2(a.x, a.y), pk2(b.x, b.y), acc);
  return fma2(pk2(a.z, a.w), pk2(b.z, b.w), acc);
}

// read-only, streaming 16 B load (vectors are immutable once written; rows are rarely re-used by the same SM)
__device__ __forceinline__ float4 ldg_f4(const float4 *p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}

__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// the same load when `ok`; otherwise nothing is loaded and the result is undefined (the empty asm only tells the compiler
// the registers are written, so no zero fill is emitted)
__device__ __forceinline__ float4 ldg_f4_if(const float4 *p, bool ok) {
  float4 v;
  asm("" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w));
  if (ok) v = ldg_f4(p);
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
  f32x2 acc = 0ull;
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
    acc = acc_dot(Q.q[c], Q.q[c], acc);
  }
  Q.qq = butterfly(sum2(acc));
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
    f32x2 acc = 0ull;
#pragma unroll
    for (int c = 0; c < CPL; c++) acc = acc_l2(r[c], Q.q[c], acc);
    return __fsqrt_rn(butterfly(sum2(acc)));
  }
  f32x2 dot2 = 0ull, na2 = 0ull;
#pragma unroll
  for (int c = 0; c < CPL; c++) {
    dot2 = acc_dot(r[c], Q.q[c], dot2);
    na2 = acc_dot(r[c], r[c], na2);
  }
  const float dot = butterfly(sum2(dot2));
  const float na = butterfly(sum2(na2));
  // 1 - dot / sqrt(|a|^2 * |b|^2)
  return __fsub_rn(1.0f, __fdiv_rn(dot, __fsqrt_rn(__fmul_rn(na, Q.qq))));
}

// generic-dimension variant (any number of chunks): lanes stride over chunks, both operands in memory
__device__ __forceinline__ float warp_distance_generic(const float4 *a, const float4 *b, int nchunks, int metric,
                                                       int lane) {
  if (metric == 0) {
    f32x2 acc = 0ull;
    for (int chunk = lane; chunk < nchunks; chunk += 32) acc = acc_l2(a[chunk], b[chunk], acc);
    return __fsqrt_rn(butterfly(sum2(acc)));
  }
  f32x2 dot2 = 0ull, na2 = 0ull, nb2 = 0ull;
  for (int chunk = lane; chunk < nchunks; chunk += 32) {
    float4 x = a[chunk], y = b[chunk];
    dot2 = acc_dot(x, y, dot2);
    na2 = acc_dot(x, x, na2);
    nb2 = acc_dot(y, y, nb2);
  }
  const float dot = butterfly(sum2(dot2));
  const float na = butterfly(sum2(na2));
  const float nb = butterfly(sum2(nb2));
  return __fsub_rn(1.0f, __fdiv_rn(dot, __fsqrt_rn(__fmul_rn(na, nb))));
}

}  // namespace b200
