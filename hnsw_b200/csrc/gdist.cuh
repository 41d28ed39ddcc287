// gdist.cuh -- EuclideanDistance / CosineDistance (distance.go:15-23) evaluated by a GROUP of G lanes (G = 8, 16, 32),
// so one warp serves 32/G queries at once.  Bit-identical to the canonical order of dist.cuh / oracle/hnsw_oracle.c:
//   chunk c (float4) belongs to chain (c mod 32); a chain is one fused-multiply-add sequence in index order; the 32
//   chain sums are combined by an xor butterfly with offsets 16, 8, 4, 2, 1.
// With G lanes, lane g owns chunks g, g+G, g+2G, ... and therefore the 32/G chains g, g+G, ...; butterfly steps with
// offset >= G combine chains held by the same lane (plain adds), the remaining steps are shuffles inside the group.
// fadd is commutative, so which lane performs an addition does not change its bits.
//
// U rows are reduced together: at each shuffle step a lane sends half of its partial sums and keeps the other half
// ("transpose reduce"), which needs U-1+log2(G/U) shuffles instead of U*log2(G).  Row u ends up in lanes
// [u*G/U, (u+1)*G/U) of the group.
#pragma once
#include "dist.cuh"

namespace b200 {

template <int G>
struct GroupGeom {
  static constexpr int NCH = 32 / G;  // chains per lane
};

// combine the chains a lane holds (butterfly offsets 16 .. G)
template <int NCH>
__device__ __forceinline__ float inlane_combine(const float (&a)[NCH]) {
  if constexpr (NCH == 1) return a[0];
  else if constexpr (NCH == 2) return __fadd_rn(a[0], a[1]);                                    // offset 16
  else return __fadd_rn(__fadd_rn(a[0], a[2]), __fadd_rn(a[1], a[3]));                           // 16, then 8
}

// transpose reduce over the lanes of a group: v[0..CNT) partial sums -> v[0] = full sum of row (gl * U / G)
template <int OFF, int CNT>
struct XReduce {
  template <int U>
  static __device__ __forceinline__ void run(float (&v)[U], int gl) {
    if constexpr (OFF >= 1) {
      if constexpr (CNT > 1) {
        constexpr int H = CNT / 2;
        const bool bit = (gl & OFF) != 0;
#pragma unroll
        for (int i = 0; i < H; i++) {
          float keep = bit ? v[H + i] : v[i];
          float send = bit ? v[i] : v[H + i];
          float recv = __shfl_xor_sync(FULL, send, OFF);
          v[i] = __fadd_rn(keep, recv);
        }
        XReduce<OFF / 2, H>::template run<U>(v, gl);
      } else {
        v[0] = __fadd_rn(v[0], __shfl_xor_sync(FULL, v[0], OFF));
        XReduce<OFF / 2, 1>::template run<U>(v, gl);
      }
    }
  }
};

template <int G, int U>
__device__ __forceinline__ void group_reduce(float (&v)[U], int gl) {
  static_assert(U <= G && (U & (U - 1)) == 0, "U must be a power of two <= G");
  XReduce<G / 2, U>::template run<U>(v, gl);
}

// query held in registers by a group: lane g keeps chunks g, g+G, ... (CH of them)
template <int CH>
struct GQuery {
  float4 q[CH];
  float qq;  // |q|^2, canonical order (cosine)
};

// `query` points at [dim] floats (any alignment).  Executed by all 32 lanes; lanes with !on load nothing.
template <int G, int CH>
__device__ __forceinline__ void group_load_query(GQuery<CH> &Q, const float *query, bool on, int dim, int nchunks,
                                                 int gl) {
  constexpr int NCH = GroupGeom<G>::NCH;
  float acc[NCH];
#pragma unroll
  for (int i = 0; i < NCH; i++) acc[i] = 0.f;
#pragma unroll
  for (int t = 0; t < CH; t++) {
    const int chunk = gl + G * t;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (on && chunk < nchunks) {
#pragma unroll
      for (int e = 0; e < 4; e++) {
        int i = chunk * 4 + e;
        if (i < dim) v[e] = __ldg(query + i);
      }
    }
    Q.q[t] = make_float4(v[0], v[1], v[2], v[3]);
    float a = acc[t % NCH];
#pragma unroll
    for (int e = 0; e < 4; e++) a = __fmaf_rn(v[e], v[e], a);
    acc[t % NCH] = a;
  }
  float s[1] = {inlane_combine<NCH>(acc)};
  group_reduce<G, 1>(s, gl);
  Q.qq = s[0];
}

// partial sums of one row against the query (before the cross-lane reduction)
template <int G, int CH>
__device__ __forceinline__ float row_partial_l2(const float4 (&r)[CH], const GQuery<CH> &Q) {
  constexpr int NCH = GroupGeom<G>::NCH;
  float acc[NCH];
#pragma unroll
  for (int i = 0; i < NCH; i++) acc[i] = 0.f;
#pragma unroll
  for (int t = 0; t < CH; t++) {
    float dx = __fsub_rn(r[t].x, Q.q[t].x), dy = __fsub_rn(r[t].y, Q.q[t].y);
    float dz = __fsub_rn(r[t].z, Q.q[t].z), dw = __fsub_rn(r[t].w, Q.q[t].w);
    float a = acc[t % NCH];
    a = __fmaf_rn(dx, dx, a);
    a = __fmaf_rn(dy, dy, a);
    a = __fmaf_rn(dz, dz, a);
    a = __fmaf_rn(dw, dw, a);
    acc[t % NCH] = a;
  }
  return inlane_combine<NCH>(acc);
}

template <int G, int CH>
__device__ __forceinline__ void row_partial_cos(const float4 (&r)[CH], const GQuery<CH> &Q, float &dot, float &na) {
  constexpr int NCH = GroupGeom<G>::NCH;
  float ad[NCH], an[NCH];
#pragma unroll
  for (int i = 0; i < NCH; i++) ad[i] = an[i] = 0.f;
#pragma unroll
  for (int t = 0; t < CH; t++) {
    float d = ad[t % NCH], n = an[t % NCH];
    d = __fmaf_rn(r[t].x, Q.q[t].x, d);
    d = __fmaf_rn(r[t].y, Q.q[t].y, d);
    d = __fmaf_rn(r[t].z, Q.q[t].z, d);
    d = __fmaf_rn(r[t].w, Q.q[t].w, d);
    n = __fmaf_rn(r[t].x, r[t].x, n);
    n = __fmaf_rn(r[t].y, r[t].y, n);
    n = __fmaf_rn(r[t].z, r[t].z, n);
    n = __fmaf_rn(r[t].w, r[t].w, n);
    ad[t % NCH] = d;
    an[t % NCH] = n;
  }
  dot = inlane_combine<NCH>(ad);
  na = inlane_combine<NCH>(an);
}

// Distances of U rows (ids[u] < 0 = no row) to the group's query.  Returns, in every lane, the distance of row
// (gl * U / G); the caller fetches row u from lane u * (G / U) of its group.
template <int G, int CH, int U>
__device__ __forceinline__ float group_distances(const float4 *vec, int row_chunks, const int (&ids)[U],
                                                 const GQuery<CH> &Q, int metric, int gl) {
  float4 r[U][CH];
  const bool full = row_chunks == G * CH;  // every lane owns CH real chunks: no per-chunk bounds test
#pragma unroll
  for (int u = 0; u < U; u++) {
    // a missing row (ids[u] < 0) reads row 0 instead of branching; its distance is never used
    const float4 *row = vec + (size_t)(ids[u] < 0 ? 0 : ids[u]) * row_chunks + gl;
#pragma unroll
    for (int t = 0; t < CH; t++)
      r[u][t] = (full || gl + G * t < row_chunks) ? ldg_f4(row + G * t) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if (metric == 0) {
    float p[U];
#pragma unroll
    for (int u = 0; u < U; u++) p[u] = row_partial_l2<G, CH>(r[u], Q);
    group_reduce<G, U>(p, gl);
    return __fsqrt_rn(p[0]);
  }
  float pd[U], pn[U];
#pragma unroll
  for (int u = 0; u < U; u++) row_partial_cos<G, CH>(r[u], Q, pd[u], pn[u]);
  group_reduce<G, U>(pd, gl);
  group_reduce<G, U>(pn, gl);
  return __fsub_rn(1.0f, __fdiv_rn(pd[0], __fsqrt_rn(__fmul_rn(pn[0], Q.qq))));
}

}  // namespace b200
