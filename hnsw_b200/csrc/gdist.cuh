// gdist.cuh -- EuclideanDistance / CosineDistance (distance.go:15-23) evaluated by a GROUP of G lanes (G = 8, 16, 32),
// so one warp serves 32/G queries at once.  Bit-identical to the canonical order of dist.cuh / oracle/hnsw_oracle.c:
//   chunk c (float4) belongs to chain pair (c mod 32): an even chain over x, z and an odd chain over y, w of its chunks
//   in index order, added at the end (FADD2/FFMA2 advance both at once); the 32 sums are combined by an xor butterfly
//   with offsets 16, 8, 4, 2, 1.
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
__device__ __forceinline__ float inlane_combine(const f32x2 (&p)[NCH]) {
  float a[NCH];
#pragma unroll
  for (int i = 0; i < NCH; i++) a[i] = sum2(p[i]);  // even + odd chain
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
  f32x2 acc[NCH];
#pragma unroll
  for (int i = 0; i < NCH; i++) acc[i] = 0ull;
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
    acc[t % NCH] = acc_dot(Q.q[t], Q.q[t], acc[t % NCH]);
  }
  float s[1] = {inlane_combine<NCH>(acc)};
  group_reduce<G, 1>(s, gl);
  Q.qq = s[0];
}

// partial sums of one row against the query (before the cross-lane reduction)
template <int G, int CH>
__device__ __forceinline__ float row_partial_l2(const float4 (&r)[CH], const GQuery<CH> &Q) {
  constexpr int NCH = GroupGeom<G>::NCH;
  f32x2 acc[NCH];
#pragma unroll
  for (int i = 0; i < NCH; i++) acc[i] = 0ull;
#pragma unroll
  for (int t = 0; t < CH; t++) acc[t % NCH] = acc_l2(r[t], Q.q[t], acc[t % NCH]);
  return inlane_combine<NCH>(acc);
}

template <int G, int CH>
__device__ __forceinline__ void row_partial_cos(const float4 (&r)[CH], const GQuery<CH> &Q, float &dot, float &na) {
  constexpr int NCH = GroupGeom<G>::NCH;
  f32x2 ad[NCH], an[NCH];
#pragma unroll
  for (int i = 0; i < NCH; i++) ad[i] = an[i] = 0ull;
#pragma unroll
  for (int t = 0; t < CH; t++) {
    ad[t % NCH] = acc_dot(r[t], Q.q[t], ad[t % NCH]);
    an[t % NCH] = acc_dot(r[t], r[t], an[t % NCH]);
  }
  dot = inlane_combine<NCH>(ad);
  na = inlane_combine<NCH>(an);
}

// ---- the three phases of evaluating U rows (ids[u] < 0 = no row), split so a caller can issue the loads of the next
//      batch between the multiply-adds of this one and its reduction (software pipelining, traverse.cuh) -----------

// phase 1: issue the 16 B loads.  A missing row loads nothing and leaves its registers undefined; its (garbage)
// distance is never used.  G == 8: rows are padded to 128 B (row_chunks % 8 == 0) and the launcher instantiates
// CH == row_chunks / 8 exactly, so every lane owns CH real chunks and no bounds test is needed.
template <int G, int CH, int U>
__device__ __forceinline__ void group_load_rows(float4 (&r)[U][CH], const float4 *vec, int row_chunks,
                                                const int (&ids)[U], int gl) {
#pragma unroll
  for (int u = 0; u < U; u++) {
    const float4 *row = vec + (size_t)ids[u] * row_chunks + gl;
    const bool ok = ids[u] >= 0;
#pragma unroll
    for (int t = 0; t < CH; t++) {
      if constexpr (G == 8) {
        r[u][t] = ldg_f4_if(row + G * t, ok);
      } else {
        r[u][t] = gl + G * t < row_chunks ? ldg_f4_if(row + G * t, ok) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
}

template <int U>
struct GPartial {
  float a[U];  // L2: sum of squared differences; cosine: dot
  float b[U];  // cosine: |row|^2
};

// phase 2: the lane's share of the sums (consumes the row registers)
template <int G, int CH, int U>
__device__ __forceinline__ void group_partials(GPartial<U> &P, const float4 (&r)[U][CH], const GQuery<CH> &Q,
                                               int metric) {
  if (metric == 0) {
#pragma unroll
    for (int u = 0; u < U; u++) P.a[u] = row_partial_l2<G, CH>(r[u], Q);
  } else {
#pragma unroll
    for (int u = 0; u < U; u++) row_partial_cos<G, CH>(r[u], Q, P.a[u], P.b[u]);
  }
}

// phase 3: cross-lane reduction and the final formula.  Returns, in every lane, the distance of row (gl * U / G);
// the caller fetches row u from lane u * (G / U) of its group.
template <int G, int CH, int U>
__device__ __forceinline__ float group_finish(GPartial<U> &P, const GQuery<CH> &Q, int metric, int gl) {
  if (metric == 0) {
    group_reduce<G, U>(P.a, gl);
    return __fsqrt_rn(P.a[0]);
  }
  group_reduce<G, U>(P.a, gl);
  group_reduce<G, U>(P.b, gl);
  return __fsub_rn(1.0f, __fdiv_rn(P.a[0], __fsqrt_rn(__fmul_rn(P.b[0], Q.qq))));
}

// all three at once
template <int G, int CH, int U>
__device__ __forceinline__ float group_distances(const float4 *vec, int row_chunks, const int (&ids)[U],
                                                 const GQuery<CH> &Q, int metric, int gl) {
  float4 r[U][CH];
  group_load_rows<G, CH, U>(r, vec, row_chunks, ids, gl);
  GPartial<U> P;
  group_partials<G, CH, U>(P, r, Q, metric);
  return group_finish<G, CH, U>(P, Q, metric, gl);
}

}  // namespace b200
