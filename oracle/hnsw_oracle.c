/*
 * hnsw_oracle.c -- CPU restatement of TFMV/hnsw's hot path.  TEST INFRASTRUCTURE ONLY
 * (see hnsw_oracle.h for the parity status and who may load this).
 *
 * Citations are file:line into the reference tree (TFMV/hnsw).  Nothing here is shipped.
 * Build: see oracle/Makefile (gcc -O3 -march=x86-64-v3 -ffp-contract=off).
 */
#define _GNU_SOURCE
#include "hnsw_oracle.h"

#include <immintrin.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ============================================================================================
 * Distances -- distance.go:15-23.  The arithmetic lives in github.com/viterin/vek v0.4.2
 * (vek32.Distance, vek32.CosineSimilarity), absent from the reference tree.  Published
 * algorithm: Distance = sqrt(sum((a_i-b_i)^2)); CosineSimilarity = dot / sqrt(|a|^2 * |b|^2).
 * ============================================================================================ */

/* ORA_FLAVOR_CANONICAL: the order the device kernels use.  The vector is cut in float4 chunks;
 * chunk c belongs to accumulator pair (c mod 32): an EVEN fused-multiply-add chain over elements
 * x then z of its chunks in index order, an ODD chain over y then w, added (even + odd) at the
 * end; the 32 sums are combined by an xor butterfly with offsets 16, 8, 4, 2, 1 (float addition
 * is commutative, so every lane holds the same bits).  Element i therefore goes to chain pair
 * (i >> 2) & 31, half i & 1. */
static inline float butterfly32(float *acc) {
  /* lane 0 of the butterfly only ever reads lanes < off, so a halving reduction gives the same bits */
  for (int off = 16; off >= 1; off >>= 1)
    for (int i = 0; i < off; i++) acc[i] = acc[i] + acc[i + off];
  return acc[0];
}

/* AVX2 form of the same order.  A block of 128 floats touches each of the 32 accumulators once, four
 * elements in a row.  For 8 consecutive chunks held in 4 registers r0..r3 (r_i = [chunk 2i | chunk 2i+1]) a
 * 4x4 transpose inside each 128-bit half gives X = [c0.x c2.x c4.x c6.x | c1.x c3.x c5.x c7.x] (likewise Y, Z,
 * W), so one register of 8 even accumulators advances by fma(X), fma(Z) and one of 8 odd accumulators by
 * fma(Y), fma(W).  Bit-identical to the scalar definition (tests/test_oracle_kat.py). */
static inline void transpose8(__m256 r0, __m256 r1, __m256 r2, __m256 r3, __m256 *x, __m256 *y, __m256 *z,
                              __m256 *w) {
  __m256 t0 = _mm256_unpacklo_ps(r0, r1), t1 = _mm256_unpackhi_ps(r0, r1);
  __m256 t2 = _mm256_unpacklo_ps(r2, r3), t3 = _mm256_unpackhi_ps(r2, r3);
  *x = _mm256_shuffle_ps(t0, t2, 0x44);
  *y = _mm256_shuffle_ps(t0, t2, 0xEE);
  *z = _mm256_shuffle_ps(t1, t3, 0x44);
  *w = _mm256_shuffle_ps(t1, t3, 0xEE);
}
/* register lane j of group g holds accumulator 8g + perm[j] */
static const int kLanePerm[8] = {0, 2, 4, 6, 1, 3, 5, 7};
static inline void unpermute(const __m256 v[4], float *acc) {
  float t[8] __attribute__((aligned(32)));
  for (int g = 0; g < 4; g++) {
    _mm256_store_ps(t, v[g]);
    for (int j = 0; j < 8; j++) acc[8 * g + kLanePerm[j]] = t[j];
  }
}
/* even + odd per chain, then the butterfly */
static inline float finish64(float *even, float *odd) {
  for (int i = 0; i < 32; i++) even[i] = even[i] + odd[i];
  return butterfly32(even);
}

static float l2sq_canonical(const float *restrict a, const float *restrict b, int d) {
  float acc[2][32] __attribute__((aligned(64)));
  __m256 ve[4], vo[4];
  for (int g = 0; g < 4; g++) ve[g] = vo[g] = _mm256_setzero_ps();
  int i = 0;
  for (; i + 128 <= d; i += 128) {
    for (int g = 0; g < 4; g++) {
      const float *pa = a + i + 32 * g, *pb = b + i + 32 * g;
      __m256 r0 = _mm256_sub_ps(_mm256_loadu_ps(pa), _mm256_loadu_ps(pb));
      __m256 r1 = _mm256_sub_ps(_mm256_loadu_ps(pa + 8), _mm256_loadu_ps(pb + 8));
      __m256 r2 = _mm256_sub_ps(_mm256_loadu_ps(pa + 16), _mm256_loadu_ps(pb + 16));
      __m256 r3 = _mm256_sub_ps(_mm256_loadu_ps(pa + 24), _mm256_loadu_ps(pb + 24));
      __m256 x, y, z, w;
      transpose8(r0, r1, r2, r3, &x, &y, &z, &w);
      ve[g] = _mm256_fmadd_ps(z, z, _mm256_fmadd_ps(x, x, ve[g]));
      vo[g] = _mm256_fmadd_ps(w, w, _mm256_fmadd_ps(y, y, vo[g]));
    }
  }
  unpermute(ve, acc[0]);
  unpermute(vo, acc[1]);
  for (; i < d; i++) {
    float diff = a[i] - b[i];
    int lane = (i >> 2) & 31;
    acc[i & 1][lane] = fmaf(diff, diff, acc[i & 1][lane]);
  }
  return finish64(acc[0], acc[1]);
}

static void dot3_canonical(const float *restrict a, const float *restrict b, int d, float *dot, float *na,
                           float *nb) {
  float accd[2][32] __attribute__((aligned(64))), acca[2][32] __attribute__((aligned(64))),
      accb[2][32] __attribute__((aligned(64)));
  __m256 vd[2][4], va[2][4], vb[2][4];
  for (int h = 0; h < 2; h++)
    for (int g = 0; g < 4; g++) vd[h][g] = va[h][g] = vb[h][g] = _mm256_setzero_ps();
  int i = 0;
  for (; i + 128 <= d; i += 128) {
    for (int g = 0; g < 4; g++) {
      const float *pa = a + i + 32 * g, *pb = b + i + 32 * g;
      __m256 ax, ay, az, aw, bx, by, bz, bw;
      transpose8(_mm256_loadu_ps(pa), _mm256_loadu_ps(pa + 8), _mm256_loadu_ps(pa + 16), _mm256_loadu_ps(pa + 24),
                 &ax, &ay, &az, &aw);
      transpose8(_mm256_loadu_ps(pb), _mm256_loadu_ps(pb + 8), _mm256_loadu_ps(pb + 16), _mm256_loadu_ps(pb + 24),
                 &bx, &by, &bz, &bw);
      vd[0][g] = _mm256_fmadd_ps(az, bz, _mm256_fmadd_ps(ax, bx, vd[0][g]));
      vd[1][g] = _mm256_fmadd_ps(aw, bw, _mm256_fmadd_ps(ay, by, vd[1][g]));
      va[0][g] = _mm256_fmadd_ps(az, az, _mm256_fmadd_ps(ax, ax, va[0][g]));
      va[1][g] = _mm256_fmadd_ps(aw, aw, _mm256_fmadd_ps(ay, ay, va[1][g]));
      vb[0][g] = _mm256_fmadd_ps(bz, bz, _mm256_fmadd_ps(bx, bx, vb[0][g]));
      vb[1][g] = _mm256_fmadd_ps(bw, bw, _mm256_fmadd_ps(by, by, vb[1][g]));
    }
  }
  for (int h = 0; h < 2; h++) {
    unpermute(vd[h], accd[h]);
    unpermute(va[h], acca[h]);
    unpermute(vb[h], accb[h]);
  }
  for (; i < d; i++) {
    int lane = (i >> 2) & 31, h = i & 1;
    accd[h][lane] = fmaf(a[i], b[i], accd[h][lane]);
    acca[h][lane] = fmaf(a[i], a[i], acca[h][lane]);
    accb[h][lane] = fmaf(b[i], b[i], accb[h][lane]);
  }
  *dot = finish64(accd[0], accd[1]);
  *na = finish64(acca[0], acca[1]);
  *nb = finish64(accb[0], accb[1]);
}

/* ORA_FLAVOR_SEQUENTIAL: vek's portable Go loops (one accumulator, index order, no fusing). */
static float l2sq_sequential(const float *a, const float *b, int d) {
  volatile float sum = 0.0f; /* volatile: forbid reassociation/vectorisation */
  for (int i = 0; i < d; i++) {
    float diff = a[i] - b[i];
    float sq = diff * diff;
    sum = sum + sq;
  }
  return sum;
}
static void dot3_sequential(const float *a, const float *b, int d, float *dot, float *na, float *nb) {
  volatile float sd = 0.0f, sa = 0.0f, sb = 0.0f;
  for (int i = 0; i < d; i++) {
    float p = a[i] * b[i], q = a[i] * a[i], r = b[i] * b[i];
    sd = sd + p;
    sa = sa + q;
    sb = sb + r;
  }
  *dot = sd;
  *na = sa;
  *nb = sb;
}

/* ORA_FLAVOR_SIMD: four 8-wide FMA accumulators, the shape of vek32's AVX2 assembly path.
 * Used only for the timed CPU baseline. */
static inline float hsum8(__m256 v) {
  __m128 lo = _mm256_castps256_ps128(v), hi = _mm256_extractf128_ps(v, 1);
  lo = _mm_add_ps(lo, hi);
  lo = _mm_add_ps(lo, _mm_movehl_ps(lo, lo));
  lo = _mm_add_ss(lo, _mm_shuffle_ps(lo, lo, 1));
  return _mm_cvtss_f32(lo);
}
static float l2sq_simd(const float *a, const float *b, int d) {
  __m256 s0 = _mm256_setzero_ps(), s1 = s0, s2 = s0, s3 = s0;
  int i = 0;
  for (; i + 32 <= d; i += 32) {
    __m256 d0 = _mm256_sub_ps(_mm256_loadu_ps(a + i), _mm256_loadu_ps(b + i));
    __m256 d1 = _mm256_sub_ps(_mm256_loadu_ps(a + i + 8), _mm256_loadu_ps(b + i + 8));
    __m256 d2 = _mm256_sub_ps(_mm256_loadu_ps(a + i + 16), _mm256_loadu_ps(b + i + 16));
    __m256 d3 = _mm256_sub_ps(_mm256_loadu_ps(a + i + 24), _mm256_loadu_ps(b + i + 24));
    s0 = _mm256_fmadd_ps(d0, d0, s0);
    s1 = _mm256_fmadd_ps(d1, d1, s1);
    s2 = _mm256_fmadd_ps(d2, d2, s2);
    s3 = _mm256_fmadd_ps(d3, d3, s3);
  }
  for (; i + 8 <= d; i += 8) {
    __m256 d0 = _mm256_sub_ps(_mm256_loadu_ps(a + i), _mm256_loadu_ps(b + i));
    s0 = _mm256_fmadd_ps(d0, d0, s0);
  }
  float s = hsum8(_mm256_add_ps(_mm256_add_ps(s0, s1), _mm256_add_ps(s2, s3)));
  for (; i < d; i++) {
    float diff = a[i] - b[i];
    s = fmaf(diff, diff, s);
  }
  return s;
}
static void dot3_simd(const float *a, const float *b, int d, float *dot, float *na, float *nb) {
  __m256 sd0 = _mm256_setzero_ps(), sd1 = sd0, sa0 = sd0, sa1 = sd0, sb0 = sd0, sb1 = sd0;
  int i = 0;
  for (; i + 16 <= d; i += 16) {
    __m256 a0 = _mm256_loadu_ps(a + i), a1 = _mm256_loadu_ps(a + i + 8);
    __m256 b0 = _mm256_loadu_ps(b + i), b1 = _mm256_loadu_ps(b + i + 8);
    sd0 = _mm256_fmadd_ps(a0, b0, sd0);
    sd1 = _mm256_fmadd_ps(a1, b1, sd1);
    sa0 = _mm256_fmadd_ps(a0, a0, sa0);
    sa1 = _mm256_fmadd_ps(a1, a1, sa1);
    sb0 = _mm256_fmadd_ps(b0, b0, sb0);
    sb1 = _mm256_fmadd_ps(b1, b1, sb1);
  }
  float sd = hsum8(_mm256_add_ps(sd0, sd1)), sa = hsum8(_mm256_add_ps(sa0, sa1)), sb = hsum8(_mm256_add_ps(sb0, sb1));
  for (; i < d; i++) {
    sd = fmaf(a[i], b[i], sd);
    sa = fmaf(a[i], a[i], sa);
    sb = fmaf(b[i], b[i], sb);
  }
  *dot = sd;
  *na = sa;
  *nb = sb;
}

float ora_distance(int metric, int flavor, const float *a, const float *b, int d) {
  if (metric == ORA_EUCLIDEAN) { /* distance.go:20-23 */
    float s = flavor == ORA_FLAVOR_CANONICAL    ? l2sq_canonical(a, b, d)
              : flavor == ORA_FLAVOR_SEQUENTIAL ? l2sq_sequential(a, b, d)
                                                : l2sq_simd(a, b, d);
    return sqrtf(s);
  }
  /* distance.go:15-17: 1 - cos; a zero vector gives 0/0 = NaN (noted at graph.go:65-66) */
  float dot, na, nb;
  if (flavor == ORA_FLAVOR_CANONICAL) dot3_canonical(a, b, d, &dot, &na, &nb);
  else if (flavor == ORA_FLAVOR_SEQUENTIAL) dot3_sequential(a, b, d, &dot, &na, &nb);
  else dot3_simd(a, b, d, &dot, &na, &nb);
  return 1.0f - dot / sqrtf(na * nb);
}

/* ============================================================================================
 * Heap -- heap/heap.go:42-95 over Go's container/heap (up/down/Push/Pop/Remove).
 * Min-heap on dist with strict '<' (graph.go:88-90).  Max() is the LAST ARRAY SLOT
 * (heap.go:89-91) and PopLast() is Remove(len-1) (heap.go:73-75), i.e. a truncation.
 * ============================================================================================ */
void ora_heap_init(ora_heap *h, int cap) {
  if (cap < 4) cap = 4;
  h->dist = (float *)malloc(sizeof(float) * (size_t)cap);
  h->id = (int32_t *)malloc(sizeof(int32_t) * (size_t)cap);
  h->len = 0;
  h->cap = cap;
}
void ora_heap_free(ora_heap *h) {
  free(h->dist);
  free(h->id);
  h->dist = NULL;
  h->id = NULL;
  h->len = h->cap = 0;
}
static inline void heap_swap(ora_heap *h, int i, int j) {
  float d = h->dist[i];
  h->dist[i] = h->dist[j];
  h->dist[j] = d;
  int32_t t = h->id[i];
  h->id[i] = h->id[j];
  h->id[j] = t;
}
/* container/heap.up */
static void heap_up(ora_heap *h, int j) {
  for (;;) {
    int i = (j - 1) / 2; /* parent; Go's integer division truncates toward zero: j=0 -> i=0 */
    if (i == j || !(h->dist[j] < h->dist[i])) break;
    heap_swap(h, i, j);
    j = i;
  }
}
/* container/heap.down; returns whether the element moved */
static int heap_down(ora_heap *h, int i0, int n) {
  int i = i0;
  for (;;) {
    int j1 = 2 * i + 1;
    if (j1 >= n || j1 < 0) break;
    int j = j1;
    int j2 = j1 + 1;
    if (j2 < n && h->dist[j2] < h->dist[j1]) j = j2;
    if (!(h->dist[j] < h->dist[i])) break;
    heap_swap(h, i, j);
    i = j;
  }
  return i > i0;
}
void ora_heap_push(ora_heap *h, float d, int32_t id) {
  if (h->len == h->cap) {
    h->cap *= 2;
    h->dist = (float *)realloc(h->dist, sizeof(float) * (size_t)h->cap);
    h->id = (int32_t *)realloc(h->id, sizeof(int32_t) * (size_t)h->cap);
  }
  h->dist[h->len] = d;
  h->id[h->len] = id;
  h->len++;
  heap_up(h, h->len - 1);
}
void ora_heap_pop(ora_heap *h, float *d, int32_t *id) {
  int n = h->len - 1;
  heap_swap(h, 0, n);
  heap_down(h, 0, n);
  if (d) *d = h->dist[n];
  if (id) *id = h->id[n];
  h->len = n;
}
void ora_heap_remove(ora_heap *h, int i, float *d, int32_t *id) {
  int n = h->len - 1;
  if (n != i) {
    heap_swap(h, i, n);
    if (!heap_down(h, i, n)) heap_up(h, i);
  }
  if (d) *d = h->dist[n];
  if (id) *id = h->id[n];
  h->len = n;
}
void ora_heap_poplast(ora_heap *h, float *d, int32_t *id) { ora_heap_remove(h, h->len - 1, d, id); }

/* ============================================================================================
 * Levels -- graph.go:370-417
 * ============================================================================================ */
int ora_max_level(double ml, int64_t num_nodes) {
  if (ml == 0) return -1;
  if (num_nodes == 0) return 1;
  double l = log((double)num_nodes);
  l /= log(1 / ml);
  return (int)round(l) + 1; /* Go math.Round: half away from zero, same as C round() */
}

int ora_random_level(double ml, int n_layers, int64_t layer0_size, const double *uniforms, int *used) {
  int max = 1;
  if (n_layers > 0) max = ora_max_level(ml, layer0_size);
  int u = 0;
  for (int level = 0; level < max; level++) {
    double r = uniforms[u++];
    if (r > ml) {
      if (used) *used = u;
      return level;
    }
  }
  if (used) *used = u;
  return max;
}
