/*
 * hnsw_oracle.c -- CPU restatement of TFMV/hnsw's hot path.  TEST INFRASTRUCTURE ONLY
 * (see hnsw_oracle.h for the parity status and who may load this).
 *
 * Citations are file:line into the reference tree (TFMV/hnsw).  Nothing here is shipped.
 * Build: see oracle/Makefile (gcc -O3 -march=native -ffp-contract=off).
 */
#define _GNU_SOURCE
#include "hnsw_oracle.h"

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
 * chunk c belongs to accumulator (c mod 32); each accumulator runs one fused-multiply-add chain
 * over its elements in index order; the 32 partial sums are combined by an xor butterfly with
 * offsets 16, 8, 4, 2, 1 (float addition is commutative, so every lane holds the same bits). */
static inline float butterfly32(float *acc) {
  for (int off = 16; off >= 1; off >>= 1) {
    float t[32];
    for (int i = 0; i < 32; i++) t[i] = acc[i] + acc[i ^ off];
    memcpy(acc, t, sizeof t);
  }
  return acc[0];
}

static float l2sq_canonical(const float *a, const float *b, int d) {
  float acc[32];
  for (int i = 0; i < 32; i++) acc[i] = 0.0f;
  for (int i = 0; i < d; i++) {
    float diff = a[i] - b[i];
    int lane = (i >> 2) & 31;
    acc[lane] = fmaf(diff, diff, acc[lane]);
  }
  return butterfly32(acc);
}

static void dot3_canonical(const float *a, const float *b, int d, float *dot, float *na, float *nb) {
  float accd[32], acca[32], accb[32];
  for (int i = 0; i < 32; i++) accd[i] = acca[i] = accb[i] = 0.0f;
  for (int i = 0; i < d; i++) {
    int lane = (i >> 2) & 31;
    accd[lane] = fmaf(a[i], b[i], accd[lane]);
    acca[lane] = fmaf(a[i], a[i], acca[lane]);
    accb[lane] = fmaf(b[i], b[i], accb[lane]);
  }
  *dot = butterfly32(accd);
  *na = butterfly32(acca);
  *nb = butterfly32(accb);
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

/* ORA_FLAVOR_SIMD: 16 independent accumulators so gcc vectorises it (AVX2/AVX-512 + FMA), the
 * shape of vek32's assembly path.  Used only for the timed CPU baseline. */
#define SIMD_W 16
static float l2sq_simd(const float *a, const float *b, int d) {
  float acc[SIMD_W] = {0};
  int i = 0;
  for (; i + SIMD_W <= d; i += SIMD_W)
    for (int j = 0; j < SIMD_W; j++) {
      float diff = a[i + j] - b[i + j];
      acc[j] = fmaf(diff, diff, acc[j]);
    }
  float s = 0.0f;
  for (; i < d; i++) {
    float diff = a[i] - b[i];
    s = fmaf(diff, diff, s);
  }
  for (int w = SIMD_W / 2; w >= 1; w >>= 1)
    for (int j = 0; j < w; j++) acc[j] += acc[j + w];
  return acc[0] + s;
}
static void dot3_simd(const float *a, const float *b, int d, float *dot, float *na, float *nb) {
  float ad[SIMD_W] = {0}, aa[SIMD_W] = {0}, ab[SIMD_W] = {0};
  int i = 0;
  for (; i + SIMD_W <= d; i += SIMD_W)
    for (int j = 0; j < SIMD_W; j++) {
      ad[j] = fmaf(a[i + j], b[i + j], ad[j]);
      aa[j] = fmaf(a[i + j], a[i + j], aa[j]);
      ab[j] = fmaf(b[i + j], b[i + j], ab[j]);
    }
  float sd = 0, sa = 0, sb = 0;
  for (; i < d; i++) {
    sd = fmaf(a[i], b[i], sd);
    sa = fmaf(a[i], a[i], sa);
    sb = fmaf(b[i], b[i], sb);
  }
  for (int w = SIMD_W / 2; w >= 1; w >>= 1)
    for (int j = 0; j < w; j++) {
      ad[j] += ad[j + w];
      aa[j] += aa[j + w];
      ab[j] += ab[j + w];
    }
  *dot = ad[0] + sd;
  *na = aa[0] + sa;
  *nb = ab[0] + sb;
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
