/*
 * ora_encode_exact.c -- CPU restatement of encode.go's v1 wire format (int keys) and of
 * hnsw-extensions/hybrid/exact.go's brute-force search.  TEST INFRASTRUCTURE ONLY (see hnsw_oracle.h).
 */
#define _GNU_SOURCE
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#include "hnsw_oracle.h"

/* ---- Go encoding/binary varints (used by encode.go:20-33, 67-71) ------------------------------ */
int ora_put_varint(uint8_t *buf, int64_t v) { /* binary.PutVarint: zig-zag then base-128 */
  uint64_t ux = (uint64_t)v << 1;
  if (v < 0) ux = ~ux;
  int i = 0;
  while (ux >= 0x80) {
    buf[i++] = (uint8_t)(ux | 0x80);
    ux >>= 7;
  }
  buf[i++] = (uint8_t)ux;
  return i;
}
int ora_get_varint(const uint8_t *buf, size_t len, int64_t *v) {
  uint64_t ux = 0;
  unsigned s = 0;
  for (size_t i = 0; i < len && i < 10; i++) {
    uint8_t b = buf[i];
    if (b < 0x80) {
      if (i == 9 && b > 1) return -2; /* overflow */
      ux |= (uint64_t)b << s;
      int64_t x = (int64_t)(ux >> 1);
      if (ux & 1) x = ~x;
      *v = x;
      return (int)i + 1;
    }
    ux |= (uint64_t)(b & 0x7f) << s;
    s += 7;
  }
  return -1; /* truncated */
}

/* ---- Export -- encode.go:133-174 --------------------------------------------------------------- */
typedef struct {
  uint8_t *buf;
  int64_t cap, len;
} wbuf;
static void w_bytes(wbuf *w, const void *p, size_t n) {
  if (w->buf && w->len + (int64_t)n <= w->cap) memcpy(w->buf + w->len, p, n);
  w->len += (int64_t)n;
}
static void w_varint(wbuf *w, int64_t v) {
  uint8_t t[10];
  int n = ora_put_varint(t, v);
  w_bytes(w, t, (size_t)n);
}

/* Returns the encoded length; writes only if it fits in cap (call with buf=NULL to size). */
int64_t ora_graph_export(const ora_graph *g, double ml, int ef, uint8_t *buf, int64_t cap) {
  wbuf w = {buf, cap, 0};
  int n_layers = ora_graph_layers(g);
  int64_t total = ora_graph_len(g);
  int M = ora_graph_M(g), metric = ora_graph_metric(g), dim = ora_graph_dim(g);
  /* header: version, M, Ml, EfSearch, distance name (encode.go:138-145) */
  w_varint(&w, 1);
  w_varint(&w, M);
  w_bytes(&w, &ml, 8); /* float64 little-endian via binary.Write (encode.go:98-104) */
  w_varint(&w, ef);
  const char *name = metric == ORA_EUCLIDEAN ? "euclidean" : "cosine";
  w_varint(&w, (int64_t)strlen(name));
  w_bytes(&w, name, strlen(name));
  w_varint(&w, n_layers); /* encode.go:149 */
  int32_t *nb = (int32_t *)malloc(sizeof(int32_t) * (size_t)(M + 1));
  /* rows sorted by key once: Go writes map order; ascending key is our deterministic stand-in */
  int32_t *order = (int32_t *)malloc(sizeof(int32_t) * (size_t)(total ? total : 1));
  for (int32_t r = 0; r < (int32_t)total; r++) order[r] = r;
  /* insertion sort is fine for fixtures; large exports use rows already in key order */
  for (int64_t i = 1; i < total; i++) {
    int32_t x = order[i];
    int64_t j = i;
    while (j > 0 && ora_graph_key(g, order[j - 1]) > ora_graph_key(g, x)) {
      order[j] = order[j - 1];
      j--;
    }
    order[j] = x;
  }
  for (int l = 0; l < n_layers; l++) {
    w_varint(&w, ora_graph_layer_size(g, l)); /* encode.go:154 */
    for (int64_t i = 0; i < total; i++) {
      int32_t r = order[i];
      if (ora_graph_level(g, r) < l) continue;
      int deg = ora_graph_neighbors(g, l, r, nb);
      w_varint(&w, ora_graph_key(g, r)); /* key: int -> varint */
      w_varint(&w, dim);                 /* []float32: len + LE data (encode.go:89-94) */
      w_bytes(&w, ora_graph_vector(g, r), sizeof(float) * (size_t)dim);
      w_varint(&w, deg);
      for (int t = 0; t < deg; t++) w_varint(&w, ora_graph_key(g, nb[t])); /* encode.go:165-170 */
    }
  }
  free(nb);
  free(order);
  return w.len;
}

/* ---- Import -- encode.go:180-262 ---------------------------------------------------------------- */
typedef struct {
  const uint8_t *p;
  int64_t len, off;
  int err;
} rbuf;
static int64_t r_varint(rbuf *r) {
  int64_t v = 0;
  int n = ora_get_varint(r->p + r->off, (size_t)(r->len - r->off), &v);
  if (n < 0) {
    r->err = 1;
    return 0;
  }
  r->off += n;
  return v;
}
static void r_bytes(rbuf *r, void *dst, size_t n) {
  if (r->off + (int64_t)n > r->len) {
    r->err = 1;
    memset(dst, 0, n);
    return;
  }
  memcpy(dst, r->p + r->off, n);
  r->off += (int64_t)n;
}

typedef struct {
  int64_t key;
  int32_t row;
} kv;
static int kv_cmp(const void *a, const void *b) {
  int64_t x = ((const kv *)a)->key, y = ((const kv *)b)->key;
  return x < y ? -1 : x > y;
}
static int32_t kv_find(const kv *t, int64_t n, int64_t key) {
  int64_t lo = 0, hi = n - 1;
  while (lo <= hi) {
    int64_t mid = (lo + hi) / 2;
    if (t[mid].key == key) return t[mid].row;
    if (t[mid].key < key) lo = mid + 1;
    else hi = mid - 1;
  }
  return -1;
}

/* Two passes over the stream: pass 1 finds every key's top layer, pass 2 builds rows (in layer-0 wire
 * order) and links.  Neighbour keys missing from their layer resolve to nil in Go (encode.go:253-257)
 * and are skipped by search (graph.go:141); they are dropped here. */
ora_graph *ora_graph_import(const uint8_t *buf, int64_t len, int flavor, double *ml, int *ef) {
  rbuf r = {buf, len, 0, 0};
  int64_t version = r_varint(&r);
  int64_t M = r_varint(&r);
  double mlv;
  r_bytes(&r, &mlv, 8);
  int64_t efv = r_varint(&r);
  int64_t sl = r_varint(&r);
  char name[64] = {0};
  if (sl < 0 || sl >= 64) return NULL;
  r_bytes(&r, name, (size_t)sl);
  if (r.err || version != 1) return NULL; /* encode.go:202-204 */
  int metric;
  if (!strcmp(name, "euclidean")) metric = ORA_EUCLIDEAN;
  else if (!strcmp(name, "cosine")) metric = ORA_COSINE;
  else return NULL; /* encode.go:194-197 */
  int64_t n_layers = r_varint(&r);
  if (r.err || n_layers < 0 || n_layers > ORA_MAX_LAYERS) return NULL;
  int64_t body = r.off;
  /* pass 1 */
  int64_t n0 = 0;
  kv *tab = NULL;
  int32_t *lvl = NULL;
  int dim = 0;
  for (int l = 0; l < n_layers; l++) {
    int64_t nn = r_varint(&r);
    if (l == 0) {
      n0 = nn;
      tab = (kv *)malloc(sizeof(kv) * (size_t)(nn ? nn : 1));
      lvl = (int32_t *)calloc((size_t)(nn ? nn : 1), sizeof(int32_t));
    }
    for (int64_t j = 0; j < nn && !r.err; j++) {
      int64_t key = r_varint(&r);
      int64_t d = r_varint(&r);
      if (l == 0 && j == 0) dim = (int)d;
      r.off += d * 4;
      int64_t nnb = r_varint(&r);
      for (int64_t t = 0; t < nnb; t++) r_varint(&r);
      if (l == 0) {
        tab[j].key = key;
        tab[j].row = (int32_t)j;
      } else {
        int32_t row = kv_find(tab, n0, key);
        if (row >= 0 && lvl[row] < l) lvl[row] = l;
      }
    }
    if (l == 0) qsort(tab, (size_t)n0, sizeof(kv), kv_cmp);
  }
  if (r.err) {
    free(tab);
    free(lvl);
    return NULL;
  }
  /* pass 2 */
  ora_graph *g = ora_graph_new(dim, (int)M, metric, flavor);
  r.off = body;
  float *vec = (float *)malloc(sizeof(float) * (size_t)(dim ? dim : 1));
  for (int l = 0; l < n_layers; l++) {
    int64_t nn = r_varint(&r);
    for (int64_t j = 0; j < nn; j++) {
      int64_t key = r_varint(&r);
      int64_t d = r_varint(&r);
      if (d != dim) r.err = 1;
      if (r.err) break;
      r_bytes(&r, vec, sizeof(float) * (size_t)d);
      int32_t row;
      if (l == 0) row = ora_graph_put_node(g, key, vec, lvl[j]);
      else row = kv_find(tab, n0, key);
      int64_t nnb = r_varint(&r);
      for (int64_t t = 0; t < nnb; t++) {
        int64_t nk = r_varint(&r);
        if (l == 0) continue; /* layer-0 links need every row to exist: third mini-pass below */
        int32_t to = kv_find(tab, n0, nk);
        if (row >= 0 && to >= 0 && ora_graph_level(g, to) >= l) ora_graph_link(g, l, row, to);
      }
    }
  }
  /* layer-0 links */
  r.off = body;
  {
    int64_t nn = r_varint(&r);
    for (int64_t j = 0; j < nn && !r.err; j++) {
      r_varint(&r);
      int64_t d = r_varint(&r);
      r.off += d * 4;
      int64_t nnb = r_varint(&r);
      for (int64_t t = 0; t < nnb; t++) {
        int64_t nk = r_varint(&r);
        int32_t to = kv_find(tab, n0, nk);
        if (to >= 0) ora_graph_link(g, 0, (int32_t)j, to);
      }
    }
  }
  free(vec);
  free(tab);
  free(lvl);
  if (r.err) {
    ora_graph_free(g);
    return NULL;
  }
  if (ml) *ml = mlv;
  if (ef) *ef = (int)efv;
  return g;
}

/* ---- ExactIndex.Search -- hybrid/exact.go:61-111 ------------------------------------------------ */
typedef struct {
  float d;
  int32_t i;
} item;

static void merge_sort_items(item *a, item *tmp, int64_t n) { /* stable, strict '>' swaps only */
  if (n < 2) return;
  int64_t h = n / 2;
  merge_sort_items(a, tmp, h);
  merge_sort_items(a + h, tmp, n - h);
  int64_t i = 0, j = h, o = 0;
  while (i < h && j < n) {
    if (a[i].d > a[j].d) tmp[o++] = a[j++]; /* only a strictly larger left element yields */
    else tmp[o++] = a[i++];
  }
  while (i < h) tmp[o++] = a[i++];
  while (j < n) tmp[o++] = a[j++];
  memcpy(a, tmp, sizeof(item) * (size_t)n);
}

int ora_exact_search(const float *vecs, int64_t n, int dim, int metric, int flavor, const float *query, int k,
                     int faithful, int32_t *out_idx, float *out_dists) {
  if (n == 0) return 0; /* exact.go:65-67 */
  item *items = (item *)malloc(sizeof(item) * (size_t)n);
  for (int64_t i = 0; i < n; i++) { /* exact.go:76-84; map order -> ascending index */
    items[i].d = ora_distance(metric, flavor, query, vecs + (size_t)i * dim, dim);
    items[i].i = (int32_t)i;
  }
  if (faithful) { /* exact.go:88-94 */
    for (int64_t i = 1; i < n; i++) {
      int64_t j = i;
      while (j > 0 && items[j - 1].d > items[j].d) {
        item t = items[j - 1];
        items[j - 1] = items[j];
        items[j] = t;
        j--;
      }
    }
  } else {
    /* NaN-free inputs: insertion sort with strict '>' == any stable ascending sort */
    item *tmp = (item *)malloc(sizeof(item) * (size_t)n);
    merge_sort_items(items, tmp, n);
    free(tmp);
  }
  int cnt = k < n ? k : (int)n; /* exact.go:97-100 */
  for (int i = 0; i < cnt; i++) {
    out_idx[i] = items[i].i;
    out_dists[i] = items[i].d;
  }
  free(items);
  return cnt;
}

typedef struct {
  const float *vecs, *queries;
  int64_t n, q0, q1;
  int dim, metric, flavor, k, faithful;
  int32_t *out_idx;
  float *out_dists;
} ex_job;
static void *ex_worker(void *arg) {
  ex_job *j = (ex_job *)arg;
  for (int64_t q = j->q0; q < j->q1; q++)
    ora_exact_search(j->vecs, j->n, j->dim, j->metric, j->flavor, j->queries + (size_t)q * j->dim, j->k,
                     j->faithful, j->out_idx + (size_t)q * j->k, j->out_dists + (size_t)q * j->k);
  return NULL;
}
void ora_exact_batch_search(const float *vecs, int64_t n, int dim, int metric, int flavor, const float *queries,
                            int64_t nq, int k, int faithful, int n_threads, int32_t *out_idx, float *out_dists) {
  if (n_threads < 1) n_threads = 1;
  if (n_threads > nq) n_threads = nq > 0 ? (int)nq : 1;
  ex_job *jobs = (ex_job *)calloc((size_t)n_threads, sizeof *jobs);
  pthread_t *th = (pthread_t *)calloc((size_t)n_threads, sizeof *th);
  for (int t = 0; t < n_threads; t++) {
    ex_job j = {vecs, queries, n, nq * t / n_threads, nq * (t + 1) / n_threads, dim, metric, flavor, k, faithful,
                out_idx, out_dists};
    jobs[t] = j;
    if (n_threads == 1) ex_worker(&jobs[t]);
    else pthread_create(&th[t], NULL, ex_worker, &jobs[t]);
  }
  if (n_threads > 1)
    for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
  free(jobs);
  free(th);
}
