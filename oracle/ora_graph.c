/*
 * ora_graph.c -- CPU restatement of graph.go's layer search and insertion.  TEST INFRASTRUCTURE ONLY
 * (see hnsw_oracle.h).  Citations are file:line into the reference tree.
 *
 * Storage differs from the reference on purpose (rows instead of pointers/maps); behaviour does not:
 *   - a node is a row id (int32, insertion order); key[row] is the reference's K (int64 here);
 *   - a node inserted with level L is a member of layers 0..L (graph.go:510-515);
 *   - layer-0 neighbour lists live in adj0[row*(M+1)..], upper layers in adjU[(upper_off[row]+l-1)*(M+1)..];
 *   - every neighbour list is kept sorted by key, which is both the order the reference sorts into before
 *     expanding (graph.go:137-138) and the documented stand-in for Go's random map iteration order.
 */
#define _GNU_SOURCE
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#include "hnsw_oracle.h"

struct ora_graph {
  int dim, M, metric, flavor, stride; /* stride = M+1 slots per list */
  int64_t n, cap;
  float *vec;
  int64_t *key;
  int32_t *level;
  uint8_t *dead; /* 1 = removed from layer.nodes by Delete (graph.go:843-866); the row and its list stay */
  int64_t n_dead;
  int build_mode; /* ORA_MODE_* of the layer search used by Add (reference: ORA_MODE_REFERENCE) */
  int32_t *adj0;
  int32_t *deg0;
  int32_t *upper_off;
  int32_t *adjU;
  int32_t *degU;
  int64_t nU, capU;
  int n_layers;
  int32_t entry[ORA_MAX_LAYERS];
  int64_t layer_size[ORA_MAX_LAYERS];
  /* visited stamps for the single-threaded paths */
  uint32_t *stamp;
  uint32_t epoch;
  ora_counters ctr;
};

typedef struct {
  uint32_t *stamp;
  uint32_t epoch;
  int64_t cap;
} visited_t;

/* Evaluation trace of a search (rows in evaluation order with the distance this flavour computed): what the flavour
 * divergence report compares.  Off unless a caller installs a buffer for its thread. */
typedef struct {
  int32_t *rows;
  float *dists;
  int64_t cap, n;
} ora_trace;
static __thread ora_trace *tls_trace;
#define TRACE_EVAL(r, d)                                     \
  do {                                                       \
    if (tls_trace) {                                         \
      if (tls_trace->n < tls_trace->cap) {                   \
        tls_trace->rows[tls_trace->n] = (r);                 \
        tls_trace->dists[tls_trace->n] = (d);                \
      }                                                      \
      tls_trace->n++;                                        \
    }                                                        \
  } while (0)

ora_graph *ora_graph_new(int dim, int M, int metric, int flavor) {
  ora_graph *g = (ora_graph *)calloc(1, sizeof *g);
  g->dim = dim;
  g->M = M;
  g->metric = metric;
  g->flavor = flavor;
  g->stride = M + 1;
  for (int i = 0; i < ORA_MAX_LAYERS; i++) g->entry[i] = -1;
  return g;
}
void ora_graph_free(ora_graph *g) {
  if (!g) return;
  free(g->vec);
  free(g->key);
  free(g->level);
  free(g->dead);
  free(g->adj0);
  free(g->deg0);
  free(g->upper_off);
  free(g->adjU);
  free(g->degU);
  free(g->stamp);
  free(g);
}
int64_t ora_graph_len(const ora_graph *g) { return g->n_layers ? g->layer_size[0] : 0; } /* graph.go:829-838 */
int ora_graph_layers(const ora_graph *g) { return g->n_layers; }
int ora_graph_M(const ora_graph *g) { return g->M; }
int ora_graph_metric(const ora_graph *g) { return g->metric; }
int ora_graph_dim(const ora_graph *g) { return g->dim; }
int64_t ora_graph_layer_size(const ora_graph *g, int l) { return (l >= 0 && l < g->n_layers) ? g->layer_size[l] : 0; }
int32_t ora_graph_entry(const ora_graph *g, int l) { return (l >= 0 && l < g->n_layers) ? g->entry[l] : -1; }
void ora_graph_set_entry(ora_graph *g, int l, int32_t row) {
  if (l >= 0 && l < ORA_MAX_LAYERS) g->entry[l] = row;
}
int ora_graph_level(const ora_graph *g, int32_t row) { return g->level[row]; }
int64_t ora_graph_key(const ora_graph *g, int32_t row) { return g->key[row]; }
const float *ora_graph_vector(const ora_graph *g, int32_t row) { return g->vec + (size_t)row * g->dim; }
void ora_graph_counters(const ora_graph *g, ora_counters *out) { *out = g->ctr; }
void ora_graph_reset_counters(ora_graph *g) { memset(&g->ctr, 0, sizeof g->ctr); }

static inline int32_t *list_ptr(const ora_graph *g, int layer, int32_t row, int32_t **deg) {
  if (layer == 0) {
    *deg = &g->deg0[row];
    return g->adj0 + (size_t)row * g->stride;
  }
  if (g->level[row] < layer) {
    *deg = NULL;
    return NULL;
  }
  int64_t slot = (int64_t)g->upper_off[row] + (layer - 1);
  *deg = &g->degU[slot];
  return g->adjU + (size_t)slot * g->stride;
}

int ora_graph_neighbors(const ora_graph *g, int layer, int32_t row, int32_t *out) {
  int32_t *deg;
  int32_t *l = list_ptr(g, layer, row, &deg);
  if (!l) return -1;
  memcpy(out, l, sizeof(int32_t) * (size_t)*deg);
  return *deg;
}

static void grow_rows(ora_graph *g) {
  int64_t cap = g->cap ? g->cap * 2 : 1024;
  g->vec = (float *)realloc(g->vec, sizeof(float) * (size_t)cap * g->dim);
  g->key = (int64_t *)realloc(g->key, sizeof(int64_t) * (size_t)cap);
  g->level = (int32_t *)realloc(g->level, sizeof(int32_t) * (size_t)cap);
  g->dead = (uint8_t *)realloc(g->dead, (size_t)cap);
  memset(g->dead + g->cap, 0, (size_t)(cap - g->cap));
  g->adj0 = (int32_t *)realloc(g->adj0, sizeof(int32_t) * (size_t)cap * g->stride);
  g->deg0 = (int32_t *)realloc(g->deg0, sizeof(int32_t) * (size_t)cap);
  g->upper_off = (int32_t *)realloc(g->upper_off, sizeof(int32_t) * (size_t)cap);
  g->stamp = (uint32_t *)realloc(g->stamp, sizeof(uint32_t) * (size_t)cap);
  memset(g->stamp + g->cap, 0, sizeof(uint32_t) * (size_t)(cap - g->cap));
  g->cap = cap;
}

/* append storage for a node that will be a member of layers 0..level; no links, no layer bookkeeping */
static int32_t new_row(ora_graph *g, int64_t key, const float *vec, int level) {
  if (g->n == g->cap) grow_rows(g);
  int32_t row = (int32_t)g->n++;
  memcpy(g->vec + (size_t)row * g->dim, vec, sizeof(float) * (size_t)g->dim);
  g->key[row] = key;
  g->level[row] = level;
  g->deg0[row] = 0;
  g->upper_off[row] = -1;
  if (level > 0) {
    while (g->nU + level > g->capU) {
      int64_t cap = g->capU ? g->capU * 2 : 1024;
      g->adjU = (int32_t *)realloc(g->adjU, sizeof(int32_t) * (size_t)cap * g->stride);
      g->degU = (int32_t *)realloc(g->degU, sizeof(int32_t) * (size_t)cap);
      g->capU = cap;
    }
    g->upper_off[row] = (int32_t)g->nU;
    for (int l = 0; l < level; l++) g->degU[g->nU + l] = 0;
    g->nU += level;
  }
  return row;
}

/* ---- sorted-by-key list primitives (Go: map insert / delete on layerNode.neighbors) --------- */
static int list_find(const ora_graph *g, const int32_t *l, int deg, int32_t row) {
  for (int i = 0; i < deg; i++)
    if (l[i] == row) return i;
  (void)g;
  return -1;
}
static void list_insert(const ora_graph *g, int32_t *l, int32_t *deg, int32_t row) {
  if (list_find(g, l, *deg, row) >= 0) return; /* map overwrite: length unchanged */
  int i = *deg;
  int64_t k = g->key[row];
  for (int j = 0; j < *deg; j++)
    if (g->key[l[j]] == k) { /* another row under the same key (a key deleted and added again): the map entry is */
      l[j] = row;            /* overwritten, n.neighbors[key] = newNode (graph.go:50) */
      return;
    }
  while (i > 0 && g->key[l[i - 1]] > k) {
    l[i] = l[i - 1];
    i--;
  }
  l[i] = row;
  (*deg)++;
}
static void list_delete(const ora_graph *g, int32_t *l, int32_t *deg, int32_t row) {
  int i = list_find(g, l, *deg, row);
  if (i < 0) return;
  for (; i + 1 < *deg; i++) l[i] = l[i + 1];
  (*deg)--;
}

static inline float gdist(const ora_graph *g, int metric, int32_t a, const float *b) {
  return ora_distance(metric, g->flavor, g->vec + (size_t)a * g->dim, b, g->dim);
}

static void replenish(ora_graph *g, int layer, int32_t n, int m);

/* layerNode.addNeighbor -- graph.go:41-81 */
static void add_neighbor(ora_graph *g, int layer, int32_t n, int32_t newn, int m, int metric) {
  int32_t *deg;
  int32_t *l = list_ptr(g, layer, n, &deg);
  if (!l) return;
  list_insert(g, l, deg, newn); /* graph.go:50 */
  if (*deg <= m) return;        /* graph.go:51-53 */
  /* graph.go:56-71: scan every neighbour (ascending key stands in for map order) for the worst */
  float worst_dist = -INFINITY;
  int32_t worst = -1;
  const float *nv = g->vec + (size_t)n * g->dim;
  for (int i = 0; i < *deg; i++) {
    float d = gdist(g, metric, l[i], nv); /* dist(neighbor.Value, n.Value) */
    g->ctr.evict_evals++;
    if (d > worst_dist || worst < 0) {
      worst_dist = d;
      worst = l[i];
    }
  }
  if (worst >= 0) { /* graph.go:73-80 */
    g->ctr.evictions++;
    list_delete(g, l, deg, worst);
    int32_t *wdeg;
    int32_t *wl = list_ptr(g, layer, worst, &wdeg);
    if (wl) list_delete(g, wl, wdeg, n);
    replenish(g, layer, worst, m);
  }
}

/* layerNode.replenish -- graph.go:172-219.  Always CosineDistance (graph.go:204, 216). */
static void replenish(ora_graph *g, int layer, int32_t n, int m) {
  int32_t *deg;
  int32_t *l = list_ptr(g, layer, n, &deg);
  if (!l || *deg >= m) return; /* graph.go:173-175 */
  ora_heap cand;
  ora_heap_init(&cand, m * 2);
  /* visited = {n} U neighbours(n) (graph.go:183-190); epoch stamps play the map[K]bool */
  if (++g->epoch == 0) {
    memset(g->stamp, 0, sizeof(uint32_t) * (size_t)g->cap);
    g->epoch = 1;
  }
  g->stamp[n] = g->epoch;
  for (int i = 0; i < *deg; i++) g->stamp[l[i]] = g->epoch;
  const float *nv = g->vec + (size_t)n * g->dim;
  for (int i = 0; i < *deg; i++) { /* graph.go:193-210 */
    int32_t *ndeg;
    int32_t *nl = list_ptr(g, layer, l[i], &ndeg);
    if (!nl) continue;
    for (int j = 0; j < *ndeg; j++) {
      int32_t c = nl[j];
      if (g->stamp[c] == g->epoch) continue;
      g->stamp[c] = g->epoch;
      float d = gdist(g, ORA_COSINE, c, nv); /* CosineDistance(candidate.Value, n.Value) */
      g->ctr.replenish_evals++;
      ora_heap_push(&cand, d, c);
    }
  }
  while (cand.len > 0 && *deg < m) { /* graph.go:213-218 */
    float d;
    int32_t best;
    ora_heap_pop(&cand, &d, &best);
    add_neighbor(g, layer, n, best, m, ORA_COSINE);
  }
  ora_heap_free(&cand);
}

/* order-preserving image of a distance for the extended mode's keys: ascending unsigned order == ascending float
 * order; -0 is taken as +0, a NaN (any sign) as the canonical positive NaN, which sorts after +inf */
static inline uint32_t ord_bits(float d) {
  if (d != d) return 0x7fc00000u | 0x80000000u;
  d = d + 0.0f;
  uint32_t u;
  memcpy(&u, &d, 4);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
static inline float ord_float(uint32_t o) {
  uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
  float d;
  memcpy(&d, &u, 4);
  return d;
}

/* ---- layerNode.search -- graph.go:94-170 ----------------------------------------------------- */
static int layer_search(const ora_graph *g, visited_t *vis, ora_counters *ctr, int layer, int32_t entry, int k,
                        int ef, const float *target, int mode, int32_t *out_rows, float *out_dists) {
  if (entry < 0) return 0; /* graph.go:101-103 */
  ora_heap cand, res;
  ora_heap_init(&cand, ef + 1);
  ora_heap_init(&res, k + 1);
  if (++vis->epoch == 0) { /* stamp wrap */
    memset(vis->stamp, 0, sizeof(uint32_t) * (size_t)vis->cap);
    vis->epoch = 1;
  }
  float d0 = gdist(g, g->metric, entry, target);
  ctr->dist_evals++;
  TRACE_EVAL(entry, d0);
  ora_heap_push(&cand, d0, entry);           /* graph.go:109-114 */
  ora_heap_push(&res, cand.dist[0], cand.id[0]); /* graph.go:122 */
  vis->stamp[entry] = vis->epoch;            /* graph.go:123 */

  if (mode == ORA_MODE_REFERENCE) {
    while (cand.len > 0) { /* graph.go:125 */
      int32_t cur;
      ora_heap_pop(&cand, NULL, &cur); /* graph.go:127 */
      int improved = 0;
      int32_t *deg;
      int32_t *l = list_ptr(g, layer, cur, &deg);
      if (!l) continue; /* graph.go:131-133 */
      ctr->expansions++;
      for (int i = 0; i < *deg; i++) { /* graph.go:137-160, keys already ascending */
        int32_t nb = l[i];
        ctr->nbr_reads++;
        if (vis->stamp[nb] == vis->epoch) continue;
        vis->stamp[nb] = vis->epoch;
        float d = gdist(g, g->metric, nb, target);
        ctr->dist_evals++;
        TRACE_EVAL(nb, d);
        improved = improved || (res.len > 0 && d < res.dist[0]); /* graph.go:147 */
        if (res.len < k) {
          ora_heap_push(&res, d, nb);
        } else if (d < res.dist[res.len - 1]) { /* result.Max() == last slot */
          ora_heap_poplast(&res, NULL, NULL);
          ora_heap_push(&res, d, nb);
        }
        ora_heap_push(&cand, d, nb);
        if (cand.len > ef) ora_heap_poplast(&cand, NULL, NULL); /* graph.go:157-159 */
      }
      if (!improved && res.len >= k) break; /* graph.go:164-166 */
    }
  } else {
    /* ORA_MODE_EXTENDED: NOT the reference.  Textbook HNSW layer search (Malkov & Yashunin, Alg. 2), stated so
     * that its result is a function of the graph and the query only (no heap-layout or arrival-order effects):
     *
     *   key(d, row) orders candidates by distance, ties by row; -0 counts as +0; a NaN distance never enters W
     *   (the entry point is always taken, a NaN distance there sorts last).
     *   W = the efk = max(ef, k) smallest keys among the nodes evaluated so far, each with an `expanded` flag.
     *   Repeat: take the smallest unexpanded element of W, mark it, evaluate its unvisited neighbours, merge them
     *   into W.  Stop when every element of W is expanded.  (Alg. 2's candidate queue C only ever adds elements
     *   that are farther than W's worst at the time they leave W, and W's worst never grows, so "nearest element
     *   of C, stop when it is farther than W's worst" visits exactly the unexpanded elements of W in key order.)
     *   Result: the first min(k, |W|) elements of W, ascending.
     * The device kernel (hnsw_b200/csrc/traverse_ext.cuh) implements the same statement and is held to it bit
     * for bit, E / H / neighbour-read counters included. */
    ora_heap_free(&cand);
    ora_heap_free(&res);
    int efk = ef > k ? ef : k;
    uint64_t *W = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)(efk + 1));
    unsigned char *ex = (unsigned char *)malloc((size_t)(efk + 1));
    int wn = 0;
#define ORD_KEY(dv, rowv) ((((uint64_t)ord_bits(dv)) << 32) | (uint32_t)(rowv))
    W[0] = ORD_KEY(d0, entry);
    ex[0] = 0;
    wn = 1;
    for (;;) {
      int ci = 0;
      while (ci < wn && ex[ci]) ci++;
      if (ci == wn) break;
      ex[ci] = 1;
      int32_t cur = (int32_t)(W[ci] & 0xffffffffu);
      int32_t *deg;
      int32_t *l = list_ptr(g, layer, cur, &deg);
      if (!l) continue;
      ctr->expansions++;
      for (int i = 0; i < *deg; i++) {
        int32_t nb = l[i];
        ctr->nbr_reads++;
        if (vis->stamp[nb] == vis->epoch) continue;
        vis->stamp[nb] = vis->epoch;
        float d = gdist(g, g->metric, nb, target);
        ctr->dist_evals++;
        TRACE_EVAL(nb, d);
        if (d != d) continue;
        uint64_t key = ORD_KEY(d, nb);
        if (wn == efk) {
          if (!(key < W[wn - 1])) continue;
          wn--;
        }
        int pos = wn;
        while (pos > 0 && W[pos - 1] > key) {
          W[pos] = W[pos - 1];
          ex[pos] = ex[pos - 1];
          pos--;
        }
        W[pos] = key;
        ex[pos] = 0;
        wn++;
      }
    }
#undef ORD_KEY
    int n = wn < k ? wn : k;
    for (int i = 0; i < n; i++) {
      out_rows[i] = (int32_t)(W[i] & 0xffffffffu);
      out_dists[i] = ord_float((uint32_t)(W[i] >> 32));
    }
    free(W);
    free(ex);
    return n;
  }
  int n = res.len; /* graph.go:169: Slice() = raw heap array */
  for (int i = 0; i < n; i++) {
    out_rows[i] = res.id[i];
    out_dists[i] = res.dist[i];
  }
  ora_heap_free(&cand);
  ora_heap_free(&res);
  return n;
}

int ora_layer_search(ora_graph *g, int layer, int32_t entry, int k, int ef, const float *target, int mode,
                     int32_t *out_rows, float *out_dists) {
  visited_t v = {g->stamp, g->epoch, g->cap};
  int n = layer_search(g, &v, &g->ctr, layer, entry, k, ef, target, mode, out_rows, out_dists);
  g->epoch = v.epoch;
  return n;
}

/* Graph.Search body -- graph.go:571-593 (== BatchSearch body, graph.go:1081-1106) */
static int graph_search(const ora_graph *g, visited_t *vis, ora_counters *ctr, const float *q, int k, int ef,
                        int mode, const int32_t *entries, int32_t *out_rows, float *out_dists) {
  int32_t elevator = -1;
  for (int layer = g->n_layers - 1; layer >= 0; layer--) {
    int32_t sp = entries ? entries[layer] : g->entry[layer];
    if (elevator >= 0) sp = g->dead[elevator] ? -1 : elevator; /* layer.nodes[*elevator]: nil once deleted */
    if (layer > 0) {
      int32_t r;
      float d;
      /* extended mode keeps the reference's greedy k=1 descent on upper layers with ef=1 */
      int n = layer_search(g, vis, ctr, layer, sp, 1, mode == ORA_MODE_EXTENDED ? 1 : ef, q, mode, &r, &d);
      if (n == 0) continue; /* graph.go:580-583 */
      elevator = r;
      continue;
    }
    return layer_search(g, vis, ctr, 0, sp, k, ef, q, mode, out_rows, out_dists);
  }
  return 0;
}

int ora_graph_search(ora_graph *g, const float *query, int k, int ef, int mode, const int32_t *entries,
                     int32_t *out_rows, float *out_dists) {
  visited_t v = {g->stamp, g->epoch, g->cap};
  int n = graph_search(g, &v, &g->ctr, query, k, ef, mode, entries, out_rows, out_dists);
  g->epoch = v.epoch;
  return n;
}

typedef struct {
  const ora_graph *g;
  const float *queries;
  int64_t q0, q1;
  int k, ef, mode;
  int32_t *out_rows;
  float *out_dists;
  int32_t *out_counts;
  ora_counters ctr;
  int32_t *trace_rows; /* [nq][trace_cap] or NULL */
  float *trace_dists;
  int64_t *trace_len;  /* [nq]: evaluations made (may exceed trace_cap: truncated) */
  int64_t trace_cap;
} bs_job;

static void *bs_worker(void *arg) {
  bs_job *j = (bs_job *)arg;
  visited_t v;
  v.cap = j->g->cap;
  v.stamp = (uint32_t *)calloc((size_t)(v.cap ? v.cap : 1), sizeof(uint32_t));
  v.epoch = 0;
  for (int64_t i = j->q0; i < j->q1; i++) {
    ora_trace tr;
    if (j->trace_rows) {
      tr.rows = j->trace_rows + (size_t)i * j->trace_cap;
      tr.dists = j->trace_dists + (size_t)i * j->trace_cap;
      tr.cap = j->trace_cap;
      tr.n = 0;
      tls_trace = &tr;
    }
    int n = graph_search(j->g, &v, &j->ctr, j->queries + (size_t)i * j->g->dim, j->k, j->ef, j->mode, NULL,
                         j->out_rows + (size_t)i * j->k, j->out_dists + (size_t)i * j->k);
    j->out_counts[i] = n;
    if (j->trace_rows) {
      j->trace_len[i] = tr.n;
      tls_trace = NULL;
    }
  }
  free(v.stamp);
  return NULL;
}

void ora_graph_batch_search(ora_graph *g, const float *queries, int64_t nq, int k, int ef, int mode, int n_threads,
                            int32_t *out_rows, float *out_dists, int32_t *out_counts, ora_counters *total) {
  ora_graph_batch_search_trace(g, queries, nq, k, ef, mode, n_threads, out_rows, out_dists, out_counts, total, NULL,
                               NULL, NULL, 0);
}

/* the same, also recording per query the rows it evaluated, in order, with their distances (flavour report) */
void ora_graph_batch_search_trace(ora_graph *g, const float *queries, int64_t nq, int k, int ef, int mode,
                                  int n_threads, int32_t *out_rows, float *out_dists, int32_t *out_counts,
                                  ora_counters *total, int32_t *trace_rows, float *trace_dists, int64_t *trace_len,
                                  int64_t trace_cap) {
  if (n_threads < 1) n_threads = 1;
  if (n_threads > nq) n_threads = nq > 0 ? (int)nq : 1;
  bs_job *jobs = (bs_job *)calloc((size_t)n_threads, sizeof *jobs);
  pthread_t *th = (pthread_t *)calloc((size_t)n_threads, sizeof *th);
  for (int t = 0; t < n_threads; t++) {
    jobs[t].g = g;
    jobs[t].queries = queries;
    jobs[t].q0 = nq * t / n_threads;
    jobs[t].q1 = nq * (t + 1) / n_threads;
    jobs[t].k = k;
    jobs[t].ef = ef;
    jobs[t].mode = mode;
    jobs[t].out_rows = out_rows;
    jobs[t].out_dists = out_dists;
    jobs[t].out_counts = out_counts;
    jobs[t].trace_rows = trace_rows;
    jobs[t].trace_dists = trace_dists;
    jobs[t].trace_len = trace_len;
    jobs[t].trace_cap = trace_cap;
    if (n_threads == 1) bs_worker(&jobs[t]);
    else pthread_create(&th[t], NULL, bs_worker, &jobs[t]);
  }
  ora_counters sum;
  memset(&sum, 0, sizeof sum);
  for (int t = 0; t < n_threads; t++) {
    if (n_threads > 1) pthread_join(th[t], NULL);
    sum.dist_evals += jobs[t].ctr.dist_evals;
    sum.expansions += jobs[t].ctr.expansions;
    sum.nbr_reads += jobs[t].ctr.nbr_reads;
  }
  if (total) *total = sum;
  free(jobs);
  free(th);
}

/* ---- Graph.Add for one node -- graph.go:445-528 ----------------------------------------------- */
int32_t ora_graph_add(ora_graph *g, int64_t key, const float *vec, int level, int ef) {
  if (level < 0 || level >= ORA_MAX_LAYERS) return -1;
  /* A layer emptied by deletions above the drawn level: the reference parks the node there without the layers
   * in between (graph.go:485-488), after which every Search loses its elevator.  Not reproduced: refused. */
  for (int i = level + 1; i < g->n_layers; i++)
    if (g->layer_size[i] == 0) return -3;
  /* The reference interleaves search and linking per layer (graph.go:475-522).  A layer's search reads that
   * layer's lists only and linking at layer i writes layer-i lists only, so running every search first gives the
   * same neighbourhoods; it lets a failed Add (deleted elevator, graph.go:501-505) leave the graph untouched,
   * where the reference leaves the node half inserted. */
  const int M = g->M;
  const int top = (g->n_layers > level + 1 ? g->n_layers : level + 1) - 1; /* graph.go:462-464 */
  int32_t *nb_rows = (int32_t *)malloc(sizeof(int32_t) * (size_t)M * (size_t)(top + 1));
  float *nb_d = (float *)malloc(sizeof(float) * (size_t)M);
  int cnts[ORA_MAX_LAYERS];
  int32_t elevator = -1;
  for (int i = top; i >= 0; i--) { /* graph.go:475 */
    cnts[i] = -1;
    if (i >= g->n_layers || g->layer_size[i] == 0) continue; /* graph.go:485-488 */
    int32_t sp = g->entry[i];                                 /* graph.go:492 */
    if (elevator >= 0) sp = g->dead[elevator] ? -1 : elevator; /* graph.go:496-498: layer.nodes[*elevator] */
    int cnt = ora_layer_search(g, i, sp, M, ef, vec, g->build_mode, nb_rows + (size_t)i * M, nb_d); /* :500 */
    if (cnt == 0) { /* graph.go:501-505 "no nodes found in neighborhood search" */
      free(nb_rows);
      free(nb_d);
      return -2;
    }
    cnts[i] = cnt;
    elevator = nb_rows[(size_t)i * M]; /* graph.go:508 */
  }
  int32_t row = new_row(g, key, vec, level);
  while (level >= g->n_layers) {
    g->entry[g->n_layers] = -1;
    g->layer_size[g->n_layers] = 0;
    g->n_layers++;
  }
  for (int i = top; i >= 0; i--) {
    if (cnts[i] < 0) { /* empty layer takes the node as its only member */
      g->entry[i] = row;
      g->layer_size[i] = 1;
      continue;
    }
    if (level >= i) {     /* graph.go:510 */
      g->layer_size[i]++; /* graph.go:515 */
      for (int t = 0; t < cnts[i]; t++) { /* graph.go:516-520, heap-array order */
        add_neighbor(g, i, nb_rows[(size_t)i * M + t], row, M, g->metric);
        add_neighbor(g, i, row, nb_rows[(size_t)i * M + t], M, g->metric);
      }
    }
  }
  free(nb_rows);
  free(nb_d);
  return row;
}

/* ---- Graph.Delete -- graph.go:843-866; layerNode.isolate -- graph.go:223-235 ------------------
 * The key leaves layer.nodes of every layer it is in (base layer first, the order of h.layers); per layer every
 * neighbour drops its link to the node and is replenished at once, in ascending key order (the stand-in for
 * Go's map order).  As in the reference the node object survives: its own list is untouched, a replenish can
 * link it again one-way through a neighbour that still lists it, traversals walk through it and can return it;
 * only layer.nodes lookups (entry point, elevator, Lookup, Len) stop seeing it. */
void ora_graph_set_build_mode(ora_graph *g, int mode) { g->build_mode = mode; }
/* summation order used from now on (the flavour report and the bench search one loaded graph under several) */
void ora_graph_set_flavor(ora_graph *g, int flavor) { g->flavor = flavor; }
int ora_graph_is_dead(const ora_graph *g, int32_t row) { return row >= 0 && row < g->n ? g->dead[row] : -1; }
int64_t ora_graph_dead_count(const ora_graph *g) { return g->n_dead; }

int ora_graph_delete_row(ora_graph *g, int32_t row) {
  if (row < 0 || row >= g->n || g->dead[row]) return 0;
  g->dead[row] = 1;
  g->n_dead++;
  int32_t *nb = (int32_t *)malloc(sizeof(int32_t) * (size_t)g->stride);
  for (int layer = 0; layer <= g->level[row] && layer < g->n_layers; layer++) {
    g->layer_size[layer]--; /* delete(layer.nodes, key) */
    if (g->entry[layer] == row) { /* entry(): lowest live member row takes over */
      int32_t e = -1;
      for (int32_t r = 0; r < (int32_t)g->n; r++)
        if (!g->dead[r] && g->level[r] >= layer) {
          e = r;
          break;
        }
      g->entry[layer] = e;
    }
    int32_t *deg;
    int32_t *l = list_ptr(g, layer, row, &deg);
    int n = *deg;
    memcpy(nb, l, sizeof(int32_t) * (size_t)n); /* the node's own map is not modified by isolate */
    for (int i = 0; i < n; i++) {
      int32_t *ndeg;
      int32_t *nl = list_ptr(g, layer, nb[i], &ndeg);
      if (!nl) continue;
      list_delete(g, nl, ndeg, row); /* graph.go:232 */
      replenish(g, layer, nb[i], g->M); /* graph.go:233 */
    }
  }
  free(nb);
  return 1;
}

int ora_graph_delete(ora_graph *g, int64_t key) {
  for (int32_t r = (int32_t)g->n - 1; r >= 0; r--)
    if (!g->dead[r] && g->key[r] == key) return ora_graph_delete_row(g, r);
  return 0;
}

int32_t ora_graph_put_node(ora_graph *g, int64_t key, const float *vec, int level) {
  int32_t row = new_row(g, key, vec, level);
  while (level >= g->n_layers) {
    g->entry[g->n_layers] = -1;
    g->layer_size[g->n_layers] = 0;
    g->n_layers++;
  }
  for (int l = 0; l <= level; l++) {
    if (g->entry[l] < 0) g->entry[l] = row;
    g->layer_size[l]++;
  }
  return row;
}

int ora_graph_link(ora_graph *g, int layer, int32_t from, int32_t to) {
  int32_t *deg;
  int32_t *l = list_ptr(g, layer, from, &deg);
  if (!l || *deg >= g->stride) return -1;
  list_insert(g, l, deg, to);
  return 0;
}

/* rows that own a list in `layer` (members, and former members removed by Delete) */
int64_t ora_graph_layer_rows(const ora_graph *g, int layer) {
  int64_t nm = 0;
  for (int32_t r = 0; r < (int32_t)g->n; r++) nm += g->level[r] >= layer;
  return nm;
}
/* bulk read-back for the test harness: CSR of one layer over the rows that own a list in it (ascending row) */
int64_t ora_graph_dump_layer(const ora_graph *g, int layer, int32_t *rows, int64_t *ptr, int32_t *nbr) {
  int64_t nm = 0, nn = 0;
  for (int32_t r = 0; r < (int32_t)g->n; r++) {
    if (g->level[r] < layer) continue;
    int32_t *deg;
    int32_t *l = list_ptr(g, layer, r, &deg);
    if (rows) rows[nm] = r;
    if (ptr) ptr[nm] = nn;
    if (nbr) memcpy(nbr + nn, l, sizeof(int32_t) * (size_t)*deg);
    nn += *deg;
    nm++;
  }
  if (ptr) ptr[nm] = nn;
  return nn;
}
void ora_graph_dump_rows(const ora_graph *g, float *vec, int64_t *keys, int32_t *levels) {
  memcpy(vec, g->vec, sizeof(float) * (size_t)g->n * g->dim);
  memcpy(keys, g->key, sizeof(int64_t) * (size_t)g->n);
  memcpy(levels, g->level, sizeof(int32_t) * (size_t)g->n);
}
/* Graph.Add for n nodes in input order (the serial loop of graph.go:445) */
int64_t ora_graph_add_many(ora_graph *g, const int64_t *keys, const float *vecs, const int32_t *levels, int64_t n,
                           int ef) {
  for (int64_t i = 0; i < n; i++)
    if (ora_graph_add(g, keys[i], vecs + (size_t)i * g->dim, levels[i], ef) < 0) return i;
  return n;
}

/* Fill an EMPTY graph from flat arrays (the shape the device library dumps): rows 0..n-1 with vectors, keys and
 * levels; per layer a CSR over its member rows (rows[], ptr[], nbr[] concatenated layer after layer, each layer's
 * ptr starting at 0).  Lists are re-sorted by key.  Returns 0 or a negative error. */
int ora_graph_load_flat(ora_graph *g, int64_t n, const float *vecs, const int64_t *keys, const int32_t *levels,
                        int n_layers, const int32_t *entries, const int64_t *layer_sizes, const int32_t *layer_rows,
                        const int64_t *layer_ptr, const int32_t *layer_nbr) {
  if (g->n != 0) return -1;
  for (int64_t i = 0; i < n; i++) new_row(g, keys[i], vecs + (size_t)i * g->dim, levels[i]);
  g->n_layers = n_layers;
  int64_t roff = 0, poff = 0, noff = 0;
  for (int l = 0; l < n_layers; l++) {
    int64_t nl = layer_sizes[l];
    g->layer_size[l] = nl;
    g->entry[l] = entries ? entries[l] : (nl ? layer_rows[roff] : -1);
    for (int64_t i = 0; i < nl; i++) {
      int32_t row = layer_rows[roff + i];
      if (row < 0 || row >= n || levels[row] < l) return -2;
      int64_t b = layer_ptr[poff + i], e = layer_ptr[poff + i + 1];
      if (e - b > g->M) return -3;
      for (int64_t t = b; t < e; t++)
        if (ora_graph_link(g, l, row, layer_nbr[noff + t])) return -4;
    }
    noff += layer_ptr[poff + nl];
    roff += nl;
    poff += nl + 1;
  }
  return 0;
}

/* ================================================================================================
 * WAVE-BATCHED ADD -- NOT the reference's algorithm.
 *
 * The reference inserts strictly one node after the other (graph.go:445, one write lock).  The device
 * library also offers a batched construction (north_star: "Batched Add evaluates candidates in parallel
 * in the same kernels and applies the same neighbor-selection heuristic"): nodes are inserted in WAVES;
 * all nodes of a wave search the graph as it was before the wave, then their links are committed in
 * deterministic phases that apply the reference's heuristic (link to every search result, evict the
 * farthest on overflow, delete the back-link, replenish the evicted node with CosineDistance).
 * This function is the CPU restatement of THAT specification, so the device path can be checked bit for
 * bit.  It is pinned to nothing in the reference; what it borrows from the reference is cited inline.
 *
 *   schedule  : ora_wave_len() below (identical formula in hnsw_b200/csrc/wave_insert.cu)
 *   phase S   : per new node, per layer top->0: search(k=M, ef) from the elevator (graph.go:475-508)
 *   phase A   : the new node's own list at each of its layers = its search results (graph.go:519)
 *   phase B   : per (layer, target): insert the new rows in input order; on overflow evict the farthest
 *               (graph.go:50-73, scan in ascending key order, rule `d > worst || worst == nil`)
 *   phase C   : per eviction (n, worst): delete n from worst's list (graph.go:75-78)
 *   phase D   : per evicted node with fewer than M links: replenish from the lists as they are after
 *               phase C (graph.go:172-219), results staged
 *   phase E   : staged lists replace the old ones
 * ================================================================================================ */
int64_t ora_wave_len(int64_t cur_rows, int n_layers, const int32_t *levels, int64_t remaining, int64_t wave_max) {
  if (remaining <= 0) return 0;
  if (n_layers == 0) return 1; /* the very first node */
  int64_t cap = cur_rows / 16;
  if (cap < 1) cap = 1;
  if (cap > wave_max) cap = wave_max;
  int64_t w = 0;
  while (w < remaining && w < cap) {
    if (levels[w] >= n_layers) { /* a node that opens new layers is inserted alone */
      if (w == 0) w = 1;
      break;
    }
    w++;
  }
  return w;
}

typedef struct {
  int32_t row, layer, cnt;
  int32_t *rows; /* [M] heap-array order */
  float *dists;  /* [M] */
} wave_slot;

typedef struct {
  ora_graph *g;
  wave_slot *slots;
  const int64_t *slot_off; /* first slot of node j */
  int32_t first_row;
  int64_t j0, j1;
  int ef, old_layers, fail;
  ora_counters ctr;
} waveS_job;

static void *waveS_worker(void *arg) {
  waveS_job *J = (waveS_job *)arg;
  ora_graph *g = J->g;
  visited_t v;
  v.cap = g->cap;
  v.stamp = (uint32_t *)calloc((size_t)v.cap, sizeof(uint32_t));
  v.epoch = 0;
  int M = g->M;
  int32_t *rows = (int32_t *)malloc(sizeof(int32_t) * (size_t)M);
  float *d = (float *)malloc(sizeof(float) * (size_t)M);
  for (int64_t j = J->j0; j < J->j1; j++) {
    int32_t row = J->first_row + (int32_t)j;
    int level = g->level[row];
    const float *vec = g->vec + (size_t)row * g->dim;
    int32_t elevator = -1;
    for (int i = J->old_layers - 1; i >= 0; i--) {
      int32_t sp = elevator >= 0 ? elevator : g->entry[i];
      int cnt = layer_search(g, &v, &J->ctr, i, sp, M, J->ef, vec, g->build_mode, rows, d);
      if (cnt == 0) {
        J->fail = 1;
        break;
      }
      elevator = rows[0];
      if (level >= i) {
        wave_slot *s = &J->slots[J->slot_off[j] + i];
        s->row = row;
        s->layer = i;
        s->cnt = cnt;
        memcpy(s->rows, rows, sizeof(int32_t) * (size_t)cnt);
        memcpy(s->dists, d, sizeof(float) * (size_t)cnt);
      }
    }
  }
  free(rows);
  free(d);
  free(v.stamp);
  return NULL;
}

typedef struct {
  int32_t layer, n, worst;
} wave_evict;

typedef struct {
  int32_t layer, row, deg; /* deg after replenish */
  int32_t *list;           /* [M+1] staged list */
} wave_rep;

typedef struct {
  ora_graph *g;
  wave_rep *reps;
  int64_t r0, r1;
  ora_counters ctr;
} waveD_job;

/* graph.go:172-219 on the lists as they stand (read-only), result written to rep->list */
static void wave_replenish(const ora_graph *g, wave_rep *rep, ora_counters *ctr, uint32_t *tab, uint32_t mask) {
  int32_t *deg;
  const int32_t *l = list_ptr(g, rep->layer, rep->row, &deg);
  int m = g->M, dg = *deg;
  memcpy(rep->list, l, sizeof(int32_t) * (size_t)dg);
  rep->deg = dg;
  if (dg >= m) return;
  memset(tab, 0xff, sizeof(uint32_t) * (size_t)(mask + 1));
#define TAB_INSERT(idv, fresh)                                   \
  do {                                                           \
    uint32_t h__ = ((uint32_t)(idv) * 2654435761u) >> 7;         \
    for (;;) {                                                   \
      h__ &= mask;                                               \
      if (tab[h__] == 0xffffffffu) {                             \
        tab[h__] = (uint32_t)(idv);                              \
        fresh = 1;                                               \
        break;                                                   \
      }                                                          \
      if (tab[h__] == (uint32_t)(idv)) {                         \
        fresh = 0;                                               \
        break;                                                   \
      }                                                          \
      h__++;                                                     \
    }                                                            \
  } while (0)
  int fresh;
  TAB_INSERT(rep->row, fresh);
  for (int i = 0; i < dg; i++) TAB_INSERT(l[i], fresh);
  ora_heap cand;
  ora_heap_init(&cand, m * 2);
  const float *nv = g->vec + (size_t)rep->row * g->dim;
  for (int i = 0; i < dg; i++) {
    int32_t *ndeg;
    const int32_t *nl = list_ptr(g, rep->layer, l[i], &ndeg);
    if (!nl) continue;
    for (int j = 0; j < *ndeg; j++) {
      int32_t c = nl[j];
      TAB_INSERT(c, fresh);
      if (!fresh) continue;
      float d = gdist(g, ORA_COSINE, c, nv);
      ctr->replenish_evals++;
      ora_heap_push(&cand, d, c);
    }
  }
#undef TAB_INSERT
  while (cand.len > 0 && rep->deg < m) {
    float d;
    int32_t best;
    ora_heap_pop(&cand, &d, &best);
    list_insert(g, rep->list, &rep->deg, best);
  }
  ora_heap_free(&cand);
}

static void *waveD_worker(void *arg) {
  waveD_job *J = (waveD_job *)arg;
  int M = J->g->M;
  uint32_t cap = 1;
  while (cap < (uint32_t)(2 * (M * M + M + 2))) cap <<= 1;
  uint32_t *tab = (uint32_t *)malloc(sizeof(uint32_t) * cap);
  for (int64_t r = J->r0; r < J->r1; r++) wave_replenish(J->g, &J->reps[r], &J->ctr, tab, cap - 1);
  free(tab);
  return NULL;
}

static void ctr_add(ora_counters *a, const ora_counters *b) {
  a->dist_evals += b->dist_evals;
  a->expansions += b->expansions;
  a->nbr_reads += b->nbr_reads;
  a->evict_evals += b->evict_evals;
  a->replenish_evals += b->replenish_evals;
  a->evictions += b->evictions;
}

static int evict_cmp(const void *a, const void *b) {
  const wave_evict *x = (const wave_evict *)a, *y = (const wave_evict *)b;
  if (x->layer != y->layer) return x->layer < y->layer ? -1 : 1;
  if (x->worst != y->worst) return x->worst < y->worst ? -1 : 1;
  return 0;
}

/* one wave: rows first_row .. first_row+w-1 are already stored (new_row) but not linked */
static int wave_commit(ora_graph *g, int32_t first_row, int64_t w, int ef, int n_threads) {
  const int M = g->M, old_layers = g->n_layers;
  int64_t *slot_off = (int64_t *)malloc(sizeof(int64_t) * (size_t)(w + 1));
  int64_t ns = 0;
  for (int64_t j = 0; j < w; j++) {
    slot_off[j] = ns;
    ns += g->level[first_row + j] + 1;
  }
  slot_off[w] = ns;
  wave_slot *slots = (wave_slot *)calloc((size_t)ns, sizeof *slots);
  int32_t *srows = (int32_t *)malloc(sizeof(int32_t) * (size_t)ns * M);
  float *sd = (float *)malloc(sizeof(float) * (size_t)ns * M);
  for (int64_t s = 0; s < ns; s++) {
    slots[s].rows = srows + (size_t)s * M;
    slots[s].dists = sd + (size_t)s * M;
    slots[s].layer = -1;
  }
  /* ---- phase S ---- */
  if (n_threads < 1) n_threads = 1;
  int nt = (int)(w < n_threads ? w : n_threads);
  waveS_job *sj = (waveS_job *)calloc((size_t)nt, sizeof *sj);
  pthread_t *th = (pthread_t *)calloc((size_t)n_threads, sizeof *th);
  int fail = 0;
  for (int t = 0; t < nt; t++) {
    sj[t].g = g;
    sj[t].slots = slots;
    sj[t].slot_off = slot_off;
    sj[t].first_row = first_row;
    sj[t].j0 = w * t / nt;
    sj[t].j1 = w * (t + 1) / nt;
    sj[t].ef = ef;
    sj[t].old_layers = old_layers;
    if (nt == 1) waveS_worker(&sj[t]);
    else pthread_create(&th[t], NULL, waveS_worker, &sj[t]);
  }
  for (int t = 0; t < nt; t++) {
    if (nt > 1) pthread_join(th[t], NULL);
    ctr_add(&g->ctr, &sj[t].ctr);
    fail |= sj[t].fail;
  }
  free(sj);
  /* ---- phase A: own lists; layer membership ---- */
  int64_t nreq = 0;
  for (int64_t s = 0; s < ns && !fail; s++) {
    wave_slot *S = &slots[s];
    if (S->layer < 0) continue; /* a layer this wave opens: no links there */
    int32_t *deg;
    int32_t *l = list_ptr(g, S->layer, S->row, &deg);
    for (int t = 0; t < S->cnt; t++) list_insert(g, l, deg, S->rows[t]);
    nreq += S->cnt;
  }
  /* ---- phase B: targets take the new rows in input order; evictions recorded ---- */
  wave_evict *ev = (wave_evict *)malloc(sizeof(wave_evict) * (size_t)(nreq + 1));
  int64_t nev = 0;
  for (int64_t s = 0; s < ns && !fail; s++) {
    wave_slot *S = &slots[s];
    if (S->layer < 0) continue;
    for (int t = 0; t < S->cnt; t++) {
      int32_t n = S->rows[t];
      int32_t *deg;
      int32_t *l = list_ptr(g, S->layer, n, &deg);
      list_insert(g, l, deg, S->row);
      if (*deg <= M) continue;
      float worst_dist = -INFINITY;
      int32_t worst = -1;
      const float *nv = g->vec + (size_t)n * g->dim;
      for (int i = 0; i < *deg; i++) {
        float d = gdist(g, g->metric, l[i], nv);
        g->ctr.evict_evals++;
        if (d > worst_dist || worst < 0) {
          worst_dist = d;
          worst = l[i];
        }
      }
      g->ctr.evictions++;
      list_delete(g, l, deg, worst);
      ev[nev].layer = S->layer;
      ev[nev].n = n;
      ev[nev].worst = worst;
      nev++;
    }
  }
  /* ---- phase C: back-links.  A row of this wave that a target evicted keeps its own (now one-directional) link,
   * as the reference's newNode.addNeighbor(n) re-adds it right after n.addNeighbor(newNode) (graph.go:518-519),
   * and is not replenished. ---- */
  {
    int64_t kept = 0;
    for (int64_t e = 0; e < nev; e++)
      if (ev[e].worst < first_row) ev[kept++] = ev[e];
    nev = kept;
  }
  for (int64_t e = 0; e < nev; e++) {
    int32_t *deg;
    int32_t *l = list_ptr(g, ev[e].layer, ev[e].worst, &deg);
    if (l) list_delete(g, l, deg, ev[e].n);
  }
  /* ---- phase D: replenish every distinct evicted node from the post-C lists, staged ---- */
  qsort(ev, (size_t)nev, sizeof *ev, evict_cmp);
  int64_t nrep = 0;
  for (int64_t e = 0; e < nev; e++)
    if (e == 0 || evict_cmp(&ev[e - 1], &ev[e]) != 0) ev[nrep++] = ev[e];
  wave_rep *reps = (wave_rep *)calloc((size_t)(nrep + 1), sizeof *reps);
  int32_t *rl = (int32_t *)malloc(sizeof(int32_t) * (size_t)(nrep + 1) * (M + 1));
  for (int64_t r = 0; r < nrep; r++) {
    reps[r].layer = ev[r].layer;
    reps[r].row = ev[r].worst;
    reps[r].list = rl + (size_t)r * (M + 1);
  }
  nt = (int)(nrep < n_threads ? nrep : n_threads);
  if (nt < 1) nt = 1;
  waveD_job *dj = (waveD_job *)calloc((size_t)nt, sizeof *dj);
  for (int t = 0; t < nt; t++) {
    dj[t].g = g;
    dj[t].reps = reps;
    dj[t].r0 = nrep * t / nt;
    dj[t].r1 = nrep * (t + 1) / nt;
    if (nt == 1) waveD_worker(&dj[t]);
    else pthread_create(&th[t], NULL, waveD_worker, &dj[t]);
  }
  for (int t = 0; t < nt; t++) {
    if (nt > 1) pthread_join(th[t], NULL);
    ctr_add(&g->ctr, &dj[t].ctr);
  }
  free(dj);
  free(th);
  /* ---- phase E ---- */
  for (int64_t r = 0; r < nrep; r++) {
    int32_t *deg;
    int32_t *l = list_ptr(g, reps[r].layer, reps[r].row, &deg);
    memcpy(l, reps[r].list, sizeof(int32_t) * (size_t)reps[r].deg);
    *deg = reps[r].deg;
  }
  free(reps);
  free(rl);
  free(ev);
  free(slots);
  free(srows);
  free(sd);
  free(slot_off);
  return fail ? -1 : 0;
}

/* Batched Add in waves (see the banner above).  Returns the number of nodes inserted. */
int64_t ora_graph_add_waves(ora_graph *g, const int64_t *keys, const float *vecs, const int32_t *levels, int64_t n,
                            int ef, int64_t wave_max, int n_threads) {
  int64_t i = 0;
  if (g->n_dead) return 0; /* after a Delete, Add goes through the sequential path (ora_graph_add) */
  while (i < n) {
    int64_t w = ora_wave_len(g->n, g->n_layers, levels + i, n - i, wave_max);
    int32_t first_row = (int32_t)g->n;
    int top = g->n_layers;
    for (int64_t j = 0; j < w; j++) {
      if (levels[i + j] < 0 || levels[i + j] >= ORA_MAX_LAYERS) return i;
      new_row(g, keys[i + j], vecs + (size_t)(i + j) * g->dim, levels[i + j]);
      if (levels[i + j] + 1 > top) top = levels[i + j] + 1;
    }
    if (g->n_layers > 0 && wave_commit(g, first_row, w, ef, n_threads)) return i;
    for (int64_t j = 0; j < w; j++)
      for (int l = 0; l <= levels[i + j]; l++) {
        if (l >= g->n_layers || g->layer_size[l] == 0) g->entry[l] = first_row + (int32_t)j;
        g->layer_size[l]++;
      }
    if (top > g->n_layers) g->n_layers = top;
    i += w;
  }
  return n;
}
