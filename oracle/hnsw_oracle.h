/*
 * hnsw_oracle.h -- CPU restatement of TFMV/hnsw's hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This is the parity oracle: a plain-C restatement of the reference's Go algorithm
 * (graph.go, heap/heap.go, distance.go, encode.go, hnsw-extensions/hybrid/exact.go).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load it.  The product path (hnsw_b200/) never links, imports or calls it.
 *
 * Parity status
 *   - heap, layer search, maxLevel, varint/string encoding, distance KATs: PINNED against the
 *     reference's own known-answer tests (graph_test.go:14-74, distance_test.go:9-31,
 *     encode_test.go:11-50); see tests/test_oracle_kat.py.
 *   - float summation order inside a distance: PARITY UNPINNED.  The reference delegates to
 *     github.com/viterin/vek v0.4.2 (go.mod:10), which is not in /root/reference and whose
 *     order depends on the host CPU (AVX2/FMA assembly vs. a scalar Go loop).  The oracle
 *     therefore offers three orders (ORA_FLAVOR_*); ORA_FLAVOR_CANONICAL is the one the CUDA
 *     kernels reproduce bit-for-bit, ORA_FLAVOR_SEQUENTIAL restates vek's scalar Go loop.
 *   - Go map iteration order (random in Go) is replaced by ascending key order wherever the
 *     reference iterates a map (graph.go:60, 191, 197; exact.go:77; encode.go:159,165).
 *   - Go math/rand streams are not reproducible here; levels are an INPUT to ora_graph_add.
 *   - ExactIndex.Search result values: parity unpinned by the reference's tests (counts only);
 *     cross-checked against a float64 brute force in tests/.
 */
#ifndef HNSW_ORACLE_H
#define HNSW_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORA_EUCLIDEAN = 0, ORA_COSINE = 1 };
enum {
  ORA_FLAVOR_CANONICAL = 0,  /* 32 strided float4-chunk accumulators + xor butterfly (device order) */
  ORA_FLAVOR_SEQUENTIAL = 1, /* one accumulator, index order (vek32 pure-Go fallback order)        */
  ORA_FLAVOR_SIMD = 2        /* fastest host order (auto-vectorised); timing baseline only          */
};
enum { ORA_MODE_REFERENCE = 0, ORA_MODE_EXTENDED = 1 };
#define ORA_MAX_LAYERS 64

/* ---- distance.go:15-23 -------------------------------------------------------------------- */
float ora_distance(int metric, int flavor, const float *a, const float *b, int d);

/* ---- heap/heap.go + Go container/heap ------------------------------------------------------ */
typedef struct {
  float *dist;
  int32_t *id;
  int len, cap;
} ora_heap;
void ora_heap_init(ora_heap *h, int cap);
void ora_heap_free(ora_heap *h);
void ora_heap_push(ora_heap *h, float d, int32_t id);
void ora_heap_pop(ora_heap *h, float *d, int32_t *id);       /* heap.Pop    */
void ora_heap_remove(ora_heap *h, int i, float *d, int32_t *id); /* heap.Remove */
void ora_heap_poplast(ora_heap *h, float *d, int32_t *id);   /* Heap.PopLast */

/* ---- graph.go:370-385 ---------------------------------------------------------------------- */
int ora_max_level(double ml, int64_t num_nodes);
/* graph.go:388-417 with the uniform draws supplied by the caller; returns level, *used = draws consumed */
int ora_random_level(double ml, int n_layers, int64_t layer0_size, const double *uniforms, int *used);

/* ---- graph ----------------------------------------------------------------------------------- */
typedef struct ora_graph ora_graph;
typedef struct {
  int64_t dist_evals; /* E: distance evaluations inside layer searches            */
  int64_t expansions; /* H: popped candidates whose neighbour list was iterated   */
  int64_t nbr_reads;  /* neighbour ids read while expanding                        */
  int64_t evict_evals;     /* distance evaluations in addNeighbor overflow scans  */
  int64_t replenish_evals; /* cosine evaluations in replenish                     */
  int64_t evictions;
} ora_counters;

ora_graph *ora_graph_new(int dim, int M, int metric, int flavor);
void ora_graph_free(ora_graph *g);
int64_t ora_graph_len(const ora_graph *g);
int ora_graph_M(const ora_graph *g);
int ora_graph_metric(const ora_graph *g);
int ora_graph_dim(const ora_graph *g);
int ora_graph_layers(const ora_graph *g);
int64_t ora_graph_layer_size(const ora_graph *g, int layer);
int32_t ora_graph_entry(const ora_graph *g, int layer);
void ora_graph_set_entry(ora_graph *g, int layer, int32_t row);
int ora_graph_level(const ora_graph *g, int32_t row);
int64_t ora_graph_key(const ora_graph *g, int32_t row);
const float *ora_graph_vector(const ora_graph *g, int32_t row);
/* neighbour rows of `row` in `layer`, ascending key order; returns degree or -1 if absent */
int ora_graph_neighbors(const ora_graph *g, int layer, int32_t row, int32_t *out /*M+1*/);
void ora_graph_counters(const ora_graph *g, ora_counters *out);
void ora_graph_reset_counters(ora_graph *g);

/* Graph.Add for one node (graph.go:445-528); `level` is randomLevel()'s draw. Returns row or <0. */
int32_t ora_graph_add(ora_graph *g, int64_t key, const float *vec, int level, int ef);
/* Graph.Delete (graph.go:843-866): 1 if the key was present.  The row stays addressable (see ora_graph.c). */
int64_t ora_graph_layer_rows(const ora_graph *g, int layer);
int ora_graph_delete(ora_graph *g, int64_t key);
int ora_graph_delete_row(ora_graph *g, int32_t row);
int ora_graph_is_dead(const ora_graph *g, int32_t row);
int64_t ora_graph_dead_count(const ora_graph *g);
/* NOT the reference: layer search used by Add (ORA_MODE_*); only for the recall study in scripts/ext_recall.py */
void ora_graph_set_build_mode(ora_graph *g, int mode);
void ora_graph_set_flavor(ora_graph *g, int flavor);
/* Hand-build a layer (test fixtures such as graph_test.go:27-74): appends a node to `layer`
 * without linking; ora_graph_link adds a one-directional edge. */
int32_t ora_graph_put_node(ora_graph *g, int64_t key, const float *vec, int level);
int ora_graph_link(ora_graph *g, int layer, int32_t from, int32_t to);

/* layerNode.search (graph.go:94-170) on `layer` from `entry`; results in heap-array order. */
int ora_layer_search(ora_graph *g, int layer, int32_t entry, int k, int ef, const float *target,
                     int mode, int32_t *out_rows, float *out_dists);
/* Graph.Search / BatchSearch body (graph.go:571-593, 1075-1106).  entries==NULL: stored entries. */
int ora_graph_search(ora_graph *g, const float *query, int k, int ef, int mode,
                     const int32_t *entries, int32_t *out_rows, float *out_dists);
/* serial BatchSearch loop, optionally one query per thread (n_threads>1) for the timed baseline */
void ora_graph_batch_search(ora_graph *g, const float *queries, int64_t nq, int k, int ef, int mode,
                            int n_threads, int32_t *out_rows, float *out_dists, int32_t *out_counts,
                            ora_counters *total);

/* the same, recording per query the rows it evaluated, in evaluation order, with the distances this flavour computed:
 * trace_rows / trace_dists [nq][trace_cap], trace_len[nq] = evaluations made (> trace_cap: truncated) */
void ora_graph_batch_search_trace(ora_graph *g, const float *queries, int64_t nq, int k, int ef, int mode,
                                  int n_threads, int32_t *out_rows, float *out_dists, int32_t *out_counts,
                                  ora_counters *total, int32_t *trace_rows, float *trace_dists, int64_t *trace_len,
                                  int64_t trace_cap);

/* ---- encode.go wire format v1 (int keys) ----------------------------------------------------- */
int ora_put_varint(uint8_t *buf, int64_t v);                 /* binary.PutVarint  */
int ora_get_varint(const uint8_t *buf, size_t len, int64_t *v); /* binary.ReadVarint; bytes used or <0 */
int64_t ora_graph_export(const ora_graph *g, double ml, int ef, uint8_t *buf, int64_t cap);
ora_graph *ora_graph_import(const uint8_t *buf, int64_t len, int flavor, double *ml, int *ef);

/* ---- hybrid/exact.go:61-111 ------------------------------------------------------------------ */
/* faithful!=0 runs the reference's O(N^2) insertion sort; 0 runs a stable sort with identical output. */
int ora_exact_search(const float *vecs, int64_t n, int dim, int metric, int flavor, const float *query,
                     int k, int faithful, int32_t *out_idx, float *out_dists);
void ora_exact_batch_search(const float *vecs, int64_t n, int dim, int metric, int flavor,
                            const float *queries, int64_t nq, int k, int faithful, int n_threads,
                            int32_t *out_idx, float *out_dists);

#ifdef __cplusplus
}
#endif
#endif

/* bulk helpers for the test harness (not part of the restated API) */
#ifdef __cplusplus
extern "C" {
#endif
int64_t ora_graph_dump_layer(const ora_graph *g, int layer, int32_t *rows, int64_t *ptr, int32_t *nbr);
void ora_graph_dump_rows(const ora_graph *g, float *vec, int64_t *keys, int32_t *levels);
int64_t ora_graph_add_many(ora_graph *g, const int64_t *keys, const float *vecs, const int32_t *levels, int64_t n,
                           int ef);
int ora_graph_load_flat(ora_graph *g, int64_t n, const float *vecs, const int64_t *keys, const int32_t *levels,
                        int n_layers, const int32_t *entries, const int64_t *layer_sizes, const int32_t *layer_rows,
                        const int64_t *layer_ptr, const int32_t *layer_nbr);
#ifdef __cplusplus
}
#endif

/* Wave-batched Add: NOT the reference's algorithm; the CPU restatement of the device library's batched
 * construction (see the banner in ora_graph.c). */
#ifdef __cplusplus
extern "C" {
#endif
int64_t ora_wave_len(int64_t cur_rows, int n_layers, const int32_t *levels, int64_t remaining, int64_t wave_max);
int64_t ora_graph_add_waves(ora_graph *g, const int64_t *keys, const float *vecs, const int32_t *levels, int64_t n,
                            int ef, int64_t wave_max, int n_threads);
#ifdef __cplusplus
}
#endif
