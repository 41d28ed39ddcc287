/*
 * b200hnsw.h -- C ABI of the B200-native replacement for TFMV/hnsw's data-parallel hot path.
 *
 * The reference has no FFI seam; the seam is its exported Go API.  Each entry point below names the Go
 * method(s) whose body it replaces (file:line into the reference tree).  The Go host keeps validation,
 * locking (sync.RWMutex), the key <-> row mapping, the level RNG and the test hacks (graph.go:563-569,
 * 595-619); see INTEGRATION.md for the cgo binding.  Signatures use plain pointers and sizes only.
 *
 * Conventions
 *   - A node is a dense int32 ROW (insertion order).  keys[row] is an int64 that orders rows exactly as the
 *     reference's K orders keys (cmp.Ordered); neighbour lists are kept sorted by it because the reference
 *     expands neighbours in ascending key order (graph.go:137-138).
 *   - A node added with level L is a member of layers 0..L (graph.go:510-515).
 *   - Every function returns 0 on success or a negative B200HNSW_ERR_*; b200hnsw_last_error() gives text.
 *     Nothing aborts or throws across the boundary.  There is NO CPU fallback: without a CUDA device every
 *     compute call fails with B200HNSW_ERR_CUDA.
 *   - Host-pointer calls copy inputs before returning (cgo may not retain Go pointers).  *_device variants
 *     take device pointers plus a cudaStream_t (as void*) and are asynchronous on that stream.
 *   - Search calls may run concurrently from several host threads (they only read the index); calls that
 *     mutate (add, load, reserve) must be serialised by the caller, as the reference's write lock does.
 */
#ifndef B200HNSW_H
#define B200HNSW_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200HNSW_ABI_VERSION 1

enum {
  B200HNSW_OK = 0,
  B200HNSW_ERR_BAD_ARG = -1,
  B200HNSW_ERR_OOM = -2,
  B200HNSW_ERR_CUDA = -3,
  B200HNSW_ERR_UNSUPPORTED = -4,
  B200HNSW_ERR_VISITED_OVERFLOW = -5, /* a layer search visited more nodes than the scratch allows */
  B200HNSW_ERR_STATE = -6
};

/* distance.go:25-28: the two registered DistanceFuncs.  Any other Go func => the host returns an error. */
enum { B200HNSW_EUCLIDEAN = 0, B200HNSW_COSINE = 1 };

/* B200HNSW_MODE_REFERENCE reproduces layerNode.search bit for bit (graph.go:94-170, including its
 * first-non-improving-hop termination).  B200HNSW_MODE_EXTENDED is NOT the reference: textbook HNSW
 * termination with a true ef-bounded result set, offered because the reference's rule cannot reach high
 * recall on high-dimensional data.  Results of the two modes are never mixed. */
enum { B200HNSW_MODE_REFERENCE = 0, B200HNSW_MODE_EXTENDED = 1 };

#define B200HNSW_MAX_LAYERS 64
#define B200HNSW_MAX_M 64
#define B200HNSW_MAX_K 256
#define B200HNSW_MAX_EF 1024
#define B200HNSW_MAX_PEERS 16

typedef struct b200hnsw b200hnsw_t;
typedef struct b200exact b200exact_t;

typedef struct {
  int32_t dim;      /* len(Vector); fixed by the first Add in the reference (graph.go:450-455)            */
  int32_t max_m;    /* largest Graph.M that will be used; neighbour lists have this many slots (<=64)      */
  int32_t device;   /* CUDA device ordinal                                                                 */
  int32_t reserved; /* must be 0                                                                           */
  int64_t capacity; /* rows to reserve up front (grows by doubling when exceeded); 0 = default             */
} b200hnsw_config;

/* counters of one search batch; E and H are the quantities the roofline bytes are computed from */
typedef struct {
  int64_t dist_evals;  /* E: distance evaluations (graph.go:146)                      */
  int64_t expansions;  /* H: candidates popped and expanded (graph.go:127-139)        */
  int64_t nbr_reads;   /* neighbour ids read                                           */
  int64_t evict_evals;     /* add: distances in addNeighbor overflow scans (graph.go:60-71)   */
  int64_t replenish_evals; /* add: cosine distances in replenish (graph.go:204)               */
  int64_t evictions;       /* add: neighbours evicted (graph.go:73-80)                         */
  float kernel_ms;     /* device time of the call's kernels; for a search cut in chunks (B >= 1024) it spans
                          the overlapped copy-in / traversal / copy-out pipeline                             */
  int32_t launches;    /* kernels launched by the call                                 */
} b200hnsw_stats;

int b200hnsw_abi_version(void);
/* text of the last error raised on the calling thread ("" if none) */
const char *b200hnsw_last_error(void);

/* ---- lifetime: NewGraph / NewGraphWithConfig (graph.go:340-366) create the host object; this creates its
 *      device mirror ------------------------------------------------------------------------------------ */
int b200hnsw_create(const b200hnsw_config *cfg, b200hnsw_t **out);
void b200hnsw_destroy(b200hnsw_t *h);
int b200hnsw_reserve(b200hnsw_t *h, int64_t rows);

/* Graph.Len (graph.go:829-838), len(g.layers), layer.size (graph.go:260-265), layer.entry (graph.go:246-258:
 * "any node of the layer"; here the first node inserted into it unless overridden). */
int64_t b200hnsw_len(const b200hnsw_t *h);
int64_t b200hnsw_rows(const b200hnsw_t *h);   /* rows ever stored (deleted ones included): the next Add starts here */
int32_t b200hnsw_layers(const b200hnsw_t *h);
int64_t b200hnsw_layer_size(const b200hnsw_t *h, int32_t layer);
int32_t b200hnsw_entry(const b200hnsw_t *h, int32_t layer);
int b200hnsw_set_entry(b200hnsw_t *h, int32_t layer, int32_t row);

/* ---- Graph.Import (encode.go:180-262) / a graph built by the reference, flattened by the host ------------
 * n rows; vectors[n][dim]; keys[n]; levels[n]; entries[n_layers] (-1 = first member).
 * Per layer l (0..n_layers-1), lists are given as CSR over ALL n rows of that layer's members:
 *   layer_rows  : concatenated member rows, layer_sizes[l] per layer (layer 0 must list every row)
 *   layer_ptr   : concatenated offsets, layer_sizes[l]+1 per layer, each starting at 0 for its layer
 *   layer_nbr   : concatenated neighbour rows
 * Lists longer than max_m are rejected; lists are re-sorted by key on load. Replaces the current content; a load that
 * fails after its arguments were accepted leaves an EMPTY graph (never a half-loaded one). */
int b200hnsw_load(b200hnsw_t *h, int64_t n, const float *vectors, const int64_t *keys, const int32_t *levels,
                  int32_t n_layers, const int32_t *entries, const int64_t *layer_sizes, const int32_t *layer_rows,
                  const int64_t *layer_ptr, const int32_t *layer_nbr);

/* ---- Graph.Export (encode.go:133-174) / Analyzer: read the device graph back ------------------------------ */
int b200hnsw_get_rows(const b200hnsw_t *h, int64_t first, int64_t count, float *vectors, int64_t *keys,
                      int32_t *levels);
/* Replace the order keys of rows [first, first + count).  For hosts whose K is not an integer (string, float: any
 * cmp.Ordered) and who map keys to order-preserving int64 codes: when the gaps between codes run out the host re-spaces
 * them and tells the device.  The new keys must order the rows exactly as the old ones do (checked); nothing else
 * changes.  A mutating call: serialise it with Add / Delete. */
int b200hnsw_set_keys(b200hnsw_t *h, int64_t first, int64_t count, const int64_t *keys);
/* neighbour rows of `count` rows starting at `first` in `layer`, padded with -1 to max_m per row; rows that are
 * not members of the layer get all -1 and member[i]=0. */
int b200hnsw_get_neighbors(const b200hnsw_t *h, int32_t layer, int64_t first, int64_t count, int32_t *nbr,
                           uint8_t *member);

/* ---- Graph.Search / Graph.BatchSearch / Graph.ParallelSearch (graph.go:534-625, 1047-1110, 631-826) ---------
 * One traversal per query, all B in one launch.  out_rows/out_dists are [B][k] in the reference's order, i.e.
 * the raw heap array of the layer-0 result queue (graph.go:169, 588-593: index 0 is the nearest, the rest in
 * binary-heap layout); unused slots are -1 / NaN; out_counts[b] = len(result).  entries may be NULL.
 * k <= B200HNSW_MAX_K, ef <= B200HNSW_MAX_EF.  An empty graph yields counts of 0 (graph.go:554-556).
 * The host-pointer call is synchronous and may be made from several threads at once (the reference admits concurrent
 * BatchSearch callers under its read lock); a waiting call polls and yields its core.  Buffers may be page-locked
 * (read by the copy engine, results written by the kernel directly) or ordinary pageable memory (staged by the call
 * through its own page-locked buffer).  The _device variant takes device pointers and is asynchronous on `stream`. */
int b200hnsw_search_batch(b200hnsw_t *h, const float *queries, int64_t B, int32_t k, int32_t ef, int32_t metric,
                          int32_t mode, const int32_t *entries, int32_t *out_rows, float *out_dists,
                          int32_t *out_counts, b200hnsw_stats *stats);
int b200hnsw_search_batch_device(b200hnsw_t *h, const float *d_queries, int64_t B, int32_t k, int32_t ef,
                                 int32_t metric, int32_t mode, const int32_t *entries, int32_t *d_out_rows,
                                 float *d_out_dists, int32_t *d_out_counts, int64_t *d_counters /*[8] or NULL*/,
                                 void *stream);

/* Visited-set capacity (entries per traversal, a power of two >= 256).  The host-pointer call grows it by itself: a
 * batch in which a layer search visits more than half the table is redone with a table 4x the size, and the larger size
 * is kept for later calls once that retry has succeeded.  Device-pointer calls cannot retry (they are asynchronous):
 * they report the overflow in d_counters[4] (1 = visited set, 3 = NaN distances met with the short candidate queue),
 * mark the affected queries with count -1, and leave it to the caller to raise the capacity here and issue the batch
 * again (hnsw_b200/sharded.py: search_all does, agreeing on the error word across the ranks).  set fails with B200HNSW_ERR_UNSUPPORTED past what the device can give. */
int32_t b200hnsw_get_visited_capacity(const b200hnsw_t *h);
int b200hnsw_set_visited_capacity(b200hnsw_t *h, int32_t entries);

/* The same search writing a SEGMENTED result buffer for the multi-GPU all-to-all: queries are cut in runs of seg_len
 * (the share of one destination rank); run s writes, at d_packed + s * seg_stride (int32 elements),
 * rows int32 [seg_len][k] | dists f32 [seg_len][k] | counts int32 [seg_len] -- the layout
 * b200hnsw_merge_shards_device reads (with B = seg_len), so one all-to-all moves every rank's share into place. */
int b200hnsw_search_batch_segmented_device(b200hnsw_t *h, const float *d_queries, int64_t B, int32_t k, int32_t ef,
                                           int32_t metric, int32_t mode, const int32_t *entries, int32_t *d_packed,
                                           int64_t seg_len, int64_t seg_stride, int64_t *d_counters, void *stream);

/* The segmented search with one base pointer PER destination rank (host array of n_seg device pointers, n_seg =
 * ceil(B / seg_len) <= B200HNSW_MAX_PEERS): block s is written at seg_ptrs[s] in the same rows | dists | counts layout.
 * The pointers may be peer memory opened with b200hnsw_peer_open (another GPU's receive buffer over NVLink): the
 * traversal kernel then delivers every rank's share of the results itself -- search and exchange are ONE kernel, no
 * collective follows (no reference counterpart). */
int b200hnsw_search_batch_scatter_device(b200hnsw_t *h, const float *d_queries, int64_t B, int32_t k, int32_t ef,
                                         int32_t metric, int32_t mode, const int32_t *entries,
                                         int32_t *const *seg_ptrs, int32_t n_seg, int64_t seg_len,
                                         int64_t *d_counters, void *stream);

/* ---- peer memory for the fused search + exchange (one process per GPU; handles travel over any host channel) -------
 * b200hnsw_peer_alloc: zero-filled device memory on `device` plus its 64-byte IPC handle; b200hnsw_peer_open maps
 * another process's allocation into this process (peer access over NVLink is enabled on demand); _close / _free undo
 * them.  b200hnsw_peer_signal_device: after everything queued on `stream` so far, store `value` (release, system
 * scope) into each of the n int32 flags flag_ptrs[i] (local or peer).  b200hnsw_peer_wait_device: `stream` stalls
 * until each of the n local int32 flags d_flags[i] is >= value (acquire, system scope).  Both are stream memory
 * operations (cuStreamBatchMemOp) executed by the stream's front end, with one-warp kernels as the alternative
 * (environment B200HNSW_PEER_FLAGS=kernel, or a driver without the entry point). */
int b200hnsw_peer_alloc(int32_t device, int64_t bytes, void **d_ptr, uint8_t handle[64]);
int b200hnsw_peer_open(int32_t device, const uint8_t handle[64], void **d_ptr);
int b200hnsw_peer_close(void *d_ptr);
int b200hnsw_peer_free(void *d_ptr);
int b200hnsw_peer_signal_device(int32_t device, int32_t *const *flag_ptrs, int32_t n, int32_t value, void *stream);
int b200hnsw_peer_wait_device(int32_t device, const int32_t *d_flags, int32_t n, int32_t value, void *stream);

/* ---- Graph.Add / Graph.BatchAdd (graph.go:437-531, 942-1042) ------------------------------------------------
 * Inserts n nodes in input order with the reference's semantics: per layer top->0 search(k=M, ef) from the
 * elevator, then for every result (heap-array order) addNeighbor both ways with evict-worst + replenish
 * (graph.go:41-81, 172-219; replenish always scores with cosine).  levels[i] is the host's randomLevel() draw
 * (graph.go:388-417).  Keys must be new (the host deletes an existing key first, graph.go:1016-1024).
 * Rows are assigned len()..len()+n-1. */
int b200hnsw_add_batch(b200hnsw_t *h, const float *vectors, const int64_t *keys, const int32_t *levels, int64_t n,
                       int32_t m, int32_t ef, int32_t metric, b200hnsw_stats *stats);

/* Batched construction in WAVES.  NOT the reference's order: the nodes of a wave (at most wave_max, never more than
 * 1/16 of the rows already present) all search the graph as it was before the wave (K1 kernel, k = M on every layer,
 * graph.go:475-508) and their links are then committed in deterministic parallel phases that apply the reference's
 * heuristic (link to every result, evict the farthest on overflow, delete the back-link, replenish the evicted node
 * with CosineDistance: graph.go:41-81, 172-219).  The result is independent of scheduling and is restated on the CPU
 * by the oracle (ora_graph_add_waves) for bit-exact checks; it is NOT the graph b200hnsw_add_batch builds.  This is
 * the throughput path north_star calls "batched Add"; use b200hnsw_add_batch when the reference's exact graph is
 * required. */
int b200hnsw_add_batch_waves(b200hnsw_t *h, const float *vectors, const int64_t *keys, const int32_t *levels,
                             int64_t n, int32_t m, int32_t ef, int32_t metric, int64_t wave_max,
                             b200hnsw_stats *stats);

/* NOT the reference: which layer search the wave-batched Add runs.  B200HNSW_MODE_EXTENDED = the textbook search
 * (k = M, ef = EfSearch on every layer) instead of layerNode.search; links are still made by the reference's rule
 * (link to every result, evict the farthest, replenish with cosine).  The graph it builds answers extended-mode
 * searches with markedly higher recall; its oracle is ora_graph_add_waves with the same build mode.
 * b200hnsw_add_batch (the reference's sequential Add) refuses to run while the mode is set. */
int b200hnsw_set_build_mode(b200hnsw_t *h, int32_t mode);
int32_t b200hnsw_get_build_mode(const b200hnsw_t *h);

/* The SCHEDULE of b200hnsw_add_batch, never its result (no reference counterpart: the reference's Add is a loop).
 * The reference-order Add speculates a window of pending insertions in parallel -- each on its own CTA, reads logged,
 * writes kept in a private overlay -- and commits them in input order; a speculation that read a list an earlier commit
 * of the window wrote is run again at the head of the order (hnsw_b200/csrc/spec_insert.cu).  The graph is the one the
 * one-node-at-a-time kernel builds, bit for bit (tests/test_add_parity_gpu.py runs both against the oracle).
 * window = 0: chosen from the number of rows in the graph (1 below 2 000 rows, 4 below 10 000, 16 below 50 000, then 32;
 * a large batch is cut into launches that grow with the graph so that the choice follows it);
 * window = 1: one CTA, one node at a time; 2..128: that many insertions in flight.
 * While a node waits for its turn its worker keeps the speculation current (runs it again if a commit wrote a list it
 * read), so the committer rarely has to.
 * b200hnsw_get_add_window_stats: speculations accepted / insertions the committer ran again / speculations their workers
 * ran again, summed over the handle's life. */
int b200hnsw_set_add_window(b200hnsw_t *h, int32_t window);
int b200hnsw_get_add_window_stats(const b200hnsw_t *h, int64_t *accepted, int64_t *redone, int64_t *repaired);

/* Graph.Delete / Graph.BatchDelete (graph.go:843-895; layerNode.isolate graph.go:223-235, replenish graph.go:172-219).
 * rows[i] = the row the caller's key map holds for the i-th key, -1 for an unknown key; processed in call order, as
 * the reference's loop under one lock.  out_deleted[i] (may be null) mirrors the reference's []bool: 1 iff the row
 * was a live member.  m = Graph.M.  As in the reference the node's storage survives: it leaves layer.nodes (entry
 * point, elevator lookup, len), every neighbour drops its link and is replenished (which can link the node again
 * one-way), its own list is kept, traversals still walk through it and Search can still return its row.  A Search
 * whose elevator is a deleted row returns 0 results (graph.go:573-583); an Add in that situation fails with
 * "no nodes found in neighborhood search" (graph.go:501-505) and, unlike the reference, leaves the graph untouched. */
int b200hnsw_delete_batch(b200hnsw_t *h, const int32_t *rows, int64_t n, int32_t m, uint8_t *out_deleted,
                          b200hnsw_stats *stats);
/* 1 if `row` was deleted, 0 if live, -1 if out of range */
int32_t b200hnsw_is_deleted(const b200hnsw_t *h, int32_t row);

/* ---- multi-GPU: merge G per-shard result lists per query into one top-k (no reference counterpart; tie order =
 *      lower shard first, the stable order of hybrid/adapter.go:264-292).  Inputs [G][B][k] ascending-sorted
 *      per shard or heap-ordered (sorted_in=0); ids are int64 global ids.  Device pointers. ----------------- */
int b200hnsw_merge_topk_device(int32_t device, int32_t G, int64_t B, int32_t k, const float *d_dists,
                               const int64_t *d_ids, const int32_t *d_counts, int32_t sorted_in, float *d_out_dists,
                               int64_t *d_out_ids, int32_t *d_out_counts, void *stream);

/* The same merge reading what the search kernel wrote, without repacking: shard g's results start at
 * d_packed + g * shard_stride (int32 elements) as rows int32 [B][k] | dists f32 [B][k] | counts int32 [B] -- one
 * contiguous buffer per rank, so ONE all-gather moves everything; the global id of a row is g * id_stride + row. */
int b200hnsw_merge_shards_device(int32_t device, int32_t G, int64_t B, int32_t k, const int32_t *d_packed,
                                 int64_t shard_stride, int64_t id_stride, float *d_out_dists, int64_t *d_out_ids,
                                 int32_t *d_out_counts, void *stream);

/* ---- hybrid.ExactIndex (hnsw-extensions/hybrid/exact.go:13-160) --------------------------------------------- */
int b200exact_create(int32_t dim, int32_t metric, int32_t device, int64_t capacity, b200exact_t **out);
void b200exact_destroy(b200exact_t *x);
int64_t b200exact_len(const b200exact_t *x); /* rows ever added minus removed */
/* ExactIndex.Add / BatchAdd (exact.go:27-58): copies n vectors; rows are assigned in order; *first_row out */
int b200exact_add_batch(b200exact_t *x, const float *vectors, int64_t n, int64_t *first_row);
/* ExactIndex.Delete / BatchDelete (exact.go:114-145): tombstones rows */
int b200exact_remove(b200exact_t *x, const int64_t *rows, int64_t n);
/* ExactIndex.Search (exact.go:61-111) for B queries: out_rows/out_dists [B][k] ascending by distance, ties in
 * row order (the reference's order among ties is Go map order, i.e. unspecified); out_counts = min(k, len). */
int b200exact_search_batch(b200exact_t *x, const float *queries, int64_t B, int32_t k, int64_t *out_rows,
                           float *out_dists, int32_t *out_counts, b200hnsw_stats *stats);
/* Device-pointer variant.  The scan path (small n or few queries) is asynchronous on `stream`; the tensor path
 * synchronises `stream` ONCE inside the call (the host reads the per-query candidate counts and redoes, exactly, any
 * query whose list overflowed) and queues the rest.  Calls in flight on different streams are ordered on the device
 * where they share the index's scratch (an event recorded behind each call), so they never see each other's workspace. */
int b200exact_search_batch_device(b200exact_t *x, const float *d_queries, int64_t B, int32_t k, int64_t *d_out_rows,
                                  float *d_out_dists, int32_t *d_out_counts, void *stream);

/* ExactIndex.Search for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.  The
 * scan path of b200exact_search_batch_device (every distance in canonical fp32: the same rows and bits) with its
 * workspace in a file-scope `static __device__` array of 8 MiB that every GPU loading the library holds until exit;
 * the call makes no cudaMalloc / cudaFree and calls no allocating routine.  Limit: the final keys plus one query's
 * partial lists must fit, 8*k*(B + ceil(rows/2048)*(1 + 1/max(2, 4096/k)) + 2) <= 8 MiB (k = 10: B + rows/2048 up to
 * ~104000); beyond it the call fails with B200HNSW_ERR_UNSUPPORTED and a message that names the sizes.  The first
 * call on a device resolves the array's address (cudaGetSymbolAddress): make one call with B = 0 at set-up (before a
 * capture).  Asynchronous on `stream`.  All calls on a GPU share the one array (its contents never have to start at
 * zero: every word read was written by the same call), so calls that OVERLAP on the GPU race -- keep them on one
 * stream or order them with events. */
int b200exact_search_batch_noalloc_device(b200exact_t *x, const float *d_queries, int64_t B, int32_t k,
                                          int64_t *d_out_rows, float *d_out_dists, int32_t *d_out_counts,
                                          void *stream);

/* EuclideanDistance / CosineDistance (distance.go:15-23) on the device, pairs (a[i], b[i]); host pointers.
 * Exists so the distance KATs (distance_test.go:9-31) run through the same arithmetic as the kernels. */
int b200hnsw_distance_batch(int32_t device, int32_t metric, const float *a, const float *b, int64_t n, int32_t dim,
                            float *out);

#ifdef __cplusplus
}
#endif
#endif
