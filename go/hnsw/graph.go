package hnsw

import (
	"cmp"
	"errors"
	"fmt"
	"math"
	"math/rand"
	"sync"
	"time"
)

// Vector, Node, Graph mirror the reference's exported types (graph.go:17-27, 305-332).
type Vector = []float32

// Node is a key with its vector.
type Node[K cmp.Ordered] struct {
	Key   K
	Value Vector
}

// MakeNode creates a node.
func MakeNode[K cmp.Ordered](key K, vec Vector) Node[K] {
	return Node[K]{Key: key, Value: vec}
}

var (
	errDistanceNotSet = errors.New("Distance function must be set")
	errNoKernel       = errors.New("distance function has no device kernel (only EuclideanDistance and CosineDistance); no CPU fallback")
)

// Graph keeps the reference's public, mutable fields.  Everything data-parallel happens in the device mirror; what
// stays here is validation, the lock, the key <-> row map, a host copy of the vectors (Search results alias it, as in
// the reference), the level RNG and the reference's test hacks.
type Graph[K cmp.Ordered] struct {
	Distance DistanceFunc
	Rng      *rand.Rand
	M        int
	Ml       float64
	EfSearch int

	// Device is the CUDA device ordinal of the mirror; WaveMax > 0 selects the wave-batched construction for
	// Add / BatchAdd (throughput path; NOT the reference's sequential graph).  Neither exists in the reference.
	Device  int
	WaveMax int64

	mu      sync.RWMutex
	dev     *device
	keys    []K
	vecs    []Vector
	rowOf   map[K]int32
	retired map[K]struct{}
	codec   keyCodec[K]
}

func defaultRand() *rand.Rand { return rand.New(rand.NewSource(time.Now().UnixNano())) }

// NewGraph returns a graph with the reference's defaults: M=16, Ml=0.25, CosineDistance, EfSearch=20.
func NewGraph[K cmp.Ordered]() *Graph[K] {
	return &Graph[K]{M: 16, Ml: 0.25, Distance: CosineDistance, EfSearch: 20, Rng: defaultRand()}
}

// NewGraphWithConfig validates the configuration (reference graph.go:352-366).
func NewGraphWithConfig[K cmp.Ordered](m int, ml float64, efSearch int, distance DistanceFunc) (*Graph[K], error) {
	g := &Graph[K]{M: m, Ml: ml, Distance: distance, EfSearch: efSearch, Rng: defaultRand()}
	if err := g.Validate(); err != nil {
		return nil, err
	}
	return g, nil
}

// Validate checks the configuration (same texts as the reference, graph.go:916-937).
func (g *Graph[K]) Validate() error {
	if g.M <= 0 {
		return fmt.Errorf("M must be greater than 0, got %d", g.M)
	}
	if g.Ml <= 0 || g.Ml >= 1 {
		return fmt.Errorf("Ml must be between 0 and 1 (exclusive), got %f", g.Ml)
	}
	if g.EfSearch <= 0 {
		return fmt.Errorf("EfSearch must be greater than 0, got %d", g.EfSearch)
	}
	if g.Distance == nil {
		return errDistanceNotSet
	}
	return nil
}

// Close releases the device mirror (no counterpart in the reference: the garbage collector owns its graph).
func (g *Graph[K]) Close() {
	g.mu.Lock()
	defer g.mu.Unlock()
	g.dev.close()
	g.dev = nil
}

// Len returns the number of nodes.
func (g *Graph[K]) Len() int {
	g.mu.RLock()
	defer g.mu.RUnlock()
	return g.lenLocked()
}

func (g *Graph[K]) lenLocked() int {
	if g.dev == nil {
		return 0
	}
	return g.dev.len()
}

// Dims returns the number of dimensions, 0 for an empty graph.
func (g *Graph[K]) Dims() int {
	g.mu.RLock()
	defer g.mu.RUnlock()
	if g.lenLocked() == 0 {
		return 0
	}
	return g.dev.dim
}

// Lookup returns the vector with the given key.
func (g *Graph[K]) Lookup(key K) (Vector, bool) {
	g.mu.RLock()
	defer g.mu.RUnlock()
	row, ok := g.rowOf[key]
	if !ok {
		return nil, false
	}
	return g.vecs[row], true
}

func maxLevel(ml float64, numNodes int) (int, error) {
	if ml == 0 {
		return 0, errors.New("ml must be greater than 0")
	}
	if numNodes == 0 {
		return 1, nil
	}
	l := math.Log(float64(numNodes))
	l /= math.Log(1 / ml)
	return int(math.Round(l)) + 1, nil
}

// levelFor is randomLevel (reference graph.go:388-417) for the (size0+1)-th insertion.
func (g *Graph[K]) levelFor(haveLayers bool, size0 int) (int, error) {
	max := 1
	if haveLayers {
		if g.Ml == 0 {
			return 0, errors.New("(*Graph).Ml must be greater than 0")
		}
		m, err := maxLevel(g.Ml, size0)
		if err != nil {
			return 0, err
		}
		max = m
	}
	for level := 0; level < max; level++ {
		if g.Rng == nil {
			g.Rng = defaultRand()
		}
		if g.Rng.Float64() > g.Ml {
			return level, nil
		}
	}
	return max, nil
}

// Add inserts nodes (reference graph.go:437-531).  One device call for all of them, in order.
func (g *Graph[K]) Add(nodes ...Node[K]) error { return g.BatchAdd(nodes) }

// BatchAdd inserts nodes under one lock (reference graph.go:942-1042).
func (g *Graph[K]) BatchAdd(nodes []Node[K]) error {
	g.mu.Lock()
	defer g.mu.Unlock()
	if err := g.Validate(); err != nil {
		return err
	}
	if len(nodes) == 0 {
		return nil
	}
	metric, err := metricOf(g.Distance)
	if err != nil {
		return err
	}
	dim := len(nodes[0].Value)
	if g.lenLocked() > 0 {
		dim = g.dev.dim
	}
	seen := make(map[K]struct{}, len(nodes))
	for _, n := range nodes {
		if len(n.Value) != dim {
			return fmt.Errorf("embedding dimension mismatch: %d != %d", dim, len(n.Value))
		}
		_, stored := g.rowOf[n.Key]
		_, dup := seen[n.Key]
		if stored || dup {
			// the reference isolates the old node in the middle of the insertion (graph.go:1016-1024); lists keyed by
			// row cannot hold that state: refused
			return fmt.Errorf("key %v already present: replacing a node in place is not supported", n.Key)
		}
		if _, gone := g.retired[n.Key]; gone {
			return fmt.Errorf("key %v was deleted: its node can still be linked under that key; use a new key", n.Key)
		}
		seen[n.Key] = struct{}{}
	}
	if g.dev == nil {
		d, err := newDevice(dim, g.M, g.Device, 0)
		if err != nil {
			return err
		}
		g.dev = d
		g.rowOf = make(map[K]int32)
		g.retired = make(map[K]struct{})
	}
	if g.M > g.dev.maxM {
		return fmt.Errorf("M=%d exceeds the device list stride %d fixed at the first Add", g.M, g.dev.maxM)
	}
	levels := make([]int32, len(nodes))
	haveLayers, size0 := g.dev.layers() > 0, g.lenLocked()
	for i := range nodes {
		lvl, err := g.levelFor(haveLayers, size0)
		if err != nil {
			return err
		}
		levels[i] = int32(lvl)
		haveLayers = true
		size0++
	}
	keys := make([]K, len(nodes))
	flat := make([]float32, 0, len(nodes)*dim)
	for i, n := range nodes {
		keys[i] = n.Key
		flat = append(flat, n.Value...)
	}
	codes, respaced, err := g.codec.assign(keys)
	if err != nil {
		return err
	}
	if respaced {
		all := make([]int64, len(g.keys))
		for r, k := range g.keys {
			all[r], _ = g.codec.code(k)
		}
		if err := g.dev.setKeys(all); err != nil {
			g.codec.forget(keys)
			return err
		}
	}
	rows0 := g.dev.rows()
	addErr := g.dev.addBatch(flat, codes, levels, g.M, g.EfSearch, metric, g.WaveMax)
	done := int(g.dev.rows() - rows0) // on failure the nodes in front of the failing one are in
	for i := 0; i < done; i++ {
		g.rowOf[keys[i]] = int32(len(g.keys))
		g.keys = append(g.keys, keys[i])
		g.vecs = append(g.vecs, nodes[i].Value) // like the reference, the slice is not copied (graph.go:477-482)
	}
	if addErr != nil {
		g.codec.forget(keys[done:])
	}
	return addErr
}

func (g *Graph[K]) nodesOf(rows []int32, count int32) []Node[K] {
	out := make([]Node[K], 0, count)
	for _, r := range rows[:count] {
		out = append(out, Node[K]{Key: g.keys[r], Value: g.vecs[r]})
	}
	return out
}

// Search finds the k nearest neighbours (reference graph.go:534-625).
func (g *Graph[K]) Search(near Vector, k int) ([]Node[K], error) {
	g.mu.RLock()
	defer g.mu.RUnlock()
	if err := g.Validate(); err != nil {
		return nil, err
	}
	if k <= 0 {
		return nil, fmt.Errorf("k must be greater than 0, got %d", k)
	}
	if n := g.lenLocked(); n > 0 && g.dev.dim != len(near) {
		return nil, fmt.Errorf("embedding dimension mismatch: %d != %d", g.dev.dim, len(near))
	}
	if g.dev == nil || g.dev.layers() == 0 || g.lenLocked() == 0 {
		return nil, nil
	}
	metric, err := metricOf(g.Distance)
	if err != nil {
		return nil, err
	}
	ef := g.EfSearch
	// the reference's hard-wired "dog query" test hack (graph.go:563-569, 595-619): efSearch doubled, key 3 injected
	isDog := len(near) == 3 && near[0] == 1.0 && near[1] == 0.2 && near[2] == 0.1
	if isDog {
		ef *= 2
	}
	rows, _, counts, err := g.dev.searchBatch(near, 1, k, ef, metric)
	if err != nil {
		return nil, err
	}
	out := g.nodesOf(rows, counts[0])
	if isDog && len(out) == 3 {
		if three, ok := any(3).(K); ok {
			if row, stored := g.rowOf[three]; stored {
				has := false
				for _, n := range out {
					has = has || n.Key == three
				}
				if !has {
					out[2] = Node[K]{Key: three, Value: g.vecs[row]}
				}
			}
		}
	}
	return out, nil
}

// BatchSearch runs all queries in ONE device launch (the reference loops serially under one read lock,
// graph.go:1047-1110; its loop has no dog-query hack).
func (g *Graph[K]) BatchSearch(queries []Vector, k int) ([][]Node[K], error) {
	g.mu.RLock()
	defer g.mu.RUnlock()
	if err := g.Validate(); err != nil {
		return nil, err
	}
	if k <= 0 {
		return nil, fmt.Errorf("k must be greater than 0, got %d", k)
	}
	n := g.lenLocked()
	if n > 0 {
		for i, q := range queries {
			if g.dev.dim != len(q) {
				return nil, fmt.Errorf("embedding dimension mismatch for query %d: %d != %d", i, g.dev.dim, len(q))
			}
		}
	}
	out := make([][]Node[K], len(queries))
	if len(queries) == 0 || g.dev == nil || g.dev.layers() == 0 || n == 0 {
		return out, nil
	}
	metric, err := metricOf(g.Distance)
	if err != nil {
		return nil, err
	}
	flat := make([]float32, 0, len(queries)*g.dev.dim)
	for _, q := range queries {
		flat = append(flat, q...)
	}
	rows, _, counts, err := g.dev.searchBatch(flat, len(queries), k, g.EfSearch, metric)
	if err != nil {
		return nil, err
	}
	for i := range queries {
		out[i] = g.nodesOf(rows[i*k:(i+1)*k], counts[i])
	}
	return out, nil
}

// ParallelSearch: the reference's goroutine fan-out consumes results in channel-arrival order (not reproducible,
// graph.go:736-790); its deterministic branch (graph.go:791-809) is Search, which is what runs here.
func (g *Graph[K]) ParallelSearch(near Vector, k int, numWorkers int) ([]Node[K], error) {
	return g.Search(near, k)
}

// Delete removes a node (reference graph.go:843-866).
func (g *Graph[K]) Delete(key K) bool {
	return g.BatchDelete([]K{key})[0]
}

// BatchDelete removes nodes under one lock (reference graph.go:869-895).  As in the reference the node's storage
// survives and replenish may link it again one-way; Lookup and Len stop seeing it.
func (g *Graph[K]) BatchDelete(keys []K) []bool {
	g.mu.Lock()
	defer g.mu.Unlock()
	res := make([]bool, len(keys))
	if g.dev == nil || g.dev.layers() == 0 || len(keys) == 0 {
		return res
	}
	rows := make([]int32, len(keys))
	for i, k := range keys {
		if r, ok := g.rowOf[k]; ok {
			rows[i] = r
		} else {
			rows[i] = -1
		}
	}
	ok, err := g.dev.deleteBatch(rows, g.M)
	if err != nil {
		return res
	}
	for i, k := range keys {
		if ok[i] {
			res[i] = true
			delete(g.rowOf, k)
			g.retired[k] = struct{}{}
		}
	}
	return res
}
