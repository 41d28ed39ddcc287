// Package hybrid: ExactIndex over the device (reference hnsw-extensions/hybrid/exact.go:13-160).  NOT compiled in this
// repository's image; the tested forms are hybrid::ExactIndex in include/b200hnsw_host.hpp and hnsw_b200/exact.py.
package hybrid

/*
#cgo CFLAGS:  -I${SRCDIR}/../../../include
#cgo LDFLAGS: -L${SRCDIR}/../../../hnsw_b200/lib -lb200hnsw -lcudart
#include "b200hnsw.h"
*/
import "C"

import (
	"cmp"
	"errors"
	"fmt"
	"sync"
	"unsafe"

	hnsw "github.com/TFMV/hnsw"
)

// ExactIndex stores vectors on the device and answers Search by brute force (tcgen05 distance GEMM + exact re-rank).
type ExactIndex[K cmp.Ordered] struct {
	mu     sync.RWMutex
	metric int
	x      *C.b200exact_t
	dim    int
	rowOf  map[K]int64
	keyOf  []K
	vecs   [][]float32
}

func lastErr(rc C.int) error {
	if rc == 0 {
		return nil
	}
	return errors.New("b200hnsw: " + C.GoString(C.b200hnsw_last_error()))
}

// NewExactIndex creates an exact index (reference exact.go:20-25).  distance must be hnsw.EuclideanDistance or
// hnsw.CosineDistance: only those have device kernels, and there is no CPU fallback.
func NewExactIndex[K cmp.Ordered](distance hnsw.DistanceFunc) (*ExactIndex[K], error) {
	m, err := hnsw.MetricOf(distance)
	if err != nil {
		return nil, err
	}
	return &ExactIndex[K]{metric: m, rowOf: make(map[K]int64)}, nil
}

// Add stores a copy of the vector (reference exact.go:27-38).
func (idx *ExactIndex[K]) Add(key K, vector []float32) error {
	return idx.BatchAdd([]K{key}, [][]float32{vector})
}

// BatchAdd (reference exact.go:41-58).  Assigning an existing key overwrites it: the old row is tombstoned.
func (idx *ExactIndex[K]) BatchAdd(keys []K, vectors [][]float32) error {
	if len(keys) != len(vectors) {
		return fmt.Errorf("number of keys (%d) does not match number of vectors (%d)", len(keys), len(vectors))
	}
	if len(keys) == 0 {
		return nil
	}
	idx.mu.Lock()
	defer idx.mu.Unlock()
	if idx.x == nil {
		idx.dim = len(vectors[0])
		if err := lastErr(C.b200exact_create(C.int32_t(idx.dim), C.int32_t(idx.metric), 0, 0, &idx.x)); err != nil {
			return err
		}
	}
	last := make(map[K]int, len(keys))
	for i, k := range keys {
		if len(vectors[i]) != idx.dim {
			return fmt.Errorf("embedding dimension mismatch: %d != %d", idx.dim, len(vectors[i]))
		}
		last[k] = i
	}
	var keep []int
	var stale []int64
	for i, k := range keys {
		if last[k] != i {
			continue
		}
		keep = append(keep, i)
		if r, ok := idx.rowOf[k]; ok {
			stale = append(stale, r)
		}
	}
	if len(stale) > 0 {
		rc := C.b200exact_remove(idx.x, (*C.int64_t)(unsafe.Pointer(&stale[0])), C.int64_t(len(stale)))
		if err := lastErr(rc); err != nil {
			return err
		}
	}
	flat := make([]float32, 0, len(keep)*idx.dim)
	for _, i := range keep {
		flat = append(flat, vectors[i]...)
	}
	var first C.int64_t
	rc := C.b200exact_add_batch(idx.x, (*C.float)(unsafe.Pointer(&flat[0])), C.int64_t(len(keep)), &first)
	if err := lastErr(rc); err != nil {
		return err
	}
	for j, i := range keep {
		idx.rowOf[keys[i]] = int64(first) + int64(j)
		idx.keyOf = append(idx.keyOf, keys[i])
		idx.vecs = append(idx.vecs, append([]float32(nil), vectors[i]...))
	}
	return nil
}

// Search returns the k nearest nodes, ascending by distance (reference exact.go:61-111).
func (idx *ExactIndex[K]) Search(query []float32, k int) ([]hnsw.Node[K], error) {
	idx.mu.RLock()
	defer idx.mu.RUnlock()
	if idx.x == nil || k <= 0 || int64(C.b200exact_len(idx.x)) == 0 {
		return []hnsw.Node[K]{}, nil
	}
	if len(query) != idx.dim {
		return nil, fmt.Errorf("embedding dimension mismatch: %d != %d", idx.dim, len(query))
	}
	rows := make([]int64, k)
	dists := make([]float32, k)
	var count C.int32_t
	rc := C.b200exact_search_batch(idx.x, (*C.float)(unsafe.Pointer(&query[0])), 1, C.int32_t(k),
		(*C.int64_t)(unsafe.Pointer(&rows[0])), (*C.float)(unsafe.Pointer(&dists[0])), &count, nil)
	if err := lastErr(rc); err != nil {
		return nil, err
	}
	out := make([]hnsw.Node[K], 0, int(count))
	for _, r := range rows[:int(count)] {
		out = append(out, hnsw.Node[K]{Key: idx.keyOf[r], Value: idx.vecs[r]})
	}
	return out, nil
}

// Delete removes a key (reference exact.go:114-126).
func (idx *ExactIndex[K]) Delete(key K) bool {
	idx.mu.Lock()
	defer idx.mu.Unlock()
	r, ok := idx.rowOf[key]
	if !ok {
		return false
	}
	delete(idx.rowOf, key)
	C.b200exact_remove(idx.x, (*C.int64_t)(unsafe.Pointer(&r)), 1)
	return true
}

// Len returns the number of stored vectors.
func (idx *ExactIndex[K]) Len() int {
	idx.mu.RLock()
	defer idx.mu.RUnlock()
	if idx.x == nil {
		return 0
	}
	return int(C.b200exact_len(idx.x))
}

// Close releases the device memory.
func (idx *ExactIndex[K]) Close() {
	idx.mu.Lock()
	defer idx.mu.Unlock()
	if idx.x != nil {
		C.b200exact_destroy(idx.x)
		idx.x = nil
	}
}
