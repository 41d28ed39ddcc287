// Package hnsw: cgo binding to libb200hnsw.so (include/b200hnsw.h).  NOT compiled in this repository's image (no Go
// toolchain there); see README.md in this directory.
package hnsw

/*
#cgo CFLAGS:  -I${SRCDIR}/../../include
#cgo LDFLAGS: -L${SRCDIR}/../../hnsw_b200/lib -lb200hnsw -lcudart
#include <stdlib.h>
#include "b200hnsw.h"
*/
import "C"

import (
	"errors"
	"unsafe"
)

// device is the mirror of one Graph in HBM.
type device struct {
	h    *C.b200hnsw_t
	dim  int
	maxM int
}

func lastErr(rc C.int) error {
	if rc == 0 {
		return nil
	}
	return errors.New("b200hnsw: " + C.GoString(C.b200hnsw_last_error()))
}

func newDevice(dim, maxM, dev int, capacity int64) (*device, error) {
	if maxM > C.B200HNSW_MAX_M {
		maxM = C.B200HNSW_MAX_M
	}
	cfg := C.b200hnsw_config{dim: C.int32_t(dim), max_m: C.int32_t(maxM), device: C.int32_t(dev),
		capacity: C.int64_t(capacity)}
	var h *C.b200hnsw_t
	if err := lastErr(C.b200hnsw_create(&cfg, &h)); err != nil {
		return nil, err
	}
	return &device{h: h, dim: dim, maxM: maxM}, nil
}

func (d *device) close() {
	if d != nil && d.h != nil {
		C.b200hnsw_destroy(d.h)
		d.h = nil
	}
}

func (d *device) len() int      { return int(C.b200hnsw_len(d.h)) }
func (d *device) rows() int64   { return int64(C.b200hnsw_rows(d.h)) }
func (d *device) layers() int   { return int(C.b200hnsw_layers(d.h)) }
func (d *device) layerSize(l int) int { return int(C.b200hnsw_layer_size(d.h, C.int32_t(l))) }

// searchBatch: Graph.Search / BatchSearch (graph.go:534-625, 1047-1110).  flat = queries laid out [B][dim].
// rows / dists come back [B][k] in the reference's heap-array order; counts[b] = len(result).
func (d *device) searchBatch(flat []float32, B, k, ef, metric int) (rows []int32, dists []float32, counts []int32, err error) {
	rows = make([]int32, B*k)
	dists = make([]float32, B*k)
	counts = make([]int32, B)
	rc := C.b200hnsw_search_batch(d.h, (*C.float)(unsafe.Pointer(&flat[0])), C.int64_t(B), C.int32_t(k),
		C.int32_t(ef), C.int32_t(metric), C.int32_t(C.B200HNSW_MODE_REFERENCE), nil,
		(*C.int32_t)(unsafe.Pointer(&rows[0])), (*C.float)(unsafe.Pointer(&dists[0])),
		(*C.int32_t)(unsafe.Pointer(&counts[0])), nil)
	return rows, dists, counts, lastErr(rc)
}

// addBatch: Graph.Add / BatchAdd (graph.go:437-531, 942-1042).  levels[i] = randomLevel() drawn in Go, in order.
// waveMax == 0: the reference's exact sequential order; > 0: the wave-batched construction.
func (d *device) addBatch(flat []float32, keys []int64, levels []int32, m, ef, metric int, waveMax int64) error {
	n := C.int64_t(len(keys))
	v := (*C.float)(unsafe.Pointer(&flat[0]))
	k := (*C.int64_t)(unsafe.Pointer(&keys[0]))
	l := (*C.int32_t)(unsafe.Pointer(&levels[0]))
	if waveMax > 0 {
		return lastErr(C.b200hnsw_add_batch_waves(d.h, v, k, l, n, C.int32_t(m), C.int32_t(ef), C.int32_t(metric),
			C.int64_t(waveMax), nil))
	}
	return lastErr(C.b200hnsw_add_batch(d.h, v, k, l, n, C.int32_t(m), C.int32_t(ef), C.int32_t(metric), nil))
}

// deleteBatch: Graph.Delete / BatchDelete (graph.go:843-895).  rows[i] = -1 for an unknown key.
func (d *device) deleteBatch(rows []int32, m int) ([]bool, error) {
	ok := make([]C.uint8_t, len(rows))
	rc := C.b200hnsw_delete_batch(d.h, (*C.int32_t)(unsafe.Pointer(&rows[0])), C.int64_t(len(rows)), C.int32_t(m),
		&ok[0], nil)
	out := make([]bool, len(rows))
	for i := range ok {
		out[i] = ok[i] != 0
	}
	return out, lastErr(rc)
}

// setKeys: new order codes for rows [0, len(codes)) after the host re-spaced them (keys.go).
func (d *device) setKeys(codes []int64) error {
	return lastErr(C.b200hnsw_set_keys(d.h, 0, C.int64_t(len(codes)), (*C.int64_t)(unsafe.Pointer(&codes[0]))))
}

// deviceDistance: one pair through the kernels' arithmetic (b200hnsw_distance_batch).
func deviceDistance(metric int, a, b []float32) (float32, error) {
	var out C.float
	rc := C.b200hnsw_distance_batch(0, C.int32_t(metric), (*C.float)(unsafe.Pointer(&a[0])),
		(*C.float)(unsafe.Pointer(&b[0])), 1, C.int32_t(len(a)), &out)
	return float32(out), lastErr(rc)
}
