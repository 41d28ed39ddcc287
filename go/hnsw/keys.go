package hnsw

import (
	"cmp"
	"fmt"
	"math"
	"reflect"
	"sort"
)

// The device keeps neighbour lists sorted by an int64 per row, because the reference expands neighbours in ascending
// key order (graph.go:137-138).  keyCodec maps K (any cmp.Ordered) to int64 CODES that order exactly as the keys do:
// integers map to themselves, floats to the monotone image of their IEEE-754 bits, every other ordered kind (strings)
// to gap-spaced codes over the sorted keys: a new key takes the midpoint between its neighbours' codes, and when a gap
// is used up all codes are re-spaced and the device is told (b200hnsw_set_keys).  Same logic as hnsw_b200/keys.py,
// which is the tested form.
const codeGap = int64(1) << 32

type keyCodec[K cmp.Ordered] struct {
	sorted []K     // table kinds only: keys ascending ...
	codes  []int64 // ... and their codes, parallel
}

// directCode returns the code of kinds that need no table.
func directCode[K cmp.Ordered](k K) (int64, bool, error) {
	v := reflect.ValueOf(k)
	switch v.Kind() {
	case reflect.Int, reflect.Int8, reflect.Int16, reflect.Int32, reflect.Int64:
		return v.Int(), true, nil
	case reflect.Uint, reflect.Uint8, reflect.Uint16, reflect.Uint32, reflect.Uint64, reflect.Uintptr:
		return int64(v.Uint() ^ (1 << 63)), true, nil // order-preserving shift of the unsigned range
	case reflect.Float32, reflect.Float64:
		f := v.Float()
		if math.IsNaN(f) {
			return 0, true, fmt.Errorf("NaN cannot be a key")
		}
		if f == 0 {
			f = 0 // -0 == 0 as a map key
		}
		u := int64(math.Float64bits(f))
		if u < 0 {
			u = math.MinInt64 - u - 1 // negative floats: reverse their order, below every positive one
		}
		return u, true, nil
	}
	return 0, false, nil
}

func (c *keyCodec[K]) find(k K) (int, bool) {
	i := sort.Search(len(c.sorted), func(i int) bool { return !cmp.Less(c.sorted[i], k) })
	return i, i < len(c.sorted) && c.sorted[i] == k
}

// code of a key already stored.
func (c *keyCodec[K]) code(k K) (int64, bool) {
	if v, direct, err := directCode(k); direct {
		return v, err == nil
	}
	i, ok := c.find(k)
	if !ok {
		return 0, false
	}
	return c.codes[i], true
}

// assign returns codes for new (distinct, not yet stored) keys; respaced is true when every stored key got a new code
// (the caller then pushes code(k) of every stored row to the device).
func (c *keyCodec[K]) assign(keys []K) (codes []int64, respaced bool, err error) {
	codes = make([]int64, len(keys))
	if len(keys) == 0 {
		return codes, false, nil
	}
	if _, direct, _ := directCode(keys[0]); direct {
		for i, k := range keys {
			v, _, e := directCode(k)
			if e != nil {
				return nil, false, e
			}
			codes[i] = v
		}
		return codes, false, nil
	}
	order := make([]int, len(keys))
	for i := range order {
		order[i] = i
	}
	sort.Slice(order, func(a, b int) bool { return cmp.Less(keys[order[a]], keys[order[b]]) })
	n := len(c.sorted) + len(keys)
	mk := make([]K, 0, n)
	mc := make([]int64, 0, n)
	isNew := make([]bool, 0, n)
	i, j := 0, 0
	for i < len(c.sorted) || j < len(order) {
		if j >= len(order) || (i < len(c.sorted) && cmp.Less(c.sorted[i], keys[order[j]])) {
			mk, mc, isNew = append(mk, c.sorted[i]), append(mc, c.codes[i]), append(isNew, false)
			i++
		} else {
			mk, mc, isNew = append(mk, keys[order[j]]), append(mc, 0), append(isNew, true)
			j++
		}
	}
	ok := true
	for p := 0; p < n && ok; {
		if !isNew[p] {
			p++
			continue
		}
		q := p
		for q < n && isNew[q] {
			q++
		}
		r := int64(q - p)
		var base, step int64
		switch {
		case p == 0 && q == n:
			base, step = 0, codeGap
		case p == 0:
			base, step = mc[q]-codeGap*(r+1), codeGap
		case q == n:
			base, step = mc[p-1], codeGap
		default:
			base, step = mc[p-1], (mc[q]-mc[p-1])/(r+1)
		}
		if step < 1 || base+step < -(1<<62) || base+step*r > (1<<62) {
			ok = false
			break
		}
		for t := int64(0); t < r; t++ {
			mc[p+int(t)] = base + step*(t+1)
		}
		p = q
	}
	if !ok { // a gap ran out: re-space everything
		start := -int64(n/2) * codeGap
		for p := range mc {
			mc[p] = start + codeGap*int64(p)
		}
		respaced = len(c.sorted) > 0
	}
	c.sorted, c.codes = mk, mc
	for i, k := range keys {
		p, _ := c.find(k)
		codes[i] = c.codes[p]
	}
	return codes, respaced, nil
}

// forget undoes assign for keys whose insertion failed.
func (c *keyCodec[K]) forget(keys []K) {
	for _, k := range keys {
		if _, direct, _ := directCode(k); direct {
			return
		}
		if p, ok := c.find(k); ok {
			c.sorted = append(c.sorted[:p], c.sorted[p+1:]...)
			c.codes = append(c.codes[:p], c.codes[p+1:]...)
		}
	}
}
