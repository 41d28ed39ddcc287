package hnsw

import "reflect"

// DistanceFunc is a function that computes the distance between two vectors (reference distance.go:11-12).
type DistanceFunc func(a, b []float32) float32

const (
	metricEuclidean = 0 // B200HNSW_EUCLIDEAN
	metricCosine    = 1 // B200HNSW_COSINE
)

// CosineDistance and EuclideanDistance select the device metric.  Called directly they evaluate ONE pair on the device
// (so even the distance known-answer tests run the kernels' arithmetic); there is no CPU arithmetic in this package.
func CosineDistance(a, b []float32) float32 {
	d, err := deviceDistance(metricCosine, a, b)
	if err != nil {
		panic(err)
	}
	return d
}

// EuclideanDistance: sqrt(sum((a-b)^2)) (reference distance.go:20-23).
func EuclideanDistance(a, b []float32) float32 {
	d, err := deviceDistance(metricEuclidean, a, b)
	if err != nil {
		panic(err)
	}
	return d
}

var distanceFuncs = map[string]DistanceFunc{
	"euclidean": EuclideanDistance,
	"cosine":    CosineDistance,
}

func distanceFuncToName(fn DistanceFunc) (string, bool) {
	for name, f := range distanceFuncs {
		if reflect.ValueOf(fn).Pointer() == reflect.ValueOf(f).Pointer() {
			return name, true
		}
	}
	return "", false
}

// RegisterDistanceFunc registers a distance function with a name (reference distance.go:44-46).  A registered custom
// function can be exported / imported by name, but only the two built-in metrics have device kernels.
func RegisterDistanceFunc(name string, fn DistanceFunc) {
	distanceFuncs[name] = fn
}

// metricOf: only the two built-in metrics have kernels; anything else is an error (no CPU fallback).
func metricOf(fn DistanceFunc) (int, error) {
	if fn == nil {
		return 0, errDistanceNotSet
	}
	p := reflect.ValueOf(fn).Pointer()
	switch p {
	case reflect.ValueOf(EuclideanDistance).Pointer():
		return metricEuclidean, nil
	case reflect.ValueOf(CosineDistance).Pointer():
		return metricCosine, nil
	}
	return 0, errNoKernel
}

// MetricOf reports the device metric a DistanceFunc selects (0 Euclidean, 1 cosine); used by the hybrid package.
func MetricOf(fn DistanceFunc) (int, error) { return metricOf(fn) }
