package gpustore

/*
#include "omnia_recall.h"
*/
import "C"

import (
	"fmt"
	"math"
	"sort"
	"unsafe"

	"gonum.org/v1/gonum/mat"
)

// PCAReduceResident is PCAReduce (ee/pkg/memory/projection/pca.go:28-67) over rows that already sit in HBM:
// the column means, the covariance and the final projection run on the GPU, the dim x dim symmetric
// eigenproblem (independent of the row count) stays in gonum.  `rows` are the store rows of the entities
// the projector kept after applyCap (project.go:99-112); ProjectionRows lists the candidates.
//
// NOT COMPILED in the build image (no Go toolchain there); the same sequence is exercised through ctypes
// by GPUMemoryStore.pca_reduce and tests/test_gpu_pca.py.
func (s *Store) PCAReduceResident(rows []int64, dim, k int) (*mat.Dense, error) {
	n := len(rows)
	if n == 0 || dim == 0 {
		return nil, fmt.Errorf("projection: empty input matrix (%dx%d)", n, dim)
	}
	if k > dim {
		k = dim
	}
	if k > n {
		k = n
	}
	if k < 1 {
		k = 1
	}
	basis := make([]float64, dim*k) // [dim][k] row-major
	if n < 2 {
		for j := 0; j < k; j++ { // firstKCols (pca.go:70-83)
			basis[j*k+j] = 1
		}
	} else {
		mean := make([]float64, dim)
		cov := make([]float64, dim*dim)
		if rc := C.omr_pca_moments(s.h, (*C.int64_t)(unsafe.Pointer(&rows[0])), C.int64_t(n),
			(*C.double)(unsafe.Pointer(&mean[0])), (*C.double)(unsafe.Pointer(&cov[0]))); rc != C.OMR_OK {
			return nil, lastErr(rc)
		}
		var es mat.EigenSym
		if ok := es.Factorize(mat.NewSymDense(dim, cov), true); !ok {
			return nil, fmt.Errorf("projection: PCA failed (%dx%d, k=%d)", n, dim, k)
		}
		vals := es.Values(nil) // ascending
		var vecs mat.Dense
		es.VectorsTo(&vecs)
		order := make([]int, dim)
		for i := range order {
			order[i] = i
		}
		sort.SliceStable(order, func(a, b int) bool { return vals[order[a]] > vals[order[b]] })
		for c := 0; c < k; c++ {
			for r := 0; r < dim; r++ {
				basis[r*k+c] = vecs.At(r, order[c])
			}
		}
	}
	out := make([]float64, n*k)
	if rc := C.omr_pca_project(s.h, (*C.int64_t)(unsafe.Pointer(&rows[0])), C.int64_t(n),
		(*C.double)(unsafe.Pointer(&basis[0])), C.int32_t(k), (*C.double)(unsafe.Pointer(&out[0]))); rc != C.OMR_OK {
		return nil, lastErr(rc)
	}
	return mat.NewDense(n, k, out), nil
}

// ProjectionRows lists one resident row per live entity in scope, the population LoadProjectionInputs
// returns (internal/memory/projection_store.go:33-62).
func (s *Store) ProjectionRows(scope map[string]string) ([]int64, error) {
	f := singleTierFilter(scope, 0, math.NaN()) // NaN = no distance cutoff
	defer f.free()
	var found C.int64_t
	if rc := C.omr_projection_rows(s.h, &f.f, nil, 0, &found); rc != C.OMR_OK {
		return nil, lastErr(rc)
	}
	rows := make([]int64, int(found))
	if found == 0 {
		return rows, nil
	}
	if rc := C.omr_projection_rows(s.h, &f.f, (*C.int64_t)(unsafe.Pointer(&rows[0])), found, &found); rc != C.OMR_OK {
		return nil, lastErr(rc)
	}
	return rows[:found], nil
}
