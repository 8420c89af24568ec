package gpustore

/*
#include "omnia_recall.h"
*/
import "C"

import (
	"context"
	"fmt"

	"github.com/altairalabs/omnia/internal/memory/consolidation"
)

// PreFilterRunner replaces the entity-duplicate arm of internal/memory/postgres.PreFilterRunner
// (consolidation_prefilter.go:176-203); the other two axes keep delegating to Postgres.
// Wired at cmd/memory-api/main.go:1627, consumed by ee/pkg/memory/consolidation/worker.go:413-451.
type PreFilterRunner struct {
	Base interface {
		RunStaleObservations(context.Context, consolidation.PreFilterOptions) ([]consolidation.Bucket, error)
		RunCrossScopeCandidates(context.Context, consolidation.PreFilterOptions) ([]consolidation.Bucket, error)
	}
	Store *Store
}

func (r *PreFilterRunner) RunEntityDuplicateCandidates(ctx context.Context,
	opts consolidation.PreFilterOptions) ([]consolidation.Bucket, error) {
	f := &cFilter{}
	defer f.free()
	o := C.omr_selfjoin_opts{
		workspace_id:      f.str(opts.WorkspaceID),
		similarity_floor:  C.float(opts.SimilarityFloor), // a Go float32, widened once inside (prefilter.go:18)
		limit:             C.int32_t(opts.MaxBucketsPerPass),
		require_same_kind: 1,
		require_mutable:   1,
	}
	n := opts.MaxBucketsPerPass
	if n <= 0 {
		n = 100 // worker.go:432-441
	}
	pairs := make([]C.omr_pair, n)
	var found C.int64_t
	if rc := C.omr_selfjoin(r.Store.h, &o, &pairs[0], C.int64_t(n), &found); rc != C.OMR_OK {
		return nil, fmt.Errorf("entity-dupe query: %w", lastErr(rc))
	}
	if int(found) < n {
		n = int(found)
	}
	out := make([]consolidation.Bucket, 0, n)
	for i := 0; i < n; i++ {
		canonical, dupe := uuidOf(&pairs[i].canonical_id), uuidOf(&pairs[i].dupe_id)
		kind := C.GoString(C.omr_kind_name(r.Store.h, pairs[i].kind))
		out = append(out, consolidation.Bucket{
			Key: fmt.Sprintf("dupe:%s:%s", canonical, dupe),
			Entries: []consolidation.BucketEntry{ // names are filled by one SELECT ... WHERE id = ANY($1)
				{ID: canonical, Scope: consolidation.Scope{WorkspaceID: opts.WorkspaceID}, Mutability: consolidation.MutabilityMutable},
				{ID: dupe, Scope: consolidation.Scope{WorkspaceID: opts.WorkspaceID}, Mutability: consolidation.MutabilityMutable},
			},
			Stats: map[string]any{"similarity": float64(pairs[i].similarity), "kind": kind},
		})
	}
	return out, nil
}
