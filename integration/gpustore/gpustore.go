// Package gpustore puts the B200 recall engine behind Omnia's memory.Store.
//
// NOT COMPILED in the build image (no Go toolchain there); see README.md.
package gpustore

/*
#cgo CFLAGS: -I${SRCDIR}/../../../third_party/omnia_recall/include
#cgo LDFLAGS: -L${SRCDIR}/../../../third_party/omnia_recall/lib -lomnia_recall
#include <stdlib.h>
#include <math.h>
#include "omnia_recall.h"
*/
import "C"

import (
	"context"
	"errors"
	"time"
	"unsafe"

	"github.com/google/uuid"

	"github.com/altairalabs/omnia/internal/memory"
	"github.com/altairalabs/omnia/internal/memory/consolidation"
	"github.com/altairalabs/omnia/internal/memory/postgres"
)

// Store is a drop-in memory.Store. Everything delegates to the embedded Postgres store
// except the scans that pgvector's `<=>` used to do.
type Store struct {
	*memory.PostgresMemoryStore
	h   *C.omr_store
	dim int
}

var _ memory.Store = (*Store)(nil) // same compile-time check as internal/memory/store.go:31

// Open replaces the vector column EnsureEmbeddingSchema creates
// (internal/memory/postgres/embedding_schema.go:125-151): dim fixes the width.
func Open(pg *memory.PostgresMemoryStore, dim int, capacity int64, device int) (*Store, error) {
	cfg := C.omr_config{dim: C.int32_t(dim), device: C.int32_t(device), capacity: C.int64_t(capacity)}
	var h *C.omr_store
	if rc := C.omr_open(&cfg, &h); rc != C.OMR_OK {
		return nil, lastErr(rc)
	}
	return &Store{PostgresMemoryStore: pg, h: h, dim: dim}, nil
}

// Close releases the GPU store; the pgx pool still belongs to the caller (store.go:123-130).
func (s *Store) Close() {
	C.omr_close(s.h)
	s.PostgresMemoryStore.Close()
}

func lastErr(rc C.int) error {
	msg := C.GoString(C.omr_last_error()) // carries the reference's "memory: " prefix already
	if rc == C.OMR_E_NOT_FOUND {
		return memory.ErrNotFound // store.go:69 -> HTTP 404 (internal/memory/api/handler.go:598-600)
	}
	return errors.New(msg)
}

// cFilter owns the C strings of one omr_filter.
type cFilter struct {
	f    C.omr_filter
	strs []unsafe.Pointer
}

func (c *cFilter) str(s string) *C.char {
	if s == "" {
		return nil
	}
	p := C.CString(s)
	c.strs = append(c.strs, unsafe.Pointer(p))
	return p
}

func (c *cFilter) free() {
	for _, p := range c.strs {
		C.free(p)
	}
}

// singleTierFilter mirrors store.go:250-258: a nil user / agent means "no constraint".
func singleTierFilter(scope map[string]string, minConfidence float64, maxDistance float64) *cFilter {
	c := &cFilter{}
	c.f.workspace_id = c.str(scope[memory.ScopeWorkspaceID])
	if u := scope[memory.ScopeVirtualUserID]; u != "" {
		c.f.virtual_user_id, c.f.user_mode = c.str(u), C.OMR_MATCH_EQ
	} else {
		c.f.user_mode = C.OMR_MATCH_ANY
	}
	if a := scope[memory.ScopeAgentID]; a != "" {
		c.f.agent_id, c.f.agent_mode = c.str(a), C.OMR_MATCH_EQ
	} else {
		c.f.agent_mode = C.OMR_MATCH_ANY
	}
	c.f.min_confidence = C.float(minConfidence)
	c.f.max_distance = C.double(maxDistance)
	c.f.now_us = C.int64_t(time.Now().UnixMicro()) // `valid_until > now()` is evaluated per query
	return c
}

// multiTierFilter mirrors retrieve_multi_tier.go:313-328: empty means IS NULL, set means IS NULL OR = x.
func multiTierFilter(req memory.MultiTierRequest) *cFilter {
	c := &cFilter{}
	c.f.workspace_id = c.str(req.WorkspaceID)
	c.f.user_mode, c.f.agent_mode = C.OMR_MATCH_NULL_ONLY, C.OMR_MATCH_NULL_ONLY
	if req.UserID != "" {
		c.f.virtual_user_id, c.f.user_mode = c.str(req.UserID), C.OMR_MATCH_NULL_OR_EQ
	}
	if req.AgentID != "" {
		c.f.agent_id, c.f.agent_mode = c.str(req.AgentID), C.OMR_MATCH_NULL_OR_EQ
	}
	c.f.min_confidence = C.float(req.MinConfidence)
	c.f.max_distance = C.double(C.NAN)
	now := req.Now
	if now.IsZero() {
		now = time.Now()
	}
	c.f.now_us = C.int64_t(now.UnixMicro())
	// req.Types / req.Purposes become c.f.kinds / c.f.purposes (arrays of C strings) the same way.
	return c
}

func uuidOf(b *[16]C.uint8_t) string {
	return uuid.UUID(*(*[16]byte)(unsafe.Pointer(b))).String()
}

// FindSimilarObservations replaces internal/memory/store_read.go:243-306.
func (s *Store) FindSimilarObservations(ctx context.Context, scope map[string]string,
	queryEmbedding []float32, k int, minSimilarity float64) ([]memory.SimilarObservation, error) {
	if scope[memory.ScopeWorkspaceID] == "" {
		return nil, errors.New("memory: workspace_id is required")
	}
	if k <= 0 {
		k = memory.DefaultDuplicateCandidateLimit
	}
	if len(queryEmbedding) == 0 {
		return nil, nil
	}
	f := singleTierFilter(scope, 0, 1.0-minSimilarity) // maxDistance := 1.0 - minSimilarity (store_read.go:260)
	defer f.free()
	hits := make([]C.omr_hit, k)
	var n C.int32_t
	// Concurrent writers are coalesced inside the library: the calls that arrive while a search runs share
	// one pass over the matrix (service_write.go:129-141 calls this once per write).
	rc := C.omr_search(s.h, (*C.float)(unsafe.Pointer(&queryEmbedding[0])), 1, C.int32_t(len(queryEmbedding)),
		&f.f, 1, nil, C.int32_t(k), &hits[0], &n)
	if rc != C.OMR_OK {
		return nil, lastErr(rc) // a dimension mismatch is an error, not a panic (hnsw_ef_search_test.go:126-144)
	}
	out := make([]memory.SimilarObservation, int(n))
	for i := range out {
		out[i].ObservationID = uuidOf(&hits[i].observation_id)
		out[i].EntityID = uuidOf(&hits[i].entity_id)
		out[i].Similarity = 1 - float64(hits[i].distance) // store_read.go:268
	}
	return out, s.fillContent(ctx, out)
}

// RankedEntity is one row of the `cosine_ranked` CTE (store.go:267-272).
type RankedEntity struct {
	EntityID string
	CosDist  float64
}

// CosineRanked is the whole cosine arm of the hybrid query (store.go:240-272). RetrieveHybrid and
// RetrieveMultiTierHybrid run the reference's own SQL with the three cosine CTEs replaced by
//
//	cosine_ranked AS (SELECT entity_id, row_number() OVER (ORDER BY cos_dist) AS cos_rn
//	                  FROM unnest($ids::uuid[], $dists::float8[]) AS t(entity_id, cos_dist))
//
// so FTS, RRF, the quality multipliers and the final eligibility re-check stay in Postgres.
func (s *Store) CosineRanked(f *cFilter, queryEmbedding []float32, fanout int) ([]RankedEntity, error) {
	hits := make([]C.omr_hit, fanout)
	var n C.int32_t
	rc := C.omr_cosine_ranked(s.h, (*C.float)(unsafe.Pointer(&queryEmbedding[0])), C.int32_t(len(queryEmbedding)),
		&f.f, C.int32_t(fanout), &hits[0], &n)
	if rc != C.OMR_OK {
		return nil, lastErr(rc)
	}
	out := make([]RankedEntity, int(n))
	for i := range out {
		out[i] = RankedEntity{EntityID: uuidOf(&hits[i].entity_id), CosDist: float64(hits[i].distance)}
	}
	return out, nil
}

// RetrieveHybrid replaces internal/memory/store_read.go:351-406 (cosine arm only).
func (s *Store) RetrieveHybrid(ctx context.Context, scope map[string]string, query string,
	queryEmbedding []float32, opts memory.RetrieveOptions) ([]*memory.Memory, error) {
	if query == "" || len(queryEmbedding) == 0 {
		return s.PostgresMemoryStore.RetrieveHybrid(ctx, scope, query, queryEmbedding, opts) // FTS fallback (:361-363)
	}
	limit := opts.Limit
	if limit <= 0 {
		limit = 50 // defaultMemoryLimit, store.go:330
	}
	fanout := limit * 5 // store_read.go:369-372
	if fanout < 50 {
		fanout = 50
	}
	f := singleTierFilter(scope, opts.MinConfidence, float64(C.NAN))
	defer f.free()
	ranked, err := s.CosineRanked(f, queryEmbedding, fanout)
	if err != nil {
		return nil, err
	}
	return s.PostgresMemoryStore.RetrieveHybridWithCosine(ctx, scope, query, ranked, opts) // spliced statement
}

// ---- keeping the HBM matrix in sync (SURVEY.md appendix A) ----------------------------------

// UpdateEmbedding replaces nothing in Postgres: it mirrors store_write.go:557-578 onto the GPU.
func (s *Store) UpdateEmbedding(ctx context.Context, entityID string, embedding []float32, model string) error {
	if err := s.PostgresMemoryStore.UpdateEmbedding(ctx, entityID, embedding, model); err != nil {
		return err
	}
	id := uuid.MustParse(entityID)
	rc := C.omr_update_embedding(s.h, (*C.uint8_t)(unsafe.Pointer(&id[0])),
		(*C.float)(unsafe.Pointer(&embedding[0])), C.int32_t(len(embedding)))
	if rc == C.OMR_E_NOT_FOUND {
		return s.upsertFromPostgres(ctx, entityID, embedding) // first vector of this entity: needs its scope columns
	}
	if rc != C.OMR_OK {
		return lastErr(rc)
	}
	return nil
}

// Delete mirrors the soft forget of store_delete.go:30-57.
func (s *Store) Delete(ctx context.Context, scope map[string]string, memoryID string) error {
	if err := s.PostgresMemoryStore.Delete(ctx, scope, memoryID); err != nil {
		return err
	}
	id := uuid.MustParse(memoryID)
	C.omr_set_entity_flags(s.h, (*C.uint8_t)(unsafe.Pointer(&id[0])), C.OMR_ROW_FORGOTTEN, 0)
	return nil
}

// DeleteAll mirrors the hard delete of store_delete.go:60-87 (cascade included: rows go with their entity).
func (s *Store) DeleteAll(ctx context.Context, scope map[string]string) error {
	if err := s.PostgresMemoryStore.DeleteAll(ctx, scope); err != nil {
		return err
	}
	_, err := s.deleteScope(scope, 0)
	return err
}

// BatchDelete mirrors store_delete.go:89-125.  Postgres picks which `limit` entities go, so the GPU
// side deletes by the ids the statement returned rather than by its own iteration order.
func (s *Store) BatchDelete(ctx context.Context, scope map[string]string, limit int) (int, error) {
	ids, err := s.PostgresMemoryStore.BatchDeleteReturningIDs(ctx, scope, limit) // DELETE ... RETURNING id
	if err != nil {
		return 0, err
	}
	for _, e := range ids {
		id := uuid.MustParse(e)
		C.omr_delete_entity(s.h, (*C.uint8_t)(unsafe.Pointer(&id[0])))
	}
	return len(ids), nil
}

func (s *Store) deleteScope(scope map[string]string, limit int64) (int64, error) {
	ws := C.CString(scope[memory.ScopeWorkspaceID])
	defer C.free(unsafe.Pointer(ws))
	var user *C.char
	if u := scope[memory.ScopeVirtualUserID]; u != "" {
		user = C.CString(u)
		defer C.free(unsafe.Pointer(user))
	}
	var n C.int64_t
	if rc := C.omr_delete_scope(s.h, ws, user, C.int64_t(limit), &n); rc != C.OMR_OK {
		return 0, lastErr(rc)
	}
	return int64(n), nil
}

// ConsolidationWriter wraps postgres.ConsolidationWriter so the scope and entity rewrites of
// internal/memory/postgres/consolidation_writes.go:63-160 reach the device-side row metadata.
type ConsolidationWriter struct {
	*postgres.ConsolidationWriter
	gpu *Store
}

func idBlock(ids []string) []byte {
	out := make([]byte, 0, 16*len(ids))
	for _, s := range ids {
		id := uuid.MustParse(s)
		out = append(out, id[:]...)
	}
	return out
}

func (w *ConsolidationWriter) Rescope(ctx context.Context, rw consolidation.RescopeWrite) error {
	if err := w.ConsolidationWriter.Rescope(ctx, rw); err != nil || len(rw.TargetIDs) == 0 {
		return err
	}
	ids := idBlock(rw.TargetIDs)
	user, agent := C.CString(rw.NewScope.UserID), C.CString(rw.NewScope.AgentID)
	defer C.free(unsafe.Pointer(user))
	defer C.free(unsafe.Pointer(agent))
	if rc := C.omr_rescope(w.gpu.h, (*C.uint8_t)(unsafe.Pointer(&ids[0])), C.int32_t(len(rw.TargetIDs)), user, agent); rc != C.OMR_OK {
		return lastErr(rc)
	}
	return nil
}

func (w *ConsolidationWriter) MergeEntities(ctx context.Context, mw consolidation.MergeWrite) error {
	if err := w.ConsolidationWriter.MergeEntities(ctx, mw); err != nil || len(mw.MergeIDs) == 0 {
		return err
	}
	canon := uuid.MustParse(mw.CanonicalID)
	ids := idBlock(mw.MergeIDs)
	if rc := C.omr_merge_entities(w.gpu.h, (*C.uint8_t)(unsafe.Pointer(&canon[0])),
		(*C.uint8_t)(unsafe.Pointer(&ids[0])), C.int32_t(len(mw.MergeIDs))); rc != C.OMR_OK {
		return lastErr(rc)
	}
	return nil
}

// Supersede, Invalidate, Discard and Rescore (consolidation_writes.go:63-73, 94-104, 129-160) map onto
// omr_set_row_flags / omr_set_valid_until / omr_set_confidence one target id at a time.
func (w *ConsolidationWriter) Supersede(ctx context.Context, sw consolidation.SupersedeWrite) error {
	if err := w.ConsolidationWriter.Supersede(ctx, sw); err != nil {
		return err
	}
	for _, t := range sw.TargetIDs {
		id := uuid.MustParse(t)
		C.omr_set_row_flags(w.gpu.h, (*C.uint8_t)(unsafe.Pointer(&id[0])), C.OMR_ROW_SUPERSEDED, 0)
	}
	return nil
}
