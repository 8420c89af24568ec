/*
 * omnia_recall.h -- C ABI of the B200 recall engine (libomnia_recall.so).
 *
 * This is the drop-in boundary for ONE path of AltairaLabs/Omnia: semantic recall
 * over the agent memory store and the near-duplicate scan of memory consolidation.
 * The reference has no FFI for this path (it is pure Go issuing pgvector SQL), so
 * every entry point below names the reference interface it stands in for; a Go
 * `gpustore` package binds them with cgo (INTEGRATION.md shows the stubs).
 *
 * Conventions
 *  - Every function returns 0 (OMR_OK) or a negative OMR_E_* code; the message for
 *    the calling thread's last failure is omr_last_error().  Nothing aborts.
 *  - Host-pointer arguments are only read/written during the call (cgo pointer rule);
 *    the library copies what it keeps.
 *  - A store handle may be used from many threads; calls on one store serialise.
 *  - uuids are 16 raw bytes in network order, compared bytewise as Postgres does.
 *  - Entry points whose name contains `_device_` take DEVICE pointers and a CUDA
 *    stream (cudaStream_t passed as void*); they are what bench.py times with the
 *    inputs already resident in HBM, and what the multi-GPU host code drives.
 *  - There is no CPU fallback: without a CUDA device omr_open fails.
 */
#ifndef OMNIA_RECALL_H_
#define OMNIA_RECALL_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define OMR_API __attribute__((visibility("default")))
#else
#define OMR_API
#endif

#define OMR_OK 0
#define OMR_E_INVALID (-1)   /* bad argument */
#define OMR_E_CUDA (-2)      /* a CUDA call failed */
#define OMR_E_DIM (-3)       /* embedding dimension mismatch (hnsw_ef_search_test.go:126-144 wants an error) */
#define OMR_E_CAPACITY (-4)  /* store or output buffer full */
#define OMR_E_NOT_FOUND (-5) /* memory.ErrNotFound, internal/memory/store.go:69 */
#define OMR_E_WORKSPACE (-6) /* errWorkspaceRequired, internal/memory/store.go:60 */
#define OMR_E_CANCELLED (-7) /* the caller's cancel flag was raised (Go: ctx.Err()) */

/* Row flag bits (the boolean columns the recall WHERE clause reads, store.go:250-258). */
#define OMR_ROW_FORGOTTEN 1u     /* memory_entities.forgotten */
#define OMR_ROW_SUPERSEDED 2u    /* memory_observations.superseded_by IS NOT NULL */
#define OMR_ROW_HAS_EMBEDDING 4u /* memory_observations.embedding IS NOT NULL */
#define OMR_ROW_MUTABLE 8u       /* mutability = 'mutable' (consolidation/prefilter.go:125-126) */

/* How a scope column is matched.  Single-tier recall uses ANY / EQ
 * (`$2::text IS NULL OR e.virtual_user_id = $2`, store.go:252-253); multi-tier uses
 * NULL_ONLY / NULL_OR_EQ (internal/memory/retrieve_multi_tier.go:313-328). */
#define OMR_MATCH_ANY 0
#define OMR_MATCH_NULL_ONLY 1
#define OMR_MATCH_EQ 2
#define OMR_MATCH_NULL_OR_EQ 3

#define OMR_NO_EXPIRY INT64_MAX /* valid_until IS NULL */
#define OMR_MAX_DIM 2000        /* internal/memory/postgres/embedding_schema.go:64-68 */

typedef struct omr_store omr_store;

/* Replaces the vector column created by EnsureEmbeddingSchema
 * (internal/memory/postgres/embedding_schema.go:125-151): `dim` fixes the width. */
typedef struct omr_config {
  int32_t dim;              /* 1..OMR_MAX_DIM */
  int32_t device;           /* CUDA ordinal */
  int64_t capacity;         /* rows this shard can hold */
  int64_t row_base;         /* global row id = row_base + local_row * row_stride */
  int64_t row_stride;       /* 0 is read as 1 */
  float candidate_margin;   /* bf16 stage error allowance in cosine units; <= 0 -> 4.5e-3, the deterministic bound
                               of the bf16 stage for any dim <= 2048; smaller values turn the
                               exactness proof into a statistical statement and are for experiments only */
  int32_t sub_bucket;       /* 0 = choose from k; else 32 / 64 / 128 / 256 */
  int32_t rescore_width;    /* 0 = choose from k; candidates rescored exactly per query */
  int32_t force_path;       /* 0 auto, 1 streaming scorer, 2 tensor-core scorer (one CTA per tile),
                               3 exact scan only, 4 tensor-core scorer with CTA pairs (cta_group::2) */
} omr_config;

OMR_API int omr_open(const omr_config* cfg, omr_store** out);
OMR_API void omr_close(omr_store* s);
OMR_API const char* omr_last_error(void);
OMR_API int64_t omr_count(const omr_store* s); /* rows in use, tombstones included */

/* ---- ingest ----------------------------------------------------------------
 * Bulk form of UpdateEmbedding / UpdateObservationEmbedding
 * (internal/memory/store_write.go:557-600) for the start-up load and the re-embed
 * worker (internal/memory/reembed_worker.go:181-218).  Struct-of-arrays with small
 * string tables so a million rows cross the boundary without a million strings.
 * Any array pointer may be NULL: ids default to the row number, entity = observation,
 * user/agent NULL, confidence 1, flags HAS_EMBEDDING|MUTABLE, no expiry, kind/purpose 0. */
typedef struct omr_bulk {
  int64_t n;
  const float* vectors;            /* [n][dim], host */
  const uint8_t* observation_ids;  /* [n][16] */
  const uint8_t* entity_ids;       /* [n][16] */
  const char* const* workspaces;   int32_t n_workspaces;  const uint32_t* workspace_idx; /* [n] */
  const char* const* users;        int32_t n_users;       const int32_t* user_idx;       /* -1 = NULL */
  const char* const* agents;       int32_t n_agents;      const int32_t* agent_idx;      /* -1 = NULL */
  const char* const* kinds;        int32_t n_kinds;       const uint16_t* kind_idx;
  const char* const* purposes;     int32_t n_purposes;    const uint16_t* purpose_idx;
  const float* confidence;
  const uint8_t* flags;
  const int64_t* valid_until_us;
} omr_bulk;
OMR_API int omr_bulk_load(omr_store* s, const omr_bulk* b);

/* One row with its scope as strings.  If observation_id is already resident the row is
 * overwritten in place (UpdateObservationEmbedding, store_write.go:580-600). */
typedef struct omr_row {
  uint8_t observation_id[16];
  uint8_t entity_id[16];
  const char* workspace_id;    /* required */
  const char* virtual_user_id; /* NULL = SQL NULL */
  const char* agent_id;        /* NULL = SQL NULL */
  const char* kind;
  const char* purpose;
  float confidence;
  uint32_t flags;
  int64_t valid_until_us;
} omr_row;
OMR_API int omr_upsert(omr_store* s, const omr_row* row, const float* vec, int32_t dim);
/* n upserts in one call: one staged copy of the vectors and one ingest launch instead of n of each (the re-embed
 * worker writes a page of rows at a time, internal/memory/reembed_worker.go:181-218).  rows[i] goes with
 * vecs[i*dim .. (i+1)*dim); an observation id given twice keeps its last vector.  A row that fails validation
 * stops the call; the rows before it stay upserted, as with n single calls. */
OMR_API int omr_upsert_batch(omr_store* s, const omr_row* rows, const float* vecs, int32_t n, int32_t dim);
/* Reclamation of hard-deleted rows (tombstone GC, internal/memory/tombstone.go:124; DeleteAll / BatchDelete;
 * re-embedding churn).  A hard-deleted row's slot goes on a free list and omr_upsert / omr_upsert_batch hand it
 * to the next NEW observation, so a store under steady churn never fills up with tombstones.  omr_compact
 * additionally squeezes every hard-deleted row out (rows behind it move down, order kept, matrices rewritten
 * in place on the device) and lowers omr_count; global row ids handed out earlier are stale afterwards.
 * *reclaimed (may be NULL) receives the number of rows removed.  omr_free_slots: slots waiting for reuse. */
OMR_API int omr_compact(omr_store* s, int64_t* reclaimed);
OMR_API int64_t omr_free_slots(const omr_store* s);
/* UpdateEmbedding (store_write.go:557-578): new vector for the entity's latest observation. */
OMR_API int omr_update_embedding(omr_store* s, const uint8_t entity_id[16], const float* vec, int32_t dim);

/* Row-state edits that change recall eligibility. */
OMR_API int omr_set_row_flags(omr_store* s, const uint8_t observation_id[16], uint32_t set_mask, uint32_t clear_mask);
OMR_API int omr_set_entity_flags(omr_store* s, const uint8_t entity_id[16], uint32_t set_mask, uint32_t clear_mask);
OMR_API int omr_set_valid_until(omr_store* s, const uint8_t observation_id[16], int64_t valid_until_us);
OMR_API int omr_set_confidence(omr_store* s, const uint8_t observation_id[16], float confidence);
OMR_API int omr_delete_entity(omr_store* s, const uint8_t entity_id[16]); /* hard delete, store_delete.go:60-103 */
/* Hard delete of one observation: tombstone GC (internal/memory/tombstone.go:124) and the
 * superseded-grace purge (internal/memory/supersession_store.go:43). */
OMR_API int omr_delete_observation(omr_store* s, const uint8_t observation_id[16]);
/* DeleteAll / BatchDelete (internal/memory/store_delete.go:60-125): hard-deletes the entities of a
 * workspace, narrowed to one user when virtual_user_id is non-empty; limit <= 0 means no limit
 * (DeleteAll), limit > 0 deletes at most that many entities (BatchDelete).  *deleted (may be NULL)
 * receives the number of entities removed. */
OMR_API int omr_delete_scope(omr_store* s, const char* workspace_id, const char* virtual_user_id, int64_t limit,
                             int64_t* deleted);
/* Consolidation Rescope (internal/memory/postgres/consolidation_writes.go:75-92): rewrites
 * virtual_user_id / agent_id of the entities that own the target observations ("" or NULL = SQL NULL,
 * the statement's NULLIF); every resident row of those entities follows.  Unknown ids match nothing. */
OMR_API int omr_rescope(omr_store* s, const uint8_t* observation_ids /*[n][16]*/, int32_t n,
                        const char* virtual_user_id, const char* agent_id);
/* Consolidation MergeEntities (consolidation_writes.go:104-127): the observations of merge_ids move to
 * canonical_id and take its entity-level columns (workspace, user, agent, kind, purpose, forgotten,
 * mutability) when the canonical entity has a resident row; the merged entities are left empty, which
 * is what their forgotten = true amounts to for recall. */
OMR_API int omr_merge_entities(omr_store* s, const uint8_t canonical_id[16], const uint8_t* merge_ids /*[n][16]*/,
                               int32_t n);
/* Drops every row and identity map; capacity, dimension and device buffers stay.  The embedding
 * dimension change (internal/memory/postgres/embedding_schema.go:125-151 drops and recreates the
 * column, discarding all vectors) is omr_close + omr_open at the new dimension; omr_clear is the
 * same flush when the dimension is unchanged (re-embed with a new model, reembed_worker.go:181-218). */
OMR_API int omr_clear(omr_store* s);

/* ---- recall ------------------------------------------------------------------
 * The WHERE clause of the recall queries (reference row a8). */
typedef struct omr_filter {
  const char* workspace_id;     /* required: e.workspace_id = $1 */
  const char* virtual_user_id;  int32_t user_mode;
  const char* agent_id;         int32_t agent_mode;
  const char* const* kinds;     int32_t n_kinds;     /* kind = ANY($types); 0 = no constraint */
  const char* const* purposes;  int32_t n_purposes;
  float min_confidence;         /* o.confidence >= $8 */
  double max_distance;          /* `o.embedding <=> $1 <= $5` (store_read.go:278); >= 2 or NaN = none */
  int64_t now_us;               /* the query's now() for `valid_until > now()` */
} omr_filter;

typedef struct omr_hit {
  uint8_t observation_id[16];
  uint8_t entity_id[16];
  double distance; /* pgvector cosine distance; similarity = 1 - distance (store_read.go:268) */
  int64_t row;     /* global row id */
} omr_hit;

/* `ORDER BY o.embedding <=> $q LIMIT k` under a filter, for a batch of queries.
 * Replaces the SQL of FindSimilarObservations (internal/memory/store_read.go:243-306) and
 * the cosine_ann CTEs (internal/memory/store.go:240-261,
 * internal/memory/retrieve_multi_tier_hybrid.go:78-96).  query_filter[i] picks the filter
 * of query i (NULL = filter 0 for all).  hits is [nq][k]; counts[i] <= k entries are valid,
 * nearest first, ties by row id.  Calls made concurrently from several threads are coalesced: the
 * calls that arrive while a search is running are answered together by ONE pass over the matrix
 * (the dedup-on-write micro-batching of internal/memory/api/service_write.go:129-141, 213-228). */
OMR_API int omr_search(omr_store* s, const float* queries, int32_t nq, int32_t dim, const omr_filter* filters,
               int32_t n_filters, const int32_t* query_filter, int32_t k, omr_hit* hits, int32_t* counts);

/* Same query as omr_search with struct-of-arrays results and no uuid lookup: rows [nq][k]
 * global row ids (-1 pad), distances [nq][k] (+inf pad).  For callers that keep their own
 * row -> id table (the spliced-SQL drop-in passes row ids straight to `unnest($ids, $dists)`). */
OMR_API int omr_search_rows(omr_store* s, const float* queries, int32_t nq, int32_t dim, const omr_filter* filters,
                            int32_t n_filters, const int32_t* query_filter, int32_t k, int64_t* rows,
                            double* distances, int32_t* counts);

/* Pipelined form of omr_search_rows for a serving loop: _begin stages the batch (host->device copy,
 * bitmaps, search, device->host read-back) on an internal stream and returns at once; _finish(ticket)
 * waits for it and fills the caller's arrays.  Up to four batches may be in flight, so the copies and the
 * select / rescore tail of one batch overlap the scorer of the next.  `queries` must stay valid until
 * _finish when it is page-locked memory (it is then copied from directly); pageable memory is copied
 * before _begin returns.  Filters are consumed before _begin returns. */
OMR_API int omr_search_rows_begin(omr_store* s, const float* queries, int32_t nq, int32_t dim,
                                  const omr_filter* filters, int32_t n_filters, const int32_t* query_filter,
                                  int32_t k, int32_t* ticket);
OMR_API int omr_search_rows_finish(omr_store* s, int32_t ticket, int64_t* rows, double* distances,
                                   int32_t* counts);

/* The whole cosine arm of the hybrid query: cosine_ann (top 4*fanout observations) ->
 * cosine (nearest observation per entity) -> cosine_ranked (LIMIT fanout)
 * (internal/memory/store.go:240-272).  hits[i] is the entity's nearest observation,
 * rank = i + 1. */
OMR_API int omr_cosine_ranked(omr_store* s, const float* query, int32_t dim, const omr_filter* filter, int32_t fanout,
                      omr_hit* hits, int32_t* count);

/* ---- consolidation near-duplicate scan ---------------------------------------
 * Replaces BuildEntityDuplicateCandidatesQuery + RunEntityDuplicateCandidates
 * (internal/memory/consolidation/prefilter.go:112-138,
 * internal/memory/postgres/consolidation_prefilter.go:176-203).  An entity is represented
 * by its latest embedded observation (the row UpdateEmbedding targets). */
typedef struct omr_selfjoin_opts {
  const char* workspace_id;
  float similarity_floor;    /* PreFilterOptions.SimilarityFloor, a Go float32 (prefilter.go:18) */
  int32_t limit;             /* MaxBucketsPerPass; <= 0 = no limit */
  int32_t require_same_kind; /* e1.kind = e2.kind */
  int32_t require_mutable;   /* both mutability = 'mutable' */
  int32_t part_index;        /* block partition of the pair space: this call scans part part_index ... */
  int32_t part_count;        /* ... of part_count (0 or 1 = the whole join).  With the matrix replicated on G
                                GPUs, rank r runs part r of G and the G pair lists are concatenated
                                (SURVEY.md section 8e); each part applies `limit` to its own pairs. */
  const volatile int32_t* cancel; /* NULL, or a flag the caller may raise from another thread while the call runs (the
                                Go binding's ctx.Done watcher): the join is launched in chunks of ~50 ms and returns
                                OMR_E_CANCELLED at the next chunk boundary.  Must stay valid until the call returns. */
} omr_selfjoin_opts;

typedef struct omr_pair {
  uint8_t canonical_id[16]; /* the smaller entity uuid (e1.id < e2.id) */
  uint8_t dupe_id[16];
  double similarity;
  int64_t canonical_row, dupe_row;
  uint32_t kind;
} omr_pair;
/* Writes min(found, cap) pairs ordered by similarity DESC (ties: by canonical row, then dupe row); *found is the
 * total.  `ORDER BY similarity DESC LIMIT n` runs on the device (running cut in the candidate pass, exact rescore,
 * radix selection of the n best), so with a limit only n pairs ever reach the host.  The store's rows cannot
 * change while a join runs (calls that would change them wait); recall calls proceed beside it. */
OMR_API int omr_selfjoin(omr_store* s, const omr_selfjoin_opts* opts, omr_pair* out, int64_t cap, int64_t* found);
/* The same join with struct-of-arrays results and no uuid lookup: global row ids of the canonical and the dupe
 * observation and their similarity.  What the multi-GPU host code gathers from the ranks (SURVEY.md section 8e:
 * counts, then payload) before the owner of the identity maps turns rows into uuids. */
OMR_API int omr_selfjoin_rows(omr_store* s, const omr_selfjoin_opts* opts, int64_t* canonical_rows, int64_t* dupe_rows,
                              double* similarities, int64_t cap, int64_t* found);

/* Identity of a resident row: the uuid of the entity that owns global row `row` and (kind may be NULL) its interned
 * kind (omr_kind_name turns it into the string).  For callers that hold rows from omr_search_rows /
 * omr_selfjoin_rows and need the uuids the reference's records carry (Bucket.Entries,
 * internal/memory/postgres/consolidation_prefilter.go:176-203). */
OMR_API int omr_row_entity(omr_store* s, int64_t row, uint8_t entity_id[16], uint32_t* kind);

/* ---- Memory Galaxy projection input: PCA over resident rows ----------
 * The reference loads one row per live entity (LoadProjectionInputs,
 * internal/memory/projection_store.go:33-62), keeps at most Cap of them and reduces their
 * embeddings to PCADims columns (PCAReduce, ee/pkg/memory/projection/pca.go:28-67: gonum centres
 * the columns, takes an SVD and multiplies the UNCENTRED matrix by the first k right singular
 * vectors, in float64).  The parts that grow with the row count run on the GPU from the fp32 master
 * rows; the D x D symmetric eigenproblem in between does not depend on the row count and is the
 * caller's (gonum EigenSym in Go, LAPACK in the Python mirror).
 *
 * omr_projection_rows: the newest eligible row of each entity that has one under `filter`
 * (single-tier scope semantics; rows without an embedding never qualify), in ascending row order.
 * Writes min(found, cap) rows. */
OMR_API int omr_projection_rows(omr_store* s, const omr_filter* filter, int64_t* rows, int64_t cap, int64_t* found);
/* mean[dim] and the sample covariance cov[dim][dim] (divisor n - 1, as gonum's stat.PC) of the
 * listed rows, float64.  n >= 2; rows are store row ids (as omr_search_rows returns with row_stride 1). */
OMR_API int omr_pca_moments(omr_store* s, const int64_t* rows, int64_t n, double* mean, double* cov);
/* out[n][k] = X * basis with X the listed fp32 rows widened to float64 (NOT centred, as the
 * reference's proj.Mul(m, vecs)); basis is [dim][k] row-major. */
OMR_API int omr_pca_project(omr_store* s, const int64_t* rows, int64_t n, const double* basis, int32_t k, double* out);

/* ---- device-pointer entry points ----------------------------------------------
 * Same recall, with queries, bitmaps and results in device memory and the work enqueued
 * on `stream`.  d_bitmaps is [n_groups][omr_bitmap_words()] (NULL = every live row),
 * d_query_group [nq] (NULL = group 0), d_max_distance [nq] (NULL = no cutoff).
 * Outputs: d_rows [nq][k] global row ids (-1 pad), d_dist [nq][k] (+inf pad), d_count [nq].
 * The call returns after the last kernel is enqueued and the (rare) exact-fallback
 * decision is read back, i.e. it synchronises the stream once. */
OMR_API int64_t omr_bitmap_words(const omr_store* s);
OMR_API int omr_device_build_bitmaps(omr_store* s, const omr_filter* filters, int32_t n_filters, uint32_t* d_bitmaps,
                             void* stream);
OMR_API int omr_device_search(omr_store* s, const float* d_queries, int32_t nq, const uint32_t* d_bitmaps, int32_t n_groups,
                      const int32_t* d_query_group, const double* d_max_distance, int32_t k, int64_t* d_rows,
                      double* d_dist, int32_t* d_count, void* stream);
/* Pipelined form of omr_device_search, for servers that keep the GPU busy with back-to-back
 * batches: _begin enqueues the batch on an internal stream (ordered after what `stream` already
 * holds) and returns at once; _finish(ticket) waits for that batch, runs the exact fallback if a
 * certificate failed, and orders `stream` after the results.  Up to four batches may be in flight, so
 * the select / rescore tail of batch i runs underneath the scorer of batch i+1.  Inputs, bitmaps and
 * outputs of a batch must stay untouched between its _begin and its _finish. */
OMR_API int omr_device_search_begin(omr_store* s, const float* d_queries, int32_t nq, const uint32_t* d_bitmaps,
                                    int32_t n_groups, const int32_t* d_query_group, const double* d_max_distance,
                                    int32_t k, int64_t* d_rows, double* d_dist, int32_t* d_count, void* stream,
                                    int32_t* ticket);
OMR_API int omr_device_search_finish(omr_store* s, int32_t ticket);
/* ---- fused candidate exchange over NVLink peer memory -------------------------------------
 * Row-sharded stores, one process per GPU.  Instead of an NCCL allgather issued by
 * the host, every rank's local answer is stored straight into every rank's exchange buffer by a kernel
 * that runs right behind the local search (st.global on CUDA-IPC mapped peer memory = NVLink writes,
 * published with a system-scope release flag), and the merge kernel waits on those flags on the device.
 *   1. every rank: omr_exchange_alloc -> omr_ipc_export of the returned buffer; the 64-byte handles are
 *      exchanged by the host code (any transport) and passed to omr_exchange_setup, [world][64] in rank order;
 *   2. per batch, on every rank, with the SAME slot sequence: omr_device_search_exchange_begin(..., slot, ...)
 *      then omr_device_search_exchange_finish(ticket, outputs): d_rows/d_dist/d_count receive the GLOBAL
 *      top-k once `stream` (the one given to _begin) drains.
 * Use at least 2 * (batches in flight) + 1 slots so a slot is never rewritten before every rank merged it.
 * A merge that waits ~4 s for a peer raises an error reported by the next call instead of hanging. */
OMR_API int omr_exchange_alloc(omr_store* s, int32_t world, int32_t nq_max, int32_t k_max, int32_t slots,
                               void** d_buf, int64_t* bytes);
OMR_API int omr_ipc_export(const void* d_ptr, uint8_t handle[64]);
OMR_API int omr_exchange_setup(omr_store* s, int32_t rank, int32_t world, const uint8_t* handles);
OMR_API int omr_device_search_exchange_begin(omr_store* s, const float* d_queries, int32_t nq,
                                             const uint32_t* d_bitmaps, int32_t n_groups,
                                             const int32_t* d_query_group, const double* d_max_distance, int32_t k,
                                             int32_t slot, void* stream, int32_t* ticket);
OMR_API int omr_device_search_exchange_finish(omr_store* s, int32_t ticket, int64_t* d_rows, double* d_dist,
                                              int32_t* d_count);
/* Row-sharded stores: merges the G per-shard answers gathered into [G][nq][k]
 * (the gather itself is an NCCL allgather done by the host code).
 * shard_stride = elements between consecutive shards' [nq][k] blocks (0 = nq*k), so rows and
 * distances packed into ONE gathered buffer [G][2][nq][k] can be merged in place. */
OMR_API int omr_device_merge_topk(omr_store* s, const int64_t* d_in_rows, const double* d_in_dist, int32_t n_shards,
                          int32_t nq, int32_t k, int64_t shard_stride, int64_t* d_rows, double* d_dist, int32_t* d_count, void* stream);

/* ---- allocation-free exact top-k ------------------------------------------------
 * No counterpart in the reference.  For hosts that cannot allocate device memory while they
 * serve (all of HBM already belongs to another framework's pool; a latency-critical thread
 * that must not take cudaMalloc's device-wide stall; a fixed start-up memory budget): the
 * exact `ORDER BY row <=> query LIMIT k` of ONE query over the caller's own fp32 matrix
 * d_rows [n][row_pitch] (first `dim` floats of each row are the vector), with the caller's
 * result buffers d_out_rows [k] (-1 pad), d_out_dist [k] (+inf pad) and d_out_count [1].
 * Rows without a cosine (zero vectors) never rank.  The one-distance-per-row scratch is a fixed
 * array inside the library (OMR_NOALLOC_MAX_ROWS * 8 + 256 bytes, resident on every GPU that
 * loads the library until exit), so n and k are capped and a larger request is refused with
 * OMR_E_CAPACITY.  The call makes no cudaMalloc / cudaFree and calls no allocating routine.
 * All calls on a GPU share that array: calls whose GPU work overlaps race, so serialise them.
 * The first call on each device resolves the array's address; make it before any stream
 * capture on that device.  Work is enqueued on `stream`; results are ready when it drains. */
#define OMR_NOALLOC_MAX_ROWS 524288
#define OMR_NOALLOC_MAX_K 128
OMR_API int omr_exact_topk_noalloc(const float* d_rows, int64_t n, int32_t dim, int64_t row_pitch,
                                   const float* d_query, int32_t k, int64_t* d_out_rows, double* d_out_dist,
                                   int32_t* d_out_count, void* stream);

/* ---- introspection ---------------------------------------------------------------*/
typedef struct omr_stats {
  int64_t kernel_launches;   /* kernels this store has enqueued */
  int64_t searches;          /* query batches */
  int64_t queries;           /* queries */
  int64_t rescued_rows;      /* rows rescored beyond the first round (emission mode) / because their sub-bucket was suspect */
  int64_t fallback_queries;  /* queries answered by the exact scan */
  int32_t last_path;         /* 1 streaming, 2 tensor-core, 3 exact */
  int32_t last_sub_bucket;
  int32_t last_rescore_width;
  int32_t num_sms;
  double scorer_ms_sum;      /* device time of the approximate scorer, summed (profiling on, key-table paths: stream
                                events around the kernel; emission mode reports scorer_kernel_ms_sum instead) */
  int64_t scorer_timed;      /* batches that sum covers */
  int64_t host_calls;        /* omr_search / omr_search_rows calls */
  int64_t host_batches;      /* searches they were coalesced into (concurrent calls share one pass) */
  int64_t listed_entries;    /* candidate-list entries the tensor-core scorers emitted (emission mode) */
  double phase_ms_sum[3];    /* unused since the seed pass, the seed thresholds and the scorer are chained by
                                programmatic dependent launch (no stream event fits between them); always 0 */
  double scorer_kernel_ms_sum; /* profiling on, emission mode: the scorer kernel's own span (first CTA start to last
                                  CTA end, %globaltimer), summed; it excludes the time the kernel queued behind
                                  another batch in flight.  The stamps travel with the batch's counters, so
                                  profiling adds no stream operation to a batch */
  int64_t scorer_kernel_timed;
  int64_t join_candidates;   /* candidate pairs the self-join's tensor-core pass emitted (before the exact rescore) */
  double scorer_start_skew_ms_sum; /* latest minus earliest CTA start of those launches, summed */
  double scorer_end_skew_ms_sum;   /* latest minus earliest CTA end, summed */
  double scorer_gap_ms_sum;        /* idle between the end of one batch's scorer and the first CTA of the next, summed */
  double post_span_ms_sum;         /* first block start to last block end of the finishing kernel, summed */
  double post_end_after_scorer_ms_sum; /* how long after its scorer's end a batch's finishing kernel ended, summed */
} omr_stats;
OMR_API int omr_get_stats(const omr_store* s, omr_stats* out);
/* Brackets the approximate scorer (the dominant kernel) with CUDA events on the search stream
 * and accumulates its device time in omr_stats; off by default.  Resets the accumulators. */
OMR_API int omr_set_profiling(omr_store* s, int32_t on);
OMR_API const char* omr_kind_name(const omr_store* s, uint32_t kind);
/* Test hook, no reference counterpart: runs query preparation and ONE approximate scorer
 * (path 1 = streaming, 2 = tensor-core, 4 = tensor-core with CTA pairs) over every live row and returns the raw packed
 * candidate keys, [nq][ceil(n/256) * (256/S) * 2] int32 on the host.  A key is the
 * order-preserving int32 image of the bf16-accurate cosine with its low 8 bits replaced by
 * the row's offset inside its S-row sub-bucket; INT32_MIN = no eligible row. */
/* Test hook: mean device time (ms) of `iters` back-to-back launches of ONE scorer (path 1, 2 or 4)
 * over every live row, the queries being the store's own first nq rows; nothing downstream runs. */
OMR_API int omr_debug_time_scorer(omr_store* s, int32_t nq, int32_t path, int32_t S, int32_t iters, float* ms_out);
/* Test hook: per-query clock64 stamps [nq][8] of the fused post-processing kernel's phases from the
 * last search (only recorded when the store was opened with OMR_POST_TIMING=1 in the environment). */
OMR_API int omr_debug_post_timing(omr_store* s, int64_t* out, int32_t nq);
OMR_API int omr_debug_candidate_keys(omr_store* s, const float* queries, int32_t nq, int32_t path, int32_t S,
                                     int32_t* out_keys, int64_t out_len);

#ifdef __cplusplus
}
#endif
#endif /* OMNIA_RECALL_H_ */
