package gpustore

// The two hybrid statements of the reference with ONLY their cosine arm replaced.
//
// Reference statements: hybridRetrieveSQL (internal/memory/store.go:221-317) and hybridMultiTierTemplate
// (internal/memory/retrieve_multi_tier_hybrid.go:57-118).  In both, the three CTEs `cosine_ann` -> `cosine` ->
// `cosine_ranked` (the pgvector `<=>` scan, the per-entity dedup and the rank list) are what the GPU engine
// computes (omr_cosine_ranked); everything else -- the FTS arm, Reciprocal Rank Fusion, the quality weights,
// the final projection -- needs the text columns and stays in Postgres, byte for byte as the reference has it.
//
// The replacement CTE reads the engine's rank list from ONE array parameter that takes the place of the
// embedding parameter, so the placeholder numbering of the reference is unchanged:
//
//	single tier   $5  was the query embedding (pgvector), is now the ranked entity ids (uuid[])
//	multi tier    $%[5]d  likewise
//
// The CTE re-applies the reference's own eligibility predicate to the ids it is given (workspace, scope columns,
// `forgotten`, an active observation with an embedding and enough confidence), so Postgres stays the authority:
// a device-side flag that lags a write costs a rank slot, never a wrong row (SURVEY.md appendix A).  Ranks are
// re-numbered after that filter, which is what `row_number() OVER (ORDER BY cos_dist)` did on the filtered scan.

// HybridRetrieveGPUSQL is hybridRetrieveSQL with the cosine arm spliced.  Parameters: $1 workspace_id,
// $2 virtual_user_id, $3 agent_id, $4 query text, $5 ranked entity ids (uuid[], nearest first, from
// Store.CosineRanked), $6 fanout, $7 limit, $8 min_confidence.
const HybridRetrieveGPUSQL = `
WITH fts AS (
    SELECT DISTINCT ON (e.id)
        e.id AS entity_id,
        ts_rank_cd(o.search_vector, websearch_to_tsquery('english', $4)) AS fts_rank
    FROM memory_entities e
    JOIN memory_observations o ON o.entity_id = e.id
        AND o.superseded_by IS NULL
        AND (o.valid_until IS NULL OR o.valid_until > now())
    WHERE e.workspace_id = $1
      AND ($2::text IS NULL OR e.virtual_user_id = $2)
      AND ($3::uuid IS NULL OR e.agent_id = $3)
      AND e.forgotten = false
      AND o.confidence >= $8
      AND o.search_vector @@ websearch_to_tsquery('english', $4)
    ORDER BY e.id, fts_rank DESC
    LIMIT $6
), fts_ranked AS (
    SELECT entity_id, row_number() OVER (ORDER BY fts_rank DESC) AS fts_rn
    FROM fts
), cosine_ranked AS (
    -- Rank list computed by the GPU engine (exact cosine top fanout*4 observations, nearest observation per
    -- entity, first fanout entities).  The predicate below is the one the reference's cosine_ann applied.
    SELECT g.entity_id, row_number() OVER (ORDER BY g.ord) AS cos_rn
    FROM unnest($5::uuid[]) WITH ORDINALITY AS g(entity_id, ord)
    JOIN memory_entities e ON e.id = g.entity_id
        AND e.workspace_id = $1
        AND ($2::text IS NULL OR e.virtual_user_id = $2)
        AND ($3::uuid IS NULL OR e.agent_id = $3)
        AND e.forgotten = false
    WHERE EXISTS (
        SELECT 1 FROM memory_observations o
        WHERE o.entity_id = g.entity_id
          AND o.superseded_by IS NULL
          AND (o.valid_until IS NULL OR o.valid_until > now())
          AND o.embedding IS NOT NULL
          AND o.confidence >= $8)
    ORDER BY g.ord
    LIMIT $6
), fused AS (
    -- Reciprocal Rank Fusion. The 60.0 constant is the k from
    -- Cormack 2009 reproduced by Weaviate, Vespa, OpenSearch,
    -- Elastic, et al. — larger flattens per-list contribution,
    -- smaller amplifies the top of each list.
    SELECT
        coalesce(f.entity_id, c.entity_id) AS entity_id,
        coalesce(1.0/(60.0 + f.fts_rn), 0)
          + coalesce(1.0/(60.0 + c.cos_rn), 0) AS rrf
    FROM fts_ranked f FULL OUTER JOIN cosine_ranked c USING (entity_id)
)
SELECT id, kind, metadata, created_at, expires_at, title,
       content, confidence, session_id, turn_range, observed_at, accessed_at,
       summary, body_size_bytes, final_score
FROM (
    SELECT DISTINCT ON (e.id)
        e.id, e.kind, e.metadata, e.created_at, e.expires_at, e.title,
        o.content, o.confidence, o.session_id, o.turn_range, o.observed_at, o.accessed_at,
        o.summary, o.body_size_bytes,
        fused.rrf
            * (CASE e.source_type
                  WHEN 'user_requested'           THEN 1.0
                  WHEN 'operator_curated'         THEN 1.0
                  WHEN 'reflection'               THEN 0.85
                  WHEN 'conversation_extraction'  THEN 0.7
                  WHEN 'system_generated'         THEN 0.5
                  ELSE 0.7 END)
            * o.confidence
            * exp(-EXTRACT(EPOCH FROM (now() - o.observed_at)) / 2592000.0) AS final_score
    FROM fused
    JOIN memory_entities e ON e.id = fused.entity_id
    JOIN memory_observations o ON o.entity_id = e.id
        AND o.superseded_by IS NULL
        AND (o.valid_until IS NULL OR o.valid_until > now())
    WHERE e.forgotten = false
    ORDER BY e.id, o.observed_at DESC
) ranked
ORDER BY final_score DESC
LIMIT $7`

// HybridMultiTierGPUTemplate is hybridMultiTierTemplate with the cosine arm spliced.  Sprintf slots are the
// reference's: %[1]s candidates WHERE (the shared tier-anchored predicate, aliases e and o), %[2]s RRF
// expression, %[3]s per-tier recency-decay expression, $%[4]d query text, $%[5]d ranked entity ids (uuid[];
// the reference passes the embedding here), $%[6]d fanout, %[7]s source-type weight CASE.  The argument list
// buildMultiTierHybridQuery assembles is reused unchanged except for that one value.
const HybridMultiTierGPUTemplate = `
WITH candidates AS (
    SELECT e.id AS entity_id, e.kind, e.metadata, e.created_at, e.expires_at,
           e.title, e.virtual_user_id, e.agent_id, e.source_type,
           o.content, o.confidence, o.session_id, o.turn_range, o.observed_at,
           o.accessed_at, o.access_count, o.summary, o.body_size_bytes,
           o.search_vector, o.embedding
    FROM memory_entities e
    JOIN memory_observations o ON o.entity_id = e.id
        AND o.superseded_by IS NULL
        AND (o.valid_until IS NULL OR o.valid_until > now())
    WHERE %[1]s
), fts AS (
    SELECT DISTINCT ON (entity_id) entity_id,
           ts_rank_cd(search_vector, websearch_to_tsquery('english', $%[4]d)) AS fts_rank
    FROM candidates
    WHERE search_vector @@ websearch_to_tsquery('english', $%[4]d)
    ORDER BY entity_id, fts_rank DESC
    LIMIT $%[6]d
), fts_ranked AS (
    SELECT entity_id, row_number() OVER (ORDER BY fts_rank DESC) AS fts_rn FROM fts
), cosine_ranked AS (
    -- Rank list computed by the GPU engine under the same tier-anchored scope (OMR_MATCH_NULL_ONLY /
    -- OMR_MATCH_NULL_OR_EQ).  The ids are re-checked against the SAME shared predicate the candidates CTE
    -- uses, so the two arms cannot drift and Postgres stays the authority on eligibility.
    SELECT g.entity_id, row_number() OVER (ORDER BY g.ord) AS cos_rn
    FROM unnest($%[5]d::uuid[]) WITH ORDINALITY AS g(entity_id, ord)
    WHERE EXISTS (
        SELECT 1
        FROM memory_entities e
        JOIN memory_observations o ON o.entity_id = e.id
            AND o.superseded_by IS NULL
            AND (o.valid_until IS NULL OR o.valid_until > now())
        WHERE e.id = g.entity_id AND %[1]s AND o.embedding IS NOT NULL)
    ORDER BY g.ord
    LIMIT $%[6]d
), fused AS (
    SELECT coalesce(f.entity_id, c.entity_id) AS entity_id, %[2]s AS rrf
    FROM fts_ranked f FULL OUTER JOIN cosine_ranked c USING (entity_id)
)
SELECT DISTINCT ON (e.entity_id)
    e.entity_id, e.kind, e.metadata, e.created_at, e.expires_at, e.title,
    e.virtual_user_id, e.agent_id,
    e.content, e.confidence, e.session_id, e.turn_range, e.observed_at,
    e.accessed_at, e.access_count, e.summary, e.body_size_bytes,
    fused.rrf * (%[7]s) * coalesce(e.confidence, 0.7) * %[3]s AS final_score
FROM candidates e JOIN fused ON fused.entity_id = e.entity_id
ORDER BY e.entity_id, e.observed_at DESC`
