"""Structure of the spliced hybrid statements, in the style of the reference's own
DB-free wiring test (internal/memory/retrieve_multi_tier_hybrid_test.go:249-289): the statements the Go
`gpustore` package runs are the reference's with ONLY the three cosine CTEs replaced, the placeholder numbering
is unchanged, and the replacement re-applies the reference's eligibility predicate."""
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
GO = (ROOT / "integration" / "gpustore" / "hybrid_sql.go").read_text()
REF = Path("/root/reference/internal/memory")


def _const(src: str, name: str) -> str:
    m = re.search(r"const\s+" + name + r"\s*=\s*`(.*?)`", src, re.S)
    assert m, f"constant {name} not found"
    return m.group(1)


def _ctes(sql: str):
    """Names of the CTEs of a WITH statement, in order."""
    return re.findall(r"(?:WITH|\),)\s+(\w+)\s+AS\s+\(", sql)


def _strip(sql: str) -> str:
    """SQL text without comments and with whitespace collapsed (for comparisons with the reference's text)."""
    return re.sub(r"\s+", " ", re.sub(r"--[^\n]*", "", sql)).strip()


def _without_cosine_arm(sql: str) -> str:
    """Drops every CTE whose name starts with `cosine` (balanced parentheses), keeps the rest."""
    out, i = "", 0
    for m in re.finditer(r",\s+(cosine\w*)\s+AS\s+\(", sql):
        if m.start() < i:
            continue
        out += sql[i:m.start()]
        depth, j = 1, m.end()
        while depth:
            depth += {"(": 1, ")": -1}.get(sql[j], 0)
            j += 1
        i = j
    return out + sql[i:]


SINGLE = _const(GO, "HybridRetrieveGPUSQL")
MULTI = _const(GO, "HybridMultiTierGPUTemplate")


def test_single_tier_statement_structure():
    assert _ctes(SINGLE) == ["fts", "fts_ranked", "cosine_ranked", "fused"]      # cosine_ann and cosine are gone
    assert "<=>" not in SINGLE and "cosine_ann" not in _strip(SINGLE)
    assert sorted(set(re.findall(r"\$(\d+)", SINGLE)), key=int) == [str(i) for i in range(1, 9)]   # $1..$8 as the reference
    assert "unnest($5::uuid[]) WITH ORDINALITY" in SINGLE                         # $5: embedding -> ranked entity ids
    cos = SINGLE[SINGLE.index("cosine_ranked AS ("):SINGLE.index("), fused AS (")]
    for want in ("e.workspace_id = $1", "($2::text IS NULL OR e.virtual_user_id = $2)",
                 "($3::uuid IS NULL OR e.agent_id = $3)", "e.forgotten = false", "o.superseded_by IS NULL",
                 "(o.valid_until IS NULL OR o.valid_until > now())", "o.embedding IS NOT NULL", "o.confidence >= $8",
                 "row_number() OVER (ORDER BY g.ord) AS cos_rn", "LIMIT $6"):
        assert want in cos, f"re-applied WHERE clause misses {want!r}"
    for want in ("websearch_to_tsquery('english', $4)", "FULL OUTER JOIN cosine_ranked c USING (entity_id)",
                 "1.0/(60.0 + c.cos_rn)", "final_score", "LIMIT $7"):
        assert want in SINGLE


def test_multi_tier_template_structure():
    assert _ctes(MULTI) == ["candidates", "fts", "fts_ranked", "cosine_ranked", "fused"]
    assert "<=>" not in MULTI and "cosine_ann" not in _strip(MULTI)
    assert sorted(set(re.findall(r"%\[(\d)\]", MULTI))) == [str(i) for i in range(1, 8)]          # Sprintf slots 1..7
    assert "unnest($%[5]d::uuid[]) WITH ORDINALITY" in MULTI
    cos = MULTI[MULTI.index("cosine_ranked AS ("):MULTI.index("), fused AS (")]
    assert "%[1]s AND o.embedding IS NOT NULL" in cos                              # the SAME shared predicate
    assert "o.superseded_by IS NULL" in cos and "o.valid_until > now()" in cos and "LIMIT $%[6]d" in cos
    # filled in the way buildMultiTierHybridQuery does, the wiring test's needles must all be there
    where = ("e.workspace_id=$1 AND e.forgotten = false AND (e.virtual_user_id IS NULL OR e.virtual_user_id=$2) "
             "AND (e.agent_id IS NULL OR e.agent_id=$3)")
    sql = MULTI
    for slot, val in ((1, where), (2, "coalesce(1.0/(60.0 + f.fts_rn), 0) + coalesce(1.0/(60.0 + c.cos_rn), 0)"),
                      (3, "exp(-EXTRACT(EPOCH FROM (now() - e.observed_at)) / (CASE WHEN e.virtual_user_id IS NOT NULL THEN $7 END))"),
                      (4, "4"), (5, "5"), (6, "6"), (7, "CASE e.source_type WHEN 'reflection' THEN 0.85 ELSE 0.7 END")):
        sql = sql.replace(f"%[{slot}]s", val).replace(f"%[{slot}]d", val)
    assert "%[" not in sql
    for want in ("websearch_to_tsquery('english'", "FULL OUTER JOIN", "e.forgotten = false",
                 "virtual_user_id IS NULL OR e.virtual_user_id=", "agent_id IS NULL OR e.agent_id=",
                 "EXTRACT(EPOCH FROM (now() - e.observed_at))", "WHEN e.virtual_user_id IS NOT NULL THEN", "final_score",
                 "unnest($5::uuid[])"):
        assert want in sql, f"SQL missing {want!r}"


@pytest.mark.skipif(not REF.exists(), reason="the reference tree is only present in the build container")
def test_everything_but_the_cosine_arm_is_the_references_text():
    ref_single = _const((REF / "store.go").read_text(), "hybridRetrieveSQL")
    ref_multi = _const((REF / "retrieve_multi_tier_hybrid.go").read_text(), "hybridMultiTierTemplate")
    assert _strip(_without_cosine_arm(SINGLE)) == _strip(_without_cosine_arm(ref_single))
    assert _strip(_without_cosine_arm(MULTI)) == _strip(_without_cosine_arm(ref_multi))
    assert _ctes(ref_single) == ["fts", "fts_ranked", "cosine_ann", "cosine", "cosine_ranked", "fused"]
