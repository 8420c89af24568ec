/*
 * oracle.c -- CPU restatement of the reference's semantic-recall arithmetic.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under omnia_b200/ may import, link or
 * execute this file; it is the checker for tests/, __graft_entry__.smoke() and
 * the cpu_baseline / --impl reference legs of bench.py.
 *
 * What it restates
 * ----------------
 * The reference (AltairaLabs/Omnia) computes every cosine score inside the
 * pgvector Postgres extension through the SQL operator `<=>`; the Go code only
 * builds SQL.  pgvector is NOT under /root/reference: the reference pins it
 * only by floating image tags (pgvector/pgvector:pg16 at
 * internal/memory/store_test.go:51, :pg17 at charts/omnia/values.yaml:1300).
 * The distance below therefore follows pgvector's published algorithm
 * (src/vector.c, cosine_distance / VectorCosineSimilarity, v0.7-v0.8 line):
 * three float accumulators, a double divide by sqrt(double*double), a clamp to
 * [-1, 1], and 1 - s.  Parity is anchored on the reference's own call sites:
 *
 *   a1  `<=>`                       every site listed in SURVEY.md section 0
 *   a3  FindSimilarObservations     internal/memory/store_read.go:243-306
 *   a4  cosine_ann / cosine_ranked  internal/memory/store.go:240-272
 *   a8  row eligibility             internal/memory/store.go:250-258,
 *                                   internal/memory/store_read.go:271-278,
 *                                   internal/memory/retrieve_multi_tier.go:313-328
 *   a11 near-duplicate self-join    internal/memory/consolidation/prefilter.go:112-138
 *
 * Parity status: ordering, threshold and scope semantics are pinned by the
 * reference tests transcribed in tests/test_oracle_kat.py.  Numeric score
 * values are "parity unpinned" (no reference test checks a score beyond
 * `> 0.99`), and so are self-join pair sets (no executing reference test).
 *
 * Two accumulation variants are provided because upstream builds its loop with
 * -fassociative-math, so its summation order is compiler dependent:
 *   variant 0  fp32, strictly sequential (this file is compiled WITHOUT
 *              fast-math, so the order below is the order executed)
 *   variant 1  fp64 accumulation (ground truth for the score tolerance)
 * oracle_vec.c holds the same fp32 loop compiled with upstream's vectorising
 * flags; it is the timed CPU baseline, not the parity checker.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define OMO_USER_ANY 0        /* $n IS NULL  -> no constraint (single-tier) */
#define OMO_USER_NULL 1       /* column IS NULL (multi-tier, empty request)  */
#define OMO_USER_EQ 2         /* column = x (single-tier, value given)       */
#define OMO_USER_NULL_OR_EQ 3 /* column IS NULL OR column = x (multi-tier)   */

#define OMO_FLAG_FORGOTTEN 1u
#define OMO_FLAG_SUPERSEDED 2u
#define OMO_FLAG_HAS_EMBEDDING 4u
#define OMO_FLAG_MUTABLE 8u

/* ---- a1: pgvector cosine distance ------------------------------------- */

double omo_cosine_distance_f32(const float *a, const float *b, int dim) {
  float dot = 0.0f, na = 0.0f, nb = 0.0f;
  for (int i = 0; i < dim; i++) {
    dot += a[i] * b[i];
    na += a[i] * a[i];
    nb += b[i] * b[i];
  }
  double s = (double)dot / sqrt((double)na * (double)nb);
  if (s > 1.0) s = 1.0;
  else if (s < -1.0) s = -1.0;
  return 1.0 - s; /* NaN when either norm is zero */
}

double omo_cosine_distance_f64(const float *a, const float *b, int dim) {
  double dot = 0.0, na = 0.0, nb = 0.0;
  for (int i = 0; i < dim; i++) {
    dot += (double)a[i] * (double)b[i];
    na += (double)a[i] * (double)a[i];
    nb += (double)b[i] * (double)b[i];
  }
  double s = dot / sqrt(na * nb);
  if (s > 1.0) s = 1.0;
  else if (s < -1.0) s = -1.0;
  return 1.0 - s;
}

double omo_cosine_distance_vec(const float *a, const float *b, int dim); /* oracle_vec.c */

static double dist_variant(const float *a, const float *b, int dim, int variant) {
  if (variant == 1) return omo_cosine_distance_f64(a, b, dim);
  if (variant == 2) return omo_cosine_distance_vec(a, b, dim);
  return omo_cosine_distance_f32(a, b, dim);
}

/* ---- a8: which rows a query may see ----------------------------------- */

typedef struct {
  const uint32_t *workspace;   /* interned id per row */
  const uint32_t *user;        /* 0 = NULL */
  const uint32_t *agent;       /* 0 = NULL */
  const uint16_t *kind;
  const uint16_t *purpose;
  const float *confidence;
  const uint8_t *flags;
  const int64_t *valid_until_us; /* INT64_MAX = NULL */
} omo_rows;

typedef struct {
  uint32_t workspace;
  uint32_t user;
  int32_t user_mode;
  uint32_t agent;
  int32_t agent_mode;
  const uint16_t *kinds;
  int32_t n_kinds; /* 0 = no kind constraint */
  const uint16_t *purposes;
  int32_t n_purposes;
  float min_confidence;
  int64_t now_us;
} omo_filter;

static int scope_ok(uint32_t col, uint32_t want, int mode) {
  switch (mode) {
    case OMO_USER_ANY: return 1;
    case OMO_USER_NULL: return col == 0;
    case OMO_USER_EQ: return col != 0 && col == want;
    case OMO_USER_NULL_OR_EQ: return col == 0 || col == want;
  }
  return 0;
}

int omo_row_eligible(const omo_rows *r, int64_t i, const omo_filter *f) {
  if (r->workspace[i] != f->workspace) return 0;
  uint8_t fl = r->flags[i];
  if (fl & OMO_FLAG_FORGOTTEN) return 0;
  if (fl & OMO_FLAG_SUPERSEDED) return 0;
  if (!(fl & OMO_FLAG_HAS_EMBEDDING)) return 0;
  if (!(r->valid_until_us[i] > f->now_us)) return 0; /* NULL is INT64_MAX */
  if (!(r->confidence[i] >= f->min_confidence)) return 0;
  if (!scope_ok(r->user[i], f->user, f->user_mode)) return 0;
  if (!scope_ok(r->agent[i], f->agent, f->agent_mode)) return 0;
  if (f->n_kinds > 0) {
    int hit = 0;
    for (int j = 0; j < f->n_kinds; j++) hit |= (r->kind[i] == f->kinds[j]);
    if (!hit) return 0;
  }
  if (f->n_purposes > 0) {
    int hit = 0;
    for (int j = 0; j < f->n_purposes; j++) hit |= (r->purpose[i] == f->purposes[j]);
    if (!hit) return 0;
  }
  return 1;
}

void omo_eligible(const omo_rows *r, int64_t n, const omo_filter *f, uint8_t *out) {
  for (int64_t i = 0; i < n; i++) out[i] = (uint8_t)omo_row_eligible(r, i, f);
}

/* ---- a3 / a4: ORDER BY d ASC LIMIT k over the eligible rows ------------ */

typedef struct {
  double d;
  int64_t row;
} hit_t;

/* "worse" = sorts later under (distance asc, row asc).  The row tie-break is
 * this oracle's choice: SQL leaves tie order unspecified. */
static int hit_worse(const hit_t *a, const hit_t *b) {
  if (a->d != b->d) return a->d > b->d;
  return a->row > b->row;
}

static void heap_sift_down(hit_t *h, int n, int i) {
  for (;;) {
    int l = 2 * i + 1, r = l + 1, m = i;
    if (l < n && hit_worse(&h[l], &h[m])) m = l;
    if (r < n && hit_worse(&h[r], &h[m])) m = r;
    if (m == i) return;
    hit_t t = h[i]; h[i] = h[m]; h[m] = t;
    i = m;
  }
}

static int hit_cmp(const void *pa, const void *pb) {
  const hit_t *a = pa, *b = pb;
  if (hit_worse(a, b)) return 1;
  if (hit_worse(b, a)) return -1;
  return 0;
}

/* Returns the number of hits written (<= k).  max_distance is applied as
 * `d <= max_distance` in float8, as at store_read.go:260,278; pass +inf for
 * the recall paths that have no cutoff.  Zero-norm rows give NaN, which
 * Postgres sorts after every number; they are dropped here (the HNSW plan
 * never returns them either). */
int omo_search_one(const float *rows, int64_t n, int dim, int64_t row_pitch, const uint8_t *eligible,
                   const float *q, int k, double max_distance, int variant, int64_t *out_rows,
                   double *out_dist) {
  if (k <= 0) return 0;
  hit_t *heap = (hit_t *)malloc(sizeof(hit_t) * (size_t)k);
  int cnt = 0;
  for (int64_t i = 0; i < n; i++) {
    if (eligible && !eligible[i]) continue;
    double d = dist_variant(rows + i * row_pitch, q, dim, variant);
    if (isnan(d)) continue;
    if (!(d <= max_distance)) continue;
    hit_t h = {d, i};
    if (cnt < k) {
      heap[cnt++] = h;
      if (cnt == k)
        for (int j = k / 2 - 1; j >= 0; j--) heap_sift_down(heap, k, j);
    } else if (hit_worse(&heap[0], &h)) {
      heap[0] = h;
      heap_sift_down(heap, k, 0);
    }
  }
  qsort(heap, (size_t)cnt, sizeof(hit_t), hit_cmp);
  for (int j = 0; j < cnt; j++) {
    out_rows[j] = heap[j].row;
    out_dist[j] = heap[j].d;
  }
  free(heap);
  return cnt;
}

/* Batch form; eligible is [n_groups][n] bytes and query_group picks the row
 * (NULL -> group 0).  threads <= 0 means "all the host offers". */
void omo_search_batch(const float *rows, int64_t n, int dim, int64_t row_pitch, const uint8_t *eligible,
                      const int32_t *query_group, const float *queries, int nq, int k,
                      double max_distance, int variant, int threads, int64_t *out_rows,
                      double *out_dist, int32_t *out_count) {
#ifdef _OPENMP
  if (threads > 0) omp_set_num_threads(threads);
#else
  (void)threads;
#endif
#pragma omp parallel for schedule(dynamic, 1)
  for (int qi = 0; qi < nq; qi++) {
    const uint8_t *el = eligible ? eligible + (int64_t)(query_group ? query_group[qi] : 0) * n : NULL;
    for (int j = 0; j < k; j++) {
      out_rows[(int64_t)qi * k + j] = -1;
      out_dist[(int64_t)qi * k + j] = INFINITY;
    }
    out_count[qi] = omo_search_one(rows, n, dim, row_pitch, el, queries + (int64_t)qi * dim, k,
                                   max_distance, variant, out_rows + (int64_t)qi * k,
                                   out_dist + (int64_t)qi * k);
  }
}

int omo_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* ---- a11: near-duplicate self-join ------------------------------------ */

/* prefilter.go:113-131: pairs (e1, e2) of one workspace with equal kind,
 * e1.id < e2.id in uuid order (Postgres compares uuids with memcmp), both
 * mutable, both embedded, and 1.0 - (e1 <=> e2) >= floor.  The floor arrives
 * as a Go float32 (prefilter.go:18) and is compared in float8, so it is
 * widened exactly once here.  `member` marks the rows that belong to the
 * workspace and carry an embedding; kind/mutable come from the per-row arrays.
 * Emits (canonical row, dupe row, similarity); returns the pair count, or
 * -(needed) if cap is too small.  ORDER BY / LIMIT is applied by the caller. */
int64_t omo_selfjoin(const float *rows, int64_t n, int dim, int64_t row_pitch, const uint8_t *member,
                     const uint16_t *kind, const uint8_t *flags, const uint8_t *ids16,
                     float floor_f32, int require_same_kind, int require_mutable, int variant,
                     int64_t cap, int64_t *out_a, int64_t *out_b, double *out_sim) {
  const double floor_d = (double)floor_f32;
  int64_t cnt = 0;
#pragma omp parallel for schedule(dynamic, 16)
  for (int64_t i = 0; i < n; i++) {
    if (member && !member[i]) continue;
    if (require_mutable && !(flags[i] & OMO_FLAG_MUTABLE)) continue;
    for (int64_t j = i + 1; j < n; j++) {
      if (member && !member[j]) continue;
      if (require_mutable && !(flags[j] & OMO_FLAG_MUTABLE)) continue;
      if (require_same_kind && kind[i] != kind[j]) continue;
      int c = memcmp(ids16 + 16 * i, ids16 + 16 * j, 16);
      if (c == 0) continue; /* e1.id < e2.id can never hold for one entity */
      double d = dist_variant(rows + i * row_pitch, rows + j * row_pitch, dim, variant);
      double sim = 1.0 - d;
      if (!(sim >= floor_d)) continue;
      int64_t slot;
#pragma omp atomic capture
      slot = cnt++;
      if (slot < cap) {
        out_a[slot] = c < 0 ? i : j;
        out_b[slot] = c < 0 ? j : i;
        out_sim[slot] = sim;
      }
    }
  }
  return cnt <= cap ? cnt : -cnt;
}
