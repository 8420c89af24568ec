// Launchers for the recall engine's kernels.  Each launcher enqueues on the given
// stream and returns 0 or a negative omr error code; none of them allocates.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace omr {

// ---- ingest / query preparation (k_ingest.cu) ------------------------------
// rows land at row0 + i, or at dst_rows[i] when given (device array); bad_flag as launch_prep_queries
int launch_ingest_rows(const float* src, int64_t n, int dim, int pitch, int64_t row0, const int64_t* dst_rows,
                       __nv_bfloat16* rows_bf16, float* rows_f32, double* norm2, int32_t* bad_flag, cudaStream_t st);
// bad_flag (nullable): set to 1 / 2 when a query holds a NaN / an infinite element
int launch_prep_queries(const float* q, int nq, int qpad, int dim, int pitch, __nv_bfloat16* qn, double* qnorm2,
                        int32_t* bad_flag, cudaStream_t st);
constexpr uint32_t kAnyWorkspaceId = 0xFFFFFFFFu;  // DeviceFilter.workspace wildcard
int launch_build_bitmaps(RowMetaView meta, int64_t n, int words_per_group, const DeviceFilter* filters,
                         int n_filters, uint32_t* bitmaps, cudaStream_t st);

// ---- approximate scoring -> per-sub-bucket top-2 keys ------------------------
// cand_keys is [qpad][n_tiles * (256 / S) * 2]; S = sub-bucket rows.
int launch_score_stream(const ShardView& sh, const __nv_bfloat16* qn, int nq, int n_tiles, int S,
                        const uint32_t* bitmaps, int words_per_group, const int32_t* query_group,
                        int32_t* cand_keys, int num_sms, cudaStream_t st);
int launch_score_umma(const CUtensorMap& tmap_q, const CUtensorMap& tmap_rows, int nq, int n_qtiles, int n_tiles,
                      int k_blocks, int S, const uint32_t* bitmaps, int words_per_group,
                      const int32_t* query_group, int32_t* cand_keys, int num_sms, cudaStream_t st);
int launch_score_umma2(const CUtensorMap& tmap_q, const CUtensorMap& tmap_rows_half, int nq, int n_qtiles256, int n_tiles,
                       int k_blocks, int S, const uint32_t* bitmaps, int words_per_group,
                       const int32_t* query_group, int32_t* cand_keys, int num_sms, cudaStream_t st);
int score_umma_smem_bytes();

// ---- threshold-filtered candidate emission (tensor-core scorers, k <= 256) ----------
// Instead of two keys per (query, sub-bucket) the scorer appends to a short per-query list every
// eligible row whose approximate score clears a per-query threshold that only ever rises:
//   thr[q]   = (a lower bound of the k-th best approximate score of query q) - band, stored as its
//              order-preserving int32 image and raised with atomicMax;
//   list[q]  = (biased score << 32 | row) entries, 0 = slot reserved but not written yet;
//   count[q] = entries appended (beyond `cap` the query is answered by the exact scan).
// band = 2 * margin: a row of the exact top-k scores at least T - 2*margin approximately (T = the k-th best
// approximate score), so a threshold that never exceeds T - band never drops an answer row.
struct EmitView {
  int32_t* thr = nullptr;
  unsigned int* count = nullptr;
  unsigned long long* list = nullptr;
  int cap = 0;      // entries per query
  int k = 0;
  int trigger = 32; // power of two: a threshold refresh is requested every `trigger` appended entries
  float band = 0.0f;
};
// Seeds thr[] from the top-2-per-32-row keys of the first `seed_tiles` row tiles and zeroes count[].
// Queries q >= nq (tile padding) and zero-norm queries get a threshold nothing clears.
// `pdl` (here and in the two scorer launchers below): launch with programmatic stream serialization, so the kernel is
// scheduled while the one before it on `st` still runs and waits on the device for that kernel's results.  Only
// for the chain seed scorer -> seed thresholds -> full-store scorer, with nothing else enqueued between them.
int launch_emit_seed(const int32_t* seed_keys, int64_t keys_per_query, const double* qnorm2, int nq, int qpad,
                     const EmitView& ev, bool pdl, cudaStream_t st);
int launch_score_umma_emit(const CUtensorMap& tmap_q, const CUtensorMap& tmap_rows, int nq, int n_qtiles, int n_tiles,
                           int k_blocks, const uint32_t* bitmaps, int words_per_group, const int32_t* query_group,
                           const EmitView& ev, unsigned long long* timer, int num_sms, bool pdl, cudaStream_t st);
int launch_score_umma2_emit(const CUtensorMap& tmap_q, const CUtensorMap& tmap_rows_half, int nq, int n_qtiles256,
                            int n_tiles, int k_blocks, const uint32_t* bitmaps, int words_per_group,
                            const int32_t* query_group, const EmitView& ev, unsigned long long* timer, int num_sms,
                            bool pdl, cudaStream_t st);
// `timer` (nullable, device, 4 words): [0] receives the earliest and [2] the latest CTA start, [1] the latest and
// [3] the earliest CTA end of the launch in %globaltimer nanoseconds ([0] and [3] must start at ~0ull, [1] and [2]
// at 0): the kernel's own duration even when it queued behind another batch's kernels, which stream events
// around it cannot tell apart, and how ragged its start and its end were.
// Per query: keep the list entries within `band` of the k-th best approximate score, rescore them exactly
// from the fp32 master (best K2 first, then whatever the exact k-th distance found so far still allows),
// ORDER BY distance LIMIT k with the cutoff, and clear the list for the lane's next batch (k_emit.cu).
struct EmitPostArgs {
  ShardView sh;
  const float* q = nullptr;
  const double* qnorm2 = nullptr;
  EmitView ev;
  int nq = 0, k = 10, K2 = 16;
  int cand = 512;                    // entries inside the band one block can rescore (emit_post_capacity)
  const double* max_dist = nullptr;
  float margin = 0.0f;
  int64_t row_base = 0, row_stride = 1;
  int64_t* out_rows = nullptr;
  double* out_dist = nullptr;
  int32_t* out_count = nullptr;
  int32_t* fail_list = nullptr;
  int32_t* fail_count = nullptr;
  int32_t* rescued_total = nullptr;  // rows rescored in the second round
  int32_t* listed_total = nullptr;   // list entries consumed (diagnostics)
  unsigned long long* timer = nullptr;  // nullable: [0] earliest block start, [1] latest block end (%globaltimer ns)
};
bool emit_supported(int k);
int emit_post_capacity(int pitch);
int launch_emit_post(const EmitPostArgs& a, cudaStream_t st);

// ---- select / exact rescore / finalize (k_select.cu) -------------------------
// Candidate slots of a query are sel_rows/sel_dist[q][K2x]: the first K2 hold the keys
// select_keys picked, the remaining K2x - K2 are "extra" slots that finalize pass 1 fills
// with every eligible row of a suspect sub-bucket.
int launch_select_keys(const int32_t* cand_keys, int nq, int64_t keys_per_query, int S, int K2, int K2x,
                       int32_t* sel_rows, float* sel_bound, cudaStream_t st);
int launch_rescore(const ShardView& sh, const float* q, const double* qnorm2, int nq, int K2, int K2x,
                   const int32_t* sel_rows, double* sel_dist, cudaStream_t st);
int launch_rescore_items(const ShardView& sh, const float* q, const double* qnorm2, int K2x, const int32_t* items,
                         const int32_t* item_count, const int32_t* sel_rows, double* sel_dist, int num_sms,
                         cudaStream_t st);

struct FinalizeArgs {
  int nq = 0, K2 = 0, K2x = 0, k = 0, S = 256;
  int32_t* sel_rows = nullptr;
  const double* sel_dist = nullptr;
  const double* max_dist = nullptr;    // per query; nullptr = no cutoff
  const float* sel_bound = nullptr;
  const int32_t* cand_keys = nullptr;  // nullptr = no certificate (exact fallback results)
  int64_t keys_per_query = 0;
  float margin = 0.0f;
  const uint32_t* bitmaps = nullptr;
  int words_per_group = 0;
  const int32_t* query_group = nullptr;
  int64_t row_base = 0;                // returned row = row_base + local_row * row_stride
  int64_t row_stride = 1;
  int64_t* out_rows = nullptr;
  double* out_dist = nullptr;
  int32_t* out_count = nullptr;
  int32_t* extra_cnt = nullptr;        // [nq]
  int32_t* extra_items = nullptr;      // flat slot indices to rescore, [nq * (K2x - K2)]
  int32_t* extra_total = nullptr;      // device counter
  int32_t* fail_list = nullptr;        // queries that need the exact fallback
  int32_t* fail_count = nullptr;
  const int32_t* qlist = nullptr;      // optional subset of queries
  int nlist = 0;
};
int launch_finalize(const FinalizeArgs& a, cudaStream_t st);
int launch_finalize_extra(const FinalizeArgs& a, cudaStream_t st);

// Fused replacement of select_keys -> rescore -> finalize -> rescore_items -> finalize_extra
// (k_post.cu): one block per query, no intermediate global buffers.  K2 <= 256.
struct PostArgs {
  ShardView sh;
  const float* q = nullptr;            // [nq][dim] fp32 queries as given
  const double* qnorm2 = nullptr;
  const int32_t* cand_keys = nullptr;
  int64_t keys_per_query = 0;
  int nq = 0, S = 256, K2 = 32, k = 10;
  int sort_cap = 2048;                 // in-block merge capacity (candidates + rescued rows); >= 384, even
  const double* max_dist = nullptr;
  float margin = 0.0f;
  const uint32_t* bitmaps = nullptr;
  int words_per_group = 0;
  const int32_t* query_group = nullptr;
  int64_t row_base = 0, row_stride = 1;
  int64_t* out_rows = nullptr;
  double* out_dist = nullptr;
  int32_t* out_count = nullptr;
  int32_t* fail_list = nullptr;
  int32_t* fail_count = nullptr;
  int32_t* rescued_total = nullptr;
  long long* timing = nullptr;          // optional [nq][8] clock64 stamps per phase (diagnostics)
};
bool post_fused_supported(int K2, int k);
int launch_post(const PostArgs& a, cudaStream_t st);

// Exact fallback: distances of `nslots` queries (indices in slot_query) to every row.
int launch_exact_scan(const ShardView& sh, const float* q, const double* qnorm2, const int32_t* slot_query,
                      int nslots, const uint32_t* bitmaps, int words_per_group, const int32_t* query_group,
                      double* dist_out, cudaStream_t st);
int launch_select_exact(const double* dist, int64_t n, const int32_t* slot_query, int nslots, int K2, int K2x, int k,
                        int32_t* sel_rows, double* sel_dist, cudaStream_t st);

// Multi-GPU: merge G gathered per-shard results [G][nq][k] into [nq][k].
int launch_merge_topk(const int64_t* in_rows, const double* in_dist, int G, int nq, int k, int64_t shard_stride,
                      int64_t* out_rows, double* out_dist, int32_t* out_count, cudaStream_t st);

// ---- fused candidate exchange over NVLink peer memory (k_exchange.cu) ---------------
constexpr int kMaxWorld = 16;
struct ExchangeView {
  void* peers[kMaxWorld];  // exchange buffer of every rank as mapped in THIS process (own buffer at [rank])
  int world = 0, rank = 0, slots = 0;
  int64_t slot_stride = 0;  // bytes of one (slot, rank) data region: 2 * nq_max * k_max * 8
  __host__ __device__ size_t data_offset(int slot, int r) const { return ((size_t)slot * world + r) * (size_t)slot_stride; }
  __host__ __device__ size_t flag_offset(int slot, int r) const {
    return (size_t)slots * world * (size_t)slot_stride + ((size_t)slot * world + r) * 8;
  }
  __host__ __device__ size_t total_bytes() const { return flag_offset(slots, 0); }
};
int launch_exchange_scatter(const ExchangeView& x, const int64_t* local, int nq, int k, int slot,
                            unsigned long long epoch, const int32_t* fail_count, unsigned int* done_blocks,
                            cudaStream_t st);
int launch_exchange_merge(const ExchangeView& x, int slot, unsigned long long epoch, int nq, int k,
                          int64_t* out_rows, double* out_dist, int32_t* out_count, unsigned int* error_word,
                          cudaStream_t st);

// ---- near-duplicate self-join (k_score_umma.cu / k_select.cu) ----------------
struct PairCand {
  int32_t a, b;
};
// One launch of the join kernel over a contiguous range of one partition's work items.
struct JoinLaunch {
  int64_t n = 0;            // member rows (compact index space)
  int k_blocks = 0;
  int part_index = 0, part_count = 1;
  float threshold = 0.0f;   // floor - margin
  const uint32_t* member_bitmap = nullptr;
  PairCand* pairs = nullptr;
  unsigned long long cap = 0;
  unsigned long long* pair_count = nullptr;
  // LIMIT mode (limit > 0): running cut from a histogram of the emitted pairs' approximate scores
  int limit = 0;
  int32_t* run_cut = nullptr;      // ordered-int image; starts at kOrderedNegInf
  unsigned int* ghist = nullptr;   // [kJoinHistBins], zeroed
  const uint16_t* kind = nullptr;  // per member row; nullptr = no kind rule
  float hist_lo = 0.0f, hist_scale = 0.0f, band = 0.0f;
  // this launch: items [chunk_first, chunk_first + chunk_items) of the partition (chunk_items < 0 = to the end)
  int64_t chunk_first = 0, chunk_items = -1;
};
constexpr int kJoinHistBins = 2048;
int64_t selfjoin_items(int64_t n, int k_blocks, int part_index, int part_count, bool pair_kernel);
int launch_selfjoin_umma(const CUtensorMap& tmap_a, const CUtensorMap& tmap_b, const JoinLaunch& j, int num_sms,
                         cudaStream_t st);
// the same join on CTA pairs (cta_group::2, A and B tiles of 256 rows); tmap_half has a 128-row box
int launch_selfjoin_umma2(const CUtensorMap& tmap_half, const JoinLaunch& j, int num_sms, cudaStream_t st);
int launch_gather_rows(const ShardView& sh, const int32_t* idx, int64_t m, __nv_bfloat16* dst, cudaStream_t st);
// dst row r = src row idx[r] for rows of row_bytes (a multiple of 8) bytes; dst must not overlap src
int launch_gather_row_bytes(const void* src, size_t row_bytes, const int32_t* idx, int64_t m, void* dst, cudaStream_t st);

// ---- PCA moments / projection over resident rows (k_pca.cu) -------------------
int pca_mean_chunks(int64_t n);
int pca_cov_splits(int64_t n, int dim, int num_sms);
// colsum_partial [n_chunks][dim], cov_partial [n_splits][dim][dim], mean [dim], cov [dim][dim]; n >= 2
int launch_pca_moments(const ShardView& sh, const int64_t* rows, int64_t n, double* colsum_partial, int n_chunks,
                       double* mean, double* cov_partial, int n_splits, double* cov, cudaStream_t st);
// out [n][k] = rows_f32[rows] * basis, basis [dim][k] row-major
int launch_pca_project(const ShardView& sh, const int64_t* rows, int64_t n, const double* basis, int k, double* out,
                       cudaStream_t st);

}  // namespace omr
